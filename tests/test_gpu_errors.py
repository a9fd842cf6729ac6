"""GPU: error behaviour of the C-ABI — what the reference's operators refuse with GGML_ASSERT / GGML_STATUS_FAILED comes back here as a
status code plus a message (`eb_last_error`), nothing is launched, and the library keeps working afterwards.  Empty inputs (no rows) are
not errors: they return EB_OK without a launch, as ggml's ops do for tensors with no elements (R/ggml/src/ggml-cpu/ggml-cpu.c, every
`ggml_compute_forward_*` starts from `ggml_nrows`)."""
import numpy as np
import pytest

import fixtures as fx
import pyoracle as po
from golden_io import load_ops

pytestmark = pytest.mark.gpu
HP = dict(E=256, H=2, F=512, V=1024, L=2)
EB_ERR_ARG, EB_ERR_UNSUPPORTED = -1, -3


def _refused(eb, rc, code, text):
    msg = eb.lib().eb_last_error().decode()
    assert rc == code, (rc, msg)
    assert text in msg, msg


def test_operator_entries_refuse_what_they_cannot_compute(eb):
    lib = eb.lib()
    x = eb.DeviceBuffer.from_host(np.ones((4, 384), np.float32))
    y = eb.DeviceBuffer(4 * 64 * 4)
    w = eb.DeviceBuffer(1 << 20)
    ws = eb.DeviceBuffer(lib.eb_mul_mat_ws_size(eb.Q4_K, 64, 512, 4))
    _refused(eb, lib.eb_mul_mat(eb.F16, w.ptr, 64, 256, x.ptr, 4, y.ptr, ws.ptr, None), EB_ERR_UNSUPPORTED, "weight type")
    _refused(eb, lib.eb_mul_mat(eb.Q4_K, w.ptr, 64, 384, x.ptr, 4, y.ptr, ws.ptr, None), EB_ERR_ARG, "not a multiple of 256")
    _refused(eb, lib.eb_mul_mat(eb.Q8_0, w.ptr, 64, 96, x.ptr, 4, y.ptr, ws.ptr, None), EB_ERR_ARG, "not a multiple of 128")
    _refused(eb, lib.eb_mul_mat(eb.Q4_K, w.ptr, 64, 256, x.ptr, 4, y.ptr, None, None), EB_ERR_ARG, "workspace required")
    assert lib.eb_mul_mat(eb.Q4_K, w.ptr, 64, 256, x.ptr, 0, y.ptr, ws.ptr, None) == 0          # no rows: nothing to do
    assert lib.eb_row_size(eb.Q4_K, 100) == 0 and lib.eb_row_size(99, 256) == 0                  # not a whole number of blocks / unknown type
    with pytest.raises(eb.EagleB200Error, match="top_k"):
        eb.top_k(np.zeros((1, 1024), np.float32), 17)
    with pytest.raises(eb.EagleB200Error, match="top_k"):
        eb.top_k(np.zeros((1, 1024), np.float32), 0)
    q = np.zeros((3, 2, 64), np.float32); kv = np.zeros((32, 2, 64), np.float16)
    with pytest.raises(eb.EagleB200Error, match="head size 64"):
        eb.flash_attn_ext(q, kv, kv, 1.0, mask_f16=np.zeros((32, 32), np.float16))
    q = np.zeros((129, 2, 128), np.float32); kv = np.zeros((256, 2, 128), np.float16)       # a tree is at most one 128-row tile
    with pytest.raises(eb.EagleB200Error, match="T=129"):                                    # (f16-mask batches of any length are chunked)
        eb.flash_attn_ext(q, kv, kv, 1.0, tree_bits=np.zeros((129, 4), np.uint32), n_prefix=100)
    with pytest.raises(eb.EagleB200Error, match="rope: D=260"):
        eb.rope_norm(np.zeros((2, 1, 260), np.float32), np.arange(2), 10000.0)


def test_model_and_session_entries_refuse_bad_arguments(eb):
    with pytest.raises(eb.EagleB200Error, match="multiples of 256"):
        eb.Model(200, 2, 512, 1024, 1)
    tgt = fx.make_model(**HP, ftype="q4_k_m", seed=31)
    head = fx.make_model(**{**HP, "L": 1}, ftype=po.Q4_0, seed=32, is_eagle=True)
    mt = fx.to_gpu_model(eb, tgt); mh = fx.to_gpu_model(eb, head, target=mt)
    t, raw = tgt["layers"][0][po.L_WQ]
    with pytest.raises(eb.EagleB200Error, match="bad tensor slot"):
        mt.set_tensor(HP["L"], po.L_WQ, t, raw)
    with pytest.raises(eb.EagleB200Error, match="size"):
        mt.get_tensor(0, po.L_WQ, raw.size - 1)
    incomplete = eb.Model(HP["E"], HP["H"], HP["F"], HP["V"], 1)
    s0 = eb.Session(incomplete, n_ctx=64, max_nodes=8, n_branch=2, n_draft=6)
    with pytest.raises(eb.EagleB200Error, match="missing"):
        s0.forward(0, [1], [0])
    for bad in (dict(max_nodes=129), dict(n_branch=17), dict(top_k=0), dict(top_k=17)):
        with pytest.raises(eb.EagleB200Error, match="bad parameters"):
            eb.Session(mt, mh, n_ctx=64, **{**dict(max_nodes=8, n_branch=2, n_draft=6), **bad})
    other = fx.to_gpu_model(eb, fx.make_model(**{**HP, "E": 512, "L": 1}, ftype=po.Q4_0, seed=33, is_eagle=True))
    with pytest.raises(eb.EagleB200Error, match="shape mismatch"):
        eb.Session(mt, other, n_ctx=64, max_nodes=8, n_branch=2, n_draft=6)

    s = eb.Session(mt, mh, n_ctx=64, max_nodes=8, n_branch=2, n_draft=6, max_depth=3, top_k=4, use_graph=False)
    with pytest.raises(eb.EagleB200Error, match=r"parent\[1\]=1"):
        s.forward(0, [1, 2], [0, 1], parent=[-1, 1])
    with pytest.raises(eb.EagleB200Error, match="T=0"):
        s.forward(0, np.zeros(0, np.int32), np.zeros(0, np.int32))
    with pytest.raises(eb.EagleB200Error, match="hidden input"):
        s.forward(1, [1], [0])
    with pytest.raises(eb.EagleB200Error, match="no draft head"):
        eb.Session(mt, n_ctx=64, max_nodes=8, n_branch=2, n_draft=6).round()
    with pytest.raises(eb.EagleB200Error, match="exceed n_ctx"):
        s.prefill(np.ones(65, np.int32))
    with pytest.raises(eb.EagleB200Error, match="set_n_past"):
        s.set_n_past(65)
    with pytest.raises(eb.EagleB200Error, match="kv: layer"):
        s.kv_read(0, 0, 0, 60, 8)

    # the cache fills up: rounds are refused once a whole tree no longer fits, forwards once the batch does not, and the session is intact
    rng = np.random.default_rng(5)
    prompt = rng.integers(3, HP["V"], 50).astype(np.int32)
    lg, ft = s.prefill(prompt)
    assert s.n_past() == 50
    s.set_root(int(lg.argmax()), ft)
    n_rounds = 0
    with pytest.raises(eb.EagleB200Error, match="KV cache full"):
        for _ in range(16):
            r = s.round(); n_rounds += 1
            assert 1 <= r.n_accept <= 4
    assert n_rounds >= 1 and s.n_past() + 8 + 2 > 64
    with pytest.raises(eb.EagleB200Error, match="KV cache full"):
        s.forward(0, np.ones(64 - s.n_past() + 1, np.int32), s.n_past() + np.arange(64 - s.n_past() + 1))
    n = s.n_past()
    lg2, _ = s.forward(0, [5], [n])                          # one more token still fits, and the answer is the oracle's business elsewhere
    assert np.isfinite(lg2).all()


def test_the_library_still_computes_after_refusals(eb):
    ops = load_ops()
    assert np.array_equal(eb.rms_norm(ops["rms_x"], 1e-5), ops["rms_y"])
