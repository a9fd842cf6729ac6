"""GPU parity tests of the operator layer, through the C-ABI with host buffers, against the CPU
oracle on the same seeded inputs and against the committed golden vectors (reference outputs).
Tolerances follow the reference's own harness (R/tests/test-backend-ops.cpp: NMSE 5e-4 for
MUL_MAT / FLASH_ATTN_EXT, 1e-7 elsewhere) — and are tighter where the arithmetic allows:
bit-exact for dequantisation, activation quantisation, index outputs."""
import numpy as np
import pytest

import fixtures as fx
import pyoracle as po
from golden_io import load_ops, nmse, rel_max

pytestmark = pytest.mark.gpu
OPS = load_ops()


@pytest.mark.parametrize("t", fx.QTYPES)
def test_dequantize_bit_exact(eb, t):
    n = po.TYPE_NAMES[t]
    got = eb.dequantize(t, OPS[f"w_{n}"], OPS[f"deq_{n}"].size)
    assert np.array_equal(got.view(np.uint32), OPS[f"deq_{n}"].view(np.uint32))
    rng = np.random.default_rng(t)
    raw = fx.random_blocks(t, 64 * 1024, rng)          # arbitrary block bytes incl. extreme scales
    assert np.array_equal(eb.dequantize(t, raw, 64 * 1024).view(np.uint32), po.oracle().dequantize(t, raw, 64 * 1024).view(np.uint32))


@pytest.mark.parametrize("t", fx.QTYPES)
def test_repack_round_trip(eb, t):
    rng = np.random.default_rng(40 + t)
    raw = rng.integers(0, 256, po.row_size(t, 8 * 1024), dtype=np.uint8)
    buf = eb.upload_weights(t, raw, 8 * 1024)
    assert np.array_equal(eb.download_weights(t, buf, 8 * 1024), raw)


@pytest.mark.parametrize("kq", [0, 1])
def test_activation_quant_bit_exact(eb, kq):
    rng = np.random.default_rng(7 + kq)
    x = rng.normal(0, 1, (5, 1024)).astype(np.float32)
    x[1, :256] = 0.0                                     # all-zero block
    x[2, 3] = 9.0; x[2, 77] = -9.0                       # tie on |max|
    x[3] *= 1e-20; x[4] *= 1e4
    qs, d, bs = eb.quantize_act(kq, x)
    o = po.oracle()
    for r in range(x.shape[0]):
        raw = o.quantize_act("q8_K" if kq else po.Q8_0, x[r])
        if kq:
            blk = raw.reshape(-1, 292)
            assert np.array_equal(blk[:, 4:260].view(np.int8).reshape(-1), qs[r])
            assert np.array_equal(blk[:, :4].copy().view(np.float32).reshape(-1).view(np.uint32), d[r].view(np.uint32))
            assert np.array_equal(blk[:, 260:].copy().view(np.int16).reshape(-1).astype(np.float32), bs[r])
        else:
            blk = raw.reshape(-1, 34)
            assert np.array_equal(blk[:, 2:].view(np.int8).reshape(-1), qs[r])
            assert np.array_equal(blk[:, :2].copy().view(np.float16).reshape(-1).astype(np.float32), d[r])


@pytest.mark.parametrize("t", fx.QTYPES)
def test_mul_mat_golden(eb, t):
    n = po.TYPE_NAMES[t]
    w = eb.Weights(t, OPS[f"w_{n}"], 32, 512)
    got = eb.mul_mat(w, OPS["mm_x"])
    assert rel_max(got, OPS[f"mm_{n}"]) < 1e-5


@pytest.mark.parametrize("t", fx.QTYPES)
@pytest.mark.parametrize("T", [1, 2, 3, 5, 8, 9, 16, 27, 33, 65, 80, 128, 130])
def test_mul_mat_vs_oracle(eb, t, T):
    rng = np.random.default_rng(100 * t + T)
    N, K = 208, 1536
    raw = fx.quantize(t, rng.normal(0, 0.02, (N, K)).astype(np.float32), rng)
    X = rng.normal(0, 1, (T, K)).astype(np.float32)
    X[0, :256] = 0
    got = eb.mul_mat(eb.Weights(t, raw, N, K), X)
    want = po.oracle().mul_mat(t, raw, N, K, X)
    assert got.shape == want.shape
    assert rel_max(got, want) < 1e-5, (po.TYPE_NAMES[t], T)       # same integer dots, f32 order differs
    assert nmse(got, want) < 1e-10


@pytest.mark.parametrize("t,N,K", [(po.Q4_K, 4096, 4096), (po.Q6_K, 1024, 4096), (po.Q4_K, 512, 11008), (po.Q4_0, 512, 4096), (po.Q8_0, 512, 4096)])
@pytest.mark.parametrize("T", [1, 65])
def test_mul_mat_model_shapes(eb, t, N, K, T):
    rng = np.random.default_rng(N + K + T)
    raw = fx.random_blocks(t, N * K, rng)
    X = rng.normal(0, 1, (T, K)).astype(np.float32)
    got = eb.mul_mat(eb.Weights(t, raw, N, K), X)
    want = po.oracle().mul_mat(t, raw, N, K, X)
    assert rel_max(got, want) < 2e-5


@pytest.mark.parametrize("t,N,K", [(po.Q4_K, 12288, 4096), (po.Q6_K, 4816, 11008), (po.Q4_K, 4096, 8192)])
@pytest.mark.parametrize("T", [4, 8])
def test_mul_mat_draft_shapes(eb, t, N, K, T):
    """T <= 8 (one draft level) on the persistent small-batch kernel: several (32-row block, K split) items per CTA with
    the weight prefetch running across item boundaries, a row count that is not a multiple of 32, K = 11008 (43 super-blocks
    over 8 K groups: uneven), random blocks (every scale / min bit pattern)."""
    rng = np.random.default_rng(N + K + T)
    raw = fx.random_blocks(t, N * K, rng)
    X = rng.normal(0, 1, (T, K)).astype(np.float32)
    X[1, 256:512] = 0
    got = eb.mul_mat(eb.Weights(t, raw, N, K), X)
    want = po.oracle().mul_mat(t, raw, N, K, X)
    assert got.shape == want.shape
    assert rel_max(got, want) < 2e-5


@pytest.mark.parametrize("t", [po.Q4_K, po.Q6_K])
@pytest.mark.parametrize("T", [1, 2, 3, 4, 7, 8])
@pytest.mark.parametrize("K", [2816, 4096, 11008])
def test_mul_mat_kquant_matvec_ragged(eb, t, T, K):
    """the dequant-dot matvec of the K-quants (mmvq.cu) at every column count it is instantiated for: a row count that no GEMM kernel
    takes (not a multiple of 16), K of 11 / 16 / 43 super-blocks (partial 16-super-block steps, a last step of 3 groups), random block
    bytes, a zero activation super-block"""
    rng = np.random.default_rng(11 * t + T + K)
    N = 203
    raw = fx.random_blocks(t, N * K, rng)
    X = rng.normal(0, 1, (T, K)).astype(np.float32)
    X[T - 1, 256:512] = 0
    got = eb.mul_mat(eb.Weights(t, raw, N, K), X)
    want = po.oracle().mul_mat(t, raw, N, K, X)
    assert got.shape == want.shape
    assert rel_max(got, want) < 2e-5


@pytest.mark.parametrize("t", [po.Q4_0, po.Q8_0])
@pytest.mark.parametrize("T", [2, 17, 40, 65, 72, 128])
def test_mul_mat_32block_gemm_ragged(eb, t, T):
    """the int8 tensor-core GEMM for the 32-element block formats (mmq32.cu): K = 11008 (43 chunks of 8 blocks: uneven K
    splits, a last split of one chunk), a row count that is not a multiple of the 64-row CTA tile or the 16-row warp tile,
    every column-tile instantiation (T 2..128), random quant bytes (the -128 corner included), zero activation blocks."""
    rng = np.random.default_rng(7 * t + T)
    N, K = 1003, 11008
    raw = fx.random_blocks(t, N * K, rng)
    X = rng.normal(0, 1, (T, K)).astype(np.float32)
    X[1, 256:512] = 0; X[T - 1, :32] = 0
    got = eb.mul_mat(eb.Weights(t, raw, N, K), X)
    want = po.oracle().mul_mat(t, raw, N, K, X)
    assert got.shape == want.shape
    assert rel_max(got, want) < 2e-5


@pytest.mark.parametrize("t", [po.Q6_K, po.Q8_0, po.Q4_0])
def test_mul_mat_linearity_full_size(eb, t):
    """size-independent property at the BASELINE shape (lm_head 32000 x 4096, T = 65; Q6_K for config 2 on the tcgen05 GEMM,
    Q8_0 / Q4_0 for configs 3 / 1 on the int8 mma.sync GEMM): a column that is an exact copy of another must give identical
    outputs, zero columns zero."""
    rng = np.random.default_rng(9)
    N, K, T = 32000, 4096, 65
    w = eb.Weights(t, fx.random_blocks(t, N * K, rng), N, K)
    X = rng.normal(0, 1, (T, K)).astype(np.float32)
    X[10] = X[3]; X[64] = X[3]; X[20] = 0
    Y = eb.mul_mat(w, X)
    assert np.array_equal(Y[10], Y[3]) and np.array_equal(Y[64], Y[3]) and not Y[20].any()
    y1 = eb.mul_mat(w, X[3:4])                                     # matvec kernel vs GEMM kernel
    assert rel_max(y1[0], Y[3]) < 1e-5


def test_rms_norm_rope_get_rows_elementwise(eb):
    o = po.oracle()
    assert np.array_equal(eb.rms_norm(OPS["rms_x"], 1e-5), OPS["rms_y"])
    rng = np.random.default_rng(2)
    x = rng.normal(0, 3, (9, 4096)).astype(np.float32); g = rng.normal(1, 0.1, 4096).astype(np.float32)
    assert rel_max(eb.rms_norm(x, 1e-5, g), o.rms_norm(x, 1e-5) * g) < 1e-6
    assert np.abs(eb.rope_norm(OPS["rope_x"], OPS["rope_pos"], 10000.0) - OPS["rope_y"]).max() < 5e-6
    q = rng.normal(0, 1, (65, 32, 128)).astype(np.float32); pos = rng.integers(0, 4096, 65).astype(np.int32)
    assert np.abs(eb.rope_norm(q, pos, 10000.0) - o.rope_norm(q, pos, 10000.0)).max() < 2e-5
    for t in fx.QTYPES + (po.F32, po.F16):
        raw = fx.quantize(t, rng.normal(0, 0.02, (300, 512)).astype(np.float32), rng)
        ids = np.array([0, 299, 7, 7, 150], np.int32)
        got = eb.get_rows(eb.Weights(t, raw, 300, 512), ids)
        assert np.array_equal(got.view(np.uint32), o.get_rows(t, raw, 512, 300, ids).view(np.uint32))
    a = rng.normal(0, 1, (7, 512)).astype(np.float32); b = rng.normal(0, 1, 512).astype(np.float32)
    assert np.array_equal(eb.add(a, b), a + b) and np.array_equal(eb.mul(a, b), a * b) and np.array_equal(eb.add(a, a), a + a)
    assert np.array_equal(eb.relu(a), np.maximum(a, 0))
    assert rel_max(eb.silu(a), a / (1 + np.exp(-a.astype(np.float64)))) < 1e-6


def test_flash_attn_golden_both_mask_forms(eb):
    par = OPS["fa_parent"]; n_past = int(OPS["fa_n_past"]); n_kv = OPS["fa_k"].shape[0]
    bits = eb.tree_bits_from_parent(par)
    got = eb.flash_attn_ext(OPS["fa_q"], OPS["fa_k"], OPS["fa_v"], 1 / np.sqrt(128), tree_bits=bits, n_prefix=n_past)
    assert nmse(got, OPS["fa_out"]) < 5e-4 and rel_max(got, OPS["fa_out"]) < 1e-2
    vis = fx.vis_matrix(par, n_past, n_kv)
    T = len(par); Tp = (T + 63) // 64 * 64
    mask = np.full((Tp, n_kv), -np.inf, np.float16); mask[:T][vis.astype(bool)] = 0
    got2 = eb.flash_attn_ext(OPS["fa_q"], OPS["fa_k"], OPS["fa_v"], 1 / np.sqrt(128), mask_f16=mask)
    assert nmse(got2, OPS["fa_out"]) < 5e-4
    # compact form = tcgen05 kernel (P rounded to f16 once, like the reference's CUDA kernel); f16-mask form = mma.sync
    # kernel (P carried as f16 hi + lo): same computation up to the f16 rounding of P
    assert np.abs(got - got2).max() < 5e-4


@pytest.mark.parametrize("T,n_past,H", [(1, 0, 2), (1, 513, 4), (8, 100, 4), (16, 1000, 3), (27, 512, 2), (65, 4031, 2), (128, 300, 1), (65, 61, 32)])
def test_flash_attn_vs_oracle(eb, T, n_past, H):
    rng = np.random.default_rng(T * 7 + n_past)
    D = 128; n_kv = (n_past + T + 255) // 256 * 256       # padded like kv_self.n (R/src/llama-kv-cache.cpp:14-17)
    par = fx.tree_parents(T, rng, 5)
    q = rng.normal(0, 1, (T, H, D)).astype(np.float32)
    k = rng.normal(0, 1, (n_kv, H, D)).astype(np.float16); v = rng.normal(0, 1, (n_kv, H, D)).astype(np.float16)
    got = eb.flash_attn_ext(q, k, v, 1 / np.sqrt(D), tree_bits=eb.tree_bits_from_parent(par), n_prefix=n_past)
    want = po.oracle().flash_attn(q, k.view(np.uint16), v.view(np.uint16), fx.vis_matrix(par, n_past, n_kv), 1 / np.sqrt(D))
    assert nmse(got, want) < 5e-4, nmse(got, want)
    assert np.isfinite(got).all()


@pytest.mark.parametrize("T,n_kv,H", [(1, 96, 2), (27, 256, 2), (65, 4096, 4), (65, 1000, 3), (128, 384, 1), (200, 512, 2), (300, 300, 2)])
def test_flash_attn_bit_mask_form_vs_oracle(eb, T, n_kv, H):
    """eb_mask_to_bits + eb_flash_attn_bits: the tcgen05 kernel under ANY 0 / -inf f16 mask — holes inside the committed prefix (cells of
    rejected branches, R/src/llama-kv-cache.cpp:368-420), an unused cache tail, n_kv that is not a multiple of the 128-cell tile, prompt
    batches beyond one 128-row tile — against the oracle's attention over the same visibility matrix"""
    rng = np.random.default_rng(T * 13 + n_kv)
    D = 128
    q = rng.normal(0, 1, (T, H, D)).astype(np.float32)
    k = rng.normal(0, 1, (n_kv, H, D)).astype(np.float16); v = rng.normal(0, 1, (n_kv, H, D)).astype(np.float16)
    vis = np.zeros((T, n_kv), bool)
    used = max(T, int(0.8 * n_kv))
    prefix = used - min(T, used)
    alive = rng.random(prefix) > 0.15                     # ~15 % of the prefix cells are empty
    vis[:, :prefix] = alive[None, :]
    for t in range(T):                                    # batch cells: causal over a random subset (tree-like)
        cells = prefix + np.arange(min(T, used))
        vis[t, cells[:t + 1]] = rng.random(min(t + 1, len(cells))) > 0.4
        vis[t, cells[min(t, len(cells) - 1)]] = True      # every row sees at least itself
    Tp = (T + 63) // 64 * 64
    mask = np.full((Tp, n_kv), -np.inf, np.float16); mask[:T][vis] = 0
    got, general = eb.flash_attn_ext_bits(q, k, v, 1 / np.sqrt(D), mask)
    assert not general
    want = po.oracle().flash_attn(q, k.view(np.uint16), v.view(np.uint16), vis.astype(np.uint8), 1 / np.sqrt(D))
    assert np.isfinite(got).all()
    assert nmse(got, want) < 5e-4, nmse(got, want)         # the bar of R/tests/test-backend-ops.cpp for FLASH_ATTN_EXT
    if T <= 128:                                          # and the generic f16-mask kernel on the same input
        got2 = eb.flash_attn_ext(q, k, v, 1 / np.sqrt(D), mask_f16=mask)
        assert np.abs(got - got2).max() < 2e-3
    # a mask with other values (ALiBi-like) must be reported, not mis-read
    mask2 = mask.copy(); mask2[0, 0] = np.float16(-0.5)
    assert eb.flash_attn_ext_bits(q, k, v, 1 / np.sqrt(D), mask2) == (None, True)


@pytest.mark.parametrize("n", [1, 8, 60, 1000, 1024])
def test_argsort_vs_numpy(eb, n):
    """GGML_OP_ARGSORT: ascending and descending, ties keep index order (a stable sort of the row, like numpy's with kind='stable')"""
    lib = eb.lib()
    rng = np.random.default_rng(n)
    x = rng.integers(-20, 20, (7, n)).astype(np.float32)              # plenty of ties
    dx, di = eb.DeviceBuffer.from_host(x), eb.DeviceBuffer(x.size * 4)
    for desc in (0, 1):
        assert lib.eb_argsort(dx.ptr, 7, n, desc, di.ptr, None) == 0, lib.eb_last_error().decode()
        assert lib.eb_synchronize(None) == 0
        got = di.to_host(np.int32, (7, n))
        want = np.argsort(-x if desc else x, axis=1, kind="stable")
        assert np.array_equal(got, want)
    assert lib.eb_argsort(dx.ptr, 1, 2000, 0, di.ptr, None) != 0       # beyond one CTA's network: refused, not mis-sorted


def test_top_k_on_two_streams_does_not_share_scratch(eb):
    """eb_top_k keeps one scratch list per stream: calls issued back to back on two streams (they may overlap on the device) give the
    answers of the same calls made alone"""
    import ctypes as C
    lib = eb.lib()
    rng = np.random.default_rng(21)
    xs = [rng.normal(0, 2, (16, 32000)).astype(np.float32) for _ in range(2)]
    want = [eb.top_k(x, 10) for x in xs]
    sts = [C.c_void_p(), C.c_void_p()]
    for s_ in sts:
        assert lib.eb_stream_create(C.byref(s_)) == 0
    bufs = []
    for x in xs:
        bufs.append((eb.DeviceBuffer.from_host(x), eb.DeviceBuffer(16 * 10 * 4), eb.DeviceBuffer(16 * 10 * 4), eb.DeviceBuffer(16 * 10 * 4)))
    for rep in range(20):
        for (dl, di, dv, dp), s_ in zip(bufs, sts):
            assert lib.eb_top_k(dl.ptr, 16, 32000, 10, di.ptr, dv.ptr, dp.ptr, s_) == 0
    for s_ in sts:
        assert lib.eb_synchronize(s_) == 0
    for (dl, di, dv, dp), w in zip(bufs, want):
        assert np.array_equal(di.to_host(np.int32, (16, 10)), w[0]) and np.array_equal(dv.to_host(np.float32, (16, 10)), w[1])
    for s_ in sts:
        assert lib.eb_stream_destroy(s_) == 0


def test_mul_mat_shared_activation_and_rope_out_of_place(eb):
    """eb_mul_mat_ex: three weight matrices on ONE activation tensor (wq / wk / wv in a ggml graph) with the quantisation done once — bit-identical
    to three independent eb_mul_mat calls; eb_rope_norm_to out of place == eb_rope_norm in place, source untouched"""
    lib = eb.lib()
    rng = np.random.default_rng(12)
    K, N, T = 1024, 384, 27
    X = rng.normal(0, 1, (T, K)).astype(np.float32)
    for types in ((po.Q4_K, po.Q4_K, po.Q6_K), (po.Q4_0, po.Q8_0, po.Q4_0)):
        ws = [eb.Weights(t, fx.random_blocks(t, N * K, rng), N, K) for t in types]
        want = [eb.mul_mat(w, X) for w in ws]
        dx = eb.DeviceBuffer.from_host(X); scratch = eb.DeviceBuffer(lib.eb_mul_mat_ws_size(types[0], N, K, T))
        got = []
        for i, w in enumerate(ws):
            dy = eb.DeviceBuffer(T * N * 4)
            assert lib.eb_mul_mat_ex(w.t, w.buf.ptr, N, K, dx.ptr, T, dy.ptr, scratch.ptr, 1 if i else 0, None) == 0, lib.eb_last_error().decode()
            assert lib.eb_synchronize(None) == 0
            got.append(dy.to_host(np.float32, (T, N)))
        for a, b in zip(got, want):
            assert np.array_equal(a, b)
    dy = eb.DeviceBuffer(200 * N * 4); dx2 = eb.DeviceBuffer.from_host(np.zeros((200, K), np.float32))
    assert lib.eb_mul_mat_ex(ws[0].t, ws[0].buf.ptr, N, K, dx2.ptr, 200, dy.ptr, scratch.ptr, 1, None) != 0       # reuse beyond one 128-row tile is refused
    x = rng.normal(0, 1, (9, 4, 128)).astype(np.float32); pos = rng.integers(0, 4000, 9).astype(np.int32)
    want = eb.rope_norm(x, pos, 10000.0)
    ds, dd, dp = eb.DeviceBuffer.from_host(x), eb.DeviceBuffer(x.nbytes), eb.DeviceBuffer.from_host(pos)
    assert lib.eb_rope_norm_to(ds.ptr, dd.ptr, dp.ptr, 128, 4, 9, 10000.0, None) == 0 and lib.eb_synchronize(None) == 0
    assert np.array_equal(dd.to_host(np.float32, x.shape), want) and np.array_equal(ds.to_host(np.float32, x.shape), x)


def test_top_k_argmax(eb):
    rng = np.random.default_rng(4)
    x = rng.normal(0, 2, (9, 32000)).astype(np.float32)
    x[3, [5, 500, 31999]] = 50.0                          # ties: lower ids first
    x[4] = -np.inf; x[4, 123] = 0.0
    ids, vals, probs = eb.top_k(x, 10)
    for r in range(x.shape[0]):
        oi, ov, op = po.top_k(x[r], 10)
        if r != 4:
            assert np.array_equal(ids[r], oi) and np.array_equal(vals[r], ov) and np.abs(probs[r] - op).max() < 1e-6
    assert ids[3, :3].tolist() == [5, 500, 31999] and ids[4, 0] == 123
    assert np.array_equal(eb.argmax(x), np.array([np.flatnonzero(r == r.max())[0] for r in x], np.int32))
    ids1, _, p1 = eb.top_k(np.log(np.array([[0.1, 0.2, 0.3, 0.4]], np.float32)), 1)   # R/tests/test-sampling.cpp:317
    assert ids1[0, 0] == 3 and abs(p1[0, 0] - 1) < 1e-6


@pytest.mark.parametrize("V,k", [(32003, 16), (1030, 7), (4096, 10)])
def test_top_k_split_slices(eb, V, k):
    """the 16-slice form (one CTA per 1/16 row, last arriver merges): ragged last slice, V % 4 != 0, heavy ties that straddle
    slice and warp boundaries, fewer finite values than k, repeated launches (the arrival counters must come back to zero)"""
    rng = np.random.default_rng(V)
    x = rng.integers(-3, 4, (8, V)).astype(np.float32)            # 7 distinct values: every pick is a tie broken by id
    x[1] = rng.normal(0, 1, V)
    x[2] = -np.inf; x[2, [V - 1, 7, V // 2]] = [1.0, 1.0, 2.0]
    x[3, :] = 5.0
    for _ in range(3):
        ids, vals, probs = eb.top_k(x, k)
        for r in range(x.shape[0]):
            oi, ov, op = po.top_k(x[r], k)
            if r != 2:
                assert np.array_equal(ids[r], oi) and np.array_equal(vals[r], ov) and np.abs(probs[r] - op).max() < 1e-6, r
        assert ids[2, :3].tolist() == [V // 2, 7, V - 1] and ids[3].tolist() == list(range(k))


def test_mul_mat_q6_K_extreme_blocks(eb):
    """Q6_K at the corners of its range: scales -128 / 127, quants 0 / 63 (|sc * (q - 32)| up to 4096) against activations that
    quantise to +-127 everywhere.  The tcgen05 path feeds the weight as two exact f16 addends (sc*ql and 16*sc*(qh-2)); the
    per-super-block integer sum reaches 1.3e8 > 2^24 here, where the f32 accumulation of exact products may round (<= 1 ulp
    per add) — the result must still match the CPU's int32 sum to f32 accuracy."""
    rng = np.random.default_rng(5)
    N, K, T = 256, 2048, 33
    nb = N * K // 256
    raw = np.zeros((nb, 210), np.uint8)
    pick = rng.integers(0, 4, nb)
    raw[:, 0:128] = np.where(pick[:, None] & 1, 0xFF, 0x00)                       # ql nibbles all 15 or all 0
    raw[:, 128:192] = np.where(pick[:, None] & 1, 0xFF, 0x00)                     # qh all 3 or all 0 -> q = 63 or 0
    raw[:, 192:208] = np.where(pick[:, None] & 2, 0x80, 0x7F)                     # scales -128 or 127
    raw[:, 208:210] = np.frombuffer(np.float16(2.0 ** -12).tobytes(), np.uint8)           # d = 2^-12
    X = np.where(rng.random((T, K)) < 0.5, -1.0, 1.0).astype(np.float32)
    X[1] = 1.0; X[2] = -1.0                                                        # every product with the same sign
    w = eb.Weights(po.Q6_K, raw.reshape(-1), N, K)
    got = eb.mul_mat(w, X)
    want = po.oracle().mul_mat(po.Q6_K, raw.reshape(-1), N, K, X)
    assert np.abs(want).max() > 1e3                                                # rows 1, 2: |isum| = 1.3e8 per super-block
    assert rel_max(got, want) < 2e-6


def test_fa_off_operator_surface(eb):
    """SOFT_MAX (ggml_soft_max_ext), MUL_MAT with an f16 src0 batched over heads (with broadcast), CPY over strides — the ops of
    llm_build_kqv without flash attention (R/src/llama.cpp:762-812) — against plain numpy restatements of the CPU kernels"""
    rng = np.random.default_rng(21)
    B, T, n = 4, 9, 301
    x = rng.normal(0, 3, (B, T, n)).astype(np.float32)
    mask = np.where(rng.random((12, n)) < 0.3, -np.inf, 0.0).astype(np.float32); mask[:, 0] = 0
    for m in (mask.astype(np.float16), mask, None):
        got = eb.soft_max_ext(x, m, 0.25)
        w = x * np.float32(0.25) + (0 if m is None else m[:T].astype(np.float32)[None])
        e = np.exp(w - w.max(-1, keepdims=True)); want = (e / e.sum(-1, keepdims=True)).astype(np.float32)
        assert np.abs(got - want).max() < 2e-6 and np.abs(got.sum(-1) - 1).max() < 1e-5
    Ba, Bx, N, K = 2, 6, 37, 128
    a = rng.normal(0, 1, (Ba, N, K)).astype(np.float16); xx = rng.normal(0, 1, (Bx, T, K)).astype(np.float32)
    got = eb.mul_mat_dense(a, xx)
    want = np.einsum("btk,bnk->btn", xx.astype(np.float16).astype(np.float32), np.repeat(a.astype(np.float32), Bx // Ba, 0))
    assert np.abs(got - want).max() < 1e-4 * np.abs(want).max()
    # transposed store: src f32 [E][T] read as its transpose [T][E] -> f16 rows of a cache laid out [E][n_ctx] at column kv_head
    E, Tt, n_ctx, kv_head = 16, 5, 12, 4
    v = rng.normal(0, 1, (Tt, E)).astype(np.float32)                      # ggml: ne = [E, T]; transpose -> ne [T, E], nb [E*4, 4]
    raw = eb.copy_strided(v, [Tt, E, 1, 1], [E * 4, 4, Tt * E * 4, Tt * E * 4], np.float16, [Tt, E, 1, 1], [2, n_ctx * 2, 0, 0], E * n_ctx * 2 + kv_head * 2)
    cache = raw[:E * n_ctx * 2].view(np.float16).reshape(E, n_ctx)
    # the call wrote relative to the buffer start: column 0..T-1
    assert np.array_equal(cache[:, :Tt], v.T.astype(np.float16)) and not cache[:, Tt:].any()
