"""CPU: pins oracle/eagle_oracle.c against the committed golden vectors (outputs of the
reference's own ggml-cpu code) and, when oracle/_ref is present, against the reference live."""
import numpy as np
import pytest

import fixtures as fx
import pyoracle as po
from golden_io import load_ops, load_tiny, nmse, rel_max

OPS = load_ops()


@pytest.mark.parametrize("t", fx.QTYPES)
def test_dequantize_bit_exact_vs_golden(t):
    n = po.TYPE_NAMES[t]
    got = po.oracle().dequantize(t, OPS[f"w_{n}"], OPS[f"deq_{n}"].size)
    assert np.array_equal(got.view(np.uint32), OPS[f"deq_{n}"].view(np.uint32))


def test_activation_quantisers_bit_exact_vs_golden():
    o = po.oracle()
    x = OPS["mm_x"][0]
    assert np.array_equal(o.quantize_act(po.Q8_0, x), OPS["act_q8_0"])
    assert np.array_equal(o.quantize_act("q8_K", x), OPS["act_q8_K"])


@pytest.mark.parametrize("t", fx.QTYPES)
def test_mul_mat_vs_golden(t):
    n = po.TYPE_NAMES[t]
    got = po.oracle().mul_mat(t, OPS[f"w_{n}"], 32, 512, OPS["mm_x"])
    assert rel_max(got, OPS[f"mm_{n}"]) < 2e-6      # integer dots are exact; f32 summation order only


def test_glue_ops_vs_golden():
    o = po.oracle()
    assert np.array_equal(o.rms_norm(OPS["rms_x"], 1e-5), OPS["rms_y"])
    assert np.abs(o.rope_norm(OPS["rope_x"], OPS["rope_pos"], 10000.0) - OPS["rope_y"]).max() < 2e-6


def test_flash_attn_vs_golden():
    par = OPS["fa_parent"]; n_past = int(OPS["fa_n_past"])
    vis = fx.vis_matrix(par, n_past, OPS["fa_k"].shape[0])
    got = po.oracle().flash_attn(OPS["fa_q"], OPS["fa_k"], OPS["fa_v"], vis, 1 / np.sqrt(128))
    assert nmse(got, OPS["fa_out"]) < 1e-6          # both keep the f16 V accumulator; expf / sum order differ


def test_fp16_round_trip_all_values():
    o = po.oracle().lib
    import ctypes as C
    o.eo_f16_to_f32.restype = C.c_float; o.eo_f16_to_f32.argtypes = [C.c_uint16]
    o.eo_f32_to_f16.restype = C.c_uint16; o.eo_f32_to_f16.argtypes = [C.c_float]
    h = np.arange(65536, dtype=np.uint16)
    f = h.view(np.float16).astype(np.float32)
    for i in list(range(0, 65536, 97)) + [0, 1, 0x3ff, 0x400, 0x7bff, 0x7c00, 0x8000, 0xfbff]:
        if np.isnan(f[i]):
            continue
        assert o.eo_f16_to_f32(int(h[i])) == f[i]
        assert o.eo_f32_to_f16(float(f[i])) == h[i]
    r = np.random.default_rng(0).normal(0, 1, 2000).astype(np.float32) * np.float32(10.0) ** np.random.default_rng(1).integers(-8, 5, 2000)
    for v in r:
        assert o.eo_f32_to_f16(float(v)) == np.float32(v).astype(np.float16).view(np.uint16)


def test_tiny_model_forward_vs_golden():
    z, tgt, head = load_tiny()
    o = po.oracle()
    n_ctx, E, V = 64, 256, 512
    mt = fx.to_cpu_model(o, tgt); mh = fx.to_cpu_model(o, head, target=mt)
    kvt, kvh = o._g("kv_new")(2, E, n_ctx), o._g("kv_new")(1, E, n_ctx)
    prompt = z["prompt"]; P = len(prompt)
    vis = np.zeros((P, n_ctx), np.uint8); vis[:, :P] = np.tril(np.ones((P, P), np.uint8))
    lg, ft = o.forward(mt, kvt, prompt, np.arange(P), None, 0, n_ctx, vis, E, V)
    assert rel_max(lg, z["tgt_prefill_logits"]) < 1e-2 and nmse(lg, z["tgt_prefill_logits"]) < 1e-5
    hid = np.zeros((P, E), np.float32); hid[1:] = z["tgt_prefill_feat"][:-1]
    lgh, fth = o.forward(mh, kvh, prompt, np.arange(P), hid, 0, n_ctx, vis, E, V)
    assert rel_max(lgh, z["head_prefill_logits"]) < 1e-2
    par = z["tree_parent"]; vis_t = fx.vis_matrix(par, P, n_ctx)
    lg_t, ft_t = o.forward(mt, kvt, z["tree_tokens"], z["tree_pos"], None, P, n_ctx, vis_t, E, V)
    assert rel_max(lg_t, z["tgt_tree_logits"]) < 1e-2 and nmse(ft_t, z["tgt_tree_feat"]) < 1e-5
    lgh_t, fth_t = o.forward(mh, kvh, z["tree_tokens"], z["tree_pos"], z["tree_hidden"], P, n_ctx, vis_t, E, V)
    assert rel_max(lgh_t, z["head_tree_logits"]) < 1e-2
    # greedy tokens identical to the reference's, ties flagged not failed
    for a, b in ((lg_t, z["tgt_tree_logits"]), (lgh_t, z["head_tree_logits"])):
        for row_a, row_b in zip(a, b):
            top = np.sort(row_b)[-2:]
            if top[1] - top[0] > 1e-3 * abs(top[1]):
                assert row_a.argmax() == row_b.argmax()


@pytest.mark.skipif(not po.have_ref(), reason="oracle/_ref not built (needs /root/reference)")
@pytest.mark.parametrize("t", fx.QTYPES)
def test_oracle_vs_live_reference(t):
    o, r = po.oracle(), po.ref()
    rng = np.random.default_rng(5 + t)
    N, K, T = 48, 768, 4
    raw = fx.quantize(t, rng.normal(0, 0.02, (N, K)).astype(np.float32))
    assert np.array_equal(o.dequantize(t, raw, N * K).view(np.uint32), r.dequantize(t, raw, N * K).view(np.uint32))
    X = rng.normal(0, 1, (T, K)).astype(np.float32)
    kind = po.Q8_0 if t in (po.Q4_0, po.Q8_0) else "q8_K"
    for row in X:
        assert np.array_equal(o.quantize_act(kind, row), r.quantize_act(kind, row))
    assert rel_max(o.mul_mat(t, raw, N, K, X), r.mul_mat(t, raw, N, K, X)) < 2e-6


def test_activation_quant_edge_cases():
    o = po.oracle()
    z = np.zeros(256, np.float32)
    q = o.quantize_act("q8_K", z)
    assert not q.any()
    x = z.copy(); x[3] = 2.0; x[100] = -2.0          # equal |max|: the FIRST one sets the sign of d (iscale = -127/max)
    raw = o.quantize_act("q8_K", x)
    q = raw.view(np.int8)
    assert raw[:4].view(np.float32)[0] < 0 and q[4 + 3] == -127 and q[4 + 100] == 127
    x = -x
    raw = o.quantize_act("q8_K", x)
    assert raw[:4].view(np.float32)[0] > 0 and raw.view(np.int8)[4 + 3] == -127
    x = np.full(32, 1e-30, np.float32)
    assert o.quantize_act(po.Q8_0, x).view(np.int8)[2:].tolist() == [127] * 32
