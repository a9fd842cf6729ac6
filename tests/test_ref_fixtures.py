"""The known-answer fixtures the reference's OWN unit tests hold for functions on this path, restated and applied to both
the oracle (CPU tests) and the CUDA path (-m gpu tests, through the C-ABI):

* R/tests/test-quantize-fns.cpp — data 0.1 + 2 cos(i + offset), 32 * 128 values: quantise -> dequantise RMSE / n below
  MAX_QUANTIZATION_TOTAL_ERROR (0.002), |vec_dot - float dot| / n below MAX_DOT_PRODUCT_ERROR (0.02), for the four weight formats of
  the path and their activation formats (Q8_0 for Q4_0 / Q8_0, Q8_K for Q4_K / Q6_K);
* R/tests/test-rope.cpp mode 0 — rotating by p0 and then by a constant shift p1 equals rotating by p0 + p1 ("context swap"):
  sum |r1 - r2| / sum |r| < 1e-4 on a [73 tokens][32 heads] tensor with positions 100.. shifted by -67;
* R/tests/test-sampling.cpp:317-320 (top-k known answers) lives in tests/test_spec_logic.py / tests/test_gpu_ops.py.

The weight quantisers themselves (quantize_row_q4_K ...) are not on the path — weights arrive quantised — so the blocks come from the
reference's own ggml_quantize_chunk (oracle/_ref), exactly what llama-quantize writes."""
import numpy as np
import pytest

import fixtures as fx
import pyoracle as po

N = 32 * 128
MAX_QUANTIZATION_TOTAL_ERROR = 0.002
MAX_DOT_PRODUCT_ERROR = 0.02
TYPES = [po.Q4_0, po.Q8_0, po.Q4_K, po.Q6_K]
NAMES = {po.Q4_0: "q4_0", po.Q8_0: "q8_0", po.Q4_K: "q4_K", po.Q6_K: "q6_K"}


def generate_data(offset, n=N):
    """R/tests/test-quantize-fns.cpp:29-33"""
    i = np.arange(n, dtype=np.float32) + np.float32(offset)
    return (0.1 + 2.0 * np.cos(i).astype(np.float64)).astype(np.float32)


def array_rmse(a, b):
    """R/tests/test-quantize-fns.cpp:35-42 (note: sqrt(sum) / n, not the textbook RMSE)"""
    d = a.astype(np.float64) - b.astype(np.float64)
    return float(np.sqrt((d * d).sum()) / a.size)


def _need_ref():
    if not po.have_ref():
        pytest.fail("oracle/_ref is missing: run __graft_entry__.build() where /root/reference is present")


def act_kind(t):
    return po.Q8_0 if t in (po.Q4_0, po.Q8_0) else "q8_K"


@pytest.mark.parametrize("t", TYPES, ids=lambda t: NAMES[t])
def test_oracle_dequantize_meets_the_reference_tests_total_error_bound(t):
    _need_ref()
    data = generate_data(0.0)
    raw = fx.quantize(t, data.reshape(1, N))
    err = array_rmse(data, po.oracle().dequantize(t, raw, N))
    print(f"[quantize-fns] {NAMES[t]} absolute quantization error (oracle dequantiser): {err:.6f}")
    assert err < MAX_QUANTIZATION_TOTAL_ERROR
    assert np.array_equal(po.oracle().dequantize(t, raw, N), po.ref().dequantize(t, raw, N))


@pytest.mark.parametrize("t", TYPES, ids=lambda t: NAMES[t])
def test_oracle_vec_dot_meets_the_reference_tests_dot_product_bound(t):
    _need_ref()
    d1, d2 = generate_data(0.0), generate_data(1.0)
    w = fx.quantize(t, d1.reshape(1, N))
    want = float((d1.astype(np.float64) * d2.astype(np.float64)).sum())
    for path in (po.oracle(), po.ref()):
        got = path.vec_dot(t, N, w, path.quantize_act(act_kind(t), d2))
        err = abs(got - want) / N
        print(f"[quantize-fns] {NAMES[t]} dot product error ({path.kind}): {err:.6f}")
        assert err < MAX_DOT_PRODUCT_ERROR
    a, b = (path.vec_dot(t, N, w, path.quantize_act(act_kind(t), d2)) for path in (po.oracle(), po.ref()))
    assert abs(a - b) <= 2e-6 * abs(b)              # integer block dots are exact; only the f32 summation order over blocks differs (AVX2 lanes)


def _rope_case():
    """R/tests/test-rope.cpp:141-168, mode 0, with the whole head rotated (n_rot = head size, the only form on this path)"""
    rng = np.random.default_rng(7)
    T, H, D = 73, 32, 128
    x = rng.uniform(-1.0, 1.0, (T, H, D)).astype(np.float32)
    p0 = 100 + np.arange(T, dtype=np.int32)
    p1 = np.full(T, 33 - 100, np.int32)
    p2 = 33 + np.arange(T, dtype=np.int32)
    return x, p0, p1, p2


def _rope_check(rope, who):
    x, p0, p1, p2 = _rope_case()
    r1 = rope(rope(x, p0), p1)
    r2 = rope(x, p2)
    diff, s0, s1 = np.abs(r1.astype(np.float64) - r2).sum(), np.abs(r1.astype(np.float64)).sum(), np.abs(r2.astype(np.float64)).sum()
    print(f"[test-rope] {who}: rel err {diff / s0:.3e}")
    assert diff / s0 < 1e-4 and diff / s1 < 1e-4
    return r2


def test_oracle_rope_context_swap():
    _rope_check(lambda x, p: po.oracle().rope_norm(x, p, 10000.0), "oracle")


@pytest.mark.gpu
@pytest.mark.parametrize("t", TYPES, ids=lambda t: NAMES[t])
def test_gpu_dequantize_meets_the_reference_tests_total_error_bound(eb, t):
    _need_ref()
    data = generate_data(0.0)
    raw = fx.quantize(t, data.reshape(1, N))
    got = eb.dequantize(t, raw, N)
    err = array_rmse(data, got)
    print(f"[quantize-fns] {NAMES[t]} absolute quantization error (CUDA dequantiser): {err:.6f}")
    assert err < MAX_QUANTIZATION_TOTAL_ERROR
    assert np.array_equal(got, po.oracle().dequantize(t, raw, N))


@pytest.mark.gpu
@pytest.mark.parametrize("T", [1, 4, 16])
@pytest.mark.parametrize("t", TYPES, ids=lambda t: NAMES[t])
def test_gpu_mul_mat_meets_the_reference_tests_dot_product_bound(eb, t, T):
    """the same bound on every element of a [256 rows] x [T columns] product (T = 1: matvec, 4: the draft-head kernel, 16: the tcgen05 /
    int8 mma GEMMs), rows and columns being the fixture's data at different offsets"""
    _need_ref()
    rows = 256
    W = np.stack([generate_data(float(r)) for r in range(rows)])
    X = np.stack([generate_data(1000.0 + 3 * c) for c in range(T)])
    raw = fx.quantize(t, W)
    got = eb.mul_mat(eb.Weights(t, raw, rows, N), X)
    want = X.astype(np.float64) @ W.astype(np.float64).T
    err = np.abs(got - want).max() / N
    print(f"[quantize-fns] {NAMES[t]} T={T}: worst dot product error over {rows * T} dots (CUDA mul_mat): {err:.6f}")
    assert err < MAX_DOT_PRODUCT_ERROR
    ref = po.oracle().mul_mat(t, raw, rows, N, X)
    assert np.abs(got - ref).max() <= 1e-4 * np.abs(ref).max()        # same integer dots as the oracle; f32 summation order only


@pytest.mark.gpu
def test_gpu_rope_context_swap(eb):
    r2 = _rope_check(lambda x, p: eb.rope_norm(x, p, 10000.0), "CUDA")
    x, _, _, p2 = _rope_case()
    want = po.oracle().rope_norm(x, p2, 10000.0)
    assert np.abs(r2 - want).max() < 2e-5
