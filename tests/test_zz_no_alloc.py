"""GPU: eb_top_k_noalloc — the allocation-free variant of eb_top_k whose scratch is a file-scope `static __device__` array
(it has no counterpart in the reference).  Called through the C-ABI with the caller's
own device buffers and checked against the existing entry point eb_top_k on the same inputs: ids and values identical, probabilities
within the 1e-6 the other top-k tests use — on several calls in a row (the array keeps the previous call's contents) and with
shrinking row counts.  Collected last on purpose (file name).  Passes alone in a fresh process and
after the other tests."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
MAX_ROWS = 32          # EB_TOP_K_NOALLOC_MAX_ROWS in include/eagle_b200.h


def _both(eb, x, k):
    """eb_top_k and eb_top_k_noalloc on the same device logits -> ((ids, vals, probs), (ids, vals, probs))"""
    lib = eb.lib()
    rows, V = x.shape
    dl = eb.DeviceBuffer.from_host(x)
    outs = []
    for fn in (lib.eb_top_k, lib.eb_top_k_noalloc):
        di, dv, dp = eb.DeviceBuffer(rows * k * 4), eb.DeviceBuffer(rows * k * 4), eb.DeviceBuffer(rows * k * 4)
        for b in (di, dv, dp):
            assert lib.eb_memset(b.ptr, 0xFF, b.nbytes, None) == 0          # stale output bytes must not pass for results
        rc = fn(dl.ptr, rows, V, k, di.ptr, dv.ptr, dp.ptr, None)
        assert rc == 0, lib.eb_last_error().decode()
        assert lib.eb_synchronize(None) == 0
        outs.append((di.to_host(np.int32, (rows, k)), dv.to_host(np.float32, (rows, k)), dp.to_host(np.float32, (rows, k))))
    return outs


def _check(a, b, tag):
    assert np.array_equal(a[0], b[0]), f"{tag}: ids differ"
    assert np.array_equal(a[1], b[1]), f"{tag}: values differ"
    assert np.abs(a[2] - b[2]).max() < 1e-6, f"{tag}: probabilities differ"
    assert (b[0] >= 0).all() and np.isfinite(b[2]).all()


def test_top_k_noalloc_matches_top_k_on_consecutive_calls(eb):
    lib = eb.lib()
    rng = np.random.default_rng(2024)
    # call 1: a full 32-row batch over the Vicuna vocabulary, Gaussian logits
    x1 = rng.normal(0, 2, (MAX_ROWS, 32000)).astype(np.float32)
    ref1, got1 = _both(eb, x1, 10)
    _check(ref1, got1, "call 1")
    assert np.array_equal(got1[0][:, 0], x1.argmax(1))
    # call 2, right after: fewer rows, other V and k, heavy ties (every pick decided by the id) — the array still holds call 1's lists
    x2 = rng.integers(-3, 4, (5, 4099)).astype(np.float32)
    x2[1] = -np.inf; x2[1, [4098, 7, 2000]] = [1.0, 1.0, 2.0]
    ref2, got2 = _both(eb, x2, 16)
    _check(ref2, got2, "call 2")
    assert got2[0][1, :3].tolist() == [2000, 7, 4098]
    # calls 3-5: the same input again and again must give the same answer (arrival counters back at zero each time)
    free0, tot = C.c_size_t(), C.c_size_t()
    x3 = rng.normal(0, 1, (8, 1030)).astype(np.float32)
    ref3, first = _both(eb, x3, 7)
    _check(ref3, first, "call 3")
    dl = eb.DeviceBuffer.from_host(x3)
    di, dv, dp = eb.DeviceBuffer(8 * 7 * 4), eb.DeviceBuffer(8 * 7 * 4), eb.DeviceBuffer(8 * 7 * 4)
    assert lib.eb_device_memory(C.byref(free0), C.byref(tot)) == 0
    for rep in range(3):
        assert lib.eb_top_k_noalloc(dl.ptr, 8, 1030, 7, di.ptr, dv.ptr, dp.ptr, None) == 0, lib.eb_last_error().decode()
        assert lib.eb_synchronize(None) == 0
        _check(ref3, (di.to_host(np.int32, (8, 7)), dv.to_host(np.float32, (8, 7)), dp.to_host(np.float32, (8, 7))), f"repeat {rep}")
    free1 = C.c_size_t()
    assert lib.eb_device_memory(C.byref(free1), C.byref(tot)) == 0
    assert free1.value == free0.value, "device memory in use changed across eb_top_k_noalloc calls"


def test_top_k_noalloc_rejects_what_does_not_fit(eb):
    lib = eb.lib()
    x = np.zeros((MAX_ROWS + 1, 2048), np.float32)
    dl = eb.DeviceBuffer.from_host(x)
    di, dv, dp = eb.DeviceBuffer(x.shape[0] * 16 * 4), eb.DeviceBuffer(x.shape[0] * 16 * 4), eb.DeviceBuffer(x.shape[0] * 16 * 4)
    assert lib.eb_top_k_noalloc(dl.ptr, MAX_ROWS + 1, 2048, 4, di.ptr, dv.ptr, dp.ptr, None) != 0
    assert b"EB_TOP_K_NOALLOC_MAX_ROWS" in lib.eb_last_error()
    assert lib.eb_top_k_noalloc(dl.ptr, 4, 512, 4, di.ptr, dv.ptr, dp.ptr, None) != 0        # below the 16-slice kernel's range
    assert lib.eb_top_k_noalloc(dl.ptr, 4, 40000, 4, di.ptr, dv.ptr, dp.ptr, None) != 0      # above it
    assert lib.eb_top_k_noalloc(dl.ptr, 4, 2048, 17, di.ptr, dv.ptr, dp.ptr, None) != 0      # k > 16
    assert lib.eb_top_k_noalloc(dl.ptr, MAX_ROWS, 2048, 4, di.ptr, dv.ptr, dp.ptr, None) == 0
    assert lib.eb_synchronize(None) == 0
    assert di.to_host(np.int32, (MAX_ROWS + 1, 16)).reshape(-1)[:4].tolist() == [0, 1, 2, 3]  # all-equal row: lowest ids
