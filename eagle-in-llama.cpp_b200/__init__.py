"""eagle_b200 — host-side mirror of the reference's operator / decode interface for the EAGLE
speculative-step hot path, over the C-ABI of libeagle_b200.so (include/eagle_b200.h).

Names follow the reference: ggml ops (mul_mat, flash_attn_ext, rms_norm, rope, get_rows, top_k …,
R/ggml/include/ggml.h) and the llama.cpp decode surface (a model with GGUF tensors, a session
with KV caches, forward over a tree-shaped batch, one speculative round).  Everything runs on
the GPU through the shared library; there is NO CPU fallback — importing works anywhere, but
any compute call raises if the library or an sm_100 device is missing.

The directory name contains '-' and '.', so import it through `load_package()` in
__graft_entry__.py / tests/conftest.py, which registers it as module `eagle_b200`.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libeagle_b200.so")

F32, F16, Q4_0, Q8_0, Q4_K, Q6_K = 0, 1, 2, 8, 12, 14
BLOCK = {F32: (1, 4), F16: (1, 2), Q4_0: (32, 18), Q8_0: (32, 34), Q4_K: (256, 144), Q6_K: (256, 210)}
T_TOK_EMBD, T_OUTPUT_NORM, T_OUTPUT, T_FC, T_FC_BIAS = range(5)
L_ATTN_NORM, L_WQ, L_WK, L_WV, L_WO, L_FFN_NORM, L_GATE, L_UP, L_DOWN = range(9)


class EagleB200Error(RuntimeError):
    pass


class SessionParams(C.Structure):
    _fields_ = [("n_ctx", C.c_int), ("max_nodes", C.c_int), ("n_branch", C.c_int), ("n_draft", C.c_int),
                ("max_depth", C.c_int), ("top_k", C.c_int), ("p_split", C.c_float), ("fixed_accept", C.c_int),
                ("use_graph", C.c_int), ("p_min", C.c_float)]


class RoundResult(C.Structure):
    _fields_ = [("n_nodes", C.c_int), ("n_accept", C.c_int), ("bonus", C.c_int32), ("tokens", C.c_int32 * 128),
                ("parent", C.c_int32 * 128), ("path", C.c_int32 * 128), ("gpu_ms", C.c_float)]


_lib = None


def lib():
    """Load libeagle_b200.so (built by `make` / __graft_entry__.build()).  Fails loudly."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise EagleB200Error(f"{LIB_PATH} is missing: run `make` (or __graft_entry__.build()); there is no CPU fallback")
    L = C.CDLL(LIB_PATH)
    vp, i32, i64, f32, sz = C.c_void_p, C.c_int, C.c_int64, C.c_float, C.c_size_t
    sig = {
        "eb_last_error": ([], C.c_char_p), "eb_version": ([], i32), "eb_device_init": ([i32], i32),
        "eb_device_sm_count": ([], i32), "eb_device_memory": ([C.POINTER(sz), C.POINTER(sz)], i32),
        "eb_copy": ([vp, vp, sz, vp], i32), "eb_copy_strided": ([i32, i32, vp, vp, vp, vp, vp, vp, vp], i32),
        "eb_mul_mat_dense": ([i32, vp, i64, i64, i64, vp, i64, i64, vp, i64, i64, i64, i64, vp], i32),
        "eb_soft_max_ext": ([vp, vp, i32, i64, vp, i64, i64, i64, f32, vp], i32), "eb_concat2": ([vp, vp, vp, i32, i32, vp], i32), "eb_row_size": ([i32, i64], sz),
        "eb_malloc": ([C.POINTER(vp), sz], i32), "eb_free": ([vp], i32), "eb_memset": ([vp, i32, sz, vp], i32),
        "eb_upload": ([vp, vp, sz, vp], i32), "eb_download": ([vp, vp, sz, vp], i32), "eb_synchronize": ([vp], i32),
        "eb_repack": ([i32, vp, vp, i64, vp], i32), "eb_unpack": ([i32, vp, vp, i64, vp], i32),
        "eb_quantize_act_planes": ([i32, vp, i64, i32, vp, vp, vp, vp], i32),
        "eb_dequantize": ([i32, vp, vp, i64, vp], i32),
        "eb_mul_mat_ws_size": ([i32, i64, i64, i64], sz),
        "eb_mul_mat": ([i32, vp, i64, i64, vp, i64, vp, vp, vp], i32),
        "eb_mul_mat_ex": ([i32, vp, i64, i64, vp, i64, vp, vp, i32, vp], i32),
        "eb_rope_norm_to": ([vp, vp, vp, i32, i32, i32, f32, vp], i32),
        "eb_rms_norm": ([vp, vp, vp, i64, i64, f32, vp], i32),
        "eb_rope_norm": ([vp, vp, i32, i32, i32, f32, vp], i32),
        "eb_get_rows_n": ([i32, vp, i64, i64, vp, i32, vp, vp], i32),
        "eb_add": ([vp, vp, vp, i64, i64, vp], i32), "eb_mul": ([vp, vp, vp, i64, i64, vp], i32),
        "eb_silu": ([vp, vp, i64, vp], i32), "eb_scale": ([vp, vp, i64, f32, vp], i32), "eb_relu": ([vp, vp, i64, vp], i32),
        "eb_cpy_f32_f16": ([vp, vp, i64, vp], i32),
        "eb_flash_attn_ws_size": ([i32, i32, i32, i32], sz),
        "eb_flash_attn": ([vp, vp, vp, i64, vp, i64, vp, i32, i32, i32, i32, i32, f32, vp, vp, vp], i32),
        "eb_mask_bits_size": ([i32, i32], sz), "eb_mask_to_bits": ([vp, i64, i32, i32, vp, C.POINTER(i32), vp], i32),
        "eb_flash_attn_bits_supported": ([i32, i32, i32, i64], i32),
        "eb_flash_attn_bits": ([vp, vp, vp, i64, vp, i32, i32, i32, i32, f32, vp, vp, vp], i32),
        "eb_malloc_host": ([C.POINTER(vp), sz], i32), "eb_free_host": ([vp], i32),
        "eb_stream_create": ([C.POINTER(vp)], i32), "eb_stream_destroy": ([vp], i32), "eb_device_synchronize": ([], i32),
        "eb_upload_async": ([vp, vp, sz, vp], i32), "eb_download_async": ([vp, vp, sz, vp], i32),
        "eb_event_create": ([C.POINTER(vp)], i32), "eb_event_destroy": ([vp], i32), "eb_event_record": ([vp, vp], i32),
        "eb_stream_wait_event": ([vp, vp], i32), "eb_event_synchronize": ([vp], i32),
        "eb_top_k": ([vp, i32, i32, i32, vp, vp, vp, vp], i32), "eb_argmax": ([vp, i32, i32, vp, vp], i32), "eb_argsort": ([vp, i32, i32, i32, vp, vp], i32),
        "eb_top_k_noalloc": ([vp, i32, i32, i32, vp, vp, vp, vp], i32),
        "eb_model_create": ([i32, i32, i32, i32, i32, f32, f32, i32], vp), "eb_model_free": ([vp], None),
        "eb_model_set_tensor": ([vp, i32, i32, i32, vp], i32),
        "eb_model_synth_tensor": ([vp, i32, i32, i32, C.c_uint64, f32], i32),
        "eb_model_get_tensor": ([vp, i32, i32, vp, sz], i32),
        "eb_model_share_lm_head": ([vp, vp], i32), "eb_model_weight_bytes": ([vp], sz),
        "eb_model_get_hparams": ([vp, C.POINTER(i32), C.POINTER(C.c_float)], i32),
        "eb_model_load_gguf": ([C.c_char_p, C.POINTER(vp), C.POINTER(C.c_double), C.POINTER(C.c_uint64)], i32),
        "eb_gguf_probe": ([C.c_char_p, C.POINTER(i32), C.POINTER(C.c_float), C.POINTER(i32), C.POINTER(C.c_uint64)], i32),
        "eb_session_create": ([vp, vp, C.POINTER(SessionParams)], vp), "eb_session_free": ([vp], None),
        "eb_session_forward": ([vp, i32, vp, vp, vp, vp, i32, vp, vp], i32),
        "eb_tree_bits_from_batch": ([i32, vp, vp, vp, vp], i32),
        "eb_session_forward_batch": ([vp, i32, vp, vp, vp, vp, vp, i32, vp, vp], i32),
        "eb_model_create_tp": ([i32, i32, i32, i32, i32, i32, f32, f32], vp),
        "eb_tp_create": ([i32, i32, i64], vp), "eb_tp_handle": ([vp, vp], i32), "eb_tp_open_peer": ([vp, i32, vp], i32),
        "eb_tp_free": ([vp], None), "eb_tp_error": ([vp], i32), "eb_tp_allreduce": ([vp, vp, i32, i64, vp, vp], i32),
        "eb_tp_bench": ([vp, vp, i32, i64, vp, i32, i32, C.POINTER(C.c_float)], i32),
        "eb_session_set_tp": ([vp, vp], i32),
        "eb_session_set_tap": ([vp, vp, sz], i32), "eb_session_tap_state": ([vp, C.POINTER(sz), C.POINTER(i32)], i32),
        "eb_session_prefill": ([vp, vp, i32, vp, vp], i32),
        "eb_session_kv_write": ([vp, i32, i32, i32, i32, i32, vp], i32), "eb_session_kv_read": ([vp, i32, i32, i32, i32, i32, vp], i32),
        "eb_session_set_n_past": ([vp, i32], i32),
        "eb_session_commit": ([vp, i32, i32], i32), "eb_session_n_past": ([vp, i32], i32),
        "eb_session_synth_kv": ([vp, i32, C.c_uint64], i32), "eb_session_set_root": ([vp, C.c_int32, vp], i32),
        "eb_session_round": ([vp, C.POINTER(RoundResult)], i32),
        "eb_session_verify_only": ([vp, i32, C.POINTER(f32)], i32),
        "eb_session_gen_draft": ([vp, vp, vp, i32, C.POINTER(i32)], i32),
        "eb_bench_mul_mat_kernel": ([i32, i64, i64, i32, i32, i32, C.POINTER(f32)], i32),
        "eb_bench_flash_attn": ([i32, i32, i32, i32, i32, i32, C.POINTER(f32)], i32),
        "eb_launch_count": ([i32], C.c_long),
    }
    for name, (args, res) in sig.items():
        fn = getattr(L, name)
        fn.argtypes, fn.restype = args, res
    _lib = L
    return L


EXPORTS = None  # filled by tests: every symbol declared in include/eagle_b200.h


def _ck(rc):
    if rc != 0:
        raise EagleB200Error(f"eagle_b200 error {rc}: {lib().eb_last_error().decode()}")


def init(device=0):
    _ck(lib().eb_device_init(device))


def row_size(t, k):
    be, bb = BLOCK[t]
    assert k % be == 0, (t, k)
    return k // be * bb


def _hp(a):
    if a is None:
        return None
    assert isinstance(a, np.ndarray) and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(C.c_void_p)


class DeviceBuffer:
    """A device allocation (ggml_backend_buffer: alloc / set_tensor / get_tensor)."""

    def __init__(self, nbytes):
        self.ptr = C.c_void_p()
        self.nbytes = int(nbytes)
        _ck(lib().eb_malloc(C.byref(self.ptr), self.nbytes))

    @classmethod
    def from_host(cls, arr):
        arr = np.ascontiguousarray(arr)
        b = cls(max(arr.nbytes, 1))
        if arr.nbytes:
            _ck(lib().eb_upload(b.ptr, _hp(arr), arr.nbytes, None))
        return b

    def to_host(self, dtype, shape):
        out = np.empty(shape, dtype)
        assert out.nbytes <= self.nbytes
        if out.nbytes:
            _ck(lib().eb_download(_hp(out), self.ptr, out.nbytes, None))
        return out

    def free(self):
        if self.ptr:
            lib().eb_free(self.ptr)
            self.ptr = C.c_void_p()

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


def upload_weights(t, raw, n_elems):
    """GGUF block bytes (host) -> device layout buffer (set_tensor)."""
    src = DeviceBuffer.from_host(np.ascontiguousarray(raw).view(np.uint8))
    if t in (Q4_0, Q8_0, Q6_K):
        dst = DeviceBuffer(src.nbytes)
        _ck(lib().eb_repack(t, src.ptr, dst.ptr, n_elems, None))
        _ck(lib().eb_synchronize(None))
        src.free()
        return dst
    return src


def download_weights(t, buf, n_elems):
    """device layout buffer -> GGUF block bytes (get_tensor)."""
    nbytes = row_size(t, n_elems)
    if t in (Q4_0, Q8_0, Q6_K):
        tmp = DeviceBuffer(nbytes)
        _ck(lib().eb_unpack(t, buf.ptr, tmp.ptr, n_elems, None))
        return tmp.to_host(np.uint8, nbytes)
    return buf.to_host(np.uint8, nbytes)


# ---- ggml ops over host buffers (each call: H2D, kernel(s), D2H) -------------------------------
def dequantize(t, raw, n_elems):
    w = upload_weights(t, raw, n_elems)
    out = DeviceBuffer(n_elems * 4)
    _ck(lib().eb_dequantize(t, w.ptr, out.ptr, n_elems, None))
    return out.to_host(np.float32, n_elems)


def quantize_act(kq, x):
    """-> (qs int8 [T][K], d f32 [T][K/B], bsum f32 [T][K/16] or None)"""
    x = np.ascontiguousarray(x, np.float32)
    T, K = x.shape
    B = 256 if kq else 32
    dx = DeviceBuffer.from_host(x)
    qs, d = DeviceBuffer(T * K), DeviceBuffer(T * (K // B) * 4)
    bs = DeviceBuffer(T * (K // 16) * 2) if kq else None
    _ck(lib().eb_quantize_act_planes(int(kq), dx.ptr, K, T, qs.ptr, d.ptr, bs.ptr if bs else None, None))
    return (qs.to_host(np.int8, (T, K)), d.to_host(np.float32, (T, K // B)),
            bs.to_host(np.float16, (T, K // 16)).astype(np.float32) if bs else None)


class Weights:
    """A quantised matrix resident on the device (rows = output features)."""

    def __init__(self, t, raw, N, K):
        self.t, self.N, self.K = t, N, K
        self.buf = upload_weights(t, raw, N * K)


def mul_mat(w, x):
    """GGML_OP_MUL_MAT: w Weights [N][K], x f32 [T][K] -> f32 [T][N]"""
    x = np.ascontiguousarray(x, np.float32).reshape(-1, w.K)
    T = x.shape[0]
    dx, dy = DeviceBuffer.from_host(x), DeviceBuffer(T * w.N * 4)
    ws = DeviceBuffer(lib().eb_mul_mat_ws_size(w.t, w.N, w.K, min(T, 128)))
    _ck(lib().eb_mul_mat(w.t, w.buf.ptr, w.N, w.K, dx.ptr, T, dy.ptr, ws.ptr, None))
    return dy.to_host(np.float32, (T, w.N))


def rms_norm(x, eps, weight=None):
    x = np.ascontiguousarray(x, np.float32)
    n = x.shape[-1]
    dx, dy = DeviceBuffer.from_host(x), DeviceBuffer(x.nbytes)
    dw = DeviceBuffer.from_host(np.ascontiguousarray(weight, np.float32)) if weight is not None else None
    _ck(lib().eb_rms_norm(dx.ptr, dw.ptr if dw else None, dy.ptr, n, x.size // n, eps, None))
    return dy.to_host(np.float32, x.shape)


def rope_norm(x, pos, freq_base):
    x = np.ascontiguousarray(x, np.float32)
    T, H, D = x.shape
    dx, dp = DeviceBuffer.from_host(x), DeviceBuffer.from_host(np.ascontiguousarray(pos, np.int32))
    _ck(lib().eb_rope_norm(dx.ptr, dp.ptr, D, H, T, freq_base, None))
    return dx.to_host(np.float32, x.shape)


def get_rows(w, ids):
    ids = np.ascontiguousarray(ids, np.int32)
    di, out = DeviceBuffer.from_host(ids), DeviceBuffer(ids.size * w.K * 4)
    _ck(lib().eb_get_rows_n(w.t, w.buf.ptr, w.N, w.K, di.ptr, ids.size, out.ptr, None))
    return out.to_host(np.float32, (ids.size, w.K))


def _binary(fn, a, b):
    a = np.ascontiguousarray(a, np.float32); b = np.ascontiguousarray(b, np.float32)
    da, db, dy = DeviceBuffer.from_host(a), DeviceBuffer.from_host(b), DeviceBuffer(a.nbytes)
    _ck(fn(da.ptr, db.ptr, dy.ptr, a.size, b.size, None))
    return dy.to_host(np.float32, a.shape)


def add(a, b):
    return _binary(lib().eb_add, a, b)


def mul(a, b):
    return _binary(lib().eb_mul, a, b)


def _unary(fn, x):
    x = np.ascontiguousarray(x, np.float32)
    dx, dy = DeviceBuffer.from_host(x), DeviceBuffer(x.nbytes)
    _ck(fn(dx.ptr, dy.ptr, x.size, None))
    return dy.to_host(np.float32, x.shape)


def silu(x):
    return _unary(lib().eb_silu, x)


def relu(x):
    return _unary(lib().eb_relu, x)


def soft_max_ext(x, mask, scale):
    """GGML_OP_SOFT_MAX: x f32 [B][T][n]; mask f16 / f32 [>= T][n] or None (shared by the B batches)"""
    x = np.ascontiguousarray(x, np.float32)
    B, T, n = x.shape
    dx, dy = DeviceBuffer.from_host(x), DeviceBuffer(x.nbytes)
    dm, mt = None, 0
    if mask is not None:
        mask = np.ascontiguousarray(mask); mt = F16 if mask.dtype == np.float16 else F32
        dm = DeviceBuffer.from_host(mask)
    _ck(lib().eb_soft_max_ext(dx.ptr, dm.ptr if dm else None, mt, n, dy.ptr, n, B * T, T, scale, None))
    return dy.to_host(np.float32, x.shape)


def mul_mat_dense(a, x):
    """GGML_OP_MUL_MAT with f16 / f32 src0: a [Ba][N][K], x f32 [B][T][K] (B % Ba == 0) -> [B][T][N]"""
    a = np.ascontiguousarray(a); x = np.ascontiguousarray(x, np.float32)
    Ba, N, K = a.shape; B, T, _ = x.shape
    da, dx, dy = DeviceBuffer.from_host(a), DeviceBuffer.from_host(x), DeviceBuffer(B * T * N * 4)
    _ck(lib().eb_mul_mat_dense(F16 if a.dtype == np.float16 else F32, da.ptr, K * a.itemsize, N * K * a.itemsize, Ba, dx.ptr, K * 4, T * K * 4,
                               dy.ptr, N, K, T, B, None))
    return dy.to_host(np.float32, (B, T, N))


def copy_strided(src, ne, nb_src, dst_dtype, dne, nb_dst, dst_bytes):
    """GGML_OP_CPY over byte strides: src = flat numpy array (f32 / f16); returns the destination buffer as bytes"""
    src = np.ascontiguousarray(src)
    ds, dd = DeviceBuffer.from_host(src), DeviceBuffer(dst_bytes)
    _ck(lib().eb_memset(dd.ptr, 0, dst_bytes, None))
    arr = lambda v: np.ascontiguousarray(v, np.int64)
    a1, a2, a3, a4 = arr(ne), arr(nb_src), arr(dne), arr(nb_dst)
    _ck(lib().eb_copy_strided(F16 if src.dtype == np.float16 else F32, F16 if dst_dtype == np.float16 else F32, ds.ptr, dd.ptr, _hp(a1), _hp(a2), _hp(a3), _hp(a4), None))
    return dd.to_host(np.uint8, dst_bytes)


def tree_bits_from_parent(parent):
    """ancestor-or-self sets, 128 bits per row (the compact tree mask)"""
    T = len(parent)
    bits = np.zeros((T, 4), np.uint32)
    for t in range(T):
        if parent[t] >= 0:
            bits[t] = bits[parent[t]]
        bits[t, t >> 5] |= np.uint32(1 << (t & 31))
    return bits


def flash_attn_ext(q, k, v, scale, mask_f16=None, tree_bits=None, n_prefix=0):
    """GGML_OP_FLASH_ATTN_EXT.  q f32 [T][H][D]; k, v f16 [n_kv][H][D]; mask_f16 [T_pad][n_kv]
    (0 / -inf) or the compact tree form (tree_bits [T][4] u32 + n_prefix)."""
    q = np.ascontiguousarray(q, np.float32)
    T, H, D = q.shape
    k = np.ascontiguousarray(k).view(np.uint16); v = np.ascontiguousarray(v).view(np.uint16)
    n_kv = k.shape[0]
    dq, dk, dv, do = DeviceBuffer.from_host(q), DeviceBuffer.from_host(k), DeviceBuffer.from_host(v), DeviceBuffer(q.nbytes)
    ws = DeviceBuffer(lib().eb_flash_attn_ws_size(T, H, D, n_kv))
    dm = db = None
    stride = 0
    if mask_f16 is not None:
        m = np.ascontiguousarray(mask_f16, np.float16)
        stride = m.shape[1]
        dm = DeviceBuffer.from_host(m.view(np.uint16))
    else:
        db = DeviceBuffer.from_host(np.ascontiguousarray(tree_bits, np.uint32))
    _ck(lib().eb_flash_attn(dq.ptr, dk.ptr, dv.ptr, H * D, dm.ptr if dm else None, stride, db.ptr if db else None,
                            n_prefix, T, H, D, n_kv, scale, do.ptr, ws.ptr, None))
    return do.to_host(np.float32, q.shape)


def flash_attn_ext_bits(q, k, v, scale, mask_f16, stream=None):
    """GGML_OP_FLASH_ATTN_EXT through eb_mask_to_bits + eb_flash_attn_bits (tcgen05 kernel, any 0 / -inf mask, any T).
    -> (out, general): general = True means the mask held other values and `out` is None (use flash_attn_ext)."""
    q = np.ascontiguousarray(q, np.float32)
    T, H, D = q.shape
    k = np.ascontiguousarray(k).view(np.uint16); v = np.ascontiguousarray(v).view(np.uint16)
    n_kv = k.shape[0]
    m = np.ascontiguousarray(mask_f16, np.float16)
    dq, dk, dv, do = DeviceBuffer.from_host(q), DeviceBuffer.from_host(k), DeviceBuffer.from_host(v), DeviceBuffer(q.nbytes)
    dm = DeviceBuffer.from_host(m.view(np.uint16))
    ws = DeviceBuffer(lib().eb_flash_attn_ws_size(T, H, D, n_kv)); bits = DeviceBuffer(lib().eb_mask_bits_size(T, n_kv))
    general = C.c_int(-1)
    _ck(lib().eb_mask_to_bits(dm.ptr, m.shape[1], T, n_kv, bits.ptr, C.byref(general), stream))
    if general.value:
        return None, True
    _ck(lib().eb_flash_attn_bits(dq.ptr, dk.ptr, dv.ptr, H * D, bits.ptr, T, H, D, n_kv, scale, do.ptr, ws.ptr, stream))
    _ck(lib().eb_synchronize(stream))
    return do.to_host(np.float32, q.shape), False


def top_k(logits, k):
    logits = np.ascontiguousarray(logits, np.float32)
    rows, V = logits.shape
    dl = DeviceBuffer.from_host(logits)
    di, dv, dp = DeviceBuffer(rows * k * 4), DeviceBuffer(rows * k * 4), DeviceBuffer(rows * k * 4)
    _ck(lib().eb_top_k(dl.ptr, rows, V, k, di.ptr, dv.ptr, dp.ptr, None))
    return di.to_host(np.int32, (rows, k)), dv.to_host(np.float32, (rows, k)), dp.to_host(np.float32, (rows, k))


def argmax(logits):
    logits = np.ascontiguousarray(logits, np.float32)
    rows, V = logits.shape
    dl, di = DeviceBuffer.from_host(logits), DeviceBuffer(rows * 4)
    _ck(lib().eb_argmax(dl.ptr, rows, V, di.ptr, None))
    return di.to_host(np.int32, rows)


# ---- engine ---------------------------------------------------------------------------------
class Model:
    """Weights of one GGUF model in HBM (llama target or eagle head)."""

    def __init__(self, E, H, F, V, L, eps=1e-5, freq_base=10000.0, is_eagle=False):
        self.E, self.H, self.F, self.V, self.L, self.is_eagle = E, H, F, V, L, bool(is_eagle)
        self.eps, self.freq_base = eps, freq_base
        self.h = lib().eb_model_create(E, H, F, V, L, eps, freq_base, int(is_eagle))
        if not self.h:
            raise EagleB200Error(lib().eb_last_error().decode())

    @classmethod
    def from_gguf(cls, path):
        """eb_model_load_gguf: a GGUF file (arch llama or eagle) streamed into HBM -> Model; .load_seconds / .load_bytes report the upload"""
        h, sec, nb = C.c_void_p(), C.c_double(), C.c_uint64()
        _ck(lib().eb_model_load_gguf(os.fsencode(path), C.byref(h), C.byref(sec), C.byref(nb)))
        hp, hf = (C.c_int32 * 6)(), (C.c_float * 2)()
        _ck(lib().eb_model_get_hparams(h, hp, hf))
        m = cls.__new__(cls)
        m.E, m.H, m.F, m.V, m.L, m.is_eagle = hp[0], hp[1], hp[2], hp[3], hp[4], bool(hp[5])
        m.eps, m.freq_base, m.h = float(hf[0]), float(hf[1]), h
        m.load_seconds, m.load_bytes = sec.value, nb.value
        return m

    def set_tensor(self, layer, which, t, raw):
        raw = np.ascontiguousarray(raw)
        _ck(lib().eb_model_set_tensor(self.h, layer, which, t, _hp(raw)))

    def synth_tensor(self, layer, which, t, seed, scale=0.02):
        _ck(lib().eb_model_synth_tensor(self.h, layer, which, t, seed, scale))

    def get_tensor(self, layer, which, nbytes):
        out = np.empty(nbytes, np.uint8)
        _ck(lib().eb_model_get_tensor(self.h, layer, which, _hp(out), nbytes))
        return out

    def share_lm_head(self, target):
        _ck(lib().eb_model_share_lm_head(self.h, target.h))

    def weight_bytes(self):
        return int(lib().eb_model_weight_bytes(self.h))

    def free(self):
        if self.h:
            lib().eb_model_free(self.h)
            self.h = None


class ShardModel(Model):
    """One rank's shard of a target model for tensor parallelism (eb_model_create_tp): n_head_local heads of size D, F_local ffn rows"""

    def __init__(self, E, D, n_head_local, F_local, V, L, eps=1e-5, freq_base=10000.0):
        self.E, self.H, self.F, self.V, self.L, self.is_eagle = E, n_head_local, F_local, V, L, False
        self.D, self.Ea = D, n_head_local * D
        self.eps, self.freq_base = eps, freq_base
        self.h = lib().eb_model_create_tp(E, D, n_head_local, F_local, V, L, eps, freq_base)
        if not self.h:
            raise EagleB200Error(lib().eb_last_error().decode())


def tp_partition(H, F, size):
    """(head0, n_heads, f0, n_f) per rank: heads in pairs (256-element blocks of the K-split wo), ffn rows in 256-row super-blocks,
    as even as the block granularity allows (13B at 8 ranks: 6,6,6,6,4,4,4,4 heads; 7,7,7,7,7,7,6,6 ffn blocks)"""
    def split(n_units):
        base, extra = divmod(n_units, size)
        cnt = [base + (r < extra) for r in range(size)]
        off = [sum(cnt[:r]) for r in range(size)]
        return off, cnt
    assert H % 2 == 0 and F % 256 == 0 and H // 2 >= size and F // 256 >= size
    ho, hc = split(H // 2); fo, fc = split(F // 256)
    return [(2 * ho[r], 2 * hc[r], 256 * fo[r], 256 * fc[r]) for r in range(size)]


def tp_shard_tensor(which, t, raw, E, F, D, part):
    """this rank's slice of one per-layer tensor (raw ggml blocks, row-major [rows][cols/block]): wq|wk|wv and gate|up are split by
    output rows (N-split: whole heads / 256-row ffn super-blocks), wo and ffn_down by input columns (K-split, block aligned), the
    norms are replicated.  part = one entry of tp_partition()."""
    h0, nh, f0, nf = part

    def rows_of(rows, cols, r0, n):
        return np.ascontiguousarray(np.asarray(raw).reshape(rows, row_size(t, cols))[r0:r0 + n]).reshape(-1)

    def cols_of(rows, cols, c0, n):
        be, bb = BLOCK[t]
        assert c0 % be == 0 and n % be == 0, (c0, n, be)
        return np.ascontiguousarray(np.asarray(raw).reshape(rows, cols // be, bb)[:, c0 // be:(c0 + n) // be]).reshape(-1)

    if which in (L_WQ, L_WK, L_WV):
        return rows_of(E, E, h0 * D, nh * D)
    if which == L_WO:
        return cols_of(E, E, h0 * D, nh * D)
    if which in (L_GATE, L_UP):
        return rows_of(F, E, f0, nf)
    if which == L_DOWN:
        return cols_of(E, F, f0, nf)
    return raw


class TpGroup:
    """peer-memory communicator (csrc/tp.cu).  exchange(fn): fn(handle_bytes) -> list of every rank's handle bytes (the host program's
    transport, e.g. torch.distributed.all_gather_object)"""

    def __init__(self, rank, size, max_elems, exchange):
        self.rank, self.size = rank, size
        self.h = lib().eb_tp_create(rank, size, int(max_elems))
        if not self.h:
            raise EagleB200Error(lib().eb_last_error().decode())
        mine = C.create_string_buffer(64)
        _ck(lib().eb_tp_handle(self.h, mine))
        handles = exchange(bytes(mine.raw))
        for r, hb in enumerate(handles):
            if r != rank:
                _ck(lib().eb_tp_open_peer(self.h, r, C.create_string_buffer(hb, 64)))

    def allreduce(self, planes):
        """planes f32 [ks][n] (host) -> sum over ranks and planes [n] (host); test helper"""
        planes = np.ascontiguousarray(planes, np.float32)
        ks, n = planes.shape
        dp, do = DeviceBuffer.from_host(planes), DeviceBuffer(n * 4)
        _ck(lib().eb_tp_allreduce(self.h, dp.ptr, ks, n, do.ptr, None))
        _ck(lib().eb_synchronize(None))
        return do.to_host(np.float32, n)

    def error(self):
        return lib().eb_tp_error(self.h)

    def free(self):
        if self.h:
            lib().eb_tp_free(self.h); self.h = None


class Session:
    """One decoding stream: target + draft head + both KV caches + device-side tree state."""

    def __init__(self, target, head=None, n_ctx=4096, max_nodes=65, n_branch=8, n_draft=64, max_depth=8, top_k=10,
                 p_split=0.1, fixed_accept=-1, use_graph=True, p_min=0.0):
        self.target, self.head = target, head
        self.params = SessionParams(n_ctx, max_nodes, n_branch, n_draft, max_depth, top_k, p_split, fixed_accept, int(use_graph), p_min)
        self.h = lib().eb_session_create(target.h, head.h if head else None, C.byref(self.params))
        if not self.h:
            raise EagleB200Error(lib().eb_last_error().decode())

    def forward(self, model, tokens, pos, parent=None, hidden=None, want_logits=True):
        """llama_decode over a tree-shaped batch (host buffers in, host buffers out)."""
        tokens = np.ascontiguousarray(tokens, np.int32); pos = np.ascontiguousarray(pos, np.int32)
        T = tokens.size
        m = self.head if model else self.target
        par = None if parent is None else np.ascontiguousarray(parent, np.int32)
        hid = None if hidden is None else np.ascontiguousarray(hidden, np.float32)
        logits = np.empty((T, m.V), np.float32) if want_logits else None
        feat = np.empty((T, m.E), np.float32)
        _ck(lib().eb_session_forward(self.h, model, _hp(tokens), _hp(pos), _hp(par), _hp(hid), T, _hp(logits), _hp(feat)))
        return logits, feat

    def forward_batch(self, model, tokens, pos, seq_ids, hidden=None, want_logits=True):
        """the same forward from a tree-shaped llama_batch: seq_ids[i] = the seq-id list of token i (llama_batch.seq_id)"""
        tokens = np.ascontiguousarray(tokens, np.int32); pos = np.ascontiguousarray(pos, np.int32)
        T = tokens.size
        m = self.head if model else self.target
        n_seq, ptrs, keep = _seq_id_arrays(seq_ids)
        hid = None if hidden is None else np.ascontiguousarray(hidden, np.float32)
        logits = np.empty((T, m.V), np.float32) if want_logits else None
        feat = np.empty((T, m.E), np.float32)
        _ck(lib().eb_session_forward_batch(self.h, model, _hp(tokens), _hp(pos), _hp(n_seq), C.cast(ptrs, C.c_void_p), _hp(hid), T, _hp(logits), _hp(feat)))
        return logits, feat

    def forward_tapped(self, model, tokens, pos, parent=None, hidden=None):
        """forward + the int8 activation plane of every quantisation site (eb_session_set_tap: the counterpart of llama.cpp's
        cb_eval callback).  -> logits, feat, [plane [T][K] int8, ...] in site order."""
        m = self.head if model else self.target
        T = int(np.asarray(tokens).size)
        cap = T * ((2 * m.E if m.is_eagle else 0) + m.L * (3 * m.E + m.F) + m.E) + 1024
        buf = DeviceBuffer(cap)
        _ck(lib().eb_session_set_tap(self.h, buf.ptr, cap))
        try:
            logits, feat = self.forward(model, tokens, pos, parent=parent, hidden=hidden)
            n, sites = C.c_size_t(), C.c_int()
            _ck(lib().eb_session_tap_state(self.h, C.byref(n), C.byref(sites)))
            raw = np.empty(n.value, np.int8)
            _ck(lib().eb_download(_hp(raw), buf.ptr, n.value, None))
        finally:
            _ck(lib().eb_session_set_tap(self.h, None, 0))
            buf.free()
        widths = ([2 * m.E] if m.is_eagle else []) + [m.E, m.E, m.E, m.F] * m.L + [m.E]
        assert sites.value == len(widths) and n.value == T * sum(widths), (sites.value, len(widths), n.value)
        planes, off = [], 0
        for w in widths:
            planes.append(raw[off:off + T * w].reshape(T, w)); off += T * w
        return logits, feat, planes

    def set_tp(self, tp):
        _ck(lib().eb_session_set_tp(self.h, tp.h if tp else None))

    def prefill(self, tokens):
        """prompt of any length through target (+ head): -> (logits [V], feature [E]) of the last token; both caches committed"""
        tokens = np.ascontiguousarray(tokens, np.int32)
        lg = np.empty(self.target.V, np.float32); ft = np.empty(self.target.E, np.float32)
        _ck(lib().eb_session_prefill(self.h, _hp(tokens), tokens.size, _hp(lg), _hp(ft)))
        return lg, ft

    def kv_write(self, model, layer, which, cell0, rows):
        rows = np.ascontiguousarray(rows).view(np.uint16)
        _ck(lib().eb_session_kv_write(self.h, model, layer, which, cell0, rows.shape[0], _hp(rows)))

    def kv_read(self, model, layer, which, cell0, n_cells):
        m = self.head if model else self.target
        out = np.empty((n_cells, m.E), np.uint16)
        _ck(lib().eb_session_kv_read(self.h, model, layer, which, cell0, n_cells, _hp(out)))
        return out

    def set_n_past(self, n):
        _ck(lib().eb_session_set_n_past(self.h, n))

    def commit(self, model, n):
        _ck(lib().eb_session_commit(self.h, model, n))

    def n_past(self):
        return lib().eb_session_n_past(self.h, 0)

    def synth_kv(self, n_past, seed=1):
        _ck(lib().eb_session_synth_kv(self.h, n_past, seed))

    def set_root(self, id_last, feat=None):
        f = None if feat is None else np.ascontiguousarray(feat, np.float32)
        _ck(lib().eb_session_set_root(self.h, int(id_last), _hp(f)))

    def round(self):
        r = RoundResult()
        _ck(lib().eb_session_round(self.h, C.byref(r)))
        return r

    def gen_draft(self):
        """the draft half alone (common_speculative_gen_draft): -> (tokens, parent) of the drafted tree, node 0 = the root"""
        tok, par, n = np.zeros(128, np.int32), np.zeros(128, np.int32), C.c_int32()
        _ck(lib().eb_session_gen_draft(self.h, _hp(tok), _hp(par), 128, C.byref(n)))
        return tok[:n.value].copy(), par[:n.value].copy()

    def verify_only(self, T):
        ms = C.c_float()
        _ck(lib().eb_session_verify_only(self.h, T, C.byref(ms)))
        return ms.value

    def free(self):
        if self.h:
            lib().eb_session_free(self.h)
            self.h = None


def bench_mul_mat_kernel(t, N, K, T, n_bufs=8, iters=50):
    ms = C.c_float()
    _ck(lib().eb_bench_mul_mat_kernel(t, N, K, T, n_bufs, iters, C.byref(ms)))
    return ms.value


def bench_flash_attn(T, H, D, n_kv, n_bufs=4, iters=50):
    ms = C.c_float()
    _ck(lib().eb_bench_flash_attn(T, H, D, n_kv, n_bufs, iters, C.byref(ms)))
    return ms.value


def _seq_id_arrays(seq_ids):
    """llama_batch.n_seq_id / llama_batch.seq_id (int32 **) from a list of per-token seq-id lists"""
    keep = [np.ascontiguousarray(x, np.int32) for x in seq_ids]
    n_seq = np.array([len(x) for x in keep], np.int32)
    ptrs = (C.c_void_p * len(keep))(*[x.ctypes.data for x in keep])
    return n_seq, ptrs, keep


def tree_bits_from_batch(pos, seq_ids):
    """visibility sets [T][4] u32 of a tree-shaped llama_batch (eb_tree_bits_from_batch; host-only, needs no GPU)"""
    pos = np.ascontiguousarray(pos, np.int32)
    n_seq, ptrs, keep = _seq_id_arrays(seq_ids)
    out = np.zeros((len(keep), 4), np.uint32)
    _ck(lib().eb_tree_bits_from_batch(len(keep), _hp(pos), _hp(n_seq), C.cast(ptrs, C.c_void_p), _hp(out)))
    return out


def gguf_probe(path):
    """header of a GGUF file, no device needed -> dict(E, H, F, V, L, is_eagle, eps, freq_base, n_tensors, bytes)"""
    hp, hf, n, nb = (C.c_int32 * 6)(), (C.c_float * 2)(), C.c_int32(), C.c_uint64()
    _ck(lib().eb_gguf_probe(os.fsencode(path), hp, hf, C.byref(n), C.byref(nb)))
    return dict(E=hp[0], H=hp[1], F=hp[2], V=hp[3], L=hp[4], is_eagle=bool(hp[5]), eps=float(hf[0]), freq_base=float(hf[1]), n_tensors=n.value, bytes=nb.value)


def launch_count(reset=False):
    return int(lib().eb_launch_count(int(reset)))
