"""ctypes bindings for the CPU oracle (oracle/libeagle_oracle.so, prefix ``eo_``) and for the
reference harness (oracle/_ref/libref_harness.so, prefix ``ref_``).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  The product package never imports this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))

F32, F16, Q4_0, Q8_0, Q4_K, Q6_K = 0, 1, 2, 8, 12, 14
TYPE_NAMES = {F32: "f32", F16: "f16", Q4_0: "q4_0", Q8_0: "q8_0", Q4_K: "q4_K", Q6_K: "q6_K"}
BLOCK = {F32: (1, 4), F16: (1, 2), Q4_0: (32, 18), Q8_0: (32, 34), Q4_K: (256, 144), Q6_K: (256, 210)}
Q8_K_BYTES = 292  # block_q8_K

# tensor slots (same numbering in eagle_oracle.h, ref_harness.cpp and include/eagle_b200.h)
T_TOK_EMBD, T_OUTPUT_NORM, T_OUTPUT, T_FC, T_FC_BIAS = range(5)
L_ATTN_NORM, L_WQ, L_WK, L_WV, L_WO, L_FFN_NORM, L_GATE, L_UP, L_DOWN = range(9)


def row_size(t, k):
    be, bb = BLOCK[t]
    assert k % be == 0
    return k // be * bb


def build(ref=True):
    """make the oracle (and oracle/_ref when the reference tree is present)"""
    subprocess.check_call(["make", "-s", "-C", HERE, "all" if ref else "libeagle_oracle.so"])


def _ptr(a):
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        assert a.flags["C_CONTIGUOUS"]
        return a.ctypes.data_as(C.c_void_p)
    return a


class CpuPath:
    """One CPU implementation of the path: prefix 'eo_' (restatement) or 'ref_' (reference)."""

    def __init__(self, lib, prefix):
        self.lib, self.p = lib, prefix
        self.kind = "port" if prefix == "eo_" else "reference"
        g = lambda n: getattr(lib, prefix + n)
        vp, i32, i64, f32 = C.c_void_p, C.c_int, C.c_int64, C.c_float
        g("dequantize_row").argtypes = [i32, vp, vp, i64]
        g("mul_mat").argtypes = [i32, vp, i64, i64, vp, i64, vp]
        g("rms_norm").argtypes = [vp, vp, i64, i64, f32]
        g("rope_norm").argtypes = [vp, vp, i32, i32, i32, f32]
        g("flash_attn").argtypes = [vp, vp, vp, vp, i32, i32, i32, i32, f32, vp]
        g("vec_dot").argtypes = [i32, i64, vp, vp]
        g("vec_dot").restype = f32
        g("model_new").argtypes = [i32, i32, i32, i32, i32, f32, f32, i32]
        g("model_new").restype = vp
        g("model_free").argtypes = [vp]
        g("model_set_global").argtypes = [vp, i32, i32, vp]
        g("model_set_layer").argtypes = [vp, i32, i32, i32, vp]
        g("model_set_lm_head_from").argtypes = [vp, vp]
        g("kv_new").argtypes = [i32, i32, i32]
        g("kv_new").restype = vp
        g("kv_free").argtypes = [vp]
        for n in ("kv_k", "kv_v"):
            g(n).argtypes = [vp, i32]
            g(n).restype = vp
        g("forward").argtypes = [vp, vp, vp, vp, vp, i32, i32, i32, vp, vp, vp]
        if prefix == "eo_":
            g("quantize_row_q8_K").argtypes = [vp, vp, i64]
            g("quantize_row_q8_0").argtypes = [vp, vp, i64]
            g("get_rows").argtypes = [i32, vp, i64, vp, i32, vp]
            g("build_kq_mask").argtypes = [vp, vp, i32, vp, vp, i32, vp]
            g("top_k").argtypes = [vp, i32, i32, vp, vp, vp]
            g("argmax").argtypes = [vp, i32]
            g("tree_vis").argtypes = [vp, i32, vp]
            g("accept_walk").argtypes = [vp, vp, i32, vp, vp, vp]
            g("kv_compact").argtypes = [vp, i32, vp, i32]
            g("set_tap").argtypes = [vp, C.c_size_t]
            g("tap_bytes").restype = C.c_size_t
        else:
            g("quantize").argtypes = [i32, vp, vp, i64, i64]
            g("quantize").restype = C.c_size_t
            g("quantize_act").argtypes = [i32, vp, vp, i64]
            g("get_rows").argtypes = [i32, vp, i64, i64, vp, i32, vp]
            g("set_threads").argtypes = [i32]
        self._g = g

    # ---- formats
    def dequantize(self, t, raw, n):
        out = np.empty(n, np.float32)
        self._g("dequantize_row")(t, _ptr(raw), _ptr(out), n)
        return out

    def quantize_act(self, kind, x):
        """kind: Q8_0 or 'q8_K' (ggml type 15)"""
        x = np.ascontiguousarray(x, np.float32)
        k = x.size
        if kind == Q8_0:
            out = np.empty(k // 32 * 34, np.uint8)
            if self.p == "eo_":
                self._g("quantize_row_q8_0")(_ptr(x), _ptr(out), k)
            else:
                self._g("quantize_act")(8, _ptr(x), _ptr(out), k)
        else:
            out = np.empty(k // 256 * Q8_K_BYTES, np.uint8)
            if self.p == "eo_":
                self._g("quantize_row_q8_K")(_ptr(x), _ptr(out), k)
            else:
                self._g("quantize_act")(15, _ptr(x), _ptr(out), k)
        return out

    def vec_dot(self, t, k, w, a):
        return float(self._g("vec_dot")(t, k, _ptr(w), _ptr(a)))

    # ---- ops
    def mul_mat(self, t, W, N, K, X):
        X = np.ascontiguousarray(X, np.float32).reshape(-1, K)
        Y = np.empty((X.shape[0], N), np.float32)
        self._g("mul_mat")(t, _ptr(W), N, K, _ptr(X), X.shape[0], _ptr(Y))
        return Y

    def rms_norm(self, x, eps):
        x = np.ascontiguousarray(x, np.float32)
        y = np.empty_like(x)
        self._g("rms_norm")(_ptr(x), _ptr(y), x.shape[-1], x.size // x.shape[-1], eps)
        return y

    def rope_norm(self, x, pos, freq_base):
        """x [T][H][D] -> rotated copy"""
        x = np.array(x, np.float32, order="C")
        pos = np.ascontiguousarray(pos, np.int32)
        T, H, D = x.shape
        self._g("rope_norm")(_ptr(x), _ptr(pos), D, H, T, freq_base)
        return x

    def get_rows(self, t, W, K, n_rows, ids):
        ids = np.ascontiguousarray(ids, np.int32)
        out = np.empty((ids.size, K), np.float32)
        if self.p == "eo_":
            self._g("get_rows")(t, _ptr(W), K, _ptr(ids), ids.size, _ptr(out))
        else:
            self._g("get_rows")(t, _ptr(W), K, n_rows, _ptr(ids), ids.size, _ptr(out))
        return out

    def flash_attn(self, q, k, v, vis, scale):
        """q [T][H][D] f32, k/v [n_kv][H][D] f16 (as uint16), vis [T][n_kv] u8"""
        q = np.ascontiguousarray(q, np.float32)
        T, H, D = q.shape
        n_kv = k.shape[0]
        vis = np.ascontiguousarray(vis, np.uint8)
        out = np.empty_like(q)
        self._g("flash_attn")(_ptr(q), _ptr(k), _ptr(v), _ptr(vis), T, H, D, n_kv, scale, _ptr(out))
        return out

    # ---- model
    def model_new(self, E, H, F, V, L, eps, freq_base, is_eagle):
        return self._g("model_new")(E, H, F, V, L, eps, freq_base, int(is_eagle))

    def forward(self, m, kv, tokens, pos, hidden_in, kv_head, n_kv, vis, E, V, want_logits=True):
        tokens = np.ascontiguousarray(tokens, np.int32)
        pos = np.ascontiguousarray(pos, np.int32)
        T = tokens.size
        vis = np.ascontiguousarray(vis, np.uint8)
        assert vis.shape == (T, n_kv)
        logits = np.empty((T, V), np.float32) if want_logits else None
        feat = np.empty((T, E), np.float32)
        hid = None if hidden_in is None else np.ascontiguousarray(hidden_in, np.float32)
        self._g("forward")(m, kv, _ptr(tokens), _ptr(pos), _ptr(hid), T, kv_head, n_kv, _ptr(vis),
                           _ptr(logits), _ptr(feat))
        return logits, feat

    def forward_tapped(self, m, kv, tokens, pos, hidden_in, kv_head, n_kv, vis, E, V, widths):
        """eo_ only: forward + the int8 activation quants [T][K] of every quantisation site (widths = K per site, in order)"""
        assert self.p == "eo_"
        T = int(np.asarray(tokens).size)
        buf = np.zeros(T * sum(widths), np.int8)
        self._g("set_tap")(_ptr(buf), buf.size)
        try:
            logits, feat = self.forward(m, kv, tokens, pos, hidden_in, kv_head, n_kv, vis, E, V)
            assert self._g("tap_bytes")() == buf.size, (self._g("tap_bytes")(), buf.size)
        finally:
            self._g("set_tap")(None, 0)
        planes, off = [], 0
        for w in widths:
            planes.append(buf[off:off + T * w].reshape(T, w)); off += T * w
        return logits, feat, planes

    def kv_arrays(self, kv, L, E, n_ctx):
        ks, vs = [], []
        for l in range(L):
            for name, dst in (("kv_k", ks), ("kv_v", vs)):
                p = self._g(name)(kv, l)
                dst.append(np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint16)), shape=(n_ctx, E)))
        return ks, vs


_cache = {}


def oracle():
    if "eo" not in _cache:
        path = os.path.join(HERE, "libeagle_oracle.so")
        if not os.path.exists(path):
            build(ref=False)
        _cache["eo"] = CpuPath(C.CDLL(path), "eo_")
    return _cache["eo"]


def have_ref():
    return os.path.exists(os.path.join(HERE, "_ref", "libref_harness.so"))


def ref():
    if "ref" not in _cache:
        C.CDLL(os.path.join(HERE, "_ref", "libggml-ref.so"), mode=C.RTLD_GLOBAL)
        _cache["ref"] = CpuPath(C.CDLL(os.path.join(HERE, "_ref", "libref_harness.so")), "ref_")
    return _cache["ref"]


def set_fattn_acc_f32(on):
    """oracle only: f32 V accumulator in flash attention (checker for the GPU deviation)"""
    oracle().lib.eo_set_fattn_acc_f32(int(on))


class fattn_acc_f32:
    def __enter__(self):
        set_fattn_acc_f32(1)

    def __exit__(self, *a):
        set_fattn_acc_f32(0)


# ---- oracle-only helpers (speculative-step pieces)
def top_k(logits, k):
    o = oracle()
    logits = np.ascontiguousarray(logits, np.float32)
    ids = np.empty(k, np.int32); vals = np.empty(k, np.float32); probs = np.empty(k, np.float32)
    o._g("top_k")(_ptr(logits), logits.size, k, _ptr(ids), _ptr(vals), _ptr(probs))
    return ids, vals, probs


def tree_vis(parent):
    parent = np.ascontiguousarray(parent, np.int32)
    T = parent.size
    out = np.empty((T, T), np.uint8)
    oracle()._g("tree_vis")(_ptr(parent), T, _ptr(out))
    return out


def accept_walk(tokens, parent, best):
    tokens = np.ascontiguousarray(tokens, np.int32); parent = np.ascontiguousarray(parent, np.int32)
    best = np.ascontiguousarray(best, np.int32)
    T = tokens.size
    path = np.empty(T, np.int32); bonus = np.zeros(1, np.int32)
    n = oracle()._g("accept_walk")(_ptr(tokens), _ptr(parent), T, _ptr(best), _ptr(path), _ptr(bonus))
    return path[:n].copy(), int(bonus[0])


def build_kq_mask(cell_pos, cell_seq, q_pos, q_seq0):
    cell_pos = np.ascontiguousarray(cell_pos, np.int32); cell_seq = np.ascontiguousarray(cell_seq, np.uint64)
    q_pos = np.ascontiguousarray(q_pos, np.int32); q_seq0 = np.ascontiguousarray(q_seq0, np.int32)
    vis = np.empty((q_pos.size, cell_pos.size), np.uint8)
    oracle()._g("build_kq_mask")(_ptr(cell_pos), _ptr(cell_seq), cell_pos.size, _ptr(q_pos), _ptr(q_seq0), q_pos.size, _ptr(vis))
    return vis
