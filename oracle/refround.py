"""ctypes bindings for oracle/_ref/libref_round.so: the REFERENCE's own libllama + common behind a C harness
(oracle/ref_round.cpp).  TEST INFRASTRUCTURE ONLY (tests/, smoke, bench.py's CPU arm)."""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.path.join(HERE, "_ref")
_lib = None


def have():
    return os.path.exists(os.path.join(REF, "libref_round.so"))


def native_ok():
    """libggml-native.so was built with -march=native on the build machine; it may only be loaded on a host whose ISA covers that
    machine's (oracle/_ref/native.cpuflags, written by oracle/Makefile)"""
    f = os.path.join(REF, "native.cpuflags")
    if not os.path.exists(f) or not os.path.exists(os.path.join(REF, "libggml-native.so")):
        return False
    need = set(open(f).read().split())
    have_ = set()
    for line in open("/proc/cpuinfo"):
        if line.startswith("flags"):
            have_ = set(line.split(":", 1)[1].split()); break
    isa = {x for x in need if x.startswith(("avx", "amx", "fma", "f16c", "bmi", "sse4", "gfni", "vaes", "vpclmul"))}
    return isa <= have_


VARIANTS = {"ref": "libggml-ref.so",          # the reference's ggml-cpu, deterministic flags: THE parity oracle
            "native": "libggml-native.so",    # same sources, -march=native + llamafile + repack: the timed CPU arm only
            "f32acc": "libggml-f32acc.so"}    # attribution instrument: f32 flash-attention V accumulator, nothing else changed
_variant = None


def lib(native=False, variant=None):
    """One ggml build per process (all three carry the soname libggml-ref.so; the first one loaded serves libllama-ref.so).
    variant: "ref" (default) | "native" | "f32acc"; the environment variable EB_REF_GGML overrides the default."""
    global _lib, _variant
    want = variant or ("native" if native else os.environ.get("EB_REF_GGML", "ref"))
    if _lib is not None and want != _variant and (variant or native):
        raise RuntimeError(f"refround: ggml variant {_variant!r} is already loaded in this process, cannot switch to {want!r}")
    if _lib is None:
        _variant = want
        C.CDLL(os.path.join(REF, VARIANTS[want]), mode=C.RTLD_GLOBAL)
        l = C.CDLL(os.path.join(REF, "libref_round.so"), mode=C.RTLD_GLOBAL)
        vp, i32, i64 = C.c_void_p, C.c_int, C.c_int64
        l.rr_load_backend.argtypes = [C.c_char_p]
        l.rr_dev_name.argtypes = [i32]; l.rr_dev_name.restype = C.c_char_p
        l.rr_model_load.argtypes = [C.c_char_p, i32]; l.rr_model_load.restype = vp
        l.rr_model_free.argtypes = [vp]
        l.rr_model_n_vocab.argtypes = [vp]
        l.rr_model_tensor_type.argtypes = [vp, C.c_char_p]
        l.rr_ctx_new.argtypes = [vp, i32, i32, i32, i32, i32, i32]; l.rr_ctx_new.restype = vp
        l.rr_ctx_free.argtypes = [vp]
        l.rr_decode.argtypes = [vp, i32, vp, vp, vp, vp, i32, vp]
        l.rr_tap_enable.argtypes = [vp, i32]; l.rr_tap_count.argtypes = [vp]
        l.rr_tap_info.argtypes = [vp, i32, C.c_char_p, i32, vp, vp]; l.rr_tap_data.argtypes = [vp, i32, vp]
        l.rr_kq_mask.argtypes = [vp, vp, i64, vp, vp]; l.rr_kq_mask.restype = i64
        l.rr_set_logits.argtypes = [vp, vp, i32]
        l.rr_draft_candidates.argtypes = [vp, i32, i32, vp, vp, vp]
        l.rr_accept_n.argtypes = [vp, vp, i32, vp]
        l.rr_kv_seq_rm.argtypes = [vp, i32, i32, i32]
        l.rr_kv_seq_cp.argtypes = [vp, i32, i32, i32, i32]
        l.rr_kv_seq_keep.argtypes = [vp, i32]
        l.rr_kv_clear.argtypes = [vp]
        l.rr_kv_defrag.argtypes = [vp]
        l.rr_kv_cells.argtypes = [vp, vp, vp, i32]
        l.rr_kv_read_row.argtypes = [vp, i32, i32, i32, vp, i32]
        l.rr_kv_write_rows.argtypes = [vp, i32, i32, i32, i32, vp, i32]; l.rr_kv_set_prefix.argtypes = [vp, i32, i32]
        l.rr_time_decode.argtypes = [vp, i32, vp, vp, vp, vp, i32]; l.rr_time_decode.restype = i64
        _lib = l
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def seq_arrays(seq_sets, max_seq=8):
    n = len(seq_sets)
    n_seq = np.array([len(s) for s in seq_sets], np.int32)
    seq = np.zeros((n, max_seq), np.int32)
    for i, s in enumerate(seq_sets):
        seq[i, :len(s)] = list(s)
    return n_seq, seq


class Model:
    def __init__(self, path, n_gpu_layers=0, native=False, variant=None):
        self.l = lib(native, variant)
        self.h = self.l.rr_model_load(path.encode(), n_gpu_layers)
        if not self.h:
            raise RuntimeError(f"reference loader rejected {path}")
        self.V = self.l.rr_model_n_vocab(self.h)

    def tensor_type(self, name):
        return self.l.rr_model_tensor_type(self.h, name.encode())

    def free(self):
        if self.h:
            self.l.rr_model_free(self.h); self.h = None


class Context:
    def __init__(self, model, n_ctx, n_batch=512, n_seq_max=8, n_threads=4, flash_attn=True, type_kv=1):
        self.m, self.l = model, model.l
        self.h = self.l.rr_ctx_new(model.h, n_ctx, n_batch, n_seq_max, n_threads, int(flash_attn), type_kv)
        if not self.h:
            raise RuntimeError("llama_init_from_model failed")
        self.max_seq = n_seq_max

    def decode(self, tokens, pos, seq_sets=None, want_logits=True):
        tokens = np.ascontiguousarray(tokens, np.int32); pos = np.ascontiguousarray(pos, np.int32)
        n = tokens.size
        n_seq, seq = seq_arrays(seq_sets if seq_sets is not None else [[0]] * n, self.max_seq)
        out = np.empty((n, self.m.V), np.float32) if want_logits else None
        rc = self.l.rr_decode(self.h, n, _p(tokens), _p(pos), _p(n_seq), _p(seq), self.max_seq, _p(out))
        if rc != 0:
            raise RuntimeError(f"llama_decode returned {rc}")
        return out

    def time_decode(self, tokens, pos, seq_sets=None):
        tokens = np.ascontiguousarray(tokens, np.int32); pos = np.ascontiguousarray(pos, np.int32)
        n = tokens.size
        n_seq, seq = seq_arrays(seq_sets if seq_sets is not None else [[0]] * n, self.max_seq)
        us = self.l.rr_time_decode(self.h, n, _p(tokens), _p(pos), _p(n_seq), _p(seq), self.max_seq)
        if us < 0:
            raise RuntimeError("llama_decode failed")
        return us

    def decode_tapped(self, tokens, pos, seq_sets=None):
        """llama_decode + the f32 activation operand of every quantised MUL_MAT (graph order): [(weight name, x [T][K])]"""
        self.l.rr_tap_enable(self.h, 1)
        try:
            logits = self.decode(tokens, pos, seq_sets)
        finally:
            n = self.l.rr_tap_count(self.h)
            taps = []
            for i in range(n):
                name = C.create_string_buffer(128); K, T = C.c_int64(), C.c_int64()
                self.l.rr_tap_info(self.h, i, name, 128, C.byref(K), C.byref(T))
                x = np.empty((T.value, K.value), np.float32)
                self.l.rr_tap_data(self.h, i, _p(x))
                taps.append((name.value.decode(), x))
            self.l.rr_tap_enable(self.h, 0)
        return logits, taps

    def kq_mask(self):
        ne0, ne1 = C.c_int64(), C.c_int64()
        n = self.l.rr_kq_mask(self.h, None, 0, C.byref(ne0), C.byref(ne1))
        out = np.empty(n, np.float32)
        self.l.rr_kq_mask(self.h, _p(out), n, C.byref(ne0), C.byref(ne1))
        return out.reshape(ne1.value, ne0.value)

    def set_logits(self, logits):
        logits = np.ascontiguousarray(logits, np.float32).reshape(-1, self.m.V)
        self.l.rr_set_logits(self.h, _p(logits), logits.shape[0])

    def draft_candidates(self, idx, k):
        ids = np.empty(k, np.int32); lg = np.empty(k, np.float32); pr = np.empty(k, np.float32)
        n = self.l.rr_draft_candidates(self.h, idx, k, _p(ids), _p(lg), _p(pr))
        return ids[:n], lg[:n], pr[:n]

    def accept_n(self, draft):
        draft = np.ascontiguousarray(draft, np.int32)
        out = np.empty(draft.size + 1, np.int32)
        n = self.l.rr_accept_n(self.h, _p(draft), draft.size, _p(out))
        return out[:n].copy()

    def kv_cells(self, cap=65536):
        pos = np.empty(cap, np.int32); seq = np.empty(cap, np.uint64)
        n = self.l.rr_kv_cells(self.h, _p(pos), _p(seq), cap)
        return pos[:n].copy(), seq[:n].copy()

    def kv_row(self, layer, which, cell, E):
        out = np.empty(E, np.uint16)
        if self.l.rr_kv_read_row(self.h, layer, which, cell, _p(out), E) != 0:
            raise RuntimeError("kv_read_row: cache is not f16 / V is transposed")
        return out

    def kv_write(self, layer, which, cell0, rows):
        rows = np.ascontiguousarray(rows).view(np.uint16)
        if self.l.rr_kv_write_rows(self.h, layer, which, cell0, rows.shape[0], _p(rows), rows.shape[1]) != 0:
            raise RuntimeError("kv_write: cache is not f16 / V is transposed")

    def kv_set_prefix(self, n, n_seq=1):
        if self.l.rr_kv_set_prefix(self.h, n, n_seq) != 0:
            raise RuntimeError("kv_set_prefix")

    def seq_rm(self, s, p0, p1): self.l.rr_kv_seq_rm(self.h, s, p0, p1)
    def seq_cp(self, a, b, p0=-1, p1=-1): self.l.rr_kv_seq_cp(self.h, a, b, p0, p1)
    def seq_keep(self, s): self.l.rr_kv_seq_keep(self.h, s)
    def kv_clear(self): self.l.rr_kv_clear(self.h)
    def defrag(self): self.l.rr_kv_defrag(self.h)

    def free(self):
        if self.h:
            self.l.rr_ctx_free(self.h); self.h = None
