"""attention in isolation at full size: GPU (compact mask, tcgen05) vs float64 numpy vs the reference ggml-cpu (f16 / f32 accumulator)"""
import sys, os, numpy as np
ROOT=os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT,'tests')); sys.path.insert(0, os.path.join(ROOT,'oracle'))
import fixtures as fx, pyoracle as po
from conftest import load_package
eb = load_package(); eb.init(0)
H, D = 32, 128
def exact(q, k, v, vis, scale):
    q16 = q.astype(np.float16).astype(np.float64); k = k.astype(np.float64); v = v.astype(np.float64)
    T = q.shape[0]; out = np.zeros((T, H, D))
    for h in range(H):
        s = q16[:, h, :] @ k[:, h, :].T * scale
        s = np.where(vis > 0, s, -np.inf)
        p = np.exp(s - s.max(1, keepdims=True)); p /= p.sum(1, keepdims=True)
        out[:, h, :] = p @ v[:, h, :]
    return out
for n_prefix, T, sigma in ((4031, 65, 3.0), (4031, 33, 3.0), (1500, 65, 3.0), (4031, 65, 1.6), (4031, 65, 5.0), (8000, 65, 4.0), (8000, 128, 3.0), (500, 17, 6.0)):
    rng = np.random.default_rng(n_prefix + T)
    n_kv = n_prefix + T
    q = (rng.standard_normal((T, H, D)) * sigma).astype(np.float32)
    k = rng.standard_normal((n_kv, H, D)).astype(np.float16); v = rng.standard_normal((n_kv, H, D)).astype(np.float16)
    par = fx.driver_like_tree(T, 8, rng)
    vis = fx.vis_matrix(par, n_prefix, n_kv)
    tv = po.tree_vis(par)
    bits = np.zeros((T, 4), np.uint32)
    for t in range(T):
        for j in range(T):
            if tv[t, j]: bits[t, j >> 5] |= np.uint32(1 << (j & 31))
    scale = 1.0 / np.sqrt(D)
    g = eb.flash_attn_ext(q, k, v, scale, tree_bits=bits, n_prefix=n_prefix)
    ex = exact(q, k, v, vis, scale)
    r16 = r32 = ex
    errs = []
    for rep in range(12):
        g = eb.flash_attn_ext(q, k, v, scale, tree_bits=bits, n_prefix=n_prefix)
        errs.append(float(np.abs(g - ex).max() / np.abs(ex).max()))
    print("   repeats:", ["%.1e" % e for e in errs])
    def err(a): return float(np.abs(a - ex).max() / np.abs(ex).max())
    rowerr = np.abs(g - ex).max(axis=(1, 2)) / np.abs(ex).max()
    print("   worst rows", np.argsort(-rowerr)[:6].tolist(), np.round(np.sort(rowerr)[::-1][:6], 4).tolist(), "heads of worst", np.argsort(-np.abs(g - ex).max(axis=(0, 2)))[:4].tolist())
    print(f"n_prefix {n_prefix} T {T} sigma {sigma}: vs float64 exact: GPU {err(g):.2e} | oracle f16-acc {err(r16):.2e} | oracle f32-acc {err(r32):.2e}", flush=True)
