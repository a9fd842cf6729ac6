import sys, numpy as np
sys.path.insert(0, "."); sys.path.insert(0, "oracle"); sys.path.insert(0, "tests")
import pyoracle as po, fixtures as fx
from __graft_entry__ import load_package
eb = load_package(); eb.init(0)
rng = np.random.default_rng(0)
po.set_fattn_acc_f32(1)
for (T, n_past, H, sc) in [(10, 0, 2, 1.0), (10, 0, 2, 0.05), (10, 54, 2, 0.05), (65, 4031, 2, 0.05)]:
    D = 128; n_kv = max(64, (n_past + T + 63) // 64 * 64)
    par = np.arange(-1, T - 1).astype(np.int32)
    q = (rng.normal(0, sc, (T, H, D))).astype(np.float32)
    k = rng.normal(0, sc, (n_kv, H, D)).astype(np.float16); v = rng.normal(0, 1, (n_kv, H, D)).astype(np.float16)
    got = eb.flash_attn_ext(q, k, v, 1 / np.sqrt(D), tree_bits=eb.tree_bits_from_parent(par), n_prefix=n_past)
    want = po.oracle().flash_attn(q, k.view(np.uint16), v.view(np.uint16), fx.vis_matrix(par, n_past, n_kv), 1 / np.sqrt(D))
    print(T, n_past, sc, "max abs err per row", np.abs(got - want).reshape(T, -1).max(1)[:12], "max|want|", np.abs(want).max())
