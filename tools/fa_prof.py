"""EB_FA_PROF=1 python tools/fa_prof.py — per-section cycles of the tcgen05 attention kernel at the 7B verify shape"""
import sys, numpy as np
sys.path.insert(0, "."); sys.path.insert(0, "tests"); sys.path.insert(0, "oracle")
from __graft_entry__ import load_package
import fixtures as fx
eb = load_package(); eb.init(0)
rng = np.random.default_rng(0)
T, H, D, n_past = 65, 32, 128, 4096
n_kv = n_past + 256
par = fx.tree_parents(T, rng, 6)
q = rng.normal(0, 1, (T, H, D)).astype(np.float32)
k = rng.normal(0, 1, (n_kv, H, D)).astype(np.float16); v = rng.normal(0, 1, (n_kv, H, D)).astype(np.float16)
for _ in range(3):
    eb.flash_attn_ext(q, k, v, 1 / np.sqrt(D), tree_bits=eb.tree_bits_from_parent(par), n_prefix=n_past)
