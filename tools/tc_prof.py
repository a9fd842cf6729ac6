import sys, numpy as np
sys.path.insert(0, "."); sys.path.insert(0, "tests"); sys.path.insert(0, "oracle")
from __graft_entry__ import load_package
import fixtures as fx, pyoracle as po
eb = load_package(); eb.init(0)
rng = np.random.default_rng(0)
for (N, K, T) in [(22016, 4096, 65), (4096, 4096, 65), (4096, 11008, 65)]:
    raw = fx.random_blocks(po.Q4_K, N * K, rng)
    w = eb.Weights(po.Q4_K, raw, N, K)
    X = rng.normal(0, 1, (T, K)).astype(np.float32)
    for _ in range(3):
        eb.mul_mat(w, X)
