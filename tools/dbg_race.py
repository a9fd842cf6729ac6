import sys, numpy as np
sys.path.insert(0, "."); sys.path.insert(0, "tests"); sys.path.insert(0, "oracle")
from __graft_entry__ import load_package
import fixtures as fx, pyoracle as po
eb = load_package(); eb.init(0)
rng = np.random.default_rng(5)
for (N, K, T) in [(512, 512, 40), (1024, 512, 40), (512, 1024, 40), (4096, 4096, 65), (2048, 512, 40)]:
    raw = fx.quantize(po.Q4_K, rng.normal(0, 0.02, (N, K)).astype(np.float32), rng)
    w = eb.Weights(po.Q4_K, raw, N, K)
    X = rng.normal(0, 1, (T, K)).astype(np.float32)
    want = po.oracle().mul_mat(po.Q4_K, raw, N, K, X)
    bad = 0; worst = 0.0
    for it in range(300):
        got = eb.mul_mat(w, X)
        e = np.abs(got - want).max() / np.abs(want).max()
        worst = max(worst, e)
        if e > 1e-5:
            bad += 1
            if bad <= 3:
                d = np.abs(got - want); r, c = np.unravel_index(d.argmax(), d.shape)
                rows = np.unique(np.nonzero(d > 1e-5 * np.abs(want).max())[0]); cols = np.unique(np.nonzero(d > 1e-5 * np.abs(want).max())[1])
                print(f"  it {it}: err {e:.2e} at token {r} row {c}; bad tokens {rows[:10]} n={len(rows)}; bad rows {cols[:10]} n={len(cols)}")
    print(f"N={N} K={K} T={T}: {bad}/300 bad, worst {worst:.2e}", flush=True)
