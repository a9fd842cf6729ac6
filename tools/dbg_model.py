import sys, numpy as np
sys.path.insert(0, "."); sys.path.insert(0, "tests"); sys.path.insert(0, "oracle")
from __graft_entry__ import load_package
import fixtures as fx, pyoracle as po
eb = load_package(); eb.init(0)
hp = dict(E=512, H=4, F=1024, V=2048, L=3)
tgt = fx.make_model(**hp, ftype="q4_k_m", seed=21)
o = po.oracle(); n_ctx = 256
mo = fx.to_cpu_model(o, tgt)
mg = fx.to_gpu_model(eb, tgt)
po.set_fattn_acc_f32(1)
# single-token rows have no attention mixing: every row of a batch of INDEPENDENT first tokens (parent = -1) is a pure GEMM chain
P = 40
flips = 0; rows = 0; errs = []
for seed in range(12):
    kv = o._g("kv_new")(hp["L"], hp["E"], n_ctx)
    s = eb.Session(mg, None, n_ctx=n_ctx, max_nodes=65, use_graph=False)
    rng = np.random.default_rng(1000 + seed)
    prompt = rng.integers(3, hp["V"], P).astype(np.int32)
    par = np.full(P, -1, np.int32)
    vis = np.zeros((P, n_ctx), np.uint8); vis[np.arange(P), np.arange(P)] = 1
    lo, fo = o.forward(mo, kv, prompt, np.zeros(P, np.int32), None, 0, n_ctx, vis, hp["E"], hp["V"])
    lg, fg = s.forward(0, prompt, np.zeros(P, np.int32), parent=par)
    err = np.abs(lg - lo).max(1) / np.abs(lo).max()
    flips += int((err > 1e-4).sum()); rows += P; errs.append(err.max())
    s.free()
print(f"rows with a visible deviation: {flips}/{rows}; worst {max(errs):.2e}; median of per-seed worst {np.median(errs):.2e}")
