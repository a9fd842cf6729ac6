"""Shape of the drafted tree in the bench configuration: nodes per depth (how many draft levels do real work)."""
import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np
import bench
from __graft_entry__ import load_package
eb = load_package(); eb.init(0)
tgt, head = bench.build_gpu_models(eb)
s = eb.Session(tgt, head, n_ctx=4096 + 256, fixed_accept=bench.FIXED_ACCEPT, use_graph=True, **bench.TREE)
s.synth_kv(bench.N_PAST0, seed=1); tok = 1
for i in range(6):
    s.set_root(tok); r = s.round(); tok = r.bonus
    par = np.array(r.parent[:r.n_nodes]); dep = np.zeros(r.n_nodes, int)
    for j in range(1, r.n_nodes): dep[j] = dep[par[j]] + 1
    print(f"round {i}: {r.n_nodes} nodes, per depth {np.bincount(dep).tolist()}, {r.gpu_ms:.3f} ms")
