"""one 7B Q4_K_M verify step (T=65, n_kv=4096) without CUDA graph — for ncu launch lists"""
import sys
sys.path.insert(0, ".")
import bench
from __graft_entry__ import load_package
eb = load_package(); eb.init(0)
tgt, head = bench.build_gpu_models(eb)
s = eb.Session(tgt, head, n_ctx=4096 + 256, fixed_accept=3, use_graph=int(sys.argv[1]) if len(sys.argv) > 1 else 0, **bench.TREE)
s.synth_kv(bench.N_PAST0, seed=3); s.set_root(1)
for _ in range(3):
    print("verify ms", s.verify_only(65))
r = s.round(); print("round ms", r.gpu_ms, "nodes", r.n_nodes)
r = s.round(); print("round ms", r.gpu_ms, "nodes", r.n_nodes)
