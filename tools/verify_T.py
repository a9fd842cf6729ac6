import sys
sys.path.insert(0, ".")
import bench
from __graft_entry__ import load_package
eb = load_package(); eb.init(0)
tgt, head = bench.build_gpu_models(eb)
s = eb.Session(tgt, head, n_ctx=4096 + 256, fixed_accept=3, use_graph=1, **bench.TREE)
s.synth_kv(bench.N_PAST0, seed=3); s.set_root(1)
for T in (8, 65):
    v = sorted(s.verify_only(T) for _ in range(5))
    print("T", T, "verify ms", v[2])
