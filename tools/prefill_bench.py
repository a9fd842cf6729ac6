"""N1: wall time of eb_session_prefill (target + head, 7B Q4_K_M synthetic weights) for a 2048-token prompt, at the given rows per pass
(EB_PREFILL_CHUNK is read once per process).  usage: EB_PREFILL_CHUNK=128 python tools/prefill_bench.py [n_tokens=2048]"""
import os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np
import bench
from __graft_entry__ import load_package
eb = load_package(); eb.init(0)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
tgt, head = bench.build_gpu_models(eb)
s = eb.Session(tgt, head, n_ctx=4096 + 256, use_graph=False, **bench.TREE)
toks = np.random.default_rng(1).integers(3, 32000, n).astype(np.int32)
ts = []
for rep in range(3):
    s.set_n_past(0)
    t0 = time.perf_counter(); s.prefill(toks); ts.append(time.perf_counter() - t0)
print(f"prefill {n} tokens, rows per pass {os.environ.get('EB_PREFILL_CHUNK', 'default')}: {min(ts) * 1e3:.1f} ms = {n / min(ts):.0f} tok/s (best of 3: {[round(t * 1e3, 1) for t in ts]})")
