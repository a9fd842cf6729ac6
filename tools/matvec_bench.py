"""T = 1 matvec (mmvq.cu) per weight format: us per launch by CUDA-graph replay over 8 weight buffers, and GB/s of weight bytes"""
import sys; sys.path.insert(0, ".")
from __graft_entry__ import load_package
eb = load_package(); eb.init(0)
BPW = {eb.Q4_K: 144/256, eb.Q6_K: 210/256, eb.Q4_0: 18/32, eb.Q8_0: 34/32}
for name, t, N, K in (("q4_K", eb.Q4_K, 22016, 4096), ("q6_K", eb.Q6_K, 32000, 4096), ("q4_0", eb.Q4_0, 22016, 4096), ("q8_0", eb.Q8_0, 22016, 4096), ("q4_K", eb.Q4_K, 4096, 11008)):
    ms = eb.bench_mul_mat_kernel(t, N, K, 1, n_bufs=8, iters=100)
    print(f"matvec T=1 {name} {N}x{K}: {ms*1e3:.2f} us, {N*K*BPW[t]/(ms*1e-3)/1e9:.0f} GB/s")
