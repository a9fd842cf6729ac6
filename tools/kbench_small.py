"""T = 8 GEMM shapes of one draft level (ncu target)"""
import sys
sys.path.insert(0, ".")
from __graft_entry__ import load_package
eb = load_package(); eb.init(0)
iters = int(sys.argv[1]) if len(sys.argv) > 1 else 20
for (t, name, N, K, T) in [(eb.Q4_K, "q4_K", 4096, 8192, 8), (eb.Q4_K, "q4_K", 12288, 4096, 8), (eb.Q4_K, "q4_K", 4096, 4096, 8), (eb.Q4_K, "q4_K", 22016, 4096, 8),
                           (eb.Q6_K, "q6_K", 4096, 11008, 8), (eb.Q4_K, "q4_K", 4096, 11008, 8), (eb.Q6_K, "q6_K", 32000, 4096, 8)]:
    ms = eb.bench_mul_mat_kernel(t, N, K, T, n_bufs=8, iters=iters)
    print(f"{name} N={N} K={K} T={T}: {ms*1e3:.1f} us/launch  {eb.row_size(t, K) * N / 1e9 / ms * 1e3:.0f} GB/s", flush=True)
