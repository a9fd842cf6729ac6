"""tc GEMM ablations: EB_TC_DBG bit0 = no B reloads, bit1 = no W reloads, bit2 = no epilogue math"""
import sys
sys.path.insert(0, ".")
from __graft_entry__ import load_package
eb = load_package(); eb.init(0)
for (N, K, T) in [(11008, 4096, 65), (4096, 4096, 65), (4096, 11008, 65), (11008, 4096, 16)]:
    ms = eb.bench_mul_mat_kernel(eb.Q4_K, N, K, T, n_bufs=8, iters=100)
    gb = eb.row_size(eb.Q4_K, K) * N / 1e9
    print(f"N={N} K={K} T={T}: {ms*1e3:.1f} us/launch  {gb/ms*1e3:.0f} GB/s", flush=True)
