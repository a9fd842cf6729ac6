"""tree attention kernel alone (7B verify shape): us per call over rotating K/V buffers; EB_FA_PROF=1 adds the per-section cycles"""
import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from __graft_entry__ import load_package
eb = load_package(); eb.init(0)
for T, n_kv in ((65, 4096), (65, 1024), (8, 4096)):
    ms = eb.bench_flash_attn(T, 32, 128, n_kv, n_bufs=4, iters=50)
    print(f"T {T} n_kv {n_kv}: {ms * 1000:.2f} us  ({n_kv * 2 * 8192 / ms / 1e6:.0f} GB/s)")
