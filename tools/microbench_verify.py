"""BASELINE config 5 — verify-only micro-bench: tree-masked attention (H 32, D 128, f16 KV) and quantised mul_mat at tree sizes
16 / 32 / 64 and KV lengths 1K-8K, CUDA-event timed (graph replay, operands rotated through buffers larger than L2), next to the
reference's ggml-cpu backend (oracle/_ref, all host threads) on the same shapes.  Prints a markdown table.
usage: python tools/microbench_verify.py [--no-cpu]"""
import os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "oracle"))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
import numpy as np
from __graft_entry__ import load_package

eb = load_package(); eb.init(0)
HBM = 6462.4
try:
    import json
    HBM = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass
cpu = None
if "--no-cpu" not in sys.argv:
    import pyoracle as po, fixtures as fx
    if po.have_ref():
        cpu = po.ref(); cpu._g("set_threads")(os.cpu_count() or 1)
TS = (16, 32, 64)

print(f"## quantised mul_mat (weights from HBM; roofline = weight bytes / {HBM:.0f} GB/s)\n")
print("| N x K | type | T | GPU us | GB/s | % HBM | ggml-cpu us |")
print("|---|---|---|---|---|---|---|")
rng = np.random.default_rng(0)
for (t, name, N, K) in [(eb.Q4_K, "Q4_K", 4096, 4096), (eb.Q4_K, "Q4_K", 11008, 4096), (eb.Q4_K, "Q4_K", 4096, 11008), (eb.Q6_K, "Q6_K", 32000, 4096)]:
    wbytes = eb.row_size(t, K) * N
    W = fx.random_blocks(t, N * K, rng) if cpu else None
    for T in TS:
        ms = eb.bench_mul_mat_kernel(t, N, K, T, n_bufs=8, iters=100)
        c_us = ""
        if cpu:
            X = rng.normal(0, 1, (T, K)).astype(np.float32); Y = np.empty((T, N), np.float32)
            best = 1e9
            for _ in range(3):
                t0 = time.perf_counter(); cpu._g("mul_mat")(t, po._ptr(W), N, K, po._ptr(X), T, po._ptr(Y)); best = min(best, time.perf_counter() - t0)
            c_us = f"{best * 1e6:.0f}"
        gbs = wbytes / ms / 1e6
        print(f"| {N} x {K} | {name} | {T} | {ms * 1e3:.1f} | {gbs:.0f} | {100 * gbs / HBM:.0f} | {c_us} |", flush=True)

print(f"\n## tree-masked flash attention, H = 32, D = 128, f16 KV (roofline = K + V bytes / {HBM:.0f} GB/s)\n")
print("| n_kv | T | GPU us | GB/s | % HBM | ggml-cpu us |")
print("|---|---|---|---|---|---|")
H, D = 32, 128
for n_kv in (1024, 2048, 4096, 8192):
    kvb = 2 * n_kv * H * D * 2
    if cpu:
        k = rng.normal(0, 1, (n_kv, H, D)).astype(np.float16); v = rng.normal(0, 1, (n_kv, H, D)).astype(np.float16)
    for T in TS:
        ms = eb.bench_flash_attn(T, H, D, n_kv, n_bufs=max(2, int(400e6 // kvb)), iters=100)
        c_us = ""
        if cpu:
            q = rng.normal(0, 1, (T, H, D)).astype(np.float32)
            vis = np.zeros((T, n_kv), np.uint8); vis[:, :n_kv - T] = 1
            for i in range(T): vis[i, n_kv - T:n_kv - T + i + 1] = 1
            best = 1e9
            for _ in range(3):
                t0 = time.perf_counter(); cpu.flash_attn(q, k.view(np.uint16), v.view(np.uint16), vis, 1 / np.sqrt(D)); best = min(best, time.perf_counter() - t0)
            c_us = f"{best * 1e6:.0f}"
        gbs = kvb / ms / 1e6
        print(f"| {n_kv} | {T} | {ms * 1e3:.1f} | {gbs:.0f} | {100 * gbs / HBM:.0f} | {c_us} |", flush=True)
if cpu:
    print(f"\nggml-cpu: oracle/_ref (the reference's own ggml-cpu sources), {cpu._g("get_threads")()} threads, best of 3, includes graph build + tensor setup of the harness call.")
