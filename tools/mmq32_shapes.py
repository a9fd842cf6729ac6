"""Q4_0 / Q8_0 tree-batched GEMM (mmq32.cu) on the 7B layer shapes: us per launch (graph replay over 8 weight buffers) and GB/s of weight bytes.
usage: python tools/mmq32_shapes.py [T ...]"""
import os, sys; sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from __graft_entry__ import load_package
eb = load_package(); eb.init(0)
SHAPES = ((12288, 4096), (4096, 4096), (22016, 4096), (4096, 11008), (32000, 4096))
print("| type | T | " + " | ".join(f"{n} x {k}" for n, k in SHAPES) + " |")
for t, name in ((eb.Q8_0, "Q8_0"), (eb.Q4_0, "Q4_0")):
    for T in [int(x) for x in sys.argv[1:]] or [16, 65]:
        cells = []
        for (N, K) in SHAPES:
            ms = eb.bench_mul_mat_kernel(t, N, K, T, n_bufs=8, iters=40)
            cells.append("%.1f us (%.0f GB/s)" % (ms * 1e3, eb.row_size(t, K) * N / ms / 1e6))
        print(f"| {name} | {T} | " + " | ".join(cells) + " |")
