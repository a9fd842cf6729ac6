import sys, time, subprocess
sys.path.insert(0, ".")
import bench
from __graft_entry__ import load_package
eb = load_package(); eb.init(0)
def smi():
    return subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,power.draw,temperature.gpu,clocks_event_reasons.sw_power_cap,clocks_event_reasons.active", "--format=csv,noheader"], capture_output=True, text=True).stdout.strip()
print("idle", smi())
tgt, head = bench.build_gpu_models(eb)
s = eb.Session(tgt, head, n_ctx=4096 + 256, fixed_accept=3, use_graph=1, **bench.TREE)
s.synth_kv(bench.N_PAST0, seed=3); s.set_root(1)
tok = 1
for i in range(40):
    s.set_root(tok); r = s.round(); tok = r.bonus
    if i % 5 == 0: print(i, "round ms", round(r.gpu_ms, 3), smi())
print("verify", [round(s.verify_only(65), 3) for _ in range(5)], smi())
for it in (20, 200):
    print("kernel iters", it, eb.bench_mul_mat_kernel(eb.Q4_K, 22016, 4096, 65, n_bufs=8, iters=it) * 1e3, "us", smi())
