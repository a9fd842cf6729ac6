"""BASELINE config 3 on one GPU: Vicuna-7B Q8_0 + EAGLE head (Q8_0), several independent request streams over ONE resident copy of the
weights (each stream: its own KV caches, tree state and CUDA graph), rounds issued round-robin.  Prints per-stream and aggregate tok/s.
usage: python tools/config3_streams.py [n_streams=8] [rounds=10]"""
import os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import bench
from __graft_entry__ import load_package

n_streams = int(sys.argv[1]) if len(sys.argv) > 1 else 8
rounds = int(sys.argv[2]) if len(sys.argv) > 2 else 10
eb = load_package(); eb.init(0)
HP = bench.HP
tgt = eb.Model(**HP)
tgt.synth_tensor(-1, eb.T_TOK_EMBD, eb.Q8_0, 1); tgt.synth_tensor(-1, eb.T_OUTPUT_NORM, eb.F32, 2); tgt.synth_tensor(-1, eb.T_OUTPUT, eb.Q8_0, 3)
for l in range(HP["L"]):
    for which in range(9):
        tgt.synth_tensor(l, which, eb.F32 if which in (eb.L_ATTN_NORM, eb.L_FFN_NORM) else eb.Q8_0, 1000 + 16 * l + which)
head = eb.Model(**{**HP, "L": 1}, is_eagle=True)
head.synth_tensor(-1, eb.T_TOK_EMBD, eb.Q8_0, 7); head.synth_tensor(-1, eb.T_FC, eb.Q8_0, 8); head.synth_tensor(-1, eb.T_FC_BIAS, eb.F32, 9)
for which in range(1, 9):
    head.synth_tensor(0, which, eb.F32 if which == eb.L_FFN_NORM else eb.Q8_0, 9000 + which)
head.share_lm_head(tgt)
streams = []
for i in range(n_streams):
    s = eb.Session(tgt, head, n_ctx=4096 + 256, fixed_accept=bench.FIXED_ACCEPT, use_graph=True, **bench.TREE)
    s.synth_kv(bench.N_PAST0, seed=100 + i); s.set_root(1 + i)
    streams.append(s)
toks = [1 + i for i in range(n_streams)]
for i, s in enumerate(streams):                       # warm-up: capture each stream's graph
    r = s.round(); toks[i] = r.bonus
for i, s in enumerate(streams):
    s.synth_kv(bench.N_PAST0, seed=100 + i)
gpu_ms = 0.0; n_tok = 0
eb.lib().eb_synchronize(None); t0 = time.perf_counter()
for _ in range(rounds):
    for i, s in enumerate(streams):
        s.set_root(toks[i]); r = s.round(); toks[i] = r.bonus; gpu_ms += r.gpu_ms; n_tok += r.n_accept
wall = time.perf_counter() - t0
print(f"config 3 (1 GPU): {n_streams} streams x {rounds} rounds, Q8_0 target {tgt.weight_bytes() / 1e9:.2f} GB: round {gpu_ms / (rounds * n_streams):.3f} ms, "
      f"{n_tok / (gpu_ms * 1e-3):.0f} tok/s device time, {n_tok / wall:.0f} tok/s wall (rounds of different streams run back to back, no batching across streams)")
v = sorted(streams[0].verify_only(65) for _ in range(5))[2]
print(f"verify step (T = 65, n_kv = 4096): {v:.3f} ms = {100 * bench.algorithmic_bytes(tgt.weight_bytes(), 4096, 65) / (v * 1e-3) / 1e9 / 6462.4:.1f} % of the HBM roofline")
