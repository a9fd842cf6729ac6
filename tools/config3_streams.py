"""BASELINE config 3 on one GPU: Vicuna-7B Q8_0 + EAGLE head (Q8_0), several independent request streams over ONE resident copy of the
weights (each stream: its own KV caches, tree state and CUDA graph), rounds issued round-robin.  Prints per-stream and aggregate tok/s.
usage: python tools/config3_streams.py [n_streams=8] [rounds=10] [type=Q8_0|Q4_0] [n_draft=64] [n_past=4096]
BASELINE config 1 (Q4_0 target, 26-node tree, 512-token prompt, one stream): python tools/config3_streams.py 1 20 Q4_0 26 512"""
import os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import bench
from __graft_entry__ import load_package

def build(eb, QT):
    HP = bench.HP
    tgt = eb.Model(**HP)
    tgt.synth_tensor(-1, eb.T_TOK_EMBD, QT, 1); tgt.synth_tensor(-1, eb.T_OUTPUT_NORM, eb.F32, 2); tgt.synth_tensor(-1, eb.T_OUTPUT, QT, 3)
    for l in range(HP["L"]):
        for which in range(9):
            tgt.synth_tensor(l, which, eb.F32 if which in (eb.L_ATTN_NORM, eb.L_FFN_NORM) else QT, 1000 + 16 * l + which)
    head = eb.Model(**{**HP, "L": 1}, is_eagle=True)
    head.synth_tensor(-1, eb.T_TOK_EMBD, QT, 7); head.synth_tensor(-1, eb.T_FC, QT, 8); head.synth_tensor(-1, eb.T_FC_BIAS, eb.F32, 9)
    for which in range(1, 9):
        head.synth_tensor(0, which, eb.F32 if which == eb.L_FFN_NORM else QT, 9000 + which)
    head.share_lm_head(tgt)
    return tgt, head


def run(eb, n_streams=8, rounds=10, QT=None, n_draft=None, n_past=None, warm=1, barrier=None, seed0=100):
    """-> dict: device ms per round (mean over streams and rounds), tokens, wall seconds of the timed region, verify-step ms"""
    QT = eb.Q8_0 if QT is None else QT
    n_draft = bench.TREE["n_draft"] if n_draft is None else n_draft
    n_past = bench.N_PAST0 if n_past is None else n_past
    TREE = {**bench.TREE, "n_draft": n_draft, "max_nodes": n_draft + 1}
    tgt, head = build(eb, QT)
    streams = []
    for i in range(n_streams):
        s = eb.Session(tgt, head, n_ctx=4096 + 256, fixed_accept=bench.FIXED_ACCEPT, use_graph=True, **TREE)
        s.synth_kv(n_past, seed=seed0 + i); s.set_root(1 + i)
        streams.append(s)
    toks = [1 + i for i in range(n_streams)]
    for _ in range(max(1, warm)):                         # warm-up: capture each stream's graph
        for i, s in enumerate(streams):
            s.synth_kv(n_past, seed=seed0 + i); s.set_root(toks[i]); r = s.round(); toks[i] = r.bonus
    for i, s in enumerate(streams):
        s.synth_kv(n_past, seed=seed0 + i)
    n0 = eb.launch_count()
    gpu_ms = 0.0; n_tok = 0
    eb.lib().eb_synchronize(None)
    if barrier:
        barrier()
    t0 = time.perf_counter()
    for _ in range(rounds):
        for i, s in enumerate(streams):
            s.set_root(toks[i]); r = s.round(); toks[i] = r.bonus; gpu_ms += r.gpu_ms; n_tok += r.n_accept
    wall = time.perf_counter() - t0
    launches = eb.launch_count() - n0
    T = n_draft + 1
    v = sorted(streams[0].verify_only(T) for _ in range(5))[2]
    res = {"n_streams": n_streams, "rounds": rounds, "round_ms": gpu_ms / (rounds * n_streams), "tokens": n_tok, "gpu_ms": gpu_ms, "wall_s": wall,
           "verify_ms": v, "weight_bytes": tgt.weight_bytes(), "launches": launches, "T": T, "n_past": n_past}
    for s in streams:
        s.free()
    return res


if __name__ == "__main__":
    n_streams = int(sys.argv[1]) if len(sys.argv) > 1 else 8
    rounds = int(sys.argv[2]) if len(sys.argv) > 2 else 10
    eb = load_package(); eb.init(0)
    QN = sys.argv[3] if len(sys.argv) > 3 else "Q8_0"
    n_draft = int(sys.argv[4]) if len(sys.argv) > 4 else bench.TREE["n_draft"]
    n_past = int(sys.argv[5]) if len(sys.argv) > 5 else bench.N_PAST0
    r = run(eb, n_streams, rounds, getattr(eb, QN), n_draft, n_past)
    print(f"{n_streams} streams x {rounds} rounds, {n_draft}-node tree, n_past {n_past}, {QN} target {r['weight_bytes'] / 1e9:.2f} GB: round {r['round_ms']:.3f} ms, "
          f"{r['tokens'] / (r['gpu_ms'] * 1e-3):.0f} tok/s device time, {r['tokens'] / r['wall_s']:.0f} tok/s wall (rounds of different streams run back to back, no batching across streams)")
    print(f"verify step (T = {r['T']}, n_kv = {n_past}): {r['verify_ms']:.3f} ms = "
          f"{100 * bench.algorithmic_bytes(r['weight_bytes'], n_past, r['T']) / (r['verify_ms'] * 1e-3) / 1e9 / 6462.4:.1f} % of the HBM roofline")
