#!/usr/bin/env python
"""bench.py — EAGLE speculative decoding on B200: Vicuna-7B Q4_K_M target + EAGLE head,
64-node draft tree (+ root = 65-row verify batch), ~4K-token KV context, batch 1, fixed accept
length a = 3 (BASELINE.json configs[1]).  One step = one speculative round (draft-head forwards +
device tree expansion, tree-verify forward of the target, accept walk, KV compaction).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

N > 1: independent request streams, one replica per GPU, no collective on the data path
(SURVEY §8e "replicas only"); launched by torchrun, NCCL only for the barrier / max-over-ranks.
`--impl reference` times the reference's own ggml-cpu implementation (oracle/_ref, compiled from
the reference sources) on the host cores, on a bounded sample of the same workload.
Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

HP = dict(E=4096, H=32, F=11008, V=32000, L=32)          # Vicuna-7B-v1.3 (R/src/llama-model.cpp:410-487)
N_PAST0 = 4096 - 65                                       # KV cells committed before the first round
TREE = dict(max_nodes=65, n_branch=8, n_draft=64, max_depth=10, top_k=10, p_split=0.1)
FIXED_ACCEPT = 3
Q4_K_BPW, Q6_K_BPW = 144 / 256, 210 / 256
ROOFLINE_TRAFFIC_BYTES = 51.65e6   # profiles/r1c_summary.md: ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum per launch, N=22016 K=4096 T=65


def q4_k_m_more_bits(l, n):                               # R/src/llama-quant.cpp:129-131
    return l < n // 8 or l >= 7 * n // 8 or (l - n // 8) % 3 == 2


def peaks():
    try:
        p = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler(threading.Thread):
    def __init__(self, gpu):
        super().__init__(daemon=True)
        self.gpu, self.rows, self.stop_flag = gpu, [], False

    def run(self):
        q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={q}", "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([c.strip() for c in out.split(",")])
            except Exception:
                pass
            time.sleep(0.1)

    def summary(self):
        self.stop_flag = True
        sm = sorted(int(r[0]) for r in self.rows if r[0].isdigit())
        mx = max([int(r[1]) for r in self.rows if r[1].isdigit()] or [0])
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None, "reasons": reasons, "samples": len(self.rows)}


# ------------------------------------------------------------------------------------------ GPU arm
def build_gpu_models(eb):
    tgt = eb.Model(**HP)
    seed = 1
    tgt.synth_tensor(-1, eb.T_TOK_EMBD, eb.Q4_K, seed); tgt.synth_tensor(-1, eb.T_OUTPUT_NORM, eb.F32, seed + 1)
    tgt.synth_tensor(-1, eb.T_OUTPUT, eb.Q6_K, seed + 2)
    for l in range(HP["L"]):
        more = q4_k_m_more_bits(l, HP["L"])
        for which in range(9):
            if which in (eb.L_ATTN_NORM, eb.L_FFN_NORM):
                t = eb.F32
            elif which in (eb.L_WV, eb.L_DOWN) and more:
                t = eb.Q6_K
            else:
                t = eb.Q4_K
            tgt.synth_tensor(l, which, t, 1000 + 16 * l + which)
    head = eb.Model(**{**HP, "L": 1}, is_eagle=True)
    head.synth_tensor(-1, eb.T_TOK_EMBD, eb.Q4_K, 7); head.synth_tensor(-1, eb.T_FC, eb.Q4_K, 8); head.synth_tensor(-1, eb.T_FC_BIAS, eb.F32, 9)
    for which in range(1, 9):
        t = eb.F32 if which == eb.L_FFN_NORM else (eb.Q6_K if which in (eb.L_WV, eb.L_DOWN) else eb.Q4_K)
        head.synth_tensor(0, which, t, 9000 + which)
    head.share_lm_head(tgt)
    return tgt, head


def algorithmic_bytes(tgt_bytes, n_kv, T):
    kv_read = 2 * n_kv * HP["E"] * 2 * HP["L"]
    kv_write = 2 * T * HP["E"] * 2 * HP["L"]
    return tgt_bytes + kv_read + kv_write


def max_over_ranks(dist, times_ms, device):
    """Replicas share nothing on the data path; the only exchange is this MAX of the per-rank
    timings (the job is as slow as its slowest replica).  dist=None -> single process."""
    if dist is None:
        return list(times_ms)
    import torch
    tt = torch.tensor(list(times_ms), device=device, dtype=torch.float64)
    dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    dist.barrier()
    return tt.tolist()


def job_tokens(steps, world):
    """whole-job tokens: every replica commits fixed_accept drafted tokens + the bonus token per round"""
    return (FIXED_ACCEPT + 1) * steps * world


def replica_seed(rank):
    """each replica fills its KV prefix from its own seed (independent sequences, weak scaling)"""
    return 3 + rank


def run_gpu(args, rank, world, local_rank):
    from __graft_entry__ import load_package
    eb = load_package()
    eb.init(local_rank)
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    tgt, head = build_gpu_models(eb)
    s = eb.Session(tgt, head, n_ctx=4096 + 256, fixed_accept=FIXED_ACCEPT, use_graph=True, **TREE)
    s.synth_kv(N_PAST0, seed=replica_seed(rank))
    s.set_root(1)
    tok = 1
    res = None
    for _ in range(max(args.warmup, 3)):
        s.set_root(tok); res = s.round(); tok = res.bonus
    # rewind the KV length so that the timed rounds sit around the 4K mark
    s.synth_kv(N_PAST0, seed=replica_seed(rank)); s.set_root(tok)

    def barrier():
        if dist is not None:
            dist.barrier()
        eb.lib().eb_synchronize(None)

    sampler = ClockSampler(local_rank); sampler.start()
    eb.launch_count(reset=True)
    barrier()
    t0 = time.perf_counter()
    gpu_ms = 0.0
    nodes = []
    for _ in range(args.steps):
        s.set_root(tok)                      # H2D: the step's input (last accepted token id), 4 bytes
        res = s.round()                      # D2H: eb_round_result (accepted ids, tree) — the step's result
        tok = res.bonus; gpu_ms += res.gpu_ms; nodes.append(res.n_nodes)
    eb.lib().eb_synchronize(None)
    t1 = time.perf_counter()
    launches = eb.launch_count()
    wall_ms = (t1 - t0) * 1e3
    wall_ms, gpu_ms = max_over_ranks(dist, [wall_ms, gpu_ms], "cuda")
    clocks = sampler.summary()
    tokens = job_tokens(args.steps, world)
    out = None
    if rank == 0:
        hbm, peak_src = peaks()
        # verify step alone (device time, CUDA events inside the library)
        ver = sorted(s.verify_only(65) for _ in range(12))[:-2]
        verify_us = 1e3 * ver[len(ver) // 2]
        step_bytes = algorithmic_bytes(tgt.weight_bytes(), 4096, 65)
        # dominant kernel: the Q4_K tree-batched GEMM at the fused ffn_gate|ffn_up shape (2 x 11008 rows x 4096, T = 65):
        # the launch that moves the most bytes per step (32 per step, 50.7 MB each)
        kN, kK, kT = 2 * HP["F"], HP["E"], 65
        k_ms = eb.bench_mul_mat_kernel(eb.Q4_K, kN, kK, kT, n_bufs=8, iters=200)
        k_bytes = kN * kK * Q4_K_BPW
        achieved = k_bytes / (k_ms * 1e-3) / 1e9
        mv_ms = eb.bench_mul_mat_kernel(eb.Q4_K, kN, kK, 1, n_bufs=8, iters=200)
        traffic = ROOFLINE_TRAFFIC_BYTES      # dram bytes per launch of the same kernel/shape from the committed ncu capture (profiles/)
        out = {
            "metric": "eagle_decode_tok_per_s", "value": tokens / (gpu_ms * 1e-3), "unit": "tok/s",
            "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": gpu_ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int8xq4_K/q6_K dot, f32 accumulate, f16 KV",
            "data": "synthetic (random-init Q4_K_M blocks, N(0,1) f16 KV prefill, fixed accept length 3)",
            "config": {"workload": "vicuna-7b Q4_K_M + EAGLE head, 64-node tree (65-row verify batch), KV ~4096, batch 1, fixed_accept=3",
                       "tree": TREE, "n_kv_at_verify": 4096, "tree_nodes_median": int(np.median(nodes)),
                       "l2": "weights (4.0 GB) + KV (2.1 GB) per step exceed L2 (126 MB): no flush needed",
                       "parallelism": f"replicas x{world}, no collective"},
            "e2e": {"value": tokens / (wall_ms * 1e-3), "unit": "tok/s", "h2d_bytes_per_step": 4, "d2h_bytes_per_step": 1552},
            "gpu_launches": launches,
            "verify_step_us": verify_us,
            "verify_step_hbm_frac": step_bytes / (verify_us * 1e-6) / 1e9 / hbm,
            "verify_step_bytes": step_bytes,
            "roofline": {"kernel": "mmq_tc_kernel<0> (Q4_K x Q8_K tcgen05 kind::f16 exact-integer GEMM, A in TMEM, TMA weight ring; ffn_gate|ffn_up fused: N=22016 K=4096 T=65)", "bound": "hbm", "achieved": achieved, "peak": hbm,
                         "unit": "GB/s", "frac": achieved / hbm, "traffic": traffic, "peak_source": peak_src, "ms_per_launch": k_ms,
                         "matvec_T1_gbs": k_bytes / (mv_ms * 1e-3) / 1e9},
            "clocks": clocks,
        }
        if world == 1 and not args.no_cpu_baseline:
            try:
                out["cpu_baseline"] = cpu_baseline(budget_s=25.0)
            except Exception as e:                     # the baseline is a reported number, never the product path
                out["cpu_baseline"] = {"error": str(e)}
    s.free()
    if dist is not None:
        dist.destroy_process_group()
    return out


# ------------------------------------------------------------------------------------------ CPU arm
def _rand_blocks(t, n_elems, rng):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import fixtures as fx
    return fx.random_blocks(t, n_elems, rng)


class CpuSlice:
    """The reference's ggml-cpu path (or the oracle port) on a slice of the 7B workload."""

    def __init__(self):
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import pyoracle as po
        self.po = po
        if po.have_ref():
            self.path = po.ref()
            flags = open("/proc/cpuinfo").read()
            self.cores = len(os.sched_getaffinity(0))
            self.path._g("set_threads")(self.cores)
            assert "avx2" in flags and "fma" in flags, "oracle/_ref needs AVX2+FMA"
        else:
            self.path = po.oracle()
            self.cores = len(os.sched_getaffinity(0))
        self.kind = self.path.kind
        rng = np.random.default_rng(0)
        E, F, V = HP["E"], HP["F"], HP["V"]
        self.keep = []

        def blocks(t, n):
            a = _rand_blocks(t, n, rng); self.keep.append(a); return a

        self.emb = blocks(po.Q4_K, V * E); self.out = blocks(po.Q6_K, V * E)
        self.norm = np.ones(E, np.float32)
        self.layer = {po.L_ATTN_NORM: (po.F32, self.norm), po.L_FFN_NORM: (po.F32, self.norm),
                      po.L_WQ: (po.Q4_K, blocks(po.Q4_K, E * E)), po.L_WK: (po.Q4_K, blocks(po.Q4_K, E * E)),
                      po.L_WV: (po.Q6_K, blocks(po.Q6_K, E * E)), po.L_WO: (po.Q4_K, blocks(po.Q4_K, E * E)),
                      po.L_GATE: (po.Q4_K, blocks(po.Q4_K, F * E)), po.L_UP: (po.Q4_K, blocks(po.Q4_K, F * E)),
                      po.L_DOWN: (po.Q4_K, blocks(po.Q4_K, E * F))}
        self.fc = blocks(po.Q4_K, 2 * E * E); self.fc_b = np.zeros(E, np.float32)
        self.n_ctx = 4096 + 64
        self.models = {}

    def model(self, L, eagle):
        key = (L, eagle)
        if key in self.models:
            return self.models[key]
        po, p = self.po, self.path
        m = p.model_new(HP["E"], HP["H"], HP["F"], HP["V"], L, 1e-5, 10000.0, eagle)
        p._g("model_set_global")(m, po.T_TOK_EMBD, po.Q4_K, po._ptr(self.emb))
        p._g("model_set_global")(m, po.T_OUTPUT, po.Q6_K, po._ptr(self.out))
        if eagle:
            p._g("model_set_global")(m, po.T_FC, po.Q4_K, po._ptr(self.fc)); p._g("model_set_global")(m, po.T_FC_BIAS, po.F32, po._ptr(self.fc_b))
        else:
            p._g("model_set_global")(m, po.T_OUTPUT_NORM, po.F32, po._ptr(self.norm))
        for l in range(L):
            for which, (t, raw) in self.layer.items():
                if eagle and which == po.L_ATTN_NORM:
                    continue
                p._g("model_set_layer")(m, l, which, t, po._ptr(raw))
        kv = p._g("kv_new")(max(L, 1), HP["E"], self.n_ctx)
        ks, vs = p.kv_arrays(kv, max(L, 1), HP["E"], self.n_ctx)
        rng = np.random.default_rng(1)
        for a in ks + vs:
            a[:N_PAST0] = rng.normal(0, 1, (N_PAST0, HP["E"])).astype(np.float16).view(np.uint16)
        self.models[key] = (m, kv)
        return self.models[key]

    def forward_s(self, L, eagle, T):
        m, kv = self.model(L, eagle)
        tokens = (np.arange(T) * 37 % 1000).astype(np.int32)
        pos = (N_PAST0 + np.arange(T)).astype(np.int32)
        vis = np.zeros((T, self.n_ctx), np.uint8); vis[:, :N_PAST0] = 1
        vis[:, N_PAST0:N_PAST0 + T] = np.tril(np.ones((T, T), np.uint8))
        hid = np.zeros((T, HP["E"]), np.float32) if eagle else None
        t0 = time.perf_counter()
        self.path.forward(m, kv, tokens, pos, hid, N_PAST0, self.n_ctx, vis, HP["E"], HP["V"])
        return time.perf_counter() - t0

    def round_seconds(self, t_layer_plus, t_overhead, t_head):
        per_layer = max(t_layer_plus - t_overhead, 1e-9)
        verify = HP["L"] * per_layer + t_overhead
        return verify + (TREE["max_depth"] + 1) * t_head, verify


def cpu_baseline(budget_s=12.0):
    """reference CPU path on the host cores: a bounded sample (about budget_s seconds of CPU work) of the same workload"""
    c = CpuSlice()
    c.forward_s(1, False, 65)                                # warm the thread pool and the page cache
    t_first = c.forward_s(1, False, 65)
    reps = int(min(64, max(3, budget_s * 0.8 / max(t_first, 1e-3))))
    t_l1 = float(np.median([c.forward_s(1, False, 65) for _ in range(reps)]))      # one decoder layer at n_kv = 4096 (+ lm_head slice)
    t_over = float(np.median([c.forward_s(0, False, 65) for _ in range(5)]))       # embeddings + final norm + lm_head
    t_head = float(np.median([c.forward_s(1, True, 8) for _ in range(5)]))         # one draft level: head forward over 8 branches
    rnd, ver = c.round_seconds(t_l1, t_over, t_head)
    return {"value": (FIXED_ACCEPT + 1) / rnd, "unit": "tok/s", "cores": c.cores, "kind": c.kind,
            "verify_step_us": ver * 1e6,
            "sample": f"1 of 32 target layers (T=65, n_kv=4096) timed {reps}x + lm_head slice 5x + one 8-row draft-head level 5x "
                      f"(medians {t_l1:.2f}s, {t_over:.2f}s, {t_head:.2f}s; {reps * t_l1 + 5 * (t_over + t_head):.0f}s of CPU work), "
                      f"scaled to 32 layers + {TREE['max_depth'] + 1} draft levels"}


def run_reference(args, rank, world):
    if rank != 0:
        return None
    c = CpuSlice()
    t_over = c.forward_s(0, False, 65)
    t_head = c.forward_s(1, True, 8)
    for _ in range(min(args.warmup, 2)):
        c.forward_s(1, False, 65)
    ts = [c.forward_s(1, False, 65) for _ in range(args.steps)]
    t_l1 = float(np.median(ts))
    rnd, ver = c.round_seconds(t_l1, t_over, t_head)
    val = (FIXED_ACCEPT + 1) / rnd
    sample = (f"each step = 1 of 32 target layers + lm_head slice at T=65, n_kv=4096 on {c.cores} host threads "
              f"(median {t_l1:.2f}s; lm_head slice {t_over:.2f}s, draft level {t_head:.2f}s), scaled to the full round")
    return {"impl": "reference", "metric": "eagle_decode_tok_per_s", "value": val, "unit": "tok/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": rnd * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "int8xq4_K/q6_K dot, f32 accumulate, f16 KV", "data": "synthetic (random-init Q4_K_M blocks)",
            "config": {"workload": "vicuna-7b Q4_K_M + EAGLE head, 64-node tree (65-row verify batch), KV ~4096, batch 1, fixed_accept=3",
                       "tree": TREE, "parallelism": "host cores only"},
            "verify_step_us": ver * 1e6,
            "cpu_baseline": {"value": val, "unit": "tok/s", "cores": c.cores, "kind": c.kind, "sample": sample},
            "e2e": {"value": val, "unit": "tok/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        out = run_reference(args, rank, world)
    else:
        out = run_gpu(args, rank, world, local)
    if rank == 0 and out is not None:
        print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
