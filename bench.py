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


def roofline_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernel at the bench shape, from the committed
    `ncu --set full` capture of the SHIPPED kernel (profiles/r2_ncu_traffic.json, written by tools/ncu_extract.py); null without it"""
    try:
        d = json.load(open(os.path.join(ROOT, "profiles", "r2_ncu_traffic.json")))
        return float(d["dram_bytes_per_launch"])
    except Exception:
        return None



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
    # room for every timed round (4 committed tokens each) after the 4K prefix, + one tree
    n_ctx = (N_PAST0 + (FIXED_ACCEPT + 1) * (args.steps + 1) + TREE["max_nodes"] + TREE["n_branch"] + 255) // 256 * 256
    s = eb.Session(tgt, head, n_ctx=max(n_ctx, 4096 + 256), fixed_accept=FIXED_ACCEPT, use_graph=True, **TREE)
    s.synth_kv(N_PAST0, seed=replica_seed(rank))
    s.set_root(1)
    tok = 1
    res = None
    for w in range(max(args.warmup, 3)):
        if w and w % 16 == 0:
            s.synth_kv(N_PAST0, seed=replica_seed(rank))     # long warm-ups stay inside the context
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
    committed = 0
    for _ in range(args.steps):
        s.set_root(tok)                      # H2D: the step's input (last accepted token id), 4 bytes
        res = s.round()                      # D2H: eb_round_result (accepted ids, tree) — the step's result
        tok = res.bonus; gpu_ms += res.gpu_ms; nodes.append(res.n_nodes)
        committed += res.n_accept            # accepted nodes incl. the root = tokens this round committed (fixed_accept drafted + bonus)
    assert committed == (FIXED_ACCEPT + 1) * args.steps, f"rounds committed {committed} tokens, expected {(FIXED_ACCEPT + 1) * args.steps}: the synthetic tree is shallower than fixed_accept"
    eb.lib().eb_synchronize(None)
    t1 = time.perf_counter()
    launches = eb.launch_count()
    wall_ms = (t1 - t0) * 1e3
    wall_ms, gpu_ms = max_over_ranks(dist, [wall_ms, gpu_ms], "cuda")
    clocks = sampler.summary()
    tokens = committed * world               # every replica runs the same fixed-acceptance schedule (asserted above on each rank)
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
        traffic = roofline_traffic()
        out = {
            "metric": "eagle_decode_tok_per_s", "value": tokens / (gpu_ms * 1e-3), "unit": "tok/s",
            "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": gpu_ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int8xq4_K/q6_K dot, f32 accumulate, f16 KV",
            "data": "synthetic (random-init Q4_K_M blocks, N(0,1) f16 KV prefill, fixed accept length 3)",
            "config": config_dict(f"replicas x{world}, no collective", int(np.median(nodes)),
                                  "weights (4.0 GB) + KV (2.1 GB) per step exceed L2 (126 MB): no flush needed"),
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
                out["cpu_baseline"] = cpu_baseline(n_rounds=3, with_plugin=True)
                # the same verify step through the reference's own llama_decode -> ggml_backend_sched -> ggml_backend_graph_compute of
                # libggml-b200.so (unfused ggml ops, host-side graph build and launches included), next to the engine's fused step
                out["plugin_verify_step_us"] = out["cpu_baseline"].pop("plugin_verify_step_us", None)
            except Exception as e:                     # the baseline is a reported number, never the product path
                out["cpu_baseline"] = {"error": str(e)}
    s.free()
    if dist is not None:
        dist.destroy_process_group()
    return out


# ------------------------------------------------------------------------------------------ CPU arm
WORKLOAD = "vicuna-7b Q4_K_M + EAGLE head, 64-node tree (65-row verify batch), KV ~4096, batch 1, fixed_accept=3"


def config_dict(parallelism, nodes_median, l2):
    """the same key set in both arms (the driver compares them)"""
    return {"workload": WORKLOAD, "tree": TREE, "n_kv_at_verify": 4096, "tree_nodes_median": nodes_median, "l2": l2, "parallelism": parallelism}


def native_cpu_ok():
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import refround as rr
    return rr.native_ok()


class CpuRound:
    """The reference's own CPU implementation of one round, measured, not extrapolated: the tree-verify forward is ONE llama_decode
    (R/src/llama.cpp:11916) of the reference's libllama over the 65-token tree batch on a 32-layer Vicuna-7B-shaped Q4_K_M GGUF with
    4031 committed KV cells; the draft side is the reference's build_eagle graph (oracle/ref_harness.cpp, same ggml-cpu) run for
    every one of the max_depth + 1 levels.  ggml build: -march=native + llamafile + repack (the reference's default cmake
    configuration) when this host's ISA allows it, else the deterministic AVX2 build — named in `kind_detail`."""

    def __init__(self, with_plugin=False):
        import tempfile
        sys.path.insert(0, os.path.join(ROOT, "oracle")); sys.path.insert(0, os.path.join(ROOT, "tests"))
        import fixtures as fx
        import fullsize_run as fr
        import gguf_fixture as gf
        import pyoracle as po
        import refround as rr
        if not rr.have():
            raise RuntimeError("oracle/_ref (the reference build) is missing")
        self.variant = "native" if native_cpu_ok() else "ref"
        self.cores = len(os.sched_getaffinity(0))
        rr.lib(variant=self.variant)                       # the ggml build of this process, loaded before anything else links ggml
        self.po, self.fx = po, fx
        t0 = time.perf_counter()
        tgt = fr.random_model(HP, "q4_k_m", 7)
        tmp = tempfile.mkdtemp(prefix="eb_bench_")
        path = gf.write_gguf(tgt, os.path.join(tmp, "vicuna7b_q4_k_m_random.gguf"), n_ctx_train=8192)
        self.m = rr.Model(path, variant=self.variant)
        rng = np.random.default_rng(11)
        self.par = fx.driver_like_tree(65, TREE["n_branch"], rng); self.depth = fx.tree_depths(self.par)
        self.seqs, n_seq = fx.tree_seq_ids(self.par)
        self.c = rr.Context(self.m, 4096 + 256, n_batch=512, n_seq_max=max(8, n_seq), n_threads=self.cores)
        for l in range(HP["L"]):
            for which in (0, 1):
                self.c.kv_write(l, which, 0, rng.standard_normal((N_PAST0, HP["E"]), np.float32).astype(np.float16))
        self.n_seq = n_seq
        self.tokens = rng.integers(3, HP["V"], 65).astype(np.int32)
        # the same GGUF through the reference's loader with every layer on libggml-b200.so (the CPU context above was created before
        # the plug-in was registered, so its scheduler holds the CPU backend only)
        self.plugin_us = None; self.plugin_err = None
        if with_plugin:
            try:
                plug = os.path.join(ROOT, "eagle-in-llama.cpp_b200", "ggml_backend", "libggml-b200.so")
                if rr.lib().rr_load_backend(plug.encode()) < 2:
                    raise RuntimeError("ggml_backend_load(libggml-b200.so) failed")
                mp = rr.Model(path, n_gpu_layers=99, variant=self.variant)
                cp = rr.Context(mp, 4096 + 256, n_batch=512, n_seq_max=max(8, n_seq), n_threads=self.cores)
                for l in range(HP["L"]):
                    for which in (0, 1):
                        cp.kv_write(l, which, 0, rng.standard_normal((N_PAST0, HP["E"]), np.float32).astype(np.float16))
                ts = []
                for _ in range(8):
                    cp.kv_set_prefix(N_PAST0, n_seq)
                    ts.append(cp.time_decode(self.tokens, N_PAST0 + self.depth, self.seqs))
                self.plugin_us = float(np.median(ts[3:]))
                cp.free(); mp.free()
            except Exception as e:                           # reported, never fatal for the bench line
                self.plugin_err = str(e)
        os.remove(path); os.rmdir(tmp)
        # draft head: build_eagle on the same ggml-cpu (lm_head = the target's output matrix)
        self.ref = po.ref(); self.ref._g("set_threads")(self.cores)
        E, F, V = HP["E"], HP["F"], HP["V"]
        self.keep = [tgt]
        hm = self.ref.model_new(E, HP["H"], F, V, 1, 1e-5, 10000.0, True)
        hr = np.random.default_rng(12)
        def blocks(t, n):
            a = fx.random_blocks(t, n, hr); self.keep.append(a); return a
        fcb = np.zeros(E, np.float32); ones = np.ones(E, np.float32); self.keep += [fcb, ones]
        g = self.ref._g
        g("model_set_global")(hm, po.T_TOK_EMBD, po.Q4_K, po._ptr(tgt["globals"][po.T_TOK_EMBD][1]))
        g("model_set_global")(hm, po.T_OUTPUT, po.Q6_K, po._ptr(tgt["globals"][po.T_OUTPUT][1]))
        g("model_set_global")(hm, po.T_FC, po.Q4_K, po._ptr(blocks(po.Q4_K, 2 * E * E))); g("model_set_global")(hm, po.T_FC_BIAS, po.F32, po._ptr(fcb))
        for which, (t, n) in {po.L_WQ: (po.Q4_K, E * E), po.L_WK: (po.Q4_K, E * E), po.L_WV: (po.Q6_K, E * E), po.L_WO: (po.Q4_K, E * E),
                              po.L_GATE: (po.Q4_K, F * E), po.L_UP: (po.Q4_K, F * E), po.L_DOWN: (po.Q6_K, E * F)}.items():
            g("model_set_layer")(hm, 0, which, t, po._ptr(blocks(t, n)))
        g("model_set_layer")(hm, 0, po.L_FFN_NORM, po.F32, po._ptr(ones))
        self.hm = hm
        self.h_ctx = 4096 + 64
        self.hkv = g("kv_new")(1, E, self.h_ctx)
        ks, vs = self.ref.kv_arrays(self.hkv, 1, E, self.h_ctx)
        for a in ks + vs:
            a[:N_PAST0] = hr.standard_normal((N_PAST0, E), np.float32).astype(np.float16).view(np.uint16)
        self.setup_s = time.perf_counter() - t0

    def verify_s(self):
        """one llama_decode of the 65-token tree batch at n_kv = 4096 (the tree's cells are released again afterwards)"""
        self.c.kv_set_prefix(N_PAST0, self.n_seq)
        us = self.c.time_decode(self.tokens, N_PAST0 + self.depth, self.seqs)
        return us * 1e-6

    def head_level_s(self):
        """one draft level: the head over n_branch rows + (host) top-k is negligible next to it"""
        T = TREE["n_branch"]
        tokens = (np.arange(T) * 37 % 1000).astype(np.int32); pos = np.full(T, N_PAST0, np.int32)
        vis = np.zeros((T, self.h_ctx), np.uint8); vis[:, :N_PAST0] = 1; vis[np.arange(T), N_PAST0 + np.arange(T)] = 1
        hid = np.zeros((T, HP["E"]), np.float32)
        t0 = time.perf_counter()
        self.ref.forward(self.hm, self.hkv, tokens, pos, hid, N_PAST0, self.h_ctx, vis, HP["E"], HP["V"])
        return time.perf_counter() - t0

    def round_s(self):
        v = self.verify_s()
        h = sum(self.head_level_s() for _ in range(TREE["max_depth"] + 1))
        return v + h, v

    def describe(self):
        return {"cores": self.cores, "kind": "reference",
                "kind_detail": ("reference libllama + ggml-cpu, -march=native + llamafile sgemm + runtime repack" if self.variant == "native"
                                else "reference libllama + ggml-cpu, deterministic AVX2/FMA build (this host's ISA does not cover the native build's)")}


def cpu_baseline(n_rounds=3, with_plugin=False):
    """reference CPU path on the host cores: n_rounds whole rounds (each ~3 s of a 16-thread host), nothing scaled"""
    c = CpuRound(with_plugin=with_plugin)
    c.round_s()                                              # warm the thread pool and the page cache
    rs = [c.round_s() for _ in range(n_rounds)]
    rnd = float(np.median([r[0] for r in rs])); ver = float(np.median([r[1] for r in rs]))
    d = c.describe()
    d.update({"value": (FIXED_ACCEPT + 1) / rnd, "unit": "tok/s", "verify_step_us": ver * 1e6,
              "sample": f"{n_rounds} whole rounds after 1 warm-up: llama_decode of the 65-token tree batch over all 32 layers at n_kv 4096 "
                        f"(median {ver:.2f} s) + {TREE['max_depth'] + 1} draft-head levels of {TREE['n_branch']} rows (median round {rnd:.2f} s); "
                        f"model + KV set-up {c.setup_s:.0f} s outside the timing"})
    if with_plugin:
        d["plugin_verify_step_us"] = c.plugin_us if c.plugin_us is not None else {"error": c.plugin_err}
    return d


def run_reference(args, rank, world):
    if rank != 0:
        return None
    c = CpuRound()
    for _ in range(max(1, min(args.warmup, 2))):
        c.round_s()
    t0 = time.perf_counter()
    rs = [c.round_s() for _ in range(args.steps)]
    wall = time.perf_counter() - t0
    rnd = float(np.mean([r[0] for r in rs])); ver = float(np.median([r[1] for r in rs]))
    val = (FIXED_ACCEPT + 1) / rnd
    d = c.describe()
    d.update({"value": val, "unit": "tok/s",
              "sample": f"each step = one whole round on {c.cores} host threads: llama_decode of the 65-token tree batch over all 32 layers at n_kv 4096 "
                        f"(median {ver:.2f} s) + {TREE['max_depth'] + 1} draft-head levels; {args.steps} steps in {wall:.0f} s, nothing scaled"})
    return {"impl": "reference", "metric": "eagle_decode_tok_per_s", "value": val, "unit": "tok/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": rnd * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "int8xq4_K/q6_K dot, f32 accumulate, f16 KV", "data": "synthetic (random-init Q4_K_M blocks, N(0,1) f16 KV prefill, fixed accept length 3)",
            "config": config_dict("host cores only", 65, "n/a (host arm)"),
            "verify_step_us": ver * 1e6, "cpu_baseline": d,
            "e2e": {"value": val, "unit": "tok/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}


def run_config3(args, rank, world, local_rank):
    """BASELINE config 3: Vicuna-7B Q8_0 target + Q8_0 EAGLE head, 64 concurrent request streams on 8 GPUs = 8 streams per GPU over ONE
    resident copy of the weights (each stream: its own KV caches, tree state and CUDA graph).  A step = one round of every stream of the
    rank, issued round-robin; the streams of a GPU are NOT batched into one weight pass (that needs a T >= 512 Q8_0 tensor-core GEMM,
    DESIGN.md section 7), so the aggregate equals the single-stream rate per GPU.  Weak scaling: 8 streams per GPU whatever N."""
    if args.impl == "reference":
        return {"impl": "reference", "unavailable": "the reference arm covers the metric's own config (2)"} if rank == 0 else None
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import config3_streams as c3
    from __graft_entry__ import load_package
    eb = load_package(); eb.init(local_rank)
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    c1 = args.config == 1                                   # config 1: the reference's CPU-runnable case — Q4_0 target + Q4_0 head, 26-node tree, 512-token prompt, one stream
    per_gpu = 1 if c1 else 8
    qt, qn = (eb.Q4_0, "Q4_0") if c1 else (eb.Q8_0, "Q8_0")
    sampler = ClockSampler(local_rank); sampler.start()
    r = c3.run(eb, per_gpu, args.steps, qt, n_draft=26 if c1 else None, n_past=512 if c1 else None, warm=max(3, args.warmup),
               barrier=(dist.barrier if dist else None), seed0=100 + 1000 * rank)
    clocks = sampler.summary()
    ms = max_over_ranks(dist, [r["gpu_ms"]], "cuda")[0] if dist else r["gpu_ms"]
    tokens = r["tokens"] * world
    if dist:
        dist.destroy_process_group()
    if rank != 0:
        return None
    return {"metric": "eagle_decode_tok_per_s", "value": tokens / (ms * 1e-3), "unit": "tok/s", "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup),
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": f"int8x{qn.lower()} dot, f32 accumulate, f16 KV", "data": f"synthetic (random-init {qn} blocks, N(0,1) f16 KV prefill, fixed accept length 3)",
            "config": {"workload": (f"config 1: vicuna-7b Q4_0 + Q4_0 EAGLE head, 26-node tree, 512-token prompt, {world} stream(s) (one per GPU)" if c1 else
                                    f"config 3: vicuna-7b Q8_0 + Q8_0 EAGLE head, {per_gpu * world} request streams ({per_gpu} per GPU, one resident weight copy per GPU), "
                                    "64-node tree, KV ~4096 per stream"),
                       "parallelism": f"{world} GPU(s) x {per_gpu} stream(s), round-robin, no batching across streams, no collective",
                       "l2": f"weights ({r['weight_bytes'] / 1e9:.1f} GB) per verify exceed L2: no flush needed"},
            "round_ms_per_stream": r["round_ms"], "verify_step_us": 1e3 * r["verify_ms"],
            "verify_step_hbm_frac": algorithmic_bytes(r["weight_bytes"], r["n_past"], r["T"]) / (r["verify_ms"] * 1e-3) / 1e9 / peaks()[0],
            "gpu_launches": r["launches"], "clocks": clocks}


def run_config4(args, rank, world):
    """BASELINE config 4: one tree-verify step of a LLaMA-2-13B-shaped Q4_K_M target (40 layers, E 5120, 40 heads, F 13824), 65 rows at
    n_kv 4096, tensor parallel over the ranks (N-split wq|wk|wv|gate|up, K-split wo|down, one peer-memory all-reduce after each
    K-split GEMM: csrc/tp.cu).  STRONG scaling: the same step on 1, 2, 4, 8 GPUs; time-like metric (tools/tp_run.py does the work)."""
    if args.impl == "reference":
        return {"impl": "reference", "unavailable": "config 4 is a latency table of this repo's tensor-parallel path; the reference arm covers the metric's own config"} if rank == 0 else None
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import tp_run
    sampler = ClockSampler(int(os.environ.get("LOCAL_RANK", "0"))); sampler.start()
    r = tp_run.bench("13b", args.steps, args.warmup, quiet=True)
    clocks = sampler.summary()
    tp_run.shutdown()
    if rank != 0:
        return None
    return {"metric": "verify_step_us", "value": r["verify_step_us"], "unit": "us", "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup),
            "ms_per_step": r["verify_step_us"] / 1e3, "higher_is_better": False, "scaling": "strong", "vs_baseline": None,
            "dtype": "int8xq4_K/q6_K dot, f32 accumulate, f16 KV", "data": "synthetic (random-init Q4_K_M blocks, N(0,1) f16 KV prefill)",
            "config": {"workload": "config 4: llama-2-13b-shaped Q4_K_M target, 65-row tree verify step at n_kv 4096, batch 1",
                       "parallelism": f"tp{world}: N-split wq|wk|wv|gate|up, K-split wo|down, {r['allreduces_per_step']} peer-memory all-reduces of [65][5120] f32 per step",
                       "l2": "weights per rank exceed L2: no flush needed"},
            "gpu_launches": r["launches"], "weight_bytes_per_rank": r["weight_bytes_per_rank"], "tp_error_word": r["tp_error_word"],
            "clocks": clocks}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--config", type=int, default=2, choices=[1, 2, 3, 4],
                    help="BASELINE.json config: 2 = the metric's own (default, the driver's run); 1 = Q4_0 target, 26-node tree, 512-token prompt, one stream; "
                         "3 = Q8_0 target, 8 request streams per GPU; 4 = LLaMA-2-13B verify step, tensor parallel over --gpus")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.config == 4:
        out = run_config4(args, rank, world)
    elif args.config in (1, 3):
        out = run_config3(args, rank, world, local)
    elif args.impl == "reference":
        out = run_reference(args, rank, world)
    else:
        out = run_gpu(args, rank, world, local)
    if rank == 0 and out is not None:
        print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
