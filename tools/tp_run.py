"""Tensor-parallel target forward (BASELINE config 4), one process per GPU, launched with torchrun:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node G --master-addr 127.0.0.1 --master-port 29511 tools/tp_run.py check
    ... tools/tp_run.py bench [13b|7b] [steps]

check: a small Q4_K_M model, full on rank 0 and sharded over the G ranks: allreduce known answer, prefix + tree forward, TP logits vs the
       single-GPU logits (rows with identical arithmetic agree to 1e-5; the K-split changes the f32 summation order, so a row may carry a
       rounding flip: bounded as in tests/parity.py).  Prints TP_CHECK_JSON.
bench: LLaMA-2-13B-shaped (or 7B) Q4_K_M target with synthetic weights, n_kv 4096, 65-row tree: verify-step time at TP = G, device-timed,
       max over ranks.  Prints TP_BENCH_JSON.  G = 1 runs the unsharded engine on one GPU.
torch.distributed is only the host-side plumbing (handle exchange, barrier, max over ranks); the data path is csrc/tp.cu."""
import ctypes as C
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "oracle"))
from __graft_entry__ import load_package  # noqa: E402

rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
eb = load_package(); eb.init(local)
dist = None
if world > 1:
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))


def exchange(hb):
    if dist is None:
        return [hb]
    out = [None] * world
    dist.all_gather_object(out, hb)
    return out


def barrier():
    if dist is not None:
        dist.barrier()


def max_over_ranks(x):
    if dist is None:
        return x
    import torch
    t = torch.tensor([x], device="cuda", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def shard_model(fxm, part, D):
    """fixtures.make_model dict -> this rank's ShardModel (N-split wq|wk|wv|gate|up, K-split wo|down, the rest replicated)"""
    hp = fxm["hparams"]
    m = eb.ShardModel(hp["E"], D, part[1], part[3], hp["V"], hp["L"], hp["eps"], hp["freq_base"])
    for which, (t, raw) in fxm["globals"].items():
        m.set_tensor(-1, which, t, raw)
    for l, d in enumerate(fxm["layers"]):
        for which, (t, raw) in d.items():
            m.set_tensor(l, which, t, eb.tp_shard_tensor(which, t, raw, hp["E"], hp["F"], D, part))
    return m


def check_allreduce(tp, E):
    """the collective by itself: known answer, three calls (both buffer parities, flags reused) -> max abs error"""
    ar_err = 0.0
    for call in range(3):
        n = 65 * E
        planes = np.stack([np.random.default_rng(100 * call + 10 * rank + s).normal(0, 1, n).astype(np.float32) for s in range(3)])
        got = tp.allreduce(planes)
        want = np.zeros(n, np.float32)
        for r in range(world):
            pr = [np.random.default_rng(100 * call + 10 * r + s).normal(0, 1, n).astype(np.float32) for s in range(3)]
            want += (pr[0] + pr[1]) + pr[2]
        ar_err = max(ar_err, float(np.abs(got - want).max()))
    return ar_err


def check_model(tp, L):
    """prompt + tree forward of an L-layer Q4_K_M model, sharded over the group vs whole on rank 0.  With L = 1 the only arithmetic
    that differs is the f32 summation order of the two K-split outputs, met by two quantisers: almost every row is free of rounding
    flips and must agree to 1e-5; deeper models accumulate flips and are bounded like any two builds."""
    import fixtures as fx
    import parity as pa
    hp = dict(E=1024, H=8, F=2816, V=2048, L=L)            # F = 11 super-blocks: uneven split (6 + 5 at two ranks)
    tgt = fx.make_model(**hp, ftype="q4_k_m", seed=77)
    D = hp["E"] // hp["H"]
    parts = eb.tp_partition(hp["H"], hp["F"], world)
    rng = np.random.default_rng(5)
    P, T, n_ctx = 40, 33, 256
    prompt = rng.integers(3, hp["V"], P).astype(np.int32)
    par = fx.driver_like_tree(T, 8, rng); depth = fx.tree_depths(par); toks = rng.integers(3, hp["V"], T).astype(np.int32)
    ms = shard_model(tgt, parts[rank], D)
    s = eb.Session(ms, None, n_ctx=n_ctx, max_nodes=65, use_graph=False)
    s.set_tp(tp)
    lp, _ = s.forward(0, prompt, np.arange(P)); s.commit(0, P)
    lt, _ = s.forward(0, toks, P + depth, parent=par)
    res = {"layers": L}
    if dist is not None:
        import torch
        # every rank must hold the same bits (the reduction adds in rank order everywhere)
        mine = torch.from_numpy(np.concatenate([lp, lt]).copy()).cuda(); ref0 = mine.clone(); dist.broadcast(ref0, 0)
        same = torch.tensor([int(torch.equal(mine, ref0))], device="cuda"); dist.all_reduce(same, op=dist.ReduceOp.MIN)
        res["ranks_bit_identical"] = bool(same.item())
    if rank == 0:
        mf = fx.to_gpu_model(eb, tgt)
        s1 = eb.Session(mf, None, n_ctx=n_ctx, max_nodes=65, use_graph=False)
        lp1, _ = s1.forward(0, prompt, np.arange(P)); s1.commit(0, P)
        lt1, _ = s1.forward(0, toks, P + depth, parent=par)
        for tag, a, b in (("prompt", lp, lp1), ("tree", lt, lt1)):
            rel = pa.row_rel(a, b)
            # greedy ids: a mismatch only counts where the single-GPU top-2 margin exceeds twice that row's deviation
            top2 = np.sort(b, axis=1)[:, -2:]; margin = (top2[:, 1] - top2[:, 0]) / np.abs(b).max(axis=1)
            diff = a.argmax(1) != b.argmax(1)
            res[tag] = {"rows": int(len(rel)), "rows_within_1e-5": int((rel <= 1e-5).sum()), "max_rel": float(rel.max()),
                        "nmse": float(((a - b) ** 2).sum() / (b ** 2).sum()), "greedy_differs": int(diff.sum()),
                        "greedy_mismatch": int((diff & (margin > 2 * rel)).sum())}
        s1.free()
    s.free(); barrier()
    return res


def check():
    tp = eb.TpGroup(rank, world, 128 * 1024, exchange)
    barrier()
    res = {"world": world, "allreduce_max_abs_err": check_allreduce(tp, 1024)}
    res["models"] = [check_model(tp, L) for L in (1, 3)]
    res["tp_error_word"] = tp.error()
    if rank == 0:
        print("TP_CHECK_JSON " + json.dumps(res), flush=True)
    else:
        assert all(m.get("ranks_bit_identical", True) for m in res["models"]), "rank %d holds different logits" % rank


def bench_allreduce(iters=200):
    """the collective alone: [65][5120] f32 with 4 split-K planes per rank; device time per call from a CUDA-graph replay of `iters` calls
    (eb_tp_bench), in the engine's form (rank partials left in the slots for the consumer) and in the operator form (rank-order sum included)"""
    E, T, ks = 5120, 65, 4
    tp = eb.TpGroup(rank, world, 128 * E, exchange)
    n = T * E
    planes = eb.DeviceBuffer.from_host(np.random.default_rng(rank).normal(0, 1, (ks, n)).astype(np.float32)); out = eb.DeviceBuffer(n * 4)
    lib = eb.lib()
    res = {}
    for form, eng in (("engine_sets", 1), ("operator_sets", 0)):
        us = C.c_float()
        barrier()
        rc = lib.eb_tp_bench(tp.h, planes.ptr, ks, n, out.ptr, eng, iters, C.byref(us))
        assert rc == 0, lib.eb_last_error().decode()
        res[form] = max_over_ranks(float(us.value))
    if rank == 0:
        print("TP_AR_JSON " + json.dumps({"tp": world, "elems": n, "planes": ks, "us_per_allreduce": res["engine_sets"],
                                          "us_per_allreduce_operator_sets": res["operator_sets"], "tp_error_word": tp.error()}), flush=True)


def bench(shape, steps, warmup=5, quiet=False):
    HPS = {"13b": dict(E=5120, H=40, F=13824, V=32000, L=40), "7b": dict(E=4096, H=32, F=11008, V=32000, L=32)}
    hp = HPS[shape]; D = 128
    more = lambda l: l < hp["L"] // 8 or l >= 7 * hp["L"] // 8 or (l - hp["L"] // 8) % 3 == 2      # R/src/llama-quant.cpp:129-131
    if world == 1:
        m = eb.Model(**hp)
    else:
        h0, nh, f0, nf = eb.tp_partition(hp["H"], hp["F"], world)[rank]
        m = eb.ShardModel(hp["E"], D, nh, nf, hp["V"], hp["L"])
    m.synth_tensor(-1, eb.T_TOK_EMBD, eb.Q4_K, 1); m.synth_tensor(-1, eb.T_OUTPUT_NORM, eb.F32, 2); m.synth_tensor(-1, eb.T_OUTPUT, eb.Q6_K, 3)
    for l in range(hp["L"]):
        for which in range(9):
            t = eb.F32 if which in (eb.L_ATTN_NORM, eb.L_FFN_NORM) else (eb.Q6_K if which in (eb.L_WV, eb.L_DOWN) and more(l) else eb.Q4_K)
            m.synth_tensor(l, which, t, 1000 + 16 * l + which)
    s = eb.Session(m, None, n_ctx=4096 + 256, max_nodes=65, use_graph=True)
    tp = None
    if world > 1:
        tp = eb.TpGroup(rank, world, 128 * hp["E"], exchange)
        s.set_tp(tp)
    s.synth_kv(4096 - 65, seed=3)
    barrier()
    for _ in range(max(3, warmup)):
        s.verify_only(65)
    barrier()
    n0 = eb.launch_count()
    ts = []
    for _ in range(steps):
        barrier()                                    # ranks enter each step together (a step is one graph replay per rank)
        ts.append(s.verify_only(65))
    launches = eb.launch_count() - n0
    ms = max_over_ranks(float(np.median(ts)))
    res = {"shape": shape, "tp": world, "verify_step_us": 1e3 * ms, "steps": steps, "weight_bytes_per_rank": m.weight_bytes(),
           "tp_error_word": tp.error() if tp else 0, "launches": launches, "allreduces_per_step": 2 * hp["L"] if tp else 0}
    if rank == 0 and not quiet:
        print("TP_BENCH_JSON " + json.dumps(res), flush=True)
    s.free()
    return res


def shutdown():
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    mode = sys.argv[1] if len(sys.argv) > 1 else "check"
    if mode == "check":
        check()
    elif mode == "ar":
        bench_allreduce()
    else:
        bench(sys.argv[2] if len(sys.argv) > 2 else "13b", int(sys.argv[3]) if len(sys.argv) > 3 else 20)
    shutdown()
