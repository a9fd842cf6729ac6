"""Tensor parallelism of the target's verify forward (BASELINE config 4; csrc/tp.cu, tools/tp_run.py).  Replaces the reference's
row-split of the weight matrices over devices (R/ggml/src/ggml-cuda/ggml-cuda.cu:735-748 split buffers, :1359-1668 the row-split
MUL_MAT with its peer copies).

CPU: the partition and the block slicing (host logic), and a world_size-2 gloo run in which each rank dequantises ITS slices with the
oracle, multiplies, and the gloo all-reduce of the K-split partial products must reproduce the whole matrices' products.
GPU (needs 2 GPUs, skipped otherwise): the 2-rank torchrun check of the real path — the peer-memory all-reduce against its known
answer, every rank holding the same bits, and the sharded model against the single-GPU one."""
import json
import os
import subprocess
import sys
import textwrap

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from __graft_entry__ import load_package  # noqa: E402


@pytest.mark.parametrize("H,F", [(40, 13824), (32, 11008), (8, 2816)])
@pytest.mark.parametrize("size", [1, 2, 4, 8])
def test_tp_partition(H, F, size):
    eb = load_package()
    if H // 2 < size:
        pytest.skip("fewer head pairs than ranks")
    parts = eb.tp_partition(H, F, size)
    assert len(parts) == size
    h = f = 0
    for h0, nh, f0, nf in parts:
        assert h0 == h and f0 == f and nh > 0 and nf > 0
        assert nh % 2 == 0 and nf % 256 == 0            # 256-element blocks of the K-split wo (D = 128) and ffn_down
        h += nh; f += nf
    assert h == H and f == F
    assert max(p[1] for p in parts) - min(p[1] for p in parts) <= 2 and max(p[3] for p in parts) - min(p[3] for p in parts) <= 256


WORKER = textwrap.dedent("""
    import os, sys, json
    import numpy as np
    sys.path.insert(0, %(root)r); sys.path.insert(0, os.path.join(%(root)r, "tests")); sys.path.insert(0, os.path.join(%(root)r, "oracle"))
    import torch, torch.distributed as dist
    from __graft_entry__ import load_package
    import fixtures as fx, pyoracle as po
    eb = load_package()
    rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"])
    dist.init_process_group("gloo")
    E, H, F, D = 512, 4, 768, 128
    tgt = fx.make_model(E=E, H=H, F=F, V=256, L=1, ftype="q4_k_m", seed=9)
    part = eb.tp_partition(H, F, world)[rank]
    h0, nh, f0, nf = part
    o = po.oracle()
    def deq(which, rows, cols, sliced):
        t, raw = tgt["layers"][0][which]
        if sliced:
            raw = eb.tp_shard_tensor(which, t, raw, E, F, D, part)
        return o.dequantize(t, raw, rows * cols).reshape(rows, cols)
    rng = np.random.default_rng(1)
    x = rng.normal(0, 1, (5, E)).astype(np.float32)
    # N-split: this rank's rows of wq / ffn_gate are exactly rows [h0*D, ...) / [f0, ...) of the whole matrix
    wq_full, wq_mine = deq(po.L_WQ, E, E, False), deq(po.L_WQ, nh * D, E, True)
    gate_full, gate_mine = deq(po.L_GATE, F, E, False), deq(po.L_GATE, nf, E, True)
    n_ok = bool(np.array_equal(wq_mine, wq_full[h0 * D:(h0 + nh) * D]) and np.array_equal(gate_mine, gate_full[f0:f0 + nf]))
    # K-split: partial products of wo / ffn_down over this rank's columns, summed over the group, equal the whole product
    att = rng.normal(0, 1, (5, E)).astype(np.float32); act = rng.normal(0, 1, (5, F)).astype(np.float32)
    wo_full, wo_mine = deq(po.L_WO, E, E, False), deq(po.L_WO, E, nh * D, True)
    dn_full, dn_mine = deq(po.L_DOWN, E, F, False), deq(po.L_DOWN, E, nf, True)
    k_ok = bool(np.array_equal(wo_mine, wo_full[:, h0 * D:(h0 + nh) * D]) and np.array_equal(dn_mine, dn_full[:, f0:f0 + nf]))
    p_wo = torch.from_numpy(att[:, h0 * D:(h0 + nh) * D].astype(np.float64) @ wo_mine.T.astype(np.float64))
    p_dn = torch.from_numpy(act[:, f0:f0 + nf].astype(np.float64) @ dn_mine.T.astype(np.float64))
    dist.all_reduce(p_wo); dist.all_reduce(p_dn)
    e_wo = float(np.abs(p_wo.numpy() - att.astype(np.float64) @ wo_full.T.astype(np.float64)).max())
    e_dn = float(np.abs(p_dn.numpy() - act.astype(np.float64) @ dn_full.T.astype(np.float64)).max())
    oks = [None] * world
    dist.all_gather_object(oks, (n_ok, k_ok, e_wo, e_dn, part))
    if rank == 0:
        print("TP_HOST_JSON " + json.dumps(oks))
    dist.destroy_process_group()
""") % {"root": ROOT}


def test_tp_shards_reduce_to_whole_gloo(tmp_path):
    w = tmp_path / "worker.py"
    w.write_text(WORKER)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                          "--master-port", "29733", str(w)], capture_output=True, text=True, timeout=300, env=env)
    assert out.returncode == 0, out.stdout[-1500:] + out.stderr[-2500:]
    line = [l for l in out.stdout.splitlines() if l.startswith("TP_HOST_JSON ")][-1]
    res = json.loads(line[len("TP_HOST_JSON "):])
    assert len(res) == 2 and res[0][4] != res[1][4]
    for n_ok, k_ok, e_wo, e_dn, part in res:
        assert n_ok and k_ok, part
        assert e_wo < 1e-9 and e_dn < 1e-9, (e_wo, e_dn)          # float64 products: only the association differs


@pytest.mark.gpu
def test_tp2_matches_single_gpu():
    """2 ranks over peer memory vs one GPU.  The all-reduce is exact against its known answer and leaves the same bits on every rank;
    the 1-layer model (two K-split outputs -> two quantisers) has >= 80 % of its rows free of rounding flips and those agree to 1e-5;
    every row of the 3-layer model is inside the flip bound of tests/parity.py with the whole-graph NMSE bar of the reference's own
    test (2e-3, R/tests/test-backend-ops.cpp:3482), and no greedy id differs where the margin exceeds twice the row's deviation."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run with gpurun --gpus 2)")
    import parity as pa
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                          "--master-port", "29735", os.path.join(ROOT, "tools", "tp_run.py"), "check"], capture_output=True, text=True, timeout=600)
    line = [l for l in out.stdout.splitlines() if l.startswith("TP_CHECK_JSON ")]
    assert out.returncode == 0 and line, out.stdout[-1500:] + out.stderr[-2500:]
    res = json.loads(line[-1][len("TP_CHECK_JSON "):])
    print("[tp] " + json.dumps(res))
    assert res["world"] == 2 and res["allreduce_max_abs_err"] == 0.0 and res["tp_error_word"] == 0
    one, three = res["models"]
    assert one["layers"] == 1 and three["layers"] == 3
    for m in (one, three):
        assert m["ranks_bit_identical"]
        for tag in ("prompt", "tree"):
            r = m[tag]
            assert r["max_rel"] <= pa.FLIP_REL and r["nmse"] <= 2e-3 and r["greedy_mismatch"] == 0, (m["layers"], tag, r)
    for tag in ("prompt", "tree"):
        assert one[tag]["rows_within_1e-5"] >= 0.8 * one[tag]["rows"], one[tag]
