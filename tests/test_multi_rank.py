"""world_size-2 gloo run of the bench's only cross-rank step (replicas only, no data-path
collective): per-rank timings are reduced with MAX, whole-job tokens scale with the world size, the
reference arm runs on rank 0 alone."""
import os
import subprocess
import sys
import textwrap

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = textwrap.dedent("""
    import os, sys, json
    sys.path.insert(0, %r)
    import torch.distributed as dist
    import bench
    rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"])
    dist.init_process_group("gloo")
    wall, gpu = bench.max_over_ranks(dist, [10.0 + 5 * rank, 7.0 - rank], "cpu")
    seeds = [None] * world
    dist.all_gather_object(seeds, bench.replica_seed(rank))
    class A: steps = 1; warmup = 0
    ref = bench.run_reference(A, rank, world) if rank != 0 else "rank0-skipped-here"
    if rank == 0:
        print(json.dumps({"wall": wall, "gpu": gpu, "seeds": seeds, "tokens": bench.job_tokens(4, world)}))
    else:
        assert ref is None          # non-zero ranks of the reference arm do no work
    dist.destroy_process_group()
""") % ROOT


def test_two_rank_gloo(tmp_path):
    w = tmp_path / "worker.py"
    w.write_text(WORKER)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                          "--master-port", "29731", str(w)], capture_output=True, text=True, timeout=300, env=env)
    assert out.returncode == 0, out.stderr[-2000:]
    import json
    line = [l for l in out.stdout.splitlines() if l.startswith("{")][-1]
    r = json.loads(line)
    assert r["wall"] == 15.0 and r["gpu"] == 7.0            # MAX over ranks, per component
    assert r["seeds"] == [3, 4]                              # distinct replicas
    import bench
    assert r["tokens"] == 2 * bench.job_tokens(4, 1)


def test_single_process_passthrough():
    sys.path.insert(0, ROOT)
    import bench
    assert bench.max_over_ranks(None, [1.5, 2.5], "cpu") == [1.5, 2.5]
