"""The N>1 host logic of bench.py (replicas, weak scaling) on CPU: two gloo ranks report different
durations, the job's duration is the maximum and the rate counts the units of both."""
import os
import subprocess
import sys
import textwrap

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = textwrap.dedent("""
    import json, os, sys
    sys.path.insert(0, %r)
    import torch.distributed as dist
    import bench
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    local = [10.0 + 5.0 * rank, 20.0 - 3.0 * rank]           # rank 0: 10, 20 ms; rank 1: 15, 17 ms
    ms = bench.job_times(local, dist, "cpu")
    dist.barrier()
    if rank == 0:
        print(json.dumps({"ms": ms, "rate": bench.whole_job_rate(world, 30, ms[0]), "world": world}))
    dist.destroy_process_group()
""") % ROOT


def test_two_ranks_take_the_maximum_and_count_both(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", OMP_NUM_THREADS="1")
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                          "--master-port", "29533", str(script)], capture_output=True, text=True, env=env, timeout=240)
    assert out.returncode == 0, out.stderr[-2000:]
    import json
    line = [l for l in out.stdout.splitlines() if l.startswith("{")][-1]
    r = json.loads(line)
    assert r["world"] == 2
    assert r["ms"] == [15.0, 20.0]                      # max over ranks, per timed region
    assert abs(r["rate"] - 2 * 30 / 0.015) < 1e-6       # both replicas' steps over the slower one's time


def test_single_rank_needs_no_process_group():
    sys.path.insert(0, ROOT)
    import bench
    assert bench.job_times([3.0, 4.0]) == [3.0, 4.0]
    assert bench.whole_job_rate(1, 128, 256.0) == 500.0
