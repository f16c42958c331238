"""CPU tests of the N > 1 plumbing of bench.py over gloo, world_size 2 (the GPU run uses the same code over NCCL).
The path does not shard inside one solve this round: bench.py --gpus N runs N independent replicas."""
import json
import os
import socket
import subprocess
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import bench
    dist.init_process_group("gloo", rank=rank, world_size=world)
    # per-rank step time (ms) and counts, as run_ours produces them
    maxima = (100.0 + 25.0 * rank, 120.0 - 5.0 * rank, 0.5 + rank)
    sums = (15 + rank, 600 + 10 * rank, 15)
    m, c = bench.aggregate_over_ranks(dist, maxima, sums, "cpu")
    cfg = bench.replica_config("cfg3", rank)
    dist.barrier()
    dist.destroy_process_group()
    json.dump({"max": m, "sum": c, "seed": cfg["seed"], "frames": cfg["n_frames"]}, open(os.path.join(out, "r%d.json" % rank), "w"))


def test_rank_aggregation_over_gloo(tmp_path):
    world, port = 2, _free_port()
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    r0, r1 = (json.load(open(tmp_path / ("r%d.json" % r))) for r in range(world))
    assert r0["max"] == r1["max"] == [125.0, 120.0, 1.5]          # slowest rank decides the time
    assert r0["sum"] == r1["sum"] == [31.0, 1210.0, 30.0]           # units are whole-job totals
    assert r0["frames"] == r1["frames"] and r1["seed"] == r0["seed"] + 1   # same shape, independent scenes
    value = r0["sum"][0] / (r0["max"][0] * 1e-3)
    assert value == pytest.approx(31 / 0.125)


def test_single_process_aggregation_is_identity():
    import bench
    m, c = bench.aggregate_over_ranks(None, (1.0, 2.0), (3, 4), "cpu")
    assert (m, c) == ([1.0, 2.0], [3, 4])


def test_reference_arm_runs_on_rank_zero_only():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    res = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0"],
                         capture_output=True, text=True, env=env, timeout=300)
    assert res.returncode == 0 and res.stdout.strip() == ""


def test_bench_refuses_to_run_without_a_gpu():
    if torch.cuda.is_available():
        pytest.skip("box has a GPU")
    res = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "1", "--warmup", "0"], capture_output=True, text=True, timeout=600)
    assert res.returncode != 0 and "no CPU fallback" in (res.stderr + res.stdout)
