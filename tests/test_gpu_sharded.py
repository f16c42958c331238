"""GPU parity of the sharded solve through the C ABI: a one-rank communicator must change nothing, and two
ranks on two GPUs must reproduce the single-GPU solve (the sum of the rank contributions is the same system)."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

import dynosam_b200 as d
from dynosam_b200 import sharded
import scenes

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu


def test_one_rank_communicator_matches_plain_solver():
    g = scenes.small_noisy().full_batch_graph()
    p = d.default_lm_params()
    p.max_iterations = 5
    one = d.GpuLevenbergMarquardt(g, p)
    lm = sharded.ShardedLevenbergMarquardt(g, 0, 1, sharded.make_unique_id(), p)
    assert lm.share.n_factors == g.n_factors and np.array_equal(lm.share.point_global, np.arange(g.n_point))
    assert lm.error() == pytest.approx(one.error(), rel=1e-12)
    pose, pts = lm.optimize()
    r_pose, r_pts = one.optimize()
    assert (lm.iterations(), lm.getInnerIterations()) == (one.iterations(), one.getInnerIterations())
    np.testing.assert_allclose(lm.error_trace, one.error_trace, rtol=1e-9)
    np.testing.assert_allclose(pose, r_pose, atol=1e-8)
    np.testing.assert_allclose(pts, r_pts, atol=1e-8)
    t = lm.timing()
    assert t["collectives"] > 0 and t["collective_bytes"] > 0      # the exchange step really ran through NCCL


def test_pcg_is_refused_when_sharded():
    g = scenes.simulator_cfg1(10).full_batch_graph()
    p = d.default_lm_params()
    p.linear_solver = 1
    lm = sharded.ShardedLevenbergMarquardt(g, 0, 1, sharded.make_unique_id(), p)
    with pytest.raises(d.DynoError):
        lm.optimize()


def test_two_ranks_match_single_gpu():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (gpurun --gpus 2)")
    nproc = 4 if torch.cuda.device_count() >= 4 else 2      # four ranks also exercise the 2 / world scaling of the front-per-GPU exchange
    env = dict(os.environ, OMP_NUM_THREADS="4")
    res = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(nproc), "--master-addr", "127.0.0.1", "--master-port", "29581",
                          os.path.join(ROOT, "tests", "sharded_worker.py")], capture_output=True, text=True, env=env, timeout=900)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    line = [l for l in res.stdout.splitlines() if l.startswith("SHARDED_REPORT ")][-1]
    report = json.loads(line[len("SHARDED_REPORT "):])
    for name, r in report.items():
        assert r["e0_rel"] < 1e-12, (name, r)
        assert r["solved"] == [True, True], (name, r)
        assert r["step_pose"] < 1e-7 and r["step_point"] < 1e-7, (name, r)      # one damped solve (same bar as the single-GPU parity test)
        assert r["iterations"][0] == r["iterations"][1] and r["inner"][0] == r["inner"][1], (name, r)
        assert r["trace_rel"] is not None and r["trace_rel"] < 1e-6, (name, r)   # north_star: final chi2 within 1e-6 relative
        assert r["pose"] < 1e-6 and r["points"] < 1e-6, (name, r)
        assert r["collectives"] > 0
        assert r["band_fronts"] == (2 if name == "medium" else 1) and r["single_front_repeats"] == 0, (name, r)
