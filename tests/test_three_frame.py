"""The reference's own hand-built graph (test/test_rgbd_backend.cc:272-664, restated by tests/golden/make_three_frame.py)
through every implementation here: the independent dense numpy LM, the CPU oracle, and (-m gpu) the CUDA path. The
reference asserts no output for it, so the three are checked against each other; the keys, the natural variable order and
the block pattern of J^T J are checked against a plain numpy statement of the key tuples."""
import json
import os

import numpy as np
import pytest

import dynosam_b200 as d
import np_ref as ref
import scenes
from oracle_lib import OracleGraph

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "three_frame.json")


@pytest.fixture(scope="module", params=["graph", "full_graph"])
def g(request):
    return scenes.three_frame(request.param)


def test_keys_match_library_encoding():
    with open(GOLDEN) as f:
        known = json.load(f)["known_keys"]
    assert int(known["X0"]) == d.CameraPoseSymbol(0)
    assert int(known["l0"]) == d.StaticLandmarkSymbol(0)
    assert int(known["m_frame2_tracklet3"]) == d.DynamicLandmarkSymbol(2, 3)
    assert int(known["H_object1_frame2"]) == d.ObjectMotionSymbol(1, 2)


def test_natural_order_and_sizes(g):
    # gtsam::Values iterates in ascending key order: H.. < X.. among poses, l.. < m.. among points
    chars = [d.key_decode(int(k))[0] for k in g.pose_keys]
    assert chars == sorted(chars) and set(chars) == {"H", "X"}
    assert np.all(np.diff(g.pose_keys.astype(np.uint64)) > 0) and np.all(np.diff(g.point_keys.astype(np.uint64)) > 0)
    assert [d.key_decode(int(k))[0] for k in g.point_keys][0] == "l"
    frames = 3 if g.n_pose == 5 else 4
    assert (g.n_pose, g.n_point) == (frames + frames - 1, 1 + 3 * frames)
    assert g.n_p2p == 4 * frames and g.n_ternary == 3 * (frames - 1) and g.n_between == frames - 1 and g.n_prior == 1
    order = np.concatenate([g.p2p_order, g.ter_order, g.btw_order, g.pri_order])
    assert sorted(order.tolist()) == list(range(g.n_factors))   # one insertion slot per factor


def test_oracle_matches_dense_numpy(g):
    ora = OracleGraph.from_graph(g)
    dg = ref.DenseGraph(g)
    assert abs(ora.error() - dg.error()) <= 1e-12 * dg.error()
    r = ora.optimize(max_iterations=4, relative_error_tol=0.0, absolute_error_tol=0.0)
    n = ref.dense_lm(g, max_iterations=4, relative_error_tol=0.0, absolute_error_tol=0.0)
    assert (r["iterations"], r["inner_iterations"]) == (n["iterations"], n["inner_iterations"])
    assert np.allclose(r["error_trace"], n["error_trace"], rtol=1e-6, atol=1e-14)
    assert r["error_trace"][-1] < 0.9 * r["error_trace"][0]   # the static observations of this hand-made graph are not consistent: chi^2 does not reach zero


def test_host_hessian_pattern_is_exact():
    """north_star: sparsity pattern, variable ordering and block-index maps bit-exact. The library's pattern (host pass
    over a flattened graph) against the key tuples in numpy, on BASELINE config 1 and a noisy scene."""
    for graph in (scenes.simulator_cfg1().full_batch_graph(), scenes.small_noisy().full_batch_graph(), scenes.small_noisy(seed=3).window_graph(5, 30, True)):
        row_ptr, col = graph.hessian_pattern()
        want_ptr, want_col = scenes.numpy_hessian_pattern(graph)
        assert np.array_equal(row_ptr, want_ptr) and np.array_equal(col, want_col)
        assert row_ptr[-1] == graph.n_pose + graph.n_point + graph.n_p2p + 3 * graph.n_ternary + graph.n_between - _duplicates(graph)


def _duplicates(g):
    # pairs named by more than one factor (e.g. two ternary factors of one object at one frame share (m, H)? no: (point, motion)
    # pairs are unique per factor; between pairs may repeat) -- counted so the closed form above stays exact
    pairs = {}
    n = g.n_pose
    for a, b in zip(g.p2p_pose, g.p2p_point):
        pairs[(int(a), n + int(b))] = pairs.get((int(a), n + int(b)), 0) + 1
    for a, b, h in zip(g.ter_prev, g.ter_cur, g.ter_motion):
        for k in ((n + int(a), n + int(b)), (n + int(a), int(h)), (n + int(b), int(h))):
            k = (min(k), max(k))
            pairs[k] = pairs.get(k, 0) + 1
    for a, b in zip(g.btw_a, g.btw_b):
        k = (min(int(a), int(b)), max(int(a), int(b)))
        pairs[k] = pairs.get(k, 0) + 1
    return sum(c - 1 for c in pairs.values())


@pytest.mark.gpu
def test_gpu_matches_oracle_on_reference_graph(g):
    p = d.default_lm_params()
    p.max_iterations = 4   # past 5 iterations the cost change is round-off and the accept decisions are noise
    p.relative_error_tol = 0.0
    p.absolute_error_tol = 0.0
    lm = d.GpuLevenbergMarquardt(None, p)
    lm.load_host_arrays(g)
    ora = OracleGraph.from_graph(g)
    assert abs(lm.error() - ora.error()) <= 1e-12 * ora.error()
    a, b = lm.linearize(), ora.linearize()
    # a residual that is zero up to round-off (the odometry measurements are pose_{k-1}.between(pose_k) exactly) has no
    # relative error: b is measured against max(|b|, whitened residual of a 1 cm offset), as in test_gpu_parity.py
    companion = {"p2p_b": "p2p_A_point", "ter_b": "ter_A_prev", "btw_b": "btw_A_b", "pri_b": "pri_A"}
    for k in b:
        if b[k].size:
            scale = np.max(np.abs(b[k]))
            if k in companion:
                scale = max(scale, 1e-2 * np.max(np.abs(b[companion[k]])))
            assert np.max(np.abs(a[k] - b[k])) <= 1e-10 * max(scale, 1e-30), k
    row_ptr, col = lm.hessian_pattern()
    want_ptr, want_col = scenes.numpy_hessian_pattern(g)
    assert np.array_equal(row_ptr, want_ptr) and np.array_equal(col, want_col)
    lm.optimize()
    r = ora.optimize(max_iterations=4, relative_error_tol=0.0, absolute_error_tol=0.0)
    assert (lm.iterations(), lm.getInnerIterations()) == (r["iterations"], r["inner_iterations"])
    assert np.allclose(lm.error_trace, r["error_trace"], rtol=1e-6, atol=1e-14)
    pose, pt = lm.values()
    opose, opt = ora.values()
    assert np.max(np.abs(pose - opose)) <= 1e-6 and np.max(np.abs(pt - opt)) <= 1e-6


@pytest.mark.gpu
def test_gpu_hessian_pattern_is_exact():
    for graph in (scenes.simulator_cfg1().full_batch_graph(), scenes.small_noisy().full_batch_graph()):
        lm = d.GpuLevenbergMarquardt(graph)
        row_ptr, col = lm.hessian_pattern()
        want_ptr, want_col = scenes.numpy_hessian_pattern(graph)
        assert np.array_equal(row_ptr, want_ptr) and np.array_equal(col, want_col)
