"""The DynoSAM text graph file (reference writer: src/backend/FactorGraphTools.cc:256-381): write -> read round trips,
the reference's line layout, error reporting; on the GPU, a graph that went through the file solves like the original."""
import numpy as np
import pytest

import dynosam_b200 as d
import scenes
from dynosam_b200 import graphfile as gf
from oracle_lib import OracleGraph

ARRAYS = ("pose_keys", "point_keys", "p2p_pose", "p2p_point", "ter_prev", "ter_cur", "ter_motion", "btw_a", "btw_b", "pri_a")
FLOATS = ("pose_Rt", "point_xyz", "p2p_z", "p2p_sigma", "ter_sigma", "btw_z", "btw_sigmas", "pri_z", "pri_sigmas")


def test_quaternion_round_trip():
    rng = np.random.default_rng(4)
    for _ in range(50):
        q = rng.normal(size=4)
        q /= np.linalg.norm(q)
        R = gf.quat_to_R(*q)
        assert np.allclose(R @ R.T, np.eye(3), atol=1e-14) and np.linalg.det(R) > 0
        q2 = gf.R_to_quat(R)
        assert np.allclose(q2, q if q[3] >= 0 else -q, atol=1e-14)


def test_write_read_round_trip(tmp_path):
    g = scenes.small_noisy(n_frames=12).full_batch_graph()
    path = tmp_path / "graph.txt"
    gf.write_graph_file(str(path), g)
    h = gf.read_graph_file(str(path), huber_k=1e-4)
    assert (h.n_pose, h.n_point, h.n_p2p, h.n_ternary, h.n_between, h.n_prior, h.n_factors) == \
           (g.n_pose, g.n_point, g.n_p2p, g.n_ternary, g.n_between, g.n_prior, g.n_factors)
    for a in ARRAYS:
        assert np.array_equal(getattr(h, a), getattr(g, a)), a
    for a in FLOATS:
        assert np.max(np.abs(getattr(h, a) - getattr(g, a))) <= 1e-14, a     # 17 significant digits; rotations via quaternions
    assert np.array_equal(h.p2p_order, g.p2p_order) and np.array_equal(h.ter_order, g.ter_order)
    # the file is an equivalent graph for the oracle too
    e_g, e_h = OracleGraph.from_graph(g).error(), OracleGraph.from_graph(h).error()
    assert abs(e_g - e_h) <= 1e-10 * e_g


def test_file_layout_matches_the_reference_writer(tmp_path):
    g = d.Scene.simulator(4).full_batch_graph()
    path = tmp_path / "graph.txt"
    gf.write_graph_file(str(path), g, precision=6)      # the reference writes with the default ostream precision
    lines = open(path).read().splitlines()
    tags = [ln.split()[0] for ln in lines]
    assert set(tags) <= {"SE3_PRIOR_FACTOR", "SE3_BETWEEN_FACTOR", "SE3_MOTION_FACTOR", "POSE_TO_POINT_FACTOR",
                         "SE3_POSE_VALUE", "SE3_MOTION_VALUE", "POINT3_STATIC_VALUE", "POINT3_DYNAMIC_VALUE"}
    n_tok = {"SE3_PRIOR_FACTOR": 1 + 1 + 7 + 21, "SE3_BETWEEN_FACTOR": 1 + 2 + 7 + 21, "SE3_MOTION_FACTOR": 1 + 3 + 3 + 6,
             "POSE_TO_POINT_FACTOR": 1 + 2 + 3 + 6, "SE3_POSE_VALUE": 2 + 7, "SE3_MOTION_VALUE": 2 + 7, "POINT3_STATIC_VALUE": 2 + 3,
             "POINT3_DYNAMIC_VALUE": 2 + 3}
    for ln in lines:
        tok = ln.split()
        assert len(tok) == n_tok[tok[0]], ln
    # a variable is written right after the first factor that mentions it, and only once
    seen, value_keys = set(), []
    for ln in lines:
        tok = ln.split()
        if tok[0].endswith("_VALUE"):
            assert int(tok[1]) in seen and int(tok[1]) not in value_keys
            value_keys.append(int(tok[1]))
        else:
            n_keys = {"SE3_PRIOR_FACTOR": 1, "SE3_BETWEEN_FACTOR": 2, "SE3_MOTION_FACTOR": 3, "POSE_TO_POINT_FACTOR": 2}[tok[0]]
            seen.update(int(t) for t in tok[1:1 + n_keys])
    assert len(value_keys) == g.n_pose + g.n_point
    # the motion factor carries a zero "measurement" and the information of sigma 0.01
    mot = [ln.split() for ln in lines if ln.startswith("SE3_MOTION_FACTOR")][0]
    assert mot[4:7] == ["0", "0", "0"] and float(mot[7]) == pytest.approx(1e4)
    # six-digit files still parse, values agree to that precision
    h = gf.read_graph_file(str(path))
    assert np.max(np.abs(h.point_xyz - g.point_xyz)) <= 1e-3 and np.max(np.abs(h.pose_Rt - g.pose_Rt)) <= 1e-3


def test_reader_reports_bad_input(tmp_path):
    p = tmp_path / "bad.txt"
    p.write_text("SE3_WHATEVER 1 2 3\n")
    with pytest.raises(ValueError, match="unknown tag"):
        gf.read_graph_file(str(p))
    X0 = d.CameraPoseSymbol(0)
    p.write_text("POSE_TO_POINT_FACTOR %d %d 1 2 3 1 0 0 1 0 1\nSE3_POSE_VALUE %d 0 0 0 0 0 0 1\n" % (X0, d.StaticLandmarkSymbol(5), X0))
    with pytest.raises(KeyError, match="no value line"):
        gf.read_graph_file(str(p))
    p.write_text("POSE_TO_POINT_FACTOR %d %d 1 2 3 1 0.5 0 1 0 1\n" % (X0, d.StaticLandmarkSymbol(5)))
    with pytest.raises(ValueError, match="diagonal noise"):
        gf.read_graph_file(str(p))


@pytest.mark.gpu
def test_graph_from_file_solves_like_the_original(tmp_path):
    g = scenes.small_noisy().full_batch_graph()
    path = tmp_path / "graph.txt"
    gf.write_graph_file(str(path), g)
    h = gf.read_graph_file(str(path), huber_k=1e-4)
    p = d.default_lm_params()
    p.max_iterations = 4
    p.relative_error_tol = 0.0
    p.absolute_error_tol = 0.0
    a = d.GpuLevenbergMarquardt(g, p)
    b = d.GpuLevenbergMarquardt(None, p)
    b.load_host_arrays(h)
    assert abs(a.error() - b.error()) <= 1e-10 * a.error()
    a.optimize(); b.optimize()
    assert a.iterations() == b.iterations()
    assert np.allclose(a.error_trace, b.error_trace, rtol=1e-8, atol=0)
