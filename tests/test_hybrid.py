"""Hybrid (object-centric) formulation, builder-defined (the fork has none): emitted through the exact
re-parametrisation T_k = H_k^-1 X_k (csrc/host/formulation_hybrid.hpp), so every factor is a reference factor type and the
oracle, the dense numpy LM and the device kernels apply unchanged. Parity is self-pinned: numpy <-> oracle <-> GPU."""
import numpy as np
import pytest

import dynosam_b200 as d
from dynosam_b200 import postsolve as ps
import np_ref as ref
from oracle_lib import OracleGraph


def tiny():
    return d.Scene.synthetic(n_frames=7, n_objects=2, static_per_frame=10, dynamic_per_object=5, static_track_max=5, dynamic_track_max=5, seed=21)


def small(seed=7):
    return d.Scene.synthetic(n_frames=40, n_objects=3, static_per_frame=50, dynamic_per_object=12, static_track_max=20, dynamic_track_max=15, seed=seed)


def test_structure():
    scene = small()
    g, w = scene.hybrid_graph(), scene.full_batch_graph()
    assert g.n_ternary == 0 and g.n_prior == 1
    chars = np.array([d.key_decode(int(k))[0] for k in g.pose_keys])
    assert set(chars) == {"L", "X"} and (chars == "X").sum() == 40
    pchars = np.array([d.key_decode(int(k))[0] for k in g.point_keys])
    dyn = pchars == "m"
    # one point per tracklet: far fewer dynamic points than the world-centric graph has, about as many factors as measurements
    assert dyn.sum() < 0.25 * (np.array([d.key_decode(int(k))[0] for k in w.point_keys]) == "m").sum()
    tracklets = [d.key_decode(int(k))[3] for k in g.point_keys[dyn]]
    assert len(set(tracklets)) == len(tracklets)
    deg = np.bincount(g.p2p_point, minlength=g.n_point)
    assert deg[dyn].min() >= 3                      # min_dynamic_obs
    # a dynamic point is observed from its object's virtual cameras, or from the camera itself at the keyframe
    pose_char = chars[g.p2p_pose]
    is_dyn = dyn[g.p2p_point]
    assert np.all(pose_char[~is_dyn] == "X") and (pose_char[is_dyn] == "L").mean() > 0.8 and (pose_char[is_dyn] == "X").any()
    # initial values: a point sits where its first observation puts it
    first = {}
    for i in np.argsort(g.p2p_order):
        first.setdefault(int(g.p2p_point[i]), i)
    for pt, i in list(first.items())[::7]:
        Rt = g.pose_Rt[g.p2p_pose[i]]
        assert np.allclose(g.point_xyz[pt], Rt[:9].reshape(3, 3) @ g.p2p_z[i] + Rt[9:], atol=1e-12)


def test_oracle_matches_dense_numpy_on_hybrid_graph():
    g = tiny().hybrid_graph()
    ora, dg = OracleGraph.from_graph(g), ref.DenseGraph(g)
    assert abs(ora.error() - dg.error()) <= 1e-11 * dg.error()
    r = ora.optimize(max_iterations=4, relative_error_tol=0.0, absolute_error_tol=0.0)
    n = ref.dense_lm(g, max_iterations=4, relative_error_tol=0.0, absolute_error_tol=0.0)
    assert (r["iterations"], r["inner_iterations"]) == (n["iterations"], n["inner_iterations"])
    assert np.allclose(r["error_trace"], n["error_trace"], rtol=1e-6, atol=1e-14)


def test_hybrid_solution_recovers_the_motions():
    scene = small()
    g = scene.hybrid_graph()
    ora = OracleGraph.from_graph(g)
    r = ora.optimize()
    assert r["final_error"] < r["initial_error"]
    pose, _ = ora.values()
    keys, Rt = ps.hybrid_as_world_motion_values(g.pose_keys, pose)
    acc = d.metrics.evaluate(scene, keys, Rt)
    keys0, Rt0 = ps.hybrid_as_world_motion_values(g.pose_keys, g.pose_Rt)
    acc0 = d.metrics.evaluate(scene, keys0, Rt0)
    assert acc["n_object_motions"] > 50
    # without a smoothing factor every frame's motion is estimated from its own dozen points (sigma 6 cm, object radius 1 m:
    # about 0.02 rad per frame, times the lever arm to the world origin in the translation of a world-frame motion); the
    # initial values are the front-end's near-exact motions, so the noisy maximum-likelihood estimate is further from the
    # ground truth than they are. The bound is that noise level, not the initial one.
    assert acc["object_motion_error_m"] < 0.5 and acc["object_motion_error_rad"] < 0.08
    assert acc["camera_ate_m"] <= acc0["camera_ate_m"] + 1e-3
    key_H, frame_H = ps.hybrid_motions(g.pose_keys, pose)
    assert all(np.allclose(H[3], [0, 0, 0, 1]) for H in frame_H.values())


@pytest.mark.gpu
@pytest.mark.parametrize("which", ["small", "window", "medium"])
def test_gpu_matches_oracle_on_hybrid_graphs(which):
    if which == "small":
        g = small().hybrid_graph()
    elif which == "window":
        g = small(seed=3).hybrid_graph(5, 30, True)
    else:
        g = d.Scene.synthetic(n_frames=120, n_objects=4, static_per_frame=120, dynamic_per_object=25, seed=11).hybrid_graph()
    p = d.default_lm_params()
    p.max_iterations = 5
    p.relative_error_tol = 0.0
    p.absolute_error_tol = 0.0
    lm = d.GpuLevenbergMarquardt(g, p)
    ora = OracleGraph.from_graph(g)
    assert abs(lm.error() - ora.error()) <= 1e-12 * ora.error()
    a, b = lm.linearize(), ora.linearize()
    for k in ("p2p_A_pose", "p2p_A_point", "btw_A_a", "btw_A_b", "pri_A"):
        assert np.max(np.abs(a[k] - b[k])) <= 1e-10 * np.max(np.abs(b[k])), k
    ok_g, dp_g, dpt_g, _ = lm.solve_once(1e-4)
    ok_c, dp_c, dpt_c = ora.solve_once(1e-4)
    assert ok_g and ok_c
    assert np.max(np.abs(dp_g - dp_c)) <= 1e-7 * np.max(np.abs(dp_c)) and np.max(np.abs(dpt_g - dpt_c)) <= 1e-7 * np.max(np.abs(dpt_c))
    pose_g, pt_g = lm.optimize()
    r = ora.optimize(max_iterations=5, relative_error_tol=0.0, absolute_error_tol=0.0)
    assert (lm.iterations(), lm.getInnerIterations()) == (r["iterations"], r["inner_iterations"])
    assert np.allclose(lm.error_trace, r["error_trace"], rtol=1e-6, atol=0)
    pose_c, pt_c = ora.values()
    assert np.max(np.abs(pose_g - pose_c)) <= 1e-6 and np.max(np.abs(pt_g - pt_c)) <= 1e-5
    # one writer per block, no atomics: the same bits twice
    lm.reset_values()
    pose_2, pt_2 = lm.optimize()
    assert np.array_equal(pose_g, pose_2) and np.array_equal(pt_g, pt_2)
