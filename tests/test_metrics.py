"""Post-solve metrics (reference: src/utils/Metrics.cc:34-55): known answers, and the oracle's optimum on seeded scenes."""
import numpy as np

import dynosam_b200 as d
import np_ref as ref
import scenes
from dynosam_b200 import metrics
from oracle_lib import OracleGraph


def test_pose_error_known_answers():
    I = ref.to_Rt(np.eye(4))
    T = ref.se3_exp(np.array([0.0, 0.0, 0.3, 1.0, 2.0, 2.0]))      # yaw of 0.3 rad
    t_e, r_e = metrics.pose_error([I], [ref.to_Rt(T)])
    assert abs(r_e[0] - 0.3) <= 1e-12 and abs(t_e[0] - np.linalg.norm(T[:3, 3])) <= 1e-12
    # error is that of E = M^-1 M_hat: invariant under a common left transform
    rng = np.random.default_rng(2)
    A, B, G = (ref.se3_exp(rng.normal(0, 0.5, 6)) for _ in range(3))
    e1 = metrics.pose_error([ref.to_Rt(A)], [ref.to_Rt(B)])
    e2 = metrics.pose_error([ref.to_Rt(G @ A)], [ref.to_Rt(G @ B)])
    assert np.allclose(e1, e2, atol=1e-12)
    E = np.linalg.inv(A) @ B
    assert abs(e1[0][0] - np.linalg.norm(E[:3, 3])) <= 1e-12
    assert abs(e1[1][0] - np.linalg.norm(ref.se3_log(E)[:3])) <= 1e-9
    # relative error of two identical increments is zero
    t_r, r_r = metrics.relative_pose_error([ref.to_Rt(A)], [ref.to_Rt(A @ G)], [ref.to_Rt(B)], [ref.to_Rt(B @ G)])
    assert t_r[0] <= 1e-12 and r_r[0] <= 1e-7


def test_noise_free_scene_is_recovered_exactly():
    scene = scenes.simulator_cfg1(20)
    g = scene.full_batch_graph()
    ora = OracleGraph.from_graph(g)
    ora.optimize()
    pose, _ = ora.values()
    m = metrics.evaluate(scene, g.pose_keys, pose)
    assert m["camera_ate_m"] <= 1e-6 and m["camera_ate_rad"] <= 1e-7
    assert m["object_motion_error_m"] <= 1e-5 and m["object_motion_error_rad"] <= 1e-6 and m["n_object_motions"] > 0


def test_optimisation_improves_a_noisy_scene():
    scene = scenes.small_noisy()
    g = scene.full_batch_graph()
    before = metrics.evaluate(scene, g.pose_keys, g.pose_Rt)
    ora = OracleGraph.from_graph(g)
    ora.optimize()
    pose, _ = ora.values()
    after = metrics.evaluate(scene, g.pose_keys, pose)
    assert after["camera_rpe_m"] < before["camera_rpe_m"]
    assert after["object_motion_error_m"] < before["object_motion_error_m"]    # motions start from identity
