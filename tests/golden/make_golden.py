"""Regenerates the golden fixtures in this directory:  python tests/golden/make_golden.py

Sources of the numbers (nothing here needs /root/reference or a GPU at test time):
  ternary_fixture.json  the inputs of the reference's test/test_factors.cc:143-203 (H = Rodrigues(-0.1, 0.2, 0.25),
                        t = (0.05, -0.10, 0.20), P1 = (0.4, 1.0, 0.8), P2 = H P1, H perturbed by a fixed tangent vector);
                        residual and Jacobians computed by tests/np_ref.py (numpy, central differences), NOT by the oracle.
  lm_traces.json        chi^2 traces, iteration counters and pose checksums of the CPU oracle on the seeded scenes the
                        parity tests use (BASELINE config 1 = the reference's simulator scenario, and noisy scenes).
                        GTSAM is not available, so these pin the oracle against regressions and give the GPU tests a
                        committed target; they are not GTSAM outputs (parity unpinned against GTSAM).
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import dynosam_b200 as d  # noqa: E402
import np_ref as ref  # noqa: E402
import scenes  # noqa: E402
from oracle_lib import OracleGraph  # noqa: E402
from scipy.spatial.transform import Rotation  # noqa: E402

PERTURB = [0.21, -0.13, 0.34, -0.08, 0.27, 0.11]   # stands in for utils::perturbWithNoise(H, 0.3)


def ternary_fixture():
    H = np.eye(4)
    H[:3, :3] = Rotation.from_rotvec([-0.1, 0.2, 0.25]).as_matrix()
    H[:3, 3] = [0.05, -0.10, 0.20]
    P1 = np.array([0.4, 1.0, 0.8])
    P2 = H[:3, :3] @ P1 + H[:3, 3]
    Hp = ref.retract(H, np.array(PERTURB))
    return {
        "H_Rt": ref.to_Rt(H).tolist(), "H_perturbed_Rt": ref.to_Rt(Hp).tolist(), "P1": P1.tolist(), "P2": P2.tolist(),
        "error_at_H": ref.ternary_residual(P1, P2, H).tolist(),
        "error_at_H_perturbed": ref.ternary_residual(P1, P2, Hp).tolist(),
        "J_prev": ref.numeric_jacobian_point(lambda p: ref.ternary_residual(p, P2, Hp), P1).tolist(),
        "J_cur": ref.numeric_jacobian_point(lambda p: ref.ternary_residual(P1, p, Hp), P2).tolist(),
        "J_motion": ref.numeric_jacobian_pose(lambda T: ref.ternary_residual(P1, P2, T), Hp).tolist(),
    }


def lm_trace(name, graph, budget):
    ora = OracleGraph.from_graph(graph)
    e0 = ora.error()
    r = ora.optimize(max_iterations=budget, relative_error_tol=0.0, absolute_error_tol=0.0)
    pose, pt = ora.values()
    return {"scene": name, "budget": budget, "n_pose": graph.n_pose, "n_point": graph.n_point, "n_factors": graph.n_factors,
            "initial_error": e0, "error_trace": r["error_trace"], "iterations": r["iterations"], "inner_iterations": r["inner_iterations"],
            "pose_translation_sum": float(np.sum(pose[:, 9:12])), "pose_abs_sum": float(np.sum(np.abs(pose))),
            "point_abs_sum": float(np.sum(np.abs(pt)))}


def main():
    json.dump(ternary_fixture(), open(os.path.join(HERE, "ternary_fixture.json"), "w"), indent=1)
    traces = [
        lm_trace("simulator_cfg1", scenes.simulator_cfg1().full_batch_graph(), 8),
        lm_trace("small_noisy", scenes.small_noisy().full_batch_graph(), 8),
        lm_trace("window", scenes.small_noisy(seed=3).window_graph(5, 30, True), 8),
        lm_trace("medium_noisy", scenes.medium_noisy().full_batch_graph(), 6),
    ]
    json.dump(traces, open(os.path.join(HERE, "lm_traces.json"), "w"), indent=1)
    print("wrote", os.listdir(HERE))


if __name__ == "__main__":
    main()
