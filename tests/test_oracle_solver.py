"""CPU tests: the oracle's linearisation, damped solve and LM loop against the dense numpy statement in np_ref.py.

The reference's tests pin no optimiser output ("LM result, chi^2, lambda sequence: parity unpinned"), and
GTSAM cannot be built here, so the oracle is cross-checked against a second implementation that shares no code with it:
central-difference Jacobians, dense normal equations, numpy Cholesky."""
import numpy as np
import pytest

import dynosam_b200 as d
import np_ref as ref
from oracle_lib import OracleGraph


def tiny_scene(seed=21):
    return d.Scene.synthetic(n_frames=7, n_objects=2, static_per_frame=10, dynamic_per_object=5, static_track_max=5,
                             dynamic_track_max=5, seed=seed)


@pytest.fixture(scope="module", params=["tiny_noisy", "simulator8", "window"])
def graph(request):
    if request.param == "tiny_noisy":
        return tiny_scene().full_batch_graph()
    if request.param == "simulator8":
        return d.Scene.simulator(8).full_batch_graph()
    return tiny_scene(22).window_graph(2, 6, True)


def oracle_normal_equations(g, lin):
    """A^T A and A^T b assembled from the oracle's per-factor whitened blocks (poses first, then points)."""
    n = 6 * g.n_pose + 3 * g.n_point
    H, grad = np.zeros((n, n)), np.zeros(n)
    np6 = 6 * g.n_pose

    def add(blocks, b):
        for oi, Ai in blocks:
            grad[oi:oi + Ai.shape[1]] += Ai.T @ b
            for oj, Aj in blocks:
                H[oi:oi + Ai.shape[1], oj:oj + Aj.shape[1]] += Ai.T @ Aj

    for k in range(g.n_p2p):
        add([(6 * int(g.p2p_pose[k]), lin["p2p_A_pose"][k].reshape(3, 6)), (np6 + 3 * int(g.p2p_point[k]), lin["p2p_A_point"][k].reshape(3, 3))],
            lin["p2p_b"][k])
    for k in range(g.n_ternary):
        add([(np6 + 3 * int(g.ter_prev[k]), lin["ter_A_prev"][k].reshape(3, 3)), (np6 + 3 * int(g.ter_cur[k]), lin["ter_A_cur"][k].reshape(3, 3)),
             (6 * int(g.ter_motion[k]), lin["ter_A_motion"][k].reshape(3, 6))], lin["ter_b"][k])
    for k in range(g.n_between):
        add([(6 * int(g.btw_a[k]), lin["btw_A_a"][k].reshape(6, 6)), (6 * int(g.btw_b[k]), lin["btw_A_b"][k].reshape(6, 6))], lin["btw_b"][k])
    for k in range(g.n_prior):
        add([(6 * int(g.pri_a[k]), lin["pri_A"][k].reshape(6, 6))], lin["pri_b"][k])
    return H, grad


def test_error_matches_dense(graph):
    ora, dg = OracleGraph.from_graph(graph), ref.DenseGraph(graph)
    e_o, e_d = ora.error(), dg.error()
    assert abs(e_o - e_d) <= 1e-11 * max(e_d, 1e-30)


def test_normal_equations_match_dense(graph):
    ora, dg = OracleGraph.from_graph(graph), ref.DenseGraph(graph)
    H_o, g_o = oracle_normal_equations(graph, ora.linearize())
    A, b = dg.linearize()
    H_d, g_d = A.T @ A, A.T @ b
    assert np.max(np.abs(H_o - H_d)) <= 1e-6 * np.max(np.abs(H_d))     # central differences: ~1e-8 relative per entry
    assert np.max(np.abs(g_o - g_d)) <= 1e-6 * max(np.max(np.abs(g_d)), 1e-30) + 1e-12


@pytest.mark.parametrize("lam", [1e-5, 1e-2, 10.0])
def test_damped_solve_matches_dense(graph, lam):
    ora = OracleGraph.from_graph(graph)
    H, grad = oracle_normal_equations(graph, ora.linearize())
    ok, dp, dpt = ora.solve_once(lam)
    assert ok
    want = np.linalg.solve(H + lam * np.eye(len(grad)), grad)
    got = np.concatenate([dp.reshape(-1), dpt.reshape(-1)])
    assert np.max(np.abs(got - want)) <= 1e-8 * max(np.max(np.abs(want)), 1e-30) + 1e-14


@pytest.mark.parametrize("budget", [2, 6])
def test_lm_matches_dense(graph, budget):
    ora = OracleGraph.from_graph(graph)
    r_o = ora.optimize(max_iterations=budget, relative_error_tol=0.0, absolute_error_tol=0.0)
    r_d = ref.dense_lm(graph, max_iterations=budget, relative_error_tol=0.0, absolute_error_tol=0.0)
    tr_o, tr_d = np.array(r_o["error_trace"]), np.array(r_d["error_trace"])
    floor = 1e-18   # noise-free simulator scene: both sit at round-off once converged
    if min(tr_o[-1], tr_d[-1]) > floor:
        assert r_o["iterations"] == r_d["iterations"]
        assert r_o["inner_iterations"] == r_d["inner_iterations"]
    n = min(len(tr_o), len(tr_d))
    assert np.all(np.abs(tr_o[:n] - tr_d[:n]) <= 1e-6 * np.abs(tr_d[:n]) + floor), (tr_o, tr_d)
    pose_o, pt_o = ora.values()
    pose_d = np.stack([ref.to_Rt(T) for T in r_d["graph"].poses])
    assert np.max(np.abs(pose_o - pose_d)) <= 1e-6
    if graph.n_point:
        assert np.max(np.abs(pt_o - np.stack(r_d["graph"].points))) <= 1e-5


def test_lm_default_parameters_are_gtsam_defaults():
    # LevenbergMarquardtParams defaults the reference relies on (it only sets verbosity: RGBDBackendModule.cc:207-214)
    p = d.default_lm_params()
    assert (p.lambda_initial, p.lambda_factor, p.lambda_upper_bound, p.lambda_lower_bound) == (1e-5, 10.0, 1e5, 0.0)
    assert (p.max_iterations, p.relative_error_tol, p.absolute_error_tol, p.error_tol) == (100, 1e-5, 1e-5, 0.0)
    assert p.min_model_fidelity == 1e-3


def test_lm_error_is_monotone_and_lambda_follows_the_schedule():
    g = tiny_scene(23).full_batch_graph()
    ora = OracleGraph.from_graph(g)
    r = ora.optimize()
    tr = np.array(r["error_trace"])
    assert np.all(np.diff(tr) <= 1e-12 * tr[:-1])   # an outer iteration never increases the error
    assert r["inner_iterations"] >= r["iterations"] >= 1
    # every accepted step divides lambda by 10, every rejected one multiplies it by 10
    rejected = r["inner_iterations"] - r["iterations"]
    assert r["lambda"] == pytest.approx(1e-5 * 10.0 ** (rejected - r["iterations"]), rel=1e-9)


def test_blocked_factorisation_matches_scalar():
    """The OpenMP-parallel blocked skyline Cholesky (the timed CPU arm) against the scalar row-by-row one: same
    damped step to round-off, same LM trajectory, on a scene with a real band."""
    import scenes
    g = scenes.medium_noisy().full_batch_graph()
    ora = OracleGraph.from_graph(g)
    for lam in (1e-5, 1.0):
        ok_b, dp_b, dpt_b = ora.solve_once(lam, blocked=True)
        ok_s, dp_s, dpt_s = ora.solve_once(lam, blocked=False)
        assert ok_b and ok_s
        assert np.max(np.abs(dp_b - dp_s)) <= 1e-9 * np.max(np.abs(dp_s))
        assert np.max(np.abs(dpt_b - dpt_s)) <= 1e-9 * np.max(np.abs(dpt_s))
    a = OracleGraph.from_graph(g).optimize(max_iterations=3, relative_error_tol=0.0, absolute_error_tol=0.0, blocked=True)
    b = OracleGraph.from_graph(g).optimize(max_iterations=3, relative_error_tol=0.0, absolute_error_tol=0.0, blocked=False)
    assert (a["iterations"], a["inner_iterations"]) == (b["iterations"], b["inner_iterations"])
    assert np.allclose(a["error_trace"], b["error_trace"], rtol=1e-9, atol=0)
    assert a["phase_s"]["factor_reduced"] > 0
