"""LLWorld formulation factors: LandmarkMotionPoseFactor and LandmarkPoseSmoothingFactor.

CPU: the oracle restatement (oracle/llworld.hpp, central-difference Jacobians as gtsam::numericalDerivative forms them)
against known answers and an independent numpy statement (closed-form derivatives, scipy expm/logm).
GPU: the device kernels through the C ABI against the oracle, <= 1e-10 relative (north_star's per-factor bar).
"""
import numpy as np
import pytest

import dynosam_b200 as d
import np_ref
import oracle_lib


def _random_pose(rng, rot=0.6, trans=3.0):
    return np_ref.se3_exp(np.concatenate([rng.normal(0, rot, 3), rng.normal(0, trans, 3)]))


def _batch(n, seed):
    rng = np.random.default_rng(seed)
    Lp = [_random_pose(rng) for _ in range(n)]
    Lc = [Lp[i] @ np_ref.se3_exp(np.concatenate([rng.normal(0, 0.05, 3), rng.normal(0, 0.3, 3)])) for i in range(n)]
    Lk = [Lc[i] @ np_ref.se3_exp(np.concatenate([rng.normal(0, 0.05, 3), rng.normal(0, 0.3, 3)])) for i in range(n)]
    mp = rng.normal(0, 5.0, (n, 3))
    mc = np.array([(Lc[i] @ np.linalg.inv(Lp[i]) @ np.append(mp[i], 1.0))[:3] for i in range(n)]) + rng.normal(0, 0.05, (n, 3))
    to = lambda Ts: np.array([np_ref.to_Rt(T) for T in Ts])
    return mp, mc, Lp, Lc, Lk, to(Lp), to(Lc), to(Lk)


def _batch_residuals(mp, mc, Lp_a, Lc_a):
    out = np.zeros_like(mp)
    for i in range(len(mp)):
        Rp, tp, Rc, tc = Lp_a[i, :9].reshape(3, 3), Lp_a[i, 9:], Lc_a[i, :9].reshape(3, 3), Lc_a[i, 9:]
        out[i] = mc[i] - (Rc @ (Rp.T @ (mp[i] - tp)) + tc)
    return out


def _motion_residual(mp, mc, Lp, Lc):
    """LandmarkMotionPoseFactor::residual (LandmarkMotionPoseFactor.cc:82-89)."""
    return mc - (Lc @ np.linalg.inv(Lp) @ np.append(mp, 1.0))[:3]


def _smoothing_residual(k2, k1, k):
    """LandmarkPoseSmoothingFactor::residual (LandmarkPoseSmoothingFactor.cc:71-81)."""
    ha, hb = k1 @ np.linalg.inv(k2), k @ np.linalg.inv(k1)
    return np_ref.se3_log(np.linalg.inv(ha) @ hb)


def test_motion_pose_oracle_matches_closed_form():
    mp, mc, Lp, Lc, _, Lp_a, Lc_a, _ = _batch(24, 3)
    sigma = 0.05
    o = oracle_lib.llworld_motion_pose(mp, mc, Lp_a, Lc_a, sigma)
    for i in range(len(mp)):
        e = _motion_residual(mp[i], mc[i], Lp[i], Lc[i])
        np.testing.assert_allclose(o["b"][i], -e / sigma, rtol=1e-12, atol=1e-12)
        Rc, Rp = Lc[i][:3, :3], Lp[i][:3, :3]
        q = Rp.T @ (mp[i] - Lp[i][:3, 3])
        Q = np_ref.skew(q)
        np.testing.assert_allclose(o["A_cur_point"][i] * sigma, np.eye(3), atol=1e-9)
        np.testing.assert_allclose(o["A_prev_point"][i] * sigma, -Rc @ Rp.T, atol=1e-9)
        np.testing.assert_allclose(o["A_cur_pose"][i] * sigma, np.hstack([Rc @ Q, -Rc]), atol=2e-8)       # central differences: O(delta^2) of the exponential
        np.testing.assert_allclose(o["A_prev_pose"][i] * sigma, np.hstack([-Rc @ Q, Rc]), atol=2e-8)
    assert o["error"] == pytest.approx(0.5 * np.sum(o["b"] ** 2), rel=1e-12)


def test_motion_pose_known_answers():
    rng = np.random.default_rng(5)
    Lp, Lc = _random_pose(rng), _random_pose(rng)
    mp = rng.normal(0, 2, 3)
    mc = (Lc @ np.linalg.inv(Lp) @ np.append(mp, 1.0))[:3]      # the point moved exactly with the object: zero residual
    o = oracle_lib.llworld_motion_pose([mp], [mc], [np_ref.to_Rt(Lp)], [np_ref.to_Rt(Lc)], 0.1)
    np.testing.assert_allclose(o["b"], 0.0, atol=1e-13)
    assert o["error"] == pytest.approx(0.0, abs=1e-24)
    # Huber: a large residual is re-weighted by sqrt(k / dist) and costs k (dist - k / 2)
    off = np.array([3.0, 0, 0])
    o = oracle_lib.llworld_motion_pose([mp], [mc + off], [np_ref.to_Rt(Lp)], [np_ref.to_Rt(Lc)], 0.1, 0.5)
    dist = 30.0
    assert o["error"] == pytest.approx(0.5 * (dist - 0.25), rel=1e-12)
    np.testing.assert_allclose(np.linalg.norm(o["b"][0]), np.sqrt(0.5 / dist) * dist, rtol=1e-12)


def test_pose_smoothing_oracle_matches_independent_statement():
    _, _, Lp, Lc, Lk, Lp_a, Lc_a, Lk_a = _batch(10, 9)
    sig = np.array([0.01, 0.02, 0.03, 0.1, 0.2, 0.3])
    o = oracle_lib.llworld_pose_smoothing(Lp_a, Lc_a, Lk_a, sig)
    for i in range(len(Lp)):
        e = _smoothing_residual(Lp[i], Lc[i], Lk[i])
        np.testing.assert_allclose(o["b"][i], -e / sig, rtol=1e-9, atol=1e-9)
        for name, which in (("A_k2", 0), ("A_k1", 1), ("A_k", 2)):
            def f(T, which=which, i=i):
                xs = [Lp[i], Lc[i], Lk[i]]
                xs[which] = T
                return _smoothing_residual(*xs)
            J = np_ref.numeric_jacobian_pose(f, [Lp[i], Lc[i], Lk[i]][which], h=1e-5)
            np.testing.assert_allclose(o[name][i], J / sig[:, None], rtol=1e-6, atol=1e-5)
    assert o["error"] == pytest.approx(0.5 * np.sum(o["b"] ** 2), rel=1e-12)


def test_pose_smoothing_constant_motion_is_zero():
    rng = np.random.default_rng(2)
    L0 = _random_pose(rng)
    H = np_ref.se3_exp(np.array([0.02, -0.01, 0.03, 0.3, 0.1, -0.2]))
    L1, L2 = H @ L0, H @ H @ L0          # the same world-frame motion twice: k-2_H_k-1 == k-1_H_k
    o = oracle_lib.llworld_pose_smoothing([np_ref.to_Rt(L0)], [np_ref.to_Rt(L1)], [np_ref.to_Rt(L2)], np.ones(6))
    np.testing.assert_allclose(o["b"], 0.0, atol=1e-12)


@pytest.mark.gpu
@pytest.mark.parametrize("huber", [0.0, 0.8])
def test_gpu_motion_pose_matches_oracle(huber):
    mp, mc, _, _, _, Lp_a, Lc_a, _ = _batch(1000, 17)
    mc[::7] += 0.5                                         # some outliers for the robust branch
    ref = oracle_lib.llworld_motion_pose(mp, mc, Lp_a, Lc_a, 0.05, huber)
    got = d.llworld.motion_pose_linearize(mp, mc, Lp_a, Lc_a, 0.05, huber)
    assert np.abs(got["b"] - ref["b"]).max() <= 1e-10 * np.abs(ref["b"]).max()      # north_star: per-factor residuals <= 1e-10 relative
    assert got["error"] == pytest.approx(ref["error"], rel=1e-12)
    # Central-difference Jacobians: a residual evaluated in another operation order (FMA contraction on the device) differs
    # by a few ulp of the coordinates involved, and the difference quotient amplifies that by 1 / (2 delta): the bar is that
    # round-off bound -- the same uncertainty GTSAM's own numericalDerivative result carries -- not 1e-10 relative.
    coord = max(np.abs(mp).max(), np.abs(mc).max(), np.abs(Lp_a[:, 9:]).max(), np.abs(Lc_a[:, 9:]).max())
    roundoff = 16 * np.finfo(float).eps * coord / (2 * 1e-5) / 0.05
    for k in ("A_prev_point", "A_cur_point", "A_prev_pose", "A_cur_pose"):
        assert np.abs(got[k] - ref[k]).max() <= roundoff * (10.0 if k.endswith("pose") else 1.0), k
    # the closed-form mode: <= 1e-10 relative against the independent numpy derivatives, and within the truncation error of
    # the central differences
    ana = d.llworld.motion_pose_linearize(mp, mc, Lp_a, Lc_a, 0.05, huber, analytic=True)
    np.testing.assert_array_equal(ana["b"], got["b"])
    sw = np.linalg.norm(got["b"], axis=1) / np.maximum(np.linalg.norm(_batch_residuals(mp, mc, Lp_a, Lc_a), axis=1), 1e-300)   # sqrt(weight) / sigma per factor
    for i in range(0, len(mp), 37):
        Rp, Rc = Lp_a[i, :9].reshape(3, 3), Lc_a[i, :9].reshape(3, 3)
        q = Rp.T @ (mp[i] - Lp_a[i, 9:])
        Q = np_ref.skew(q)
        for k, J in (("A_cur_point", np.eye(3)), ("A_prev_point", -Rc @ Rp.T), ("A_cur_pose", np.hstack([Rc @ Q, -Rc])), ("A_prev_pose", np.hstack([-Rc @ Q, Rc]))):
            assert np.abs(ana[k][i] - sw[i] * J).max() <= 1e-10 * np.abs(sw[i] * J).max(), (k, i)
    for k in ("A_prev_point", "A_cur_point", "A_prev_pose", "A_cur_pose"):
        assert np.abs(ana[k] - ref[k]).max() <= 1e-6 * np.abs(ref[k]).max(), k


@pytest.mark.gpu
def test_gpu_pose_smoothing_matches_oracle():
    _, _, _, _, _, Lp_a, Lc_a, Lk_a = _batch(500, 23)
    sig = np.array([0.01, 0.02, 0.03, 0.1, 0.2, 0.3])
    ref = oracle_lib.llworld_pose_smoothing(Lp_a, Lc_a, Lk_a, sig)
    got = d.llworld.pose_smoothing_linearize(Lp_a, Lc_a, Lk_a, sig)
    assert np.abs(got["b"] - ref["b"]).max() <= 1e-10 * np.abs(ref["b"]).max()
    assert got["error"] == pytest.approx(ref["error"], rel=1e-12)
    # central differences of a residual that passes through compose / inverse / Logmap: round-off of the translations
    # (a few ulp of their size) amplified by 1 / (2 delta) and by the whitening (see the motion factor test)
    coord = max(np.abs(Lp_a[:, 9:]).max(), np.abs(Lk_a[:, 9:]).max())
    roundoff = 64 * np.finfo(float).eps * coord / (2 * 1e-5)
    for k in ("A_k2", "A_k1", "A_k"):
        assert np.abs((got[k] - ref[k]) * sig[None, :, None]).max() <= roundoff, k


@pytest.mark.gpu
def test_gpu_llworld_edge_cases():
    empty = np.zeros((0, 3)), np.zeros((0, 12))
    o = d.llworld.motion_pose_linearize(empty[0], empty[0], empty[1], empty[1], 0.1)
    assert o["b"].shape == (0, 3) and o["error"] == 0.0
    with pytest.raises(d.DynoError):
        d.llworld.motion_pose_linearize(np.zeros((1, 3)), np.zeros((1, 3)), np.zeros((1, 12)), np.zeros((1, 12)), 0.0)
