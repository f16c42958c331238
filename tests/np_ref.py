"""Independent numpy restatement of the SE(3) and factor arithmetic, used only to pin the C++ oracle.

Nothing here shares code with oracle/ or dynosam_b200/: rotations come from matrix exponentials / scipy,
Jacobians from central differences on x * Expmap(delta) (the way gtsam::numericalDerivative does it, which is what
the reference's own factor test uses: test/test_factors.cc:143-183)."""
import numpy as np
from scipy.linalg import expm, logm


def skew(w):
    return np.array([[0, -w[2], w[1]], [w[2], 0, -w[0]], [-w[1], w[0], 0]], dtype=float)


def se3_exp(xi):
    """Pose3::Expmap, tangent [omega, v]: matrix exponential of the 4x4 twist."""
    T = np.zeros((4, 4))
    T[:3, :3] = skew(xi[:3])
    T[:3, 3] = xi[3:]
    return expm(T)


def se3_log(T):
    L = np.real(logm(T))
    return np.array([L[2, 1], L[0, 2], L[1, 0], L[0, 3], L[1, 3], L[2, 3]])


def to_Rt(T):
    return np.concatenate([T[:3, :3].reshape(-1), T[:3, 3]])


def from_Rt(a):
    T = np.eye(4)
    T[:3, :3] = np.asarray(a[:9]).reshape(3, 3)
    T[:3, 3] = a[9:12]
    return T


def retract(T, d):
    return T @ se3_exp(np.asarray(d, dtype=float))


def numeric_jacobian_pose(f, T, h=1e-5):
    """d f(T * Exp(d)) / d d at 0 by central differences (gtsam::numericalDerivative, delta = 1e-5)."""
    cols = []
    for k in range(6):
        d = np.zeros(6)
        d[k] = h
        cols.append((f(retract(T, d)) - f(retract(T, -d))) / (2 * h))
    return np.stack(cols, axis=1)


def numeric_jacobian_point(f, p, h=1e-5):
    cols = []
    for k in range(3):
        d = np.zeros(3)
        d[k] = h
        cols.append((f(p + d) - f(p - d)) / (2 * h))
    return np.stack(cols, axis=1)


# ---- the four factor residuals (unwhitened), as the reference states them -------------------------------------------

def ternary_residual(m_prev, m_cur, H):
    """LandmarkMotionTernaryFactor::evaluateError (src/factors/LandmarkMotionTernaryFactor.cc:37-73):
    e = m_{k-1} - H^-1 m_k."""
    Hi = np.linalg.inv(H)
    return m_prev - (Hi[:3, :3] @ m_cur + Hi[:3, 3])


def p2p_residual(X, p, z):
    """gtsam::PoseToPointFactor<Pose3, Point3>: e = X^-1 p - z (typedef BackendDefinitions.hpp:54)."""
    Xi = np.linalg.inv(X)
    return Xi[:3, :3] @ p + Xi[:3, 3] - z


def between_residual(x1, x2, z):
    """gtsam::BetweenFactor<Pose3>: e = Logmap(z^-1 (x1^-1 x2))."""
    return se3_log(np.linalg.inv(z) @ np.linalg.inv(x1) @ x2)


def prior_residual(x, z):
    """gtsam::PriorFactor<Pose3>: e = Logmap(z^-1 x)."""
    return se3_log(np.linalg.inv(z) @ x)


def huber_loss(dist, k):
    """gtsam::noiseModel::mEstimator::Huber::loss on the whitened norm `dist`."""
    return 0.5 * dist * dist if dist <= k else k * (dist - 0.5 * k)


def huber_sqrt_weight(dist, k):
    return 1.0 if dist <= k else np.sqrt(k / dist)


# ---- a dense Levenberg-Marquardt over a flattened graph (dynosam_b200.Graph arrays) -----------------------------------
# Second, independent statement of the loop the reference runs (gtsam::LevenbergMarquardtOptimizer with default
# LevenbergMarquardtParams, RGBDBackendModule.cc:207-221): dense normal equations, numpy solve, Jacobians by central
# differences except for the prior factor (GTSAM 4.2.0 uses the identity there). Only for graphs of a few hundred factors.

class DenseGraph:
    def __init__(self, g):
        self.g = g
        self.poses = [from_Rt(r) for r in np.asarray(g.pose_Rt, dtype=float)]
        self.points = [np.array(p, dtype=float) for p in np.asarray(g.point_xyz, dtype=float)]
        self.n_pose, self.n_point = len(self.poses), len(self.points)
        self.dim = 6 * self.n_pose + 3 * self.n_point

    def _factors(self, poses, points):
        """yield (whitened residual, huber k, [(column offset, unwhitened-by-sigma Jacobian thunk)])."""
        g = self.g
        np6 = 6 * self.n_pose
        for k in range(g.n_p2p):
            i, j, z, s = int(g.p2p_pose[k]), int(g.p2p_point[k]), np.asarray(g.p2p_z[k], float), float(g.p2p_sigma[k])
            X, p = poses[i], points[j]
            yield (p2p_residual(X, p, z) / s, float(g.p2p_huber[k]),
                   [(6 * i, lambda X=X, p=p, z=z, s=s: numeric_jacobian_pose(lambda T: p2p_residual(T, p, z), X) / s),
                    (np6 + 3 * j, lambda X=X, p=p, z=z, s=s: numeric_jacobian_point(lambda q: p2p_residual(X, q, z), p) / s)])
        for k in range(g.n_ternary):
            a, b, h, s = int(g.ter_prev[k]), int(g.ter_cur[k]), int(g.ter_motion[k]), float(g.ter_sigma[k])
            mp, mc, H = points[a], points[b], poses[h]
            yield (ternary_residual(mp, mc, H) / s, float(g.ter_huber[k]),
                   [(np6 + 3 * a, lambda mp=mp, mc=mc, H=H, s=s: numeric_jacobian_point(lambda q: ternary_residual(q, mc, H), mp) / s),
                    (np6 + 3 * b, lambda mp=mp, mc=mc, H=H, s=s: numeric_jacobian_point(lambda q: ternary_residual(mp, q, H), mc) / s),
                    (6 * h, lambda mp=mp, mc=mc, H=H, s=s: numeric_jacobian_pose(lambda T: ternary_residual(mp, mc, T), H) / s)])
        for k in range(g.n_between):
            a, b, z, s = int(g.btw_a[k]), int(g.btw_b[k]), from_Rt(np.asarray(g.btw_z[k], float)), np.asarray(g.btw_sigmas[k], float)
            x1, x2 = poses[a], poses[b]
            yield (between_residual(x1, x2, z) / s, 0.0,
                   [(6 * a, lambda x1=x1, x2=x2, z=z, s=s: numeric_jacobian_pose(lambda T: between_residual(T, x2, z), x1, 1e-6) / s[:, None]),
                    (6 * b, lambda x1=x1, x2=x2, z=z, s=s: numeric_jacobian_pose(lambda T: between_residual(x1, T, z), x2, 1e-6) / s[:, None])])
        for k in range(g.n_prior):
            a, z, s = int(g.pri_a[k]), from_Rt(np.asarray(g.pri_z[k], float)), np.asarray(g.pri_sigmas[k], float)
            yield (prior_residual(poses[a], z) / s, 0.0, [(6 * a, lambda s=s: np.eye(6) / s[:, None])])

    def error(self, poses=None, points=None):
        poses = self.poses if poses is None else poses
        points = self.points if points is None else points
        total = 0.0
        for r, k, _ in self._factors(poses, points):
            dist = float(np.linalg.norm(r))
            total += huber_loss(dist, k) if k > 0 else 0.5 * dist * dist
        return total

    def linearize(self):
        rows = []
        for r, k, blocks in self._factors(self.poses, self.points):
            w = huber_sqrt_weight(float(np.linalg.norm(r)), k) if k > 0 else 1.0
            rows.append((-w * r, [(off, w * J()) for off, J in blocks]))
        m = sum(len(b) for b, _ in rows)
        A, bvec = np.zeros((m, self.dim)), np.zeros(m)
        at = 0
        for b, blocks in rows:
            n = len(b)
            bvec[at:at + n] = b
            for off, J in blocks:
                A[at:at + n, off:off + J.shape[1]] += J
            at += n
        return A, bvec

    def retracted(self, delta):
        poses = [retract(T, delta[6 * i:6 * i + 6]) for i, T in enumerate(self.poses)]
        base = 6 * self.n_pose
        points = [p + delta[base + 3 * j:base + 3 * j + 3] for j, p in enumerate(self.points)]
        return poses, points


def dense_lm(g, max_iterations=100, relative_error_tol=1e-5, absolute_error_tol=1e-5, error_tol=0.0, lambda_initial=1e-5,
             lambda_factor=10.0, lambda_upper=1e5, lambda_lower=0.0, min_model_fidelity=1e-3):
    dg = DenseGraph(g)
    lam, error = lambda_initial, dg.error()
    out = {"iterations": 0, "inner_iterations": 0, "error_trace": [error], "lambda_trace": []}
    new_error = error
    if error <= error_tol or max_iterations <= 0:
        out["graph"] = dg
        return out
    while True:
        current_error = new_error
        A, b = dg.linearize()
        AtA, Atb = A.T @ A, A.T @ b
        while True:   # tryLambda until a step is accepted or lambda runs out
            out["lambda_trace"].append(lam)
            ok, stop = False, False
            try:
                L = np.linalg.cholesky(AtA + lam * np.eye(dg.dim))
                delta = np.linalg.solve(L.T, np.linalg.solve(L, Atb))
            except np.linalg.LinAlgError:
                delta = None
            if delta is not None:
                old_lin = 0.5 * float(b @ b)
                res = A @ delta - b
                lin_change = old_lin - 0.5 * float(res @ res)
                if lin_change >= 0:
                    poses, points = dg.retracted(delta)
                    trial = dg.error(poses, points)
                    cost_change = error - trial
                    if lin_change > np.finfo(float).eps * old_lin:
                        ok = cost_change / lin_change > min_model_fidelity
                    if abs(cost_change) < relative_error_tol * error:
                        stop = True
            if ok:
                dg.poses, dg.points, error = poses, points, trial
                lam = max(lambda_lower, lam / lambda_factor)
                out["iterations"] += 1
                out["inner_iterations"] += 1
                break
            if stop:
                break
            lam *= lambda_factor
            out["inner_iterations"] += 1
            if lam >= lambda_upper:
                break
        new_error = error
        out["error_trace"].append(new_error)
        if new_error <= error_tol:
            break
        decrease = current_error - new_error
        converged = (relative_error_tol and decrease / current_error <= relative_error_tol) or decrease <= absolute_error_tol
        if out["iterations"] >= max_iterations or converged or not np.isfinite(current_error):
            break
    out["graph"] = dg
    out["final_error"] = error
    return out
