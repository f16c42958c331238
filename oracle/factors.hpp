// ORACLE (test infrastructure, not product code).
//
// CPU restatement of every factor on the DynoSAM backend LM path plus the noise
// model semantics GTSAM 4.2.0 applies to them.  Reference anchors:
//   F5 LandmarkMotionTernaryFactor  reference/dynosam/src/factors/LandmarkMotionTernaryFactor.cc:37-73
//   F3/F4 PoseToPointFactor         typedef include/dynosam/backend/BackendDefinitions.hpp:54,
//                                   emitted Formulation-impl.hpp:237-240,272-274 and
//                                   src/backend/rgbd/WorldMotionEstimator.cc:173-178,200-205
//                                   (arithmetic: gtsam_unstable/slam/PoseToPointFactor.h + Pose3::transformTo, 4.2.0)
//   F2/F6 BetweenFactor<Pose3>      Formulation-impl.hpp:155, WorldMotionEstimator.cc:296-301
//   F1 PriorFactor<Pose3>           Formulation-impl.hpp:100-101 (addPrior)
//   noise                           src/backend/BackendModule.cc:56-85, src/backend/RGBDBackendModule.cc:89-113
// GTSAM (4.2.0) source is not under /root/reference: PARITY UNPINNED for the
// GTSAM-owned factors; the ternary factor is pinned by the reference test
// test/test_factors.cc:143-203 (restated in tests/).
#pragma once
#include "se3.hpp"

namespace oracle {

// ---------------------------------------------------------------------------
// Noise: Diagonal/Isotropic sigmas, optionally wrapped in Robust(Huber(k)).
// GTSAM semantics: whitened r = e / sigma; Robust::WhitenSystem whitens then
// scales every block and b by sqrt(w), w = 1 if |r| <= k else k/|r|
// (mEstimator::Huber::weight, block reweighting); factor error =
// loss(|r|): 0.5 |r|^2 (Gaussian) or Huber loss.
// ---------------------------------------------------------------------------
inline double huber_weight(double dist, double k) { return dist <= k ? 1.0 : k / dist; }
inline double huber_loss(double dist, double k) {
  const double a = std::fabs(dist);
  return a <= k ? 0.5 * dist * dist : k * (a - 0.5 * k);
}
// loss of a whitened residual norm under (optional) Huber, huber_k <= 0: Gaussian
inline double robust_loss(double dist, double huber_k) {
  return huber_k > 0 ? huber_loss(dist, huber_k) : 0.5 * dist * dist;
}
inline double robust_sqrt_weight(double dist, double huber_k) {
  return huber_k > 0 ? std::sqrt(huber_weight(dist, huber_k)) : 1.0;
}

// ---------------------------------------------------------------------------
// PoseToPointFactor: e = X.transformTo(p) - z ; H_X = [ [q]x | -I ], H_p = R^T.
// ---------------------------------------------------------------------------
inline Vec3 pose_to_point_error(const Pose3& X, Vec3 p, Vec3 z, double* HX /*3x6 row-major*/,
                                double* Hp /*3x3*/) {
  const Mat3 Rt = transpose(X.R);
  const Vec3 q = Rt * (p - X.t);
  if (HX) {
    const double wx = q.x, wy = q.y, wz = q.z;
    const double J[18] = {0, -wz, +wy, -1, 0, 0,  //
                          +wz, 0, -wx, 0, -1, 0,  //
                          -wy, +wx, 0, 0, 0, -1};
    std::memcpy(HX, J, sizeof J);
  }
  if (Hp) std::memcpy(Hp, Rt.m, sizeof Rt.m);
  return q - z;
}

// ---------------------------------------------------------------------------
// LandmarkMotionTernaryFactor (reference cc:37-73): e = m_prev - H^-1 m_cur.
// J1 = I, J2 = -R_H^T, J3 = [ -[q]x | I ] with q = H^-1 m_cur.
// ---------------------------------------------------------------------------
inline Vec3 motion_ternary_error(Vec3 m_prev, Vec3 m_cur, const Pose3& H, double* J1, double* J2,
                                 double* J3 /*3x6*/) {
  const Pose3 Hinv = inverse(H);
  const Vec3 q = transform_from(Hinv, m_cur);
  if (J1) { const Mat3 I = Mat3::Identity(); std::memcpy(J1, I.m, sizeof I.m); }
  if (J2) { const Mat3 M = -1.0 * Hinv.R; std::memcpy(J2, M.m, sizeof M.m); }
  if (J3) {
    double J[18];
    for (double& v : J) v = 0;
    J[0 * 6 + 3] = 1; J[1 * 6 + 4] = 1; J[2 * 6 + 5] = 1;
    J[0 * 6 + 1] = q.z;  J[0 * 6 + 2] = -q.y;
    J[1 * 6 + 0] = -q.z; J[1 * 6 + 2] = q.x;
    J[2 * 6 + 0] = q.y;  J[2 * 6 + 1] = -q.x;
    std::memcpy(J3, J, sizeof J);
  }
  return m_prev - q;
}

// ---------------------------------------------------------------------------
// BetweenFactor<Pose3>: hx = x1^-1 x2 ; e = Logmap(z^-1 hx);
// H1 = LogmapDerivative(z^-1 hx) * (-Adjoint(hx^-1)), H2 = LogmapDerivative(z^-1 hx).
// ---------------------------------------------------------------------------
inline void between_error(const Pose3& x1, const Pose3& x2, const Pose3& z, double e[6], Mat6* H1,
                          Mat6* H2) {
  const Pose3 hx = between(x1, x2);
  const Pose3 d = between(z, hx);
  se3_logmap(d, e);
  if (H1 || H2) {
    const Mat6 Hlocal = se3_logmap_derivative(d);
    if (H1) {
      Mat6 A = adjoint_map(inverse(hx));
      for (double& v : A.m) v = -v;
      *H1 = Hlocal * A;
    }
    if (H2) *H2 = Hlocal;
  }
}

// ---------------------------------------------------------------------------
// PriorFactor<Pose3> (gtsam/nonlinear/PriorFactor.h, 4.2.0):
// e = -Local(x, prior) = -Logmap(x^-1 prior); H = Identity ("TODO(ASL) Add Jacobians").
// ---------------------------------------------------------------------------
inline void prior_error(const Pose3& x, const Pose3& prior, double e[6], Mat6* H) {
  double l[6];
  se3_logmap(between(x, prior), l);
  for (int i = 0; i < 6; ++i) e[i] = -l[i];
  if (H) *H = Mat6::Identity();
}

}  // namespace oracle
