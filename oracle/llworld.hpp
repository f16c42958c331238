// ORACLE (test infrastructure, not product code).
// LLWorld formulation factors, restated from the reference:
//   LandmarkMotionPoseFactor      dynosam/src/factors/LandmarkMotionPoseFactor.cc:32-89
//   LandmarkPoseSmoothingFactor   dynosam/src/factors/LandmarkPoseSmoothingFactor.cc:29-83
// Both take their Jacobians from gtsam::numericalDerivative4x / 3x (GTSAM 4.2.0, base/numericalDerivative.h: central
// differences, delta = 1e-5, x perturbed through traits<X>::Retract -- vector addition for Point3, x * Expmap(dx) for
// Pose3 with GTSAM_POSE3_EXPMAP -- and the result differenced in the vector space of the residual). GTSAM itself is not
// in this image: parity unpinned against it; pinned here by the analytic derivatives (tests/test_llworld_factors.py).
#pragma once
#include "factors.hpp"

namespace oracle {

constexpr double kNumericalDelta = 1e-5;   // gtsam::numericalDerivativeXY default

// residual (cc:82-89): m_k - (L_k L_{k-1}^-1) m_{k-1}
inline Vec3 landmark_motion_pose_residual(Vec3 m_prev, Vec3 m_cur, const Pose3& L_prev, const Pose3& L_cur) {
  return m_cur - transform_from(compose(L_cur, inverse(L_prev)), m_prev);
}

// evaluateError (cc:32-79). J1, J2: 3x3, J3, J4: 3x6, row-major; any may be null.
inline Vec3 landmark_motion_pose_error(Vec3 m_prev, Vec3 m_cur, const Pose3& L_prev, const Pose3& L_cur, double* J1, double* J2, double* J3, double* J4) {
  const double f = 1.0 / (2.0 * kNumericalDelta);
  auto col3 = [&](double* J, int ncol, int j, Vec3 plus, Vec3 minus) {
    const Vec3 d = f * (plus - minus);
    J[0 * ncol + j] = d.x; J[1 * ncol + j] = d.y; J[2 * ncol + j] = d.z;
  };
  auto bump = [](Vec3 p, int j, double d) { if (j == 0) p.x += d; else if (j == 1) p.y += d; else p.z += d; return p; };
  for (int j = 0; j < 3; ++j) {
    if (J1) col3(J1, 3, j, landmark_motion_pose_residual(bump(m_prev, j, kNumericalDelta), m_cur, L_prev, L_cur),
                 landmark_motion_pose_residual(bump(m_prev, j, -kNumericalDelta), m_cur, L_prev, L_cur));
    if (J2) col3(J2, 3, j, landmark_motion_pose_residual(m_prev, bump(m_cur, j, kNumericalDelta), L_prev, L_cur),
                 landmark_motion_pose_residual(m_prev, bump(m_cur, j, -kNumericalDelta), L_prev, L_cur));
  }
  for (int j = 0; j < 6; ++j) {
    double dp[6] = {0, 0, 0, 0, 0, 0}, dm[6] = {0, 0, 0, 0, 0, 0};
    dp[j] = kNumericalDelta; dm[j] = -kNumericalDelta;
    if (J3) col3(J3, 6, j, landmark_motion_pose_residual(m_prev, m_cur, retract(L_prev, dp), L_cur), landmark_motion_pose_residual(m_prev, m_cur, retract(L_prev, dm), L_cur));
    if (J4) col3(J4, 6, j, landmark_motion_pose_residual(m_prev, m_cur, L_prev, retract(L_cur, dp)), landmark_motion_pose_residual(m_prev, m_cur, L_prev, retract(L_cur, dm)));
  }
  return landmark_motion_pose_residual(m_prev, m_cur, L_prev, L_cur);
}

// residual (cc:71-81): Local(Identity, Between(L_{k-1} L_{k-2}^-1, L_k L_{k-1}^-1)) = Logmap((k-2_H_k-1)^-1 (k-1_H_k))
inline void landmark_pose_smoothing_residual(const Pose3& k2, const Pose3& k1, const Pose3& k, double e[6]) {
  const Pose3 h_a = compose(k1, inverse(k2));
  const Pose3 h_b = compose(k, inverse(k1));
  se3_logmap(between(h_a, h_b), e);
}

// evaluateError (cc:29-68). J1..J3: 6x6 row-major; any may be null.
inline void landmark_pose_smoothing_error(const Pose3& k2, const Pose3& k1, const Pose3& k, double e[6], double* J1, double* J2, double* J3) {
  const double f = 1.0 / (2.0 * kNumericalDelta);
  double* Js[3] = {J1, J2, J3};
  for (int which = 0; which < 3; ++which) {
    if (!Js[which]) continue;
    for (int j = 0; j < 6; ++j) {
      double dp[6] = {0, 0, 0, 0, 0, 0}, dm[6] = {0, 0, 0, 0, 0, 0};
      dp[j] = kNumericalDelta; dm[j] = -kNumericalDelta;
      double ep[6], em[6];
      landmark_pose_smoothing_residual(which == 0 ? retract(k2, dp) : k2, which == 1 ? retract(k1, dp) : k1, which == 2 ? retract(k, dp) : k, ep);
      landmark_pose_smoothing_residual(which == 0 ? retract(k2, dm) : k2, which == 1 ? retract(k1, dm) : k1, which == 2 ? retract(k, dm) : k, em);
      for (int r = 0; r < 6; ++r) Js[which][6 * r + j] = f * (ep[r] - em[r]);
    }
  }
  landmark_pose_smoothing_residual(k2, k1, k, e);
}

}  // namespace oracle
