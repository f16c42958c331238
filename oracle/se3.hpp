// ORACLE (test infrastructure, not product code).
//
// CPU restatement of the SO(3)/SE(3) operations GTSAM 4.2.0 performs on the
// DynoSAM backend hot path. GTSAM is an un-vendored dependency of the reference
// (pinned tags/4.2.0 with GTSAM_POSE3_EXPMAP=ON, GTSAM_ROT3_EXPMAP=ON,
// reference docker/Dockerfile:79-90) and is absent from /root/reference, so the
// formulas below follow GTSAM's published algorithm (gtsam/geometry/SO3.cpp,
// Rot3M.cpp, Pose3.cpp at tag 4.2.0) and are anchored on the reference's call
// sites: Pose3::Expmap/Logmap (test/internal/simulator.hpp:203-205),
// Pose3::inverse / operator* (src/factors/LandmarkMotionTernaryFactor.cc:43),
// BetweenFactor<Pose3> (include/dynosam/backend/Formulation-impl.hpp:155,
// src/backend/rgbd/WorldMotionEstimator.cc:296), PriorFactor<Pose3>
// (Formulation-impl.hpp:100).  PARITY UNPINNED against GTSAM itself (no GTSAM
// build or golden output exists offline); pinned by the analytic-vs-numeric and
// sympy/numpy cross-checks in tests/.
//
// Conventions: tangent vectors are [omega(3), v(3)] (rotation first); a pose is
// stored as R (row-major 3x3) followed by t; retraction is x * Expmap(delta).
#pragma once
#include <cmath>
#include <cstring>
#include <limits>

namespace oracle {

struct Vec3 {
  double x, y, z;
  double& operator[](int i) { return (&x)[i]; }
  const double& operator[](int i) const { return (&x)[i]; }
};
inline Vec3 operator+(Vec3 a, Vec3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline Vec3 operator-(Vec3 a, Vec3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline Vec3 operator*(double s, Vec3 a) { return {s * a.x, s * a.y, s * a.z}; }
inline double dot(Vec3 a, Vec3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline Vec3 cross(Vec3 a, Vec3 b) {
  return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
inline double norm(Vec3 a) { return std::sqrt(dot(a, a)); }

struct Mat3 {
  double m[9];  // row-major
  double& operator()(int r, int c) { return m[3 * r + c]; }
  const double& operator()(int r, int c) const { return m[3 * r + c]; }
  static Mat3 Identity() { return {{1, 0, 0, 0, 1, 0, 0, 0, 1}}; }
  static Mat3 Zero() { return {{0, 0, 0, 0, 0, 0, 0, 0, 0}}; }
};
inline Mat3 operator*(const Mat3& a, const Mat3& b) {
  Mat3 c;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) {
      double s = 0;
      for (int k = 0; k < 3; ++k) s += a(i, k) * b(k, j);
      c(i, j) = s;
    }
  return c;
}
inline Vec3 operator*(const Mat3& a, Vec3 v) {
  return {a(0, 0) * v.x + a(0, 1) * v.y + a(0, 2) * v.z,
          a(1, 0) * v.x + a(1, 1) * v.y + a(1, 2) * v.z,
          a(2, 0) * v.x + a(2, 1) * v.y + a(2, 2) * v.z};
}
inline Mat3 operator+(const Mat3& a, const Mat3& b) {
  Mat3 c;
  for (int i = 0; i < 9; ++i) c.m[i] = a.m[i] + b.m[i];
  return c;
}
inline Mat3 operator-(const Mat3& a, const Mat3& b) {
  Mat3 c;
  for (int i = 0; i < 9; ++i) c.m[i] = a.m[i] - b.m[i];
  return c;
}
inline Mat3 operator*(double s, const Mat3& a) {
  Mat3 c;
  for (int i = 0; i < 9; ++i) c.m[i] = s * a.m[i];
  return c;
}
inline Mat3 transpose(const Mat3& a) {
  Mat3 c;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) c(i, j) = a(j, i);
  return c;
}
// skewSymmetric(w): [w]x
inline Mat3 skew(Vec3 w) { return {{0, -w.z, w.y, w.z, 0, -w.x, -w.y, w.x, 0}}; }

// ---- SO(3) -------------------------------------------------------------
// gtsam::so3::ExpmapFunctor (SO3.cpp, 4.2.0): nearZero when theta^2 <= eps.
inline Mat3 so3_expmap(Vec3 w) {
  const double theta2 = dot(w, w);
  const Mat3 W = skew(w);
  if (theta2 <= std::numeric_limits<double>::epsilon()) return Mat3::Identity() + W;
  const double theta = std::sqrt(theta2);
  const double sin_theta = std::sin(theta);
  const double s2 = std::sin(theta / 2.0);
  const double one_minus_cos = 2.0 * s2 * s2;
  const Mat3 K = (1.0 / theta) * W;
  const Mat3 KK = K * K;
  return Mat3::Identity() + sin_theta * K + one_minus_cos * KK;
}

// gtsam::SO3::Logmap (SO3.cpp, 4.2.0).
inline Vec3 so3_logmap(const Mat3& R) {
  const double R11 = R(0, 0), R12 = R(0, 1), R13 = R(0, 2);
  const double R21 = R(1, 0), R22 = R(1, 1), R23 = R(1, 2);
  const double R31 = R(2, 0), R32 = R(2, 1), R33 = R(2, 2);
  const double tr = R11 + R22 + R33;
  Vec3 omega;
  if (tr + 1.0 < 1e-3) {
    // theta near +-pi: use the largest diagonal entry (GTSAM 4.2.0 branch).
    if (R33 > R22 && R33 > R11) {
      const double W = R21 - R12, Q1 = 2.0 + 2.0 * R33, Q2 = R31 + R13, Q3 = R23 + R32;
      const double r = std::sqrt(Q1), one_over_r = 1 / r;
      const double nrm = std::sqrt(Q1 * Q1 + Q2 * Q2 + Q3 * Q3 + W * W);
      const double sgn_w = W < 0 ? -1.0 : 1.0;
      const double mag = M_PI - (2 * sgn_w * W) / nrm;
      const double scale = 0.5 * one_over_r * mag;
      omega = (sgn_w * scale) * Vec3{Q2, Q3, Q1};
    } else if (R22 > R11) {
      const double W = R13 - R31, Q1 = 2.0 + 2.0 * R22, Q2 = R23 + R32, Q3 = R12 + R21;
      const double r = std::sqrt(Q1), one_over_r = 1 / r;
      const double nrm = std::sqrt(Q1 * Q1 + Q2 * Q2 + Q3 * Q3 + W * W);
      const double sgn_w = W < 0 ? -1.0 : 1.0;
      const double mag = M_PI - (2 * sgn_w * W) / nrm;
      const double scale = 0.5 * one_over_r * mag;
      omega = (sgn_w * scale) * Vec3{Q3, Q1, Q2};
    } else {
      const double W = R32 - R23, Q1 = 2.0 + 2.0 * R11, Q2 = R12 + R21, Q3 = R31 + R13;
      const double r = std::sqrt(Q1), one_over_r = 1 / r;
      const double nrm = std::sqrt(Q1 * Q1 + Q2 * Q2 + Q3 * Q3 + W * W);
      const double sgn_w = W < 0 ? -1.0 : 1.0;
      const double mag = M_PI - (2 * sgn_w * W) / nrm;
      const double scale = 0.5 * one_over_r * mag;
      omega = (sgn_w * scale) * Vec3{Q1, Q2, Q3};
    }
  } else {
    double magnitude;
    const double tr_3 = tr - 3.0;
    if (tr_3 < -1e-6) {
      const double theta = std::acos((tr - 1.0) / 2.0);
      magnitude = theta / (2.0 * std::sin(theta));
    } else {
      // Taylor expansion near theta = 0 (borglab/gtsam issue 746).
      magnitude = 0.5 - tr_3 / 12.0 + tr_3 * tr_3 / 60.0;
    }
    omega = magnitude * Vec3{R32 - R23, R13 - R31, R21 - R12};
  }
  return omega;
}

// gtsam::SO3::LogmapDerivative (SO3.cpp, 4.2.0).
inline Mat3 so3_logmap_derivative(Vec3 w) {
  const double theta2 = dot(w, w);
  if (theta2 <= std::numeric_limits<double>::epsilon()) return Mat3::Identity();
  const double theta = std::sqrt(theta2);
  const Mat3 W = skew(w);
  const double c = 1.0 / (theta * theta) - (1.0 + std::cos(theta)) / (2.0 * theta * std::sin(theta));
  return Mat3::Identity() + 0.5 * W + c * (W * W);
}

// ---- SE(3) -------------------------------------------------------------
struct Pose3 {
  Mat3 R;
  Vec3 t;
  static Pose3 Identity() { return {Mat3::Identity(), {0, 0, 0}}; }
};
inline Pose3 compose(const Pose3& a, const Pose3& b) { return {a.R * b.R, a.R * b.t + a.t}; }
inline Pose3 inverse(const Pose3& a) {
  const Mat3 Rt = transpose(a.R);
  return {Rt, -1.0 * (Rt * a.t)};
}
inline Vec3 transform_from(const Pose3& a, Vec3 p) { return a.R * p + a.t; }
inline Vec3 transform_to(const Pose3& a, Vec3 p) { return transpose(a.R) * (p - a.t); }
inline Pose3 between(const Pose3& a, const Pose3& b) { return compose(inverse(a), b); }

// gtsam::Pose3::Expmap (Pose3.cpp, 4.2.0).
inline Pose3 se3_expmap(const double xi[6]) {
  const Vec3 w{xi[0], xi[1], xi[2]}, v{xi[3], xi[4], xi[5]};
  const Mat3 R = so3_expmap(w);
  const double theta2 = dot(w, w);
  if (theta2 > std::numeric_limits<double>::epsilon()) {
    const Vec3 t_parallel = dot(w, v) * w;
    const Vec3 w_cross_v = cross(w, v);
    const Vec3 t = (1.0 / theta2) * (w_cross_v - R * w_cross_v + t_parallel);
    return {R, t};
  }
  return {R, v};
}

// gtsam::Pose3::Logmap (Pose3.cpp, 4.2.0).
inline void se3_logmap(const Pose3& p, double xi[6]) {
  const Vec3 w = so3_logmap(p.R);
  const Vec3 T = p.t;
  const double t = norm(w);
  if (t < 1e-10) {
    xi[0] = w.x; xi[1] = w.y; xi[2] = w.z; xi[3] = T.x; xi[4] = T.y; xi[5] = T.z;
    return;
  }
  const Mat3 W = skew((1.0 / t) * w);
  const double Tan = std::tan(0.5 * t);
  const Vec3 WT = W * T;
  const Vec3 u = T - (0.5 * t) * WT + (1 - t / (2. * Tan)) * (W * WT);
  xi[0] = w.x; xi[1] = w.y; xi[2] = w.z; xi[3] = u.x; xi[4] = u.y; xi[5] = u.z;
}

struct Mat6 {
  double m[36];  // row-major
  double& operator()(int r, int c) { return m[6 * r + c]; }
  const double& operator()(int r, int c) const { return m[6 * r + c]; }
  static Mat6 Zero() { Mat6 a; std::memset(a.m, 0, sizeof a.m); return a; }
  static Mat6 Identity() { Mat6 a = Zero(); for (int i = 0; i < 6; ++i) a(i, i) = 1; return a; }
  void set_block(int r0, int c0, const Mat3& b) {
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) (*this)(r0 + i, c0 + j) = b(i, j);
  }
};
inline Mat6 operator*(const Mat6& a, const Mat6& b) {
  Mat6 c;
  for (int i = 0; i < 6; ++i)
    for (int j = 0; j < 6; ++j) {
      double s = 0;
      for (int k = 0; k < 6; ++k) s += a(i, k) * b(k, j);
      c(i, j) = s;
    }
  return c;
}

// gtsam::Pose3::AdjointMap: [R 0; [t]x R  R].
inline Mat6 adjoint_map(const Pose3& p) {
  Mat6 A = Mat6::Zero();
  A.set_block(0, 0, p.R);
  A.set_block(3, 0, skew(p.t) * p.R);
  A.set_block(3, 3, p.R);
  return A;
}

// gtsam::Pose3::ComputeQforExpmapDerivative (Pose3.cpp, 4.2.0; Barfoot eq. 102).
inline Mat3 compute_Q(const double xi[6], double near_zero = 1e-5) {
  const Vec3 w{xi[0], xi[1], xi[2]}, v{xi[3], xi[4], xi[5]};
  const Mat3 V = skew(v), W = skew(w);
  const double phi = norm(w);
  const Mat3 WVW = W * V * W;
  if (std::fabs(phi) > near_zero) {
    const double s = std::sin(phi), c = std::cos(phi);
    const double phi2 = phi * phi, phi3 = phi2 * phi, phi4 = phi3 * phi, phi5 = phi4 * phi;
    return -0.5 * V + ((phi - s) / phi3) * (W * V + V * W - WVW) +
           ((1 - phi2 / 2 - c) / phi4) * (W * W * V + V * W * W - 3.0 * WVW) -
           (0.5 * ((1 - phi2 / 2 - c) / phi4 - 3 * (phi - s - phi3 / 6.) / phi5)) *
               (WVW * W + W * WVW);
  }
  return -0.5 * V + (1. / 6.) * (W * V + V * W - WVW) -
         (1. / 24.) * (W * W * V + V * W * W - 3.0 * WVW) +
         (1. / 120.) * (WVW * W + W * WVW);
}

// gtsam::Pose3::LogmapDerivative (Pose3.cpp, 4.2.0).
inline Mat6 se3_logmap_derivative(const Pose3& p) {
  double xi[6];
  se3_logmap(p, xi);
  const Vec3 w{xi[0], xi[1], xi[2]};
  const Mat3 Jw = so3_logmap_derivative(w);
  const Mat3 Q = compute_Q(xi);
  const Mat3 Q2 = -1.0 * (Jw * Q * Jw);
  Mat6 J = Mat6::Zero();
  J.set_block(0, 0, Jw);
  J.set_block(3, 0, Q2);
  J.set_block(3, 3, Jw);
  return J;
}

// retract: x * Expmap(delta)  (GTSAM_POSE3_EXPMAP=ON).
inline Pose3 retract(const Pose3& x, const double d[6]) { return compose(x, se3_expmap(d)); }

}  // namespace oracle
