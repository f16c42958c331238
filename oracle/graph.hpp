// ORACLE (test infrastructure, not product code).
//
// Flat factor-graph container consumed by the CPU oracle, and the two whole
// graph passes GTSAM runs on the LM path: NonlinearFactorGraph::error(values)
// (reference call sites src/backend/RGBDBackendModule.cc:216,222,363,383) and
// NonlinearFactorGraph::linearize(values) (inside
// LevenbergMarquardtOptimizer::iterate, reference RGBDBackendModule.cc:220-221).
// Variables are addressed by dense index: poses (H.., L.., X.. in ascending key
// order) then points (l.., m..), i.e. gtsam::Values' natural order.
#pragma once
#include <cstdint>
#include <vector>

#include "factors.hpp"

namespace oracle {

struct FlatGraph {
  // variables
  std::vector<Pose3> poses;
  std::vector<Vec3> points;
  // F3/F4 PoseToPoint
  std::vector<int32_t> p2p_pose, p2p_point;
  std::vector<Vec3> p2p_z;
  std::vector<double> p2p_sigma, p2p_huber;
  // F5 LandmarkMotionTernary (m_prev, m_cur, H)
  std::vector<int32_t> ter_prev, ter_cur, ter_motion;
  std::vector<double> ter_sigma, ter_huber;
  // F2/F6 Between
  std::vector<int32_t> btw_a, btw_b;
  std::vector<Pose3> btw_z;
  std::vector<double> btw_sigmas;  // 6 per factor
  // F1 Prior
  std::vector<int32_t> pri_a;
  std::vector<Pose3> pri_z;
  std::vector<double> pri_sigmas;  // 6 per factor

  size_t n_factors() const {
    return p2p_pose.size() + ter_prev.size() + btw_a.size() + pri_a.size();
  }
};

// One linearised factor, GTSAM JacobianFactor style: whitened and robust
// reweighted blocks A_j (dim x var_dim, row-major) and b = -(whitened residual).
struct LinFactor {
  int nvar;        // 1..3
  int dim;         // 3 or 6
  int var[3];      // global variable id: pose i -> i ; point j -> n_pose + j
  int vdim[3];     // 6 or 3
  double A[3][36]; // row-major dim x vdim
  double b[6];
};

inline double graph_error(const FlatGraph& g, const std::vector<Pose3>& poses,
                          const std::vector<Vec3>& points) {
  double total = 0;
  const int64_t n1 = (int64_t)g.p2p_pose.size();
#pragma omp parallel for reduction(+ : total) schedule(static)
  for (int64_t f = 0; f < n1; ++f) {
    const Vec3 e = pose_to_point_error(poses[g.p2p_pose[f]], points[g.p2p_point[f]], g.p2p_z[f], nullptr, nullptr);
    total += robust_loss(norm(e) / g.p2p_sigma[f], g.p2p_huber[f]);
  }
  const int64_t n2 = (int64_t)g.ter_prev.size();
#pragma omp parallel for reduction(+ : total) schedule(static)
  for (int64_t f = 0; f < n2; ++f) {
    const Vec3 e = motion_ternary_error(points[g.ter_prev[f]], points[g.ter_cur[f]], poses[g.ter_motion[f]], nullptr, nullptr, nullptr);
    total += robust_loss(norm(e) / g.ter_sigma[f], g.ter_huber[f]);
  }
  for (size_t f = 0; f < g.btw_a.size(); ++f) {
    double e[6], s = 0;
    between_error(poses[g.btw_a[f]], poses[g.btw_b[f]], g.btw_z[f], e, nullptr, nullptr);
    for (int i = 0; i < 6; ++i) { const double r = e[i] / g.btw_sigmas[6 * f + i]; s += r * r; }
    total += 0.5 * s;
  }
  for (size_t f = 0; f < g.pri_a.size(); ++f) {
    double e[6], s = 0;
    prior_error(poses[g.pri_a[f]], g.pri_z[f], e, nullptr);
    for (int i = 0; i < 6; ++i) { const double r = e[i] / g.pri_sigmas[6 * f + i]; s += r * r; }
    total += 0.5 * s;
  }
  return total;
}

// Factor order in the returned vector: p2p, ternary, between, prior (type-major;
// the per-type order is the flat graph's order).
inline void graph_linearize(const FlatGraph& g, const std::vector<Pose3>& poses,
                            const std::vector<Vec3>& points, std::vector<LinFactor>& out) {
  const int np = (int)poses.size();
  const int64_t n1 = (int64_t)g.p2p_pose.size(), n2 = (int64_t)g.ter_prev.size();
  const int64_t n3 = (int64_t)g.btw_a.size(), n4 = (int64_t)g.pri_a.size();
  out.resize(n1 + n2 + n3 + n4);
#pragma omp parallel for schedule(static)
  for (int64_t f = 0; f < n1; ++f) {
    LinFactor& L = out[f];
    L.nvar = 2; L.dim = 3;
    L.var[0] = g.p2p_pose[f]; L.vdim[0] = 6;
    L.var[1] = np + g.p2p_point[f]; L.vdim[1] = 3;
    const Vec3 e = pose_to_point_error(poses[g.p2p_pose[f]], points[g.p2p_point[f]], g.p2p_z[f], L.A[0], L.A[1]);
    const double inv_sigma = 1.0 / g.p2p_sigma[f];
    const double dist = norm(e) * inv_sigma;  // |whitened b|
    const double s = inv_sigma * robust_sqrt_weight(dist, g.p2p_huber[f]);
    for (int i = 0; i < 18; ++i) L.A[0][i] *= s;
    for (int i = 0; i < 9; ++i) L.A[1][i] *= s;
    for (int i = 0; i < 3; ++i) L.b[i] = -e[i] * s;
  }
#pragma omp parallel for schedule(static)
  for (int64_t f = 0; f < n2; ++f) {
    LinFactor& L = out[n1 + f];
    L.nvar = 3; L.dim = 3;
    L.var[0] = np + g.ter_prev[f]; L.vdim[0] = 3;
    L.var[1] = np + g.ter_cur[f]; L.vdim[1] = 3;
    L.var[2] = g.ter_motion[f]; L.vdim[2] = 6;
    const Vec3 e = motion_ternary_error(points[g.ter_prev[f]], points[g.ter_cur[f]], poses[g.ter_motion[f]], L.A[0], L.A[1], L.A[2]);
    const double inv_sigma = 1.0 / g.ter_sigma[f];
    const double dist = norm(e) * inv_sigma;
    const double s = inv_sigma * robust_sqrt_weight(dist, g.ter_huber[f]);
    for (int i = 0; i < 9; ++i) { L.A[0][i] *= s; L.A[1][i] *= s; }
    for (int i = 0; i < 18; ++i) L.A[2][i] *= s;
    for (int i = 0; i < 3; ++i) L.b[i] = -e[i] * s;
  }
  for (int64_t f = 0; f < n3; ++f) {
    LinFactor& L = out[n1 + n2 + f];
    L.nvar = 2; L.dim = 6;
    L.var[0] = g.btw_a[f]; L.var[1] = g.btw_b[f]; L.vdim[0] = L.vdim[1] = 6;
    double e[6]; Mat6 H1, H2;
    between_error(poses[g.btw_a[f]], poses[g.btw_b[f]], g.btw_z[f], e, &H1, &H2);
    for (int r = 0; r < 6; ++r) {
      const double inv_sigma = 1.0 / g.btw_sigmas[6 * f + r];
      for (int c = 0; c < 6; ++c) { L.A[0][6 * r + c] = H1(r, c) * inv_sigma; L.A[1][6 * r + c] = H2(r, c) * inv_sigma; }
      L.b[r] = -e[r] * inv_sigma;
    }
  }
  for (int64_t f = 0; f < n4; ++f) {
    LinFactor& L = out[n1 + n2 + n3 + f];
    L.nvar = 1; L.dim = 6; L.var[0] = g.pri_a[f]; L.vdim[0] = 6;
    double e[6]; Mat6 H;
    prior_error(poses[g.pri_a[f]], g.pri_z[f], e, &H);
    for (int r = 0; r < 6; ++r) {
      const double inv_sigma = 1.0 / g.pri_sigmas[6 * f + r];
      for (int c = 0; c < 6; ++c) L.A[0][6 * r + c] = H(r, c) * inv_sigma;
      L.b[r] = -e[r] * inv_sigma;
    }
  }
}

}  // namespace oracle
