// ORACLE (test infrastructure, not product code).
// C entry points so tests/ (ctypes) and bench.py's cpu_baseline / --impl
// reference leg can drive the CPU restatement.  Nothing under dynosam_b200/
// may link or call this library.
#include <cstdint>
#include <cstring>

#include "llworld.hpp"
#include "lm.hpp"

using namespace oracle;

namespace {
Pose3 pose_from(const double* p) {
  Pose3 x;
  std::memcpy(x.R.m, p, 9 * sizeof(double));
  x.t = {p[9], p[10], p[11]};
  return x;
}
void pose_to(const Pose3& x, double* p) {
  std::memcpy(p, x.R.m, 9 * sizeof(double));
  p[9] = x.t.x; p[10] = x.t.y; p[11] = x.t.z;
}
}  // namespace

extern "C" {

struct oracle_graph { FlatGraph g; };

oracle_graph* oracle_graph_create(int32_t n_pose, const double* pose_Rt, int32_t n_point, const double* point_xyz) {
  oracle_graph* h = new oracle_graph();
  h->g.poses.resize(n_pose); h->g.points.resize(n_point);
  for (int i = 0; i < n_pose; ++i) h->g.poses[i] = pose_from(pose_Rt + 12 * (size_t)i);
  for (int i = 0; i < n_point; ++i) h->g.points[i] = {point_xyz[3 * (size_t)i], point_xyz[3 * (size_t)i + 1], point_xyz[3 * (size_t)i + 2]};
  return h;
}
void oracle_graph_destroy(oracle_graph* h) { delete h; }

void oracle_graph_add_p2p(oracle_graph* h, int64_t n, const int32_t* pose_idx, const int32_t* point_idx, const double* z, double sigma, double huber_k) {
  for (int64_t i = 0; i < n; ++i) {
    h->g.p2p_pose.push_back(pose_idx[i]); h->g.p2p_point.push_back(point_idx[i]);
    h->g.p2p_z.push_back({z[3 * i], z[3 * i + 1], z[3 * i + 2]});
    h->g.p2p_sigma.push_back(sigma); h->g.p2p_huber.push_back(huber_k);
  }
}
void oracle_graph_add_ternary(oracle_graph* h, int64_t n, const int32_t* prev, const int32_t* cur, const int32_t* motion, double sigma, double huber_k) {
  for (int64_t i = 0; i < n; ++i) {
    h->g.ter_prev.push_back(prev[i]); h->g.ter_cur.push_back(cur[i]); h->g.ter_motion.push_back(motion[i]);
    h->g.ter_sigma.push_back(sigma); h->g.ter_huber.push_back(huber_k);
  }
}
void oracle_graph_add_between(oracle_graph* h, int64_t n, const int32_t* a, const int32_t* b, const double* z_Rt, const double* sigmas6) {
  for (int64_t i = 0; i < n; ++i) {
    h->g.btw_a.push_back(a[i]); h->g.btw_b.push_back(b[i]); h->g.btw_z.push_back(pose_from(z_Rt + 12 * i));
    for (int k = 0; k < 6; ++k) h->g.btw_sigmas.push_back(sigmas6[k]);
  }
}
void oracle_graph_add_prior(oracle_graph* h, int64_t n, const int32_t* a, const double* z_Rt, const double* sigmas6) {
  for (int64_t i = 0; i < n; ++i) {
    h->g.pri_a.push_back(a[i]); h->g.pri_z.push_back(pose_from(z_Rt + 12 * i));
    for (int k = 0; k < 6; ++k) h->g.pri_sigmas.push_back(sigmas6[k]);
  }
}

double oracle_graph_error(oracle_graph* h) { return graph_error(h->g, h->g.poses, h->g.points); }

void oracle_graph_set(oracle_graph* h, const double* pose_Rt, const double* point_xyz) {
  for (size_t i = 0; i < h->g.poses.size(); ++i) h->g.poses[i] = pose_from(pose_Rt + 12 * i);
  for (size_t i = 0; i < h->g.points.size(); ++i) h->g.points[i] = {point_xyz[3 * i], point_xyz[3 * i + 1], point_xyz[3 * i + 2]};
}
void oracle_graph_get(oracle_graph* h, double* pose_Rt, double* point_xyz) {
  for (size_t i = 0; i < h->g.poses.size(); ++i) pose_to(h->g.poses[i], pose_Rt + 12 * i);
  for (size_t i = 0; i < h->g.points.size(); ++i) { point_xyz[3 * i] = h->g.points[i].x; point_xyz[3 * i + 1] = h->g.points[i].y; point_xyz[3 * i + 2] = h->g.points[i].z; }
}

// Per-factor whitened, robust-reweighted Jacobian blocks and b (GTSAM JacobianFactor content).
// p2p: A_pose n x 18, A_point n x 9, b n x 3 ; ternary: A_prev, A_cur n x 9, A_motion n x 18, b n x 3 ;
// between: A_a, A_b n x 36, b n x 6 ; prior: A n x 36, b n x 6.  Any pointer may be null.
void oracle_graph_linearize(oracle_graph* h, double* p2p_Apose, double* p2p_Apoint, double* p2p_b,
                            double* ter_Aprev, double* ter_Acur, double* ter_Amotion, double* ter_b,
                            double* btw_Aa, double* btw_Ab, double* btw_b, double* pri_A, double* pri_b) {
  std::vector<LinFactor> lin;
  graph_linearize(h->g, h->g.poses, h->g.points, lin);
  const size_t n1 = h->g.p2p_pose.size(), n2 = h->g.ter_prev.size(), n3 = h->g.btw_a.size(), n4 = h->g.pri_a.size();
  for (size_t f = 0; f < n1; ++f) {
    const LinFactor& L = lin[f];
    if (p2p_Apose) std::memcpy(p2p_Apose + 18 * f, L.A[0], 18 * sizeof(double));
    if (p2p_Apoint) std::memcpy(p2p_Apoint + 9 * f, L.A[1], 9 * sizeof(double));
    if (p2p_b) std::memcpy(p2p_b + 3 * f, L.b, 3 * sizeof(double));
  }
  for (size_t f = 0; f < n2; ++f) {
    const LinFactor& L = lin[n1 + f];
    if (ter_Aprev) std::memcpy(ter_Aprev + 9 * f, L.A[0], 9 * sizeof(double));
    if (ter_Acur) std::memcpy(ter_Acur + 9 * f, L.A[1], 9 * sizeof(double));
    if (ter_Amotion) std::memcpy(ter_Amotion + 18 * f, L.A[2], 18 * sizeof(double));
    if (ter_b) std::memcpy(ter_b + 3 * f, L.b, 3 * sizeof(double));
  }
  for (size_t f = 0; f < n3; ++f) {
    const LinFactor& L = lin[n1 + n2 + f];
    if (btw_Aa) std::memcpy(btw_Aa + 36 * f, L.A[0], 36 * sizeof(double));
    if (btw_Ab) std::memcpy(btw_Ab + 36 * f, L.A[1], 36 * sizeof(double));
    if (btw_b) std::memcpy(btw_b + 6 * f, L.b, 6 * sizeof(double));
  }
  for (size_t f = 0; f < n4; ++f) {
    const LinFactor& L = lin[n1 + n2 + n3 + f];
    if (pri_A) std::memcpy(pri_A + 36 * f, L.A[0], 36 * sizeof(double));
    if (pri_b) std::memcpy(pri_b + 6 * f, L.b, 6 * sizeof(double));
  }
}

// One exact damped solve at the current values: d = (A^T A + lambda I)^-1 A^T b. Returns 1 on success.
static int g_blocked = 1;
void oracle_set_blocked(int32_t on) { g_blocked = on; }   // factorisation used by oracle_solve_once (oracle_lm takes it in params[10])
int32_t oracle_solve_once(oracle_graph* h, double lambda, const int32_t* pose_order_hint, double* d_pose, double* d_point) {
  std::vector<LinFactor> lin;
  graph_linearize(h->g, h->g.poses, h->g.points, lin);
  SchurSolver s;
  s.use_blocked = g_blocked != 0;
  s.analyse((int)h->g.poses.size(), (int)h->g.points.size(), lin, pose_order_hint);
  std::vector<double> dp, dpt;
  if (!s.solve(lin, lambda, dp, dpt)) return 0;
  std::memcpy(d_pose, dp.data(), dp.size() * sizeof(double));
  std::memcpy(d_point, dpt.data(), dpt.size() * sizeof(double));
  return 1;
}

// params: [lambda_initial, lambda_factor, lambda_upper, lambda_lower, min_model_fidelity,
//          max_iterations, relative_error_tol, absolute_error_tol, error_tol, verbose]
//          [10] (optional, read when params[10] is given as n_params > 10) blocked factorisation on/off
// out_stats: [iterations, inner_iterations, initial_error, final_error, lambda, n_error_trace,
//             t_linearize, t_assemble, t_eliminate, t_factor, t_substitute, t_error]   (16 slots)
void oracle_lm(oracle_graph* h, const double* params, const int32_t* pose_order_hint, double* out_stats, double* err_trace, int32_t cap) {
  LMParams P;
  P.lambda_initial = params[0]; P.lambda_factor = params[1]; P.lambda_upper = params[2]; P.lambda_lower = params[3];
  P.min_model_fidelity = params[4]; P.max_iterations = (int)params[5]; P.relative_error_tol = params[6];
  P.absolute_error_tol = params[7]; P.error_tol = params[8]; P.verbose = (int)params[9];
  P.blocked_factorisation = params[10] != 0.0;
  LMResult R = levenberg_marquardt(h->g, P, pose_order_hint);
  out_stats[0] = R.iterations; out_stats[1] = R.inner_iterations; out_stats[2] = R.initial_error;
  out_stats[3] = R.final_error; out_stats[4] = R.lambda; out_stats[5] = (double)R.error_per_iteration.size();
  out_stats[6] = R.t_linearize; out_stats[7] = R.t_assemble; out_stats[8] = R.t_eliminate; out_stats[9] = R.t_factor; out_stats[10] = R.t_substitute; out_stats[11] = R.t_error;
  for (int i = 0; i < cap && i < (int)R.error_per_iteration.size(); ++i) err_trace[i] = R.error_per_iteration[i];
}

// ---- small math exports for unit tests ---------------------------------
void oracle_se3_expmap(const double* xi, double* Rt) { pose_to(se3_expmap(xi), Rt); }
void oracle_se3_logmap(const double* Rt, double* xi) { se3_logmap(pose_from(Rt), xi); }
void oracle_se3_logmap_derivative(const double* Rt, double* J36) { Mat6 J = se3_logmap_derivative(pose_from(Rt)); std::memcpy(J36, J.m, sizeof J.m); }
void oracle_se3_compose(const double* a, const double* b, double* out) { pose_to(compose(pose_from(a), pose_from(b)), out); }
void oracle_se3_inverse(const double* a, double* out) { pose_to(inverse(pose_from(a)), out); }
void oracle_se3_retract(const double* a, const double* d, double* out) { pose_to(retract(pose_from(a), d), out); }
void oracle_between_error(const double* x1, const double* x2, const double* z, double* e6, double* H1, double* H2) {
  Mat6 A, B; between_error(pose_from(x1), pose_from(x2), pose_from(z), e6, &A, &B);
  if (H1) std::memcpy(H1, A.m, sizeof A.m);
  if (H2) std::memcpy(H2, B.m, sizeof B.m);
}
void oracle_prior_error(const double* x, const double* z, double* e6, double* H) {
  Mat6 A; prior_error(pose_from(x), pose_from(z), e6, &A);
  if (H) std::memcpy(H, A.m, sizeof A.m);
}
void oracle_p2p_error(const double* X, const double* p, const double* z, double* e3, double* HX, double* Hp) {
  Vec3 e = pose_to_point_error(pose_from(X), {p[0], p[1], p[2]}, {z[0], z[1], z[2]}, HX, Hp);
  e3[0] = e.x; e3[1] = e.y; e3[2] = e.z;
}
void oracle_ternary_error(const double* mp, const double* mc, const double* H, double* e3, double* J1, double* J2, double* J3) {
  Vec3 e = motion_ternary_error({mp[0], mp[1], mp[2]}, {mc[0], mc[1], mc[2]}, pose_from(H), J1, J2, J3);
  e3[0] = e.x; e3[1] = e.y; e3[2] = e.z;
}
// LLWorld factors, batched: whitened + robust-reweighted blocks and b = -whitened residual (NoiseModelFactor::linearize),
// *error = sum of the robust losses
void oracle_llworld_motion_pose(int64_t n, const double* prev_point, const double* cur_point, const double* prev_pose, const double* cur_pose, double sigma, double huber_k,
                                double* A_prev_point, double* A_cur_point, double* A_prev_pose, double* A_cur_pose, double* b, double* error) {
  double total = 0;
  for (int64_t i = 0; i < n; ++i) {
    double J1[9], J2[9], J3[18], J4[18];
    const Vec3 mp = {prev_point[3 * i], prev_point[3 * i + 1], prev_point[3 * i + 2]}, mc = {cur_point[3 * i], cur_point[3 * i + 1], cur_point[3 * i + 2]};
    const Vec3 e = landmark_motion_pose_error(mp, mc, pose_from(prev_pose + 12 * i), pose_from(cur_pose + 12 * i), J1, J2, J3, J4);
    const double dist = norm(e) / sigma;
    total += robust_loss(dist, huber_k);
    const double sw = robust_sqrt_weight(dist, huber_k) / sigma;
    for (int k = 0; k < 9; ++k) { if (A_prev_point) A_prev_point[9 * i + k] = sw * J1[k]; if (A_cur_point) A_cur_point[9 * i + k] = sw * J2[k]; }
    for (int k = 0; k < 18; ++k) { if (A_prev_pose) A_prev_pose[18 * i + k] = sw * J3[k]; if (A_cur_pose) A_cur_pose[18 * i + k] = sw * J4[k]; }
    if (b) { b[3 * i] = -sw * e.x; b[3 * i + 1] = -sw * e.y; b[3 * i + 2] = -sw * e.z; }
  }
  if (error) *error = total;
}
void oracle_llworld_pose_smoothing(int64_t n, const double* pose_k2, const double* pose_k1, const double* pose_k, const double* sigmas6,
                                   double* A_k2, double* A_k1, double* A_k, double* b, double* error) {
  double total = 0;
  for (int64_t i = 0; i < n; ++i) {
    double e[6], J1[36], J2[36], J3[36];
    landmark_pose_smoothing_error(pose_from(pose_k2 + 12 * i), pose_from(pose_k1 + 12 * i), pose_from(pose_k + 12 * i), e, J1, J2, J3);
    double d2 = 0;
    for (int r = 0; r < 6; ++r) { const double w = e[r] / sigmas6[r]; d2 += w * w; }
    total += 0.5 * d2;
    for (int r = 0; r < 6; ++r) {
      const double is = 1.0 / sigmas6[r];
      for (int c = 0; c < 6; ++c) {
        if (A_k2) A_k2[36 * i + 6 * r + c] = is * J1[6 * r + c];
        if (A_k1) A_k1[36 * i + 6 * r + c] = is * J2[6 * r + c];
        if (A_k) A_k[36 * i + 6 * r + c] = is * J3[6 * r + c];
      }
      if (b) b[6 * i + r] = -is * e[r];
    }
  }
  if (error) *error = total;
}

double oracle_robust_loss(double dist, double k) { return robust_loss(dist, k); }
double oracle_robust_sqrt_weight(double dist, double k) { return robust_sqrt_weight(dist, k); }

}  // extern "C"
