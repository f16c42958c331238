// ORACLE (test infrastructure, not product code).
//
// Restatement of gtsam::LevenbergMarquardtOptimizer (GTSAM 4.2.0:
// nonlinear/LevenbergMarquardtOptimizer.cpp iterate()/tryLambda(),
// NonlinearOptimizer.cpp defaultOptimize()/checkConvergence(),
// internal/LevenbergMarquardtState.h increaseLambda()/decreaseLambda()) as the
// reference drives it with default LevenbergMarquardtParams at
// reference/dynosam/src/backend/RGBDBackendModule.cc:207-221 and :364-376.
// GTSAM's source is not under /root/reference: PARITY UNPINNED against GTSAM;
// the control flow below is the published algorithm (lambda0 1e-5, factor 10
// fixed, lambdaUpper 1e5, lambdaLower 0, minModelFidelity 1e-3, lambda*I
// damping, maxIterations 100, relative/absolute tol 1e-5, errorTol 0).
#pragma once
#include <chrono>
#include <cstdio>

#include "solver.hpp"

namespace oracle {

struct LMParams {
  bool blocked_factorisation = true;   // false: the scalar skyline Cholesky (the two agree to round-off)
  double lambda_initial = 1e-5;
  double lambda_factor = 10.0;
  double lambda_upper = 1e5;
  double lambda_lower = 0.0;
  double min_model_fidelity = 1e-3;
  int max_iterations = 100;
  double relative_error_tol = 1e-5;
  double absolute_error_tol = 1e-5;
  double error_tol = 0.0;
  int verbose = 0;
};

struct LMResult {
  int iterations = 0;        // LevenbergMarquardtOptimizer::iterations()
  int inner_iterations = 0;  // getInnerIterations()
  double initial_error = 0, final_error = 0, lambda = 0;
  std::vector<double> error_per_iteration;  // error() after each outer iterate()
  std::vector<double> lambda_trace;         // lambda of every tryLambda solve
  // wall-clock seconds per phase: linearise, assemble, point elimination, reduced factorisation, substitutions, error passes
  double t_linearize = 0, t_assemble = 0, t_eliminate = 0, t_factor = 0, t_substitute = 0, t_error = 0;
};

inline void retract_all(const std::vector<Pose3>& poses, const std::vector<Vec3>& points,
                        const std::vector<double>& dp, const std::vector<double>& dpt,
                        std::vector<Pose3>& out_poses, std::vector<Vec3>& out_points) {
  out_poses.resize(poses.size()); out_points.resize(points.size());
  for (size_t i = 0; i < poses.size(); ++i) out_poses[i] = retract(poses[i], &dp[6 * i]);
  for (size_t i = 0; i < points.size(); ++i)
    out_points[i] = {points[i].x + dpt[3 * i], points[i].y + dpt[3 * i + 1], points[i].z + dpt[3 * i + 2]};
}

// GaussianFactorGraph::error(delta) = 0.5 sum |A d - b|^2
inline double linear_error(const std::vector<LinFactor>& lin, int np, const std::vector<double>* dp,
                           const std::vector<double>* dpt) {
  double total = 0;
  const int64_t n = (int64_t)lin.size();
#pragma omp parallel for reduction(+ : total) schedule(static)
  for (int64_t f = 0; f < n; ++f) {
    const LinFactor& L = lin[f];
    double r[6];
    for (int i = 0; i < L.dim; ++i) r[i] = -L.b[i];
    if (dp)
      for (int a = 0; a < L.nvar; ++a) {
        const int v = L.var[a], d = L.vdim[a];
        const double* x = v < np ? &(*dp)[6 * (size_t)v] : &(*dpt)[3 * (size_t)(v - np)];
        for (int i = 0; i < L.dim; ++i) for (int j = 0; j < d; ++j) r[i] += L.A[a][i * d + j] * x[j];
      }
    double s = 0;
    for (int i = 0; i < L.dim; ++i) s += r[i] * r[i];
    total += 0.5 * s;
  }
  return total;
}

// Runs LM in place on g.poses / g.points.
inline LMResult levenberg_marquardt(FlatGraph& g, const LMParams& P, const int* pose_order_hint) {
  LMResult R;
  const int np = (int)g.poses.size(), npt = (int)g.points.size();
  double lambda = P.lambda_initial;
  double current_factor = P.lambda_factor;
  double error = graph_error(g, g.poses, g.points);
  R.initial_error = error;
  std::vector<LinFactor> lin;
  SchurSolver solver;
  bool analysed = false;
  std::vector<double> dp, dpt;
  std::vector<Pose3> new_poses; std::vector<Vec3> new_points;

  // NonlinearOptimizer::defaultOptimize
  double currentError = error;
  if (currentError <= P.error_tol || R.iterations >= P.max_iterations) { R.final_error = error; R.lambda = lambda; return R; }
  double newError = currentError;
  do {
    currentError = newError;
    // ---- LevenbergMarquardtOptimizer::iterate ----
    { const auto t0 = std::chrono::steady_clock::now();
      graph_linearize(g, g.poses, g.points, lin);
      R.t_linearize += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count(); }
    if (!analysed) { solver.analyse(np, npt, lin, pose_order_hint); solver.use_blocked = P.blocked_factorisation; analysed = true; }
    while (true) {
      // ---- tryLambda ----
      R.lambda_trace.push_back(lambda);
      bool step_is_successful = false, stop_searching_lambda = false;
      double trialError = 0, modelFidelity = 0;
      const bool solved = solver.solve(lin, lambda, dp, dpt);
      if (solved) {
        const double oldLinearizedError = linear_error(lin, np, nullptr, nullptr);
        const double newLinearizedError = linear_error(lin, np, &dp, &dpt);
        const double linearizedCostChange = oldLinearizedError - newLinearizedError;
        if (linearizedCostChange >= 0) {
          retract_all(g.poses, g.points, dp, dpt, new_poses, new_points);
          { const auto t0 = std::chrono::steady_clock::now();
            trialError = graph_error(g, new_poses, new_points);
            R.t_error += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count(); }
          const double costChange = error - trialError;
          if (linearizedCostChange > std::numeric_limits<double>::epsilon() * oldLinearizedError) {
            modelFidelity = costChange / linearizedCostChange;
            step_is_successful = modelFidelity > P.min_model_fidelity;
          }
          const double minAbsoluteTolerance = P.relative_error_tol * error;
          if (std::fabs(costChange) < minAbsoluteTolerance) stop_searching_lambda = true;
          if (P.verbose)
            std::fprintf(stderr, "[oracle-lm] it %d lambda %.3e err %.12e -> %.12e lin %.6e fid %.4f\n", R.iterations, lambda, error, trialError, linearizedCostChange, modelFidelity);
        }
      } else if (P.verbose) {
        std::fprintf(stderr, "[oracle-lm] it %d lambda %.3e: system not positive definite\n", R.iterations, lambda);
      }
      if (step_is_successful) {
        // decreaseLambda (useFixedLambdaFactor = true)
        g.poses.swap(new_poses); g.points.swap(new_points);
        error = trialError;
        lambda = std::max(P.lambda_lower, lambda / current_factor);
        R.iterations += 1; R.inner_iterations += 1;
        break;
      } else if (!stop_searching_lambda) {
        lambda *= current_factor; R.inner_iterations += 1;  // increaseLambda
        if (lambda >= P.lambda_upper) break;                 // giving up
      } else {
        break;  // cost change too small to be worth larger lambdas
      }
    }
    newError = error;
    R.error_per_iteration.push_back(newError);
    // checkConvergence
    bool converged;
    if (newError <= P.error_tol) converged = true;
    else {
      const double absoluteDecrease = currentError - newError;
      const double relativeDecrease = absoluteDecrease / currentError;
      converged = (P.relative_error_tol && relativeDecrease <= P.relative_error_tol) || (absoluteDecrease <= P.absolute_error_tol);
    }
    if (R.iterations >= P.max_iterations || converged || !std::isfinite(currentError)) break;
  } while (true);
  R.final_error = error; R.lambda = lambda;
  R.t_assemble = solver.t_assemble; R.t_eliminate = solver.t_eliminate; R.t_factor = solver.t_factor; R.t_substitute = solver.t_substitute;
  return R;
}

}  // namespace oracle
