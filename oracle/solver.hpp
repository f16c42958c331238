// ORACLE (test infrastructure, not product code).
//
// Exact linear solve of one damped Gauss-Newton system, restating what
// gtsam::LevenbergMarquardtOptimizer::tryLambda does with
// buildDampedSystem + MULTIFRONTAL_CHOLESKY (GTSAM 4.2.0; reference call site
// src/backend/RGBDBackendModule.cc:220-221): solve (A^T A + lambda I) d = A^T b.
// GTSAM's COLAMD elimination order is library-internal and only changes
// round-off; here the points are eliminated first (dense Cholesky per connected
// point cluster: a static landmark, or a dynamic tracklet chain) and the reduced
// pose system is factorised with an envelope (skyline) Cholesky in a caller
// supplied time-like pose order.  Not positive definite -> returns false, the
// analogue of IndeterminantLinearSystemException (a failed lambda trial).
#pragma once
#include <algorithm>
#include <chrono>
#include <numeric>

#include "graph.hpp"

namespace oracle {

struct SchurSolver {
  int np = 0, npt = 0;
  // ---- symbolic ----
  std::vector<int> pos_of_pose, pose_at_pos;     // permutation of poses
  std::vector<int> cluster_ptr, cluster_pts;     // CSR of point clusters
  std::vector<int> local_of_point;               // index of a point inside its cluster
  std::vector<int> w_ptr, w_pose;                // CSR point -> adjacent poses
  std::vector<int> pp_ptr, pp_other;             // CSR point -> adjacent points (other, lower index only)
  std::vector<int64_t> row_ptr;                  // skyline: start of scalar row r (length 6np+1)
  std::vector<int> row_first;                    // first stored column of scalar row r
  // ---- numeric ----
  std::vector<double> w_blk;                     // 18 per (point, pose) adjacency: 3x6 row-major  J_pt^T J_pose
  std::vector<double> pp_blk;                    // 9 per (point, other) adjacency: block (point, other)
  std::vector<double> V, gpt;                    // 9 / 3 per point
  std::vector<double> S, gp;                     // skyline reduced matrix, pose gradient (permuted scalar order)
  std::vector<double> Hpp_diag;                  // unused outside debugging

  static int find_root(std::vector<int>& uf, int i) {
    while (uf[i] != i) { uf[i] = uf[uf[i]]; i = uf[i]; }
    return i;
  }

  int find_w(int pt, int pose) const {
    for (int k = w_ptr[pt]; k < w_ptr[pt + 1]; ++k) if (w_pose[k] == pose) return k;
    return -1;
  }
  int find_pp(int pt, int other) const {
    for (int k = pp_ptr[pt]; k < pp_ptr[pt + 1]; ++k) if (pp_other[k] == other) return k;
    return -1;
  }

  // order_hint: one integer per pose (e.g. frame id); poses are ordered by it (stable).
  void analyse(int n_pose, int n_point, const std::vector<LinFactor>& lin, const int* order_hint) {
    np = n_pose; npt = n_point;
    pose_at_pos.resize(np);
    std::iota(pose_at_pos.begin(), pose_at_pos.end(), 0);
    if (order_hint)
      std::stable_sort(pose_at_pos.begin(), pose_at_pos.end(), [&](int a, int b) { return order_hint[a] < order_hint[b]; });
    pos_of_pose.resize(np);
    for (int p = 0; p < np; ++p) pos_of_pose[pose_at_pos[p]] = p;

    // adjacency (point -> poses), (point -> lower points), clusters
    std::vector<std::vector<int>> adj_pose(npt), adj_pt(npt);
    std::vector<int> uf(npt);
    std::iota(uf.begin(), uf.end(), 0);
    std::vector<int> first_pos(np);
    for (int p = 0; p < np; ++p) first_pos[p] = p;
    for (const LinFactor& L : lin) {
      int pts[3], nps = 0, pss[3], nss = 0;
      for (int j = 0; j < L.nvar; ++j) {
        if (L.var[j] >= np) pts[nps++] = L.var[j] - np; else pss[nss++] = L.var[j];
      }
      for (int i = 0; i < nps; ++i) {
        for (int j = 0; j < nss; ++j)
          if (std::find(adj_pose[pts[i]].begin(), adj_pose[pts[i]].end(), pss[j]) == adj_pose[pts[i]].end())
            adj_pose[pts[i]].push_back(pss[j]);
        for (int j = 0; j < nps; ++j) {
          if (pts[j] < pts[i] && std::find(adj_pt[pts[i]].begin(), adj_pt[pts[i]].end(), pts[j]) == adj_pt[pts[i]].end())
            adj_pt[pts[i]].push_back(pts[j]);
          if (i != j) { int a = find_root(uf, pts[i]), b = find_root(uf, pts[j]); if (a != b) uf[a] = b; }
        }
      }
      if (nps == 0)  // pose-pose factor: envelope must cover the pair
        for (int i = 0; i < nss; ++i) for (int j = 0; j < nss; ++j) {
          const int a = pos_of_pose[pss[i]], b = pos_of_pose[pss[j]];
          if (b < first_pos[a]) first_pos[a] = b;
        }
    }
    w_ptr.assign(npt + 1, 0); pp_ptr.assign(npt + 1, 0);
    for (int i = 0; i < npt; ++i) { w_ptr[i + 1] = w_ptr[i] + (int)adj_pose[i].size(); pp_ptr[i + 1] = pp_ptr[i] + (int)adj_pt[i].size(); }
    w_pose.resize(w_ptr[npt]); pp_other.resize(pp_ptr[npt]);
    for (int i = 0; i < npt; ++i) {
      std::copy(adj_pose[i].begin(), adj_pose[i].end(), w_pose.begin() + w_ptr[i]);
      std::copy(adj_pt[i].begin(), adj_pt[i].end(), pp_other.begin() + pp_ptr[i]);
    }
    // clusters in order of their smallest point
    std::vector<int> cid(npt, -1);
    int nc = 0;
    std::vector<int> root_c(npt, -1);
    for (int i = 0; i < npt; ++i) { int r = find_root(uf, i); if (root_c[r] < 0) root_c[r] = nc++; cid[i] = root_c[r]; }
    cluster_ptr.assign(nc + 1, 0);
    for (int i = 0; i < npt; ++i) cluster_ptr[cid[i] + 1]++;
    for (int c = 0; c < nc; ++c) cluster_ptr[c + 1] += cluster_ptr[c];
    cluster_pts.resize(npt); local_of_point.resize(npt);
    { std::vector<int> fill(cluster_ptr.begin(), cluster_ptr.end() - 1);
      for (int i = 0; i < npt; ++i) { local_of_point[i] = fill[cid[i]] - cluster_ptr[cid[i]]; cluster_pts[fill[cid[i]]++] = i; } }
    // envelope: every pose of a cluster reaches back to the cluster's first pose position
    for (int c = 0; c < nc; ++c) {
      int mn = np;
      for (int k = cluster_ptr[c]; k < cluster_ptr[c + 1]; ++k)
        for (int q = w_ptr[cluster_pts[k]]; q < w_ptr[cluster_pts[k] + 1]; ++q) mn = std::min(mn, pos_of_pose[w_pose[q]]);
      for (int k = cluster_ptr[c]; k < cluster_ptr[c + 1]; ++k)
        for (int q = w_ptr[cluster_pts[k]]; q < w_ptr[cluster_pts[k] + 1]; ++q) {
          int& f = first_pos[pos_of_pose[w_pose[q]]];
          f = std::min(f, mn);
        }
    }
    row_ptr.assign(6 * (size_t)np + 1, 0); row_first.resize(6 * (size_t)np);
    for (int p = 0; p < np; ++p)
      for (int r = 0; r < 6; ++r) {
        const int row = 6 * p + r;
        row_first[row] = 6 * first_pos[p];
        row_ptr[row + 1] = row_ptr[row] + (row - row_first[row] + 1);
      }
    S.resize(row_ptr.back());
    w_blk.resize(18 * (size_t)w_pose.size()); pp_blk.resize(9 * (size_t)pp_other.size());
    V.resize(9 * (size_t)npt); gpt.resize(3 * (size_t)npt); gp.resize(6 * (size_t)np);
    // variable -> (factor, slot) incidences in factor order: assemble() gives every variable's rows to one thread
    inc_ptr.assign((size_t)np + npt + 1, 0);
    for (const LinFactor& L : lin) for (int a = 0; a < L.nvar; ++a) inc_ptr[L.var[a] + 1]++;
    for (size_t v = 0; v < (size_t)np + npt; ++v) inc_ptr[v + 1] += inc_ptr[v];
    inc_fac.resize(inc_ptr.back()); inc_slot.resize(inc_ptr.back());
    { std::vector<int64_t> fill(inc_ptr.begin(), inc_ptr.end() - 1);
      for (size_t f = 0; f < lin.size(); ++f) for (int a = 0; a < lin[f].nvar; ++a) { const int64_t at = fill[lin[f].var[a]]++; inc_fac[at] = (int)f; inc_slot[at] = (unsigned char)a; } }
    // pose position range of every cluster (eliminate_points hands out the reduced system by row ranges)
    cluster_lo.assign(nc, np); cluster_hi.assign(nc, -1);
    for (int c = 0; c < nc; ++c)
      for (int k = cluster_ptr[c]; k < cluster_ptr[c + 1]; ++k)
        for (int q = w_ptr[cluster_pts[k]]; q < w_ptr[cluster_pts[k] + 1]; ++q) {
          cluster_lo[c] = std::min(cluster_lo[c], pos_of_pose[w_pose[q]]); cluster_hi[c] = std::max(cluster_hi[c], pos_of_pose[w_pose[q]]);
        }
  }
  std::vector<int64_t> inc_ptr;
  std::vector<int> inc_fac, cluster_lo, cluster_hi;
  std::vector<unsigned char> inc_slot;

  inline double& Sat(int row, int col) { return S[row_ptr[row] + (col - row_first[row])]; }

  // accumulate A^T A and A^T b
  void assemble(const std::vector<LinFactor>& lin, double lambda) {
    std::fill(S.begin(), S.end(), 0.0); std::fill(gp.begin(), gp.end(), 0.0);
    std::fill(V.begin(), V.end(), 0.0); std::fill(gpt.begin(), gpt.end(), 0.0);
    std::fill(w_blk.begin(), w_blk.end(), 0.0); std::fill(pp_blk.begin(), pp_blk.end(), 0.0);
    // every entry below is owned by the row variable va (its gradient, its diagonal block, its W / point-point blocks,
    // its row of the reduced system): one thread per variable, incidences in factor order => deterministic sums
    const int64_t nvar_total = (int64_t)np + npt;
#pragma omp parallel for schedule(dynamic, 64)
    for (int64_t va64 = 0; va64 < nvar_total; ++va64) {
      const int va = (int)va64;
      for (int64_t at = inc_ptr[va]; at < inc_ptr[va + 1]; ++at) {
        const LinFactor& L = lin[inc_fac[at]];
        const int a = inc_slot[at], da = L.vdim[a];
        // gradient
        for (int i = 0; i < da; ++i) {
          double s = 0;
          for (int r = 0; r < L.dim; ++r) s += L.A[a][r * da + i] * L.b[r];
          if (va < np) gp[6 * pos_of_pose[va] + i] += s; else gpt[3 * (va - np) + i] += s;
        }
        for (int b = 0; b < L.nvar; ++b) {
          const int vb = L.var[b], db = L.vdim[b];
          double blk[36];
          for (int i = 0; i < da; ++i) for (int j = 0; j < db; ++j) {
            double s = 0;
            for (int r = 0; r < L.dim; ++r) s += L.A[a][r * da + i] * L.A[b][r * db + j];
            blk[i * db + j] = s;
          }
          if (va < np && vb < np) {
            const int pa = pos_of_pose[va], pb = pos_of_pose[vb];
            if (pa > pb) { for (int i = 0; i < 6; ++i) for (int j = 0; j < 6; ++j) Sat(6 * pa + i, 6 * pb + j) += blk[i * 6 + j]; }
            else if (pa == pb) { for (int i = 0; i < 6; ++i) for (int j = 0; j <= i; ++j) Sat(6 * pa + i, 6 * pa + j) += blk[i * 6 + j]; }
          } else if (va >= np && vb < np) {
            double* w = &w_blk[18 * (size_t)find_w(va - np, vb)];
            for (int i = 0; i < 18; ++i) w[i] += blk[i];
          } else if (va >= np && vb >= np) {
            if (va == vb) { double* v = &V[9 * (size_t)(va - np)]; for (int i = 0; i < 9; ++i) v[i] += blk[i]; }
            else if (vb < va) { double* v = &pp_blk[9 * (size_t)find_pp(va - np, vb - np)]; for (int i = 0; i < 9; ++i) v[i] += blk[i]; }
          }
        }
      }
    }
    for (int p = 0; p < np; ++p) for (int i = 0; i < 6; ++i) Sat(6 * p + i, 6 * p + i) += lambda;
    for (int i = 0; i < npt; ++i) for (int d = 0; d < 3; ++d) V[9 * (size_t)i + 4 * d] += lambda;
  }

  // dense in-place Cholesky + inverse of an n x n SPD matrix (row-major); false if not PD
  static bool dense_spd_inverse(std::vector<double>& M, int n, std::vector<double>& tmp) {
    for (int j = 0; j < n; ++j) {
      double d = M[j * n + j];
      for (int k = 0; k < j; ++k) d -= M[j * n + k] * M[j * n + k];
      if (!(d > 0)) return false;
      d = std::sqrt(d); M[j * n + j] = d;
      for (int i = j + 1; i < n; ++i) {
        double s = M[i * n + j];
        for (int k = 0; k < j; ++k) s -= M[i * n + k] * M[j * n + k];
        M[i * n + j] = s / d;
      }
    }
    // inverse of L (lower) into tmp
    tmp.assign((size_t)n * n, 0.0);
    for (int c = 0; c < n; ++c) {
      tmp[c * n + c] = 1.0 / M[c * n + c];
      for (int i = c + 1; i < n; ++i) {
        double s = 0;
        for (int k = c; k < i; ++k) s -= M[i * n + k] * tmp[k * n + c];
        tmp[i * n + c] = s / M[i * n + i];
      }
    }
    // M^-1 = L^-T L^-1
    for (int i = 0; i < n; ++i) for (int j = 0; j <= i; ++j) {
      double s = 0;
      for (int k = i; k < n; ++k) s += tmp[k * n + i] * tmp[k * n + j];
      M[i * n + j] = s; M[j * n + i] = s;
    }
    return true;
  }

  std::vector<std::vector<double>> Minv;  // per cluster dense inverse (kept for back-substitution)

  bool eliminate_points() {
    const int nc = (int)cluster_ptr.size() - 1;
    Minv.resize(nc);
    bool ok = true;
    // phase 1: dense inverse of every cluster's point block
#pragma omp parallel for schedule(dynamic, 64)
    for (int c = 0; c < nc; ++c) {
      const int n = cluster_ptr[c + 1] - cluster_ptr[c], N = 3 * n;
      std::vector<double>& M = Minv[c];
      std::vector<double> tmp;
      M.assign((size_t)N * N, 0.0);
      for (int li = 0; li < n; ++li) {
        const int i = cluster_pts[cluster_ptr[c] + li];
        for (int r = 0; r < 3; ++r) for (int q = 0; q < 3; ++q) M[(3 * li + r) * N + 3 * li + q] = V[9 * (size_t)i + 3 * r + q];
        for (int k = pp_ptr[i]; k < pp_ptr[i + 1]; ++k) {
          const int lj = local_of_point[pp_other[k]];
          for (int r = 0; r < 3; ++r) for (int q = 0; q < 3; ++q) {
            M[(3 * li + r) * N + 3 * lj + q] = pp_blk[9 * (size_t)k + 3 * r + q];
            M[(3 * lj + q) * N + 3 * li + r] = pp_blk[9 * (size_t)k + 3 * r + q];
          }
        }
      }
      if (!dense_spd_inverse(M, N, tmp)) {
#pragma omp atomic write
        ok = false;
      }
    }
    if (!ok) return false;
    // phase 2: S -= W^T Minv W ; gp -= W^T Minv gpt. The rows of the reduced system are cut into ranges of pose
    // positions; a range is one task that walks the clusters reaching into it (in cluster order) and writes only its
    // own rows: no atomics, and sums that do not depend on the number of threads.
    const int kRange = 8;
    const int nrange = (np + kRange - 1) / kRange;
    std::vector<std::vector<int>> range_clusters(nrange);
    for (int c = 0; c < nc; ++c)
      if (cluster_hi[c] >= 0) for (int rg = cluster_lo[c] / kRange; rg <= cluster_hi[c] / kRange; ++rg) range_clusters[rg].push_back(c);
#pragma omp parallel for schedule(dynamic, 1)
    for (int rg = 0; rg < nrange; ++rg) {
      const int p_lo = rg * kRange, p_hi = std::min(np, p_lo + kRange);
      for (int c : range_clusters[rg]) {
      const int n = cluster_ptr[c + 1] - cluster_ptr[c], N = 3 * n;
      const std::vector<double>& M = Minv[c];
      for (int li = 0; li < n; ++li) {
        const int i = cluster_pts[cluster_ptr[c] + li];
        bool mine = false;
        for (int ka = w_ptr[i]; ka < w_ptr[i + 1]; ++ka) { const int pa = pos_of_pose[w_pose[ka]]; if (pa >= p_lo && pa < p_hi) mine = true; }
        if (!mine) continue;
        for (int lj = 0; lj < n; ++lj) {
          const int j = cluster_pts[cluster_ptr[c] + lj];
          double Mij[9];
          for (int r = 0; r < 3; ++r) for (int q = 0; q < 3; ++q) Mij[3 * r + q] = M[(3 * li + r) * N + 3 * lj + q];
          for (int ka = w_ptr[i]; ka < w_ptr[i + 1]; ++ka) {
            const int pa = pos_of_pose[w_pose[ka]];
            if (pa < p_lo || pa >= p_hi) continue;
            const double* Wa = &w_blk[18 * (size_t)ka];  // 3x6
            double T[18];                                // Wa^T Mij : 6x3
            for (int r = 0; r < 6; ++r) for (int q = 0; q < 3; ++q) {
              double s = 0;
              for (int k = 0; k < 3; ++k) s += Wa[k * 6 + r] * Mij[3 * k + q];
              T[r * 3 + q] = s;
            }
            for (int r = 0; r < 6; ++r) {
              double s = 0;
              for (int q = 0; q < 3; ++q) s += T[r * 3 + q] * gpt[3 * (size_t)j + q];
              gp[6 * pa + r] -= s;
            }
            for (int kb = w_ptr[j]; kb < w_ptr[j + 1]; ++kb) {
              const int pb = pos_of_pose[w_pose[kb]];
              if (pb > pa) continue;
              const double* Wb = &w_blk[18 * (size_t)kb];
              for (int r = 0; r < 6; ++r) for (int q = 0; q < 6; ++q) {
                if (pa == pb && q > r) continue;
                double s = 0;
                for (int k = 0; k < 3; ++k) s += T[r * 3 + k] * Wb[k * 6 + q];
                Sat(6 * pa + r, 6 * pb + q) -= s;
              }
            }
          }
        }
      }
      }
    }
    return true;
  }

  // skyline Cholesky, in place (lower, row-wise)
  bool factor_reduced() {
    const int n = 6 * np;
    for (int i = 0; i < n; ++i) {
      const int fi = row_first[i];
      double* Li = &S[row_ptr[i]] - fi;  // Li[col]
      for (int j = fi; j < i; ++j) {
        const int fj = row_first[j];
        const double* Lj = &S[row_ptr[j]] - fj;
        const int k0 = std::max(fi, fj);
        double s = Li[j];
        for (int k = k0; k < j; ++k) s -= Li[k] * Lj[k];
        Li[j] = s / Lj[j];
      }
      double d = Li[i];
      for (int k = fi; k < i; ++k) d -= Li[k] * Li[k];
      if (!(d > 0)) return false;
      Li[i] = std::sqrt(d);
    }
    return true;
  }

  // ---- blocked, OpenMP-parallel skyline Cholesky ------------------------------------------------------------
  // Same factor as factor_reduced() up to the order of the floating-point sums; this is the CPU arm that is TIMED as
  // the baseline (bench.py cpu_baseline / --impl reference) and the one the full-size parity tests run, because the
  // scalar row-by-row version above is serial.  Right-looking over panels of kNB columns: the panel's diagonal block is
  // factored by one thread, the rows below it are solved against it in parallel (each row is its own task), the panel
  // rows are packed into a dense (rows x kNB) array, and the trailing update A[i, c] -= P_i . P_c runs in parallel
  // over row strips with a 4 x 4 register tile.  Every row is written by exactly one thread per phase, so the result
  // does not depend on the number of threads.
  static constexpr int kNB = 48;
  double t_factor_diag = 0, t_factor_trsm = 0, t_factor_update = 0;
  double t_assemble = 0, t_eliminate = 0, t_factor = 0, t_substitute = 0;   // seconds, accumulated over solve() calls

  // C[q][0..15] = sum_k A[q][k] * Bt[k][0..15] for four rows q: the inner loop over the 16 columns is unit stride in Bt
  // (the panel packed column-major), so the compiler turns it into broadcast-FMA on 256/512-bit vectors.
  __attribute__((target_clones("avx512f", "avx2", "default")))
  static void tile4x16(const double* a0, const double* a1, const double* a2, const double* a3, const double* bt, size_t ldb, int k, double* out /*4 x 16*/) {
    double s0[16] = {0}, s1[16] = {0}, s2[16] = {0}, s3[16] = {0};
    for (int m = 0; m < k; ++m) {
      const double* b = bt + (size_t)m * ldb;
      const double x0 = a0[m], x1 = a1[m], x2 = a2[m], x3 = a3[m];
#pragma GCC ivdep
      for (int j = 0; j < 16; ++j) { s0[j] += x0 * b[j]; s1[j] += x1 * b[j]; s2[j] += x2 * b[j]; s3[j] += x3 * b[j]; }
    }
    for (int j = 0; j < 16; ++j) { out[j] = s0[j]; out[16 + j] = s1[j]; out[32 + j] = s2[j]; out[48 + j] = s3[j]; }
  }

  bool factor_reduced_blocked() {
    const int n = 6 * np;
    if (n == 0) return true;
    // last[c]: last row whose envelope reaches column c (prefix maximum), so a panel knows how far down it acts
    std::vector<int> last(n, 0);
    for (int i = 0; i < n; ++i) last[row_first[i]] = std::max(last[row_first[i]], i);
    for (int c = 1; c < n; ++c) last[c] = std::max(last[c], last[c - 1]);
    std::vector<double> Pt;  // the same, column-major (padded to 16 rows)
    std::vector<double> P;   // packed panel rows, [row - i_begin][kNB], zero where a row starts right of a column
    bool ok = true;
    for (int j0 = 0; j0 < n && ok; j0 += kNB) {
      const int j1 = std::min(n, j0 + kNB), nb = j1 - j0;
      auto tick = std::chrono::steady_clock::now();
      auto lap = [&](double& acc) { const auto now = std::chrono::steady_clock::now(); acc += std::chrono::duration<double>(now - tick).count(); tick = now; };
      // diagonal block: rows j0..j1-1, columns j0..row (all inside the envelope or structurally zero)
      for (int i = j0; i < j1; ++i) {
        const int fi = row_first[i];
        double* Li = &S[row_ptr[i]] - fi;
        for (int j = std::max(j0, fi); j < i; ++j) {
          const int fj = row_first[j];
          const double* Lj = &S[row_ptr[j]] - fj;
          double s = Li[j];
          for (int k = std::max({j0, fi, fj}); k < j; ++k) s -= Li[k] * Lj[k];
          Li[j] = s / Lj[j];
        }
        double d = Li[i];
        for (int k = std::max(j0, fi); k < i; ++k) d -= Li[k] * Li[k];
        if (!(d > 0)) { ok = false; break; }
        Li[i] = std::sqrt(d);
      }
      if (!ok) break;
      lap(t_factor_diag);
      const int i_begin = j1, i_end = last[j1 - 1] + 1;   // rows below the panel that reach into it
      if (i_end <= i_begin) continue;
      const int nr = i_end - i_begin;
      P.assign((size_t)nr * kNB, 0.0);
      // panel solve L_iJ = A_iJ L_JJ^-T, row by row in parallel, and packing
#pragma omp parallel for schedule(static)
      for (int i = i_begin; i < i_end; ++i) {
        const int fi = row_first[i];
        if (fi >= j1) continue;
        double* Li = &S[row_ptr[i]] - fi;
        double* Pi = &P[(size_t)(i - i_begin) * kNB];
        for (int j = std::max(j0, fi); j < j1; ++j) {
          const int fj = row_first[j];
          const double* Lj = &S[row_ptr[j]] - fj;
          double s = Li[j];
          for (int k = std::max({j0, fi, fj}); k < j; ++k) s -= Li[k] * Lj[k];
          Li[j] = s / Lj[j];
          Pi[j - j0] = Li[j];
        }
      }
      lap(t_factor_trsm);
      // trailing update: strips of 4 rows against 16 columns at a time; a row never writes left of its envelope
      const size_t ldb = ((size_t)nr + 15) / 16 * 16;
      Pt.assign((size_t)kNB * ldb, 0.0);
#pragma omp parallel for schedule(static)
      for (int r = 0; r < nr; ++r) for (int m = 0; m < nb; ++m) Pt[(size_t)m * ldb + r] = P[(size_t)r * kNB + m];
      const int nstrip = (nr + 3) / 4;
#pragma omp parallel for schedule(dynamic, 4)
      for (int st = nstrip - 1; st >= 0; --st) {
        const int r0 = st * 4;
        const double* a[4];
        int fmin = n;
        bool any = false;
        for (int q = 0; q < 4; ++q) {
          const int r = std::min(r0 + q, nr - 1);
          a[q] = &P[(size_t)r * kNB];
          const int fi = row_first[i_begin + r];
          if (fi < j1) any = true;
          fmin = std::min(fmin, fi);
        }
        if (!any) continue;
        const int c_lo = std::max(0, fmin - i_begin) / 16 * 16;
        for (int c0 = c_lo; c0 <= r0 + 3; c0 += 16) {
          double out[64];
          tile4x16(a[0], a[1], a[2], a[3], &Pt[c0], ldb, nb, out);
          for (int q = 0; q < 4 && r0 + q < nr; ++q) {
            const int i = i_begin + r0 + q, fi = row_first[i];
            double* Li = &S[row_ptr[i]] - fi;
            const int t_lo = std::max(0, fi - (i_begin + c0)), t_hi = std::min(16, i - (i_begin + c0) + 1);
            for (int t = t_lo; t < t_hi; ++t) Li[i_begin + c0 + t] -= out[16 * q + t];
          }
        }
      }
      lap(t_factor_update);
    }
    return ok;
  }

  bool use_blocked = true;   // oracle_set_blocked(0) selects the scalar factorisation (tests compare the two)

  void solve_reduced(std::vector<double>& x) const {  // x: rhs in, solution out (permuted scalar order)
    const int n = 6 * np;
    for (int i = 0; i < n; ++i) {
      const double* Li = &S[row_ptr[i]] - row_first[i];
      double s = x[i];
      for (int k = row_first[i]; k < i; ++k) s -= Li[k] * x[k];
      x[i] = s / Li[i];
    }
    for (int i = n - 1; i >= 0; --i) {
      const double* Li = &S[row_ptr[i]] - row_first[i];
      x[i] /= Li[i];
      const double xi = x[i];
      for (int k = row_first[i]; k < i; ++k) x[k] -= Li[k] * xi;
    }
  }

  // Full solve. d_pose: 6*np in natural pose order, d_point: 3*npt.
  bool solve(const std::vector<LinFactor>& lin, double lambda, std::vector<double>& d_pose, std::vector<double>& d_point) {
    auto tick = std::chrono::steady_clock::now();
    auto lap = [&](double& acc) { const auto now = std::chrono::steady_clock::now(); acc += std::chrono::duration<double>(now - tick).count(); tick = now; };
    assemble(lin, lambda);
    lap(t_assemble);
    if (!eliminate_points()) return false;
    lap(t_eliminate);
    if (!(use_blocked ? factor_reduced_blocked() : factor_reduced())) return false;
    lap(t_factor);
    std::vector<double> x(gp);
    solve_reduced(x);
    lap(t_substitute);
    d_pose.resize(6 * (size_t)np); d_point.resize(3 * (size_t)npt);
    for (int p = 0; p < np; ++p) for (int i = 0; i < 6; ++i) d_pose[6 * (size_t)p + i] = x[6 * (size_t)pos_of_pose[p] + i];
    const int nc = (int)cluster_ptr.size() - 1;
#pragma omp parallel for schedule(dynamic, 64)
    for (int c = 0; c < nc; ++c) {
      const int n = cluster_ptr[c + 1] - cluster_ptr[c], N = 3 * n;
      std::vector<double> rhs(N);
      for (int li = 0; li < n; ++li) {
        const int i = cluster_pts[cluster_ptr[c] + li];
        for (int r = 0; r < 3; ++r) {
          double s = gpt[3 * (size_t)i + r];
          for (int k = w_ptr[i]; k < w_ptr[i + 1]; ++k) {
            const double* W = &w_blk[18 * (size_t)k];
            const double* dp = &d_pose[6 * (size_t)w_pose[k]];
            for (int q = 0; q < 6; ++q) s -= W[r * 6 + q] * dp[q];
          }
          rhs[3 * li + r] = s;
        }
      }
      const std::vector<double>& M = Minv[c];
      for (int li = 0; li < n; ++li) {
        const int i = cluster_pts[cluster_ptr[c] + li];
        for (int r = 0; r < 3; ++r) {
          double s = 0;
          for (int k = 0; k < N; ++k) s += M[(size_t)(3 * li + r) * N + k] * rhs[k];
          d_point[3 * (size_t)i + r] = s;
        }
      }
    }
    return true;
  }
};

}  // namespace oracle
