// ba_cpu.cpp -- multi-threaded CPU run of the bundle-adjustment path, for TIMING ONLY (bench.py's cpu_baseline /
// --impl reference leg) and for tests that check it against the numpy oracle.
//
// TEST / BENCHMARK INFRASTRUCTURE: nothing under mpsfm_b200/ links, loads or calls this file.
//
// What it is: the reference's solver stack (pycolmap + Ceres 2.1 through pyceres.solve,
// /root/reference/mpsfm/sfm/mapper/bundle_adjustment.py:104, :163-176, :285-293) cannot be built here (COLMAP fork,
// Ceres and SuiteSparse are not in the tree), so the CPU number next to the GPU one is a PORT:
// the same algorithm -- Ceres' trust-region Levenberg-Marquardt with a Schur-complement linear solve -- in C++17 +
// OpenMP over the host cores, on the same array layout the GPU engine consumes (struct mpba_problem of
// include/mpba.h, here with HOST pointers) and with the per-observation arithmetic of ba_math.cuh compiled for the
// host.  The independent restatement used for PARITY stays the numpy oracle (oracle/lm.py); tests/test_cpu_port.py
// pins this port to it (same iteration count, costs to 1e-9).
//
// The reduced camera system is solved EXACTLY, as Ceres' SPARSE_SCHUR does with CHOLMOD: an envelope (skyline)
// Cholesky of S under a reverse Cuthill-McKee ordering of the images (SkylineCholesky below; the ordering and the
// envelope are computed once per solve, like Ceres' symbolic factorisation).  ba_cpu_set_linear_solver(1) switches
// to the block-Jacobi PCG (to 1e-10) the first round timed.  Everything else (evaluation, Schur elimination,
// back-substitution, step evaluation) has the same operation count per observation as the reference's path.
#include <float.h>
#include <math.h>
#include <omp.h>
#include <string.h>

#include <algorithm>
#include <queue>
#include <vector>

#include "../../include/mpba.h"
#include "../../mpsfm_b200/csrc/ba_math.cuh"

using namespace mpba;

namespace {

struct ImgC {
  double pose[8], R[9], b, es;
  const double* prm;
  int model;
};

inline void load_img(const mpba_problem* p, const double* pose, int img, ImgC& ic) {
  memcpy(ic.pose, pose + 8 * (size_t)img, 8 * sizeof(double));
  quat_to_rotmat(ic.pose, ic.R);
  const int cam = p->img_camera[img];
  ic.model = p->cam_model[cam];
  ic.prm = p->cam_params + (size_t)MPBA_MAX_CAMERA_PARAMS * cam;
  ic.b = p->img_shift_scale[2 * (size_t)img];
  ic.es = ic.pose[7];
}

inline int upper_idx(int i, int j) { return i * (13 - i) / 2 + (j - i); }

struct Lm {
  double radius, lo, hi;
};

// ---- image order: Hcc (21 upper), gc (6), cost ---------------------------------------------------------------
void image_pass(const mpba_problem* p, const LossSpec& ls, const double* pose, const double* X, double* Hc, double* gc,
                double& cost_free, double& cost_fixed) {
  const int C = p->num_images;
  const size_t N = (size_t)p->num_obs;
  double cf = 0.0, cx = 0.0;
#pragma omp parallel for schedule(dynamic, 4) reduction(+ : cf, cx)
  for (int i = 0; i < C; ++i) {
    ImgC ic;
    load_img(p, pose, i, ic);
    const unsigned mask = p->img_free_mask[i];
    double acc[27] = {0};
    for (int k = p->img_ptr[i]; k < p->img_ptr[i + 1]; ++k) {
      const int pt = p->cm_pt[k];
      ObsLin o;
      linearize_obs<true, true>(ic.pose, ic.R, X + 4 * (size_t)pt, ic.model, ic.prm, p->cm_xy[2 * (size_t)k],
                                p->cm_xy[2 * (size_t)k + 1], p->cm_depth[k], p->cm_depth[N + k], p->cm_depth[2 * N + k],
                                ic.b, ic.es, ls, o, p->cm_plane ? p->cm_plane + 4 * (size_t)k : nullptr);
      double H[21], g[6];
      gt_n_g_upper(o.prot, o.N, H);
      gt_vec(o.prot, o.h, g);
      for (int j = 0; j < 21; ++j) acc[j] += H[j];
      for (int j = 0; j < 6; ++j) acc[21 + j] += g[j];
      if (mask == 0 && p->pt_const[pt]) cx += o.rho; else cf += o.rho;
    }
    for (int r = 0, v = 0; r < 6; ++r)
      for (int c = r; c < 6; ++c, ++v) Hc[(size_t)i * 21 + v] = (((mask >> r) & 1u) && ((mask >> c) & 1u)) ? acc[v] : 0.0;
    for (int r = 0; r < 6; ++r) gc[(size_t)i * 6 + r] = ((mask >> r) & 1u) ? acc[21 + r] : 0.0;
  }
  cost_free = cf;
  cost_fixed = cx;
}

// ---- track order: Hpp, gp, damped Cholesky, Schur factors  fac[k] = { Z = N R L^-T (9), p_rot (3), e (3), 0 } ----
void points_pass(const mpba_problem* p, const LossSpec& ls, const double* pose, const double* X, Lm lm, double* c_pt,
                 bool compute_scaling, bool jacobi, double* Hpp, double* gp, double* Linv, double* Hinv, double* fac,
                 bool write_factors, double& gmax, long& bad) {
  const int P = p->num_points;
  const size_t N = (size_t)p->num_obs;
  double gm = 0.0;
  long nbad = 0;
#pragma omp parallel for schedule(dynamic, 256) reduction(max : gm) reduction(+ : nbad)
  for (int pt = 0; pt < P; ++pt) {
    const int beg = p->pt_ptr[pt], end = p->pt_ptr[pt + 1];
    const bool active = p->pt_const[pt] == 0 && end > beg;
    double Hp[6] = {0}, g[3] = {0};
    std::vector<double> Qs;  // N R and p_rot of the track's observations, kept for the factor sweep
    if (active) {
      Qs.resize((size_t)(end - beg) * 12);
      for (int k = beg; k < end; ++k) {
        ImgC ic;
        load_img(p, pose, p->pm_img[k], ic);
        ObsLin o;
        linearize_obs<true, false>(ic.pose, ic.R, X + 4 * (size_t)pt, ic.model, ic.prm, p->pm_xy[2 * (size_t)k],
                                   p->pm_xy[2 * (size_t)k + 1], p->pm_depth[k], p->pm_depth[N + k],
                                   p->pm_depth[2 * N + k], ic.b, ic.es, ls, o,
                                   p->pm_plane ? p->pm_plane + 4 * (size_t)k : nullptr);
        double* Q = &Qs[(size_t)(k - beg) * 12];
        sym3_mul_mat(o.N, ic.R, Q);
        const double* R = ic.R;
        Hp[0] += R[0] * Q[0] + R[3] * Q[3] + R[6] * Q[6];
        Hp[1] += R[0] * Q[1] + R[3] * Q[4] + R[6] * Q[7];
        Hp[2] += R[0] * Q[2] + R[3] * Q[5] + R[6] * Q[8];
        Hp[3] += R[1] * Q[1] + R[4] * Q[4] + R[7] * Q[7];
        Hp[4] += R[1] * Q[2] + R[4] * Q[5] + R[7] * Q[8];
        Hp[5] += R[2] * Q[2] + R[5] * Q[5] + R[8] * Q[8];
        g[0] += R[0] * o.h[0] + R[3] * o.h[1] + R[6] * o.h[2];
        g[1] += R[1] * o.h[0] + R[4] * o.h[1] + R[7] * o.h[2];
        g[2] += R[2] * o.h[0] + R[5] * o.h[1] + R[8] * o.h[2];
        Q[9] = o.prot[0]; Q[10] = o.prot[1]; Q[11] = o.prot[2];
      }
    }
    double c[3] = {0, 0, 0};
    if (compute_scaling) {
      if (active)
        for (int j = 0; j < 3; ++j) c[j] = jacobi ? 1.0 / (1.0 + sqrt(Hp[j == 0 ? 0 : (j == 1 ? 3 : 5)])) : 1.0;
      for (int j = 0; j < 3; ++j) c_pt[(size_t)pt * 3 + j] = c[j];
    } else {
      for (int j = 0; j < 3; ++j) c[j] = c_pt[(size_t)pt * 3 + j];
    }
    double Li[6] = {0}, Hi[6] = {0}, e[3] = {0};
    if (active) {
      double Hd[6] = {Hp[0], Hp[1], Hp[2], Hp[3], Hp[4], Hp[5]};
      Hd[0] += lm_damping(Hp[0], c[0], lm.radius, lm.lo, lm.hi);
      Hd[3] += lm_damping(Hp[3], c[1], lm.radius, lm.lo, lm.hi);
      Hd[5] += lm_damping(Hp[5], c[2], lm.radius, lm.lo, lm.hi);
      if (sym3_chol_inverse(Hd, Li)) {
        sym3_from_chol_inverse(Li, Hi);
      } else {
        for (int j = 0; j < 6; ++j) Li[j] = 0.0;
        ++nbad;
      }
      e[0] = Li[0] * g[0];
      e[1] = Li[1] * g[0] + Li[2] * g[1];
      e[2] = Li[3] * g[0] + Li[4] * g[1] + Li[5] * g[2];
    }
    for (int j = 0; j < 6; ++j) {
      Hpp[(size_t)pt * 6 + j] = Hp[j];
      Linv[(size_t)pt * 9 + j] = Li[j];
      if (Hinv) Hinv[(size_t)pt * 6 + j] = Hi[j];
    }
    for (int j = 0; j < 3; ++j) {
      gp[(size_t)pt * 3 + j] = g[j];
      Linv[(size_t)pt * 9 + 6 + j] = e[j];
    }
    gm = std::max(gm, std::max(fabs(g[0]), std::max(fabs(g[1]), fabs(g[2]))));
    if (write_factors && active) {
      for (int k = beg; k < end; ++k) {
        const double* Q = &Qs[(size_t)(k - beg) * 12];
        double* f = fac + (size_t)k * 16;
        mul_chol_inverse_t(Q, Li, f);
        f[9] = Q[9]; f[10] = Q[10]; f[11] = Q[11];
        f[12] = e[0]; f[13] = e[1]; f[14] = e[2]; f[15] = 0.0;
      }
    }
  }
  gmax = gm;
  bad = nbad;
}

// ---- reduced camera system: one upper block per work unit (u_block / u_item_ptr / item_a / item_b) ----------------
void schur_blocks(const mpba_problem* p, const double* fac, const double* Hc, const double* gc, const double* c_img,
                  Lm lm, double* S, double* rhs) {
  const int U = p->num_upper;
#pragma omp parallel for schedule(dynamic, 8)
  for (int u = 0; u < U; ++u) {
    const int blk = p->u_block[u];
    const int i = p->s_row[blk], j = p->s_col[blk];
    const bool diag = i == j;
    double M[36] = {0}, rv[6] = {0};
    for (int it = p->u_item_ptr[u]; it < p->u_item_ptr[u + 1]; ++it) {
      const double* fa = fac + (size_t)p->item_a[it] * 16;
      const double* fb = fac + (size_t)p->item_b[it] * 16;
      const bool same = fa == fb;
      double K[9];
      for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) K[3 * r + c] = fa[3 * r] * fb[3 * c] + fa[3 * r + 1] * fb[3 * c + 1] + fa[3 * r + 2] * fb[3 * c + 2];
      double Mi[36];
      gt_k_g(fa + 9, K, fb + 9, Mi);
      if (!diag || same) {
        for (int k = 0; k < 36; ++k) M[k] += Mi[k];
      } else {
        for (int r = 0; r < 6; ++r)
          for (int c = 0; c < 6; ++c) M[6 * r + c] += Mi[6 * r + c] + Mi[6 * c + r];
      }
      if (same) {
        const double z[3] = {fa[0] * fa[12] + fa[1] * fa[13] + fa[2] * fa[14], fa[3] * fa[12] + fa[4] * fa[13] + fa[5] * fa[14],
                             fa[6] * fa[12] + fa[7] * fa[13] + fa[8] * fa[14]};
        double g6[6];
        gt_vec(fa + 9, z, g6);
        for (int k = 0; k < 6; ++k) rv[k] += g6[k];
      }
    }
    const unsigned mi = p->img_free_mask[i], mj = p->img_free_mask[j];
    for (int r = 0; r < 6; ++r)
      for (int c = 0; c < 6; ++c) {
        const bool fr = (mi >> r) & 1u, fc = (mj >> c) & 1u;
        double val;
        if (diag) {
          if (fr && fc) {
            const int ui = upper_idx(std::min(r, c), std::max(r, c));
            val = Hc[(size_t)i * 21 + ui] - 0.5 * (M[6 * r + c] + M[6 * c + r]);
            if (r == c) val += lm_damping(Hc[(size_t)i * 21 + ui], c_img[(size_t)i * 6 + r], lm.radius, lm.lo, lm.hi);
          } else {
            val = r == c ? 1.0 : 0.0;
          }
          S[(size_t)blk * 36 + 6 * r + c] = val;
        } else {
          val = (fr && fc) ? -M[6 * r + c] : 0.0;
          S[(size_t)blk * 36 + 6 * r + c] = val;
          S[(size_t)p->s_transpose[blk] * 36 + 6 * c + r] = val;
        }
      }
    if (diag)
      for (int r = 0; r < 6; ++r) rhs[(size_t)i * 6 + r] = ((mi >> r) & 1u) ? gc[(size_t)i * 6 + r] - rv[r] : 0.0;
  }
}

// 6x6 SPD inverse through Cholesky (block-Jacobi preconditioner); identity if not positive definite
void inv6(const double* A_in, double* out) {
  double A[36], Li[36] = {0};
  memcpy(A, A_in, sizeof(A));
  bool ok = true;
  for (int j = 0; j < 6 && ok; ++j) {
    double d = A[7 * j];
    for (int k = 0; k < j; ++k) d -= A[6 * j + k] * A[6 * j + k];
    if (!(d > 0.0)) { ok = false; break; }
    d = sqrt(d);
    A[7 * j] = d;
    for (int r = j + 1; r < 6; ++r) {
      double s = A[6 * r + j];
      for (int k = 0; k < j; ++k) s -= A[6 * r + k] * A[6 * j + k];
      A[6 * r + j] = s / d;
    }
  }
  if (!ok) {
    for (int k = 0; k < 36; ++k) out[k] = (k % 7 == 0) ? 1.0 : 0.0;
    return;
  }
  for (int c = 0; c < 6; ++c) {
    Li[7 * c] = 1.0 / A[7 * c];
    for (int r = c + 1; r < 6; ++r) {
      double s = 0.0;
      for (int k = c; k < r; ++k) s -= A[6 * r + k] * Li[6 * k + c];
      Li[6 * r + c] = s / A[7 * r];
    }
  }
  for (int r = 0; r < 6; ++r)
    for (int c = 0; c < 6; ++c) {
      double s = 0.0;
      for (int k = std::max(r, c); k < 6; ++k) s += Li[6 * k + r] * Li[6 * k + c];
      out[6 * r + c] = s;
    }
}

// block-Jacobi PCG on the block-CSR system; returns iterations, flag 0 converged / 1 cap / 2 breakdown
int pcg(const mpba_problem* p, const double* S, const double* rhs, double* x, double tol, int max_iter, int& flag) {
  const int C = p->num_images, n = 6 * C;
  const bool par = C >= 2048;  // below that a fork/join per PCG step costs more than the step
  std::vector<double> Minv((size_t)C * 36), r(rhs, rhs + n), z(n), q(n), d(n);
#pragma omp parallel for
  for (int i = 0; i < C; ++i) inv6(S + (size_t)p->s_diag[i] * 36, &Minv[(size_t)i * 36]);
  auto precond = [&](const std::vector<double>& in, std::vector<double>& out) {
    double dot = 0.0;
#pragma omp parallel for reduction(+ : dot) if (par)
    for (int i = 0; i < C; ++i)
      for (int a = 0; a < 6; ++a) {
        double s = 0.0;
        for (int b = 0; b < 6; ++b) s += Minv[(size_t)i * 36 + 6 * a + b] * in[6 * i + b];
        out[6 * i + a] = s;
        dot += s * in[6 * i + a];
      }
    return dot;
  };
  std::fill(x, x + n, 0.0);
  const double rz0 = precond(r, z);
  flag = 1;
  if (!(rz0 > 0.0) || !std::isfinite(rz0)) {
    flag = rz0 == 0.0 ? 0 : 2;
    return 0;
  }
  d = z;
  double rz = rz0;
  int it = 0;
  for (; it < max_iter;) {
    double dq = 0.0;
#pragma omp parallel for reduction(+ : dq) if (par)
    for (int i = 0; i < C; ++i) {
      double acc[6] = {0};
      for (int b = p->s_row_ptr[i]; b < p->s_row_ptr[i + 1]; ++b) {
        const double* Sb = S + (size_t)b * 36;
        const double* dj = &d[6 * (size_t)p->s_col[b]];
        for (int a = 0; a < 6; ++a)
          for (int c = 0; c < 6; ++c) acc[a] += Sb[6 * a + c] * dj[c];
      }
      for (int a = 0; a < 6; ++a) {
        q[6 * i + a] = acc[a];
        dq += acc[a] * d[6 * i + a];
      }
    }
    if (!(dq > 0.0) || !std::isfinite(dq)) { flag = 2; break; }
    const double alpha = rz / dq;
#pragma omp parallel for if (par)
    for (int k = 0; k < n; ++k) {
      x[k] += alpha * d[k];
      r[k] -= alpha * q[k];
    }
    const double rz_new = precond(r, z);
    ++it;
    if (!std::isfinite(rz_new)) { flag = 2; break; }
    if (sqrt(fabs(rz_new) / rz0) <= tol) { flag = 0; break; }
    const double beta = rz_new / rz;
    rz = rz_new;
#pragma omp parallel for if (par)
    for (int k = 0; k < n; ++k) d[k] = z[k] + beta * d[k];
  }
  return it;
}


// ---- exact solve of the reduced camera system ------------------------------------------------------------------
// Reverse Cuthill-McKee on the block graph of S, then a scalar envelope Cholesky: row i keeps L[i][first[i] .. i].
// Fixed tangent coordinates are identity rows of S (schur_blocks), so S is positive definite whenever Ceres' is.
struct SkylineCholesky {
  int C = 0, n = 0;
  std::vector<int> order, pos;      // order[new block] = old block, pos[old] = new
  std::vector<int> first;           // per scalar row (new numbering): first column of its envelope
  std::vector<size_t> rowp;         // start of row i in L
  std::vector<double> L, b;

  void analyze(const mpba_problem* p) {
    C = p->num_images;
    n = 6 * C;
    std::vector<int> deg(C);
    for (int i = 0; i < C; ++i) deg[i] = p->s_row_ptr[i + 1] - p->s_row_ptr[i];
    std::vector<char> seen(C, 0);
    order.clear();
    order.reserve(C);
    auto bfs = [&](int root, std::vector<int>& out, std::vector<char>& mark) {
      out.clear();
      std::queue<int> q;
      q.push(root);
      mark[root] = 1;
      std::vector<int> nb;
      while (!q.empty()) {
        const int v = q.front();
        q.pop();
        out.push_back(v);
        nb.clear();
        for (int e = p->s_row_ptr[v]; e < p->s_row_ptr[v + 1]; ++e) {
          const int w = p->s_col[e];
          if (!mark[w]) { mark[w] = 1; nb.push_back(w); }
        }
        std::sort(nb.begin(), nb.end(), [&](int a, int c) { return deg[a] != deg[c] ? deg[a] < deg[c] : a < c; });
        for (int w : nb) q.push(w);
      }
    };
    std::vector<int> comp;
    for (int s0 = 0; s0 < C; ++s0) {
      if (seen[s0]) continue;
      // pseudo-peripheral start: the last vertex of a BFS from the component's first vertex
      std::vector<char> tmp(seen);
      bfs(s0, comp, tmp);
      const int start = comp.back();
      bfs(start, comp, seen);
      order.insert(order.end(), comp.begin(), comp.end());
    }
    std::reverse(order.begin(), order.end());
    pos.assign(C, 0);
    for (int k = 0; k < C; ++k) pos[order[k]] = k;
    first.assign(n, 0);
    for (int k = 0; k < C; ++k) {
      const int i = order[k];
      int lo = k;
      for (int e = p->s_row_ptr[i]; e < p->s_row_ptr[i + 1]; ++e) lo = std::min(lo, pos[p->s_col[e]]);
      for (int r = 0; r < 6; ++r) first[6 * k + r] = 6 * lo;
    }
    rowp.assign(n + 1, 0);
    for (int i = 0; i < n; ++i) rowp[i + 1] = rowp[i] + (size_t)(i - first[i] + 1);
    L.assign(rowp[n], 0.0);
    b.assign(n, 0.0);
  }

  // returns false when S is not positive definite (Ceres: linear solver failure -> invalid step)
  bool solve(const mpba_problem* p, const double* S, const double* rhs, double* x) {
    std::fill(L.begin(), L.end(), 0.0);
    for (int k = 0; k < C; ++k) {
      const int i = order[k];
      for (int e = p->s_row_ptr[i]; e < p->s_row_ptr[i + 1]; ++e) {
        const int kj = pos[p->s_col[e]];
        if (kj > k) continue;
        const double* Sb = S + (size_t)e * 36;
        for (int r = 0; r < 6; ++r)
          for (int c = 0; c < 6; ++c) {
            const int row = 6 * k + r, col = 6 * kj + c;
            if (col <= row) L[rowp[row] + (size_t)(col - first[row])] = Sb[6 * r + c];
          }
      }
      for (int r = 0; r < 6; ++r) b[6 * k + r] = rhs[6 * i + r];
    }
    for (int i = 0; i < n; ++i) {
      double* Li = &L[rowp[i]] - first[i];
      for (int j = first[i]; j <= i; ++j) {
        const double* Lj = &L[rowp[j]] - first[j];
        const int k0 = std::max(first[i], first[j]);
        double s = Li[j];
        for (int k = k0; k < j; ++k) s -= Li[k] * Lj[k];
        if (j < i) {
          Li[j] = s / Lj[j];
        } else {
          if (!(s > 0.0) || !std::isfinite(s)) return false;
          Li[i] = sqrt(s);
        }
      }
    }
    for (int i = 0; i < n; ++i) {  // L z = b
      const double* Li = &L[rowp[i]] - first[i];
      double s = b[i];
      for (int k = first[i]; k < i; ++k) s -= Li[k] * b[k];
      b[i] = s / Li[i];
    }
    for (int i = n - 1; i >= 0; --i) {  // L^T y = z
      const double* Li = &L[rowp[i]] - first[i];
      const double v = b[i] / Li[i];
      b[i] = v;
      for (int k = first[i]; k < i; ++k) b[k] -= Li[k] * v;
    }
    for (int k = 0; k < C; ++k)
      for (int r = 0; r < 6; ++r) x[6 * order[k] + r] = b[6 * k + r];
    return true;
  }
};

int g_linear_solver = 0;  // 0: envelope Cholesky (exact, like SPARSE_SCHUR), 1: block-Jacobi PCG

// y_p = L^-T (e - sum Z^T G y_c), candidate points, point part of the model cost change, step norms
void back_substitute(const mpba_problem* p, const double* X, const double* y, const double* Hpp, const double* gp,
                     const double* Linv, const double* c_pt, const double* fac, bool use_factors, Lm lm, double* Xc,
                     double& mcc, double& step2, double& xn2) {
  const int P = p->num_points;
  double m = 0.0, s2 = 0.0, x2 = 0.0;
#pragma omp parallel for schedule(dynamic, 256) reduction(+ : m, s2, x2)
  for (int pt = 0; pt < P; ++pt) {
    const int beg = p->pt_ptr[pt], end = p->pt_ptr[pt + 1];
    const bool active = p->pt_const[pt] == 0 && end > beg;
    double yp[3] = {0, 0, 0};
    if (active) {
      double zg[3] = {0, 0, 0};
      if (use_factors)
        for (int k = beg; k < end; ++k) {
          const double* f = fac + (size_t)k * 16;
          double ga[3];
          g_vec(f + 9, y + 6 * (size_t)p->pm_img[k], ga);
          zg[0] += f[0] * ga[0] + f[3] * ga[1] + f[6] * ga[2];
          zg[1] += f[1] * ga[0] + f[4] * ga[1] + f[7] * ga[2];
          zg[2] += f[2] * ga[0] + f[5] * ga[1] + f[8] * ga[2];
        }
      const double* Li = Linv + (size_t)pt * 9;
      const double u[3] = {Li[6] - zg[0], Li[7] - zg[1], Li[8] - zg[2]};
      yp[0] = Li[0] * u[0] + Li[1] * u[1] + Li[3] * u[2];
      yp[1] = Li[2] * u[1] + Li[4] * u[2];
      yp[2] = Li[5] * u[2];
      const double d0 = lm_damping(Hpp[(size_t)pt * 6], c_pt[(size_t)pt * 3], lm.radius, lm.lo, lm.hi),
                   d1 = lm_damping(Hpp[(size_t)pt * 6 + 3], c_pt[(size_t)pt * 3 + 1], lm.radius, lm.lo, lm.hi),
                   d2 = lm_damping(Hpp[(size_t)pt * 6 + 5], c_pt[(size_t)pt * 3 + 2], lm.radius, lm.lo, lm.hi);
      m += 0.5 * (gp[(size_t)pt * 3] * yp[0] + gp[(size_t)pt * 3 + 1] * yp[1] + gp[(size_t)pt * 3 + 2] * yp[2] +
                  d0 * yp[0] * yp[0] + d1 * yp[1] * yp[1] + d2 * yp[2] * yp[2]);
      s2 += yp[0] * yp[0] + yp[1] * yp[1] + yp[2] * yp[2];
      const double* Xp = X + 4 * (size_t)pt;
      x2 += Xp[0] * Xp[0] + Xp[1] * Xp[1] + Xp[2] * Xp[2];
    }
    for (int j = 0; j < 3; ++j) Xc[4 * (size_t)pt + j] = X[4 * (size_t)pt + j] - yp[j];
    Xc[4 * (size_t)pt + 3] = 0.0;
  }
  mcc = m;
  step2 = s2;
  xn2 = x2;
}

// q (+) -y, t - y; image part of the model cost change and of the step norms
void cam_update(const mpba_problem* p, const double* pose, const double* y, const double* Hc, const double* gc,
                const double* c_img, Lm lm, double* pc, double& mcc, double& step2, double& xn2) {
  double m = 0.0, s2 = 0.0, x2 = 0.0;
  for (int i = 0; i < p->num_images; ++i) {
    const double* ps = pose + 8 * (size_t)i;
    const double* yi = y + 6 * (size_t)i;
    const unsigned mask = p->img_free_mask[i];
    const bool used = p->img_ptr[i + 1] > p->img_ptr[i];
    const double dl[3] = {-yi[0], -yi[1], -yi[2]};
    double qn[4];
    quat_plus(ps, dl, qn);
    double* o = pc + 8 * (size_t)i;
    for (int k = 0; k < 4; ++k) o[k] = qn[k];
    for (int k = 0; k < 3; ++k) o[4 + k] = ps[4 + k] - yi[3 + k];
    o[7] = ps[7];
    if (used && (mask & 7u))
      for (int k = 0; k < 4; ++k) { s2 += (qn[k] - ps[k]) * (qn[k] - ps[k]); x2 += ps[k] * ps[k]; }
    if (used && (mask & 0x38u))
      for (int k = 0; k < 3; ++k) { s2 += yi[3 + k] * yi[3 + k]; x2 += ps[4 + k] * ps[4 + k]; }
    for (int r = 0; r < 6; ++r)
      if ((mask >> r) & 1u) {
        const double dj = lm_damping(Hc[(size_t)i * 21 + upper_idx(r, r)], c_img[(size_t)i * 6 + r], lm.radius, lm.lo, lm.hi);
        m += 0.5 * (gc[(size_t)i * 6 + r] * yi[r] + dj * yi[r] * yi[r]);
      }
  }
  mcc = m;
  step2 = s2;
  xn2 = x2;
}

}  // namespace

// Ceres 2.1 TrustRegionMinimizer + LevenbergMarquardtStrategy over the passes above.  `prob` carries HOST pointers;
// img_pose / pt_xyz are updated in place to the best state visited.  summary as in mpba_solve.
extern "C" int ba_cpu_solve(const mpba_problem* prob, const mpba_options* opt_in, mpba_summary* sum, int num_threads) {
  if (!prob || !sum) return MPBA_E_INVALID;
  if (num_threads > 0) omp_set_num_threads(num_threads);
  mpba_options opt;
  if (opt_in) {
    opt = *opt_in;
  } else {
    memset(&opt, 0, sizeof(opt));
    opt.max_num_iterations = 50; opt.max_num_consecutive_invalid_steps = 5; opt.jacobi_scaling = 1;
    opt.pcg_max_iterations = 20000; opt.initial_trust_region_radius = 1e4; opt.max_trust_region_radius = 1e16;
    opt.min_trust_region_radius = 1e-32; opt.min_relative_decrease = 1e-3; opt.min_lm_diagonal = 1e-6;
    opt.max_lm_diagonal = 1e32; opt.function_tolerance = 1e-6; opt.gradient_tolerance = 1e-10;
    opt.parameter_tolerance = 1e-8; opt.pcg_tolerance = 1e-10;
  }
  const mpba_problem* p = prob;
  const size_t C = p->num_images, P = p->num_points, N = p->num_obs, B = p->num_blocks;
  const bool structure_only = (p->flags & MPBA_FLAG_STRUCTURE_ONLY) != 0;
  LossSpec ls{p->reproj_loss, p->depth_loss, p->reproj_loss_scale, p->reproj_loss_weight};
  ls.plane_scale = p->pm_plane ? p->plane_loss_scale : 1.0;
  std::vector<double> pose[2] = {std::vector<double>(C * 8), std::vector<double>(C * 8)};
  std::vector<double> X[2] = {std::vector<double>(P * 4), std::vector<double>(P * 4)};
  std::vector<double> Hc[2] = {std::vector<double>(C * 21), std::vector<double>(C * 21)};
  std::vector<double> gc[2] = {std::vector<double>(C * 6), std::vector<double>(C * 6)};
  std::vector<double> c_img(C * 6), c_pt(P * 3), Hpp(P * 6), gp(P * 3), Linv(P * 9), fac(structure_only ? 0 : N * 16),
      S(B * 36), rhs(C * 6), y(C * 6, 0.0);
  SkylineCholesky chol;
  int cur = 0;
  for (size_t i = 0; i < C; ++i) {  // COLMAP normalises the quaternions; slot 7 = exp(log-scale) of the depth prior
    const double* s = p->img_pose + 8 * i;
    const double n = sqrt(s[0] * s[0] + s[1] * s[1] + s[2] * s[2] + s[3] * s[3]);
    double* o = &pose[cur][8 * i];
    for (int k = 0; k < 4; ++k) o[k] = s[k] / n;
    for (int k = 4; k < 7; ++k) o[k] = s[k];
    o[7] = exp(p->img_shift_scale[2 * i + 1]);
  }
  memcpy(X[cur].data(), p->pt_xyz, P * 4 * sizeof(double));

  mpba_iteration* its = sum->iterations;
  const int cap = its ? sum->capacity : 0;
  int nrows = 0;
  mpba_iteration last;
  memset(&last, 0, sizeof(last));
  auto push = [&](const mpba_iteration& it) {
    if (nrows < cap) its[nrows] = it;
    ++nrows;
    last = it;
  };
  auto jacobi_scales = [&](int k) {
    for (size_t i = 0; i < C; ++i)
      for (int r = 0; r < 6; ++r) {
        const double v = Hc[k][i * 21 + upper_idx(r, r)];
        c_img[i * 6 + r] = ((p->img_free_mask[i] >> r) & 1u) ? (opt.jacobi_scaling ? 1.0 / (1.0 + sqrt(v)) : 1.0) : 0.0;
      }
  };
  auto gmax_of = [&](const std::vector<double>& g) {
    double m = 0.0;
    for (double v : g) m = std::max(m, fabs(v));
    return m;
  };

  double cf, cx;
  image_pass(p, ls, pose[cur].data(), X[cur].data(), Hc[cur].data(), gc[cur].data(), cf, cx);
  jacobi_scales(cur);
  double x_cost = 0.5 * cf;
  const double initial_cost_free = x_cost, fixed_cost = 0.5 * cx;
  double gmax_c = gmax_of(gc[cur]);
  double radius = opt.initial_trust_region_radius, decrease_factor = 2.0, minimum_cost = x_cost;
  int n_ok = 0, n_bad = 0, invalid_run = 0, termination = -1, iteration = 0;
  long long total_pcg = 0;
  const char* msg = "";
  bool need_gmax = true, first = true;
  {
    mpba_iteration it0;
    memset(&it0, 0, sizeof(it0));
    it0.cost = x_cost + fixed_cost;
    it0.trust_region_radius = radius;
    push(it0);
  }
  if (!std::isfinite(x_cost)) return MPBA_E_NUMERIC;
  if (iteration >= opt.max_num_iterations) { termination = MPBA_NO_CONVERGENCE; msg = "Maximum number of iterations reached."; }

  while (termination < 0) {
    const int cand = cur ^ 1;
    const Lm lm{radius, opt.min_lm_diagonal, opt.max_lm_diagonal};
    double gmax_p;
    long bad;
    points_pass(p, ls, pose[cur].data(), X[cur].data(), lm, c_pt.data(), first, opt.jacobi_scaling != 0, Hpp.data(),
                gp.data(), Linv.data(), nullptr, fac.data(), !structure_only, gmax_p, bad);
    first = false;
    int pcg_it = 0, pcg_flag = 0;
    if (!structure_only) {
      schur_blocks(p, fac.data(), Hc[cur].data(), gc[cur].data(), c_img.data(), lm, S.data(), rhs.data());
      if (g_linear_solver == 1) {
        pcg_it = pcg(p, S.data(), rhs.data(), y.data(), opt.pcg_tolerance, opt.pcg_max_iterations, pcg_flag);
      } else {
        if (chol.n == 0) chol.analyze(p);
        pcg_flag = chol.solve(p, S.data(), rhs.data(), y.data()) ? 0 : 2;
      }
    }
    double mcc_c, s2c, x2c, mcc_p, s2p, x2p;
    cam_update(p, pose[cur].data(), y.data(), Hc[cur].data(), gc[cur].data(), c_img.data(), lm, pose[cand].data(), mcc_c,
               s2c, x2c);
    back_substitute(p, X[cur].data(), y.data(), Hpp.data(), gp.data(), Linv.data(), c_pt.data(), fac.data(),
                    !structure_only, lm, X[cand].data(), mcc_p, s2p, x2p);
    image_pass(p, ls, pose[cand].data(), X[cand].data(), Hc[cand].data(), gc[cand].data(), cf, cx);

    if (need_gmax) {
      const double g = std::max(gmax_c, gmax_p);
      last.gradient_max_norm = g;
      if (nrows - 1 < cap && nrows > 0) its[nrows - 1].gradient_max_norm = g;
      need_gmax = false;
      if (g <= opt.gradient_tolerance) { termination = MPBA_CONVERGENCE; msg = "Gradient tolerance reached."; break; }
    }
    mpba_iteration it;
    memset(&it, 0, sizeof(it));
    it.cost = last.cost;
    it.gradient_max_norm = last.gradient_max_norm;
    it.pcg_iterations = pcg_it;
    total_pcg += pcg_it;
    ++iteration;
    const double mcc = mcc_c + mcc_p;
    const bool ok = pcg_flag != 2 && bad == 0 && std::isfinite(mcc) && std::isfinite(s2p) && std::isfinite(s2c) && mcc > 0.0;
    it.step_is_valid = ok ? 1 : 0;
    if (!ok) {
      ++invalid_run;
      radius /= decrease_factor;
      decrease_factor *= 2.0;
      it.trust_region_radius = radius;
      ++n_bad;
      push(it);
      if (invalid_run >= opt.max_num_consecutive_invalid_steps) {
        termination = MPBA_FAILURE;
        msg = "Number of consecutive invalid steps more than max_num_consecutive_invalid_steps.";
        break;
      }
    } else {
      invalid_run = 0;
      double cand_cost = 0.5 * cf;
      if (!std::isfinite(cand_cost)) cand_cost = DBL_MAX;
      it.step_norm = sqrt(s2p + s2c);
      const double x_norm = sqrt(x2p + x2c);
      if (it.step_norm <= opt.parameter_tolerance * (x_norm + opt.parameter_tolerance)) {
        termination = MPBA_CONVERGENCE; msg = "Parameter tolerance reached."; push(it); break;
      }
      it.cost_change = x_cost - cand_cost;
      if (fabs(it.cost_change) <= opt.function_tolerance * x_cost) {
        termination = MPBA_CONVERGENCE; msg = "Function tolerance reached."; push(it); break;
      }
      it.relative_decrease = it.cost_change / mcc;
      if (it.relative_decrease > opt.min_relative_decrease) {
        cur = cand;
        x_cost = cand_cost;
        gmax_c = gmax_of(gc[cur]);
        need_gmax = true;
        it.cost = x_cost + fixed_cost;
        it.step_is_successful = 1;
        const double t = 2.0 * it.relative_decrease - 1.0;
        radius = fmin(opt.max_trust_region_radius, radius / fmax(1.0 / 3.0, 1.0 - t * t * t));
        decrease_factor = 2.0;
        ++n_ok;
        minimum_cost = std::min(minimum_cost, x_cost);
      } else {
        radius /= decrease_factor;
        decrease_factor *= 2.0;
        ++n_bad;
      }
      it.trust_region_radius = radius;
      push(it);
    }
    if (iteration >= opt.max_num_iterations) { termination = MPBA_NO_CONVERGENCE; msg = "Maximum number of iterations reached."; }
    else if (radius <= opt.min_trust_region_radius) { termination = MPBA_CONVERGENCE; msg = "Minimum trust region radius reached."; }
  }
  for (size_t i = 0; i < C; ++i) memcpy(p->img_pose + 8 * i, &pose[cur][8 * i], 7 * sizeof(double));
  for (size_t i = 0; i < P; ++i) memcpy(p->pt_xyz + 4 * i, &X[cur][4 * i], 3 * sizeof(double));
  sum->initial_cost = initial_cost_free + fixed_cost;
  sum->final_cost = minimum_cost + fixed_cost;
  sum->fixed_cost = fixed_cost;
  sum->termination = termination;
  sum->num_iterations = iteration;
  sum->num_successful_steps = n_ok;
  sum->num_unsuccessful_steps = n_bad;
  sum->total_pcg_iterations = total_pcg;
  sum->kernel_launches = 0;
  sum->solve_ms = 0.0;
  strncpy(sum->message, msg, sizeof(sum->message) - 1);
  sum->message[sizeof(sum->message) - 1] = 0;
  return MPBA_OK;
}

extern "C" int ba_cpu_max_threads(void) { return omp_get_max_threads(); }
extern "C" void ba_cpu_set_linear_solver(int which) { g_linear_solver = which == 1 ? 1 : 0; }
