// integrate.cu -- depth-map integration with normal priors.
//
// Restates, as ONE kernel launch for a batch of images, what the reference does per image with cupy csr_matrix
// products and cupyx.scipy.sparse.linalg.cg calls from Python
// (/root/reference/mpsfm/sfm/scene/image/integration.py:383-520 Integration._integrate; helpers :139-376, :631-680):
// an IRLS loop on the log-depth map whose every step solves a five-point-stencil SPD system with Jacobi-preconditioned
// CG.  oracle/integration.py is the readable statement of the same arithmetic (pinned to the reference's own class by
// tests/golden/integration.npz); this file follows it operation for operation.
//
// B200 mapping: one thread-block cluster per image (16 CTAs = one GPC when the device allows it, else 8).  An image is
// ~1e5 unknowns and ~20 per-pixel arrays = a few MB: it lives in the 126 MB L2 for the whole solve, so the solve is
// bound by cluster-barrier latency and L2 bandwidth, not HBM.  The operator is applied matrix-free from three
// coefficient arrays (diag, e_r = A(i,i+1), e_d = A(i,i+W)); the two dot products of a CG iteration are reduced
// through distributed shared memory (every CTA deposits its partial into every other CTA's slot, one cluster barrier,
// fixed summation order -> bitwise identical scalars in all CTAs, no atomics, deterministic); control flow (CG
// convergence, IRLS termination rules) is therefore uniform across the cluster without any host round trip.
// There is no GEMM here: no tensor-core path.
#include <cooperative_groups.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include <string>
#include <vector>

#include "../../include/mpba.h"
#include "ba_kernels.cuh"

namespace cg = cooperative_groups;

namespace mpba {
cudaStream_t handle_stream(mpba_handle* h);
int handle_device(mpba_handle* h);
int handle_fail(mpba_handle* h, int code, const std::string& msg);
void handle_count_launches(mpba_handle* h, int n);
}  // namespace mpba

static_assert(sizeof(mpba_integration_conf) == 88 && sizeof(mpba_integration_job) == 152 &&
                  sizeof(mpba_integration_result) == 416, "integration structs: keep mpsfm_b200/integration.py in step");

namespace {

constexpr int IT = 512;        // threads per CTA
constexpr int IWARPS = IT / 32;
constexpr int MAXCL = 16;
constexpr unsigned FULLM = 0xffffffffu;

// per-pixel work arrays of one image, n doubles each, in this order inside the image's workspace slice
enum Arr {
  A_Z = 0, A_ZPRIOR, A_DPREC, A_NX, A_NY, A_NZU, A_NZV, A_PU, A_PV,  // state + constants of the solve (oracle: Prepared)
  A_R, A_P, A_Q, A_B, A_DIAG, A_ER, A_ED, A_MINV,                      // CG vectors and stencil coefficients
  A_SPA, A_SPB,                                                        // lambda2 prec / lambda2 prec log-depth of the LAST sparse hit
  A_WU, A_WV,                                                          // working copy of the weights
  A_COUNT
};

struct DevJob {
  mpba_integration_job j;
  double* ws;        // A_COUNT * n doubles, then ints: last[n], then sparse: sp_id[ns] (int), then doubles sp_prec[ns], sp_logd[ns]
  int32_t* last;     // [n] index of the last sparse entry that hit the pixel (-1: none)
  int32_t* sp_id;    // [ns] pixel of the entry, -1 when the scale filter dropped it
  double* sp_prec;   // [ns]
  double* sp_logd;   // [ns]
};

struct Conf {
  mpba_integration_conf c;
  int mode;          // 0: integrate, 1: variances
  int ignore_depths;
  double var_tol;
};

__device__ __forceinline__ double sigmoid_k(double x, double k) {
  double t = -k * x;
  t = t < -709.0 ? -709.0 : (t > 709.0 ? 709.0 : t);
  return 1.0 / (1.0 + exp(t));
}

// Cluster-wide sums of NV values: block tree, deposit into every CTA's slot through DSMEM, one cluster barrier, sum
// in rank order.  Two slot buffers alternate: a CTA may start depositing reduction k+1 while a slower CTA still reads
// reduction k (it cannot reach k+2 before everyone passed barrier k+1, i.e. finished reading k).
template <int NV>
struct ClusterSum {
  double (*warp)[NV];        // [IWARPS][NV]
  double (*slot)[MAXCL][NV]; // [2][MAXCL][NV]
  int parity;
  __device__ void run(cg::cluster_group& cl, double (&v)[NV]) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      double x = v[k];
      for (int off = 16; off > 0; off >>= 1) x += __shfl_down_sync(FULLM, x, off);
      if (lane == 0) warp[w][k] = x;
    }
    __syncthreads();
    const int ncta = (int)cl.num_blocks(), rank = (int)cl.block_rank();
    if (threadIdx.x < NV * ncta) {
      const int k = threadIdx.x % NV, dst = threadIdx.x / NV;
      double s = 0.0;
#pragma unroll
      for (int ww = 0; ww < IWARPS; ++ww) s += warp[ww][k];
      double* remote = cl.map_shared_rank(&slot[parity][rank][k], dst);
      *remote = s;
    }
    cl.sync();
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      double s = 0.0;
      for (int c = 0; c < ncta; ++c) s += slot[parity][c][k];
      v[k] = s;
    }
    parity ^= 1;
  }
};

// everything one cluster needs to know about its image
struct Img {
  int H, W, n, i0, i1;  // grid, own pixel range of this CTA
  double step, k;
  double *Z, *ZP, *DP, *NX, *NY, *NZU, *NZV, *PU, *PV, *R, *P, *Q, *B, *DG, *ER, *ED, *MI, *SPA, *SPB, *WU, *WV;
  const int32_t* last;
};

// the four one-sided differences of z at pixel i (oracle: Stencil.apply; reference: generate_dx_dy :631-680)
__device__ __forceinline__ void diffs(const Img& g, const double* __restrict__ Z, int i, int r, int c, double& a1, double& a2,
                                      double& a3, double& a4) {
  const double nu = __ldcg(g.NZU + i) / g.step, nv = __ldcg(g.NZV + i) / g.step;
  const double z = __ldcg(Z + i);
  a1 = r >= 1 ? nu * __ldcg(Z + i - g.W) + (-nu) * z : 0.0;
  a2 = r <= g.H - 2 ? (-nu) * __ldcg(Z + i + g.W) + nu * z : 0.0;
  a3 = c <= g.W - 2 ? (-nv) * z + nv * __ldcg(Z + i + 1) : 0.0;
  a4 = c >= 1 ? (-nv) * __ldcg(Z + i - 1) + nv * z : 0.0;
}

// data terms of calc_Amat (:182-185) at pixel j (zero where the neighbour does not exist)
__device__ __forceinline__ double t_top(const Img& g, int j) {
  if (j < g.W) return 0.0;
  const double nz = __ldcg(g.NZU + j);
  return (__ldcg(g.WU + j) * __ldcg(g.PU + j)) * (nz * nz);
}
__device__ __forceinline__ double t_bottom(const Img& g, int j) {
  if (j >= g.n - g.W) return 0.0;
  const double nz = __ldcg(g.NZU + j);
  return ((1.0 - __ldcg(g.WU + j)) * __ldcg(g.PU + j)) * (nz * nz);
}
__device__ __forceinline__ double t_left(const Img& g, int j, int c) {
  if (c < 1) return 0.0;
  const double nz = __ldcg(g.NZV + j);
  return ((1.0 - __ldcg(g.WV + j)) * __ldcg(g.PV + j)) * (nz * nz);
}
__device__ __forceinline__ double t_right(const Img& g, int j, int c) {
  if (c > g.W - 2) return 0.0;
  const double nz = __ldcg(g.NZV + j);
  return (__ldcg(g.WV + j) * __ldcg(g.PV + j)) * (nz * nz);
}
// the four vectors of b_vec (:445-450) at pixel j: A^T is applied to nu * (w * (-n))
__device__ __forceinline__ double b_t1(const Img& g, int j) { return (__ldcg(g.NZU + j) / g.step) * ((__ldcg(g.WU + j) * __ldcg(g.PU + j)) * (-__ldcg(g.NX + j))); }
__device__ __forceinline__ double b_t2(const Img& g, int j) { return (__ldcg(g.NZU + j) / g.step) * (((1.0 - __ldcg(g.WU + j)) * __ldcg(g.PU + j)) * (-__ldcg(g.NX + j))); }
__device__ __forceinline__ double b_t3(const Img& g, int j) { return (__ldcg(g.NZV + j) / g.step) * ((__ldcg(g.WV + j) * __ldcg(g.PV + j)) * (-__ldcg(g.NY + j))); }
__device__ __forceinline__ double b_t4(const Img& g, int j) { return (__ldcg(g.NZV + j) / g.step) * (((1.0 - __ldcg(g.WV + j)) * __ldcg(g.PV + j)) * (-__ldcg(g.NY + j))); }

// A p at pixel i in the CSR column order of the reference's matrix: i-W, i-1, i, i+1, i+W
__device__ __forceinline__ double apply_row(const Img& g, const double* __restrict__ V, int i, int r, int c) {
  double q = 0.0;
  if (r >= 1) q += __ldcg(g.ED + i - g.W) * __ldcg(V + i - g.W);
  if (c >= 1) q += __ldcg(g.ER + i - 1) * __ldcg(V + i - 1);
  q += __ldcg(g.DG + i) * __ldcg(V + i);
  if (c <= g.W - 2) q += __ldcg(g.ER + i) * __ldcg(V + i + 1);
  if (r <= g.H - 2) q += __ldcg(g.ED + i) * __ldcg(V + i + g.W);
  return q;
}

}  // namespace

namespace {

// Jacobi-preconditioned CG exactly as scipy.sparse.linalg.cg runs it for the reference (warm start, atol = rtol |b|,
// convergence test on |r| BEFORE every iteration).  X is updated in place; returns the iteration count.  Uniform
// control flow across the cluster: every branch depends on cluster-reduced scalars only.
__device__ int cluster_cg(cg::cluster_group& cl, ClusterSum<3>& red, const Img& g, double* __restrict__ X, int maxiter,
                          double rtol) {
  const int W = g.W;
  double v[3] = {0.0, 0.0, 0.0};
  for (int i = g.i0 + (int)threadIdx.x; i < g.i1; i += IT) {
    const double b = g.B[i];
    v[0] += b * b;
  }
  red.run(cl, v);  // the barrier inside also publishes the stencil coefficients written just before
  const double bnrm2 = sqrt(v[0]);
  const double atol = rtol * bnrm2;
  if (bnrm2 == 0.0) {
    for (int i = g.i0 + (int)threadIdx.x; i < g.i1; i += IT) X[i] = 0.0;
    cl.sync();
    return 0;
  }
  for (int i = g.i0 + (int)threadIdx.x; i < g.i1; i += IT) g.R[i] = g.B[i] - apply_row(g, X, i, i / W, i % W);
  double rho_prev = 0.0;
  int it = 0;
  for (; it < maxiter; ++it) {
    v[0] = v[1] = v[2] = 0.0;
    for (int i = g.i0 + (int)threadIdx.x; i < g.i1; i += IT) {
      const double r = g.R[i], zr = g.MI[i] * r;
      v[0] += r * r;
      v[1] += r * zr;
    }
    red.run(cl, v);
    if (sqrt(v[0]) < atol) break;
    const double rho = v[1];
    if (it > 0) {
      const double beta = rho / rho_prev;
      for (int i = g.i0 + (int)threadIdx.x; i < g.i1; i += IT) g.P[i] = g.P[i] * beta + g.MI[i] * g.R[i];
    } else {
      for (int i = g.i0 + (int)threadIdx.x; i < g.i1; i += IT) g.P[i] = g.MI[i] * g.R[i];
    }
    cl.sync();  // the neighbours' p
    v[0] = v[1] = v[2] = 0.0;
    for (int i = g.i0 + (int)threadIdx.x; i < g.i1; i += IT) {
      const double q = apply_row(g, g.P, i, i / W, i % W);
      g.Q[i] = q;
      v[0] += g.P[i] * q;
    }
    red.run(cl, v);
    const double alpha = rho / v[0];
    for (int i = g.i0 + (int)threadIdx.x; i < g.i1; i += IT) {
      X[i] += alpha * g.P[i];
      g.R[i] -= alpha * g.Q[i];
    }
    rho_prev = rho;
  }
  cl.sync();  // the neighbours' x before anybody differentiates it
  return it;
}

// update_W (:358-364) on the own pixels, or keep the cached weights
__device__ void pass_weights(const Img& g) {
  for (int i = g.i0 + (int)threadIdx.x; i < g.i1; i += IT) {
    double a1, a2, a3, a4;
    diffs(g, g.Z, i, i / g.W, i % g.W, a1, a2, a3, a4);
    g.WU[i] = sigmoid_k(a2 * a2 - a1 * a1, g.k);
    g.WV[i] = sigmoid_k(a4 * a4 - a3 * a3, g.k);
  }
}

// calc_energy (:139-165) with calc_Wpm (:366-376); the cluster barrier of the reduction also publishes WU / WV
__device__ double pass_energy(cg::cluster_group& cl, ClusterSum<3>& red, const Img& g, const DevJob& J, double lambda1,
                              double lambda2) {
  double v[3] = {0.0, 0.0, 0.0};
  for (int i = g.i0 + (int)threadIdx.x; i < g.i1; i += IT) {
    double a1, a2, a3, a4;
    diffs(g, g.Z, i, i / g.W, i % g.W, a1, a2, a3, a4);
    const double wu = g.WU[i], wv = g.WV[i], pu = g.PU[i], pv = g.PV[i], nx = g.NX[i], ny = g.NY[i];
    const double wup = wu * pu, wum = (1.0 - wu) * pu, wvp = wv * pv, wvm = (1.0 - wv) * pv;
    const double t1 = a1 + nx, t2 = a2 + nx, t3 = a3 + ny, t4 = a4 + ny;
    v[0] += ((wup * (t1 * t1) + wum * (t2 * t2)) + wvp * (t3 * t3)) + wvm * (t4 * t4);
    const double dz = g.ZP[i] - g.Z[i];
    v[1] += (lambda1 * g.DP[i]) * (dz * dz);
  }
  const int ncta = (int)cl.num_blocks();
  for (int k = (int)cl.block_rank() * IT + (int)threadIdx.x; k < J.j.num_sparse; k += ncta * IT) {
    const int id = J.sp_id[k];
    if (id >= 0) {
      const double dz = J.sp_logd[k] - __ldcg(g.Z + id);
      v[2] += (lambda2 * J.sp_prec[k]) * (dz * dz);
    }
  }
  red.run(cl, v);
  return (v[0] + v[1]) + v[2];
}

// calc_Amat (:167-234) as (diag, e_r, e_d), the Jacobi preconditioner (:455-462) and b_vec (:445-453) on the own pixels
__device__ void pass_coefficients(const Img& g, double lambda1, bool use_sparse, bool unit_rhs) {
  const int W = g.W, H = g.H;
  for (int i = g.i0 + (int)threadIdx.x; i < g.i1; i += IT) {
    const int r = i / W, c = i % W;
    const double left_i = t_left(g, i, c), right_i = t_right(g, i, c), top_i = t_top(g, i), bottom_i = t_bottom(g, i);
    const double left_n = c <= W - 2 ? t_left(g, i + 1, c + 1) : 0.0;    // the right neighbour's "left" term
    const double right_n = c >= 1 ? t_right(g, i - 1, c - 1) : 0.0;      // the left neighbour's "right" term
    const double top_n = r <= H - 2 ? t_top(g, i + W) : 0.0;             // the bottom neighbour's "top" term
    const double bottom_n = r >= 1 ? t_bottom(g, i - W) : 0.0;           // the top neighbour's "bottom" term
    double d = 0.0;  // the reference's eight += in its order (:191-198); absent neighbours add exact zeros
    d += left_i; d += left_n; d += right_i; d += right_n; d += top_i; d += top_n; d += bottom_i; d += bottom_n;
    d = d + lambda1 * g.DP[i];
    const bool hit = use_sparse && g.last[i] >= 0;
    if (hit) d += g.SPA[i];
    g.DG[i] = d;
    g.ER[i] = c <= W - 2 ? (-right_i) + (-left_n) : 0.0;
    g.ED[i] = r <= H - 2 ? (-top_n) + (-bottom_i) : 0.0;
    g.MI[i] = 1.0 / fmax(d, 1e-5);
    if (unit_rhs) {
      g.B[i] = 1.0;
    } else {
      double b = 0.0;  // oracle: Stencil.transpose_sum, same order of the eight updates
      if (r >= 1) b -= b_t1(g, i);
      if (r <= H - 2) b += b_t1(g, i + W);
      if (r <= H - 2) b += b_t2(g, i);
      if (r >= 1) b -= b_t2(g, i - W);
      if (c <= W - 2) b -= b_t3(g, i);
      if (c >= 1) b += b_t3(g, i - 1);
      if (c >= 1) b += b_t4(g, i);
      if (c <= W - 2) b -= b_t4(g, i + 1);
      b = b + (lambda1 * g.DP[i]) * g.ZP[i];
      if (hit) b += g.SPB[i];
      g.B[i] = b;
    }
  }
}

__global__ void __launch_bounds__(IT, 1) k_integrate(const DevJob* __restrict__ jobs, const Conf cf,
                                                     mpba_integration_result* __restrict__ results) {
  cg::cluster_group cl = cg::this_cluster();
  const int ncta = (int)cl.num_blocks(), rank = (int)cl.block_rank();
  const int job = blockIdx.x / ncta;
  const DevJob& J = jobs[job];
  __shared__ double s_warp[IWARPS][3];
  __shared__ double s_slot[2][MAXCL][3];
  ClusterSum<3> red{s_warp, s_slot, 0};
  const mpba_integration_conf& c = cf.c;

  Img g;
  g.H = J.j.H; g.W = J.j.W; g.n = g.H * g.W;
  const int per = (g.n + ncta - 1) / ncta;
  g.i0 = min(g.n, rank * per);
  g.i1 = min(g.n, g.i0 + per);
  g.step = c.step_size; g.k = c.k;
  {
    double* w = J.ws;
    const size_t n = (size_t)g.n;
    g.Z = w + A_Z * n; g.ZP = w + A_ZPRIOR * n; g.DP = w + A_DPREC * n; g.NX = w + A_NX * n; g.NY = w + A_NY * n;
    g.NZU = w + A_NZU * n; g.NZV = w + A_NZV * n; g.PU = w + A_PU * n; g.PV = w + A_PV * n; g.R = w + A_R * n;
    g.P = w + A_P * n; g.Q = w + A_Q * n; g.B = w + A_B * n; g.DG = w + A_DIAG * n; g.ER = w + A_ER * n;
    g.ED = w + A_ED * n; g.MI = w + A_MINV * n; g.SPA = w + A_SPA * n; g.SPB = w + A_SPB * n; g.WU = w + A_WU * n;
    g.WV = w + A_WV * n;
    g.last = J.last;
  }
  const bool variances = cf.mode == 1;
  const bool cached_w = !variances && J.j.integrated != 0;

  // ---- setup: process_depth_prior (:236-260), process_normals_prior (:262-279), load_depth_checkpoint (:290-301),
  //      init_int_vars (:314-356) -- oracle: prepare()
  {
    const double k0 = J.j.K[0], k1 = J.j.K[1], k2 = J.j.K[2], k3 = J.j.K[3];
    const double inv_m = 1.0 / c.normals_magnitude_multiplier;
    for (int i = g.i0 + (int)threadIdx.x; i < g.i1; i += IT) {
      const int r = i / g.W, cc = i % g.W;
      const double prior = __ldg(J.j.depth_prior + i);
      double dp = c.depth_magnitude_multiplier * (1.0 / (__ldg(J.j.uncertainty + i) + 1e-6));
      dp = dp * (prior * prior);
      g.DP[i] = dp;
      g.ZP[i] = log(prior);
      const double* nm = J.j.normals + 3 * (size_t)i;
      const double nx = __ldg(nm + 1), ny = __ldg(nm), nz = -__ldg(nm + 2);
      const bool ok = __ldg(J.j.valid + i) != 0;
      const double* nu = J.j.normals_uncertainty + 9 * (size_t)i;
      const double vnx = inv_m * (ok ? __ldg(nu + 4) : c.large_number), vny = inv_m * (ok ? __ldg(nu) : c.large_number),
                   vnz = inv_m * (ok ? __ldg(nu + 8) : c.large_number);
      g.Z[i] = log(J.j.depth[i]);
      const double uu = (double)(g.H - 1 - r) - k2, vv = (double)cc - k3;
      const double nzu = (uu * nx + vv * ny) + k0 * nz, nzv = (uu * nx + vv * ny) + k1 * nz;
      const double duz = -nx / nzu, dvz = -ny / nzv;
      const double e1 = uu * duz + 1.0, e2 = vv * duz, f1 = uu * dvz, f2 = vv * dvz + 1.0;
      g.PU[i] = 1.0 / ((vnx * (e1 * e1) + vny * (e2 * e2)) + ((k0 * k0) * vnz) * (duz * duz));
      g.PV[i] = 1.0 / ((vnx * (f1 * f1) + vny * (f2 * f2)) + ((k1 * k1) * vnz) * (dvz * dvz));
      g.NX[i] = nx; g.NY[i] = ny; g.NZU[i] = nzu; g.NZV[i] = nzv;
      g.SPA[i] = 0.0; g.SPB[i] = 0.0;
      J.last[i] = -1;
      if (cached_w) { g.WU[i] = J.j.wu[i]; g.WV[i] = J.j.wv[i]; }
      if (variances) J.j.variance[i] = 0.0;
    }
    // process_sparse_depth (:281-288) + scale filter (:391-397; calculate_hessian has none, :542)
    const double sff = variances ? 0.0 : c.scale_filter_factor;
    for (int k = rank * IT + (int)threadIdx.x; k < J.j.num_sparse; k += ncta * IT) {
      const int x = __ldg(J.j.sparse_xy + 2 * k), y = __ldg(J.j.sparse_xy + 2 * k + 1);
      const int id = y * g.W + x;
      const double d3 = __ldg(J.j.sparse_depth + k);
      const double prec = (1.0 / __ldg(J.j.sparse_zvar + k)) * (d3 * d3);
      const double logd = log(d3);
      bool ok = x >= 0 && x < g.W && y >= 0 && y < g.H;
      if (ok && sff > 0.0) {
        const double div = exp(logd) / exp(log(__ldg(J.j.depth_prior + id)));
        ok = div < sff && div > 1.0 / sff;
      }
      J.sp_id[k] = ok ? id : -1;
      J.sp_prec[k] = prec;
      J.sp_logd[k] = logd;
    }
    cl.sync();
    // numpy's `a[ids] += v` keeps only the LAST entry of a repeated pixel (:203, :453)
    for (int k = rank * IT + (int)threadIdx.x; k < J.j.num_sparse; k += ncta * IT) {
      const int id = J.sp_id[k];
      if (id >= 0) atomicMax(J.last + id, k);
    }
    cl.sync();
    for (int k = rank * IT + (int)threadIdx.x; k < J.j.num_sparse; k += ncta * IT) {
      const int id = J.sp_id[k];
      if (id >= 0 && __ldcg(J.last + id) == k) {
        const double a = c.lambda2 * J.sp_prec[k];
        g.SPA[id] = a;
        g.SPB[id] = a * J.sp_logd[k];
      }
    }
    cl.sync();
  }

  mpba_integration_result* res = results + job;
  const bool writer = rank == 0 && threadIdx.x == 0;

  if (variances) {
    // calculate_hessian (:522-574) at the current depth with fresh weights, then y = H^-1 1
    pass_weights(g);
    cl.sync();
    pass_coefficients(g, c.lambda1, cf.ignore_depths == 0, true);
    const int its = cluster_cg(cl, red, g, J.j.variance, c.cg_max_iter, cf.var_tol);
    if (writer) {
      res->updated = 0; res->integrated = J.j.integrated; res->irls_steps = 0; res->cg_iterations = its;
      res->energy_old = J.j.energy_old;
    }
    cl.sync();
    return;
  }

  // ---- Integration._integrate (:399-520)
  if (!cached_w) pass_weights(g);
  double energy = pass_energy(cl, red, g, J, c.lambda1, c.lambda2);
  if (writer) res->energy[0] = energy;
  if (J.j.integrated != 0 && !(fabs(energy - J.j.energy_old) / J.j.energy_old > c.tol)) {
    // "Energy hasn't changed. Skipping this frame." (:430-434)
    if (writer) {
      res->updated = 0; res->integrated = 1; res->irls_steps = 0; res->cg_iterations = 0; res->energy_old = J.j.energy_old;
    }
    cl.sync();
    return;
  }
  const double energy_0 = energy;
  double min_energy = energy;
  int steps = 0, cg_total = 0, outcome = 1;  // 1: depth updated, 0: "Energy increased" (:504-508)
  for (int s = 0; s < c.max_iter; ++s) {
    pass_coefficients(g, c.lambda1, true, false);
    const int its = cluster_cg(cl, red, g, g.Z, c.cg_max_iter, c.cg_tol);
    pass_weights(g);
    const double energy_old = energy;
    min_energy = fmin(energy, min_energy);
    energy = pass_energy(cl, red, g, J, c.lambda1, c.lambda2);
    ++steps;
    cg_total += its;
    if (writer) { res->energy[steps] = energy; res->cg_steps[steps - 1] = its; }
    const double rel = fabs(energy - energy_old) / energy_old, rel_min = fabs(energy - min_energy) / min_energy;
    if (((rel < c.tol && (energy_old - energy) > 0.0) || (rel_min < c.tol && (min_energy - energy) > 0.0)) && energy < energy_0)
      break;
    if (energy > energy_0) { outcome = 0; break; }
  }
  for (int i = g.i0 + (int)threadIdx.x; i < g.i1; i += IT) {
    if (outcome) J.j.depth[i] = exp(g.Z[i]);
    J.j.wu[i] = g.WU[i];  // self.wu / self.wv follow every update_W (:469), also on the "energy increased" exit
    J.j.wv[i] = g.WV[i];
  }
  if (writer) {
    res->updated = outcome; res->integrated = 1; res->irls_steps = steps; res->cg_iterations = cg_total;
    res->energy_old = outcome ? energy : energy_0;
  }
  cl.sync();
}

}  // namespace

// ================================================================================================
// host side of the C ABI
// ================================================================================================
namespace {

size_t up256(size_t x) { return (x + 255) & ~(size_t)255; }

// bytes of one job's slice: A_COUNT per-pixel doubles, last[n], sp_id[ns], sp_prec[ns], sp_logd[ns]
size_t job_bytes(const mpba_integration_job& j) {
  const size_t n = (size_t)j.H * (size_t)j.W, ns = (size_t)(j.num_sparse > 0 ? j.num_sparse : 0);
  return up256(A_COUNT * n * sizeof(double)) + up256(n * sizeof(int32_t)) + up256(ns * sizeof(int32_t)) +
         2 * up256(ns * sizeof(double));
}

int check_jobs(mpba_handle* h, int32_t num_jobs, const mpba_integration_job* jobs, const mpba_integration_conf* conf,
               bool variances) {
  if (num_jobs < 0 || (num_jobs > 0 && !jobs)) return mpba::handle_fail(h, MPBA_E_INVALID, "integration: jobs is NULL");
  if (!conf) return mpba::handle_fail(h, MPBA_E_INVALID, "integration: conf is NULL");
  if (conf->max_iter < 0 || conf->max_iter > MPBA_INT_MAX_IRLS || conf->cg_max_iter < 0 || !(conf->step_size != 0.0))
    return mpba::handle_fail(h, MPBA_E_INVALID, "integration: max_iter must be in [0, MPBA_INT_MAX_IRLS], step_size non-zero");
  for (int32_t k = 0; k < num_jobs; ++k) {
    const mpba_integration_job& j = jobs[k];
    if (j.H < 1 || j.W < 1 || (int64_t)j.H * j.W > (int64_t)1 << 30 || j.num_sparse < 0)
      return mpba::handle_fail(h, MPBA_E_INVALID, "integration: bad grid / sparse count in job " + std::to_string(k));
    if (!j.depth_prior || !j.uncertainty || !j.valid || !j.normals || !j.normals_uncertainty || !j.depth ||
        (!variances && (!j.wu || !j.wv)) || (variances && !j.variance) ||
        (j.num_sparse > 0 && (!j.sparse_xy || !j.sparse_depth || !j.sparse_zvar)))
      return mpba::handle_fail(h, MPBA_E_INVALID, "integration: NULL map pointer in job " + std::to_string(k));
  }
  return MPBA_OK;
}

int run_jobs(mpba_handle* h, int32_t num_jobs, const mpba_integration_job* jobs, const mpba_integration_conf* conf,
             int mode, int ignore_depths, double var_tol, void* workspace, size_t workspace_bytes,
             mpba_integration_result* results) {
  if (!h) return MPBA_E_INVALID;
  int rc = check_jobs(h, num_jobs, jobs, conf, mode == 1);
  if (rc) return rc;
  if (num_jobs == 0) return MPBA_OK;
  if (!results) return mpba::handle_fail(h, MPBA_E_INVALID, "integration: results is NULL");
  size_t need = 0;
  mpba_integration_workspace_bytes(num_jobs, jobs, &need);
  if (!workspace || workspace_bytes < need)
    return mpba::handle_fail(h, MPBA_E_WORKSPACE, "integration workspace too small: need " + std::to_string(need) + " bytes");
  cudaError_t e = cudaSetDevice(mpba::handle_device(h));
  if (e != cudaSuccess) return mpba::handle_fail(h, MPBA_E_CUDA, cudaGetErrorString(e));
  cudaStream_t st = mpba::handle_stream(h);

  // workspace: [DevJob x num_jobs][results x num_jobs][job slices ...]
  char* base = static_cast<char*>(workspace);
  DevJob* d_jobs = reinterpret_cast<DevJob*>(base);
  size_t off = up256(sizeof(DevJob) * (size_t)num_jobs);
  mpba_integration_result* d_res = reinterpret_cast<mpba_integration_result*>(base + off);
  off += up256(sizeof(mpba_integration_result) * (size_t)num_jobs);
  std::vector<DevJob> hj((size_t)num_jobs);
  for (int32_t k = 0; k < num_jobs; ++k) {
    const mpba_integration_job& j = jobs[k];
    const size_t n = (size_t)j.H * (size_t)j.W, ns = (size_t)j.num_sparse;
    DevJob& d = hj[(size_t)k];
    d.j = j;
    char* p = base + off;
    d.ws = reinterpret_cast<double*>(p);
    p += up256(A_COUNT * n * sizeof(double));
    d.last = reinterpret_cast<int32_t*>(p);
    p += up256(n * sizeof(int32_t));
    d.sp_id = reinterpret_cast<int32_t*>(p);
    p += up256(ns * sizeof(int32_t));
    d.sp_prec = reinterpret_cast<double*>(p);
    p += up256(ns * sizeof(double));
    d.sp_logd = reinterpret_cast<double*>(p);
    off += job_bytes(j);
  }
  e = cudaMemcpyAsync(d_jobs, hj.data(), sizeof(DevJob) * (size_t)num_jobs, cudaMemcpyHostToDevice, st);
  if (e == cudaSuccess) e = cudaMemsetAsync(d_res, 0, sizeof(mpba_integration_result) * (size_t)num_jobs, st);
  if (e != cudaSuccess) return mpba::handle_fail(h, MPBA_E_CUDA, cudaGetErrorString(e));

  // one cluster per image: 16 CTAs (one GPC, non-portable size) when the device can co-schedule them, else 8
  static int cluster_size = 0;
  if (!cluster_size) {
    cluster_size = 8;
    if (cudaFuncSetAttribute(k_integrate, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) == cudaSuccess) {
      cudaLaunchConfig_t q = {};
      q.gridDim = dim3(16);
      q.blockDim = dim3(IT);
      cudaLaunchAttribute a;
      a.id = cudaLaunchAttributeClusterDimension;
      a.val.clusterDim.x = 16; a.val.clusterDim.y = 1; a.val.clusterDim.z = 1;
      q.attrs = &a;
      q.numAttrs = 1;
      int nclusters = 0;
      if (cudaOccupancyMaxActiveClusters(&nclusters, k_integrate, &q) == cudaSuccess && nclusters >= 1) cluster_size = 16;
    }
    cudaGetLastError();
    if (const char* env = getenv("MPBA_INT_CLUSTER")) {
      const int v = atoi(env);
      if (v == 1 || v == 2 || v == 4 || v == 8 || v == 16) cluster_size = v;
    }
  }
  Conf cf;
  cf.c = *conf;
  cf.mode = mode;
  cf.ignore_depths = ignore_depths;
  cf.var_tol = var_tol;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)num_jobs * (unsigned)cluster_size);
  cfg.blockDim = dim3(IT);
  cfg.stream = st;
  cudaLaunchAttribute attr;
  attr.id = cudaLaunchAttributeClusterDimension;
  attr.val.clusterDim.x = (unsigned)cluster_size; attr.val.clusterDim.y = 1; attr.val.clusterDim.z = 1;
  cfg.attrs = &attr;
  cfg.numAttrs = 1;
  const DevJob* dj = d_jobs;
  e = cudaLaunchKernelEx(&cfg, k_integrate, dj, cf, d_res);
  mpba::handle_count_launches(h, 1);
  if (e == cudaSuccess) e = cudaMemcpyAsync(results, d_res, sizeof(mpba_integration_result) * (size_t)num_jobs, cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  if (e != cudaSuccess) return mpba::handle_fail(h, MPBA_E_CUDA, std::string("mpba_integrate_depth: ") + cudaGetErrorString(e));
  return MPBA_OK;
}

}  // namespace

extern "C" {

void mpba_integration_default_conf(mpba_integration_conf* c) {
  if (!c) return;
  memset(c, 0, sizeof(*c));
  c->max_iter = 10; c->cg_max_iter = 5000; c->tol = 5e-2; c->step_size = 1.0; c->cg_tol = 1e-3; c->lambda1 = 1.0;
  c->lambda2 = 1.0; c->k = 1.0; c->large_number = 1e6; c->depth_magnitude_multiplier = 1.0;
  c->normals_magnitude_multiplier = 1.0; c->scale_filter_factor = 1.5;
}

int mpba_integration_workspace_bytes(int32_t num_jobs, const mpba_integration_job* jobs, size_t* bytes) {
  if (!bytes || num_jobs < 0 || (num_jobs > 0 && !jobs)) return MPBA_E_INVALID;
  size_t total = up256(sizeof(DevJob) * (size_t)num_jobs) + up256(sizeof(mpba_integration_result) * (size_t)num_jobs);
  for (int32_t k = 0; k < num_jobs; ++k) {
    if (jobs[k].H < 1 || jobs[k].W < 1 || jobs[k].num_sparse < 0) return MPBA_E_INVALID;
    total += job_bytes(jobs[k]);
  }
  *bytes = total;
  return MPBA_OK;
}

int mpba_integrate_depth(mpba_handle* h, int32_t num_jobs, const mpba_integration_job* jobs,
                         const mpba_integration_conf* conf, void* workspace, size_t workspace_bytes,
                         mpba_integration_result* results) {
  return run_jobs(h, num_jobs, jobs, conf, 0, 0, 0.0, workspace, workspace_bytes, results);
}

int mpba_integration_variances(mpba_handle* h, int32_t num_jobs, const mpba_integration_job* jobs,
                               const mpba_integration_conf* conf, int32_t ignore_depths, double var_tol, void* workspace,
                               size_t workspace_bytes, mpba_integration_result* results) {
  if (!(var_tol > 0.0)) return mpba::handle_fail(h, MPBA_E_INVALID, "mpba_integration_variances: var_tol must be positive");
  return run_jobs(h, num_jobs, jobs, conf, 1, ignore_depths ? 1 : 0, var_tol, workspace, workspace_bytes, results);
}

}  // extern "C"
