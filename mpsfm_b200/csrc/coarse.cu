// coarse.cu -- coarse space of the two-level preconditioner of the reduced camera system.
//
// Block-Jacobi PCG on S needs O(chain length) iterations on long camera chains: the slow modes of S are
// slowly varying GAUGE transformations (S without damping and gauge fixing is singular exactly on the 7
// similarity modes of the scene).  The coarse space therefore takes, for every group g of consecutive
// images, the 7 gauge modes restricted to the group,
//     translation k : (d_rot, d_t) = (0, -R e_k)        rotation k : (-R e_k / 2, 0)        scale : (0, t)
// in the tangent parameterisation of ba_math.cuh (half-angle left rotation, additive translation; they follow
// from  R' X' + t' = s (R X + t)  under  X' = s Q X + d), and the preconditioner is additive two-level:
//     M^-1 = blockdiag(S)^-1 + Z (Z^T S Z)^-1 Z^T .
// Measured on the synthetic ring scenes (numpy, tolerance 1e-12): 1 000 images 1 062 -> 138 PCG iterations with
// 16 groups, 100 images 149 -> 31.
//
// Three small kernels per linear solve:  k_coarse_modes (Z_i, 6x7 per image),  k_coarse_assemble (A_c = Z^T S Z,
// one CTA per coarse block, fixed summation order),  k_coarse_invert (dense in-place Gauss-Jordan in registers,
// one CTA up to 128 modes, a four-CTA cluster up to 224; a vanishing pivot drops that mode instead of failing).  The application lives in pcg_cluster.cu.
#include <cooperative_groups.h>

#include "ba_kernels.cuh"

namespace cg = cooperative_groups;

namespace mpba {

namespace {

__global__ void k_coarse_modes(int C, const double* __restrict__ pose, const uint8_t* __restrict__ free_mask,
                               double* __restrict__ Zc) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= C) return;
  const double* ps = pose + (size_t)i * 8;
  double R[9];
  quat_to_rotmat(ps, R);
  const unsigned mask = free_mask[i];
  double Z[42];
#pragma unroll
  for (int k = 0; k < 42; ++k) Z[k] = 0.0;
#pragma unroll
  for (int r = 0; r < 3; ++r) {
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      Z[(3 + r) * 7 + k] = -R[3 * r + k];       // translation mode k
      Z[r * 7 + 3 + k] = -0.5 * R[3 * r + k];   // rotation mode k
    }
    Z[(3 + r) * 7 + 6] = ps[4 + r];             // scale mode
  }
  double* dst = Zc + (size_t)i * 42;
#pragma unroll
  for (int r = 0; r < 6; ++r)
#pragma unroll
    for (int k = 0; k < 7; ++k) dst[r * 7 + k] = ((mask >> r) & 1u) ? Z[r * 7 + k] : 0.0;
}

// A_c = Z^T S Z.  Grid (G*G, COARSE_SPLIT): CTA (cb, sp) handles every COARSE_SPLIT-th run of fine blocks of coarse
// block cb; inside the CTA 8 slices of 64 lanes (49 active: lane = 7 a + b) walk the list with stride 8 and the
// slices are folded in shared memory in slice order; the COARSE_SPLIT partial matrices are added by
// k_coarse_invert in split order -> the summation order is fixed, no atomics.
constexpr int CA_SLICES = 8;
__global__ void __launch_bounds__(64 * CA_SLICES) k_coarse_assemble(int G, const int32_t* __restrict__ cb_ptr,
                                                                    const int32_t* __restrict__ cb_fine,
                                                                    const int32_t* __restrict__ s_row,
                                                                    const int32_t* __restrict__ s_col,
                                                                    const double* __restrict__ S,
                                                                    const double* __restrict__ Zc,
                                                                    double* __restrict__ Ac_part) {
  __shared__ double part[CA_SLICES][64];
  const int cb = blockIdx.x, sp = blockIdx.y, gi = cb / G, gj = cb % G;
  const int slice = threadIdx.x >> 6, l = threadIdx.x & 63;
  const int a = l / 7, b = l % 7;
  double acc = 0.0;
  if (l < 49) {
    const int k0 = cb_ptr[cb], k1 = cb_ptr[cb + 1];
    for (int k = k0 + sp * CA_SLICES + slice; k < k1; k += CA_SLICES * COARSE_SPLIT) {
      const int blk = __ldg(cb_fine + k);
      const double* Sb = S + (size_t)blk * 36;
      const double* Zi = Zc + (size_t)__ldg(s_row + blk) * 42;
      const double* Zj = Zc + (size_t)__ldg(s_col + blk) * 42;
      double zi[6], zj[6];
#pragma unroll
      for (int r = 0; r < 6; ++r) {
        zi[r] = Zi[r * 7 + a];
        zj[r] = Zj[r * 7 + b];
      }
      double s = 0.0;
#pragma unroll
      for (int c = 0; c < 6; ++c) {
        double t = 0.0;
#pragma unroll
        for (int r = 0; r < 6; ++r) t += zi[r] * Sb[6 * r + c];
        s += t * zj[c];
      }
      acc += s;
    }
  }
  part[slice][l] = acc;
  __syncthreads();
  if (slice == 0 && l < 49) {
    double t = 0.0;
#pragma unroll
    for (int q = 0; q < CA_SLICES; ++q) t += part[q][l];
    const int nc = 7 * G;
    Ac_part[(size_t)sp * nc * nc + (size_t)(gi * 7 + a) * nc + gj * 7 + b] = t;
  }
}

// In-place Gauss-Jordan inversion of the SPD coarse matrix (no pivoting) by ONE CTA with the matrix IN REGISTERS:
// thread (warp w, lane l) owns the entries (w + 32 qi, l + 32 qj), qi, qj < GJ_Q (25 doubles).  Per pivot p the warp
// owning row p scales it and publishes it through a double-buffered shared-memory row (one CTA barrier per
// pivot); the pivot column comes from the owning lane by shuffle.  Shared memory sees 5 loads per thread and
// pivot instead of the 2 n^2 accesses of a shared-memory resident update.  A pivot that has lost all its weight
// (dependent mode: a group with fewer free coordinates than modes) is dropped: its row and column of the
// inverse are zero.  The reciprocal is a float seed + two Newton steps (this is a preconditioner).  The result
// is symmetrised on the way out.
constexpr int GJ_SINGLE_MAX = 128;  // largest system the one-CTA kernel holds (4 x 4 register tile per thread)
constexpr int GJ_Q = GJ_SINGLE_MAX / 32;
__device__ __forceinline__ double fast_rcp(double a) {
  double x = (double)(1.0f / (float)a);
  x = x * (2.0 - a * x);
  return x * (2.0 - a * x);
}
// No CTA-wide barrier in the elimination (round 2; the first version had one __syncthreads per pivot, ~0.7 us each):
// every pivot row has its own slot in shared memory and its own mbarrier; the owner warp publishes and arrives, the
// others wait on it (mbarrier.try_wait suspends, no spinning on issue slots) and run ahead as far as the data allows.
__device__ __forceinline__ void mbar_init(uint64_t* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"((unsigned)__cvta_generic_to_shared(bar)), "r"(count)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {  // release at CTA scope
  asm volatile("{\n .reg .b64 st;\n mbarrier.arrive.shared::cta.b64 st, [%0];\n}" ::"r"(
                   (unsigned)__cvta_generic_to_shared(bar))
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, unsigned parity) {  // acquire at CTA scope
  const unsigned addr = (unsigned)__cvta_generic_to_shared(bar);
  unsigned done;
  do {
    asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
                 : "=r"(done)
                 : "r"(addr), "r"(parity)
                 : "memory");
  } while (!done);
}
__global__ void __launch_bounds__(1024) k_coarse_invert(int n, const double* __restrict__ Ac_part, double* __restrict__ Ainv) {
  // [n][32 GJ_Q] scaled pivot rows during the elimination, then (after a barrier) [n][n | 1] for the symmetrised write-out
  extern __shared__ double T[];
  __shared__ double s_inv[GJ_SINGLE_MAX];
  __shared__ uint64_t s_bar[GJ_SINGLE_MAX];
  constexpr int W = 32 * GJ_Q;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid < n) mbar_init(&s_bar[tid], 32);
  __syncthreads();
  double a[GJ_Q][GJ_Q], d0[GJ_Q];
#pragma unroll
  for (int qi = 0; qi < GJ_Q; ++qi) {
    d0[qi] = 0.0;
#pragma unroll
    for (int qj = 0; qj < GJ_Q; ++qj) {
      const int i = warp + 32 * qi, j = lane + 32 * qj;
      double v = 0.0;
      if (i < n && j < n) {
#pragma unroll
        for (int sp = 0; sp < COARSE_SPLIT; ++sp) v += Ac_part[(size_t)sp * n * n + (size_t)i * n + j];
      }
      a[qi][qj] = v;
    }
  }
  // original diagonal of the rows this warp owns (every lane gets it): threshold of the pivot test
#pragma unroll
  for (int qi = 0; qi < GJ_Q; ++qi) {
    const int i = warp + 32 * qi;
    double dv = 0.0;
#pragma unroll
    for (int qj = 0; qj < GJ_Q; ++qj)
      if (lane + 32 * qj == i) dv = a[qi][qj];
    d0[qi] = __shfl_sync(0xffffffffu, dv, i & 31);
  }
  // pivots in chunks of 32 so that the register slot PQ of the pivot row / column is a compile-time constant
#pragma unroll
  for (int PQ = 0; PQ < GJ_Q; ++PQ) {
    for (int pw = 0; pw < 32; ++pw) {
      const int p = 32 * PQ + pw;
      if (p >= n) break;
      double* pr = T + (size_t)p * W;
      if (warp == pw) {  // owner of row p: scale and publish it
        const double piv = __shfl_sync(0xffffffffu, a[PQ][PQ], pw);
        const double dp = d0[PQ];
        const double inv = (piv > 1e-13 * dp && dp > 0.0) ? fast_rcp(piv) : 0.0;
#pragma unroll
        for (int qj = 0; qj < GJ_Q; ++qj) pr[lane + 32 * qj] = a[PQ][qj] * inv;
        if (lane == 0) s_inv[p] = inv;
        mbar_arrive(&s_bar[p]);
      }
      double f[GJ_Q];
#pragma unroll
      for (int qi = 0; qi < GJ_Q; ++qi) f[qi] = __shfl_sync(0xffffffffu, a[qi][PQ], pw);  // column p of my rows
      if (warp == pw) f[PQ] = 0.0;  // row p itself is replaced below
      mbar_wait(&s_bar[p], 0);
      const double inv = s_inv[p];
      double rp[GJ_Q];
#pragma unroll
      for (int qj = 0; qj < GJ_Q; ++qj) rp[qj] = pr[lane + 32 * qj];
#pragma unroll
      for (int qi = 0; qi < GJ_Q; ++qi)
#pragma unroll
        for (int qj = 0; qj < GJ_Q; ++qj) a[qi][qj] -= f[qi] * rp[qj];
      if (lane == pw) {  // column p: -f / pivot
#pragma unroll
        for (int qi = 0; qi < GJ_Q; ++qi) a[qi][PQ] = -f[qi] * inv;
      }
      if (warp == pw) {  // row p: the scaled pivot row, 1 / pivot on the diagonal
#pragma unroll
        for (int qj = 0; qj < GJ_Q; ++qj) a[PQ][qj] = rp[qj];
        if (lane == pw) a[PQ][PQ] = inv;
      }
    }
  }
  const int ld = n | 1;
  __syncthreads();  // every warp is through with the pivot rows: the buffer is reused
#pragma unroll
  for (int qi = 0; qi < GJ_Q; ++qi)
#pragma unroll
    for (int qj = 0; qj < GJ_Q; ++qj) {
      const int i = warp + 32 * qi, j = lane + 32 * qj;
      if (i < n && j < n) T[(size_t)i * ld + j] = a[qi][qj];
    }
  __syncthreads();
  for (int i = warp; i < n; i += 32)
    for (int j = lane; j < n; j += 32) Ainv[(size_t)i * n + j] = 0.5 * (T[(size_t)i * ld + j] + T[(size_t)j * ld + i]);
}

// The same elimination for 128 < n <= 224 on a CLUSTER of four CTAs: CTA r owns the row slots {2r, 2r+1} (rows
// w + 32 (2r + qi)), 2 x 7 entries per thread, so every warp still holds whole rows and the pivot column is a
// shuffle.  The owner of a pivot row scales it and stores it into the double-buffered rows of ALL four CTAs through
// distributed shared memory; pivots go in pairs, one cluster barrier per pair.  The raw inverse goes to global memory and
// k_coarse_symmetrize averages it with its transpose.
constexpr int GJC_CTAS = 4, GJC_QR = 2, GJC_QC = 7;
static_assert(GJC_CTAS * GJC_QR * 32 >= 224 && GJC_QC * 32 >= 224, "cluster inverse too small for 32 groups");
__global__ void __cluster_dims__(GJC_CTAS, 1, 1) __launch_bounds__(1024)
    k_coarse_invert_cluster(int n, const double* Ac_part, double* raw) {  // raw may alias Ac_part (see the launcher)
  __shared__ double prow[2][2][32 * GJC_QC];  // [buffer][pivot of the pair][column]
  __shared__ double s_inv[2][2];
  cg::cluster_group cluster = cg::this_cluster();
  const int rank = (int)cluster.block_rank();
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  double a[GJC_QR][GJC_QC], d0[GJC_QR];
#pragma unroll
  for (int qi = 0; qi < GJC_QR; ++qi)
#pragma unroll
    for (int qj = 0; qj < GJC_QC; ++qj) {
      const int i = warp + 32 * (rank * GJC_QR + qi), j = lane + 32 * qj;
      double v = 0.0;
      if (i < n && j < n) {
#pragma unroll
        for (int sp = 0; sp < COARSE_SPLIT; ++sp) v += Ac_part[(size_t)sp * n * n + (size_t)i * n + j];
      }
      a[qi][qj] = v;
    }
#pragma unroll
  for (int qi = 0; qi < GJC_QR; ++qi) {  // original diagonal of the owned rows: threshold of the pivot test
    const int i = warp + 32 * (rank * GJC_QR + qi);
    double dv = 0.0;
#pragma unroll
    for (int qj = 0; qj < GJC_QC; ++qj)
      if (lane + 32 * qj == i) dv = a[qi][qj];
    d0[qi] = __shfl_sync(0xffffffffu, dv, i & 31);
  }
  cluster.sync();  // every CTA has read its inputs (raw may alias them) and is resident
#pragma unroll
  for (int PQ = 0; PQ < GJC_QC; ++PQ) {
    const int owner = PQ / GJC_QR;
    const int lq = PQ % GJC_QR;  // compile-time after unrolling
    // Pivots are taken in PAIRS (p, p + 1: rows of two neighbouring warps of the same CTA) with ONE cluster barrier
    // per pair: the owner of row p leaves its scaled row in its own CTA's buffer first, the owner of row p + 1 (met
    // at a two-warp named barrier) applies pivot p to its row privately, scales it, and both rows go out together.
    // After the barrier every thread runs the two elimination steps exactly as in the one-by-one algorithm.
    for (int pw = 0; pw < 32; pw += 2) {
      const int p = 32 * PQ + pw;
      if (p >= n) break;
      const bool has_q = p + 1 < n;
      const int buf = (p >> 1) & 1;
      const bool mine_p = rank == owner && warp == pw;
      const bool mine_q = has_q && rank == owner && warp == pw + 1;
      if (mine_p) {
        const double piv = __shfl_sync(0xffffffffu, a[lq][PQ], pw);
        const double dp = d0[lq];
        const double inv = (piv > 1e-13 * dp && dp > 0.0) ? fast_rcp(piv) : 0.0;
#pragma unroll
        for (int qj = 0; qj < GJC_QC; ++qj) prow[buf][0][lane + 32 * qj] = a[lq][qj] * inv;
        if (lane == 0) s_inv[buf][0] = inv;
        if (has_q) asm volatile("bar.sync 1, 64;" ::: "memory");  // row p is in place for the next warp
#pragma unroll 1
        for (int r = 0; r < GJC_CTAS; ++r) {  // (not unrolled: the remote addresses would be kept live)
          if (r == rank) continue;
          double* dst = cluster.map_shared_rank(&prow[buf][0][0], r);
#pragma unroll
          for (int qj = 0; qj < GJC_QC; ++qj) dst[lane + 32 * qj] = a[lq][qj] * inv;
          if (lane == 0) *cluster.map_shared_rank(&s_inv[buf][0], r) = inv;
        }
      }
      if (mine_q) {
        asm volatile("bar.sync 1, 64;" ::: "memory");
        const double inv_p = s_inv[buf][0];
        const double fq = __shfl_sync(0xffffffffu, a[lq][PQ], pw);
        double t[GJC_QC];  // row p + 1 after pivot p
#pragma unroll
        for (int qj = 0; qj < GJC_QC; ++qj) t[qj] = a[lq][qj] - fq * prow[buf][0][lane + 32 * qj];
        if (lane == pw) t[PQ] = -fq * inv_p;
        const double piv = __shfl_sync(0xffffffffu, t[PQ], pw + 1);
        const double dp = d0[lq];
        const double inv = (piv > 1e-13 * dp && dp > 0.0) ? fast_rcp(piv) : 0.0;
#pragma unroll 1
        for (int r = 0; r < GJC_CTAS; ++r) {
          double* dst = cluster.map_shared_rank(&prow[buf][1][0], r);
#pragma unroll
          for (int qj = 0; qj < GJC_QC; ++qj) dst[lane + 32 * qj] = t[qj] * inv;
          if (lane == 0) *cluster.map_shared_rank(&s_inv[buf][1], r) = inv;
        }
      }
      cluster.sync();
      auto eliminate = [&](int set, int pl, bool mine) {
        double f[GJC_QR];
#pragma unroll
        for (int qi = 0; qi < GJC_QR; ++qi) f[qi] = __shfl_sync(0xffffffffu, a[qi][PQ], pl);  // pivot column of my rows
        if (mine) f[lq] = 0.0;  // the pivot row itself is replaced below
        const double inv = s_inv[buf][set];
        const double* rp = &prow[buf][set][lane];  // read where used: 2 x 7 entries per thread leave no room for a copy
#pragma unroll
        for (int qj = 0; qj < GJC_QC; ++qj) {
          const double r = rp[32 * qj];
#pragma unroll
          for (int qi = 0; qi < GJC_QR; ++qi) a[qi][qj] -= f[qi] * r;
          if (mine) a[lq][qj] = r;  // pivot row: the scaled row
        }
        if (lane == pl) {  // pivot column: -f / pivot, 1 / pivot on the diagonal
#pragma unroll
          for (int qi = 0; qi < GJC_QR; ++qi) a[qi][PQ] = -f[qi] * inv;
          if (mine) a[lq][PQ] = inv;
        }
      };
      eliminate(0, pw, mine_p);
      if (has_q) eliminate(1, pw + 1, mine_q);
    }
  }
  cluster.sync();  // nobody leaves while a peer may still write its shared memory
#pragma unroll
  for (int qi = 0; qi < GJC_QR; ++qi)
#pragma unroll
    for (int qj = 0; qj < GJC_QC; ++qj) {
      const int i = warp + 32 * (rank * GJC_QR + qi), j = lane + 32 * qj;
      if (i < n && j < n) raw[(size_t)i * n + j] = a[qi][qj];
    }
}

// 224 < n <= 448 (up to 64 groups): the same elimination on a cluster of SIXTEEN CTAs, one row slot per warp (row
// i = warp + 32 rank, 14 column slots per lane), one cluster barrier per pivot, the scaled pivot row stored into all
// sixteen CTAs through distributed shared memory.  ~1 us per pivot: this kernel only ever runs on the side stream of
// the lagged setup (ba_solver.cu), where its latency is hidden behind the stream kernels of the LM iteration.
constexpr int GJB_CTAS = 16, GJB_QC = 14;
static_assert(GJB_CTAS * 32 >= COARSE_MAX_DIM && GJB_QC * 32 >= COARSE_MAX_DIM && GJB_QC <= GJB_CTAS, "big cluster inverse too small");
__global__ void __cluster_dims__(GJB_CTAS, 1, 1) __launch_bounds__(1024)
    k_coarse_invert_big(int n, const double* Ac_part, double* raw) {  // raw may alias Ac_part
  __shared__ double prow[2][32 * GJB_QC];
  __shared__ double s_inv[2];
  cg::cluster_group cluster = cg::this_cluster();
  const int rank = (int)cluster.block_rank();
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int i = warp + 32 * rank;
  double a[GJB_QC];
#pragma unroll
  for (int qj = 0; qj < GJB_QC; ++qj) {
    const int j = lane + 32 * qj;
    double v = 0.0;
    if (i < n && j < n) {
#pragma unroll
      for (int sp = 0; sp < COARSE_SPLIT; ++sp) v += Ac_part[(size_t)sp * n * n + (size_t)i * n + j];
    }
    a[qj] = v;
  }
  double dv = 0.0;  // original diagonal of the own row: threshold of the pivot test
#pragma unroll
  for (int qj = 0; qj < GJB_QC; ++qj)
    if (qj == rank && lane == warp) dv = a[qj];
  const double d0 = __shfl_sync(0xffffffffu, dv, warp);
  cluster.sync();  // every CTA has read its inputs (raw may alias them) and is resident
#pragma unroll
  for (int PQ = 0; PQ < GJB_QC; ++PQ) {  // pivots 32 PQ .. 32 PQ + 31: rows of CTA PQ, column slot PQ (compile time)
    for (int pw = 0; pw < 32; ++pw) {
      const int p = 32 * PQ + pw;
      if (p >= n) break;
      const int buf = p & 1;
      const bool mine = rank == PQ && warp == pw;
      if (mine) {
        const double piv = __shfl_sync(0xffffffffu, a[PQ], pw);
        const double inv = (piv > 1e-13 * d0 && d0 > 0.0) ? fast_rcp(piv) : 0.0;
#pragma unroll 1
        for (int r = 0; r < GJB_CTAS; ++r) {
          double* dst = cluster.map_shared_rank(&prow[buf][0], r);
#pragma unroll
          for (int qj = 0; qj < GJB_QC; ++qj) dst[lane + 32 * qj] = a[qj] * inv;
          if (lane == 0) *cluster.map_shared_rank(&s_inv[buf], r) = inv;
        }
      }
      double f = __shfl_sync(0xffffffffu, a[PQ], pw);  // column p of my row
      if (mine) f = 0.0;
      cluster.sync();
      const double inv = s_inv[buf];
      const double* rp = &prow[buf][lane];
#pragma unroll
      for (int qj = 0; qj < GJB_QC; ++qj) {
        const double r = rp[32 * qj];
        a[qj] -= f * r;
        if (mine) a[qj] = r;
      }
      if (lane == pw) a[PQ] = mine ? inv : -f * inv;
    }
  }
  cluster.sync();  // nobody leaves while a peer may still write its shared memory
#pragma unroll
  for (int qj = 0; qj < GJB_QC; ++qj) {
    const int j = lane + 32 * qj;
    if (i < n && j < n) raw[(size_t)i * n + j] = a[qj];
  }
}

__global__ void __launch_bounds__(256) k_coarse_symmetrize(int n, const double* __restrict__ raw, double* __restrict__ Ainv) {
  const int k = blockIdx.x * 256 + threadIdx.x;
  if (k >= n * n) return;
  const int i = k / n, j = k - i * n;
  Ainv[k] = 0.5 * (raw[k] + raw[(size_t)j * n + i]);
}

}  // namespace

size_t coarse_invert_smem_bytes(int n) { return (size_t)n * 128 * sizeof(double); }  // [n][32 GJ_Q] >= [n][n | 1]

cudaError_t launch_coarse_modes(int C, const double* pose, const uint8_t* free_mask, double* Zc, cudaStream_t st) {
  k_coarse_modes<<<(C + 127) / 128, 128, 0, st>>>(C, pose, free_mask, Zc);
  return cudaGetLastError();
}

cudaError_t launch_coarse_setup(const CoarsePlan& cp, int C, const double* pose, const uint8_t* free_mask,
                                const int32_t* s_row, const int32_t* s_col, const double* S, double* Zc, double* Ac,
                                double* Ainv, cudaStream_t st, cudaEvent_t after_assemble) {
  static bool attr_set = false;
  if (!attr_set) {
    cudaFuncSetAttribute(k_coarse_invert, cudaFuncAttributeMaxDynamicSharedMemorySize,
                         (int)coarse_invert_smem_bytes(GJ_SINGLE_MAX));
    cudaFuncSetAttribute(k_coarse_invert_big, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    attr_set = true;
  }
  const int n = 7 * cp.G;
  k_coarse_modes<<<(C + 127) / 128, 128, 0, st>>>(C, pose, free_mask, Zc);
  k_coarse_assemble<<<dim3(cp.G * cp.G, COARSE_SPLIT), 64 * CA_SLICES, 0, st>>>(cp.G, cp.cb_ptr, cp.cb_fine, s_row, s_col, S, Zc, Ac);
  if (after_assemble) cudaEventRecord(after_assemble, st);  // S and the poses are not read past this point
  if (n <= GJ_SINGLE_MAX) {
    k_coarse_invert<<<1, 1024, coarse_invert_smem_bytes(n), st>>>(n, Ac, Ainv);
  } else {
    // the partial matrices are read into registers before the first cluster barrier: slice 0 is reused for the raw inverse
    if (n <= GJC_CTAS * GJC_QR * 32 && n <= GJC_QC * 32)
      k_coarse_invert_cluster<<<GJC_CTAS, 1024, 0, st>>>(n, Ac, Ac);
    else
      k_coarse_invert_big<<<GJB_CTAS, 1024, 0, st>>>(n, Ac, Ac);
    k_coarse_symmetrize<<<(n * n + 255) / 256, 256, 0, st>>>(n, Ac, Ainv);
  }
  return cudaGetLastError();
}

}  // namespace mpba
