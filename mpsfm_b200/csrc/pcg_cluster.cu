// pcg_cluster.cu -- block-Jacobi PCG on the reduced camera system inside ONE thread-block cluster.
//
// The reduced camera system S y = b of a bundle-adjustment problem with up to a few thousand images
// is small (a few MB of 6x6 blocks) and its conjugate-gradient solve is a chain of ~10^2..10^3
// dependent iterations: it is bound by the latency of one "exchange + reduce" step, not by bandwidth.
// The grid-wide variant (k_pcg in ba_kernels.cu) pays ~5 L2 round trips (~0.35 us each) per barrier;
// here the whole solve runs in a cluster of up to 16 CTAs on one GPC, where an exchange is a
// distributed-shared-memory access (~215 cycles) and a barrier is barrier.cluster (~380 cycles):
//
//   * S is symmetric: only the upper blocks (i <= j) are kept, ONE BLOCK PER THREAD IN REGISTERS
//     (36 doubles) for the whole solve (further blocks of a CTA spill to shared memory).  Per iteration a
//     thread forms S_ij u_j for row i and S_ij^T u_i for row j and deposits both 6-vectors into fixed
//     slots of the row owners' shared memory (local or remote store) -- no atomics, the row sums are
//     taken in slot (= column) order, so the result is deterministic,
//   * the recurrences are the single-reduction form of PCG (Chronopoulos & Gear): one SpMV and ONE
//     fused reduction (gamma = r.u, delta = w.u) per iteration, three cluster barriers in total:
//         A: u published        -> halo gather, block products, deposits
//         B: deposits complete  -> w = S u (row sums), partial gamma / delta to every CTA
//         C: partials complete  -> alpha, beta; p, s, x, r, u = M^-1 r updates (row local)
//   * every CTA computes the same alpha / beta from the same partials in the same order, so the
//     termination decision is uniform without a broadcast,
//   * the coarse term of the preconditioner needs Z^T r of ALL groups on every CTA.  Computing it from the updated r
//     costs a fourth cluster barrier per iteration; instead every CTA keeps Z^T r and Z^T s by the recurrences of r
//     and s (s = w + beta s, r -= alpha s  =>  Z^T s = Z^T w + beta Z^T s, Z^T r -= alpha Z^T s): the partial Z^T w
//     of the owned rows travels with the dot-product partials before barrier C.  (Timed build, cfg3: barrier D cost
//     ~1.9 k of ~18 k cycles per iteration.)
//
// Iterates coincide with textbook PCG in exact arithmetic; the iteration count semantics (number of
// x updates, relative residual sqrt(r'M^-1 r / b'M^-1 b)) are those of k_pcg.
#include <cooperative_groups.h>
#include <stdio.h>
#include <stdlib.h>

#include "ba_kernels.cuh"

namespace cg = cooperative_groups;

namespace mpba {

namespace {

constexpr int T = PCGC_THREADS;
constexpr unsigned FULL = 0xffffffffu;

// Cluster barrier with release / acquire semantics, every thread of the CTA in converged control flow (.aligned).
// cg::cluster_group::sync() is the same pair without .aligned, for which the compiler adds a divergence check per
// warp and a CTA barrier behind the wait; the loop below runs three of them per iteration.
__device__ __forceinline__ void cluster_sync_aligned() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// the two halves, for work that touches only this CTA's own data while the barrier completes
__device__ __forceinline__ void cluster_arrive_aligned() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
}
__device__ __forceinline__ void cluster_wait_aligned() {
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}

// a = S v (row product), t = S^T v2 (column product) of one 6x6 row-major block whose entries 32..35 are read from
// shared memory at their use: 32 + 24 + 12 doubles in flight instead of 36 + 24 + 12 (the compiler spilled the last
// registers of the block to local memory: an L2 round trip per iteration, because every cluster barrier ends with an
// L1 invalidation)
__device__ __forceinline__ void block_products_tail(const double (&S)[32], const double2* tail, const double (&uj)[6],
                                                    const double (&ui)[6], double (&a)[6], double (&t)[6]) {
#pragma unroll
  for (int r = 0; r < 5; ++r) {
    double acc = 0.0;
#pragma unroll
    for (int c = 0; c < 6; ++c) acc += S[6 * r + c] * uj[c];
    a[r] = acc;
  }
#pragma unroll
  for (int c = 0; c < 6; ++c) {
    double acc = 0.0;
#pragma unroll
    for (int r = 0; r < 5; ++r) acc += S[6 * r + c] * ui[r];
    t[c] = acc;
  }
  const double2 s23 = tail[0], s45 = tail[T];  // S[32], S[33] | S[34], S[35]  ([2][threads]: conflict free)
  a[5] = S[30] * uj[0] + S[31] * uj[1] + s23.x * uj[2] + s23.y * uj[3] + s45.x * uj[4] + s45.y * uj[5];
  t[0] += S[30] * ui[5];
  t[1] += S[31] * ui[5];
  t[2] += s23.x * ui[5];
  t[3] += s23.y * ui[5];
  t[4] += s45.x * ui[5];
  t[5] += s45.y * ui[5];
}

__device__ __forceinline__ void store6(double* dst, const double (&v)[6]) {
  double2* d = reinterpret_cast<double2*>(dst);
  d[0] = make_double2(v[0], v[1]);
  d[1] = make_double2(v[2], v[3]);
  d[2] = make_double2(v[4], v[5]);
}
__device__ __forceinline__ void load6(const double* src, double (&v)[6]) {
  const double2* s = reinterpret_cast<const double2*>(src);
  const double2 a = s[0], b = s[1], c = s[2];
  v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y; v[4] = c.x; v[5] = c.y;
}

// u = M^-1 r on the owned rows, r and u in shared memory.  M^-1 = blockdiag(S)^-1 + Z (Z^T S Z)^-1 Z^T (coarse.cu);
// without a coarse space (A.G == 0) only the block-Jacobi term.  Z^T r of all groups comes from a direct restriction
// (first application: one cluster barrier, every CTA must call it at the same point) or from the recurrence kept
// beside r; each CTA then multiplies its own rows of A_c^-1.
#ifdef MPBA_PCG_TIMING
// per-thread cycle sums of the preconditioner's phases: barrier | Z^T r | sync D | A^-1 rows | barrier | combine
#define PRE_ARG , long long (&g_pre_cyc)[6]
#define PRE_PASS , pre_cyc
#define PRE_TICK(k)                                        \
  do {                                                     \
    const long long now_ = clock64();                      \
    g_pre_cyc[k] += now_ - t_;                             \
    t_ = now_;                                             \
  } while (0)
#else
#define PRE_TICK(k)
#define PRE_ARG
#define PRE_PASS
#endif

// Z^T v of the owned groups (v = r or w in shared memory at offset voff), one warp per (group, mode) in fixed order,
// stored into rcall of EVERY CTA of the cluster; complete after the next cluster barrier.
__device__ __forceinline__ void restrict_to_all(double* smem, const int* smi, const PcgClusterArgs& A,
                                                cg::cluster_group& cluster, int rank, int CL, int voff) {
  const PcgClusterCarve& m = A.m;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int ntask = A.gpc * 7;
  for (int task = warp; task < ntask; task += T / 32) {
    const int gl = task / 7, a = task - 7 * gl;
    const int i0 = smi[m.glrp + gl] * 6, i1 = smi[m.glrp + gl + 1] * 6;
    double acc = 0.0, acc1 = 0.0, acc2 = 0.0, acc3 = 0.0;  // independent chains: the loop is shared-memory latency bound
    int idx = i0 + lane;
    for (; idx + 96 < i1; idx += 128) {
      acc += smem[m.zloc + idx * 7 + a] * smem[voff + idx];
      acc1 += smem[m.zloc + (idx + 32) * 7 + a] * smem[voff + idx + 32];
      acc2 += smem[m.zloc + (idx + 64) * 7 + a] * smem[voff + idx + 64];
      acc3 += smem[m.zloc + (idx + 96) * 7 + a] * smem[voff + idx + 96];
    }
    if (idx < i1) acc += smem[m.zloc + idx * 7 + a] * smem[voff + idx];  // the tail on alternating chains
    if (idx + 32 < i1) acc1 += smem[m.zloc + (idx + 32) * 7 + a] * smem[voff + idx + 32];
    if (idx + 64 < i1) acc2 += smem[m.zloc + (idx + 64) * 7 + a] * smem[voff + idx + 64];
    acc = (acc + acc1) + (acc2 + acc3);
    for (int off = 16; off > 0; off >>= 1) acc += __shfl_xor_sync(FULL, acc, off);
    if (lane < CL) {
      double* dst = ((lane == rank) ? smem : cluster.map_shared_rank(smem, lane)) + m.rcall;
      dst[(rank * A.gpc + gl) * 7 + a] = acc;
    }
  }
}

// DIRECT: Z^T r is formed from r (one cluster barrier; the first application of a solve, which also seeds the
// recurrence); otherwise it is read from the recurrence (m.zr) and the call holds no cluster barrier.
template <bool DIRECT>
__device__ __forceinline__ void apply_precond(double* smem, const int* smi, const PcgClusterArgs& A,
                                              cg::cluster_group& cluster, int rank, int CL, int nrows,
                                              int gl_first /* group of row tid / 6 */ PRE_ARG) {
  const PcgClusterCarve& m = A.m;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
#ifdef MPBA_PCG_TIMING
  long long t_ = clock64();
#endif
  __syncthreads();  // r is complete
  PRE_TICK(0);
  if (A.G) {
    const int ntask = A.gpc * 7;
    if (DIRECT) {
      restrict_to_all(smem, smi, A, cluster, rank, CL, m.r);
      PRE_TICK(1);
      cluster_sync_aligned();  // D: Z^T r of every group is everywhere
      PRE_TICK(2);
      for (int k = tid; k < A.nc; k += T) {
        smem[m.zr + k] = smem[m.rcall + k];
        smem[m.zs + k] = 0.0;
      }
    }
    const int zsrc = DIRECT ? m.rcall : m.zr;
    for (int task = warp; task < ntask; task += T / 32) {
      double acc = 0.0, acc1 = 0.0, acc2 = 0.0, acc3 = 0.0;  // four chains: the dot is shared-memory latency bound
      const double* arow = smem + m.ainv + task * A.nc;
      int k = lane;
      for (; k + 96 < A.nc; k += 128) {
        acc += arow[k] * smem[zsrc + k];
        acc1 += arow[k + 32] * smem[zsrc + k + 32];
        acc2 += arow[k + 64] * smem[zsrc + k + 64];
        acc3 += arow[k + 96] * smem[zsrc + k + 96];
      }
      if (k < A.nc) acc += arow[k] * smem[zsrc + k];
      if (k + 32 < A.nc) acc1 += arow[k + 32] * smem[zsrc + k + 32];
      if (k + 64 < A.nc) acc2 += arow[k + 64] * smem[zsrc + k + 64];
      acc = (acc + acc1) + (acc2 + acc3);
      for (int off = 16; off > 0; off >>= 1) acc += __shfl_xor_sync(FULL, acc, off);
      if (lane == 0) smem[m.ec + task] = acc;
    }
    PRE_TICK(3);
    __syncthreads();
    PRE_TICK(4);
  }
  for (int q = tid; q < nrows * 6; q += T) {
    const int lr = q / 6, c = q - 6 * lr;
    double z = 0.0, z1 = 0.0, zc0 = 0.0, zc1 = 0.0;
#pragma unroll
    for (int j = 0; j < 6; j += 2) {
      z += smem[m.minv + lr * 36 + c * 6 + j] * smem[m.r + lr * 6 + j];
      z1 += smem[m.minv + lr * 36 + c * 6 + j + 1] * smem[m.r + lr * 6 + j + 1];
    }
    if (A.G) {
      int gl = gl_first;  // looked up once per solve for the first pass (a chain of dependent loads otherwise)
      if (q != tid) {
        gl = 0;
        while (lr >= smi[m.glrp + gl + 1]) ++gl;
      }
#pragma unroll
      for (int a = 0; a < 6; a += 2) {
        zc0 += smem[m.zloc + q * 7 + a] * smem[m.ec + gl * 7 + a];
        zc1 += smem[m.zloc + q * 7 + a + 1] * smem[m.ec + gl * 7 + a + 1];
      }
      zc0 += smem[m.zloc + q * 7 + 6] * smem[m.ec + gl * 7 + 6];
    }
    smem[m.u + q] = (z + z1) + (zc0 + zc1);
  }
  PRE_TICK(5);
}

// row items: warp w, lane l < 30 -> row slot 5 w + l / 6 (+ 5 * warps per pass), component l % 6, so the six
// components of a row sit in one warp and M^-1 r is a shuffle product
constexpr int ROWS_PER_PASS = 5 * (T / 32);

__global__ void __launch_bounds__(T, 1) k_pcg_cluster(PcgClusterArgs A) {
  extern __shared__ __align__(16) double smem[];
  cg::cluster_group cluster = cg::this_cluster();
  const int rank = (int)cluster.block_rank();
  const int CL = (int)cluster.num_blocks();
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const PcgClusterCarve& m = A.m;  // offsets sit in the constant bank: no registers
  int* const smi = reinterpret_cast<int*>(smem);

  const int row0 = A.cta_row_ptr[rank], row1 = A.cta_row_ptr[rank + 1];
  const int nrows = row1 - row0;
  const int ub0 = A.cta_blk_ptr[rank], nub = A.cta_blk_ptr[rank + 1] - ub0;
  const int h0 = A.cta_halo_ptr[rank], nhalo = A.cta_halo_ptr[rank + 1] - h0;
  const int full0 = A.row_ptr[row0];

  // ---- one-time loads: my register block, spilled blocks, diagonal blocks, M^-1 rows, r = b
  double Sreg[32];
  double2* const stail = reinterpret_cast<double2*>(smem + m.stail) + tid;
  const bool have = tid < nub;
  if (have) {
    const double2* src = reinterpret_cast<const double2*>(A.S + (size_t)A.ub_sblk[ub0 + tid] * 36);
#pragma unroll
    for (int e = 0; e < 16; ++e) {
      const double2 v = src[e];
      Sreg[2 * e] = v.x;
      Sreg[2 * e + 1] = v.y;
    }
    stail[0] = src[16];
    stail[T] = src[17];
  } else {
#pragma unroll
    for (int e = 0; e < 32; ++e) Sreg[e] = 0.0;
    stail[0] = stail[T] = make_double2(0.0, 0.0);
  }
  const int nextra = nub > T ? nub - T : 0;
  for (int k = tid; k < nextra * 36; k += T) {
    const int b = k / 36, e = k % 36;
    smem[m.extra + (size_t)e * A.max_extra + b] = A.S[(size_t)A.ub_sblk[ub0 + T + b] * 36 + e];
  }
  for (int k = tid; k < nrows * 36; k += T) {
    smem[m.minv + k] = A.Minv[(size_t)row0 * 36 + k];
    smem[m.sdiag + k] = A.S[(size_t)A.s_diag[row0 + k / 36] * 36 + k % 36];
  }
  for (int k = tid; k <= nrows; k += T) smi[m.rp + k] = A.row_ptr[row0 + k] - full0;
  for (int k = tid; k < nhalo; k += T) smi[m.hdesc + k] = (A.halo_cta[h0 + k] << 20) | A.halo_rowl[h0 + k];
  for (int k = tid; k < nub; k += T) {  // metadata of register blocks (k < T) and spilled blocks alike
    const int g = ub0 + k;
    smi[m.xmeta + 5 * k] = A.ub_rowl[g];
    smi[m.xmeta + 5 * k + 1] = A.ub_halo[g];
    smi[m.xmeta + 5 * k + 2] = A.ub_depA[g];
    smi[m.xmeta + 5 * k + 3] = A.ub_tcta[g];
    smi[m.xmeta + 5 * k + 4] = A.ub_tidx[g];
  }
  for (int k = tid; k < A.max_full * 6; k += T) smem[m.dep + k] = 0.0;  // the diagonal slots stay zero
  for (int k = tid; k < nrows * 6; k += T) {
    smem[m.r + k] = A.rhs[(size_t)row0 * 6 + k];
    smem[m.x + k] = 0.0;
    smem[m.p + k] = 0.0;
    smem[m.s + k] = 0.0;
  }
  if (A.G) {
    for (int k = tid; k < nrows * 42; k += T) smem[m.zloc + k] = A.Zc[(size_t)row0 * 42 + k];
    for (int k = tid; k < A.gpc * 7 * A.nc; k += T) smem[m.ainv + k] = A.Ainv[(size_t)rank * A.gpc * 7 * A.nc + k];
    for (int k = tid; k <= A.gpc; k += T) smi[m.glrp + k] = A.grp_row_ptr[rank * A.gpc + k] - row0;
  }
  const int sub = lane / 6, comp = lane % 6;
  const bool row_lane = lane < 30;
#ifdef MPBA_PCG_TIMING
  long long pre_cyc[6] = {0, 0, 0, 0, 0, 0};
  long long upd_cyc[3] = {0, 0, 0};  // fold + alpha | x r p s | zs zr
  long long row_cyc[3] = {0, 0, 0};  // row sums + warp reduce | barrier | restriction of w
#endif
  int gl_first = 0;
  __syncthreads();  // the one-time loads above (glrp among them) are in place
  if (A.G && tid < nrows * 6)
    while (tid / 6 >= smi[m.glrp + gl_first + 1]) ++gl_first;
  apply_precond<true>(smem, smi, A, cluster, rank, CL, nrows, gl_first PRE_PASS);

  // The scalars carried from one iteration to the next live in shared memory (written by thread 0, CTA barriers lie
  // between every write and the reads): as registers they were spilled to local memory, and every cluster barrier
  // ends with an L1 invalidation (CCTL.IVALL is how barrier.cluster.wait.acquire compiles) -- the reload of
  // 1 / gamma_prev sat on the critical path of every iteration as an L2 round trip.
  enum { SC_GAMMA0 = 0, SC_TOL2G0, SC_INV_GAMMA, SC_INV_ALPHA, SC_LAST_GAMMA };
  if (tid == 0)
    for (int e = 0; e < 8; ++e) smem[m.sc + e] = 0.0;
  int iters = 0, flag = 1;
#ifdef MPBA_PCG_TIMING
  long long tA = 0, tH = 0, tP = 0, tB = 0, tR = 0, tC = 0, tU = 0;
#define TICK(v) const long long v = clock64()
#else
#define TICK(v)
#endif
  for (int k = 0;; ++k) {
    TICK(c0);
    cluster_sync_aligned();  // A: every CTA's u is in place
    TICK(c1);
    // halo: the u rows this CTA's blocks multiply with, pulled once per iteration
    for (int q = tid; q < nhalo * 3; q += T) {
      const int hs = q / 3, part = q % 3;
      const int hd = smi[m.hdesc + hs];
      const int owner = hd >> 20, rl = hd & 0xfffff;
      const double* src = ((owner == rank) ? smem : cluster.map_shared_rank(smem, owner)) + m.u;
      reinterpret_cast<double2*>(smem + m.halo + (size_t)hs * 6)[part] =
          reinterpret_cast<const double2*>(src + (size_t)rl * 6)[part];
    }
    __syncthreads();
    TICK(c2);
    // off-diagonal block products and deposits: S_ij u_j -> row i (local), S_ij^T u_i -> row j
    if (have) {
      double uj[6], ui[6], a[6], t[6];
      const int* md = smi + m.xmeta + 5 * tid;
      load6(smem + m.halo + md[1] * 6, uj);
      load6(smem + m.u + md[0] * 6, ui);
      block_products_tail(Sreg, stail, uj, ui, a, t);
      store6(smem + m.dep + md[2] * 6, a);
      const int tc = md[3];
      double* dst = ((tc == rank) ? smem : cluster.map_shared_rank(smem, tc)) + m.dep;
      store6(dst + md[4] * 6, t);
    }
    for (int b = tid; b < nextra; b += T) {
      double uj[6], ui[6], a[6], t[6];
      const int* md = smi + m.xmeta + 5 * (T + b);
      load6(smem + m.halo + md[1] * 6, uj);
      load6(smem + m.u + md[0] * 6, ui);
#pragma unroll
      for (int e = 0; e < 6; ++e) a[e] = t[e] = 0.0;
#pragma unroll
      for (int r = 0; r < 6; ++r)
#pragma unroll
        for (int cc = 0; cc < 6; ++cc) {
          const double sv = smem[m.extra + (size_t)(6 * r + cc) * A.max_extra + b];  // one read feeds both products
          a[r] += sv * uj[cc];
          t[cc] += sv * ui[r];
        }
      store6(smem + m.dep + md[2] * 6, a);
      const int tc = md[3];
      double* dst = ((tc == rank) ? smem : cluster.map_shared_rank(smem, tc)) + m.dep;
      store6(dst + md[4] * 6, t);
    }
    TICK(c3);
    cluster_arrive_aligned();  // B: my deposits are on their way ...
    // ... and while the others' land: the part of the first pass of row sums that needs only this CTA's own data
    // (diagonal block times u, r.u, the row's slot range)
    double dg_pre = 0.0, uu_pre = 0.0, ru_pre = 0.0;
    int s0_pre = 0, s1_pre = 0;
    {
      const int lr = 5 * warp + sub;
      if (row_lane && lr < nrows) {
        s0_pre = smi[m.rp + lr];
        s1_pre = smi[m.rp + lr + 1];
#pragma unroll
        for (int j = 0; j < 6; ++j) dg_pre += smem[m.sdiag + lr * 36 + comp * 6 + j] * smem[m.u + lr * 6 + j];
        uu_pre = smem[m.u + lr * 6 + comp];
        ru_pre = smem[m.r + lr * 6 + comp] * uu_pre;
      }
    }
    cluster_wait_aligned();  // B: all deposits landed
    TICK(c4);
    // w = S u on the owned rows (diagonal block from shared memory + deposits in column order); dot partials
    double pg = 0.0, pd = 0.0;
    for (int lr0 = 0; lr0 < nrows; lr0 += ROWS_PER_PASS) {
      const int lr = lr0 + 5 * warp + sub;
      if (row_lane && lr < nrows) {
        const bool first = lr0 == 0;
        const int s0 = first ? s0_pre : smi[m.rp + lr], s1 = first ? s1_pre : smi[m.rp + lr + 1];
        double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
        int sl = s0;
        {
          // eight slots per round on eight chains (a row has ~27 deposits: the sum is a chain of dependent adds)
          double b0 = 0.0, b1 = 0.0, b2 = 0.0, b3 = 0.0;
          const double* dp = smem + m.dep + comp;
          for (; sl + 7 < s1; sl += 8) {
            a0 += dp[(size_t)sl * 6];
            a1 += dp[(size_t)(sl + 1) * 6];
            a2 += dp[(size_t)(sl + 2) * 6];
            a3 += dp[(size_t)(sl + 3) * 6];
            b0 += dp[(size_t)(sl + 4) * 6];
            b1 += dp[(size_t)(sl + 5) * 6];
            b2 += dp[(size_t)(sl + 6) * 6];
            b3 += dp[(size_t)(sl + 7) * 6];
          }
          a0 += b0; a1 += b1; a2 += b2; a3 += b3;
        }
        for (; sl + 3 < s1; sl += 4) {
          a0 += smem[m.dep + (size_t)sl * 6 + comp];
          a1 += smem[m.dep + (size_t)(sl + 1) * 6 + comp];
          a2 += smem[m.dep + (size_t)(sl + 2) * 6 + comp];
          a3 += smem[m.dep + (size_t)(sl + 3) * 6 + comp];
        }
        if (sl < s1) a0 += smem[m.dep + (size_t)sl * 6 + comp];
        if (sl + 1 < s1) a1 += smem[m.dep + (size_t)(sl + 1) * 6 + comp];
        if (sl + 2 < s1) a2 += smem[m.dep + (size_t)(sl + 2) * 6 + comp];
        double dg = dg_pre, uu = uu_pre, ru = ru_pre;
        if (!first) {
          dg = 0.0;
#pragma unroll
          for (int j = 0; j < 6; ++j) dg += smem[m.sdiag + lr * 36 + comp * 6 + j] * smem[m.u + lr * 6 + j];
          uu = smem[m.u + lr * 6 + comp];
          ru = smem[m.r + lr * 6 + comp] * uu;
        }
        const double acc = ((a0 + a1) + (a2 + a3)) + dg;
        smem[m.w + lr * 6 + comp] = acc;
        pg += ru;
        pd += acc * uu;
      }
    }
    for (int off = 16; off > 0; off >>= 1) {
      pg += __shfl_down_sync(FULL, pg, off);
      pd += __shfl_down_sync(FULL, pd, off);
    }
    if (lane == 0) reinterpret_cast<double2*>(smem + m.red)[warp] = make_double2(pg, pd);
#ifdef MPBA_PCG_TIMING
    const long long r1 = clock64();
    row_cyc[0] += r1 - c4;
#endif
    __syncthreads();  // w and the warp partials are complete
#ifdef MPBA_PCG_TIMING
    const long long r2 = clock64();
    row_cyc[1] += r2 - r1;
#endif
    if (A.G) restrict_to_all(smem, smi, A, cluster, rank, CL, m.w);  // Z^T w of the owned groups to every CTA
#ifdef MPBA_PCG_TIMING
    row_cyc[2] += clock64() - r2;
#endif
    // the LAST warp folds the warp partials (fixed tree) and lane k delivers the result to CTA k: it owns no rows and
    // no restriction task at the cluster sizes in use, so the fold runs beside the restrictions instead of behind one
    if (warp == T / 32 - 1) {
      double2 v = lane < T / 32 ? reinterpret_cast<const double2*>(smem + m.red)[lane] : make_double2(0.0, 0.0);
      for (int off = 8; off > 0; off >>= 1) {
        v.x += __shfl_down_sync(FULL, v.x, off);
        v.y += __shfl_down_sync(FULL, v.y, off);
      }
      v.x = __shfl_sync(FULL, v.x, 0);
      v.y = __shfl_sync(FULL, v.y, 0);
      if (lane < CL) {
        double* dst = ((lane == rank) ? smem : cluster.map_shared_rank(smem, lane)) + m.parts;
        reinterpret_cast<double2*>(dst)[rank] = v;
      }
    }
    TICK(c5);
    cluster_sync_aligned();  // C: partials of every CTA are here
    TICK(c6);
#ifdef MPBA_PCG_TIMING
    tA += c1 - c0; tH += c2 - c1; tP += c3 - c2; tB += c4 - c3; tR += c5 - c4; tC += c6 - c5;
#endif
    double gamma = 0.0, delta = 0.0;
    {
      double2 v = lane < CL ? reinterpret_cast<const double2*>(smem + m.parts)[lane] : make_double2(0.0, 0.0);
      for (int off = 8; off > 0; off >>= 1) {
        v.x += __shfl_down_sync(FULL, v.x, off);
        v.y += __shfl_down_sync(FULL, v.y, off);
      }
      gamma = __shfl_sync(FULL, v.x, 0);
      delta = __shfl_sync(FULL, v.y, 0);
    }
    if (k == 0) {
      if (tid == 0) {
        smem[m.sc + SC_GAMMA0] = gamma;
        smem[m.sc + SC_TOL2G0] = A.tol * A.tol * gamma;
      }
      if (!(gamma > 0.0) || !isfinite(gamma)) {
        flag = (gamma == 0.0) ? 0 : 2;
        break;  // SC_LAST_GAMMA stays 0
      }
    }
    if (tid == 0) smem[m.sc + SC_LAST_GAMMA] = gamma;
    if (!isfinite(gamma)) {
      flag = 2;
      break;
    }
    iters = k;
    if (k > 0 && fabs(gamma) <= smem[m.sc + SC_TOL2G0]) {
      flag = 0;
      break;
    }
    if (k >= A.max_iter) {
      flag = 1;
      break;
    }
    // beta = gamma / gamma_prev, alpha = gamma / (delta - beta gamma / alpha_prev): one division on the
    // critical path, the reciprocals for the next iteration are taken off it
    const double beta = gamma * smem[m.sc + SC_INV_GAMMA];
    const double denom = delta - beta * gamma * smem[m.sc + SC_INV_ALPHA];
    if (!(denom > 0.0) || !isfinite(denom)) {
      flag = 2;
      break;
    }
    const double alpha = gamma / denom;
#ifdef MPBA_PCG_TIMING
    const long long u1 = clock64();
    upd_cyc[0] += u1 - c6;
#endif
    for (int lr0 = 0; lr0 < nrows; lr0 += ROWS_PER_PASS) {
      const int lr = lr0 + 5 * warp + sub;
      const bool act = row_lane && lr < nrows;
      if (act) {
        const int o = lr * 6 + comp;
        const double pn = smem[m.u + o] + beta * smem[m.p + o];
        const double sn = smem[m.w + o] + beta * smem[m.s + o];
        smem[m.p + o] = pn;
        smem[m.s + o] = sn;
        smem[m.x + o] += alpha * pn;
        smem[m.r + o] -= alpha * sn;
      }
    }
#ifdef MPBA_PCG_TIMING
    const long long u2 = clock64();
    upd_cyc[1] += u2 - u1;
#endif
    // the same recurrences for the coarse restrictions (identical on every CTA): Z^T s, Z^T r
    for (int t = tid; t < A.nc; t += T) {
      const double zsn = smem[m.rcall + t] + beta * smem[m.zs + t];
      smem[m.zs + t] = zsn;
      smem[m.zr + t] -= alpha * zsn;
    }
#ifdef MPBA_PCG_TIMING
    upd_cyc[2] += clock64() - u2;
#endif
    apply_precond<false>(smem, smi, A, cluster, rank, CL, nrows, gl_first PRE_PASS);
    if (tid == 0) {  // every thread is past its reads of the two (CTA barriers inside apply_precond)
      const double ig = 1.0 / gamma;
      smem[m.sc + SC_INV_GAMMA] = ig;
      smem[m.sc + SC_INV_ALPHA] = denom * ig;
    }
#ifdef MPBA_PCG_TIMING
    tU += clock64() - c6;
#endif
  }
  for (int q = tid; q < nrows * 6; q += T) A.x[(size_t)row0 * 6 + q] = smem[m.x + q];
  if (rank == 0 && tid == 0) {
    A.scal[S_PCG_ITERS] = (double)iters;
    A.scal[S_PCG_FLAG] = (double)flag;
    const double gamma0 = smem[m.sc + SC_GAMMA0], last_gamma = smem[m.sc + SC_LAST_GAMMA];
    A.scal[S_PCG_RES] = gamma0 > 0.0 ? sqrt(fabs(last_gamma) / gamma0) : 0.0;
#ifdef MPBA_PCG_TIMING
    printf("[pcg precond cycles / iteration] barrier %lld  ZTr %lld  syncD %lld  Ainv %lld  barrier %lld  combine %lld  | update total %lld (x/r/p/s part %lld)\n",
           pre_cyc[0] / (iters + 1), pre_cyc[1] / (iters + 1), pre_cyc[2] / (iters + 1), pre_cyc[3] / (iters + 1),
           pre_cyc[4] / (iters + 1), pre_cyc[5] / (iters + 1), tU / (iters + 1),
           (tU - pre_cyc[0] - pre_cyc[1] - pre_cyc[2] - pre_cyc[3] - pre_cyc[4] - pre_cyc[5]) / (iters + 1));
    printf("[pcg update cycles / iteration] fold+alpha %lld  xrps %lld  zszr %lld | rowsum %lld  barrier %lld  restrict %lld\n",
           upd_cyc[0] / (iters + 1), upd_cyc[1] / (iters + 1), upd_cyc[2] / (iters + 1), row_cyc[0] / (iters + 1),
           row_cyc[1] / (iters + 1), row_cyc[2] / (iters + 1));
    A.scal[S_PCG_CYC_SPMV] = (double)(tH + tP);
    A.scal[S_PCG_CYC_REDUCE] = (double)(tA + tB + tC);
    A.scal[S_PCG_CYC_UPDATE] = (double)(tR + tU);
    A.scal[S_PCG_CYC_DETAIL] = (double)tA;
    A.scal[S_PCG_CYC_DETAIL + 1] = (double)tH;
    A.scal[S_PCG_CYC_DETAIL + 2] = (double)tP;
    A.scal[S_PCG_CYC_DETAIL + 3] = (double)tB;
    A.scal[S_PCG_CYC_DETAIL + 4] = (double)tR;
    A.scal[S_PCG_CYC_DETAIL + 5] = (double)tC;
    A.scal[S_PCG_CYC_DETAIL + 6] = (double)tU;
#else
    for (int e = 0; e < 3; ++e) A.scal[S_PCG_CYC_SPMV + e] = 0.0;
#endif
  }
  cluster.sync();  // no CTA may exit while a peer can still touch its shared memory
}

}  // namespace

size_t pcg_cluster_carve(PcgClusterArgs& A) {
  PcgClusterCarve& c = A.m;
  int o = 0;
  c.dep = o;   o += A.max_full * 6;
  c.x = o;     o += A.max_rows * 6;
  c.r = o;     o += A.max_rows * 6;
  c.u = o;     o += A.max_rows * 6;
  c.p = o;     o += A.max_rows * 6;
  c.s = o;     o += A.max_rows * 6;
  c.w = o;     o += A.max_rows * 6;
  c.minv = o;  o += A.max_rows * 36;
  c.sdiag = o; o += A.max_rows * 36;
  c.halo = o;  o += A.max_halo * 6;
  c.extra = o; o += A.max_extra * 36;
  c.parts = o; o += 2 * 16;
  c.red = o;   o += 2 * (T / 32) + 2;
  c.zloc = o;  o += A.G ? A.max_rows * 42 : 0;
  c.ainv = o;  o += A.gpc * 7 * A.nc;
  c.rcall = o; o += A.nc;
  c.ec = o;    o += A.gpc * 7;
  c.zs = o;    o += A.nc;
  c.zr = o;    o += A.nc;
  c.sc = o;    o += 8;
  o += o & 1;  // double2 access
  c.stail = o; o += 4 * T;
  o += o & 1;
  c.glrp = 2 * o;
  o += (A.gpc + 4) / 2;
  c.rp = 2 * o;
  c.hdesc = c.rp + A.max_rows + 2;
  c.xmeta = c.hdesc + A.max_halo;
  const size_t ints = (size_t)c.xmeta + 5 * ((size_t)T + A.max_extra) + 2;
  return ints * sizeof(int) + 16;
}

int pcg_cluster_max_smem(int device) {
  static int optin = 0;
  if (!optin) {
    cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, device);
    cudaFuncSetAttribute(k_pcg_cluster, cudaFuncAttributeMaxDynamicSharedMemorySize, optin);
    cudaFuncSetAttribute(k_pcg_cluster, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
  }
  return optin;
}

// can a cluster of `cl` CTAs with this much shared memory be co-scheduled on this device?
bool pcg_cluster_fits(int device, int cl, size_t smem_bytes) {
  if ((int)smem_bytes > pcg_cluster_max_smem(device)) return false;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(cl);
  cfg.blockDim = dim3(T);
  cfg.dynamicSmemBytes = smem_bytes;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = cl;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  int n = 0;
  const cudaError_t e = cudaOccupancyMaxActiveClusters(&n, k_pcg_cluster, &cfg);
  if (e != cudaSuccess) {
    if (getenv("MPBA_DEBUG")) fprintf(stderr, "[mpba] cluster of %d, %zu B smem: %s\n", cl, smem_bytes, cudaGetErrorString(e));
    cudaGetLastError();
    return false;
  }
  if (getenv("MPBA_DEBUG")) fprintf(stderr, "[mpba] cluster of %d, %zu B smem: %d active clusters\n", cl, smem_bytes, n);
  return n >= 1;
}

cudaError_t launch_pcg_cluster(const PcgClusterArgs& a, int cl, size_t smem_bytes, cudaStream_t st) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(cl);
  cfg.blockDim = dim3(T);
  cfg.dynamicSmemBytes = smem_bytes;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = cl;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, k_pcg_cluster, a);
}

}  // namespace mpba
