// dense_solve.cu -- direct solve of a SMALL reduced camera system (6 C + 1 <= 128: up to 21 images) by one CTA.
//
// The mapper's local bundle adjustments (mpsfm/sfm/mapper/base.py:442-474) couple 6-20 images: S is a dense
// 36..126-square SPD matrix.  Conjugate gradients on it cost 20-30 dependent iterations of several microseconds
// each (84 us of a 157 us LM iteration at 6 images, round-2 profile); Ceres itself factors systems of this size
// densely.  Here: the augmented matrix [S | rhs] lives IN REGISTERS of one 1024-thread CTA (thread (warp w, lane l)
// owns the entries (w + 32 qi, l + 32 qj), 16 doubles) and is reduced by Gauss-Jordan elimination without pivoting
// (S is SPD: no growth).  There is NO CTA-wide barrier in the elimination: the warp that owns pivot row p publishes
// the scaled row in shared memory (every pivot has its own row: no buffer is ever reused) and arrives on that
// pivot's mbarrier; the other warps wait on it (mbarrier.try_wait suspends them, they do not spin on issue slots)
// and run ahead as far as the data allows -- the chain owner(p) -> owner(p + 1) is the only serialisation left.
// Measured against the one-__syncthreads-per-pivot version of this kernel: 97 -> see profiles/r02_small_bundles.md.
// The elimination is done in place, so the registers end up holding S^-1 next to the solution column, and one step
// of iterative refinement (r = rhs - S y from the block-CSR original, y += S^-1 r) removes the elimination's rounding.
// The register tile is NQ x NQ slots per thread, NQ = ceil((6 C + 1) / 32): 2 for the reference's 6-image local
// bundles.  The camera step that follows every linear solve (k_cam_update: one CTA as well) runs as the epilogue.
#include "ba_kernels.cuh"
#include "step_common.cuh"

namespace mpba {

namespace {

constexpr int DS_DIM = 128;  // up to 4 x 4 register slots: 128 x 128 over 32 warps x 32 lanes
constexpr unsigned FULLMASK = 0xffffffffu;
static_assert(6 * DENSE_SOLVE_MAX_IMAGES + 1 <= DS_DIM, "register tile too small");

struct CamStep {  // arguments of the fused camera step (cam_update_body); pose_cand == nullptr: solve only
  ProblemView pv;
  const double *pose, *Hc, *gc, *c_img;
  LmDiag lm;
  double* pose_cand;
};

__device__ __forceinline__ double ds_rcp(double a) {  // float seed + two Newton steps: 2^-23 -> 2^-46 -> full double
  double x = (double)(1.0f / (float)a);
  x = x * (2.0 - a * x);
  return x * (2.0 - a * x);
}

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

// NW warps; thread (warp w, lane l) owns the entries (w + NW qi, l + 32 qj), qi < RQ, qj < CQ.  Systems up to 63 unknowns
// (10 images: the reference's local bundles) run on 8 warps -- 8 x 2 slots per thread, a quarter of the shuffles and
// shared-memory traffic per pivot of the 32-warp tile; larger ones on 32 warps with a 3 x 3 or 4 x 4 tile.
template <int NW, int RQ, int CQ>
__global__ void __launch_bounds__(32 * NW) k_dense_solve(int C, const int32_t* __restrict__ row_ptr,
                                                      const int32_t* __restrict__ col, const double* __restrict__ S,
                                                      const double* __restrict__ rhs, double* y,
                                                      double* scal, CamStep cs) {
  __shared__ int tbl[DENSE_SOLVE_MAX_IMAGES * DENSE_SOLVE_MAX_IMAGES];
  extern __shared__ double prow[];  // [n][W]: the scaled pivot rows, one per pivot
  __shared__ double s_inv[DS_DIM];
  __shared__ uint64_t s_bar[DS_DIM];  // one mbarrier per pivot: 32 arrivals = the owner warp has published the row
  __shared__ double xs[DS_DIM], rs[DS_DIM];
  constexpr int W = 32 * CQ, NT = 32 * NW;
  __shared__ int s_bad;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int n = 6 * C;  // column n carries the right-hand side
  for (int k = tid; k < C * C; k += NT) tbl[k] = -1;
  if (tid < n) mbar_init(&s_bar[tid], 32);
  if (tid == 0) s_bad = 0;
  __syncthreads();
  for (int i = tid; i < C; i += NT)
    for (int b = row_ptr[i]; b < row_ptr[i + 1]; ++b) tbl[i * C + col[b]] = b;
  __syncthreads();
  double a[RQ][CQ];
#pragma unroll
  for (int qi = 0; qi < RQ; ++qi)
#pragma unroll
    for (int qj = 0; qj < CQ; ++qj) {
      const int i = warp + NW * qi, j = lane + 32 * qj;
      double v = 0.0;
      if (i < n && j < n) {
        const int b = tbl[(i / 6) * C + j / 6];
        if (b >= 0) v = S[(size_t)b * 36 + (i % 6) * 6 + j % 6];
      } else if (i < n && j == n) {
        v = rhs[i];
      }
      a[qi][qj] = v;
    }
#pragma unroll
  for (int PQ = 0; PQ < RQ; ++PQ) {  // pivots in chunks of NW: the register slots of the pivot row / column are static
    constexpr int dummy = 0;
    (void)dummy;
    const int CS = (NW * PQ) >> 5;  // column slot of the chunk's pivots (NW divides 32: one slot per chunk)
    for (int pw = 0; pw < NW; ++pw) {
      const int p = NW * PQ + pw;
      if (p >= n) break;
      const int lp = p & 31;  // lane holding column p
      double* pr = prow + (size_t)p * W;
      if (warp == pw) {  // owner of row p: scale and publish it
        double piv = 0.0;
#pragma unroll
        for (int qj = 0; qj < CQ; ++qj)
          if (qj == CS) piv = __shfl_sync(FULLMASK, a[PQ][qj], lp);
        const bool ok = piv > 0.0 && isfinite(piv);
        const double inv = ok ? ds_rcp(piv) : 0.0;
#pragma unroll
        for (int qj = 0; qj < CQ; ++qj) pr[lane + 32 * qj] = a[PQ][qj] * inv;
        if (lane == 0) {
          s_inv[p] = inv;
          if (!ok) s_bad = 1;  // not positive definite: the LM loop treats the step as invalid
        }
        mbar_arrive(&s_bar[p]);  // every lane releases its own part of the row
      }
      double f[RQ];
#pragma unroll
      for (int qi = 0; qi < RQ; ++qi) {
        f[qi] = 0.0;
#pragma unroll
        for (int qj = 0; qj < CQ; ++qj)
          if (qj == CS) f[qi] = __shfl_sync(FULLMASK, a[qi][qj], lp);  // column p of my rows
      }
      if (warp == pw) f[PQ] = 0.0;  // row p itself is replaced below
      mbar_wait(&s_bar[p], 0);
      const double inv = s_inv[p];
      double rp[CQ];
#pragma unroll
      for (int qj = 0; qj < CQ; ++qj) rp[qj] = pr[lane + 32 * qj];
#pragma unroll
      for (int qi = 0; qi < RQ; ++qi)
#pragma unroll
        for (int qj = 0; qj < CQ; ++qj) a[qi][qj] -= f[qi] * rp[qj];
      if (lane == lp) {  // column p of the inverse: -f / pivot
#pragma unroll
        for (int qi = 0; qi < RQ; ++qi)
#pragma unroll
          for (int qj = 0; qj < CQ; ++qj)
            if (qj == CS) a[qi][qj] = -f[qi] * inv;
      }
      if (warp == pw) {  // row p: the scaled pivot row, 1 / pivot on the diagonal
#pragma unroll
        for (int qj = 0; qj < CQ; ++qj) a[PQ][qj] = (qj == CS && lane == lp) ? inv : rp[qj];
      }
    }
  }
  // solution column -> shared memory
#pragma unroll
  for (int qi = 0; qi < RQ; ++qi)
#pragma unroll
    for (int qj = 0; qj < CQ; ++qj) {
      const int i = warp + NW * qi, j = lane + 32 * qj;
      if (i < n && j == n) xs[i] = a[qi][qj];
    }
  __syncthreads();
  // one refinement step: r = rhs - S x (block-CSR original), x += S^-1 r (the registers)
  if (tid < n) {
    const int bi = tid / 6, r = tid % 6;
    double acc = rhs[tid];
    for (int b = row_ptr[bi]; b < row_ptr[bi + 1]; ++b) {
      const double* Sb = S + (size_t)b * 36 + r * 6;
      const double* xj = xs + 6 * col[b];
#pragma unroll
      for (int c = 0; c < 6; ++c) acc -= Sb[c] * xj[c];
    }
    rs[tid] = acc;
  }
  __syncthreads();
#pragma unroll
  for (int qi = 0; qi < RQ; ++qi) {
    const int i = warp + NW * qi;
    double acc = 0.0;
#pragma unroll
    for (int qj = 0; qj < CQ; ++qj) {
      const int j = lane + 32 * qj;
      if (j < n) acc += a[qi][qj] * rs[j];
    }
    for (int off = 16; off > 0; off >>= 1) acc += __shfl_xor_sync(FULLMASK, acc, off);
    if (i < n && lane == 0) y[i] = s_bad ? 0.0 : xs[i] + acc;
  }
  if (warp == 0) {  // |r| / |rhs| before the refinement, for the summary
    double r2 = 0.0, b2 = 0.0;
    for (int i = lane; i < n; i += 32) {
      r2 += rs[i] * rs[i];
      b2 += rhs[i] * rhs[i];
    }
    for (int off = 16; off > 0; off >>= 1) {
      r2 += __shfl_xor_sync(FULLMASK, r2, off);
      b2 += __shfl_xor_sync(FULLMASK, b2, off);
    }
    if (lane == 0) {
      scal[S_PCG_ITERS] = 0.0;  // a direct solve: no conjugate-gradient iterations
      scal[S_PCG_FLAG] = s_bad ? 2.0 : 0.0;
      scal[S_PCG_RES] = b2 > 0.0 ? sqrt(r2 / b2) : 0.0;
    }
  }
  if (cs.pose_cand != nullptr) {
    __shared__ double sm[NW + 1];
    __shared__ double s_out[3];
    __syncthreads();  // y of this CTA is complete and visible to it
    lm_resolve(cs.lm);
    cam_update_range<NT>(cs.pv, cs.pose, y, cs.Hc, cs.gc, cs.c_img, cs.lm, cs.pose_cand, 0, cs.pv.C, s_out, sm);
    if (tid == 0) {
      scal[S_STEP2_C] = s_out[0];
      scal[S_XNORM2_C] = s_out[1];
      scal[S_MCC_C] = s_out[2];
    }
  }
}

}  // namespace

bool dense_solve_fits(int num_images) { return num_images >= 1 && num_images <= DENSE_SOLVE_MAX_IMAGES; }

cudaError_t launch_dense_solve(int C, const int32_t* row_ptr, const int32_t* col, const double* S, const double* rhs,
                               double* y, double* scal, const ProblemView* pv, const double* pose, const double* Hc,
                               const double* gc, const double* c_img, LmDiag lm, double* pose_cand, cudaStream_t st) {
  CamStep cs = {};
  if (pv != nullptr && pose_cand != nullptr) {
    cs.pv = *pv;
    cs.pose = pose; cs.Hc = Hc; cs.gc = gc; cs.c_img = c_img;
    cs.lm = lm;
    cs.pose_cand = pose_cand;
  }
  const int cols = 6 * C + 1;  // unknowns + the right-hand side
  static bool attr_set = false;
  if (!attr_set) {
    cudaFuncSetAttribute(k_dense_solve<32, 3, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 96 * 8);
    cudaFuncSetAttribute(k_dense_solve<32, 4, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 128 * 128 * 8);
    attr_set = true;
  }
  const int cq = (cols + 31) / 32;
  const size_t smem = (size_t)6 * C * 32 * cq * sizeof(double);  // one published row per pivot: <= 126 x 128 doubles
  if (cols <= 32)
    k_dense_solve<8, 4, 1><<<1, 256, smem, st>>>(C, row_ptr, col, S, rhs, y, scal, cs);
  else if (cols <= 64)
    k_dense_solve<8, 8, 2><<<1, 256, smem, st>>>(C, row_ptr, col, S, rhs, y, scal, cs);
  else if (cols <= 96)
    k_dense_solve<32, 3, 3><<<1, 1024, smem, st>>>(C, row_ptr, col, S, rhs, y, scal, cs);
  else
    k_dense_solve<32, 4, 4><<<1, 1024, smem, st>>>(C, row_ptr, col, S, rhs, y, scal, cs);
  return cudaGetLastError();
}

}  // namespace mpba
