// step_common.cuh -- device helpers shared by the kernels of ba_kernels.cu and dense_solve.cu.
#pragma once

#include "ba_kernels.cuh"

namespace mpba {

constexpr unsigned FULL = 0xffffffffu;

// index of (i, j), i <= j, in the 21 stored entries of a symmetric 6 x 6 block
__device__ __forceinline__ int upper_idx(int i, int j) { return i * (13 - i) / 2 + (j - i); }

// deterministic block sum (fixed shuffle tree, then warps in order); result valid in every thread
template <int NWARP>
__device__ __forceinline__ double block_sum(double v, double* sm /* [NWARP+1] */) {
  for (int off = 16; off > 0; off >>= 1) v += __shfl_down_sync(FULL, v, off);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __syncthreads();
  if (lane == 0) sm[warp] = v;
  __syncthreads();
  if (threadIdx.x == 0) {
    double s = 0.0;
    for (int w = 0; w < NWARP; ++w) s += sm[w];
    sm[NWARP] = s;
  }
  __syncthreads();
  return sm[NWARP];
}
template <int NWARP>
__device__ __forceinline__ double block_max(double v, double* sm) {
  for (int off = 16; off > 0; off >>= 1) v = fmax(v, __shfl_down_sync(FULL, v, off));
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __syncthreads();
  if (lane == 0) sm[warp] = v;
  __syncthreads();
  if (threadIdx.x == 0) {
    double s = 0.0;
    for (int w = 0; w < NWARP; ++w) s = fmax(s, sm[w]);
    sm[NWARP] = s;
  }
  __syncthreads();
  return sm[NWARP];
}

// device-controlled LM loop: the trust-region radius lives in device memory (see LmDiag)
__device__ __forceinline__ void lm_resolve(LmDiag& lm) {
  if (lm.dev != nullptr) {
    lm.radius = __ldcg(lm.dev + LD_RADIUS);
    lm.inv_radius = __ldcg(lm.dev + LD_INV_RADIUS);
  }
}

// Camera step of one LM iteration for the images [ibeg, iend) by one CTA of NT threads: q (+) -y, t - y, step norms,
// image part of the model cost change (g'y + y'Dy) / 2.  out[0..2] = {step^2, |x|^2, model cost change} (thread 0).
// Used by k_cam_update (one CTA, all images), the epilogue of the dense direct solve (dense_solve.cu) and the extra
// CTAs of k_back_substitute (128 images each; the slices are added by k_pt_scalars_step).
template <int NT>
__device__ __forceinline__ void cam_update_range(const ProblemView& pv, const double* __restrict__ pose,
                                                 const double* y /* may have been written by this kernel */,
                                                 const double* __restrict__ Hc, const double* __restrict__ gc,
                                                 const double* __restrict__ c_img, const LmDiag& lm,
                                                 double* __restrict__ pose_cand, int ibeg, int iend, double* out,
                                                 double* sm /* [NT / 32 + 1] */) {
  double step2 = 0.0, xn2 = 0.0, mcc = 0.0;
  for (int i = ibeg + threadIdx.x; i < iend; i += NT) {
    const double* ps = pose + (size_t)i * 8;
    const double* yi = y + (size_t)i * 6;
    const unsigned mask = pv.img_free_mask[i];
    // step norms of a sharded solve are summed over the ranks: every image counts at its owner only
    const bool used = i >= pv.i0 && i < pv.i1 && pv.img_ptr[i + 1] > pv.img_ptr[i];
    const double q[4] = {ps[0], ps[1], ps[2], ps[3]};
    const double d[3] = {-yi[0], -yi[1], -yi[2]};
    double qn[4];
    quat_plus(q, d, qn);
    double* out = pose_cand + (size_t)i * 8;
    out[0] = qn[0]; out[1] = qn[1]; out[2] = qn[2]; out[3] = qn[3];
    const double t0 = ps[4] - yi[3], t1 = ps[5] - yi[4], t2 = ps[6] - yi[5];
    out[4] = t0; out[5] = t1; out[6] = t2; out[7] = ps[7];
    if (used && (mask & 7u)) {
      step2 += (qn[0] - q[0]) * (qn[0] - q[0]) + (qn[1] - q[1]) * (qn[1] - q[1]) + (qn[2] - q[2]) * (qn[2] - q[2]) +
               (qn[3] - q[3]) * (qn[3] - q[3]);
      xn2 += q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3];
    }
    if (used && (mask & 0x38u)) {
      step2 += (t0 - ps[4]) * (t0 - ps[4]) + (t1 - ps[5]) * (t1 - ps[5]) + (t2 - ps[6]) * (t2 - ps[6]);
      xn2 += ps[4] * ps[4] + ps[5] * ps[5] + ps[6] * ps[6];
    }
    // image part of the model cost change (g'y + y'Dy) / 2 over the free tangent coordinates (all images,
    // replicated in a sharded solve: gc / Hc are complete on every rank)
#pragma unroll
    for (int r = 0; r < 6; ++r) {
      if ((mask >> r) & 1u) {
        const double dj = lm_damping(Hc[(size_t)i * 21 + upper_idx(r, r)], c_img[(size_t)i * 6 + r], lm.radius, lm.lo, lm.hi);
        mcc += 0.5 * (gc[(size_t)i * 6 + r] * yi[r] + dj * yi[r] * yi[r]);
      }
    }
  }
  step2 = block_sum<NT / 32>(step2, sm);
  xn2 = block_sum<NT / 32>(xn2, sm);
  mcc = block_sum<NT / 32>(mcc, sm);
  if (threadIdx.x == 0) {
    out[0] = step2;
    out[1] = xn2;
    out[2] = mcc;
  }
}

// all images by one CTA of 1024 threads, sums straight into the scalar block
__device__ __forceinline__ void cam_update_body(const ProblemView& pv, const double* __restrict__ pose, const double* y,
                                                const double* __restrict__ Hc, const double* __restrict__ gc,
                                                const double* __restrict__ c_img, const LmDiag& lm,
                                                double* __restrict__ pose_cand, double* scal, double* sm /* [33] */) {
  __shared__ double s_out[3];
  cam_update_range<1024>(pv, pose, y, Hc, gc, c_img, lm, pose_cand, 0, pv.C, s_out, sm);
  if (threadIdx.x == 0) {
    scal[S_STEP2_C] = s_out[0];
    scal[S_XNORM2_C] = s_out[1];
    scal[S_MCC_C] = s_out[2];
  }
}


}  // namespace mpba
