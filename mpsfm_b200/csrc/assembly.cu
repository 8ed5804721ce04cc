// assembly.cu -- problem-assembly kernels (SURVEY.md section 8, row f "next-2"): the step right before the
// bundle-adjustment solve in the reference.
//
// mpba_sample_maps restates PriorUtils._data_at_kps
// (/root/reference/mpsfm/sfm/scene/image/mixins/priorutils.py:49-62): fp64 bilinear
// torch.nn.functional.grid_sample(mode="bilinear", padding_mode="zeros", align_corners=True) of an H x W
// map at keypoints scaled by (camera.sx, camera.sy).  The reference calls it three times per image and BA
// round from a Python loop (bundle_adjustment.py:132-135, 143, 215); here one launch samples every keypoint
// of every image of the bundle.  One thread per sample, four taps gathered from the image's map.
//
// The arithmetic follows the reference operation for operation WITHOUT fused multiply-adds
// (__dmul_rn / __dadd_rn): valid_at_kps tests the interpolated mask for "== 1", so the weights have to
// round exactly as torch's do.
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#include "../../include/mpba.h"

namespace {

__device__ __forceinline__ double tap(const double* __restrict__ map, int H, int W, double xx, double yy) {
  // zero padding: a tap outside the map contributes 0
  if (!(xx >= 0.0 && xx <= (double)(W - 1) && yy >= 0.0 && yy <= (double)(H - 1))) return 0.0;
  return __ldg(map + (size_t)(int)yy * W + (int)xx);
}

__global__ void __launch_bounds__(256) k_sample_maps(int n, const double* __restrict__ kps,
                                                     const int32_t* __restrict__ map_of,
                                                     const int64_t* __restrict__ map_offset,
                                                     const int32_t* __restrict__ map_hw,
                                                     const double* __restrict__ map_scale,
                                                     const double* __restrict__ maps, double* __restrict__ out) {
  const int k = blockIdx.x * 256 + threadIdx.x;
  if (k >= n) return;
  const int m = __ldg(map_of + k);
  const int H = __ldg(map_hw + 2 * m), W = __ldg(map_hw + 2 * m + 1);
  const double sx = __ldg(map_scale + 2 * m), sy = __ldg(map_scale + 2 * m + 1);
  const double* map = maps + __ldg(map_offset + m);
  const double2 kp = __ldg(reinterpret_cast<const double2*>(kps) + k);
  const double wx = (double)(W - 1), wy = (double)(H - 1);
  // priorutils.py:52-56: scale, normalise to [-1, 1]
  const double x = __dmul_rn(kp.x, sx), y = __dmul_rn(kp.y, sy);
  const double xn = __dadd_rn(__dmul_rn(__ddiv_rn(x, wx), 2.0), -1.0);
  const double yn = __dadd_rn(__dmul_rn(__ddiv_rn(y, wy), 2.0), -1.0);
  // grid_sample, align_corners=True: un-normalise
  const double ix = __dmul_rn(__ddiv_rn(__dadd_rn(xn, 1.0), 2.0), wx);
  const double iy = __dmul_rn(__ddiv_rn(__dadd_rn(yn, 1.0), 2.0), wy);
  const double x0 = floor(ix), y0 = floor(iy);
  const double x1 = __dadd_rn(x0, 1.0), y1 = __dadd_rn(y0, 1.0);
  const double w_nw = __dmul_rn(__dadd_rn(x1, -ix), __dadd_rn(y1, -iy));
  const double w_ne = __dmul_rn(__dadd_rn(ix, -x0), __dadd_rn(y1, -iy));
  const double w_sw = __dmul_rn(__dadd_rn(x1, -ix), __dadd_rn(iy, -y0));
  const double w_se = __dmul_rn(__dadd_rn(ix, -x0), __dadd_rn(iy, -y0));
  double acc = __dmul_rn(tap(map, H, W, x0, y0), w_nw);
  acc = __dadd_rn(acc, __dmul_rn(tap(map, H, W, x1, y0), w_ne));
  acc = __dadd_rn(acc, __dmul_rn(tap(map, H, W, x0, y1), w_sw));
  acc = __dadd_rn(acc, __dmul_rn(tap(map, H, W, x1, y1), w_se));
  out[k] = acc;
}

// image-order copies of the per-observation inputs from the track-order ones: dst[i] = src[perm[i]]
__global__ void __launch_bounds__(256) k_gather_observations(int n, const int32_t* __restrict__ perm,
                                                             const double2* __restrict__ src_xy,
                                                             const double* __restrict__ src_depth,
                                                             double2* __restrict__ dst_xy, double* __restrict__ dst_depth) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= n) return;
  const int j = __ldg(perm + i);
  dst_xy[i] = __ldg(src_xy + j);
#pragma unroll
  for (int r = 0; r < 3; ++r) dst_depth[(size_t)r * n + i] = __ldg(src_depth + (size_t)r * n + j);
}

}  // namespace

extern "C" int mpba_gather_observations(mpba_handle* h, int32_t n, const int32_t* perm, const double* src_xy,
                                        const double* src_depth, double* dst_xy, double* dst_depth, void* cuda_stream) {
  (void)h;
  if (n < 0) return MPBA_E_INVALID;
  if (n == 0) return MPBA_OK;
  if (!perm || !src_xy || !src_depth || !dst_xy || !dst_depth) return MPBA_E_INVALID;
  cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
  k_gather_observations<<<(n + 255) / 256, 256, 0, st>>>(n, perm, reinterpret_cast<const double2*>(src_xy), src_depth,
                                                         reinterpret_cast<double2*>(dst_xy), dst_depth);
  return cudaGetLastError() == cudaSuccess ? MPBA_OK : MPBA_E_CUDA;
}

extern "C" int mpba_sample_maps(mpba_handle* h, int32_t n, const double* kps, const int32_t* map_of, int32_t num_maps,
                                const int64_t* map_offset, const int32_t* map_hw, const double* map_scale,
                                const double* maps, double* out, void* cuda_stream) {
  (void)h;
  if (n < 0 || num_maps < 0) return MPBA_E_INVALID;
  if (n == 0) return MPBA_OK;
  if (!kps || !map_of || !map_offset || !map_hw || !map_scale || !maps || !out || num_maps == 0) return MPBA_E_INVALID;
  cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
  k_sample_maps<<<(n + 255) / 256, 256, 0, st>>>(n, kps, map_of, map_offset, map_hw, map_scale, maps, out);
  if (cudaGetLastError() != cudaSuccess) return MPBA_E_CUDA;
  if (cudaStreamSynchronize(st) != cudaSuccess) return MPBA_E_CUDA;
  return MPBA_OK;
}
