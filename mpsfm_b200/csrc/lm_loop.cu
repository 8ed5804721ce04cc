// lm_loop.cu -- the Levenberg-Marquardt controller ON THE DEVICE, for the small bundles of the mapper.
//
// The reference runs dozens of local bundle adjustments per registered image on 6-50 images
// (mpsfm/sfm/mapper/base.py:442-474, 567-606).  At that size one LM iteration is ~13 dependent kernels of a few
// microseconds each; with the controller on the host every iteration pays 13 launches (~4 us each when the stream
// runs dry) plus a device-to-host copy and a stream synchronisation: 167 us per iteration measured in round 1 for
// ~40 us of kernels.  Here the accept / reject / trust-region / termination logic of ba_solver.cu's host loop (Ceres
// 2.1 TrustRegionMinimizer, see oracle/lm.py) is ONE single-thread kernel that reads the iteration's scalar block,
// updates the LM state in device memory and sets the condition of a CUDA-graph WHILE node: the whole LM loop is one
// graph launch and one synchronisation.  The kernels of the iteration read {radius, 1 / radius} from the state
// (LmDiag::dev) instead of from their argument block, the current state always lives in buffer set 0 and the
// candidate in set 1 (k_lm_accept copies 1 -> 0 after an accepted step: a few hundred KB at these sizes).
//
// The arithmetic is the host loop's, operation by operation (no FMA contraction where the host has none), so the two
// controllers produce the same trajectory bit for bit -- tests/test_gpu_lm_loop.py holds them against each other.
#include <cfloat>

#include "ba_kernels.cuh"
#include "lm_control.cuh"

namespace mpba {

namespace {

__global__ void k_lm_init(double* __restrict__ ld, mpba_iteration* __restrict__ rows, double radius, double x_cost,
                          double fixed_cost, double gmax_c, double* __restrict__ scal) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  for (int k = 0; k < LD_COUNT; ++k) ld[k] = 0.0;
  ld[LD_RADIUS] = radius;
  ld[LD_INV_RADIUS] = 1.0 / radius;
  ld[LD_FIRST] = 1.0;
  ld[LD_DECREASE] = 2.0;
  ld[LD_X_COST] = x_cost;
  ld[LD_MIN_COST] = x_cost;
  ld[LD_FIXED_COST] = fixed_cost;
  ld[LD_GMAX_C] = gmax_c;
  ld[LD_NEED_GMAX] = 1.0;
  ld[LD_TERMINATION] = -1.0;
  ld[LD_NROWS] = 1.0;
  mpba_iteration it0 = {};
  it0.cost = x_cost + fixed_cost;
  it0.trust_region_radius = radius;
  rows[0] = it0;
  scal[S_BAD_POINTS] = 0.0;
}

__global__ void k_lm_control(double* ld, mpba_iteration* rows, double* scal, LmControl o,
                             cudaGraphConditionalHandle handle, int use_handle) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  const bool more = lm_control_step(ld, rows, scal, o);
  if (use_handle) cudaGraphSetConditional(handle, more ? 1u : 0u);
}

// candidate -> current after an accepted step (buffer set 1 -> set 0): poses, points, image blocks and gradients
__global__ void __launch_bounds__(256) k_lm_accept(const double* __restrict__ ld, LmAcceptCopy cp) {
  if (__ldcg(ld + LD_ACCEPT) == 0.0) return;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  const size_t t0 = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const double2* __restrict__ src = reinterpret_cast<const double2*>(cp.src[k]);
    double2* __restrict__ dst = reinterpret_cast<double2*>(cp.dst[k]);
    const size_t n2 = cp.n[k] >> 1;
    for (size_t i = t0; i < n2; i += stride) dst[i] = __ldcg(src + i);
    if (t0 == 0 && (cp.n[k] & 1)) cp.dst[k][cp.n[k] - 1] = cp.src[k][cp.n[k] - 1];
  }
}

}  // namespace

void launch_lm_init(double* ld, mpba_iteration* rows, double radius, double x_cost, double fixed_cost, double gmax_c,
                    double* scal, cudaStream_t st) {
  k_lm_init<<<1, 32, 0, st>>>(ld, rows, radius, x_cost, fixed_cost, gmax_c, scal);
}

void launch_lm_control(double* ld, mpba_iteration* rows, double* scal, const LmControl& o,
                       cudaGraphConditionalHandle handle, int use_handle, cudaStream_t st) {
  k_lm_control<<<1, 32, 0, st>>>(ld, rows, scal, o, handle, use_handle);
}

void launch_lm_accept(const double* ld, const LmAcceptCopy& cp, cudaStream_t st) {
  size_t most = 0;
  for (int k = 0; k < 4; ++k) most = most > cp.n[k] ? most : cp.n[k];
  int grid = (int)((most / 2 + 255) / 256);
  grid = grid < 1 ? 1 : (grid > 592 ? 592 : grid);
  k_lm_accept<<<grid, 256, 0, st>>>(ld, cp);
}

const char* lm_message(int code) {
  switch (code) {
    case LM_MSG_GRADIENT: return "Gradient tolerance reached.";
    case LM_MSG_PARAMETER: return "Parameter tolerance reached.";
    case LM_MSG_FUNCTION: return "Function tolerance reached.";
    case LM_MSG_INVALID: return "Number of consecutive invalid steps more than max_num_consecutive_invalid_steps.";
    case LM_MSG_MAX_ITERATIONS: return "Maximum number of iterations reached.";
    case LM_MSG_MIN_RADIUS: return "Minimum trust region radius reached.";
    default: return "";
  }
}

}  // namespace mpba
