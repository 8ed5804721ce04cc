// ba_kernels.cu -- hand-written sm_100a kernels of the bundle-adjustment hot path.
//
// The workload is gather / segmented-reduce / small-block algebra in fp64 at 1-2 flop per byte:
// HBM- (or, for the reduced camera system, L2-) bound, so the kernels are plain SIMT with
// vectorised 16-byte loads, SoA observation streams read once per pass, register-resident block
// algebra, warp-shuffle reductions and fixed-order two-stage sums.  No tensor-core path: there is
// no GEMM here ("Why no tcgen05").
//
// Two observation orders feed the two reduction keys (north star item 2):
//   image order  -> Hcc, gc, cost      (k_img_chunks: one CTA per <=1024-observation chunk of one
//                                       image, shuffle + shared-memory reduction, no atomics)
//   track order  -> Hpp, gp, Hpp^-1, Schur complement, back-substitution (one thread per track)
// Jacobians are never stored: every pass recomputes N, h, p_rot (ba_math.cuh) from the 52-byte
// observation record and the cache-resident pose, which is cheaper than streaming 144-byte W blocks.
#include <cooperative_groups.h>
#include <stdlib.h>

#include "ba_kernels.cuh"
#include "step_common.cuh"
#include "lm_control.cuh"

namespace cg = cooperative_groups;

namespace mpba {

namespace {



struct ImgConst {
  double pose[8];
  double R[9];
  const double* prm;
  int model;
  double b, es;  // depth prior shift, exp(log-scale) (slot 7 of the pose record)
};

__device__ __forceinline__ void load_img(const ProblemView& pv, const double* __restrict__ pose, int img, ImgConst& ic) {
  const double2* pp = reinterpret_cast<const double2*>(pose + 8 * (size_t)img);
  const double2 a = __ldg(pp), b = __ldg(pp + 1), c = __ldg(pp + 2), d = __ldg(pp + 3);
  ic.pose[0] = a.x; ic.pose[1] = a.y; ic.pose[2] = b.x; ic.pose[3] = b.y;
  ic.pose[4] = c.x; ic.pose[5] = c.y; ic.pose[6] = d.x; ic.pose[7] = d.y;
  quat_to_rotmat(ic.pose, ic.R);
  const int cam = __ldg(pv.img_camera + img);
  ic.model = __ldg(pv.cam_model + cam);
  ic.prm = pv.cam_params + (size_t)MPBA_MAX_PRM * cam;
  ic.b = __ldg(pv.img_shift_scale + 2 * (size_t)img);
  ic.es = d.y;
}

__device__ __forceinline__ void load_point(const double* __restrict__ X, int p, double out[3]) {
  const double2* xp = reinterpret_cast<const double2*>(X + 4 * (size_t)p);
  const double2 a = __ldg(xp), b = __ldg(xp + 1);
  out[0] = a.x; out[1] = a.y; out[2] = b.x;
}

// ================================================================================================
// image order
// ================================================================================================
#ifndef MPBA_IMG_THREADS
#define MPBA_IMG_THREADS 128
#endif
constexpr int IMG_THREADS = MPBA_IMG_THREADS;
// one 128-byte line per thread slot `tid - slot`; slot advances by the number of lines of the range
__device__ __forceinline__ void l2_prefetch_range(const void* p, size_t bytes, int& slot, int tid) {
  if (bytes == 0) return;  // never touch a line past the end of an array
  const uintptr_t a0 = reinterpret_cast<uintptr_t>(p) & ~(uintptr_t)127, a1 = reinterpret_cast<uintptr_t>(p) + bytes;
  const int lines = (int)((a1 - a0 + 127) >> 7);
  const int i = tid - slot;
  if (i >= 0 && i < lines) asm volatile("prefetch.global.L2 [%0];" ::"l"(a0 + ((uintptr_t)i << 7)));
  slot += lines;
}
constexpr int IMG_WARPS = IMG_THREADS / 32;
#ifndef MPBA_IMG_MINB
#define MPBA_IMG_MINB 3
#endif
#ifndef MPBA_IMG_PREFETCH_DEFAULT
#define MPBA_IMG_PREFETCH_DEFAULT 740  // one wave: 148 SMs x 5 CTAs
#endif
// PLANE: the problem carries normal-prior rows (extension, ba_math.cuh: linearize_obs); the default instantiation
// is the kernel without them, instruction for instruction
// SACC (the default; MPBA_IMG_SMEM_ACC=0 selects the register variant for comparison): the 29 per-thread accumulators
// live in shared memory ([29][threads], conflict free) instead of 58 registers and the next observation's operands are
// not double-buffered in registers: 96 registers, five CTAs per SM instead of three at 168 -- the kernel is bound by
// the latency of its FP64 chains (ncu: fp64 pipe 33 % busy, 12 warps per SM), so the extra warps pay: 247 -> 202 us at
// cfg3 (profiles/r02_img_chunks_variants.md; 4 CTAs with the double buffer 214, 6 CTAs 224, 7 CTAs 270)
template <bool PLANE, bool SACC>
__global__ void __launch_bounds__(IMG_THREADS, SACC ? 5 : MPBA_IMG_MINB) k_img_chunks(ProblemView pv, const double* __restrict__ pose,
                                                    const double* __restrict__ X, double* __restrict__ chunk_part,
                                                    int prefetch_ahead) {
  const int c = blockIdx.x;
  const int img = pv.chunk_img[c];
  const int beg = pv.chunk_begin[c], end = pv.chunk_end[c];
  // the image's constants (pose, rotation matrix, prior shift / scale) are CTA-uniform: they live in shared
  // memory (broadcast reads) instead of ~40 registers of every thread
  __shared__ ImgConst s_ic;
  if (threadIdx.x == 0) load_img(pv, pose, img, s_ic);
  __syncthreads();
  const ImgConst& ic = s_ic;
  const bool img_const = pv.img_free_mask[img] == 0;
  const double2* xy = reinterpret_cast<const double2*>(pv.cm_xy);
  const double* dd = pv.cm_depth;
  const size_t N = (size_t)pv.N;

  double acc[29];
  __shared__ double s_acc[SACC ? 29 : 1][SACC ? IMG_THREADS : 1];
#pragma unroll
  for (int j = 0; j < 29; ++j) {
    acc[j] = 0.0;
    if (SACC) s_acc[SACC ? j : 0][SACC ? threadIdx.x : 0] = 0.0;
  }

  // Software pipeline: the operands of the thread's NEXT observation (index -> point gather, xy, depth row) are
  // in flight while the current one is linearised -- with 12 warps per SM (168 registers, three 128-thread CTAs)
  // nothing else hides the DRAM + L2 round trip of the dependent gather.  Slot 3 of a point record carries its constant flag.
  struct Operands {
    double2 xa, xb, o2;
    double d, w, a;
    double pl[PLANE ? 4 : 1];
  };
  const double2* pln = reinterpret_cast<const double2*>(pv.cm_plane);
  auto fetch = [&](int k, Operands& op) {
    const int pt = __ldg(pv.cm_pt + k);
    const double2* xp = reinterpret_cast<const double2*>(X + 4 * (size_t)pt);
    op.xa = __ldg(xp);
    op.xb = __ldg(xp + 1);
    op.o2 = __ldg(xy + k);
    op.d = __ldg(dd + k);
    op.w = __ldg(dd + N + k);
    op.a = __ldg(dd + 2 * N + k);
    if (PLANE) {
      const double2 p0 = __ldg(pln + 2 * (size_t)k), p1 = __ldg(pln + 2 * (size_t)k + 1);
      op.pl[0] = p0.x; op.pl[1] = p0.y; op.pl[PLANE ? 2 : 0] = p1.x; op.pl[PLANE ? 3 : 0] = p1.y;
    }
  };
  constexpr bool PIPE = !SACC;  // SACC: no register double buffer, five CTAs per SM hide the gathers
  int k = beg + (int)threadIdx.x;
  Operands cur, nxt;
  if (PIPE && k < end) fetch(k, cur);
  for (; k < end; k += IMG_THREADS) {
    const bool more = PIPE && k + IMG_THREADS < end;
    if (!PIPE) fetch(k, cur);
    if (more) fetch(k + IMG_THREADS, nxt);
    const double Xp[3] = {cur.xa.x, cur.xa.y, cur.xb.x};
    ObsLin o;
    linearize_obs<true, true>(ic.pose, ic.R, Xp, ic.model, ic.prm, cur.o2.x, cur.o2.y, cur.d, cur.w, cur.a, ic.b, ic.es,
                              pv.loss, o, PLANE ? cur.pl : nullptr);
    double H[21], g[6];
    gt_n_g_upper(o.prot, o.N, H);
    gt_vec(o.prot, o.h, g);
    const bool fixed = img_const && (cur.xb.y != 0.0);
    if (SACC) {
      const int t = SACC ? threadIdx.x : 0;
#pragma unroll
      for (int j = 0; j < 21; ++j) s_acc[SACC ? j : 0][t] += H[j];
#pragma unroll
      for (int j = 0; j < 6; ++j) s_acc[SACC ? 21 + j : 0][t] += g[j];
      s_acc[SACC ? (fixed ? 28 : 27) : 0][t] += o.rho;
    } else {
#pragma unroll
      for (int j = 0; j < 21; ++j) acc[j] += H[j];
#pragma unroll
      for (int j = 0; j < 6; ++j) acc[21 + j] += g[j];
      if (fixed) acc[28] += o.rho; else acc[27] += o.rho;
    }
    if (more) cur = nxt;
  }
  if (SACC) {
#pragma unroll
    for (int j = 0; j < 29; ++j) acc[j] = s_acc[SACC ? j : 0][SACC ? threadIdx.x : 0];
  }

  __shared__ double sm[IMG_WARPS][IMG_PART];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int j = 0; j < 29; ++j) {
    double v = acc[j];
    for (int off = 16; off > 0; off >>= 1) v += __shfl_down_sync(FULL, v, off);
    if (lane == 0) sm[warp][j] = v;
  }
  __syncthreads();
  if (threadIdx.x < 29) {
    double s = 0.0;
#pragma unroll
    for (int w = 0; w < IMG_WARPS; ++w) s += sm[w][threadIdx.x];
    chunk_part[(size_t)c * IMG_PART + threadIdx.x] = s;
  }
  // L2 prefetch of the sequential streams of the chunk prefetch_ahead CTAs later (see prefetch_track_cta)
  if (prefetch_ahead > 0 && c + prefetch_ahead < (int)gridDim.x) {
    const size_t b2 = (size_t)__ldg(pv.chunk_begin + c + prefetch_ahead);
    const size_t n2 = (size_t)__ldg(pv.chunk_end + c + prefetch_ahead) - b2;
    for (int t = threadIdx.x; t < 384; t += IMG_THREADS) {
      int slot = 0;
      l2_prefetch_range(pv.cm_pt + b2, n2 * 4, slot, t);
      l2_prefetch_range(pv.cm_xy + 2 * b2, n2 * 16, slot, t);
      l2_prefetch_range(dd + b2, n2 * 8, slot, t);
      l2_prefetch_range(dd + N + b2, n2 * 8, slot, t);
      l2_prefetch_range(dd + 2 * N + b2, n2 * 8, slot, t);
    }
  }
}

// one thread per (image, value): sums the image's chunk partials in chunk order
// value v (0..20 Hcc upper, 21..26 gc, 27..28 costs) of image i: chunk partials summed in chunk order
__device__ __forceinline__ void img_finalize_one(const ProblemView& pv, const double* __restrict__ chunk_part, double* Hc,
                                                 double* gc, double* img_part, int i, int v) {
  double s = 0.0;
  for (int ch = pv.img_chunk_ptr[i]; ch < pv.img_chunk_ptr[i + 1]; ++ch) s += chunk_part[(size_t)ch * IMG_PART + v];
  const unsigned mask = pv.img_free_mask[i];
  if (v < 21) {
    // (r, c) of upper index v
    int r = 0, base = 0;
    while (v >= base + (6 - r)) { base += 6 - r; ++r; }
    const int c = r + (v - base);
    const bool fr = (mask >> r) & 1u, fc = (mask >> c) & 1u;
    const double val = (fr && fc) ? s : 0.0;
    Hc[(size_t)i * 21 + v] = val;
  } else if (v < 27) {
    const int r = v - 21;
    gc[(size_t)i * 6 + r] = ((mask >> r) & 1u) ? s : 0.0;
  } else {
    img_part[(size_t)i * 2 + (v - 27)] = s;
  }
}

__global__ void __launch_bounds__(256) k_img_finalize(ProblemView pv, const double* __restrict__ chunk_part,
                                                      double* __restrict__ Hc, double* __restrict__ gc,
                                                      double* __restrict__ img_part) {
  const int gid = blockIdx.x * 256 + threadIdx.x;
  const int i = gid >> 5, v = gid & 31;
  if (i >= pv.C || v >= 29) return;
  img_finalize_one(pv, chunk_part, Hc, gc, img_part, i, v);
}

__device__ __forceinline__ void img_scalars_body(int C, const uint8_t* __restrict__ free_mask, const double* img_part,
                                                 const double* Hc, const double* gc, double* c_img,
                                                 int compute_scaling, int jacobi_scaling, double* scal, double* sm) {
  double cf = 0.0, cx = 0.0, gm = 0.0;
  for (int i = threadIdx.x; i < C; i += 1024) {
    cf += img_part[2 * (size_t)i];
    cx += img_part[2 * (size_t)i + 1];
#pragma unroll
    for (int j = 0; j < 6; ++j) gm = fmax(gm, fabs(gc[(size_t)i * 6 + j]));
    if (compute_scaling) {
      // Jacobi column scales c_j = 1 / (1 + |J_j|) of the image blocks (Ceres: fixed at iteration 0)
      const unsigned mask = free_mask[i];
#pragma unroll
      for (int r = 0; r < 6; ++r) {
        const double val = Hc[(size_t)i * 21 + upper_idx(r, r)];
        c_img[(size_t)i * 6 + r] = ((mask >> r) & 1u) ? (jacobi_scaling ? 1.0 / (1.0 + sqrt(val)) : 1.0) : 0.0;
      }
    }
  }
  cf = block_sum<32>(cf, sm);
  cx = block_sum<32>(cx, sm);
  gm = block_max<32>(gm, sm);
  if (threadIdx.x == 0) {
    scal[S_COST_FREE] = cf;
    scal[S_COST_FIXED] = cx;
    scal[S_GMAX_C] = gm;
  }
}

__global__ void __launch_bounds__(1024) k_img_scalars(int C, const uint8_t* __restrict__ free_mask,
                                                      const double* __restrict__ img_part,
                                                      const double* __restrict__ Hc, const double* __restrict__ gc,
                                                      double* __restrict__ c_img, int compute_scaling,
                                                      int jacobi_scaling, double* __restrict__ scal) {
  __shared__ double sm[33];
  img_scalars_body(C, free_mask, img_part, Hc, gc, c_img, compute_scaling, jacobi_scaling, scal, sm);
}

// Small bundles: the two kernels above as one CTA (same sums, same order), then -- device-controlled LM loop -- the
// LM decision by thread 0 (lm_control.cuh) and the condition of the graph's WHILE node.
__global__ void __launch_bounds__(1024) k_img_tail(ProblemView pv, const double* __restrict__ chunk_part, double* Hc,
                                                   double* gc, double* img_part, double* c_img, int compute_scaling,
                                                   int jacobi_scaling, double* scal, LmTail tail, int with_tail,
                                                   PtScalarsArgs ps) {
  __shared__ double sm[33];
  for (int gid = threadIdx.x; gid < pv.C * 32; gid += 1024) {
    const int i = gid >> 5, v = gid & 31;
    if (v < 29) img_finalize_one(pv, chunk_part, Hc, gc, img_part, i, v);
  }
  __syncthreads();  // Hc / gc / img_part of this CTA are complete and visible to it
  img_scalars_body(pv.C, pv.img_free_mask, img_part, Hc, gc, c_img, compute_scaling, jacobi_scaling, scal, sm);
  if (ps.nblk >= 0) {
    // the point scalars of the step (k_pt_scalars_step's work: a few hundred rows at these sizes) and the slice sums of
    // the camera step ride along: one launch less per LM iteration
    double a = 0.0, b = 0.0, c = 0.0, gm = 0.0;
    for (int i = threadIdx.x; i < ps.nblk; i += 1024) {
      const double4 v = *reinterpret_cast<const double4*>(ps.pt_part + (size_t)i * PT_PART);
      gm = fmax(gm, v.x);
      a += v.y; b += v.z; c += v.w;
    }
    a = block_sum<32>(a, sm);
    b = block_sum<32>(b, sm);
    c = block_sum<32>(c, sm);
    gm = block_max<32>(gm, sm);
    if (threadIdx.x == 0) {
      scal[S_MCC] = a;
      scal[S_STEP2_P] = b;
      scal[S_XNORM2_P] = c;
      if (ps.with_gmax) scal[S_GMAX_P] = gm;
    }
    if (threadIdx.x < 3 && ps.cam_rows > 0) {
      double acc = 0.0;
      for (int g = 0; g < ps.cam_rows; ++g) acc += ps.cam_stage[3 * g + threadIdx.x];
      scal[threadIdx.x == 0 ? S_STEP2_C : (threadIdx.x == 1 ? S_XNORM2_C : S_MCC_C)] = acc;
    }
    __syncthreads();  // the controller below reads what threads 1 and 2 wrote
  }
  if (with_tail && threadIdx.x == 0) {
    // scal[S_COST_FREE..S_GMAX_C] were written by this very thread; the rest by earlier kernels
    const bool more = lm_control_step(tail.ld, tail.rows, scal, tail.ctl);
    cudaGraphSetConditional(tail.handle, more ? 1u : 0u);
  }
}

// ================================================================================================
// track order
// ================================================================================================
// sum of v[j] over the 32 lanes for j = 0..31, delivered to lane j, in 31 exchanges
__device__ __forceinline__ double warp_transpose_reduce32(double (&v)[36], int lane) {
#pragma unroll
  for (int w = 16; w >= 1; w >>= 1) {
    const bool up = (lane & w) != 0;
#pragma unroll
    for (int i = 0; i < w; ++i) {
      const double send = up ? v[i] : v[i + w];
      const double keep = up ? v[i + w] : v[i];
      v[i] = keep + __shfl_xor_sync(FULL, send, w);
    }
  }
  return v[0];
}

template <bool PLANE>
__device__ __forceinline__ void track_obs(const ProblemView& pv, const double* __restrict__ pose, const double Xp[3],
                                          int k, ObsLin& o, ImgConst& ic) {
  const int img = __ldg(pv.pm_img + k);
  load_img(pv, pose, img, ic);
  const double2 o2 = __ldg(reinterpret_cast<const double2*>(pv.pm_xy) + k);
  const size_t N = (size_t)pv.N;
  const double d = __ldg(pv.pm_depth + k), w = __ldg(pv.pm_depth + N + k), a = __ldg(pv.pm_depth + 2 * N + k);
  if (PLANE) {
    const double2* pp = reinterpret_cast<const double2*>(pv.pm_plane) + 2 * (size_t)k;
    const double2 p0 = __ldg(pp), p1 = __ldg(pp + 1);
    const double pl[4] = {p0.x, p0.y, p1.x, p1.y};
    linearize_obs<true, false>(ic.pose, ic.R, Xp, ic.model, ic.prm, o2.x, o2.y, d, w, a, ic.b, ic.es, pv.loss, o, pl);
  } else {
    linearize_obs<true, false>(ic.pose, ic.R, Xp, ic.model, ic.prm, o2.x, o2.y, d, w, a, ic.b, ic.es, pv.loss, o);
  }
}

constexpr int PT_THREADS = 128;  // k_evaluate: one thread per track
#ifndef MPBA_PW_THREADS
#define MPBA_PW_THREADS 128
#endif
constexpr int PW_THREADS = MPBA_PW_THREADS;  // track-order passes: one thread per OBSERVATION, one warp per packed group of tracks
constexpr int PW_WARPS = PW_THREADS / 32;
#ifndef MPBA_POINTS_MINB
#define MPBA_POINTS_MINB 4
#endif
#ifndef MPBA_POINTS_BULK_DEFAULT
#define MPBA_POINTS_BULK_DEFAULT 0
#endif
#ifndef MPBA_POINTS_PREFETCH_DEFAULT
#define MPBA_POINTS_PREFETCH_DEFAULT 592  // one wave: 148 SMs x 4 CTAs (sweep: profiles/r01_prefetch_sweep_cfg3.json)
#endif
#ifndef MPBA_BACKSUB_PREFETCH_DEFAULT
#define MPBA_BACKSUB_PREFETCH_DEFAULT 333
#endif
constexpr int LINV_STRIDE = 9;  // per point: L^-1 (6, lower) and e = L^-1 gp (3)

// ---- one warp = a packed group of whole tracks (mpba_pack_tracks: <= 32 observations and <= 32 tracks, or one
// single longer track).  Lane l is observation slot base + l AND (for l < number of tracks) track t0 + l.
struct WarpTracks {
  int t0, nt, base, n;  // first track, tracks, first slot, observations
  // per track lane
  int start, len;       // first observation lane / observations of track t0 + lane
  bool cst;             // constant point
  // per observation lane
  bool has;             // lane < n
  int tl;               // lane holding this observation's track
  int seg_end;          // one past the last observation lane of the track
  bool tcst;            // the track is a constant point
};

__device__ __forceinline__ void warp_tracks_setup(const ProblemView& pv, int w, int lane, WarpTracks& wt,
                                                  int* scratch /* 32 ints of shared memory of this warp */) {
  wt.t0 = __ldg(pv.tw_first + w);
  wt.nt = __ldg(pv.tw_first + w + 1) - wt.t0;
  wt.base = __ldg(pv.pt_ptr + wt.t0);
  wt.n = __ldg(pv.pt_ptr + wt.t0 + wt.nt) - wt.base;
  wt.start = 0; wt.len = 0; wt.cst = true;
  if (lane < wt.nt) {
    const int a = __ldg(pv.pt_ptr + wt.t0 + lane), b = __ldg(pv.pt_ptr + wt.t0 + lane + 1);
    wt.start = a - wt.base;
    wt.len = b - a;
    wt.cst = pv.pt_const[wt.t0 + lane] != 0;
  }
  wt.has = lane < wt.n;
  wt.tl = 0; wt.seg_end = wt.n; wt.tcst = true;
  if (wt.n <= 32) {
    const bool nonempty = wt.len > 0;
    const unsigned heads = __reduce_or_sync(FULL, nonempty ? (1u << wt.start) : 0u);  // first observation lanes
    if (nonempty) scratch[wt.start] = lane;  // a track lane tells its first observation lane who it is
    __syncwarp();
    const unsigned le = heads & (0xffffffffu >> (31 - lane));
    wt.tl = (wt.has && le) ? scratch[31 - __clz(le)] : 0;
    __syncwarp();
    const unsigned above = (lane == 31) ? 0u : (heads & (0xffffffffu << (lane + 1)));
    wt.seg_end = above ? (__ffs(above) - 1) : wt.n;
    wt.tcst = __shfl_sync(FULL, wt.cst ? 1 : 0, wt.tl) != 0;
  }
}

// sum over the observation lanes of every track, delivered to the track's first observation lane (fixed tree)
template <int NV>
__device__ __forceinline__ void segmented_reduce(double (&v)[NV], int lane, int seg_end) {
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    const bool take = lane + off < seg_end;
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const double t = __shfl_down_sync(FULL, v[j], off);
      if (take) v[j] += t;
    }
  }
}

template <int NV>
__device__ __forceinline__ void warp_allreduce(double (&v)[NV]) {
#pragma unroll
  for (int off = 16; off > 0; off >>= 1)
#pragma unroll
    for (int j = 0; j < NV; ++j) v[j] += __shfl_xor_sync(FULL, v[j], off);
}

// Hpp / gp contributions of one linearised observation: v = [R^T N R (6 upper), R^T h (3)], Q = N R
__device__ __forceinline__ void point_terms(const ObsLin& o, const double* R, double Q[9], double v[9]) {
  sym3_mul_mat(o.N, R, Q);
  v[0] = R[0] * Q[0] + R[3] * Q[3] + R[6] * Q[6];
  v[1] = R[0] * Q[1] + R[3] * Q[4] + R[6] * Q[7];
  v[2] = R[0] * Q[2] + R[3] * Q[5] + R[6] * Q[8];
  v[3] = R[1] * Q[1] + R[4] * Q[4] + R[7] * Q[7];
  v[4] = R[1] * Q[2] + R[4] * Q[5] + R[7] * Q[8];
  v[5] = R[2] * Q[2] + R[5] * Q[5] + R[8] * Q[8];
  v[6] = R[0] * o.h[0] + R[3] * o.h[1] + R[6] * o.h[2];
  v[7] = R[1] * o.h[0] + R[4] * o.h[1] + R[7] * o.h[2];
  v[8] = R[2] * o.h[0] + R[5] * o.h[1] + R[8] * o.h[2];
}

// Per point, from the complete sums: Jacobi scale, LM damping, Cholesky of Hpp + D, Hpp^-1; writes the point's
// rows of c_pt / Hpp / gp / Hinv / Linv and returns L^-1 (Li), e = L^-1 gp.
__device__ __forceinline__ void finalize_point(int p, bool active, const double Hp[6], const double g[3], LmDiag lm,
                                               double* __restrict__ c_pt, int compute_scaling, int jacobi_scaling,
                                               double* __restrict__ Hpp, double* __restrict__ gp,
                                               double* __restrict__ Hinv, double* __restrict__ Linv,
                                               double* __restrict__ scal, double Li[6], double e[3],
                                               const double* cpre = nullptr /* c_pt row loaded ahead by the caller */) {
  double Hi[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
#pragma unroll
  for (int j = 0; j < 6; ++j) Li[j] = 0.0;
  double c[3] = {0.0, 0.0, 0.0};
  if (compute_scaling) {
    if (active) {
      c[0] = jacobi_scaling ? 1.0 / (1.0 + sqrt(Hp[0])) : 1.0;
      c[1] = jacobi_scaling ? 1.0 / (1.0 + sqrt(Hp[3])) : 1.0;
      c[2] = jacobi_scaling ? 1.0 / (1.0 + sqrt(Hp[5])) : 1.0;
    }
    c_pt[(size_t)p * 3] = c[0]; c_pt[(size_t)p * 3 + 1] = c[1]; c_pt[(size_t)p * 3 + 2] = c[2];
  } else if (cpre != nullptr) {
    c[0] = cpre[0]; c[1] = cpre[1]; c[2] = cpre[2];
  } else {
    c[0] = c_pt[(size_t)p * 3]; c[1] = c_pt[(size_t)p * 3 + 1]; c[2] = c_pt[(size_t)p * 3 + 2];
  }
  if (active) {
    double Hd[6] = {Hp[0], Hp[1], Hp[2], Hp[3], Hp[4], Hp[5]};
    Hd[0] += lm_damping_fast(Hp[0], c[0], lm.radius, lm.inv_radius, lm.lo, lm.hi);
    Hd[3] += lm_damping_fast(Hp[3], c[1], lm.radius, lm.inv_radius, lm.lo, lm.hi);
    Hd[5] += lm_damping_fast(Hp[5], c[2], lm.radius, lm.inv_radius, lm.lo, lm.hi);
    if (sym3_chol_inverse(Hd, Li)) {
      if (Hinv) sym3_from_chol_inverse(Li, Hi);
    } else {
#pragma unroll
      for (int j = 0; j < 6; ++j) Li[j] = 0.0;
      atomicAdd(scal + S_BAD_POINTS, 1.0);  // failure counter, not arithmetic
    }
  }
#pragma unroll
  for (int j = 0; j < 6; ++j) {
    Hpp[(size_t)p * 6 + j] = Hp[j];
    if (Hinv) Hinv[(size_t)p * 6 + j] = Hi[j];  // only the covariance / linearize entry points ask for it
    Linv[(size_t)p * LINV_STRIDE + j] = Li[j];
  }
#pragma unroll
  for (int j = 0; j < 3; ++j) gp[(size_t)p * 3 + j] = g[j];
  e[0] = Li[0] * g[0];
  e[1] = Li[1] * g[0] + Li[2] * g[1];
  e[2] = Li[3] * g[0] + Li[4] * g[1] + Li[5] * g[2];
#pragma unroll
  for (int j = 0; j < 3; ++j) Linv[(size_t)p * LINV_STRIDE + 6 + j] = e[j];
}

__device__ __forceinline__ void pack_factor(const double Q[9], const double Li[6], const double prot[3],
                                            const double e[3], double2 rec[8]) {
  double Z[9];
  mul_chol_inverse_t(Q, Li, Z);
  rec[0] = make_double2(Z[0], Z[1]);
  rec[1] = make_double2(Z[2], Z[3]);
  rec[2] = make_double2(Z[4], Z[5]);
  rec[3] = make_double2(Z[6], Z[7]);
  rec[4] = make_double2(Z[8], prot[0]);
  rec[5] = make_double2(prot[1], prot[2]);
  rec[6] = make_double2(e[0], e[1]);
  rec[7] = make_double2(e[2], 0.0);
}

// ---- warp path of the track-order pass: the tracks of ONE warp group (used when a CTA's groups hold a track
// longer than 32 observations).  Sums by segmented shuffle tree (<= 32 observations) or whole-warp rounds.
template <bool PLANE>
__device__ __noinline__ double points_warp_path(const ProblemView pv, const double* __restrict__ pose,
                                                const double* __restrict__ X, LmDiag lm, double* __restrict__ c_pt,
                                                int compute_scaling, int jacobi_scaling, double* __restrict__ Hpp,
                                                double* __restrict__ gp, double* __restrict__ Hinv,
                                                double* __restrict__ Linv, double* __restrict__ fac,
                                                double* __restrict__ scal, int write_factors, double2* rows_w, int w,
                                                int lane) {
  double gmax = 0.0;
  if (w < pv.num_track_warps) {
    WarpTracks wt;
    warp_tracks_setup(pv, w, lane, wt, reinterpret_cast<int*>(rows_w));
    double2* fac2 = reinterpret_cast<double2*>(fac);
    if (wt.n <= 32) {
      // ---- observation lanes: linearise, reduce over the track
      const bool act = wt.has && !wt.tcst;
      double v[9] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0}, Q[9], prot[3] = {0.0, 0.0, 0.0};
      if (act) {
        double Xp[3];
        load_point(X, wt.t0 + wt.tl, Xp);
        ObsLin o;
        ImgConst ic;
        track_obs<PLANE>(pv, pose, Xp, wt.base + lane, o, ic);
        point_terms(o, ic.R, Q, v);
        prot[0] = o.prot[0]; prot[1] = o.prot[1]; prot[2] = o.prot[2];
      }
      segmented_reduce<9>(v, lane, wt.seg_end);
      // ---- track lanes: fetch the sums from the first observation lane, factor, write the point's rows
      double Hp[6], g[3];
#pragma unroll
      for (int j = 0; j < 6; ++j) Hp[j] = __shfl_sync(FULL, v[j], wt.start);
#pragma unroll
      for (int j = 0; j < 3; ++j) g[j] = __shfl_sync(FULL, v[6 + j], wt.start);
      double Li[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0}, e[3] = {0.0, 0.0, 0.0};
      if (lane < wt.nt) {
        const bool active = !wt.cst && wt.len > 0;
        if (!active) {
#pragma unroll
          for (int j = 0; j < 6; ++j) Hp[j] = 0.0;
          g[0] = g[1] = g[2] = 0.0;
        }
        finalize_point(wt.t0 + lane, active, Hp, g, lm, c_pt, compute_scaling, jacobi_scaling, Hpp, gp, Hinv, Linv,
                       scal, Li, e);
        gmax = fmax(fabs(g[0]), fmax(fabs(g[1]), fabs(g[2])));
      }
      if (write_factors) {
        // ---- back to the observation lanes: L^-1 and e of their track, one factor record each
        double Lo[6], eo[3];
#pragma unroll
        for (int j = 0; j < 6; ++j) Lo[j] = __shfl_sync(FULL, Li[j], wt.tl);
#pragma unroll
        for (int j = 0; j < 3; ++j) eo[j] = __shfl_sync(FULL, e[j], wt.tl);
        double2* rows = rows_w;
        if (act) {
          double2 rec[8];
          pack_factor(Q, Lo, prot, eo, rec);
          const int sw = lane & 7;
#pragma unroll
          for (int c = 0; c < 8; ++c) rows[lane * 8 + (c ^ sw)] = rec[c];
        }
        const unsigned actmask = __ballot_sync(FULL, act);  // (also orders the shared-memory writes in the warp)
        const int g4 = lane >> 3, c = lane & 7;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int item = 4 * j + g4;
          if ((actmask >> item) & 1u) fac2[(size_t)(wt.base + item) * 8 + c] = rows[item * 8 + (c ^ (item & 7))];
        }
      }
    } else {
      // ---- one long track (> 32 observations): rounds of 32, whole-warp sums, second sweep for the factors
      const int p = wt.t0;
      const bool active = pv.pt_const[p] == 0;
      double Xp[3] = {0.0, 0.0, 0.0};
      if (active) load_point(X, p, Xp);
      double v[9] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
      if (active) {
        for (int k = lane; k < wt.n; k += 32) {
          ObsLin o;
          ImgConst ic;
          track_obs<PLANE>(pv, pose, Xp, wt.base + k, o, ic);
          double Q[9], t[9];
          point_terms(o, ic.R, Q, t);
#pragma unroll
          for (int j = 0; j < 9; ++j) v[j] += t[j];
        }
      }
      warp_allreduce<9>(v);
      double Li[6], e[3];
      if (lane == 0) {
        finalize_point(p, active, v, v + 6, lm, c_pt, compute_scaling, jacobi_scaling, Hpp, gp, Hinv, Linv, scal, Li, e);
        gmax = fmax(fabs(v[6]), fmax(fabs(v[7]), fabs(v[8])));
      }
#pragma unroll
      for (int j = 0; j < 6; ++j) Li[j] = __shfl_sync(FULL, Li[j], 0);
#pragma unroll
      for (int j = 0; j < 3; ++j) e[j] = __shfl_sync(FULL, e[j], 0);
      if (write_factors && active) {
        for (int k = lane; k < wt.n; k += 32) {
          ObsLin o;
          ImgConst ic;
          track_obs<PLANE>(pv, pose, Xp, wt.base + k, o, ic);
          double Q[9], t[9];
          point_terms(o, ic.R, Q, t);
          double2 rec[8];
          pack_factor(Q, Li, o.prot, e, rec);
          double2* dst = fac2 + (size_t)(wt.base + k) * 8;
#pragma unroll
          for (int c = 0; c < 8; ++c) dst[c] = rec[c];
        }
      }
    }
  }
  return gmax;
}

// ---- L2 prefetch of a later CTA's operands.  The track-order streams are sequential, but every CTA starts with a
// chain of dependent loads (descriptor -> slot indices -> point / pose records) that each miss to DRAM; a CTA that
// runs one wave earlier asks L2 for those lines, so the chain of the later CTA costs L2 hits instead.
__device__ __forceinline__ void prefetch_track_cta(const ProblemView& pv, const double* X, int4 d, int tid) {
  const size_t T0 = (size_t)d.x, nt = (size_t)d.y, base = (size_t)d.z, n = (size_t)d.w, N = (size_t)pv.N;
  int slot = 0;
  l2_prefetch_range(pv.pm_xy + 2 * base, n * 16, slot, tid);
  l2_prefetch_range(pv.pm_depth + base, n * 8, slot, tid);
  l2_prefetch_range(pv.pm_depth + N + base, n * 8, slot, tid);
  l2_prefetch_range(pv.pm_depth + 2 * N + base, n * 8, slot, tid);
  l2_prefetch_range(X + 4 * T0, nt * 32, slot, tid);
  l2_prefetch_range(pv.pm_pt + base, n * 4, slot, tid);
  l2_prefetch_range(pv.pm_img + base, n * 4, slot, tid);
  l2_prefetch_range(pv.pt_ptr + T0, (nt + 1) * 4, slot, tid);
  l2_prefetch_range(pv.pt_const + T0, nt, slot, tid);
}

// Track-order pass, one thread per OBSERVATION, one CTA per four consecutive warp groups of whole tracks
// (mpba_pack_tracks: <= 128 observations, <= 128 tracks, contiguous slots).  Every observation is linearised ONCE:
//   1. track threads publish the track of every observation slot of the CTA (shared memory),
//   2. observation threads linearise and leave R^T N R (6) and R^T h (3) in shared memory, keeping Q = N R, p_rot,
//   3. track threads add their observations in slot order (deterministic), damp, factor Hpp + D = L L^T, write the
//      point's rows (c_pt, Hpp, gp, Hpp^-1, L^-1, e = L^-1 gp) and publish L^-1, e,
//   4. observation threads write one 128-byte Schur factor each,
//        fac[k] = { Z = N R L^-T (9), p_rot (3), e (3), 0 },
//      from which k_schur_blocks forms  W_a Hpp^-1 W_b^T = G_a^T (Z_a Z_b^T) G_b  and  W_a Hpp^-1 gp = G_a^T Z_a e
//      without ever storing the 6x3 W blocks, and k_back_substitute the point steps.  The records of a warp go
//      through swizzled shared-memory rows so that the stores leave as whole 128-byte lines.
// The reductions are keyed by point without atomics (north star item 2); the per-point algebra runs on densely
// packed track threads instead of a few lanes of every warp.
template <bool PLANE>
__global__ void __launch_bounds__(PW_THREADS, MPBA_POINTS_MINB) k_points(ProblemView pv, const double* __restrict__ pose,
                                                       const double* __restrict__ X, LmDiag lm,
                                                       double* __restrict__ c_pt, int compute_scaling,
                                                       int jacobi_scaling, double* __restrict__ Hpp,
                                                       double* __restrict__ gp, double* __restrict__ Hinv,
                                                       double* __restrict__ Linv, double* __restrict__ fac,
                                                       double* __restrict__ pt_part, double* __restrict__ scal,
                                                       int write_factors, int force_warp_path, int prefetch_ahead) {
  __shared__ __align__(16) double2 rows_all[PW_WARPS][256];
  __shared__ double s_v[PW_THREADS][9];
  __shared__ double s_le[PW_THREADS][9];
  __shared__ __align__(16) int4 s_next;  // descriptor of the CTA prefetch_ahead blocks later
  __shared__ double s_gw[PW_WARPS];
  bool gmax_done = false;
  if (compute_scaling < 0) compute_scaling = __ldcg(lm.dev + LD_FIRST) != 0.0 ? 1 : 0;
  lm_resolve(lm);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int g0 = blockIdx.x * PW_WARPS, g1 = min(g0 + PW_WARPS, pv.num_track_warps);
  const int4 desc = __ldg(pv.cta_desc + blockIdx.x);  // one load instead of the tw_first -> pt_ptr chain
  const int T0 = desc.x, nt = desc.y, base = desc.z, n = desc.w;
  double gmax = 0.0;
  if (n <= PW_THREADS && !force_warp_path) {
    const bool pf = prefetch_ahead > 0 && blockIdx.x + prefetch_ahead < gridDim.x;
    if (pf && tid == PW_THREADS - 1) {  // asynchronous: lands in shared memory while the CTA linearises
      const unsigned dst = (unsigned)__cvta_generic_to_shared(&s_next);
      asm volatile("cp.async.ca.shared.global [%0], [%1], 16;\n" ::"r"(dst), "l"(pv.cta_desc + blockIdx.x + prefetch_ahead)
                   : "memory");
      asm volatile("cp.async.commit_group;\n" ::: "memory");
    }
    // track threads: their slot range and constant flag (independent loads, needed after the first barrier)
    int start = 0, len = 0;
    bool cst = true;
    double cpre[3] = {0.0, 0.0, 0.0};  // Jacobi scales of the track's point: in flight while the observations linearise
    if (tid < nt) {
      const int a = __ldg(pv.pt_ptr + T0 + tid), b = __ldg(pv.pt_ptr + T0 + tid + 1);
      start = a - base;
      len = b - a;
      cst = pv.pt_const[T0 + tid] != 0;
      if (!compute_scaling) {
        const double* cp = c_pt + 3 * (size_t)(T0 + tid);
        cpre[0] = cp[0]; cpre[1] = cp[1]; cpre[2] = cp[2];
      }
    }
    // ---- 1. observations: point of the slot (pm_pt), gather, linearise, leave R^T N R / R^T h in shared memory
    int tr = -1;
    double Q[9], prot[3] = {0.0, 0.0, 0.0};
    if (tid < n) {
      const int pt = __ldg(pv.pm_pt + base + tid);
      const double2* xp = reinterpret_cast<const double2*>(X + 4 * (size_t)pt);
      const double2 xa = __ldg(xp), xb = __ldg(xp + 1);
      if (xb.y == 0.0) {  // slot 3: constant flag
        tr = pt - T0;
        const double Xp[3] = {xa.x, xa.y, xb.x};
        ObsLin o;
        ImgConst ic;
        track_obs<PLANE>(pv, pose, Xp, base + tid, o, ic);
        double v[9];
        point_terms(o, ic.R, Q, v);
#pragma unroll
        for (int j = 0; j < 9; ++j) s_v[tid][j] = v[j];
        prot[0] = o.prot[0]; prot[1] = o.prot[1]; prot[2] = o.prot[2];
      }
    }
    const bool act = tr >= 0;
    if (pf && tid == PW_THREADS - 1) asm volatile("cp.async.wait_group 0;\n" ::: "memory");
    __syncthreads();
    if (pf && tid < 64) prefetch_track_cta(pv, X, s_next, tid);  // ~60 lines per CTA: two warps cover them
    // ---- 2. points
    if (tid < nt) {
      const bool active = !cst && len > 0;
      double v[9] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
      if (active) {
        for (int k = 0; k < len; ++k) {
#pragma unroll
          for (int j = 0; j < 9; ++j) v[j] += s_v[start + k][j];
        }
      }
      double Li[6], e[3];
      finalize_point(T0 + tid, active, v, v + 6, lm, c_pt, compute_scaling, jacobi_scaling, Hpp, gp, Hinv, Linv, scal,
                     Li, e, cpre);
#pragma unroll
      for (int j = 0; j < 6; ++j) s_le[tid][j] = Li[j];
#pragma unroll
      for (int j = 0; j < 3; ++j) s_le[tid][6 + j] = e[j];
      gmax = fmax(fabs(v[6]), fmax(fabs(v[7]), fabs(v[8])));
    }
    // max |gp| of the CTA rides on the barrier that follows (no block reduction of its own: three barriers less)
    for (int off = 16; off > 0; off >>= 1) gmax = fmax(gmax, __shfl_xor_sync(FULL, gmax, off));
    if (lane == 0) s_gw[warp] = gmax;
    gmax_done = true;
    if (write_factors) {
      __syncthreads();
      if (tid == 0) {
        double m = s_gw[0];
#pragma unroll
        for (int w = 1; w < PW_WARPS; ++w) m = fmax(m, s_gw[w]);
        pt_part[(size_t)blockIdx.x * PT_PART] = m;
      }
      // ---- 3. Schur factors
      double2* rows = rows_all[warp];
      if (act) {
        double Lo[6], eo[3];
#pragma unroll
        for (int j = 0; j < 6; ++j) Lo[j] = s_le[tr][j];
#pragma unroll
        for (int j = 0; j < 3; ++j) eo[j] = s_le[tr][6 + j];
        double2 rec[8];
        pack_factor(Q, Lo, prot, eo, rec);
        const int sw = lane & 7;
#pragma unroll
        for (int c = 0; c < 8; ++c) rows[lane * 8 + (c ^ sw)] = rec[c];
      }
      const unsigned actmask = __ballot_sync(FULL, act);  // (also orders the shared-memory writes in the warp)
      double2* fac2 = reinterpret_cast<double2*>(fac);
      const int g4 = lane >> 3, c = lane & 7;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int item = 4 * j + g4;
        if ((actmask >> item) & 1u)
          fac2[(size_t)(base + 32 * warp + item) * 8 + c] = rows[item * 8 + (c ^ (item & 7))];
      }
    }
  } else {
    const int w = g0 + warp;
    if (w < g1)
      gmax = points_warp_path<PLANE>(pv, pose, X, lm, c_pt, compute_scaling, jacobi_scaling, Hpp, gp, Hinv, Linv, fac, scal,
                              write_factors, rows_all[warp], w, lane);
  }
  if (!(gmax_done && write_factors)) {  // warp path, or no factor pass: the plain block reduction
    __shared__ double smx[PW_WARPS + 1];
    const double m = block_max<PW_WARPS>(gmax, smx);
    if (threadIdx.x == 0) pt_part[(size_t)blockIdx.x * PT_PART] = m;
  }
}

// ---- k_points with a bulk-copy pipeline (experiment of round 2) -----------------------------
// Same arithmetic as k_points' CTA path, group by group, but PERSISTENT (a few CTAs per SM walk the groups with a
// grid stride) and fed by cp.async.bulk: everything a group reads contiguously -- keypoints, the three depth-prior
// arrays, pm_pt, pm_img, the coordinates / pt_ptr / constant flags of its tracks -- lands in a two-stage shared
// memory ring, completion signalled on an mbarrier (complete_tx), one group ahead of the arithmetic.  What stays a
// gather is the image record behind pm_img (pose, camera: L1 / L2 resident).  The L2 prefetch hints of k_points
// (one wave ahead, issued by two warps of every CTA) disappear.  Source ranges are widened to 16-byte boundaries
// (cp.async.bulk's alignment rule); the in-stage offset of the first element follows from the group descriptor.
struct PtStage {
  static constexpr int XY = 0, D0 = 2064, D1 = D0 + 1040, D2 = D1 + 1040, PT = D2 + 1040, IMG = PT + 528, X = IMG + 528,
                       PTR = X + 4096, CST = PTR + 544, CPT = CST + 144, BYTES = CPT + 3088;
};
static_assert(PtStage::BYTES % 16 == 0, "stage size");
constexpr int PB_SMEM = 2 * PtStage::BYTES + PW_THREADS * 9 * 8 + PW_WARPS * 256 * 16;  // stages | s_le | s_v / rows

__device__ __forceinline__ void pb_mbar_init(uint64_t* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"((unsigned)__cvta_generic_to_shared(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void pb_mbar_expect(uint64_t* bar, unsigned bytes) {
  asm volatile("{\n .reg .b64 st;\n mbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n}" ::"r"(
                   (unsigned)__cvta_generic_to_shared(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void pb_mbar_wait(uint64_t* bar, unsigned parity) {
  const unsigned addr = (unsigned)__cvta_generic_to_shared(bar);
  unsigned done;
  do {
    asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
                 : "=r"(done) : "r"(addr), "r"(parity) : "memory");
  } while (!done);
}
// bulk copy of the 16-byte aligned cover of [src, src + bytes) to dst; returns the bytes moved
__device__ __forceinline__ unsigned pb_bulk(void* dst, const void* src, size_t bytes, uint64_t* bar) {
  const uintptr_t s0 = reinterpret_cast<uintptr_t>(src) & ~(uintptr_t)15;
  const uintptr_t s1 = (reinterpret_cast<uintptr_t>(src) + bytes + 15) & ~(uintptr_t)15;
  const unsigned n = (unsigned)(s1 - s0);
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   (unsigned)__cvta_generic_to_shared(dst)), "l"(s0), "r"(n), "r"((unsigned)__cvta_generic_to_shared(bar))
               : "memory");
  return n;
}
__device__ __forceinline__ int pb_skew(const void* src) { return (int)(reinterpret_cast<uintptr_t>(src) & 15); }

__global__ void __launch_bounds__(PW_THREADS, MPBA_POINTS_MINB) k_points_bulk(
    ProblemView pv, const double* __restrict__ pose, const double* __restrict__ X, LmDiag lm, double* __restrict__ c_pt,
    int compute_scaling, int jacobi_scaling, double* __restrict__ Hpp, double* __restrict__ gp, double* __restrict__ Hinv,
    double* __restrict__ Linv, double* __restrict__ fac, double* __restrict__ pt_part, double* __restrict__ scal,
    int write_factors, int num_groups) {
  extern __shared__ __align__(128) unsigned char pb_smem[];
  __shared__ uint64_t s_bar[2];
  __shared__ __align__(16) int4 s_desc[2];
  __shared__ double smx2[2][PW_WARPS];  // warp maxima of |gp|, by ring parity: thread 0 reads them one barrier later
  unsigned char* stage0 = pb_smem;
  double (*s_le)[9] = reinterpret_cast<double (*)[9]>(pb_smem + 2 * PtStage::BYTES);
  unsigned char* uni = pb_smem + 2 * PtStage::BYTES + PW_THREADS * 9 * 8;
  double (*s_v)[9] = reinterpret_cast<double (*)[9]>(uni);       // phases 1-2
  double2* rows_all = reinterpret_cast<double2*>(uni);            // phase 3 (after a CTA barrier)
  if (compute_scaling < 0) compute_scaling = __ldcg(lm.dev + LD_FIRST) != 0.0 ? 1 : 0;
  lm_resolve(lm);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const size_t N = (size_t)pv.N;
  if (tid == 0) {
    pb_mbar_init(&s_bar[0], 1);
    pb_mbar_init(&s_bar[1], 1);
  }
  __syncthreads();
  // thread 0: descriptor of group g into s_desc[st] and the group's bulk copies into stage st
  auto issue = [&](int g, int st) {
    const int4 d = __ldg(pv.cta_desc + g);
    s_desc[st] = d;
    if (d.w > PW_THREADS) {  // oversized group: the warp path reads global memory itself
      pb_mbar_expect(&s_bar[st], 0);
      return;
    }
    const size_t T0 = (size_t)d.x, nt = (size_t)d.y, base = (size_t)d.z, n = (size_t)d.w;
    unsigned char* sb = stage0 + (size_t)st * PtStage::BYTES;
    const void* src[10] = {pv.pm_xy + 2 * base, pv.pm_depth + base, pv.pm_depth + N + base, pv.pm_depth + 2 * N + base,
                           pv.pm_pt + base,     pv.pm_img + base,   X + 4 * T0,             pv.pt_ptr + T0,
                           pv.pt_const + T0,    c_pt + 3 * T0};
    const size_t len[10] = {n * 16, n * 8, n * 8, n * 8, n * 4, n * 4, nt * 32, (nt + 1) * 4, nt,
                            compute_scaling ? 0 : nt * 24};  // the Jacobi scales exist after the first iteration
    const int off[10] = {PtStage::XY, PtStage::D0, PtStage::D1, PtStage::D2, PtStage::PT, PtStage::IMG, PtStage::X,
                         PtStage::PTR, PtStage::CST, PtStage::CPT};
    unsigned total = 0;
#pragma unroll
    for (int k = 0; k < 10; ++k) {
      const uintptr_t a = reinterpret_cast<uintptr_t>(src[k]);
      if (len[k]) total += (unsigned)(((a + len[k] + 15) & ~(uintptr_t)15) - (a & ~(uintptr_t)15));
    }
    pb_mbar_expect(&s_bar[st], total);
#pragma unroll
    for (int k = 0; k < 10; ++k)
      if (len[k]) pb_bulk(sb + off[k], src[k], len[k], &s_bar[st]);
  };
  int g = blockIdx.x;
  if (tid == 0 && g < num_groups) issue(g, 0);
  for (int it = 0; g < num_groups; ++it, g += gridDim.x) {
    const int st = it & 1;
    double* smx = smx2[st];
    // the stage of the NEXT group was last read two groups ago: every thread is past that (barrier at the loop's end)
    if (tid == 0 && g + (int)gridDim.x < num_groups) issue(g + gridDim.x, st ^ 1);
    pb_mbar_wait(&s_bar[st], (unsigned)(it >> 1) & 1u);  // stage st completes its (it / 2)-th phase
    const int4 desc = s_desc[st];
    const int T0 = desc.x, nt = desc.y, base = desc.z, n = desc.w;
    const unsigned char* sb = stage0 + (size_t)st * PtStage::BYTES;
    double gmax = 0.0;
    if (n <= PW_THREADS) {
      const double2* s_xy = reinterpret_cast<const double2*>(sb + PtStage::XY + pb_skew(pv.pm_xy + 2 * (size_t)base));
      const double* s_d0 = reinterpret_cast<const double*>(sb + PtStage::D0 + pb_skew(pv.pm_depth + base));
      const double* s_d1 = reinterpret_cast<const double*>(sb + PtStage::D1 + pb_skew(pv.pm_depth + N + base));
      const double* s_d2 = reinterpret_cast<const double*>(sb + PtStage::D2 + pb_skew(pv.pm_depth + 2 * N + base));
      const int* s_pt = reinterpret_cast<const int*>(sb + PtStage::PT + pb_skew(pv.pm_pt + base));
      const int* s_img = reinterpret_cast<const int*>(sb + PtStage::IMG + pb_skew(pv.pm_img + base));
      const double2* s_X = reinterpret_cast<const double2*>(sb + PtStage::X);  // 32-byte records: no skew
      const int* s_ptr = reinterpret_cast<const int*>(sb + PtStage::PTR + pb_skew(pv.pt_ptr + T0));
      const uint8_t* s_cst = sb + PtStage::CST + pb_skew(pv.pt_const + T0);
      const double* s_cpt = reinterpret_cast<const double*>(sb + PtStage::CPT + pb_skew(c_pt + 3 * (size_t)T0));
      int start = 0, len = 0;
      bool cst = true;
      if (tid < nt) {
        const int a = s_ptr[tid], b = s_ptr[tid + 1];
        start = a - base;
        len = b - a;
        cst = s_cst[tid] != 0;
      }
      // ---- 1. observations
      int tr = -1;
      double Q[9], prot[3] = {0.0, 0.0, 0.0};
      if (tid < n) {
        const int pt = s_pt[tid];
        const double2 xa = s_X[2 * (pt - T0)], xb = s_X[2 * (pt - T0) + 1];
        if (xb.y == 0.0) {  // slot 3: constant flag
          tr = pt - T0;
          const double Xp[3] = {xa.x, xa.y, xb.x};
          ObsLin o;
          ImgConst ic;
          load_img(pv, pose, s_img[tid], ic);
          const double2 o2 = s_xy[tid];
          linearize_obs<true, false>(ic.pose, ic.R, Xp, ic.model, ic.prm, o2.x, o2.y, s_d0[tid], s_d1[tid], s_d2[tid], ic.b,
                                     ic.es, pv.loss, o);
          double v[9];
          point_terms(o, ic.R, Q, v);
#pragma unroll
          for (int j = 0; j < 9; ++j) s_v[tid][j] = v[j];
          prot[0] = o.prot[0]; prot[1] = o.prot[1]; prot[2] = o.prot[2];
        }
      }
      const bool act = tr >= 0;
      __syncthreads();
      // ---- 2. points
      if (tid < nt) {
        const bool active = !cst && len > 0;
        double v[9] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
        if (active) {
          for (int k = 0; k < len; ++k) {
#pragma unroll
            for (int j = 0; j < 9; ++j) v[j] += s_v[start + k][j];
          }
        }
        double Li[6], e[3];
        finalize_point(T0 + tid, active, v, v + 6, lm, c_pt, compute_scaling, jacobi_scaling, Hpp, gp, Hinv, Linv, scal,
                       Li, e, compute_scaling ? nullptr : s_cpt + 3 * tid);
#pragma unroll
        for (int j = 0; j < 6; ++j) s_le[tid][j] = Li[j];
#pragma unroll
        for (int j = 0; j < 3; ++j) s_le[tid][6 + j] = e[j];
        gmax = fmax(fabs(v[6]), fmax(fabs(v[7]), fabs(v[8])));
      }
      for (int off = 16; off > 0; off >>= 1) gmax = fmax(gmax, __shfl_xor_sync(FULL, gmax, off));
      if (lane == 0) smx[warp] = gmax;
      if (write_factors) {
        __syncthreads();  // s_le complete; s_v no longer read: its bytes become the record rows
        // ---- 3. Schur factors
        double2* rows = rows_all + warp * 256;
        if (act) {
          double Lo[6], eo[3];
#pragma unroll
          for (int j = 0; j < 6; ++j) Lo[j] = s_le[tr][j];
#pragma unroll
          for (int j = 0; j < 3; ++j) eo[j] = s_le[tr][6 + j];
          double2 rec[8];
          pack_factor(Q, Lo, prot, eo, rec);
          const int sw = lane & 7;
#pragma unroll
          for (int c = 0; c < 8; ++c) rows[lane * 8 + (c ^ sw)] = rec[c];
        }
        const unsigned actmask = __ballot_sync(FULL, act);
        double2* fac2 = reinterpret_cast<double2*>(fac);
        const int g4 = lane >> 3, c = lane & 7;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int item = 4 * j + g4;
          if ((actmask >> item) & 1u) fac2[(size_t)(base + 32 * warp + item) * 8 + c] = rows[item * 8 + (c ^ (item & 7))];
        }
      }
    } else {
      const int w = g * PW_WARPS + warp;
      if (w < min((g + 1) * PW_WARPS, pv.num_track_warps))
        gmax = points_warp_path<false>(pv, pose, X, lm, c_pt, compute_scaling, jacobi_scaling, Hpp, gp, Hinv, Linv, fac, scal,
                                write_factors, rows_all + warp * 256, w, lane);
    }
    if (n > PW_THREADS) {
      for (int off = 16; off > 0; off >>= 1) gmax = fmax(gmax, __shfl_xor_sync(FULL, gmax, off));
      if (lane == 0) smx[warp] = gmax;
    }
    __syncthreads();  // the warp maxima are in place; also the loop-end barrier of the ring (stage, s_v / rows, s_le)
    if (tid == 0) {
      double m = smx[0];
#pragma unroll
      for (int w = 1; w < PW_WARPS; ++w) m = fmax(m, smx[w]);
      pt_part[(size_t)g * PT_PART] = m;
    }
  }
}

__global__ void __launch_bounds__(1024) k_pt_scalars_gmax(const double* __restrict__ pt_part, int nblk,
                                                          double* __restrict__ scal) {
  __shared__ double sm[33];
  double gm = 0.0;
  // four independent loads in flight per thread: the kernel is one CTA and nothing but load latency
  int i = threadIdx.x;
  for (; i + 3 * 1024 < nblk; i += 4 * 1024) {
    const double a = pt_part[(size_t)i * PT_PART], b = pt_part[(size_t)(i + 1024) * PT_PART],
                 c = pt_part[(size_t)(i + 2048) * PT_PART], d = pt_part[(size_t)(i + 3072) * PT_PART];
    gm = fmax(fmax(gm, a), fmax(b, fmax(c, d)));
  }
  for (; i < nblk; i += 1024) gm = fmax(gm, pt_part[(size_t)i * PT_PART]);
  gm = block_max<32>(gm, sm);
  if (threadIdx.x == 0) scal[S_GMAX_P] = gm;
}

// ================================================================================================
// reduced camera system
// ================================================================================================
// Inverse of the SPD 6x6 block in Mm (shared memory, 36 doubles; destroyed) by ONE warp: Gauss-Jordan on [M | N], N (36
// doubles of shared memory) starting as the identity; lane = entries lane and lane + 32, six steps of ~100 cycles.
// The symmetrised result goes to dst; a pivot that is not positive (block not positive definite) makes it the identity.
__device__ __forceinline__ void warp_inverse6(double* Mm, double* Nn, int lane, double* __restrict__ dst) {
  bool ok = true;
#pragma unroll
  for (int p = 0; p < 6; ++p) {
    const double piv = Mm[7 * p];
    ok = ok && (piv > 0.0);
    const double inv = 1.0 / (ok ? piv : 1.0);
    double m2[2], n2[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int e = lane + 32 * h;
      m2[h] = 0.0; n2[h] = 0.0;
      if (e < 36) {
        const int r = e / 6, c = e % 6;
        const double f = Mm[6 * r + p];
        const double mp = Mm[6 * p + c] * inv, np_ = Nn[6 * p + c] * inv;
        m2[h] = (r == p) ? mp : Mm[e] - f * mp;
        n2[h] = (r == p) ? np_ : Nn[e] - f * np_;
      }
    }
    __syncwarp();
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int e = lane + 32 * h;
      if (e < 36) {
        Mm[e] = m2[h];
        Nn[e] = n2[h];
      }
    }
    __syncwarp();
  }
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const int e = lane + 32 * h;
    if (e < 36) {
      const int r = e / 6, c = e % 6;
      dst[e] = ok ? 0.5 * (Nn[6 * r + c] + Nn[6 * c + r]) : ((r == c) ? 1.0 : 0.0);
    }
  }
}

constexpr int SB_THREADS = 128;
// per warp and buffer: 32 items x (record a: 8 chunks, record b: 8 chunks) of 16 bytes = 8 KB; two buffers, 4 warps
constexpr int SB_WARP_CHUNKS = 512;
constexpr size_t SB_STAGE_BYTES = 2 * (SB_THREADS / 32) * SB_WARP_CHUNKS * sizeof(double2);

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gmem_src) {
  const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(d), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory");
}

// Gather of 128-byte factor records by a warp: lane = (record-in-group-of-4 << 3 | 16-byte chunk), so one
// instruction moves four whole 128-byte lines (4 L1 wavefronts, every 32-byte sector requested once) instead
// of 32 lanes touching 32 different lines for 16 bytes each (32 wavefronts, every sector requested twice --
// what the first version of this kernel did, and what bound it: ncu l1tex 87 %, 5.7 GB of L2 reads for 2.9 GB
// of records).  Record i of the warp lands in shared memory at row i with its chunks XOR-swizzled by (i & 7):
// the copies of one instruction fill four rows, and lane i reading chunk c of ITS row hits 8 distinct 16-byte
// columns per quarter warp -> no bank conflicts either way.
// The destination of copy j of a lane is rows + item * 8 + (c ^ (item & 7)) with item = 4 j + g; since g < 4 that is
// (8 g + (c ^ g)) + 32 j for even j and (8 g + ((c ^ g) ^ 4)) + 32 j for odd j: two lane constants and an immediate
// (the kernel is bound by instruction issue; the XOR form cost three integer instructions per copy).
__device__ __forceinline__ void warp_gather_records(const double2* __restrict__ fac2, int slot, int nchunks,
                                                    double2* rows /* [32][8] */, int lane) {
  const int g = lane >> 3, c = lane & 7;
  double2* const even = rows + 8 * g + (c ^ g);
  double2* const odd = rows + 8 * g + ((c ^ g) ^ 4);
  const bool want = c < nchunks;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int s = __shfl_sync(FULL, slot, 4 * j + g);
    if (s >= 0 && want) cp_async16(((j & 1) ? odd : even) + 32 * j, fac2 + (size_t)s * 8 + c);
  }
}

// One CTA per upper block (i <= j) of the reduced camera system, no atomics, fixed summation order:
//   T(i,j) = sum over the tracks seen by both images of G_a^T (Z_a Z_b^T) G_b
// is accumulated in registers over the block's item list (pairs of factor slots, sorted by track),
// reduced across the CTA (shuffle transpose + shared memory), and written once as
//   S(i,i) = Hcc_i + D_i - T(i,i),  S(i,j) = -T(i,j),  S(j,i) = S(i,j)^T,  rhs_i = gc_i - sum G_a^T Z_a e
// with identity rows / zero rhs on fixed tangent coordinates (gauge, bundle_adjustment.py:114-121).
// BULK (MPBA_SCHUR_BULK=1, a measured experiment, NOT the default): every lane fetches the two records of ITS item with
// one cp.async.bulk each (128 B and 96 B, completion on a per-warp mbarrier) instead of the warp issuing sixteen 16-byte
// cp.async per lane with a shuffle each -- the gather is a third of the instructions of an issue-bound kernel.  Rows are
// padded to 9 resp. 7 chunks so that lane i reading chunk c of its own row is conflict free without the XOR swizzle.
// Same results, but 714 us against 309 us at cfg3: 30 M bulk copies of ~112 B are bound by the rate of the copy unit
// (one per ~6.7 cycles per SM), not by the issue slots they free (profiles/r02_schur_blocks_bulk.md).
template <bool BULK>
__global__ void __launch_bounds__(SB_THREADS) k_schur_blocks(ProblemView pv, const double* __restrict__ fac,
                                                             const double* __restrict__ Hc,
                                                             const double* __restrict__ gc,
                                                             const double* __restrict__ c_img, LmDiag lm,
                                                             double* __restrict__ S, double* __restrict__ rhs,
                                                             int partial, double* __restrict__ Minv) {
  lm_resolve(lm);
  const int u = blockIdx.x;
  const int blk = pv.u_block[u];
  const int i = pv.s_row[blk], j = pv.s_col[blk];
  const bool diag = (i == j);
  const int it0 = pv.u_item_ptr[u], it1 = pv.u_item_ptr[u + 1];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;

  double M[36];
#pragma unroll
  for (int k = 0; k < 36; ++k) M[k] = 0.0;
  double rv[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};

  // Every thread owns one item per 128-item tile; the two factor records of the items of a warp are gathered
  // by the warp together (warp_gather_records) into its own shared-memory rows with cp.async, two tiles deep,
  // so the loads of tile t+1 are in flight while tile t is multiplied.  Only warp-level synchronisation.
  extern __shared__ __align__(16) double2 stage[];  // [warp][2 buffers][a rows 32x8 | b rows 32x8]
  const double2* fac2 = reinterpret_cast<const double2*>(fac);
  double2* wst = stage + (size_t)warp * 2 * SB_WARP_CHUNKS;
  const int tid = threadIdx.x;
  __shared__ uint64_t s_bar[SB_THREADS / 32][2];
  if (BULK) {
    if (lane == 0) {
      pb_mbar_init(&s_bar[warp][0], 1);
      pb_mbar_init(&s_bar[warp][1], 1);
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncwarp();
  }
  auto issue = [&](int buf, int ka, int kb) {
    double2* rows = wst + buf * SB_WARP_CHUNKS;
    if (BULK) {
      // rows a: [32][9] chunks (288 of the 512), rows b: [32][7] chunks from chunk 288
      const bool hb = kb >= 0 && kb != ka;
      const unsigned na = __popc(__ballot_sync(FULL, ka >= 0)), nb = __popc(__ballot_sync(FULL, hb));
      if (na == 0) return;  // past the end of the item list: no phase is opened (and none is waited for)
      if (lane == 0) pb_mbar_expect(&s_bar[warp][buf], na * 128u + nb * 96u);
      __syncwarp();
      const unsigned bar = (unsigned)__cvta_generic_to_shared(&s_bar[warp][buf]);
      if (ka >= 0)
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], 128, [%2];" ::"r"(
                         (unsigned)__cvta_generic_to_shared(rows + lane * 9)), "l"(fac2 + (size_t)ka * 8), "r"(bar)
                     : "memory");
      if (hb)
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], 96, [%2];" ::"r"(
                         (unsigned)__cvta_generic_to_shared(rows + 288 + lane * 7)), "l"(fac2 + (size_t)kb * 8), "r"(bar)
                     : "memory");
      return;
    }
    warp_gather_records(fac2, ka, 8, rows, lane);
    warp_gather_records(fac2, kb == ka ? -1 : kb, 6, rows + 256, lane);  // Z and p_rot of b; e is not needed
  };
  int it = it0 + tid;
  int ka = -1, kb = -1, ka1 = -1, kb1 = -1;
  if (it < it1) {
    ka = __ldg(pv.item_a + it);
    kb = __ldg(pv.item_b + it);
  }
  issue(0, ka, kb);
  if (!BULK) cp_async_commit();
  if (it + SB_THREADS < it1) {
    ka1 = __ldg(pv.item_a + it + SB_THREADS);
    kb1 = __ldg(pv.item_b + it + SB_THREADS);
  }
  for (int tile = 0;; ++tile) {
    const int base = it0 + tile * SB_THREADS;
    if (base >= it1) break;
    const bool cur = it < it1;
    __syncwarp();  // the rows about to be overwritten were read two tiles ago by every lane
    issue((tile + 1) & 1, ka1, kb1);
    if (!BULK) cp_async_commit();
    int ka2 = -1, kb2 = -1;
    if (it + 2 * SB_THREADS < it1) {
      ka2 = __ldg(pv.item_a + it + 2 * SB_THREADS);
      kb2 = __ldg(pv.item_b + it + 2 * SB_THREADS);
    }
    if (BULK) {
      // the warp has items in this tile iff its first lane has one (items of a tile are contiguous over the threads)
      if (base + warp * 32 < it1) pb_mbar_wait(&s_bar[warp][tile & 1], (unsigned)(tile >> 1) & 1u);
    } else {
      cp_async_wait<1>();
      __syncwarp();  // every lane's copies of this tile have landed
    }
    if (cur) {
      const double2* src = wst + (tile & 1) * SB_WARP_CHUNKS + (BULK ? lane * 9 : lane * 8);
      const double2* srcb = wst + (tile & 1) * SB_WARP_CHUNKS + (BULK ? 288 + lane * 7 : 256 + lane * 8);
      const int sw = BULK ? 0 : (lane & 7);
      double za[9], zb[9], qa[3], qb[3];
      {
        const double2 c0 = src[0 ^ sw], c1 = src[1 ^ sw], c2 = src[2 ^ sw], c3 = src[3 ^ sw], c4 = src[4 ^ sw],
                      c5 = src[5 ^ sw];
        za[0] = c0.x; za[1] = c0.y; za[2] = c1.x; za[3] = c1.y; za[4] = c2.x; za[5] = c2.y; za[6] = c3.x; za[7] = c3.y;
        za[8] = c4.x; qa[0] = 2.0 * c4.y; qa[1] = 2.0 * c5.x; qa[2] = 2.0 * c5.y;
      }
      const bool same = (ka == kb);
      if (same) {
        const double2 c6 = src[6 ^ sw], c7 = src[7 ^ sw];
        const double e0 = c6.x, e1 = c6.y, e2 = c7.x;
        // rhs: G_a^T (Z_a e)
        const double z0 = za[0] * e0 + za[1] * e1 + za[2] * e2, z1 = za[3] * e0 + za[4] * e1 + za[5] * e2,
                     z2 = za[6] * e0 + za[7] * e1 + za[8] * e2;
        rv[0] += qa[1] * z2 - qa[2] * z1;
        rv[1] += qa[2] * z0 - qa[0] * z2;
        rv[2] += qa[0] * z1 - qa[1] * z0;
        rv[3] += z0; rv[4] += z1; rv[5] += z2;
#pragma unroll
        for (int k = 0; k < 9; ++k) zb[k] = za[k];
        qb[0] = qa[0]; qb[1] = qa[1]; qb[2] = qa[2];
      } else {
        const double2 c0 = srcb[0 ^ sw], c1 = srcb[1 ^ sw], c2 = srcb[2 ^ sw], c3 = srcb[3 ^ sw], c4 = srcb[4 ^ sw],
                      c5 = srcb[5 ^ sw];
        zb[0] = c0.x; zb[1] = c0.y; zb[2] = c1.x; zb[3] = c1.y; zb[4] = c2.x; zb[5] = c2.y; zb[6] = c3.x; zb[7] = c3.y;
        zb[8] = c4.x; qb[0] = 2.0 * c4.y; qb[1] = 2.0 * c5.x; qb[2] = 2.0 * c5.y;
      }
      // K = Za Zb^T;  M += [Pa^T K Pb, Pa^T K; K Pb, K]  with P = -2 [p]x, q = 2 p:
      //   (K P)[r]   = ( q_y K[r][2] - q_z K[r][1],  q_z K[r][0] - q_x K[r][2],  q_x K[r][1] - q_y K[r][0] )
      //   (P^T Y)[.][c] = ( q_y Y[2][c] - q_z Y[1][c],  q_z Y[0][c] - q_x Y[2][c],  q_x Y[1][c] - q_y Y[0][c] )
      double K[9], KP[9];
#pragma unroll
      for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c)
          K[3 * r + c] = za[3 * r] * zb[3 * c] + za[3 * r + 1] * zb[3 * c + 1] + za[3 * r + 2] * zb[3 * c + 2];
#pragma unroll
      for (int r = 0; r < 3; ++r) {
        KP[3 * r] = qb[1] * K[3 * r + 2] - qb[2] * K[3 * r + 1];
        KP[3 * r + 1] = qb[2] * K[3 * r] - qb[0] * K[3 * r + 2];
        KP[3 * r + 2] = qb[0] * K[3 * r + 1] - qb[1] * K[3 * r];
      }
      if (!diag || same) {
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          // rows 0..2: P_a^T [KP | K], rows 3..5: [KP | K]
          // two fused multiply-adds into the accumulator instead of (product, fused difference, add): a third fewer
          // instructions on the 18 cross-product entries of an item in a kernel bound by instruction issue
          M[c] = fma(qa[1], KP[6 + c], fma(-qa[2], KP[3 + c], M[c]));
          M[6 + c] = fma(qa[2], KP[c], fma(-qa[0], KP[6 + c], M[6 + c]));
          M[12 + c] = fma(qa[0], KP[3 + c], fma(-qa[1], KP[c], M[12 + c]));
          M[3 + c] = fma(qa[1], K[6 + c], fma(-qa[2], K[3 + c], M[3 + c]));
          M[9 + c] = fma(qa[2], K[c], fma(-qa[0], K[6 + c], M[9 + c]));
          M[15 + c] = fma(qa[0], K[3 + c], fma(-qa[1], K[c], M[15 + c]));
          M[18 + c] += KP[c];
          M[24 + c] += KP[3 + c];
          M[30 + c] += KP[6 + c];
          M[21 + c] += K[c];
          M[27 + c] += K[3 + c];
          M[33 + c] += K[6 + c];
        }
      } else {
        // one image observing a track twice: both orientations land on the diagonal block (rare)
        double Mi[36];
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          Mi[c] = qa[1] * KP[6 + c] - qa[2] * KP[3 + c];
          Mi[6 + c] = qa[2] * KP[c] - qa[0] * KP[6 + c];
          Mi[12 + c] = qa[0] * KP[3 + c] - qa[1] * KP[c];
          Mi[3 + c] = qa[1] * K[6 + c] - qa[2] * K[3 + c];
          Mi[9 + c] = qa[2] * K[c] - qa[0] * K[6 + c];
          Mi[15 + c] = qa[0] * K[3 + c] - qa[1] * K[c];
          Mi[18 + c] = KP[c];
          Mi[24 + c] = KP[3 + c];
          Mi[30 + c] = KP[6 + c];
          Mi[21 + c] = K[c];
          Mi[27 + c] = K[3 + c];
          Mi[33 + c] = K[6 + c];
        }
#pragma unroll
        for (int r = 0; r < 6; ++r)
#pragma unroll
          for (int c = 0; c < 6; ++c) M[6 * r + c] += Mi[6 * r + c] + Mi[6 * c + r];
      }
    }
    it += SB_THREADS;
    ka = ka1; kb = kb1;
    ka1 = ka2; kb1 = kb2;
  }
  if (!BULK) cp_async_wait<0>();

  // ---- CTA reduction: lane e of every warp ends with entry e (e < 32), entries 32..35 by butterfly
  __shared__ double sm[SB_THREADS / 32][44];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    double t = M[32 + k];
    for (int off = 16; off > 0; off >>= 1) t += __shfl_xor_sync(FULL, t, off);
    if (lane == k) sm[warp][32 + k] = t;
  }
  if (diag) {
#pragma unroll
    for (int k = 0; k < 6; ++k) {
      double t = rv[k];
      for (int off = 16; off > 0; off >>= 1) t += __shfl_xor_sync(FULL, t, off);
      if (lane == k) sm[warp][36 + k] = t;
    }
  }
  sm[warp][lane] = warp_transpose_reduce32(M, lane);
  __syncthreads();
  if (threadIdx.x < 42) {
    double t = 0.0;
#pragma unroll
    for (int w = 0; w < SB_THREADS / 32; ++w) t += sm[w][threadIdx.x];
    sm[0][threadIdx.x] = t;  // only this thread touches column threadIdx.x after the barrier
  }
  __syncthreads();

  // ---- epilogue
  const unsigned mi = pv.img_free_mask[i], mj = pv.img_free_mask[j];
  if (threadIdx.x < 36) {
    const int e = threadIdx.x, r = e / 6, c = e % 6;
    const bool fr = (mi >> r) & 1u, fc = (mj >> c) & 1u;
    double val;
    if (diag && partial) {
      // this rank's tracks only; k_schur_finish completes the block after the all-reduce
      S[(size_t)blk * 36 + e] = (fr && fc) ? -0.5 * (sm[0][e] + sm[0][c * 6 + r]) : 0.0;
    } else if (diag) {
      if (fr && fc) {
        const int uidx = upper_idx(r < c ? r : c, r < c ? c : r);
        val = Hc[(size_t)i * 21 + uidx] - 0.5 * (sm[0][e] + sm[0][c * 6 + r]);
        if (r == c) val += lm_damping(Hc[(size_t)i * 21 + uidx], c_img[(size_t)i * 6 + r], lm.radius, lm.lo, lm.hi);
      } else {
        val = (r == c) ? 1.0 : 0.0;
      }
      S[(size_t)blk * 36 + e] = val;
    } else {
      val = (fr && fc) ? -sm[0][e] : 0.0;
      S[(size_t)blk * 36 + e] = val;
      S[(size_t)pv.s_transpose[blk] * 36 + c * 6 + r] = val;
    }
  } else if (diag && threadIdx.x >= 64 && threadIdx.x < 70) {
    const int r = threadIdx.x - 64;
    rhs[(size_t)i * 6 + r] = ((mi >> r) & 1u) ? (partial ? 0.0 : gc[(size_t)i * 6 + r]) - sm[0][36 + r] : 0.0;
  }
  if (diag && !partial && Minv != nullptr) {  // CTA-uniform: the block-Jacobi inverse of the block just written
    // Gauss-Jordan on [M | N] (N starts as the identity) by ONE warp in shared memory, lane = entries lane and lane + 32:
    // six steps of ~100 cycles.  (One thread running the Cholesky-based block_inverse6 here put its work arrays in
    // local memory at this kernel's 162 registers and cost the diagonal CTAs 16 us.)  A pivot that is not positive
    // (the block is not positive definite) makes the block's preconditioner the identity, as k_block_inverse does.
    double* Mm = sm[1];
    double* Nn = sm[2];
    if (threadIdx.x < 36) {
      Mm[threadIdx.x] = S[(size_t)blk * 36 + threadIdx.x];  // own store: reads back its value
      Nn[threadIdx.x] = (threadIdx.x % 7 == 0) ? 1.0 : 0.0;
    }
    __syncthreads();
    if (warp == 0) warp_inverse6(Mm, Nn, lane, Minv + (size_t)i * 36);
  }
}

// Sharded solve, after the all-reduce of the partial blocks: S(i,i) += Hcc_i + D_i, rhs_i += gc_i, identity rows
// on fixed tangent coordinates.  One thread per entry of a diagonal block (+ 6 for rhs).
__global__ void __launch_bounds__(64) k_schur_finish(ProblemView pv, const double* __restrict__ Hc,
                                                     const double* __restrict__ gc, const double* __restrict__ c_img,
                                                     LmDiag lm, double* __restrict__ S, double* __restrict__ rhs,
                                                     double* __restrict__ Minv) {
  __shared__ double Mm[36], Nn[36];
  lm_resolve(lm);
  const int i = blockIdx.x;
  const unsigned mi = pv.img_free_mask[i];
  const int blk = pv.s_diag[i];
  if (threadIdx.x < 36) {
    const int e = threadIdx.x, r = e / 6, c = e % 6;
    const bool fr = (mi >> r) & 1u, fc = (mi >> c) & 1u;
    double val;
    if (fr && fc) {
      const int uidx = upper_idx(r < c ? r : c, r < c ? c : r);
      val = Hc[(size_t)i * 21 + uidx] + S[(size_t)blk * 36 + e];
      if (r == c) val += lm_damping(Hc[(size_t)i * 21 + uidx], c_img[(size_t)i * 6 + r], lm.radius, lm.lo, lm.hi);
    } else {
      val = (r == c) ? 1.0 : 0.0;
    }
    S[(size_t)blk * 36 + e] = val;
    Mm[e] = val;
    Nn[e] = (r == c) ? 1.0 : 0.0;
  } else if (threadIdx.x < 42) {
    const int r = threadIdx.x - 36;
    rhs[(size_t)i * 6 + r] = ((mi >> r) & 1u) ? gc[(size_t)i * 6 + r] + rhs[(size_t)i * 6 + r] : 0.0;
  }
  __syncthreads();
  // block-Jacobi inverse of the completed diagonal block (the single-GPU solve does this in k_schur_blocks)
  if (threadIdx.x < 32 && Minv != nullptr) warp_inverse6(Mm, Nn, threadIdx.x, Minv + (size_t)i * 36);
}

// Per-rank scalar blocks, all-gathered -> one block, identical on every rank (fixed rank order).
__global__ void k_combine_scalars(const double* __restrict__ scal_all, int world, double* __restrict__ scal) {
  const int k = threadIdx.x;
  if (k >= S_COUNT) return;
  const bool is_sum = k == S_MCC || k == S_STEP2_P || k == S_XNORM2_P || k == S_STEP2_C || k == S_XNORM2_C ||
                      k == S_BAD_POINTS;
  const bool is_max = k == S_GMAX_P;
  if (!is_sum && !is_max) return;  // cost, |gc|, PCG scalars are computed from replicated data
  double v = 0.0;
  for (int r = 0; r < world; ++r) {
    const double x = scal_all[(size_t)r * S_COUNT + k];
    v = is_sum ? v + x : fmax(v, x);
  }
  scal[k] = v;
}

// ---- persistent cooperative PCG on the block-CSR reduced camera system --------------------------
// Block-Jacobi preconditioned conjugate gradients, one launch per linear solve.  The system is small
// (6C unknowns, a few MB of 6x6 blocks) and the solve is latency bound, so the kernel is built around
// the cost of a grid-wide exchange:
//   * each CTA owns a contiguous range of block rows and keeps everything row-local in shared memory
//     for the whole solve: its blocks of S and their column indices (when they fit; L2 otherwise), its
//     rows of M^-1 and of x, r, q.  Per iteration only z and p cross CTAs (L2, ld.cg / st.cg),
//   * two grid barriers per iteration instead of three: p <- z + beta p is never published, every CTA
//     rebuilds the entries of p it gathers from z and the previous p (double buffered),
//   * a barrier is one atomic arrive + spin on a monotonic counter; the dot-product partials ride on
//     it through per-CTA slots that every CTA then sums in the same fixed order (deterministic),
//   * (round 2) the product gathers its operand ONCE per iteration: the distinct block columns of the CTA's rows
//     (a bitmap over the images, ranked at kernel start) are pulled from L2 into shared memory by all threads in
//     one step, and the block products read shared memory.  Before, every block product issued its own two L2
//     loads behind the column index: 13.7 k of the 35.6 k cycles of an iteration at 5 000 images (r2 cycle counters).
struct PcgArgs {
  int C, rows_per_cta, max_local_blocks;
  const int32_t* row_ptr;
  const int32_t* col;
  const double* S;
  const double* Minv;
  const double* rhs;
  double* x;
  double* z;
  double* p0;
  double* p1;
  double* partial;  // [2][gridDim.x]
  unsigned* bar;    // zeroed before the launch
  double* scal;
  double tol;
  int max_iter;
  // two-level preconditioner (coarse.cu): G groups of consecutive CTAs' rows, nc = 7 G gauge modes; G == 0: off
  int G, nc;
  const int32_t* grp_row_ptr;  // [G+1] first row of every group (CTA boundaries)
  const double* Zc;            // [C][6][7]
  const double* Ainv;          // [nc][nc] (Z^T S Z)^-1, symmetric
  double* tpart;               // [gridDim.x][8] per-CTA Z^T r of its rows (one group per CTA)
  // halo staging of the product (see k_pcg): capacity in distinct block columns per CTA, 0 = off
  int halo_cap, halo_words;
};

constexpr int PCG_THREADS = 512;  // 16 warps: one block row per warp and round, ~2 rounds at 5 000 images
constexpr int PCG_MAX_NC = COARSE_MAX_DIM;

__device__ __forceinline__ double grid_reduce(double part, double* slots, unsigned* bar, unsigned& gen, double* sm) {
  const double b = block_sum<PCG_THREADS / 32>(part, sm);
  gen += gridDim.x;
  if (threadIdx.x == 0) {
    __stcg(slots + blockIdx.x, b);
    __threadfence();
    atomicAdd(bar, 1u);
    while (*reinterpret_cast<volatile unsigned*>(bar) < gen) {
    }
  }
  __syncthreads();
  double v = 0.0;
  for (int i = threadIdx.x; i < (int)gridDim.x; i += PCG_THREADS) v += __ldcg(slots + i);
  return block_sum<PCG_THREADS / 32>(v, sm);
}

__device__ __forceinline__ void ldcg6(const double* p, double (&v)[6]) {
  const double2* s = reinterpret_cast<const double2*>(p);
  const double2 a = __ldcg(s), b = __ldcg(s + 1), c = __ldcg(s + 2);
  v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y; v[4] = c.x; v[5] = c.y;
}

template <bool SMEM>
__global__ void __launch_bounds__(PCG_THREADS) k_pcg(PcgArgs A) {
  extern __shared__ double s_dyn[];
  __shared__ double sm[PCG_THREADS / 32 + 1];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  constexpr int WPB = PCG_THREADS / 32;
  const int row0 = min(A.C, (int)blockIdx.x * A.rows_per_cta);
  const int row1 = min(A.C, row0 + A.rows_per_cta);
  const int nrows = row1 - row0;
  const int blk_base = A.row_ptr[row0];
  const int nblk = A.row_ptr[row1] - blk_base;
  // shared memory carve: Minv | x | r | q | p | rowptr | [S | col]
  double* s_minv = s_dyn;
  double* s_x = s_minv + (size_t)A.rows_per_cta * 36;
  double* s_r = s_x + (size_t)A.rows_per_cta * 6;
  double* s_q = s_r + (size_t)A.rows_per_cta * 6;
  double* s_p = s_q + (size_t)A.rows_per_cta * 6;
  int* s_rowptr = reinterpret_cast<int*>(s_p + (size_t)A.rows_per_cta * 6);
  double* s_S = reinterpret_cast<double*>(s_rowptr + ((A.rows_per_cta + 4) & ~3));  // 16-byte steps: s_ph is read as double2
  int* s_col = reinterpret_cast<int*>(s_S + (SMEM ? (size_t)A.max_local_blocks * 36 : 0));
  // halo staging (A.halo_cap > 0): p of the distinct columns | column of a slot | slot of a local block | bitmap | ranks
  double* s_ph = reinterpret_cast<double*>(s_col + (SMEM ? ((A.max_local_blocks + 3) & ~3) : 0));
  int* s_ucol = reinterpret_cast<int*>(s_ph + (size_t)A.halo_cap * 6);
  int* s_bslot = s_ucol + A.halo_cap;
  unsigned* s_bits = reinterpret_cast<unsigned*>(s_bslot + (A.halo_cap > 0 ? A.max_local_blocks : 0));
  int* s_pref = reinterpret_cast<int*>(s_bits + A.halo_words);
  __shared__ int s_nuniq;

  for (int k = threadIdx.x; k < nrows * 36; k += PCG_THREADS) s_minv[k] = A.Minv[(size_t)row0 * 36 + k];
  for (int k = threadIdx.x; k <= nrows; k += PCG_THREADS) s_rowptr[k] = A.row_ptr[row0 + k] - blk_base;
  if (SMEM) {
    const double* src = A.S + (size_t)blk_base * 36;
    for (int k = threadIdx.x; k < nblk * 36; k += PCG_THREADS) s_S[k] = src[k];
    for (int k = threadIdx.x; k < nblk; k += PCG_THREADS) s_col[k] = A.col[blk_base + k];
  }
  __syncthreads();
  const double* Sbase = SMEM ? s_S : (A.S + (size_t)blk_base * 36);
  const int* colbase = SMEM ? s_col : (A.col + blk_base);
  bool halo = A.halo_cap > 0;
  int nuniq = 0;
  if (halo) {
    for (int w = threadIdx.x; w < A.halo_words; w += PCG_THREADS) s_bits[w] = 0u;
    __syncthreads();
    for (int k = threadIdx.x; k < nblk; k += PCG_THREADS) {
      const int c = colbase[k];
      atomicOr(&s_bits[c >> 5], 1u << (c & 31));
    }
    __syncthreads();
    if (warp == 0) {  // exclusive prefix of the word popcounts: rank of a column = s_pref[word] + popc(lower bits)
      int run = 0;
      for (int w0 = 0; w0 < A.halo_words; w0 += 32) {
        const int w = w0 + lane;
        const int cnt = w < A.halo_words ? __popc(s_bits[w]) : 0;
        int scan = cnt;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
          const int t = __shfl_up_sync(FULL, scan, off);
          if (lane >= off) scan += t;
        }
        if (w < A.halo_words) s_pref[w] = run + scan - cnt;
        run += __shfl_sync(FULL, scan, 31);
      }
      if (lane == 0) s_nuniq = run;
    }
    __syncthreads();
    nuniq = s_nuniq;
    halo = nuniq <= A.halo_cap;  // uniform in the CTA; a CTA over capacity keeps the per-block loads (same numbers)
    if (halo) {
      for (int w = threadIdx.x; w < A.halo_words; w += PCG_THREADS) {
        unsigned bits = s_bits[w];
        int slot = s_pref[w];
        while (bits) {
          s_ucol[slot++] = 32 * w + (__ffs(bits) - 1);
          bits &= bits - 1;
        }
      }
      for (int k = threadIdx.x; k < nblk; k += PCG_THREADS) {
        const int c = colbase[k];
        s_bslot[k] = s_pref[c >> 5] + __popc(s_bits[c >> 5] & ((1u << (c & 31)) - 1u));
      }
    }
    __syncthreads();
  }

  double* part0 = A.partial;
  double* part1 = A.partial + gridDim.x;
  const int r6 = lane % 6, kb0 = lane / 6;
  unsigned gen = 0;
  long long t_spmv = 0, t_red = 0, t_upd = 0;

  // ---- two-level preconditioner  z = M^-1 r + Z (Z^T S Z)^-1 Z^T r  (coarse.cu).  Every CTA's rows lie in ONE
  // group, so its Z^T r is 7 numbers; they travel with the r'M^-1 r partial through the same grid barrier, then
  // every CTA forms t = Z^T r (all groups, fixed order), e = A^-1 t (redundantly: nc^2 multiply-adds), and
  //     r'z = r'M^-1 r + t'e,        z_i = (M^-1 r)_i + Z_i e_g      for its own rows,
  // publishes z and meets the others at one more barrier before the next product.  That is the FIRST application;
  // inside the loop t follows the recurrence of r (r -= alpha q  =>  t -= alpha Z^T q): the seven numbers a CTA
  // contributes to Z^T q travel with the p'q partial, so an iteration has two grid barriers, not three.
  __shared__ double s_t[PCG_MAX_NC], s_e[PCG_MAX_NC], s_tw[PCG_THREADS / 32][8];
  __shared__ double s_ainv[7 * PCG_MAX_NC];  // the own group's seven rows of (Z^T S Z)^-1
  const bool coarse = A.G > 0;
  int my_group = 0;
  if (coarse)
    for (int g = 0; g < A.G; ++g)
      if (row0 >= A.grp_row_ptr[g]) my_group = g;
  if (coarse)
    for (int k = threadIdx.x; k < 7 * A.nc; k += PCG_THREADS) s_ainv[k] = A.Ainv[(size_t)my_group * 7 * A.nc + k];
  double* s_zj = s_q;  // M^-1 r of the own rows lives in the q rows between the update and the coarse step
  // per-CTA Z^T r: lanes 0..6 of every warp accumulate over the warp's rows, warps folded in order
  auto coarse_publish_t = [&](const double (&tacc)[1]) {
    if (lane < 7) s_tw[warp][lane] = tacc[0];
    __syncthreads();
    if (threadIdx.x < 7) {
      double t = 0.0;
#pragma unroll
      for (int w = 0; w < PCG_THREADS / 32; ++w) t += s_tw[w][threadIdx.x];
      __stcg(A.tpart + (size_t)blockIdx.x * 8 + threadIdx.x, t);
    }
  };
  // after the exchange: t (all groups), e of the own group; returns this CTA's share of t'e (thread 0 only)
  auto coarse_solve = [&]() -> double {
    const int nc = A.nc;
    for (int k = threadIdx.x; k < nc; k += PCG_THREADS) {
      const int g = k / 7, m = k % 7;
      // CTAs of group g: those whose first row lies in [grp_row_ptr[g], grp_row_ptr[g+1])
      const int c0 = (A.grp_row_ptr[g] + A.rows_per_cta - 1) / A.rows_per_cta;
      const int c1 = min((int)gridDim.x, (A.grp_row_ptr[g + 1] + A.rows_per_cta - 1) / A.rows_per_cta);
      double t = 0.0, t1 = 0.0;  // fixed order, two loads in flight
      int cta = c0;
      for (; cta + 1 < c1; cta += 2) {
        t += __ldcg(A.tpart + (size_t)cta * 8 + m);
        t1 += __ldcg(A.tpart + (size_t)(cta + 1) * 8 + m);
      }
      if (cta < c1) t += __ldcg(A.tpart + (size_t)cta * 8 + m);
      s_t[k] = t + t1;
    }
    __syncthreads();
    // e_g = rows of A^-1 of the OWN group (shared memory) times t: warp m forms e_g[m], lanes split k
    if (warp < 7) {
      double acc = 0.0;
      for (int k = lane; k < nc; k += 32) acc += s_ainv[warp * nc + k] * s_t[k];
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) acc += __shfl_xor_sync(FULL, acc, off);
      if (lane == 0) s_e[my_group * 7 + warp] = acc;
    }
    __syncthreads();
    // this group's share of t'e, counted once (by the group's first CTA); the sum rides on the next exchange
    double te = 0.0;
    if (threadIdx.x == 0 && row0 == A.grp_row_ptr[my_group]) {
#pragma unroll
      for (int m = 0; m < 7; ++m) te += s_t[my_group * 7 + m] * s_e[my_group * 7 + m];
    }
    return te;
  };
  // inside the loop: t -= alpha Z^T q from the partials published before the p'q exchange (all groups, on every CTA)
  auto coarse_recur_t = [&](double alpha) {
    const int nc = A.nc;
    for (int k = threadIdx.x; k < nc; k += PCG_THREADS) {
      const int g = k / 7, m = k % 7;
      const int c0 = (A.grp_row_ptr[g] + A.rows_per_cta - 1) / A.rows_per_cta;
      const int c1 = min((int)gridDim.x, (A.grp_row_ptr[g + 1] + A.rows_per_cta - 1) / A.rows_per_cta);
      double t = 0.0, t1 = 0.0;
      int cta = c0;
      for (; cta + 1 < c1; cta += 2) {
        t += __ldcg(A.tpart + (size_t)cta * 8 + m);
        t1 += __ldcg(A.tpart + (size_t)(cta + 1) * 8 + m);
      }
      if (cta < c1) t += __ldcg(A.tpart + (size_t)cta * 8 + m);
      s_t[k] -= alpha * (t + t1);
    }
  };
  // e of the own group from the current t; returns this CTA's share of t'e (thread 0 of the group's first CTA)
  auto coarse_e = [&]() -> double {
    const int nc = A.nc;
    __syncthreads();  // t is complete
    if (warp < 7) {
      double acc = 0.0;
      for (int k = lane; k < nc; k += 32) acc += s_ainv[warp * nc + k] * s_t[k];
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) acc += __shfl_xor_sync(FULL, acc, off);
      if (lane == 0) s_e[my_group * 7 + warp] = acc;
    }
    __syncthreads();
    double te = 0.0;
    if (threadIdx.x == 0 && row0 == A.grp_row_ptr[my_group]) {
#pragma unroll
      for (int m = 0; m < 7; ++m) te += s_t[my_group * 7 + m] * s_e[my_group * 7 + m];
    }
    return te;
  };
  // z_i = zj_i + Z_i e_g for the own rows, published for the other CTAs' products
  auto coarse_apply_z = [&]() {
    for (int lr = warp; lr < nrows; lr += WPB) {
      const int row = row0 + lr;
      if (lane < 6) {
        const double* Zi = A.Zc + (size_t)row * 42 + lane * 7;
        double zz = s_zj[lr * 6 + lane];
#pragma unroll
        for (int m = 0; m < 7; ++m) zz += __ldg(Zi + m) * s_e[my_group * 7 + m];
        __stcg(A.z + (size_t)row * 6 + lane, zz);
      }
    }
  };

  // x = 0, r = b, z = M^-1 r, p_old = 0
  double part = 0.0, tacc0 = 0.0;
  for (int lr = warp; lr < nrows; lr += WPB) {
    const int row = row0 + lr;
    const double rr = lane < 6 ? A.rhs[(size_t)row * 6 + lane] : 0.0;
    double zz = 0.0;
#pragma unroll
    for (int c = 0; c < 6; ++c) {
      const double rc = __shfl_sync(FULL, rr, c);
      if (lane < 6) zz += s_minv[lr * 36 + lane * 6 + c] * rc;
    }
    if (lane < 6) {
      s_x[lr * 6 + lane] = 0.0;
      s_r[lr * 6 + lane] = rr;
      if (coarse) s_zj[lr * 6 + lane] = zz; else __stcg(A.z + (size_t)row * 6 + lane, zz);
      __stcg(A.p0 + (size_t)row * 6 + lane, 0.0);
      part += rr * zz;
    }
    if (coarse) {
      double a = 0.0;
#pragma unroll
      for (int r = 0; r < 6; ++r) {
        const double rc = __shfl_sync(FULL, rr, r);
        if (lane < 7) a += __ldg(A.Zc + (size_t)row * 42 + r * 7 + lane) * rc;
      }
      tacc0 += a;
    }
  }
  if (coarse) {
    const double tt[1] = {tacc0};
    coarse_publish_t(tt);
  }
  double rz0 = grid_reduce(part, part1, A.bar, gen, sm);
  if (coarse) {
    const double te0 = coarse_solve();
    coarse_apply_z();
    rz0 += grid_reduce(te0, part0, A.bar, gen, sm);  // z of every CTA is visible; t'e summed over the groups
  }
  double rz = rz0, beta = 0.0;
  int iters = 0, flag = 1;
  double res = 1.0;
  if (!(rz0 > 0.0) || !isfinite(rz0)) {
    flag = (rz0 == 0.0) ? 0 : 2;
    res = 0.0;
  } else {
    double* pold = A.p0;
    double* pnew = A.p1;
    for (int it = 0; it < A.max_iter; ++it) {
      const long long c0 = clock64();
      // p = z + beta p_old (rebuilt where gathered); q = S p; p'q
      part = 0.0;
      double tqacc = 0.0;  // lanes 0..6: Z^T q over the warp's rows
      if (halo) {
        double2* ph2 = reinterpret_cast<double2*>(s_ph);
        for (int u = threadIdx.x; u < nuniq * 3; u += PCG_THREADS) {
          const int j = s_ucol[u / 3], h = u - 3 * (u / 3);
          const double2 zz = __ldcg(reinterpret_cast<const double2*>(A.z + (size_t)j * 6) + h);
          const double2 pp = __ldcg(reinterpret_cast<const double2*>(pold + (size_t)j * 6) + h);
          ph2[u] = make_double2(zz.x + beta * pp.x, zz.y + beta * pp.y);
        }
        __syncthreads();
      }
      for (int lr = warp; lr < nrows; lr += WPB) {
        const int row = row0 + lr;
        const int b0 = s_rowptr[lr], b1 = s_rowptr[lr + 1];
        double acc = 0.0;
        if (lane < 30) {
          if (halo) {
#pragma unroll 4
            for (int blk = b0 + kb0; blk < b1; blk += 5) {
              const double* pc = s_ph + s_bslot[blk] * 6;
              const double* Sb = Sbase + (size_t)blk * 36 + r6 * 6;
#pragma unroll
              for (int c = 0; c < 6; ++c) acc += Sb[c] * pc[c];
            }
          } else {
#pragma unroll 4
            for (int blk = b0 + kb0; blk < b1; blk += 5) {
              const int col = colbase[blk];
              double zc[6], pc[6];
              ldcg6(A.z + (size_t)col * 6, zc);
              ldcg6(pold + (size_t)col * 6, pc);
              const double* Sb = Sbase + (size_t)blk * 36 + r6 * 6;
#pragma unroll
              for (int c = 0; c < 6; ++c) acc += Sb[c] * (zc[c] + beta * pc[c]);
            }
          }
        }
        const double a1 = __shfl_sync(FULL, acc, r6 + 6), a2 = __shfl_sync(FULL, acc, r6 + 12),
                     a3 = __shfl_sync(FULL, acc, r6 + 18), a4 = __shfl_sync(FULL, acc, r6 + 24);
        const double qv = acc + a1 + a2 + a3 + a4;  // meaningful in lanes 0..5
        if (coarse) {
          double a = 0.0;
#pragma unroll
          for (int r = 0; r < 6; ++r) {
            const double qc = __shfl_sync(FULL, qv, r);
            if (lane < 7) a += __ldg(A.Zc + (size_t)row * 42 + r * 7 + lane) * qc;
          }
          tqacc += a;
        }
        if (lane < 6) {
          const size_t o = (size_t)row * 6 + lane;
          double pi;
          if (halo)
            pi = s_ph[(s_pref[row >> 5] + __popc(s_bits[row >> 5] & ((1u << (row & 31)) - 1u))) * 6 + lane];
          else
            pi = __ldcg(A.z + o) + beta * __ldcg(pold + o);
          __stcg(pnew + o, pi);
          s_p[lr * 6 + lane] = pi;
          s_q[lr * 6 + lane] = qv;
          part += pi * qv;
        }
      }
      if (coarse) {
        const double tt[1] = {tqacc};
        coarse_publish_t(tt);
      }
      const long long c1 = clock64();
      const double pq = grid_reduce(part, part0, A.bar, gen, sm);
      const long long c2 = clock64();
      if (!(pq > 0.0) || !isfinite(pq)) {
        flag = 2;
        break;
      }
      const double alpha = rz / pq;
      if (coarse) coarse_recur_t(alpha);  // L2 loads in flight while the rows are updated
      // x += alpha p, r -= alpha q, z = M^-1 r
      part = 0.0;
      for (int lr = warp; lr < nrows; lr += WPB) {
        const int row = row0 + lr;
        double rr = 0.0;
        if (lane < 6) {
          s_x[lr * 6 + lane] += alpha * s_p[lr * 6 + lane];
          rr = s_r[lr * 6 + lane] - alpha * s_q[lr * 6 + lane];
          s_r[lr * 6 + lane] = rr;
        }
        double zz = 0.0;
#pragma unroll
        for (int c = 0; c < 6; ++c) {
          const double rc = __shfl_sync(FULL, rr, c);
          if (lane < 6) zz += s_minv[lr * 36 + lane * 6 + c] * rc;
        }
        if (lane < 6) {
          if (coarse) s_zj[lr * 6 + lane] = zz; else __stcg(A.z + (size_t)row * 6 + lane, zz);
          part += rr * zz;
        }
      }
      if (coarse) {  // r'z = r'M^-1 r + t'e; z = M^-1 r + Z e is published before the one exchange of this half
        part += coarse_e();
        coarse_apply_z();
      }
      const long long c3 = clock64();
      const double rz_new = grid_reduce(part, part1, A.bar, gen, sm);
      const long long c4 = clock64();
      t_spmv += c1 - c0;
      t_upd += c3 - c2;
      t_red += (c2 - c1) + (c4 - c3);
      iters = it + 1;
      res = sqrt(fabs(rz_new) / rz0);
      if (!isfinite(rz_new)) {
        flag = 2;
        break;
      }
      if (res <= A.tol) {
        flag = 0;
        break;
      }
      beta = rz_new / rz;
      rz = rz_new;
      double* t = pold;
      pold = pnew;
      pnew = t;
    }
  }
  for (int lr = warp; lr < nrows; lr += WPB)
    if (lane < 6) A.x[(size_t)(row0 + lr) * 6 + lane] = s_x[lr * 6 + lane];
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    A.scal[S_PCG_ITERS] = (double)iters;
    A.scal[S_PCG_FLAG] = (double)flag;
    A.scal[S_PCG_RES] = res;
    A.scal[S_PCG_CYC_SPMV] = (double)t_spmv;
    A.scal[S_PCG_CYC_REDUCE] = (double)t_red;
    A.scal[S_PCG_CYC_UPDATE] = (double)t_upd;
  }
}

// ================================================================================================
// step
// ================================================================================================
__global__ void __launch_bounds__(1024) k_cam_update(ProblemView pv, const double* __restrict__ pose,
                                                    const double* __restrict__ y, const double* __restrict__ Hc,
                                                    const double* __restrict__ gc, const double* __restrict__ c_img,
                                                    LmDiag lm, double* __restrict__ pose_cand,
                                                    double* __restrict__ scal) {
  __shared__ double sm[33];
  lm_resolve(lm);
  cam_update_body(pv, pose, y, Hc, gc, c_img, lm, pose_cand, scal, sm);
}

// The same step on CTAs of 128 images each (1 000 images by ONE CTA took 21 us): slice sums into stage[slice][3], added
// in slice order by k_pt_scalars_step.  (Fusing these CTAs into k_back_substitute's launch was measured: its registers
// went 56 -> 72 and the launch lost 27 us.)
__global__ void __launch_bounds__(PW_THREADS) k_cam_update_slices(ProblemView pv, LmDiag lm, const double* __restrict__ y,
                                                                  CamStepArgs cam) {
  __shared__ double sm[PW_THREADS / 32 + 1];
  lm_resolve(lm);
  const int k = blockIdx.x;
  cam_update_range<PW_THREADS>(pv, cam.pose, y, cam.Hc, cam.gc, cam.c_img, lm, cam.pose_cand, k * PW_THREADS,
                               min(pv.C, (k + 1) * PW_THREADS), cam.stage + 3 * k, sm);
}

// Back-substitution from the stored Schur factors, one thread per observation (same warp packing as k_points).
// With Hpp + D = L L^T, Z_a = N_a R_a L^-T, e = L^-1 gp and g_a = G_a y_c(a) (p_rot is in the record):
//     y_p = (Hpp + D)^-1 (gp - sum_a W_a^T y_c(a)) = L^-T (e - sum_a Z_a^T g_a),      X_cand = X - y_p.
// No re-linearisation: the pass reads one 128-byte record per observation (warp-cooperative gather) and 48 bytes of
// y_c.  The model cost change uses the normal equations the step solves, (H + D) y = g:
//     g'y - y'Hy/2 = (g'y + y'Dy) / 2
// (exact up to the PCG residual, 1e-12 relative by default); the point part is summed here, the image part in
// k_cam_update.
__device__ __forceinline__ void point_step(int p, bool active, double2 xa, double2 xb, const double zg[3],
                                           const double* __restrict__ Linv, const double* __restrict__ Hpp,
                                           const double* __restrict__ gp, const double* __restrict__ c_pt, LmDiag lm,
                                           double* __restrict__ X_cand, double& mcc, double& step2, double& xn2) {
  const double Xp[3] = {xa.x, xa.y, xb.x};
  double yp[3] = {0.0, 0.0, 0.0};
  if (active) {
    double Li[6];
#pragma unroll
    for (int j = 0; j < 6; ++j) Li[j] = Linv[(size_t)p * LINV_STRIDE + j];
    const double u[3] = {Linv[(size_t)p * LINV_STRIDE + 6] - zg[0], Linv[(size_t)p * LINV_STRIDE + 7] - zg[1],
                         Linv[(size_t)p * LINV_STRIDE + 8] - zg[2]};
    // y_p = L^-T u,  u = e - sum Z^T g  (Li = L^-1, lower: 00, 10, 11, 20, 21, 22)
    yp[0] = Li[0] * u[0] + Li[1] * u[1] + Li[3] * u[2];
    yp[1] = Li[2] * u[1] + Li[4] * u[2];
    yp[2] = Li[5] * u[2];
    const double h0 = Hpp[(size_t)p * 6], h3 = Hpp[(size_t)p * 6 + 3], h5 = Hpp[(size_t)p * 6 + 5];
    const double d0 = lm_damping_fast(h0, c_pt[(size_t)p * 3], lm.radius, lm.inv_radius, lm.lo, lm.hi),
                 d1 = lm_damping_fast(h3, c_pt[(size_t)p * 3 + 1], lm.radius, lm.inv_radius, lm.lo, lm.hi),
                 d2 = lm_damping_fast(h5, c_pt[(size_t)p * 3 + 2], lm.radius, lm.inv_radius, lm.lo, lm.hi);
    mcc = 0.5 * (gp[(size_t)p * 3] * yp[0] + gp[(size_t)p * 3 + 1] * yp[1] + gp[(size_t)p * 3 + 2] * yp[2] +
                 d0 * yp[0] * yp[0] + d1 * yp[1] * yp[1] + d2 * yp[2] * yp[2]);
    step2 = yp[0] * yp[0] + yp[1] * yp[1] + yp[2] * yp[2];
    xn2 = Xp[0] * Xp[0] + Xp[1] * Xp[1] + Xp[2] * Xp[2];
  }
  double2* out = reinterpret_cast<double2*>(X_cand + 4 * (size_t)p);
  out[0] = make_double2(Xp[0] - yp[0], Xp[1] - yp[1]);
  out[1] = make_double2(Xp[2] - yp[2], xb.y);  // slot 3: the point's constant flag travels with it
}

// s = Z^T (G y_c) of one observation from its record (Z 9, p_rot 3) and the image's y (6)
__device__ __forceinline__ void obs_zt_g(const double Z[9], const double prot[3], const double* __restrict__ y, int img,
                                         double s[3]) {
  const double2* yi = reinterpret_cast<const double2*>(y + (size_t)img * 6);
  const double2 a = __ldg(yi), b = __ldg(yi + 1), c = __ldg(yi + 2);
  const double yc[6] = {a.x, a.y, b.x, b.y, c.x, c.y};
  double ga[3];
  g_vec(prot, yc, ga);
  s[0] = Z[0] * ga[0] + Z[3] * ga[1] + Z[6] * ga[2];
  s[1] = Z[1] * ga[0] + Z[4] * ga[1] + Z[7] * ga[2];
  s[2] = Z[2] * ga[0] + Z[5] * ga[1] + Z[8] * ga[2];
}

__global__ void __launch_bounds__(PW_THREADS) k_back_substitute(ProblemView pv, const double* __restrict__ X,
                                                                const double* __restrict__ y,
                                                                const double* __restrict__ Hpp,
                                                                const double* __restrict__ gp,
                                                                const double* __restrict__ Linv,
                                                                const double* __restrict__ c_pt,
                                                                const double* __restrict__ fac, int use_factors,
                                                                LmDiag lm, double* __restrict__ X_cand,
                                                                double* __restrict__ pt_part, int prefetch_ahead) {
  __shared__ __align__(16) double2 rows_all[PW_WARPS][256];
  lm_resolve(lm);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int bid = (int)blockIdx.x;
  const int w = bid * PW_WARPS + warp;
  double mcc = 0.0, step2 = 0.0, xn2 = 0.0;
  if (w < pv.num_track_warps) {
    const int4 desc = __ldg(pv.warp_desc + w);  // {first track, tracks, first slot, observations}
    // L2 prefetch (see prefetch_track_cta) of the group this warp position handles prefetch_ahead CTAs later
    const int wf = w + prefetch_ahead * PW_WARPS;
    if (prefetch_ahead > 0 && wf < pv.num_track_warps) {
      const int4 fd = __ldg(pv.warp_desc + wf);
      const size_t T0 = (size_t)fd.x, nt = (size_t)fd.y, fb = (size_t)fd.z, fn = (size_t)min(fd.w, 64);
#pragma unroll
      for (int r = 0; r < 3; ++r) {
        int slot = 0;
        const int t = lane + 32 * r;
        if (use_factors) l2_prefetch_range(fac + FAC_STRIDE * fb, fn * FAC_STRIDE * 8, slot, t);
        l2_prefetch_range(X + 4 * T0, nt * 32, slot, t);
        l2_prefetch_range(Linv + LINV_STRIDE * T0, nt * LINV_STRIDE * 8, slot, t);
        l2_prefetch_range(pv.pm_pt + fb, fn * 4, slot, t);
        l2_prefetch_range(pv.pm_img + fb, fn * 4, slot, t);
      }
    }
    const double2* fac2 = reinterpret_cast<const double2*>(fac);
    if (desc.w <= 32) {
      // Lane = observation slot.  Tracks are runs of equal point index: heads by comparing with the left lane;
      // the head lane of a track also does the track's point step.  (Points without observations are never
      // touched: both point buffers hold their coordinates from the start of the solve.)
      const int base = desc.z, n = desc.w;
      const bool has = lane < n;
      const int pt = has ? __ldg(pv.pm_pt + base + lane) : -1;
      const int img = has ? __ldg(pv.pm_img + base + lane) : 0;
      const int left = __shfl_up_sync(FULL, pt, 1);
      const bool head = has && (lane == 0 || pt != left);
      const unsigned heads = __ballot_sync(FULL, head);
      double2 xa = make_double2(0.0, 0.0), xb = make_double2(0.0, 1.0);
      if (head) {
        const double2* xp = reinterpret_cast<const double2*>(X + 4 * (size_t)pt);
        xa = __ldg(xp);
        xb = __ldg(xp + 1);
      }
      const unsigned le = heads & (0xffffffffu >> (31 - lane));
      const int seg_start = le ? 31 - __clz(le) : 0;
      const unsigned above = (lane == 31) ? 0u : (heads & (0xffffffffu << (lane + 1)));
      const int seg_end = above ? (__ffs(above) - 1) : n;
      const bool free_pt = __shfl_sync(FULL, xb.y == 0.0 ? 1 : 0, seg_start) != 0;  // slot 3: constant flag
      const bool act = has && free_pt && use_factors;  // no factors: all poses constant, y_c = 0
      double2* rows = rows_all[warp];
      warp_gather_records(fac2, act ? base + lane : -1, 6, rows, lane);
      cp_async_commit();
      cp_async_wait<0>();
      __syncwarp();
      double v[3] = {0.0, 0.0, 0.0};  // Z^T g of this observation
      if (act) {
        const double2* src = rows + lane * 8;
        const int sw = lane & 7;
        const double2 c0 = src[0 ^ sw], c1 = src[1 ^ sw], c2 = src[2 ^ sw], c3 = src[3 ^ sw], c4 = src[4 ^ sw],
                      c5 = src[5 ^ sw];
        const double Z[9] = {c0.x, c0.y, c1.x, c1.y, c2.x, c2.y, c3.x, c3.y, c4.x};
        const double prot[3] = {c4.y, c5.x, c5.y};
        obs_zt_g(Z, prot, y, img, v);
      }
      segmented_reduce<3>(v, lane, seg_end);
      if (head) point_step(pt, free_pt, xa, xb, v, Linv, Hpp, gp, c_pt, lm, X_cand, mcc, step2, xn2);
    } else {
      WarpTracks wt;
      warp_tracks_setup(pv, w, lane, wt, reinterpret_cast<int*>(rows_all[warp]));
      const int p = wt.t0;
      const bool active = pv.pt_const[p] == 0;
      double s3[3] = {0.0, 0.0, 0.0};
      if (active && use_factors) {
        for (int k = lane; k < wt.n; k += 32) {
          const double2* src = fac2 + (size_t)(wt.base + k) * 8;
          const double2 c0 = src[0], c1 = src[1], c2 = src[2], c3 = src[3], c4 = src[4], c5 = src[5];
          const double Z[9] = {c0.x, c0.y, c1.x, c1.y, c2.x, c2.y, c3.x, c3.y, c4.x};
          const double prot[3] = {c4.y, c5.x, c5.y};
          double t[3];
          obs_zt_g(Z, prot, y, __ldg(pv.pm_img + wt.base + k), t);
          s3[0] += t[0]; s3[1] += t[1]; s3[2] += t[2];
        }
      }
      warp_allreduce<3>(s3);
      if (lane == 0) {
        const double2* xp = reinterpret_cast<const double2*>(X + 4 * (size_t)p);
        point_step(p, active, __ldg(xp), __ldg(xp + 1), s3, Linv, Hpp, gp, c_pt, lm, X_cand, mcc, step2, xn2);
      }
    }
  }
  __shared__ double sm[PW_WARPS + 1];
  mcc = block_sum<PW_WARPS>(mcc, sm);
  step2 = block_sum<PW_WARPS>(step2, sm);
  xn2 = block_sum<PW_WARPS>(xn2, sm);
  if (threadIdx.x == 0) {
    pt_part[(size_t)bid * PT_PART + 1] = mcc;
    pt_part[(size_t)bid * PT_PART + 2] = step2;
    pt_part[(size_t)bid * PT_PART + 3] = xn2;
  }
}

// Sums (and max |gp|) of the per-CTA partials of the track-order passes.  One CTA walking all rows took 31 us at cfg3
// (39 k rows): now up to 64 CTAs reduce contiguous slices (fixed order inside a slice) into a stage row each, and the
// CTA that arrives last (ticket) adds the stage rows in slice order -- deterministic, one launch.
constexpr int PTS_THREADS = 256, PTS_MAX_CTAS = 64;
__global__ void __launch_bounds__(PTS_THREADS) k_pt_scalars_step(const double* __restrict__ pt_part, int nblk,
                                                                 double* __restrict__ scal, int with_gmax,
                                                                 double* stage /* [gridDim.x][4] */, unsigned* ticket,
                                                                 const double* cam_stage, int cam_rows) {
  __shared__ double sm[PTS_THREADS / 32 + 1];
  __shared__ unsigned s_last;
  const int G = gridDim.x;
  const int per = (nblk + G - 1) / G;
  const int r0 = blockIdx.x * per, r1 = min(nblk, r0 + per);
  double a = 0.0, b = 0.0, c = 0.0, gm = 0.0;
  for (int i = r0 + threadIdx.x; i < r1; i += PTS_THREADS) {
    const double4 v = *reinterpret_cast<const double4*>(pt_part + (size_t)i * PT_PART);
    if (with_gmax) gm = fmax(gm, v.x);
    a += v.y; b += v.z; c += v.w;
  }
  a = block_sum<PTS_THREADS / 32>(a, sm);
  b = block_sum<PTS_THREADS / 32>(b, sm);
  c = block_sum<PTS_THREADS / 32>(c, sm);
  gm = block_max<PTS_THREADS / 32>(gm, sm);
  if (threadIdx.x == 0) {
    double* row = stage + 4 * blockIdx.x;
    __stcg(row, gm); __stcg(row + 1, a); __stcg(row + 2, b); __stcg(row + 3, c);
    __threadfence();
    s_last = atomicAdd(ticket, 1u) == (unsigned)(G - 1) ? 1u : 0u;
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  if (threadIdx.x < 4) {
    double acc = 0.0;
    for (int g = 0; g < G; ++g) {
      const double v = __ldcg(stage + 4 * g + threadIdx.x);
      acc = threadIdx.x == 0 ? fmax(acc, v) : acc + v;
    }
    if (threadIdx.x == 0) {
      if (with_gmax) scal[S_GMAX_P] = acc;
      *ticket = 0u;  // ready for the next launch
    } else {
      scal[threadIdx.x == 1 ? S_MCC : (threadIdx.x == 2 ? S_STEP2_P : S_XNORM2_P)] = acc;
    }
  } else if (threadIdx.x >= 32 && threadIdx.x < 35 && cam_rows > 0) {
    // camera step slices written by the extra CTAs of k_back_substitute (complete: that kernel ran before this one)
    const int q = threadIdx.x - 32;
    double acc = 0.0;
    for (int g = 0; g < cam_rows; ++g) acc += cam_stage[3 * g + q];
    scal[q == 0 ? S_STEP2_C : (q == 1 ? S_XNORM2_C : S_MCC_C)] = acc;
  }
}

// ================================================================================================
// misc
// ================================================================================================
__global__ void __launch_bounds__(PT_THREADS) k_evaluate(ProblemView pv, const double* __restrict__ pose,
                                                         const double* __restrict__ X, double* __restrict__ r_reproj,
                                                         double* __restrict__ r_depth) {
  const int p = pv.p0 + blockIdx.x * PT_THREADS + threadIdx.x;
  if (p >= pv.p1) return;
  double Xp[3];
  load_point(X, p, Xp);
  const size_t N = (size_t)pv.N;
  for (int k = pv.pt_ptr[p]; k < pv.pt_ptr[p + 1]; ++k) {
    const int img = __ldg(pv.pm_img + k);
    ImgConst ic;
    load_img(pv, pose, img, ic);
    const double2 o2 = __ldg(reinterpret_cast<const double2*>(pv.pm_xy) + k);
    const double d = __ldg(pv.pm_depth + k), w = __ldg(pv.pm_depth + N + k), a = __ldg(pv.pm_depth + 2 * N + k);
    ObsLin o;
    linearize_obs<false, false>(ic.pose, ic.R, Xp, ic.model, ic.prm, o2.x, o2.y, d, w, a, ic.b, ic.es, pv.loss, o);
    if (r_reproj) reinterpret_cast<double2*>(r_reproj)[k] = make_double2(o.r[0], o.r[1]);
    if (r_depth) r_depth[k] = o.r[2];
  }
}

// Post-BA track filter (ba_math.cuh: filter_track), one thread per track like k_evaluate: the pass runs once per
// bundle, not per LM iteration.  Unselected points are reported as kept and not evaluated (error -1).  The count of
// removed observations is an integer sum, so the double atomicAdd is exact and order independent.
__global__ void __launch_bounds__(PT_THREADS) k_filter_tracks(ProblemView pv, const double* __restrict__ pose,
                                                              const double* __restrict__ X, double max_err2,
                                                              double min_angle, const uint8_t* __restrict__ pt_select,
                                                              uint8_t* obs_keep, uint8_t* __restrict__ pt_keep,
                                                              double* __restrict__ pt_error,
                                                              double* __restrict__ scal) {
  const int p = pv.p0 + blockIdx.x * PT_THREADS + threadIdx.x;
  int filtered = 0;
  if (p < pv.p1) {
    const int k0 = pv.pt_ptr[p], k1 = pv.pt_ptr[p + 1];
    bool alive = true;
    double err = -1.0;
    if (pt_select == nullptr || pt_select[p] != 0) {
      double Xp[3];
      load_point(X, p, Xp);
      alive = filter_track(pose, pv.img_camera, pv.cam_model, pv.cam_params, MPBA_MAX_PRM, pv.pm_img, pv.pm_xy, k0, k1,
                           Xp, max_err2, min_angle, obs_keep, err, filtered);
    } else {
      for (int k = k0; k < k1; ++k) obs_keep[k] = 1;
    }
    pt_keep[p] = alive ? 1 : 0;
    pt_error[p] = err;
  }
  __shared__ double sm[PT_THREADS / 32 + 1];
  const double total = block_sum<PT_THREADS / 32>((double)filtered, sm);
  if (threadIdx.x == 0 && total != 0.0) atomicAdd(scal + S_FILTERED, total);
}

// Observations whose point sits at z < epsilon in the camera frame (the reprojection blocks the COLMAP >= 3.11 rule
// zeroes, ba_math.cuh: linearize_obs): one thread per owned track like k_evaluate, once per solve.  Integer count, so
// the double atomicAdd is exact and order independent.
__global__ void __launch_bounds__(PT_THREADS) k_count_behind(ProblemView pv, const double* __restrict__ pose,
                                                             const double* __restrict__ X, double* __restrict__ out) {
  const int p = pv.p0 + blockIdx.x * PT_THREADS + threadIdx.x;
  int behind = 0;
  if (p < pv.p1) {
    double Xp[3];
    load_point(X, p, Xp);
    for (int k = pv.pt_ptr[p]; k < pv.pt_ptr[p + 1]; ++k) {
      const double* ps = pose + 8 * (size_t)__ldg(pv.pm_img + k);
      double R[9];
      quat_to_rotmat(ps, R);
      const double pz = R[6] * Xp[0] + R[7] * Xp[1] + R[8] * Xp[2] + ps[6];
      behind += pz < DBL_EPSILON ? 1 : 0;
    }
  }
  __shared__ double sm[PT_THREADS / 32 + 1];
  const double total = block_sum<PT_THREADS / 32>((double)behind, sm);
  if (threadIdx.x == 0 && total != 0.0) atomicAdd(out, total);
}

// COLMAP normalises every quaternion before it enters the problem
__global__ void k_normalize_poses(int C, const double* __restrict__ in, const double* __restrict__ shift_scale,
                                  double* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= C) return;
  const double* s = in + (size_t)i * 8;
  const double n = sqrt(s[0] * s[0] + s[1] * s[1] + s[2] * s[2] + s[3] * s[3]);
  double* o = out + (size_t)i * 8;
  o[0] = s[0] / n; o[1] = s[1] / n; o[2] = s[2] / n; o[3] = s[3] / n;
  o[4] = s[4]; o[5] = s[5]; o[6] = s[6];
  o[7] = exp(shift_scale[2 * (size_t)i + 1]);  // exp(log-scale) of the depth prior: constant block, one exp per image
}

// Once per call: the point of every (owned) track-order slot and the CTA descriptors of the track-order passes.
__global__ void k_track_setup(ProblemView pv, int32_t* __restrict__ pm_pt, int4* __restrict__ cta_desc,
                              int4* __restrict__ warp_desc) {
  const int gid = blockIdx.x * blockDim.x + threadIdx.x;
  const int p = pv.p0 + gid;
  if (p < pv.p1)
    for (int k = pv.pt_ptr[p]; k < pv.pt_ptr[p + 1]; ++k) pm_pt[k] = p;
  const int nb = (pv.num_track_warps + PW_WARPS - 1) / PW_WARPS;
  if (gid < nb) {
    const int g0 = gid * PW_WARPS, g1 = min(g0 + PW_WARPS, pv.num_track_warps);
    const int T0 = pv.tw_first[g0], T1 = pv.tw_first[g1];
    const int base = pv.pt_ptr[T0];
    cta_desc[gid] = make_int4(T0, T1 - T0, base, pv.pt_ptr[T1] - base);
  }
  if (gid < pv.num_track_warps) {
    const int T0 = pv.tw_first[gid], T1 = pv.tw_first[gid + 1];
    const int base = pv.pt_ptr[T0];
    warp_desc[gid] = make_int4(T0, T1 - T0, base, pv.pt_ptr[T1] - base);
  }
}

// slot 3 of every point record <- 1.0 for a constant point, 0.0 otherwise (read by k_img_chunks with the gather)
__global__ void k_flag_points(int P, const uint8_t* __restrict__ pt_const, double* __restrict__ X) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p < P) X[4 * (size_t)p + 3] = pt_const[p] ? 1.0 : 0.0;
}

__global__ void k_expand_hcc(int C, const double* __restrict__ Hc, double* __restrict__ H36) {
  const int gid = blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= C * 36) return;
  const int i = gid / 36, e = gid % 36, r = e / 6, c = e % 6;
  H36[gid] = Hc[(size_t)i * 21 + upper_idx(r < c ? r : c, r < c ? c : r)];
}

}  // namespace

// ================================================================================================
// launchers
// ================================================================================================
void launch_img_chunks(const ProblemView& pv, const double* pose, const double* X, double* chunk_part, cudaStream_t st) {
  // CTAs of lookahead of the L2 prefetch (0: off); the environment is read once per process
  static const int prefetch_ahead = getenv("MPBA_IMG_PREFETCH") ? atoi(getenv("MPBA_IMG_PREFETCH")) : MPBA_IMG_PREFETCH_DEFAULT;
  static const bool sacc = !(getenv("MPBA_IMG_SMEM_ACC") && atoi(getenv("MPBA_IMG_SMEM_ACC")) == 0);
  if (pv.num_chunks <= 0) return;
  if (pv.cm_plane && sacc)
    k_img_chunks<true, true><<<pv.num_chunks, IMG_THREADS, 0, st>>>(pv, pose, X, chunk_part, prefetch_ahead);
  else if (pv.cm_plane)
    k_img_chunks<true, false><<<pv.num_chunks, IMG_THREADS, 0, st>>>(pv, pose, X, chunk_part, prefetch_ahead);
  else if (sacc)
    k_img_chunks<false, true><<<pv.num_chunks, IMG_THREADS, 0, st>>>(pv, pose, X, chunk_part, prefetch_ahead);
  else
    k_img_chunks<false, false><<<pv.num_chunks, IMG_THREADS, 0, st>>>(pv, pose, X, chunk_part, prefetch_ahead);
}
void launch_img_finalize(const ProblemView& pv, const double* chunk_part, double* Hc, double* gc, double* img_part,
                         cudaStream_t st) {
  const int n = pv.C * 32;
  k_img_finalize<<<(n + 255) / 256, 256, 0, st>>>(pv, chunk_part, Hc, gc, img_part);
}
void launch_img_tail(const ProblemView& pv, const double* chunk_part, double* Hc, double* gc, double* img_part,
                     double* c_img, int compute_scaling, int jacobi_scaling, double* scal, const LmTail* tail,
                     const PtScalarsArgs* pts, cudaStream_t st) {
  LmTail t = {};
  if (tail) t = *tail;
  PtScalarsArgs ps = {nullptr, -1, 0, nullptr, 0};
  if (pts) ps = *pts;
  k_img_tail<<<1, 1024, 0, st>>>(pv, chunk_part, Hc, gc, img_part, c_img, compute_scaling, jacobi_scaling, scal, t,
                                 tail ? 1 : 0, ps);
}
void launch_img_scalars(const ProblemView& pv, const double* img_part, const double* Hc, const double* gc,
                        double* c_img, int compute_scaling, int jacobi_scaling, double* scal, cudaStream_t st) {
  k_img_scalars<<<1, 1024, 0, st>>>(pv.C, pv.img_free_mask, img_part, Hc, gc, c_img, compute_scaling, jacobi_scaling,
                                    scal);
}
int points_grid(int P) { return (P + PT_THREADS - 1) / PT_THREADS; }
int track_warp_grid(int num_track_warps) { return (num_track_warps + PW_WARPS - 1) / PW_WARPS; }
void launch_track_setup(const ProblemView& pv, int32_t* pm_pt, int4* cta_desc, int4* warp_desc, cudaStream_t st) {
  const int n = max(pv.p1 - pv.p0, pv.num_track_warps);
  if (n > 0) k_track_setup<<<(n + 255) / 256, 256, 0, st>>>(pv, pm_pt, cta_desc, warp_desc);
}
void launch_points(const ProblemView& pv, const double* pose, const double* X, LmDiag lm, double* c_pt,
                   int compute_scaling, int jacobi_scaling, double* Hpp, double* gp, double* Hinv, double* Linv,
                   double* fac, double* pt_part, double* scal, int write_factors, cudaStream_t st) {
  static const int force_warp = getenv("MPBA_POINTS_WARP") ? atoi(getenv("MPBA_POINTS_WARP")) : 0;
  static const int prefetch_ahead = getenv("MPBA_POINTS_PREFETCH") ? atoi(getenv("MPBA_POINTS_PREFETCH")) : MPBA_POINTS_PREFETCH_DEFAULT;
  // MPBA_POINTS_BULK=1: the cp.async.bulk / mbarrier pipelined variant (k_points_bulk), persistent CTAs
  static const int bulk = getenv("MPBA_POINTS_BULK") ? atoi(getenv("MPBA_POINTS_BULK")) : MPBA_POINTS_BULK_DEFAULT;
  static int bulk_grid = 0;
  if (bulk && !bulk_grid) {
    int dev = 0, sms = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaFuncSetAttribute(k_points_bulk, cudaFuncAttributeMaxDynamicSharedMemorySize, PB_SMEM);
    bulk_grid = sms * MPBA_POINTS_MINB;
  }
  const int groups = track_warp_grid(pv.num_track_warps);
  if (pv.num_track_warps > 0 && bulk && !force_warp && !pv.pm_plane)
    k_points_bulk<<<groups < bulk_grid ? groups : bulk_grid, PW_THREADS, PB_SMEM, st>>>(
        pv, pose, X, lm, c_pt, compute_scaling, jacobi_scaling, Hpp, gp, Hinv, Linv, fac, pt_part, scal, write_factors,
        groups);
  else if (pv.num_track_warps > 0 && pv.pm_plane)
    k_points<true><<<groups, PW_THREADS, 0, st>>>(pv, pose, X, lm, c_pt, compute_scaling, jacobi_scaling, Hpp, gp, Hinv,
                                                  Linv, fac, pt_part, scal, write_factors, force_warp, prefetch_ahead);
  else if (pv.num_track_warps > 0)
    k_points<false><<<groups, PW_THREADS, 0, st>>>(pv, pose, X, lm, c_pt, compute_scaling, jacobi_scaling, Hpp, gp, Hinv,
                                                   Linv, fac, pt_part, scal, write_factors, force_warp, prefetch_ahead);
}
void launch_pt_scalars_gmax(const double* pt_part, int nblk, double* scal, cudaStream_t st) {
  k_pt_scalars_gmax<<<1, 1024, 0, st>>>(pt_part, nblk, scal);
}
void launch_schur_blocks(const ProblemView& pv, const double* fac, const double* Hc, const double* gc,
                         const double* c_img, LmDiag lm, double* S, double* rhs, int partial, double* Minv,
                         cudaStream_t st) {
  static bool attr_set = false;
  static const bool bulk = getenv("MPBA_SCHUR_BULK") && atoi(getenv("MPBA_SCHUR_BULK")) != 0;
  if (!attr_set) {
    cudaFuncSetAttribute(k_schur_blocks<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SB_STAGE_BYTES);
    cudaFuncSetAttribute(k_schur_blocks<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SB_STAGE_BYTES);
    attr_set = true;
  }
  if (bulk)
    k_schur_blocks<true><<<pv.U, SB_THREADS, SB_STAGE_BYTES, st>>>(pv, fac, Hc, gc, c_img, lm, S, rhs, partial, Minv);
  else
    k_schur_blocks<false><<<pv.U, SB_THREADS, SB_STAGE_BYTES, st>>>(pv, fac, Hc, gc, c_img, lm, S, rhs, partial, Minv);
}
void launch_schur_finish(const ProblemView& pv, const double* Hc, const double* gc, const double* c_img, LmDiag lm,
                         double* S, double* rhs, double* Minv, cudaStream_t st) {
  k_schur_finish<<<pv.C, 64, 0, st>>>(pv, Hc, gc, c_img, lm, S, rhs, Minv);
}
namespace {
__global__ void k_stage_scalars(const double* __restrict__ scal, double* __restrict__ rows, int rank, int world) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= world * S_COUNT) return;
  const int r = k / S_COUNT, j = k - r * S_COUNT;
  double v = 0.0;
  if (r == rank) v = scal[j];
  rows[k] = v;
}
}  // namespace
void launch_stage_scalars(const double* scal, double* rows, int rank, int world, cudaStream_t st) {
  const int n = world * S_COUNT;
  k_stage_scalars<<<(n + 255) / 256, 256, 0, st>>>(scal, rows, rank, world);
}
void launch_combine_scalars(const double* scal_all, int world, double* scal, cudaStream_t st) {
  k_combine_scalars<<<1, 32, 0, st>>>(scal_all, world, scal);
}
namespace {
__global__ void __launch_bounds__(1024) k_plan_hash(const int32_t* __restrict__ region, size_t n, unsigned long long* out) {
  __shared__ unsigned long long sm[32];
  unsigned long long acc = 0;
  for (size_t i = threadIdx.x; i < n; i += 1024) acc += plan_hash_term(__ldcg(region + i), i);
  for (int off = 16; off > 0; off >>= 1) acc += __shfl_down_sync(FULL, acc, off);
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    acc = sm[threadIdx.x];
    for (int off = 16; off > 0; off >>= 1) acc += __shfl_down_sync(FULL, acc, off);
    if (threadIdx.x == 0) *out = acc;
  }
}
}  // namespace
void launch_plan_hash(const int32_t* region, size_t n, unsigned long long* out, cudaStream_t st) {
  k_plan_hash<<<1, 1024, 0, st>>>(region, n, out);
}
PcgPlan pcg_plan(int device, int C, int max_row_blocks) {
  static int sms = 0, smem_optin = 0;
  if (!sms) {
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    cudaDeviceGetAttribute(&smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, device);
    cudaFuncSetAttribute(k_pcg<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_optin - 40960);
    cudaFuncSetAttribute(k_pcg<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_optin - 40960);
  }
  PcgPlan pl;
  int grid = (C + 7) / 8;          // at least eight block rows per CTA
  if (grid > sms) grid = sms;      // one CTA per SM: co-residency for the grid barrier
  if (grid < 1) grid = 1;
  pl.rows_per_cta = (C + grid - 1) / grid;
  pl.grid = (C + pl.rows_per_cta - 1) / pl.rows_per_cta;
  const size_t rows = (size_t)pl.rows_per_cta;
  const size_t base = rows * (36 + 24) * sizeof(double) + ((rows + 4) & ~(size_t)3) * sizeof(int);
  pl.max_local_blocks = (int)(rows * (size_t)(max_row_blocks > 0 ? max_row_blocks : 1 << 16));
  // halo staging of the product: distinct columns of a CTA's rows (capacity), the slot of every local block, a bitmap
  // over the images and its ranks; taken when it fits next to the row-local vectors, S in shared memory after that
  const size_t budget = (size_t)(smem_optin - 40960);
  pl.halo_cap = 0;
  pl.halo_words = 0;
  size_t halo_bytes = 0;
  if (max_row_blocks > 0 && C <= (1 << 17)) {
    const int cap = std::min(std::min(C, pl.max_local_blocks), 2048);
    const int words = (C + 31) / 32;
    halo_bytes = (size_t)cap * (6 * sizeof(double) + sizeof(int)) + (size_t)pl.max_local_blocks * sizeof(int) +
                 2 * (size_t)words * sizeof(int) + 16;
    if (base + halo_bytes + 16 <= budget) {
      pl.halo_cap = cap;
      pl.halo_words = words;
    } else {
      halo_bytes = 0;
    }
  }
  const size_t with_s = base + (size_t)pl.max_local_blocks * 36 * sizeof(double) +
                        (size_t)((pl.max_local_blocks + 3) & ~3) * sizeof(int) + 16;
  pl.use_smem = (max_row_blocks > 0 && with_s + halo_bytes <= budget) ? 1 : 0;
  pl.smem_bytes = (pl.use_smem ? with_s : base + 16) + halo_bytes;
  return pl;
}
cudaError_t launch_pcg(const ProblemView& pv, const PcgPlan& pl, const double* S, const double* Minv, const double* rhs,
                       double* y, double* vec_z, double* vec_p0, double* vec_p1, double* partial, unsigned* bar,
                       double* scal, double tol, int max_iter, const PcgGridCoarse& co, cudaStream_t st) {
  PcgArgs a;
  a.G = co.G; a.nc = 7 * co.G; a.grp_row_ptr = co.grp_row_ptr; a.Zc = co.Zc; a.Ainv = co.Ainv; a.tpart = co.tpart;
  a.C = pv.C; a.rows_per_cta = pl.rows_per_cta; a.max_local_blocks = pl.max_local_blocks;
  a.row_ptr = pv.s_row_ptr; a.col = pv.s_col; a.S = S; a.Minv = Minv; a.rhs = rhs; a.x = y;
  a.z = vec_z; a.p0 = vec_p0; a.p1 = vec_p1; a.partial = partial; a.bar = bar; a.scal = scal;
  a.tol = tol; a.max_iter = max_iter;
  a.halo_cap = pl.halo_cap; a.halo_words = pl.halo_words;
  cudaError_t e = cudaMemsetAsync(bar, 0, sizeof(unsigned), st);
  if (e != cudaSuccess) return e;
  void* args[] = {&a};
  if (pl.use_smem)
    return cudaLaunchCooperativeKernel((void*)k_pcg<true>, dim3(pl.grid), dim3(PCG_THREADS), args, pl.smem_bytes, st);
  return cudaLaunchCooperativeKernel((void*)k_pcg<false>, dim3(pl.grid), dim3(PCG_THREADS), args, pl.smem_bytes, st);
}
void launch_cam_update(const ProblemView& pv, const double* pose, const double* y, const double* Hc, const double* gc,
                       const double* c_img, LmDiag lm, double* pose_cand, double* scal, cudaStream_t st) {
  k_cam_update<<<1, 1024, 0, st>>>(pv, pose, y, Hc, gc, c_img, lm, pose_cand, scal);
}
void launch_back_substitute(const ProblemView& pv, const double* X, const double* y, const double* Hpp,
                            const double* gp, const double* Linv, const double* c_pt, const double* fac,
                            int use_factors, LmDiag lm, double* X_cand, double* pt_part, const CamStepArgs* cam,
                            cudaStream_t st) {
  static const int prefetch_ahead = getenv("MPBA_BACKSUB_PREFETCH") ? atoi(getenv("MPBA_BACKSUB_PREFETCH")) : MPBA_BACKSUB_PREFETCH_DEFAULT;
  if (cam) k_cam_update_slices<<<cam_step_rows(pv.C), PW_THREADS, 0, st>>>(pv, lm, y, *cam);
  const int groups = track_warp_grid(pv.num_track_warps);
  if (groups > 0)
    k_back_substitute<<<groups, PW_THREADS, 0, st>>>(pv, X, y, Hpp, gp, Linv, c_pt, fac, use_factors, lm, X_cand, pt_part,
                                                     prefetch_ahead);
}
int cam_step_rows(int num_images) { return (num_images + PW_THREADS - 1) / PW_THREADS; }
void launch_pt_scalars_step(const double* pt_part, int nblk, double* scal, int with_gmax, double* stage,
                            const double* cam_stage, int cam_rows, cudaStream_t st) {
  int grid = (nblk + 4 * PTS_THREADS - 1) / (4 * PTS_THREADS);  // at least four rows per thread
  grid = grid < 1 ? 1 : (grid > PTS_MAX_CTAS ? PTS_MAX_CTAS : grid);
  // stage: [PTS_MAX_CTAS][4] doubles, then the ticket (zero at rest)
  k_pt_scalars_step<<<grid, PTS_THREADS, 0, st>>>(pt_part, nblk, scal, with_gmax, stage,
                                                   reinterpret_cast<unsigned*>(stage + 4 * PTS_MAX_CTAS), cam_stage,
                                                   cam_rows);
}
void launch_evaluate(const ProblemView& pv, const double* pose, const double* X, double* r_reproj, double* r_depth,
                     cudaStream_t st) {
  if (pv.p1 > pv.p0) k_evaluate<<<points_grid(pv.p1 - pv.p0), PT_THREADS, 0, st>>>(pv, pose, X, r_reproj, r_depth);
}
void launch_filter_tracks(const ProblemView& pv, const double* pose, const double* X, double max_err2, double min_angle,
                          const uint8_t* pt_select, uint8_t* obs_keep, uint8_t* pt_keep, double* pt_error, double* scal,
                          cudaStream_t st) {
  if (pv.p1 > pv.p0)
    k_filter_tracks<<<points_grid(pv.p1 - pv.p0), PT_THREADS, 0, st>>>(pv, pose, X, max_err2, min_angle, pt_select,
                                                                       obs_keep, pt_keep, pt_error, scal);
}
void launch_count_behind(const ProblemView& pv, const double* pose, const double* X, double* out, cudaStream_t st) {
  if (pv.p1 > pv.p0) k_count_behind<<<points_grid(pv.p1 - pv.p0), PT_THREADS, 0, st>>>(pv, pose, X, out);
}
void launch_normalize_poses(int C, const double* in, const double* shift_scale, double* out, cudaStream_t st) {
  if (C > 0) k_normalize_poses<<<(C + 127) / 128, 128, 0, st>>>(C, in, shift_scale, out);
}
void launch_flag_points(int P, const uint8_t* pt_const, double* X, cudaStream_t st) {
  if (P > 0) k_flag_points<<<(P + 255) / 256, 256, 0, st>>>(P, pt_const, X);
}
void launch_expand_hcc(int C, const double* Hc, double* Hcc36, cudaStream_t st) {
  if (C > 0) k_expand_hcc<<<(C * 36 + 255) / 256, 256, 0, st>>>(C, Hc, Hcc36);
}

}  // namespace mpba
