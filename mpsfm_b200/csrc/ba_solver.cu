// ba_solver.cu -- C ABI (include/mpba.h) and the host side of the Levenberg-Marquardt loop.
//
// The loop restates Ceres 2.1's TrustRegionMinimizer + LevenbergMarquardtStrategy, which is what
// the reference runs through pyceres.solve (/root/reference/mpsfm/sfm/mapper/bundle_adjustment.py:285-293:
// SPARSE_SCHUR, otherwise Ceres defaults).  Order of events per iteration:
//   linear solve -> invalid step? -> candidate cost -> parameter tolerance -> function tolerance ->
//   accept / reject -> radius update -> max-iterations / gradient / min-radius checks.
// Everything numeric runs in the kernels of ba_kernels.cu on the handle's stream; the host only
// reads back one block of 16 scalars per LM iteration to take the accept / reject decision.
#include <cuda_runtime.h>
#include <float.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <string>
#include <vector>

#include <nvtx3/nvToolsExt.h>  // header-only NVTX 3: no link dependency, a pointer test per call when no tool is attached

#include "../../include/mpba.h"
#include "ba_kernels.cuh"
#include "nccl_dyn.h"

using namespace mpba;

enum KernelSlot {
  KS_IMG_CHUNKS = 0, KS_IMG_FINALIZE, KS_POINTS_SCHUR, KS_S_FINALIZE, KS_PCG, KS_CAM_UPDATE, KS_BACK_SUBSTITUTE,
  KS_REDUCE, KS_EVALUATE, KS_MISC, KS_COARSE, KS_BLOCK_INVERSE, KS_COLLECTIVE, KS_COARSE_SIDE
};
static const char* kKernelNames[MPBA_NUM_KERNELS] = {
    "img_chunks", "img_finalize", "points", "schur_blocks", "pcg", "cam_update", "back_substitute",
    "reduce", "evaluate", "misc", "coarse_setup", "block_inverse", "nccl_collectives", "coarse_setup_side_stream"};

struct TimedEvent {
  cudaEvent_t a, b;
  int slot;
};

struct mpba_handle {
  int device = 0;
  cudaStream_t stream = nullptr;
  std::string err;
  double* h_scal = nullptr;  // pinned
  std::vector<TimedEvent> pool;
  size_t used = 0;
  bool profile = false;
  double ms[MPBA_NUM_KERNELS] = {0};
  int64_t launches[MPBA_NUM_KERNELS] = {0};
  int64_t launch_count = 0;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  int pcg_grid_sms = 0;
  int32_t* h_stage = nullptr;  // pinned staging for the PCG cluster plan
  size_t h_stage_ints = 0;
  // what this handle last uploaded to the plan region of the workspace: an identical plan (same bundle structure solved
  // again, the mapper's and the bench's common case) is not sent again -- a host -> device copy inside the solve would
  // queue behind whatever the caller has in flight on the H2D copy engine (e.g. the upload of the NEXT problem on
  // another stream: measured, it delayed the first PCG by the whole 8 ms upload at cfg3)
  std::vector<int32_t> plan_copy;
  const int32_t* plan_dev = nullptr;
  bool plan_valid = false;
  unsigned long long plan_hash = 0;        // hash of plan_copy (plan_hash_term, ba_kernels.cuh)
  unsigned long long* plan_hash_seen_pinned = nullptr;  // pinned: hash of what the plan region holds NOW (this call)
  bool plan_hash_fresh = false;
  // side stream of the lagged coarse-space setup: the dense inverse of iteration i is built while the
  // stream kernels of iteration i run and is used by the PCG of iteration i + 1
  cudaStream_t side = nullptr;
  cudaEvent_t ev_s_ready = nullptr, ev_asm_done = nullptr, ev_inv_done[2] = {nullptr, nullptr};
  // device-controlled LM loop (lm_loop.cu): pinned read-back of the LM state + iteration rows, and the instantiated
  // WHILE graph of the last solve, reused when the next solve has the same signature (same buffers, sizes, options)
  double* h_lm = nullptr;
  cudaGraph_t lm_graph = nullptr;
  cudaGraphExec_t lm_exec = nullptr;
  std::vector<unsigned char> lm_signature;
  int64_t lm_body_launches = 0;
  int64_t lm_body_slot[MPBA_NUM_KERNELS] = {0};
  // sharded solve (mpba_comm_init)
  ncclComm_t comm = nullptr;
  int rank = 0, world = 1;
  int* d_agree = nullptr;  // one int of device memory: the ranks' agreement on a call's local validation result
};

// accessors for the other translation units of the library (integrate.cu): the handle's layout stays private here
namespace mpba {
cudaStream_t handle_stream(mpba_handle* h) { return h->stream; }
int handle_device(mpba_handle* h) { return h->device; }
int handle_fail(mpba_handle* h, int code, const std::string& msg) {
  if (h) h->err = msg;
  return code;
}
void handle_count_launches(mpba_handle* h, int n) { h->launch_count += n; }
}  // namespace mpba

namespace {

// Coarse groups of the two-level preconditioner: beyond 18 groups (126 modes) the dense inverse of the coarse matrix
// moves from one CTA to a four-CTA cluster (one cluster barrier per pivot, ~270 us for 224 modes), which only pays
// when the PCG iterations it saves are many: large systems.
// With the lagged setup (coarse_lag) the inverse is off the solve stream and systems beyond one cluster (grid PCG) take
// 48 groups (336 modes, sixteen-CTA cluster inverse).  Measured (r2, profiles/r02_coarse_groups.md): target10m 32 -> 48
// groups 2872 -> 1944 PCG iterations per solve, 237 -> 262 LM it/s, 64 groups 1897 / 257; cfg3 (cluster PCG) 32 -> 48
// groups 2024 -> 1531 iterations but every iteration pays the larger A^-1 product and the side stream's bigger inverse
// takes SMs from the stream kernels: 481 -> 475 LM it/s, so the cluster path stays at 32.
// MPBA_COARSE_GROUPS overrides the cap for sweeps.
int coarse_max_groups(int num_images, int lagged, bool cluster) {
  static const int env = getenv("MPBA_COARSE_GROUPS") ? atoi(getenv("MPBA_COARSE_GROUPS")) : 0;
  if (env > 0) return std::min(env, (int)COARSE_MAX_GROUPS);
  // mid-size bundles (profiles/r02_small_bundles.md): 50 images 8 groups 4.69 ms per solve, 18 groups 5.28 ms -- the
  // single-CTA inverse of the coarse matrix costs n^3 and is not hidden (no lag below 600 images)
  if (num_images < 64) return 8;
  if (num_images < 600) return 18;
  return (lagged && !cluster) ? 48 : 32;
}

int fail(mpba_handle* h, int code, const std::string& msg) {
  if (h) h->err = msg;
  return code;
}

#define NC(call)                                                                                       \
  do {                                                                                                 \
    ncclResult_t e_ = (call);                                                                          \
    if (e_ != ncclSuccess)                                                                             \
      return fail(h, MPBA_E_NCCL, std::string(#call) + ": " + nccl_api().GetErrorString(e_));          \
  } while (0)

#define CU(call)                                                                                       \
  do {                                                                                                 \
    cudaError_t e_ = (call);                                                                           \
    if (e_ != cudaSuccess)                                                                             \
      return fail(h, MPBA_E_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_));                 \
  } while (0)

// The workspace is the caller's: between two calls anything may have written to it.  So the skip of the plan upload is
// decided on what the region HOLDS: before the structure download of this call a one-CTA kernel hashes the region
// (only if it is the region this handle uploaded to last and it lies inside the workspace passed now), the 8 bytes
// come back with that download, and the copy is skipped only if the staged plan equals the last upload AND the
// region still hashes to it.
int enqueue_plan_hash(mpba_handle* h, const int32_t* plan_region, size_t capacity_ints, double* dev_scratch) {
  h->plan_hash_fresh = false;
  if (!h->plan_valid || h->plan_dev != plan_region || h->plan_copy.empty() || h->plan_copy.size() > capacity_ints)
    return MPBA_OK;
  launch_plan_hash(plan_region, h->plan_copy.size(), reinterpret_cast<unsigned long long*>(dev_scratch), h->stream);
  CU(cudaMemcpyAsync(&h->plan_hash_seen_pinned[0], dev_scratch, sizeof(unsigned long long), cudaMemcpyDeviceToHost, h->stream));
  h->plan_hash_fresh = true;  // valid after the caller's stream synchronisation
  return MPBA_OK;
}

unsigned long long plan_hash_of(const int32_t* v, size_t n) {
  unsigned long long acc = 0;
  for (size_t i = 0; i < n; ++i) acc += plan_hash_term(v[i], i);
  return acc;
}

// host -> device copy of the staged plan arrays, skipped when the region already holds exactly these values
int upload_plan(mpba_handle* h, int32_t* dev, const int32_t* staged, size_t n) {
  if (h->plan_valid && h->plan_hash_fresh && h->plan_dev == dev && h->plan_copy.size() == n && n > 0 &&
      memcmp(h->plan_copy.data(), staged, n * sizeof(int32_t)) == 0 && h->plan_hash_seen_pinned[0] == h->plan_hash)
    return MPBA_OK;
  h->plan_valid = false;
  CU(cudaMemcpyAsync(dev, staged, n * sizeof(int32_t), cudaMemcpyHostToDevice, h->stream));
  h->plan_copy.assign(staged, staged + n);
  h->plan_hash = plan_hash_of(staged, n);
  h->plan_dev = dev;
  h->plan_valid = true;
  return MPBA_OK;
}

// NVTX range for the lifetime of a scope (domain-less, named): "mpba_solve", "lm_iteration", and one per stage
// through Timer below -- what `ncu --nvtx --nvtx-include` and timeline tools filter on
struct NvtxRange {
  explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
  ~NvtxRange() { nvtxRangePop(); }
  NvtxRange(const NvtxRange&) = delete;
  NvtxRange& operator=(const NvtxRange&) = delete;
};

struct Timer {
  mpba_handle* h;
  int slot;
  size_t idx = (size_t)-1;
  cudaStream_t st;
  NvtxRange range;
  Timer(mpba_handle* h_, int slot_, int nkernels = 1, cudaStream_t st_ = nullptr, bool other = false)
      : h(h_), slot(slot_), st(other ? st_ : h_->stream), range(kKernelNames[slot_]) {
    h->launch_count += nkernels;
    h->launches[slot] += nkernels;
    if (!h->profile) return;
    if (h->used == h->pool.size()) {
      TimedEvent t;
      cudaEventCreate(&t.a);
      cudaEventCreate(&t.b);
      t.slot = slot;
      h->pool.push_back(t);
    }
    idx = h->used++;
    h->pool[idx].slot = slot;
    cudaEventRecord(h->pool[idx].a, st);
  }
  ~Timer() {
    if (idx != (size_t)-1) cudaEventRecord(h->pool[idx].b, st);
  }
};

// call after a stream synchronize
void flush_timers(mpba_handle* h) {
  for (size_t i = 0; i < h->used; ++i) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, h->pool[i].a, h->pool[i].b) == cudaSuccess) h->ms[h->pool[i].slot] += ms;
  }
  h->used = 0;
}

size_t align_up(size_t x) { return (x + 255) & ~(size_t)255; }

struct Workspace {
  double *pose[2], *X[2], *chunk_part, *Hc[2], *gc[2], *ip[2], *c_img, *c_pt, *Hpp, *gp, *Hinv, *fac, *S,
      *rhs, *Minv, *y, *vr, *vz, *vp0, *vp1, *vq, *partial, *pt_part, *scal, *sr[2], *Linv, *tpart, *pts_stage, *cam_stage;
  double *Zc = nullptr, *Ac = nullptr, *Ainv = nullptr;  // coarse space (coarse.cu)
  double *Zc2 = nullptr, *Ainv2 = nullptr;               // lagged setup: the side stream's own Z, the second inverse
  int32_t* pm_pt = nullptr;   // point of every track-order slot (launch_track_setup)
  int4* cta_desc = nullptr;   // per CTA of the track-order passes
  int4* warp_desc = nullptr;  // per warp group
  int32_t* plan = nullptr;  // PCG cluster plan (pcg_cluster.cu)
  size_t plan_ints = 0;
  size_t img_span = 0;  // doubles of [Hc | gc | per-image costs] of one buffer set
  double* ld = nullptr;            // device-controlled LM loop: LD_COUNT state doubles, then the iteration rows
  mpba_iteration* rows = nullptr;
  size_t bytes = 0;
};
constexpr size_t LM_DEV_DOUBLES = LD_COUNT + (size_t)(LM_DEVICE_MAX_ITERATIONS + 2) * sizeof(mpba_iteration) / sizeof(double);

size_t plan_capacity_ints(const mpba_problem* p) {
  if (p->flags & MPBA_FLAG_STRUCTURE_ONLY) return 0;
  return 8 * (size_t)p->num_upper + 3 * 17 + 64 + (size_t)p->num_blocks + COARSE_MAX_GROUPS * COARSE_MAX_GROUPS + COARSE_MAX_GROUPS + 8;
}

// bump allocation; base == nullptr just measures.  sh != nullptr: sharded solve (per-point partial buffers cover
// the owned points only, the point buffers are padded to world * pt_stride for the all-gather).
void carve(const mpba_problem* p, const mpba_shard* sh, char* base, Workspace& w) {
  size_t off = 0;
  auto take = [&](size_t n_doubles) -> double* {
    double* ptr = base ? reinterpret_cast<double*>(base + off) : nullptr;
    off += align_up(n_doubles * sizeof(double));
    return ptr;
  };
  const size_t C = p->num_images, P = p->num_points, B = p->num_blocks;
  const size_t nblk = (size_t)track_warp_grid(p->num_track_warps);
  const size_t Ppad = sh ? std::max(P, (size_t)sh->world * (size_t)sh->pt_stride) : P;
  for (int k = 0; k < 2; ++k) {
    w.pose[k] = take(C * 8);
    w.X[k] = take(Ppad * 4);
    // image blocks, gradients and per-image costs of one buffer set are ONE span: a sharded solve all-reduces it in
    // one call (img_span doubles from Hc[k])
    // (sharded: + one row of S_COUNT scalars per rank -- every rank fills its own row, zeros elsewhere, so the SUM
    // all-reduce of the span also gathers the per-rank scalar blocks: one collective instead of two)
    const size_t hc = (C * 21 + 1) & ~(size_t)1;
    const size_t rows = sh ? (size_t)sh->world * S_COUNT : 0;
    w.Hc[k] = take(hc + C * 6 + C * 2 + rows);
    w.gc[k] = w.Hc[k] ? w.Hc[k] + hc : nullptr;
    w.ip[k] = w.Hc[k] ? w.gc[k] + C * 6 : nullptr;
    w.sr[k] = w.Hc[k] ? w.ip[k] + C * 2 : nullptr;
    w.img_span = hc + C * 6 + C * 2 + rows;
  }
  w.chunk_part = take((size_t)p->num_chunks * IMG_PART);
  w.c_img = take(C * 6);
  w.c_pt = take(P * 3);
  w.Hpp = take(P * 6);
  w.gp = take(P * 3);
  w.Hinv = take(P * 6);
  w.Linv = take(P * 9);
  w.fac = take((p->flags & MPBA_FLAG_STRUCTURE_ONLY) ? 0 : (size_t)p->num_obs * FAC_STRIDE);
  w.S = take(B * 36 + C * 6);  // S | rhs: one span, one all-reduce of a sharded solve
  w.rhs = w.S ? w.S + B * 36 : nullptr;
  w.Minv = take(C * 36);
  w.y = take(C * 6);
  w.vr = take(C * 6);
  w.vz = take(C * 6);
  w.vp0 = take(C * 6);
  w.vp1 = take(C * 6);
  w.vq = take(C * 6);
  w.partial = take(2 * 2048 + 32);
  w.tpart = take(8 * 1024);
  w.pt_part = take(nblk * PT_PART + 8);
  w.scal = take(S_COUNT);
  w.pts_stage = take(PT_SCALARS_STAGE);
  w.cam_stage = take(3 * (size_t)cam_step_rows((int)C) + 2);
  {
    const bool so = (p->flags & MPBA_FLAG_STRUCTURE_ONLY) != 0;
    w.Zc = take(so ? 0 : C * 42);
    w.Ac = take(so ? 0 : (size_t)COARSE_SPLIT * COARSE_MAX_DIM * COARSE_MAX_DIM);
    w.Ainv = take(so ? 0 : (size_t)COARSE_MAX_DIM * COARSE_MAX_DIM);
    w.Zc2 = take(so ? 0 : C * 42);
    w.Ainv2 = take(so ? 0 : (size_t)COARSE_MAX_DIM * COARSE_MAX_DIM);
  }
  w.pm_pt = reinterpret_cast<int32_t*>(take(((size_t)p->num_obs + 1) / 2 + 1));
  w.cta_desc = reinterpret_cast<int4*>(take(2 * nblk + 2));
  w.warp_desc = reinterpret_cast<int4*>(take(2 * (size_t)p->num_track_warps + 2));
  w.plan_ints = plan_capacity_ints(p);
  w.plan = reinterpret_cast<int32_t*>(take((w.plan_ints + 1) / 2));
  w.ld = take(LM_DEV_DOUBLES);
  w.rows = w.ld ? reinterpret_cast<mpba_iteration*>(w.ld + LD_COUNT) : nullptr;
  w.bytes = off;
}

int check_problem(mpba_handle* h, const mpba_problem* p) {
  if (!p) return fail(h, MPBA_E_INVALID, "problem is NULL");
  if (p->num_images <= 0 || p->num_points < 0 || p->num_obs < 0 || p->num_cameras <= 0)
    return fail(h, MPBA_E_INVALID, "problem sizes must be positive (num_images, num_cameras) / non-negative");
  if (!p->img_pose || !p->img_camera || !p->img_free_mask || !p->img_shift_scale || !p->cam_model || !p->cam_params)
    return fail(h, MPBA_E_INVALID, "image / camera arrays must not be NULL");
  if (p->num_points > 0 && (!p->pt_xyz || !p->pt_const || !p->pt_ptr))
    return fail(h, MPBA_E_INVALID, "point arrays must not be NULL");
  if (p->num_obs > 0 && (!p->pm_img || !p->pm_xy || !p->pm_depth || !p->cm_pt || !p->cm_xy || !p->cm_depth))
    return fail(h, MPBA_E_INVALID, "observation arrays must not be NULL");
  if (!p->img_ptr || !p->img_chunk_ptr || (p->num_chunks > 0 && (!p->chunk_img || !p->chunk_begin || !p->chunk_end)))
    return fail(h, MPBA_E_INVALID, "image-order chunk arrays must not be NULL");
  if (p->num_blocks < p->num_images || !p->s_row_ptr || !p->s_col || !p->s_diag || !p->s_transpose)
    return fail(h, MPBA_E_INVALID, "reduced camera system structure missing (needs at least the diagonal blocks)");
  if (!(p->flags & MPBA_FLAG_STRUCTURE_ONLY) &&
      (!p->s_row || p->num_upper < p->num_images || !p->u_block || !p->u_item_ptr || (p->num_obs > 0 && (!p->item_a || !p->item_b))))
    return fail(h, MPBA_E_INVALID, "Schur work lists (s_row, u_block, u_item_ptr, item_a, item_b) missing");
  if (p->num_points > 0 && (p->num_track_warps <= 0 || !p->track_warp_first))
    return fail(h, MPBA_E_INVALID, "track_warp_first missing (mpba_pack_tracks)");
  if (p->reproj_loss < 0 || p->reproj_loss > 3 || p->depth_loss < 0 || p->depth_loss > 3)
    return fail(h, MPBA_E_INVALID, "unknown loss type");
  if ((p->pm_plane == nullptr) != (p->cm_plane == nullptr))
    return fail(h, MPBA_E_INVALID, "normal-prior rows: pm_plane and cm_plane must both be given or both be NULL");
  if (p->pm_plane && !(p->plane_loss_scale > 0.0))
    return fail(h, MPBA_E_INVALID, "normal-prior rows: plane_loss_scale must be > 0");
  return MPBA_OK;
}

ProblemView view_of(const mpba_problem* p) {
  ProblemView v;
  v.C = p->num_images; v.K = p->num_cameras; v.P = p->num_points; v.N = p->num_obs; v.B = p->num_blocks;
  v.num_chunks = p->num_chunks;
  v.img_camera = p->img_camera; v.img_free_mask = p->img_free_mask; v.img_shift_scale = p->img_shift_scale;
  v.cam_model = p->cam_model; v.cam_params = p->cam_params; v.pt_const = p->pt_const; v.pt_ptr = p->pt_ptr;
  v.pm_img = p->pm_img; v.pm_xy = p->pm_xy; v.pm_depth = p->pm_depth; v.img_ptr = p->img_ptr; v.cm_pt = p->cm_pt;
  v.cm_xy = p->cm_xy; v.cm_depth = p->cm_depth; v.chunk_img = p->chunk_img; v.chunk_begin = p->chunk_begin;
  v.chunk_end = p->chunk_end; v.img_chunk_ptr = p->img_chunk_ptr; v.s_row_ptr = p->s_row_ptr; v.s_col = p->s_col;
  v.s_diag = p->s_diag; v.s_transpose = p->s_transpose; v.s_row = p->s_row; v.U = p->num_upper;
  v.u_block = p->u_block; v.u_item_ptr = p->u_item_ptr; v.item_a = p->item_a; v.item_b = p->item_b;
  v.loss.reproj_loss = p->reproj_loss; v.loss.depth_loss = p->depth_loss;
  v.loss.reproj_scale = p->reproj_loss_scale; v.loss.reproj_weight = p->reproj_loss_weight;
  v.loss.behind_strict = 0;
  v.pm_plane = p->pm_plane; v.cm_plane = p->cm_plane;
  v.loss.plane_scale = p->pm_plane ? p->plane_loss_scale : 1.0;
  v.p0 = 0; v.p1 = v.P; v.i0 = 0; v.i1 = v.C;
  v.num_track_warps = p->num_track_warps; v.tw_first = p->track_warp_first;
  return v;
}

struct Ctx {
  mpba_handle* h;
  ProblemView pv;
  Workspace w;
  int nblk;
  int structure_only;
  int max_row_blocks;
  // cluster PCG plan for this call (built by build_pcg_cluster_plan); cl == 0: use the grid kernel
  int cl = 0;
  size_t cl_smem = 0;
  PcgClusterArgs cla;
  CoarsePlan coarse;  // coarse.G == 0: block-Jacobi only
  PcgGridCoarse gridco;  // grid kernel: groups of CTAs (build_grid_coarse_plan)
  const mpba_shard* sh = nullptr;  // sharded solve
  bool want_hinv = false;          // Hpp^-1 rows requested (mpba_point_covariances)
  bool defer_gmax = false;         // LM loop: max |gp| is reduced together with the step scalars (one launch less)
  const LmTail* tail = nullptr;    // device LM loop: the controller rides on the next fused image tail kernel
  const PtScalarsArgs* tail_pts = nullptr;  // small bundles: the step's point scalars are reduced by that kernel too
  bool tail_done = false;
  const double* scal_rows = nullptr;  // sharded: the gathered per-rank scalar rows of the last image pass
  bool dense = false;              // reduced system solved directly by one CTA (dense_solve.cu): <= 21 images
  bool device_lm = false;          // LM controller on the device: no memset nodes in the loop body (lm_loop.cu)
  int coarse_lag = 0;              // 1: the PCG of linear solve i uses the coarse inverse built from S of solve i - 1
  int solves = 0;                  // linear solves enqueued so far in this call
};

int check_shard(mpba_handle* h, const mpba_problem* p, const mpba_shard* sh) {
  if (!sh) return MPBA_OK;
  if (!h->comm || h->world != sh->world || h->rank != sh->rank)
    return fail(h, MPBA_E_NCCL, "sharded solve: call mpba_comm_init with the same world / rank first");
  if (sh->world < 1 || sh->rank < 0 || sh->rank >= sh->world) return fail(h, MPBA_E_INVALID, "bad shard rank / world");
  if (sh->pt_begin < 0 || sh->pt_end < sh->pt_begin || sh->pt_end > p->num_points || sh->pt_stride < 0 ||
      sh->pt_begin != std::min<int64_t>((int64_t)sh->rank * sh->pt_stride, p->num_points) ||
      sh->pt_end - sh->pt_begin > sh->pt_stride ||
      (int64_t)sh->world * sh->pt_stride < p->num_points)
    return fail(h, MPBA_E_INVALID, "shard point range must be [rank * pt_stride, ...) inside num_points, world * pt_stride >= num_points");
  if (sh->img_begin < 0 || sh->img_end < sh->img_begin || sh->img_end > p->num_images)
    return fail(h, MPBA_E_INVALID, "shard image range outside num_images");
  return MPBA_OK;
}

int begin(mpba_handle* h, const mpba_problem* prob, void* workspace, size_t workspace_bytes, Ctx& c,
          const mpba_shard* sh = nullptr) {
  if (!h) return MPBA_E_INVALID;
  int rc = check_problem(h, prob);
  if (rc) return rc;
  rc = check_shard(h, prob, sh);
  if (rc) return rc;
  Workspace probe;
  carve(prob, sh, nullptr, probe);
  if (!workspace || workspace_bytes < probe.bytes)
    return fail(h, MPBA_E_WORKSPACE, "workspace too small: need " + std::to_string(probe.bytes) + " bytes");
  CU(cudaSetDevice(h->device));
  c.h = h;
  c.pv = view_of(prob);
  carve(prob, sh, static_cast<char*>(workspace), c.w);
  c.sh = sh;
  if (sh) {
    c.pv.p0 = sh->pt_begin; c.pv.p1 = sh->pt_end; c.pv.i0 = sh->img_begin; c.pv.i1 = sh->img_end;
  }
  c.nblk = track_warp_grid(c.pv.num_track_warps);
  c.pv.pm_pt = c.w.pm_pt;
  c.pv.cta_desc = c.w.cta_desc;
  c.pv.warp_desc = c.w.warp_desc;
  if (c.pv.num_track_warps > 0) launch_track_setup(c.pv, c.w.pm_pt, c.w.cta_desc, c.w.warp_desc, h->stream);
  c.structure_only = prob->flags & MPBA_FLAG_STRUCTURE_ONLY;
  c.max_row_blocks = prob->s_max_row_blocks;
  h->launch_count = 0;
  return MPBA_OK;
}

// working copy of the caller's points, slot 3 of every record = constant flag (read by the image-order pass)
int stage_points(Ctx& c, const mpba_problem* prob, double* X) {
  mpba_handle* h = c.h;
  if (c.pv.P > 0) {
    CU(cudaMemcpyAsync(X, prob->pt_xyz, (size_t)c.pv.P * 4 * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
    launch_flag_points(c.pv.P, c.pv.pt_const, X, h->stream);
    h->launch_count += 1;
  }
  return MPBA_OK;
}

// cost / Hcc / gc of state (pose, X) into buffer set k.  Sharded: every rank reduces the chunks of its own images
// (zeros elsewhere), the image blocks are then all-reduced -- a sum with exact zeros, so the result is the owner's
// value bit for bit on every rank.
int enqueue_image_pass(Ctx& c, const double* pose, const double* X, int k, int compute_scaling, int jacobi) {
  mpba_handle* h = c.h;
  {
    Timer t(h, KS_IMG_CHUNKS);
    launch_img_chunks(c.pv, pose, X, c.w.chunk_part, h->stream);
  }
  if (c.sh) {
    {
      Timer t(h, KS_IMG_FINALIZE, 1);
      launch_img_finalize(c.pv, c.w.chunk_part, c.w.Hc[k], c.w.gc[k], c.w.ip[k], h->stream);
    }
    Timer t(h, KS_COLLECTIVE);
    NcclApi& nc = nccl_api();
    launch_stage_scalars(c.w.scal, c.w.sr[k], c.sh->rank, c.sh->world, h->stream);
    c.scal_rows = c.w.sr[k];
    NC(nc.AllReduce(c.w.Hc[k], c.w.Hc[k], c.w.img_span, ncclDouble, ncclSum, h->comm, h->stream));
  } else if (c.pv.C <= IMG_TAIL_MAX_IMAGES) {
    // small bundle: finalize + scalars (+ the LM decision of the device loop) in one single-CTA kernel
    Timer t(h, KS_IMG_FINALIZE, 1);
    launch_img_tail(c.pv, c.w.chunk_part, c.w.Hc[k], c.w.gc[k], c.w.ip[k], c.w.c_img, compute_scaling, jacobi,
                    c.w.scal, c.tail, c.tail_pts, h->stream);
    if (c.tail) c.tail_done = true;
    return MPBA_OK;
  } else {
    Timer t(h, KS_IMG_FINALIZE, 1);
    launch_img_finalize(c.pv, c.w.chunk_part, c.w.Hc[k], c.w.gc[k], c.w.ip[k], h->stream);
  }
  {
    Timer t(h, KS_IMG_FINALIZE, 1);
    launch_img_scalars(c.pv, c.w.ip[k], c.w.Hc[k], c.w.gc[k], c.w.c_img, compute_scaling, jacobi, c.w.scal,
                       h->stream);
  }
  return MPBA_OK;
}

// sharded: the per-rank sums of the point passes and of the camera step -> one block, identical on every rank
int combine_scalars(Ctx& c) {
  mpba_handle* h = c.h;
  if (!c.sh) return MPBA_OK;
  // the rows came with the all-reduce of the image pass that just ran (enqueue_image_pass)
  if (c.scal_rows == nullptr) return fail(h, MPBA_E_INVALID, "combine_scalars without an image pass");
  Timer t(h, KS_COLLECTIVE, 1);
  launch_combine_scalars(c.scal_rows, c.sh->world, c.w.scal, h->stream);
  return MPBA_OK;
}

int fetch_scalars(Ctx& c) {
  mpba_handle* h = c.h;
  CU(cudaMemcpyAsync(h->h_scal, c.w.scal, S_COUNT * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
  CU(cudaStreamSynchronize(h->stream));
  CU(cudaGetLastError());
  if (h->profile) {
    if (h->side) CU(cudaStreamSynchronize(h->side));  // its events are read below
    flush_timers(h);
  }
  return MPBA_OK;
}

// Partition the block rows of S over the CTAs of one cluster and lay out the deposit / halo slots of
// the cluster-resident PCG (pcg_cluster.cu).  Runs once per C-ABI call on the host: it needs the
// block-CSR structure (two small device-to-host copies), costs O(num_blocks) and uploads ~8 ints per
// upper block.  Leaves c.cl == 0 when the system does not fit one cluster (the grid kernel is used).
int build_pcg_cluster_plan(Ctx& c, const mpba_problem* prob, int pcg_kernel, int pcg_coarse) {
  mpba_handle* h = c.h;
  c.cl = 0;
  c.coarse.G = 0;
  // small systems (the mapper's local bundles): one CTA solves S y = rhs directly, no plan to build
  c.dense = !c.structure_only && (pcg_kernel == 0 || pcg_kernel == 3) && dense_solve_fits(prob->num_images);
  if (c.structure_only || pcg_kernel == 1 || c.dense) return MPBA_OK;
  const int C = prob->num_images, B = prob->num_blocks;
  std::vector<int32_t> rp(C + 1), col(B);
  {
    const int hrc = enqueue_plan_hash(h, c.w.plan, c.w.plan_ints, c.w.scal);
    if (hrc) return hrc;
  }
  CU(cudaMemcpyAsync(rp.data(), prob->s_row_ptr, (C + 1) * sizeof(int32_t), cudaMemcpyDeviceToHost, h->stream));
  CU(cudaMemcpyAsync(col.data(), prob->s_col, (size_t)B * sizeof(int32_t), cudaMemcpyDeviceToHost, h->stream));
  CU(cudaStreamSynchronize(h->stream));
  if (rp[0] != 0 || rp[C] != B) return fail(h, MPBA_E_INVALID, "s_row_ptr does not span num_blocks");
  std::vector<int> nup(C, 0);
  int U = 0;
  for (int i = 0; i < C; ++i) {
    for (int b = rp[i]; b < rp[i + 1]; ++b) {
      if (col[b] < 0 || col[b] >= C) return fail(h, MPBA_E_INVALID, "s_col out of range");
      if (col[b] > i) ++nup[i];  // strictly upper: the diagonal block is applied by the row owner
    }
    U += nup[i];
  }
  if (plan_capacity_ints(prob) > c.w.plan_ints || (size_t)(8 * (size_t)U + 3 * 17 + 64) > c.w.plan_ints) return MPBA_OK;

  const int max_smem = pcg_cluster_max_smem(h->device);
  static const int env_min_cl = getenv("MPBA_PCG_MIN_CLUSTER") ? atoi(getenv("MPBA_PCG_MIN_CLUSTER")) : 0;
  // at least four CTAs: one CTA applying the whole preconditioner (17 groups' rows of A^-1) made the 50-image solve
  // 7.3 ms, four CTAs 4.7 ms (r2 sweep); the dense direct solve already took the systems below 22 images
  int first_cl = C >= 16 ? 4 : 1;
  if (env_min_cl > 0) first_cl = 1;
  while (first_cl < 16 && first_cl < env_min_cl) first_cl *= 2;
  for (int cl = first_cl; cl <= 16; cl *= 2) {
    // contiguous rows, balanced by upper-block count
    std::vector<int> crow(cl + 1, C);
    crow[0] = 0;
    {
      const double target = (double)(U + C) / cl;  // weight of a row: its upper blocks + 1 (row work)
      int k = 1;
      long long acc = 0;
      for (int i = 0; i < C && k < cl; ++i) {
        acc += nup[i] + 1;
        while (k < cl && acc >= target * k) crow[k++] = i + 1;
      }
    }
    std::vector<int> owner(C);
    for (int k = 0; k < cl; ++k)
      for (int i = crow[k]; i < crow[k + 1]; ++i) owner[i] = k;
    int max_rows = 1, max_full = 1, max_blk = 0;
    for (int k = 0; k < cl; ++k) {
      max_rows = std::max(max_rows, crow[k + 1] - crow[k]);
      max_full = std::max(max_full, rp[crow[k + 1]] - rp[crow[k]]);
      int nb = 0;
      for (int i = crow[k]; i < crow[k + 1]; ++i) nb += nup[i];
      max_blk = std::max(max_blk, nb);
    }
    if (max_blk > PCGC_THREADS && cl < 16) continue;  // prefer every block in registers
    if (max_rows >= (1 << 20)) return MPBA_OK;
    // halo lists and block records
    std::vector<int32_t> cta_blk(cl + 1, 0), cta_halo(cl + 1, 0);
    std::vector<int32_t> ub_sblk, ub_rowl, ub_halo, ub_depA, ub_tcta, ub_tidx, halo_cta, halo_rowl;
    ub_sblk.reserve(U);
    std::vector<int> slot_of(C, -1), touched;
    int max_halo = 1;
    for (int k = 0; k < cl; ++k) {
      touched.clear();
      const int full0 = rp[crow[k]];
      for (int i = crow[k]; i < crow[k + 1]; ++i) {
        for (int b = rp[i]; b < rp[i + 1]; ++b) {
          const int j = col[b];
          if (j <= i) continue;
          if (slot_of[j] < 0) {
            slot_of[j] = (int)touched.size();
            touched.push_back(j);
            halo_cta.push_back(owner[j]);
            halo_rowl.push_back(j - crow[owner[j]]);
          }
          ub_sblk.push_back(b);
          ub_rowl.push_back(i - crow[k]);
          ub_halo.push_back(slot_of[j]);
          ub_depA.push_back(b - full0);
          {
            // slot of column i in the full row j
            const int32_t* lo = col.data() + rp[j];
            const int32_t* hi = col.data() + rp[j + 1];
            const int32_t* it = std::lower_bound(lo, hi, (int32_t)i);
            if (it == hi || *it != i) return fail(h, MPBA_E_INVALID, "reduced camera system structure is not symmetric");
            ub_tcta.push_back(owner[j]);
            ub_tidx.push_back((int)(it - col.data()) - rp[crow[owner[j]]]);
          }
        }
      }
      for (int j : touched) slot_of[j] = -1;
      max_halo = std::max(max_halo, (int)touched.size());
      cta_blk[k + 1] = (int32_t)ub_sblk.size();
      cta_halo[k + 1] = (int32_t)halo_cta.size();
    }
    PcgClusterArgs a;
    memset(&a, 0, sizeof(a));
    a.C = C;
    a.max_rows = max_rows;
    a.max_full = max_full;
    a.max_halo = max_halo;
    a.max_extra = std::max(1, max_blk - PCGC_THREADS);
    // coarse space: gpc groups of consecutive images per CTA, at least two images per group
    int min_rows = C;
    for (int k = 0; k < cl; ++k) min_rows = std::min(min_rows, crow[k + 1] - crow[k]);
    int gpc = (pcg_coarse == 1 || C < 8) ? 0 : std::min(coarse_max_groups(C, c.coarse_lag, true) / cl, std::max(min_rows / 2, 0));
    size_t smem = 0;
    for (;; --gpc) {
      a.gpc = gpc > 0 ? gpc : 0;
      a.G = a.gpc * cl;
      a.nc = 7 * a.G;
      smem = pcg_cluster_carve(a);
      if (a.gpc == 0 || (int)smem <= max_smem) break;
    }
    static const bool plan_dbg = getenv("MPBA_PLAN_DEBUG") != nullptr;
    if (plan_dbg)
      fprintf(stderr, "[mpba] cluster plan cl %d: rows<=%d full<=%d halo<=%d blk<=%d min_rows %d gpc %d -> %zu B of %d\n", cl,
              max_rows, max_full, max_halo, max_blk, min_rows, a.gpc, smem, max_smem);
    if ((int)smem > max_smem || !pcg_cluster_fits(h->device, cl, smem)) {
      if (cl == 16) return MPBA_OK;
      continue;
    }
    std::vector<int32_t> grp_row, cb_ptr, cb_fine;
    if (a.G) {
      const int G = a.G;
      grp_row.resize(G + 1);
      std::vector<int> grp_of(C);
      for (int k = 0; k < cl; ++k) {
        const int r0 = crow[k], nr = crow[k + 1] - crow[k];
        for (int g = 0; g <= a.gpc; ++g) grp_row[k * a.gpc + g] = r0 + (int)((long long)nr * g / a.gpc);
      }
      grp_row[G] = C;
      for (int g = 0; g < G; ++g)
        for (int i = grp_row[g]; i < grp_row[g + 1]; ++i) grp_of[i] = g;
      cb_ptr.assign((size_t)G * G + 1, 0);
      for (int i = 0; i < C; ++i)
        for (int b = rp[i]; b < rp[i + 1]; ++b) ++cb_ptr[(size_t)grp_of[i] * G + grp_of[col[b]] + 1];
      for (size_t k = 0; k < (size_t)G * G; ++k) cb_ptr[k + 1] += cb_ptr[k];
      cb_fine.resize(B);
      std::vector<int32_t> fill(cb_ptr.begin(), cb_ptr.end() - 1);
      for (int i = 0; i < C; ++i)
        for (int b = rp[i]; b < rp[i + 1]; ++b) cb_fine[fill[(size_t)grp_of[i] * G + grp_of[col[b]]]++] = b;
    }
    // pack, stage, upload
    const size_t nU = ub_sblk.size(), nH = halo_cta.size();
    const size_t total = 3 * (size_t)(cl + 1) + 6 * nU + 2 * nH + grp_row.size() + cb_ptr.size() + cb_fine.size();
    if (total > c.w.plan_ints) return MPBA_OK;
    if (h->h_stage_ints < total) {
      if (h->h_stage) cudaFreeHost(h->h_stage);
      h->h_stage = nullptr;
      h->h_stage_ints = 0;
      CU(cudaMallocHost(&h->h_stage, total * sizeof(int32_t)));
      h->h_stage_ints = total;
    }
    int32_t* st = h->h_stage;
    int32_t* dev = c.w.plan;
    size_t off = 0;
    auto put = [&](const std::vector<int32_t>& v) -> const int32_t* {
      if (!v.empty()) memcpy(st + off, v.data(), v.size() * sizeof(int32_t));
      const int32_t* d = dev + off;
      off += v.size();
      return d;
    };
    std::vector<int32_t> crow32(crow.begin(), crow.end());
    a.cta_row_ptr = put(crow32);
    a.cta_blk_ptr = put(cta_blk);
    a.cta_halo_ptr = put(cta_halo);
    a.ub_sblk = put(ub_sblk);
    a.ub_rowl = put(ub_rowl);
    a.ub_halo = put(ub_halo);
    a.ub_depA = put(ub_depA);
    a.ub_tcta = put(ub_tcta);
    a.ub_tidx = put(ub_tidx);
    a.halo_cta = put(halo_cta);
    a.halo_rowl = put(halo_rowl);
    a.grp_row_ptr = put(grp_row);
    c.coarse.G = a.G;
    c.coarse.cb_ptr = put(cb_ptr);
    c.coarse.cb_fine = put(cb_fine);
    a.Zc = c.w.Zc;
    a.Ainv = c.w.Ainv;
    {
      const int urc = upload_plan(h, dev, st, off);
      if (urc) return urc;
    }
    a.row_ptr = prob->s_row_ptr;
    a.s_diag = prob->s_diag;
    c.cla = a;
    c.cl = cl;
    c.cl_smem = smem;
    return MPBA_OK;
  }
  return MPBA_OK;
}

// Coarse space for the grid-wide PCG (systems beyond one cluster): G groups of consecutive CTAs' rows, so that every
// CTA contributes to exactly one group.  Same coarse kernels (coarse.cu) as the cluster path.
int build_grid_coarse_plan(Ctx& c, const mpba_problem* prob, int pcg_coarse) {
  mpba_handle* h = c.h;
  c.gridco = PcgGridCoarse();
  if (c.structure_only || c.cl > 0 || c.dense || pcg_coarse == 1) return MPBA_OK;
  const int C = prob->num_images, B = prob->num_blocks;
  const PcgPlan pl = pcg_plan(h->device, C, c.max_row_blocks);
  const int G = std::min(coarse_max_groups(C, c.coarse_lag, false), pl.grid);
  if (G < 2 || pl.grid > 1024) return MPBA_OK;
  std::vector<int32_t> rp(C + 1), col(B);
  {
    const int hrc = enqueue_plan_hash(h, c.w.plan, c.w.plan_ints, c.w.scal);
    if (hrc) return hrc;
  }
  CU(cudaMemcpyAsync(rp.data(), prob->s_row_ptr, (C + 1) * sizeof(int32_t), cudaMemcpyDeviceToHost, h->stream));
  CU(cudaMemcpyAsync(col.data(), prob->s_col, (size_t)B * sizeof(int32_t), cudaMemcpyDeviceToHost, h->stream));
  CU(cudaStreamSynchronize(h->stream));
  std::vector<int32_t> grp_row(G + 1), grp_of(C);
  for (int g = 0; g <= G; ++g) grp_row[g] = std::min(C, (int)((long long)pl.grid * g / G) * pl.rows_per_cta);
  grp_row[G] = C;
  for (int g = 0; g < G; ++g)
    for (int i = grp_row[g]; i < grp_row[g + 1]; ++i) grp_of[i] = g;
  std::vector<int32_t> cb_ptr((size_t)G * G + 1, 0), cb_fine(B);
  for (int i = 0; i < C; ++i)
    for (int b = rp[i]; b < rp[i + 1]; ++b) ++cb_ptr[(size_t)grp_of[i] * G + grp_of[col[b]] + 1];
  for (size_t k = 0; k < (size_t)G * G; ++k) cb_ptr[k + 1] += cb_ptr[k];
  std::vector<int32_t> fill(cb_ptr.begin(), cb_ptr.end() - 1);
  for (int i = 0; i < C; ++i)
    for (int b = rp[i]; b < rp[i + 1]; ++b) cb_fine[fill[(size_t)grp_of[i] * G + grp_of[col[b]]]++] = b;
  const size_t total = grp_row.size() + cb_ptr.size() + cb_fine.size();
  if (total > c.w.plan_ints) return MPBA_OK;
  if (h->h_stage_ints < total) {
    if (h->h_stage) cudaFreeHost(h->h_stage);
    h->h_stage = nullptr;
    h->h_stage_ints = 0;
    CU(cudaMallocHost(&h->h_stage, total * sizeof(int32_t)));
    h->h_stage_ints = total;
  }
  int32_t* st = h->h_stage;
  size_t off = 0;
  auto put = [&](const std::vector<int32_t>& v) -> const int32_t* {
    memcpy(st + off, v.data(), v.size() * sizeof(int32_t));
    const int32_t* d = c.w.plan + off;
    off += v.size();
    return d;
  };
  c.gridco.grp_row_ptr = put(grp_row);
  c.coarse.cb_ptr = put(cb_ptr);
  c.coarse.cb_fine = put(cb_fine);
  {
    const int urc = upload_plan(h, c.w.plan, st, off);
    if (urc) return urc;
  }
  c.coarse.G = G;
  c.gridco.G = G;
  c.gridco.Zc = c.w.Zc;
  c.gridco.Ainv = c.w.Ainv;
  c.gridco.tpart = c.w.tpart;
  return MPBA_OK;
}

int enqueue_points_and_system(Ctx& c, const double* pose, const double* X, int kset, LmDiag lm, int compute_scaling,
                              const mpba_options* opt, bool want_system, double* pose_cand = nullptr) {
  mpba_handle* h = c.h;
  const bool schur = want_system && !c.structure_only;
  if (!c.device_lm) CU(cudaMemsetAsync(c.w.scal + S_BAD_POINTS, 0, sizeof(double), h->stream));
  {
    Timer t(h, KS_POINTS_SCHUR);
    launch_points(c.pv, pose, X, lm, c.w.c_pt, compute_scaling, opt->jacobi_scaling, c.w.Hpp, c.w.gp,
                  c.want_hinv ? c.w.Hinv : nullptr, c.w.Linv, c.w.fac,
                  c.w.pt_part, c.w.scal, schur ? 1 : 0, h->stream);
  }
  if (!(c.defer_gmax && want_system)) {
    Timer t(h, KS_REDUCE);
    launch_pt_scalars_gmax(c.w.pt_part, c.nblk, c.w.scal, h->stream);
  }
  if (!want_system) return MPBA_OK;
  if (schur) {
    // the side stream's coarse assembly of the previous solve reads S and the poses: it must be through (it is, by
    // hundreds of microseconds) before S is rewritten
    if (c.coarse_lag != 0 && h->side != nullptr && c.coarse.G > 0 && c.solves > 0)
      CU(cudaStreamWaitEvent(h->stream, h->ev_asm_done, 0));
    {
      Timer t(h, KS_S_FINALIZE);
      launch_schur_blocks(c.pv, c.w.fac, c.w.Hc[kset], c.w.gc[kset], c.w.c_img, lm, c.w.S, c.w.rhs, c.sh ? 1 : 0,
                          (c.sh || c.dense) ? nullptr : c.w.Minv, h->stream);
    }
    if (c.sh) {
      // partial reduced camera system of this rank's tracks -> all-reduce -> + Hcc + D, + gc (north star:
      // "the partial reduced-camera-system blocks ... are combined with NCCL allreduce")
      Timer t(h, KS_COLLECTIVE, 2);
      NcclApi& nc = nccl_api();
      NC(nc.AllReduce(c.w.S, c.w.S, (size_t)c.pv.B * 36 + (size_t)c.pv.C * 6, ncclDouble, ncclSum, h->comm, h->stream));
      // ... and the block-Jacobi inverses of the completed diagonal blocks (unless the dense solve follows)
      launch_schur_finish(c.pv, c.w.Hc[kset], c.w.gc[kset], c.w.c_img, lm, c.w.S, c.w.rhs, c.dense ? nullptr : c.w.Minv,
                          h->stream);
    }
    if (c.dense) {
      Timer t(h, KS_PCG);
      // pose_cand given (LM iteration): the camera step runs as the epilogue of the same one-CTA kernel
      cudaError_t e = launch_dense_solve(c.pv.C, c.pv.s_row_ptr, c.pv.s_col, c.w.S, c.w.rhs, c.w.y, c.w.scal, &c.pv, pose,
                                         c.w.Hc[kset], c.w.gc[kset], c.w.c_img, lm, pose_cand, h->stream);
      if (e != cudaSuccess) return fail(h, MPBA_E_CUDA, std::string("dense solve launch: ") + cudaGetErrorString(e));
      ++c.solves;
      return MPBA_OK;
    }
    // (the block-Jacobi inverses were written by the diagonal CTAs of k_schur_blocks / k_schur_finish)
    // Coarse space.  Synchronous (coarse_lag = 0): modes, Z^T S Z and its dense inverse on the solve stream, right
    // before the PCG -- 279 us per LM iteration at cfg3, 232 us of it the latency-bound Gauss-Jordan on four CTAs.
    // Lagged (default): the inverse is a PRECONDITIONER, any SPD matrix serves; the one built from S of the previous
    // linear solve is used, and this solve's own is built on the side stream, where it overlaps the stream kernels
    // that follow (back-substitution, image pass, the next k_points: all of them leave SMs idle at their tails).
    // Z itself (six numbers per image and mode) is always the current one.  The first solve of a call waits for its
    // own inverse.  (Rebuilding only every k-th iteration ON the solve stream was measured in round 1: k = 2 gains
    // nothing, k = 4 costs 30 % more PCG iterations.)
    const double* ainv_use = c.w.Ainv;
    if (c.coarse.G > 0) {
      const bool lag = c.coarse_lag != 0 && h->side != nullptr;
      if (!lag) {
        Timer t(h, KS_COARSE, 3);
        cudaError_t e = launch_coarse_setup(c.coarse, c.pv.C, pose, c.pv.img_free_mask, c.pv.s_row, c.pv.s_col, c.w.S,
                                            c.w.Zc, c.w.Ac, c.w.Ainv, h->stream);
        if (e != cudaSuccess) return fail(h, MPBA_E_CUDA, std::string("coarse setup: ") + cudaGetErrorString(e));
      } else {
        double* abuf[2] = {c.w.Ainv, c.w.Ainv2};
        const int mine = c.solves & 1;
        {
          Timer t(h, KS_COARSE, 1);
          cudaError_t e = launch_coarse_modes(c.pv.C, pose, c.pv.img_free_mask, c.w.Zc, h->stream);
          if (e != cudaSuccess) return fail(h, MPBA_E_CUDA, std::string("coarse modes: ") + cudaGetErrorString(e));
        }
        CU(cudaEventRecord(h->ev_s_ready, h->stream));
        CU(cudaStreamWaitEvent(h->side, h->ev_s_ready, 0));
        {
          Timer t(h, KS_COARSE_SIDE, 3, h->side, true);
          cudaError_t e = launch_coarse_setup(c.coarse, c.pv.C, pose, c.pv.img_free_mask, c.pv.s_row, c.pv.s_col, c.w.S,
                                              c.w.Zc2, c.w.Ac, abuf[mine], h->side, h->ev_asm_done);
          if (e != cudaSuccess) return fail(h, MPBA_E_CUDA, std::string("coarse setup: ") + cudaGetErrorString(e));
        }
        CU(cudaEventRecord(h->ev_inv_done[mine], h->side));
        const int use = c.solves == 0 ? mine : (mine ^ 1);
        CU(cudaStreamWaitEvent(h->stream, h->ev_inv_done[use], 0));
        ainv_use = abuf[use];
      }
    }
    {
      Timer t(h, KS_PCG);
      cudaError_t e;
      if (c.cl > 0) {
        PcgClusterArgs a = c.cla;
        a.Ainv = ainv_use;
        a.S = c.w.S; a.Minv = c.w.Minv; a.rhs = c.w.rhs; a.x = c.w.y; a.scal = c.w.scal;
        a.tol = opt->pcg_tolerance; a.max_iter = opt->pcg_max_iterations;
        e = launch_pcg_cluster(a, c.cl, c.cl_smem, h->stream);
      } else {
        const PcgPlan pl = pcg_plan(h->device, c.pv.C, c.max_row_blocks);
        PcgGridCoarse gco = c.gridco;
        gco.Ainv = ainv_use;
        e = launch_pcg(c.pv, pl, c.w.S, c.w.Minv, c.w.rhs, c.w.y, c.w.vz, c.w.vp0, c.w.vp1, c.w.partial,
                       reinterpret_cast<unsigned*>(c.w.partial + 2 * 2048), c.w.scal, opt->pcg_tolerance,
                       opt->pcg_max_iterations, gco, h->stream);
      }
      if (e != cudaSuccess) return fail(h, MPBA_E_CUDA, std::string("PCG launch: ") + cudaGetErrorString(e));
    }
    ++c.solves;
  } else {
    // all poses constant (refine_3d_points, bundle_adjustment.py:276-283): no reduced system, y_c = 0
    CU(cudaMemsetAsync(c.w.y, 0, (size_t)c.pv.C * 6 * sizeof(double), h->stream));
    CU(cudaMemsetAsync(c.w.scal + S_PCG_ITERS, 0, 3 * sizeof(double), h->stream));
  }
  return MPBA_OK;
}

// One LM iteration from the state in buffer set `cur`: linear solve, candidate into set `cand`, cost and image blocks at
// the candidate.  Everything the controller (host loop or k_lm_control) reads afterwards is in c.w.scal.
int enqueue_lm_iteration(Ctx& c, int cur, int cand, LmDiag lm, int compute_scaling, const mpba_options& opt) {
  mpba_handle* h = c.h;
  int rc = enqueue_points_and_system(c, c.w.pose[cur], c.w.X[cur], cur, lm, compute_scaling, &opt, true,
                                     c.dense ? c.w.pose[cand] : nullptr);
  if (rc) return rc;
  // the camera step: epilogue of the dense solve (<= 21 images) or extra CTAs of the back-substitution launch
  CamStepArgs cam;
  cam.pose = c.w.pose[cur]; cam.Hc = c.w.Hc[cur]; cam.gc = c.w.gc[cur]; cam.c_img = c.w.c_img;
  cam.pose_cand = c.w.pose[cand];
  cam.stage = c.w.cam_stage;
  const int cam_rows = c.dense ? 0 : cam_step_rows(c.pv.C);
  {
    Timer t(h, KS_BACK_SUBSTITUTE);
    launch_back_substitute(c.pv, c.w.X[cur], c.w.y, c.w.Hpp, c.w.gp, c.w.Linv, c.w.c_pt, c.w.fac,
                           c.structure_only ? 0 : 1, lm, c.w.X[cand], c.w.pt_part, c.dense ? nullptr : &cam, h->stream);
  }
  // small bundles: the fused image tail of the candidate pass reduces the point scalars as well
  const bool pts_in_tail = !c.sh && c.pv.C <= IMG_TAIL_MAX_IMAGES && c.nblk <= IMG_TAIL_MAX_PT_ROWS;
  PtScalarsArgs pts = {c.w.pt_part, (int)c.nblk, c.defer_gmax ? 1 : 0, c.w.cam_stage, cam_rows};
  if (!pts_in_tail) {
    Timer t(h, KS_REDUCE);
    launch_pt_scalars_step(c.w.pt_part, c.nblk, c.w.scal, c.defer_gmax ? 1 : 0, c.w.pts_stage, c.w.cam_stage, cam_rows,
                           h->stream);
  }
  if (c.sh) {
    // every rank wrote the candidates of its own points: all-gather in place (equal counts, pt_stride per rank)
    Timer t(h, KS_COLLECTIVE);
    const size_t cnt = (size_t)c.sh->pt_stride * 4;
    NC(nccl_api().AllGather(c.w.X[cand] + (size_t)c.sh->rank * cnt, c.w.X[cand], cnt, ncclDouble, h->comm, h->stream));
  }
  // cost (and, in case the step is accepted, Hcc / gc) at the candidate
  c.tail_pts = pts_in_tail ? &pts : nullptr;
  rc = enqueue_image_pass(c, c.w.pose[cand], c.w.X[cand], cand, 0, opt.jacobi_scaling);
  c.tail_pts = nullptr;
  if (rc) return rc;
  return combine_scalars(c);
}

// ---- device-controlled LM loop (lm_loop.cu) -----------------------------------------------------------------------
// When: small and medium bundles (< 600 images, the cluster-resident PCG, single GPU, no per-kernel profiling) -- the
// mapper's local and early global bundle adjustments, where the host controller's launches and synchronisations cost
// more than the kernels.  lm_loop = 1 forces the host loop, 2 asks for the device loop wherever it is legal.
bool device_lm_legal(const Ctx& c, const mpba_options& opt) {
  return !c.sh && !c.h->profile && opt.verbose == 0 && !c.structure_only && (c.cl > 0 || c.dense) && c.coarse_lag == 0 &&
         opt.max_num_iterations <= LM_DEVICE_MAX_ITERATIONS && c.w.ld != nullptr && c.h->h_lm != nullptr &&
         c.h->side != nullptr;
}

struct DeviceLmResult {
  int termination = -1, msg = 0, iteration = 0, n_ok = 0, n_bad = 0, nrows = 0;
  int64_t total_pcg = 0;
  double minimum_cost = 0.0;
};

int run_device_lm(Ctx& c, const mpba_problem* prob, const mpba_options& opt, double radius, double x_cost,
                  double fixed_cost, double gmax_c, DeviceLmResult& out) {
  mpba_handle* h = c.h;
  c.device_lm = true;
  LmDiag lm = make_lm(radius, opt.min_lm_diagonal, opt.max_lm_diagonal);
  lm.dev = c.w.ld;
  LmControl ctl;
  ctl.function_tolerance = opt.function_tolerance;
  ctl.gradient_tolerance = opt.gradient_tolerance;
  ctl.parameter_tolerance = opt.parameter_tolerance;
  ctl.min_relative_decrease = opt.min_relative_decrease;
  ctl.max_trust_region_radius = opt.max_trust_region_radius;
  ctl.min_trust_region_radius = opt.min_trust_region_radius;
  ctl.max_num_iterations = opt.max_num_iterations;
  ctl.max_consecutive_invalid = opt.max_num_consecutive_invalid_steps;
  LmAcceptCopy cp;
  const size_t C = (size_t)c.pv.C, P = (size_t)c.pv.P;
  cp.src[0] = c.w.pose[1]; cp.dst[0] = c.w.pose[0]; cp.n[0] = C * 8;
  cp.src[1] = c.w.X[1];    cp.dst[1] = c.w.X[0];    cp.n[1] = P * 4;
  cp.src[2] = c.w.Hc[1];   cp.dst[2] = c.w.Hc[0];   cp.n[2] = C * 21;
  cp.src[3] = c.w.gc[1];   cp.dst[3] = c.w.gc[0];   cp.n[3] = C * 6;

  // everything a kernel node of the body bakes in: the problem (pointers and sizes), the workspace, the options, the
  // PCG cluster plan and its coarse space
  std::vector<unsigned char> sig;
  auto add = [&](const void* p, size_t n) { sig.insert(sig.end(), (const unsigned char*)p, (const unsigned char*)p + n); };
  add(prob, sizeof(*prob));
  add(&c.w, sizeof(c.w));
  add(&opt, sizeof(opt));
  add(&c.cla, sizeof(c.cla));
  add(&c.coarse, sizeof(c.coarse));
  add(&c.cl, sizeof(c.cl));
  add(&c.cl_smem, sizeof(c.cl_smem));
  add(&c.pv.loss, sizeof(c.pv.loss));

  launch_lm_init(c.w.ld, c.w.rows, radius, x_cost, fixed_cost, gmax_c, c.w.scal, h->stream);
  h->launch_count += 1;
  if (h->lm_exec == nullptr || sig != h->lm_signature) {
    if (h->lm_exec) cudaGraphExecDestroy(h->lm_exec);
    if (h->lm_graph) cudaGraphDestroy(h->lm_graph);
    h->lm_exec = nullptr;
    h->lm_graph = nullptr;
    h->lm_signature.clear();
    CU(cudaGraphCreate(&h->lm_graph, 0));
    cudaGraphConditionalHandle handle;
    CU(cudaGraphConditionalHandleCreate(&handle, h->lm_graph, 1, cudaGraphCondAssignDefault));
    cudaGraphNodeParams np = {};
    np.type = cudaGraphNodeTypeConditional;
    np.conditional.handle = handle;
    np.conditional.type = cudaGraphCondTypeWhile;
    np.conditional.size = 1;
    cudaGraphNode_t node;
    CU(cudaGraphAddNode(&node, h->lm_graph, nullptr, 0, &np));
    cudaGraph_t body = np.conditional.phGraph_out[0];
    const int64_t lc0 = h->launch_count;
    int64_t slot0[MPBA_NUM_KERNELS];
    memcpy(slot0, h->launches, sizeof(slot0));
    // captured on the handle's own stream (the caller's may be the legacy default stream, which cannot capture);
    // nothing runs during capture, the graph itself is launched on the caller's stream
    if (h->side == nullptr) return fail(h, MPBA_E_CUDA, "device LM loop: no capture stream");
    CU(cudaStreamSynchronize(h->side));
    CU(cudaStreamBeginCaptureToGraph(h->side, body, nullptr, nullptr, 0, cudaStreamCaptureModeRelaxed));
    cudaStream_t callers = h->stream;
    h->stream = h->side;
    LmTail tail;
    tail.ld = c.w.ld;
    tail.rows = c.w.rows;
    tail.ctl = ctl;
    tail.handle = handle;
    c.tail = &tail;
    c.tail_done = false;
    int rc = enqueue_lm_iteration(c, 0, 1, lm, -1, opt);
    c.tail = nullptr;
    if (rc == MPBA_OK) {
      if (!c.tail_done) {  // beyond the fused image tail's size: the controller is a kernel of its own
        launch_lm_control(c.w.ld, c.w.rows, c.w.scal, ctl, handle, 1, h->stream);
        h->launch_count += 1;
      }
      launch_lm_accept(c.w.ld, cp, h->stream);
      h->launch_count += 1;
    }
    h->stream = callers;
    const cudaError_t ce = cudaStreamEndCapture(h->side, nullptr);  // always: the stream must leave capture mode
    h->lm_body_launches = h->launch_count - lc0;
    for (int k = 0; k < MPBA_NUM_KERNELS; ++k) h->lm_body_slot[k] = h->launches[k] - slot0[k];
    h->launch_count = lc0;
    memcpy(h->launches, slot0, sizeof(slot0));
    if (rc) return rc;
    if (ce != cudaSuccess) return fail(h, MPBA_E_CUDA, std::string("LM graph capture: ") + cudaGetErrorString(ce));
    CU(cudaGraphInstantiate(&h->lm_exec, h->lm_graph, 0));
    h->lm_signature = sig;
  }
  CU(cudaGraphLaunch(h->lm_exec, h->stream));
  CU(cudaMemcpyAsync(h->h_lm, c.w.ld, LM_DEV_DOUBLES * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
  CU(cudaStreamSynchronize(h->stream));
  CU(cudaGetLastError());
  const double* ld = h->h_lm;
  out.termination = (int)ld[LD_TERMINATION];
  out.msg = (int)ld[LD_MSG];
  out.iteration = (int)ld[LD_ITERATION];
  out.n_ok = (int)ld[LD_N_OK];
  out.n_bad = (int)ld[LD_N_BAD];
  out.nrows = (int)ld[LD_NROWS];
  out.total_pcg = (int64_t)ld[LD_TOTAL_PCG];
  out.minimum_cost = ld[LD_MIN_COST];
  // the body ran once per LM iteration (a gradient-tolerance exit ends inside an iteration that is not counted)
  const int64_t bodies = out.iteration + ((out.msg == LM_MSG_GRADIENT) ? 1 : 0);
  h->launch_count += h->lm_body_launches * bodies;
  for (int k = 0; k < MPBA_NUM_KERNELS; ++k) h->launches[k] += h->lm_body_slot[k] * bodies;
  if (out.termination < 0 || out.nrows < 1 || out.nrows > LM_DEVICE_MAX_ITERATIONS + 2)
    return fail(h, MPBA_E_CUDA, "device LM loop ended without a termination state");
  return MPBA_OK;
}

}  // namespace

// =================================================================================================
extern "C" {

int mpba_version(void) { return MPBA_VERSION; }

const char* mpba_kernel_name(int slot) { return (slot >= 0 && slot < MPBA_NUM_KERNELS) ? kKernelNames[slot] : ""; }

void mpba_default_options(mpba_options* o) {
  if (!o) return;
  memset(o, 0, sizeof(*o));
  o->max_num_iterations = 50;
  o->max_num_consecutive_invalid_steps = 5;
  o->jacobi_scaling = 1;
  o->pcg_max_iterations = 2000;
  o->initial_trust_region_radius = 1e4;
  o->max_trust_region_radius = 1e16;
  o->min_trust_region_radius = 1e-32;
  o->min_relative_decrease = 1e-3;
  o->min_lm_diagonal = 1e-6;
  o->max_lm_diagonal = 1e32;
  o->function_tolerance = 1e-6;
  o->gradient_tolerance = 1e-10;
  o->parameter_tolerance = 1e-8;
  o->pcg_tolerance = 1e-10;
  o->profile = 0;
  o->verbose = 0;
  o->pcg_kernel = 0;
  o->pcg_coarse = 0;
  o->behind_camera = 0;
  o->coarse_lag = 1;
  o->lm_loop = 0;
  o->reserved0 = 0;
}

int mpba_create(mpba_handle** out, int device, void* cuda_stream) {
  if (!out) return MPBA_E_INVALID;
  *out = nullptr;
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n <= 0 || device < 0 || device >= n) return MPBA_E_CUDA;  // no CPU fallback
  mpba_handle* h = new mpba_handle();
  h->device = device;
  h->stream = static_cast<cudaStream_t>(cuda_stream);
  if (cudaSetDevice(device) != cudaSuccess || cudaMallocHost(&h->h_scal, S_COUNT * sizeof(double)) != cudaSuccess ||
      cudaMallocHost(&h->h_lm, LM_DEV_DOUBLES * sizeof(double)) != cudaSuccess ||
      cudaMallocHost(&h->plan_hash_seen_pinned, 2 * sizeof(unsigned long long)) != cudaSuccess ||
      cudaEventCreate(&h->ev0) != cudaSuccess || cudaEventCreate(&h->ev1) != cudaSuccess) {
    delete h;
    return MPBA_E_CUDA;
  }
  // side stream of the lagged coarse setup; without it (creation failed) the setup stays on the solve stream
  if (cudaStreamCreateWithFlags(&h->side, cudaStreamNonBlocking) != cudaSuccess) h->side = nullptr;
  if (h->side) {
    bool ok = cudaEventCreateWithFlags(&h->ev_s_ready, cudaEventDisableTiming) == cudaSuccess &&
              cudaEventCreateWithFlags(&h->ev_asm_done, cudaEventDisableTiming) == cudaSuccess &&
              cudaEventCreateWithFlags(&h->ev_inv_done[0], cudaEventDisableTiming) == cudaSuccess &&
              cudaEventCreateWithFlags(&h->ev_inv_done[1], cudaEventDisableTiming) == cudaSuccess;
    if (!ok) {
      cudaStreamDestroy(h->side);
      h->side = nullptr;
    }
  }
  cudaGetLastError();
  *out = h;
  return MPBA_OK;
}

void mpba_destroy(mpba_handle* h) {
  if (!h) return;
  cudaSetDevice(h->device);
  for (auto& t : h->pool) {
    cudaEventDestroy(t.a);
    cudaEventDestroy(t.b);
  }
  if (h->ev0) cudaEventDestroy(h->ev0);
  if (h->ev1) cudaEventDestroy(h->ev1);
  if (h->side) {
    cudaStreamSynchronize(h->side);
    cudaStreamDestroy(h->side);
  }
  for (cudaEvent_t e : {h->ev_s_ready, h->ev_asm_done, h->ev_inv_done[0], h->ev_inv_done[1]})
    if (e) cudaEventDestroy(e);
  if (h->lm_exec) cudaGraphExecDestroy(h->lm_exec);
  if (h->lm_graph) cudaGraphDestroy(h->lm_graph);
  if (h->h_scal) cudaFreeHost(h->h_scal);
  if (h->h_lm) cudaFreeHost(h->h_lm);
  if (h->plan_hash_seen_pinned) cudaFreeHost(h->plan_hash_seen_pinned);
  if (h->h_stage) cudaFreeHost(h->h_stage);
  if (h->comm) nccl_api().CommDestroy(h->comm);
  if (h->d_agree) cudaFree(h->d_agree);
  delete h;
}

const char* mpba_last_error(const mpba_handle* h) { return h ? h->err.c_str() : "invalid handle"; }

int mpba_pack_tracks(const int32_t* pt_ptr, int32_t pt_begin, int32_t pt_end, int32_t* first, int32_t* num_warps) {
  if (!pt_ptr || !first || !num_warps || pt_begin < 0 || pt_end < pt_begin) return MPBA_E_INVALID;
  int32_t nw = 0;
  int32_t t = pt_begin;
  while (t < pt_end) {
    first[nw++] = t;
    const int64_t s0 = pt_ptr[t];
    int32_t e = t + 1;  // a track longer than 32 observations is a group of its own
    while (e < pt_end && e - t < 32 && (int64_t)pt_ptr[e + 1] - s0 <= 32) ++e;
    if ((int64_t)pt_ptr[t + 1] - s0 > 32) e = t + 1;
    t = e;
  }
  first[nw] = pt_end;
  *num_warps = nw;
  return MPBA_OK;
}

int mpba_workspace_bytes(const mpba_problem* prob, size_t* bytes) {
  if (!prob || !bytes) return MPBA_E_INVALID;
  Workspace w;
  carve(prob, nullptr, nullptr, w);
  *bytes = w.bytes;
  return MPBA_OK;
}

int mpba_workspace_bytes_sharded(const mpba_problem* prob, const mpba_shard* shard, size_t* bytes) {
  if (!prob || !bytes || !shard) return MPBA_E_INVALID;
  Workspace w;
  carve(prob, shard, nullptr, w);
  *bytes = w.bytes;
  return MPBA_OK;
}

int mpba_comm_unique_id(void* id, const char* nccl_library) {
  if (!id) return MPBA_E_INVALID;
  static_assert(sizeof(ncclUniqueId) <= MPBA_UNIQUE_ID_BYTES, "ncclUniqueId larger than MPBA_UNIQUE_ID_BYTES");
  NcclApi& nc = nccl_api();
  if (!nc.load(nccl_library).empty()) return MPBA_E_NCCL;
  ncclUniqueId uid;
  if (nc.GetUniqueId(&uid) != ncclSuccess) return MPBA_E_NCCL;
  memset(id, 0, MPBA_UNIQUE_ID_BYTES);
  memcpy(id, &uid, sizeof(uid));
  return MPBA_OK;
}

int mpba_comm_init(mpba_handle* h, int world, int rank, const void* id, const char* nccl_library) {
  if (!h) return MPBA_E_INVALID;
  if (!id || world < 1 || rank < 0 || rank >= world) return fail(h, MPBA_E_INVALID, "mpba_comm_init: bad world / rank / id");
  NcclApi& nc = nccl_api();
  const std::string err = nc.load(nccl_library);
  if (!err.empty()) return fail(h, MPBA_E_NCCL, err);
  CU(cudaSetDevice(h->device));
  if (h->comm) {
    nc.CommDestroy(h->comm);
    h->comm = nullptr;
  }
  ncclUniqueId uid;
  memcpy(&uid, id, sizeof(uid));
  NC(nc.CommInitRank(&h->comm, world, uid, rank));
  h->world = world;
  h->rank = rank;
  if (!h->d_agree) CU(cudaMalloc(&h->d_agree, sizeof(int)));
  return MPBA_OK;
}

// Sharded calls: a rank whose local checks failed (bad arguments, workspace too small) must not leave the others
// blocked in their first collective.  Every rank contributes its local code; the minimum (codes are <= 0) is what all
// of them return.  One 4-byte all-reduce and one synchronisation per sharded call.
static int agree_on_rc(mpba_handle* h, int world, int rc_local) {
  if (!h || !h->comm || world <= 1 || !h->d_agree) return rc_local;
  int v = rc_local;
  if (cudaSetDevice(h->device) != cudaSuccess) return rc_local;
  if (cudaMemcpyAsync(h->d_agree, &v, sizeof(int), cudaMemcpyHostToDevice, h->stream) != cudaSuccess) return rc_local;
  if (nccl_api().AllReduce(h->d_agree, h->d_agree, 1, ncclInt, ncclMin, h->comm, h->stream) != ncclSuccess) return rc_local;
  if (cudaMemcpyAsync(&v, h->d_agree, sizeof(int), cudaMemcpyDeviceToHost, h->stream) != cudaSuccess) return rc_local;
  if (cudaStreamSynchronize(h->stream) != cudaSuccess) return rc_local;
  if (v != MPBA_OK && rc_local == MPBA_OK)
    return fail(h, v, "sharded call abandoned: another rank failed its local checks with code " + std::to_string(v));
  return rc_local != MPBA_OK ? rc_local : v;
}

int mpba_kernel_times(const mpba_handle* h, double* ms, int64_t* launches) {
  if (!h) return MPBA_E_INVALID;
  for (int i = 0; i < MPBA_NUM_KERNELS; ++i) {
    if (ms) ms[i] = h->ms[i];
    if (launches) launches[i] = h->launches[i];
  }
  return MPBA_OK;
}

static void reset_profile(mpba_handle* h, int profile) {
  h->profile = profile != 0;
  for (int i = 0; i < MPBA_NUM_KERNELS; ++i) {
    h->ms[i] = 0.0;
    h->launches[i] = 0;
  }
  h->used = 0;
}

int mpba_evaluate(mpba_handle* h, const mpba_problem* prob, void* workspace, size_t workspace_bytes, double* r_reproj,
                  double* r_depth, double* cost) {
  Ctx c;
  int rc = begin(h, prob, workspace, workspace_bytes, c);
  if (rc) return rc;
  reset_profile(h, 0);
  {
    Timer t(h, KS_MISC);
    launch_normalize_poses(c.pv.C, prob->img_pose, prob->img_shift_scale, c.w.pose[0], h->stream);
  }
  rc = stage_points(c, prob, c.w.X[0]);
  if (rc) return rc;
  if (r_reproj || r_depth) {
    Timer t(h, KS_EVALUATE);
    launch_evaluate(c.pv, c.w.pose[0], c.w.X[0], r_reproj, r_depth, h->stream);
  }
  rc = enqueue_image_pass(c, c.w.pose[0], c.w.X[0], 0, 0, 1);
  if (rc) return rc;
  rc = fetch_scalars(c);
  if (rc) return rc;
  if (cost) {
    cost[0] = 0.5 * h->h_scal[S_COST_FREE];
    cost[1] = 0.5 * h->h_scal[S_COST_FIXED];
  }
  return MPBA_OK;
}

int mpba_filter_tracks(mpba_handle* h, const mpba_problem* prob, void* workspace, size_t workspace_bytes,
                       double max_reproj_error, double min_tri_angle_deg, const uint8_t* pt_select, uint8_t* obs_keep,
                       uint8_t* pt_keep, double* pt_error, int64_t* num_filtered) {
  Ctx c;
  int rc = begin(h, prob, workspace, workspace_bytes, c);
  if (rc) return rc;
  if (!obs_keep || !pt_keep || !pt_error)
    return fail(h, MPBA_E_INVALID, "mpba_filter_tracks: obs_keep, pt_keep and pt_error are required");
  reset_profile(h, 0);
  {
    Timer t(h, KS_MISC);
    launch_normalize_poses(c.pv.C, prob->img_pose, prob->img_shift_scale, c.w.pose[0], h->stream);
  }
  rc = stage_points(c, prob, c.w.X[0]);
  if (rc) return rc;
  CU(cudaMemsetAsync(c.w.scal, 0, S_COUNT * sizeof(double), h->stream));
  {
    Timer t(h, KS_EVALUATE);
    const double deg = 3.14159265358979323846 / 180.0;
    launch_filter_tracks(c.pv, c.w.pose[0], c.w.X[0], max_reproj_error * max_reproj_error, min_tri_angle_deg * deg,
                         pt_select, obs_keep, pt_keep, pt_error, c.w.scal, h->stream);
  }
  rc = fetch_scalars(c);
  if (rc) return rc;
  if (num_filtered) *num_filtered = (int64_t)h->h_scal[S_FILTERED];
  return MPBA_OK;
}

int mpba_linearize(mpba_handle* h, const mpba_problem* prob, void* workspace, size_t workspace_bytes, double* Hcc,
                   double* gc, double* Hpp, double* gp, double* cost) {
  Ctx c;
  int rc = begin(h, prob, workspace, workspace_bytes, c);
  if (rc) return rc;
  reset_profile(h, 0);
  mpba_options opt;
  mpba_default_options(&opt);
  launch_normalize_poses(c.pv.C, prob->img_pose, prob->img_shift_scale, c.w.pose[0], h->stream);
  rc = stage_points(c, prob, c.w.X[0]);
  if (rc) return rc;
  rc = enqueue_image_pass(c, c.w.pose[0], c.w.X[0], 0, 1, 1);
  if (rc) return rc;
  const LmDiag lm = make_lm(opt.initial_trust_region_radius, opt.min_lm_diagonal, opt.max_lm_diagonal);
  rc = enqueue_points_and_system(c, c.w.pose[0], c.w.X[0], 0, lm, 1, &opt, false);
  if (rc) return rc;
  if (Hcc) launch_expand_hcc(c.pv.C, c.w.Hc[0], Hcc, h->stream);
  if (gc) CU(cudaMemcpyAsync(gc, c.w.gc[0], (size_t)c.pv.C * 6 * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
  if (Hpp) CU(cudaMemcpyAsync(Hpp, c.w.Hpp, (size_t)c.pv.P * 6 * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
  if (gp) CU(cudaMemcpyAsync(gp, c.w.gp, (size_t)c.pv.P * 3 * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
  rc = fetch_scalars(c);
  if (rc) return rc;
  if (cost) {
    cost[0] = 0.5 * h->h_scal[S_COST_FREE];
    cost[1] = 0.5 * h->h_scal[S_COST_FIXED];
  }
  return MPBA_OK;
}

int mpba_point_covariances(mpba_handle* h, const mpba_problem* prob, void* workspace, size_t workspace_bytes,
                           double* cov) {
  Ctx c;
  int rc = begin(h, prob, workspace, workspace_bytes, c);
  if (rc) return rc;
  if (!cov) return fail(h, MPBA_E_INVALID, "cov is NULL");
  c.want_hinv = true;
  reset_profile(h, 0);
  mpba_options opt;
  mpba_default_options(&opt);
  opt.jacobi_scaling = 0;
  launch_normalize_poses(c.pv.C, prob->img_pose, prob->img_shift_scale, c.w.pose[0], h->stream);
  rc = stage_points(c, prob, c.w.X[0]);
  if (rc) return rc;
  // undamped: an infinite trust-region radius makes the LM diagonal vanish
  const LmDiag lm = make_lm(INFINITY, opt.min_lm_diagonal, opt.max_lm_diagonal);
  rc = enqueue_points_and_system(c, c.w.pose[0], c.w.X[0], 0, lm, 1, &opt, false);
  if (rc) return rc;
  CU(cudaMemcpyAsync(cov, c.w.Hinv, (size_t)c.pv.P * 6 * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
  CU(cudaStreamSynchronize(h->stream));
  CU(cudaGetLastError());
  return MPBA_OK;
}

int mpba_reduced_system(mpba_handle* h, const mpba_problem* prob, void* workspace, size_t workspace_bytes,
                        const mpba_options* opt_in, double radius, double* S, double* rhs, double* y,
                        int32_t* pcg_iterations) {
  Ctx c;
  int rc = begin(h, prob, workspace, workspace_bytes, c);
  if (rc) return rc;
  reset_profile(h, 0);
  mpba_options opt;
  if (opt_in) opt = *opt_in; else mpba_default_options(&opt);
  rc = build_pcg_cluster_plan(c, prob, opt.pcg_kernel, opt.pcg_coarse);
  if (rc) return rc;
  rc = build_grid_coarse_plan(c, prob, opt.pcg_coarse);
  if (rc) return rc;
  launch_normalize_poses(c.pv.C, prob->img_pose, prob->img_shift_scale, c.w.pose[0], h->stream);
  rc = stage_points(c, prob, c.w.X[0]);
  if (rc) return rc;
  rc = enqueue_image_pass(c, c.w.pose[0], c.w.X[0], 0, 1, opt.jacobi_scaling);
  if (rc) return rc;
  const LmDiag lm = make_lm(radius, opt.min_lm_diagonal, opt.max_lm_diagonal);
  rc = enqueue_points_and_system(c, c.w.pose[0], c.w.X[0], 0, lm, 1, &opt, true);
  if (rc) return rc;
  if (c.structure_only) return fail(h, MPBA_E_INVALID, "no reduced camera system: all poses are constant");
  if (S) CU(cudaMemcpyAsync(S, c.w.S, (size_t)c.pv.B * 36 * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
  if (rhs) CU(cudaMemcpyAsync(rhs, c.w.rhs, (size_t)c.pv.C * 6 * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
  if (y) CU(cudaMemcpyAsync(y, c.w.y, (size_t)c.pv.C * 6 * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
  rc = fetch_scalars(c);
  if (rc) return rc;
  if (pcg_iterations) *pcg_iterations = (int32_t)h->h_scal[S_PCG_ITERS];
  return MPBA_OK;
}

static int solve_impl(mpba_handle* h, const mpba_problem* prob, const mpba_shard* shard, void* workspace,
                      size_t workspace_bytes, const mpba_options* opt_in, mpba_summary* sum) {
  NvtxRange nvtx_solve(shard ? "mpba_solve_sharded" : "mpba_solve");
  Ctx c;
  int rc = begin(h, prob, workspace, workspace_bytes, c, shard);
  if (shard) rc = agree_on_rc(h, shard->world, rc);
  if (rc) return rc;
  if (!sum) return fail(h, MPBA_E_INVALID, "summary is NULL");
  mpba_options opt;
  if (opt_in) opt = *opt_in; else mpba_default_options(&opt);
  reset_profile(h, opt.profile);
  c.pv.loss.behind_strict = opt.behind_camera == 1 ? 1 : 0;
  // the lag pays where the dense inverse is the multi-CTA one (>= 600 images); small systems keep the synchronous
  // setup (a ~20 us single-CTA inverse) and with it the trajectory of the exact preconditioner
  c.coarse_lag = (opt.coarse_lag != 0 && prob->num_images >= 600) ? 1 : 0;
  c.defer_gmax = true;
  CU(cudaMemsetAsync(c.w.pts_stage + 4 * 64, 0, 2 * sizeof(double), h->stream));  // arrival ticket of k_pt_scalars_step
  struct SideDrain {  // whatever the exit path: nothing of this call is left running on the side stream
    mpba_handle* h;
    ~SideDrain() { if (h->side) cudaStreamSynchronize(h->side); }
  } side_drain{h};
  rc = build_pcg_cluster_plan(c, prob, opt.pcg_kernel, opt.pcg_coarse);
  if (rc) return rc;
  rc = build_grid_coarse_plan(c, prob, opt.pcg_coarse);
  if (rc) return rc;
  sum->num_behind_camera = 0;
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
  auto set_last_gmax = [&](double g) {
    last.gradient_max_norm = g;
    if (nrows - 1 < cap && nrows > 0) its[nrows - 1].gradient_max_norm = g;
  };

  CU(cudaEventRecord(h->ev0, h->stream));
  int cur = 0;
  launch_normalize_poses(c.pv.C, prob->img_pose, prob->img_shift_scale, c.w.pose[cur], h->stream);
  h->launch_count += 1;
  rc = stage_points(c, prob, c.w.X[cur]);
  if (rc) return rc;
  // points without observations are never written by the back-substitution: both buffers carry them
  if (c.pv.P > 0)
    CU(cudaMemcpyAsync(c.w.X[cur ^ 1], c.w.X[cur], (size_t)c.pv.P * 4 * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));

  // ---- iteration zero
  rc = enqueue_image_pass(c, c.w.pose[cur], c.w.X[cur], cur, 1, opt.jacobi_scaling);
  if (rc) return rc;
  rc = fetch_scalars(c);
  if (rc) return rc;
  double x_cost = 0.5 * h->h_scal[S_COST_FREE];
  const double initial_cost_free = x_cost;
  const double fixed_cost = 0.5 * h->h_scal[S_COST_FIXED];
  double gmax_c = h->h_scal[S_GMAX_C];
  if (!isfinite(x_cost)) {
    // Ceres: "Residual and Jacobian evaluation failed" at iteration 0 -> FAILURE, user parameters untouched, and
    // pyceres.solve does not raise (bundle_adjustment.py:285-293 only logs the summary).  Same here; the loud
    // return is kept for behind_camera = 1.
    snprintf(sum->message, sizeof(sum->message), "Initial cost is not finite.");
    sum->termination = MPBA_FAILURE;
    sum->initial_cost = sum->final_cost = x_cost;
    sum->fixed_cost = fixed_cost;
    sum->num_iterations = 0;
    sum->num_successful_steps = sum->num_unsuccessful_steps = 0;
    sum->total_pcg_iterations = 0;
    sum->kernel_launches = h->launch_count;
    sum->solve_ms = 0.0;
    if (nrows == 0) {
      mpba_iteration it0;
      memset(&it0, 0, sizeof(it0));
      it0.cost = x_cost;
      push(it0);
    }
    if (opt.behind_camera == 1)
      return fail(h, MPBA_E_NUMERIC, "initial cost is not finite (point behind a camera or bad input)");
    h->err = "initial cost is not finite: solve skipped, state untouched";
    return MPBA_OK;
  }
  double radius = opt.initial_trust_region_radius, decrease_factor = 2.0;
  double minimum_cost = x_cost;
  int n_ok = 0, n_bad = 0, invalid_run = 0;
  int64_t total_pcg = 0;
  int termination = -1;
  const char* msg = "";
  bool need_gmax = true, first = true;
  {
    mpba_iteration it0;
    memset(&it0, 0, sizeof(it0));
    it0.cost = x_cost + fixed_cost;
    it0.trust_region_radius = radius;
    push(it0);
  }
  int iteration = 0;
  if (iteration >= opt.max_num_iterations) {
    termination = MPBA_NO_CONVERGENCE;
    msg = "Maximum number of iterations reached.";
  }

  const bool device_lm = termination < 0 && opt.lm_loop != 1 && device_lm_legal(c, opt) &&
                         (opt.lm_loop == 2 || prob->num_images < 600);
  if (device_lm) {
    // the state sits in buffer set 0 (cur == 0 here); the candidate of every iteration in set 1
    DeviceLmResult dl;
    rc = run_device_lm(c, prob, opt, radius, x_cost, fixed_cost, gmax_c, dl);
    if (rc) return rc;
    const mpba_iteration* rows = reinterpret_cast<const mpba_iteration*>(h->h_lm + LD_COUNT);
    nrows = 0;
    for (int k = 0; k < dl.nrows; ++k) push(rows[k]);
    termination = dl.termination;
    msg = lm_message(dl.msg);
    iteration = dl.iteration;
    n_ok = dl.n_ok;
    n_bad = dl.n_bad;
    total_pcg = dl.total_pcg;
    minimum_cost = dl.minimum_cost;
  }

  while (termination < 0) {
    NvtxRange nvtx_iteration("lm_iteration");
    const int cand = cur ^ 1;
    const LmDiag lm = make_lm(radius, opt.min_lm_diagonal, opt.max_lm_diagonal);
    rc = enqueue_lm_iteration(c, cur, cand, lm, first ? 1 : 0, opt);
    if (rc) return rc;
    first = false;
    rc = fetch_scalars(c);
    if (rc) return rc;
    const double* s = h->h_scal;

    if (need_gmax) {
      const double g = fmax(gmax_c, s[S_GMAX_P]);
      set_last_gmax(g);
      need_gmax = false;
      if (g <= opt.gradient_tolerance) {
        termination = MPBA_CONVERGENCE;
        msg = "Gradient tolerance reached.";
        break;
      }
    }

    mpba_iteration it;
    memset(&it, 0, sizeof(it));
    it.cost = last.cost;
    it.gradient_max_norm = last.gradient_max_norm;
    it.pcg_iterations = (int32_t)s[S_PCG_ITERS];
    total_pcg += it.pcg_iterations;
    ++iteration;
    const double mcc = s[S_MCC] + s[S_MCC_C];  // points + images: (g'y + y'Dy) / 2
    bool ok = s[S_PCG_FLAG] != 2.0 && s[S_BAD_POINTS] == 0.0 && isfinite(mcc) && isfinite(s[S_STEP2_P]) &&
              isfinite(s[S_STEP2_C]) && mcc > 0.0;
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
      double cand_cost = 0.5 * s[S_COST_FREE];
      if (!isfinite(cand_cost)) cand_cost = DBL_MAX;
      it.step_norm = sqrt(s[S_STEP2_P] + s[S_STEP2_C]);
      const double x_norm = sqrt(s[S_XNORM2_P] + s[S_XNORM2_C]);
      if (it.step_norm <= opt.parameter_tolerance * (x_norm + opt.parameter_tolerance)) {
        termination = MPBA_CONVERGENCE;
        msg = "Parameter tolerance reached.";
        push(it);
        break;
      }
      it.cost_change = x_cost - cand_cost;
      if (fabs(it.cost_change) <= opt.function_tolerance * x_cost) {
        termination = MPBA_CONVERGENCE;
        msg = "Function tolerance reached.";
        push(it);
        break;
      }
      it.relative_decrease = it.cost_change / mcc;
      if (it.relative_decrease > opt.min_relative_decrease) {
        cur = cand;
        x_cost = cand_cost;
        gmax_c = s[S_GMAX_C];
        need_gmax = true;
        it.cost = x_cost + fixed_cost;
        it.step_is_successful = 1;
        const double t = 2.0 * it.relative_decrease - 1.0;
        radius = fmin(opt.max_trust_region_radius, radius / fmax(1.0 / 3.0, 1.0 - t * t * t));
        decrease_factor = 2.0;
        ++n_ok;
        if (x_cost < minimum_cost) minimum_cost = x_cost;
      } else {
        radius /= decrease_factor;
        decrease_factor *= 2.0;
        ++n_bad;
      }
      it.trust_region_radius = radius;
      push(it);
    }
    if (opt.verbose > 1) {
      fprintf(stderr, "[mpba]     pcg cycles on CTA 0: spmv %.0f reduce %.0f update %.0f (%d iterations)\n",
              s[S_PCG_CYC_SPMV], s[S_PCG_CYC_REDUCE], s[S_PCG_CYC_UPDATE], (int)s[S_PCG_ITERS]);
      if (c.cl > 0)
        fprintf(stderr, "[mpba]     cluster of %d: syncA %.0f halo %.0f products %.0f syncB %.0f rowsum %.0f syncC %.0f update %.0f\n",
                c.cl, s[16], s[17], s[18], s[19], s[20], s[21], s[22]);
    }
    if (opt.verbose)
      fprintf(stderr, "[mpba] %3d cost %.9e dcost %.3e |step| %.3e rho %.3e radius %.3e pcg %d (%s)\n", iteration,
              last.cost, last.cost_change, last.step_norm, last.relative_decrease, last.trust_region_radius,
              last.pcg_iterations, last.step_is_successful ? "ok" : (last.step_is_valid ? "rejected" : "invalid"));
    if (iteration >= opt.max_num_iterations) {
      termination = MPBA_NO_CONVERGENCE;
      msg = "Maximum number of iterations reached.";
    } else if (radius <= opt.min_trust_region_radius) {
      termination = MPBA_CONVERGENCE;
      msg = "Minimum trust region radius reached.";
    }
  }

  // monotonic steps: the current state is the best state visited
  CU(cudaMemcpyAsync(prob->img_pose, c.w.pose[cur], (size_t)c.pv.C * 8 * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
  if (c.pv.P > 0)
    CU(cudaMemcpyAsync(prob->pt_xyz, c.w.X[cur], (size_t)c.pv.P * 4 * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
  // observations the COLMAP >= 3.11 rule zeroed at the returned state (one launch per solve)
  CU(cudaMemsetAsync(c.w.scal, 0, sizeof(double), h->stream));
  launch_count_behind(c.pv, c.w.pose[cur], c.w.X[cur], c.w.scal, h->stream);
  h->launch_count += 1;
  if (c.sh && c.sh->world > 1) NC(nccl_api().AllReduce(c.w.scal, c.w.scal, 1, ncclDouble, ncclSum, h->comm, h->stream));
  CU(cudaMemcpyAsync(h->h_scal, c.w.scal, sizeof(double), cudaMemcpyDeviceToHost, h->stream));
  CU(cudaEventRecord(h->ev1, h->stream));
  CU(cudaStreamSynchronize(h->stream));
  if (h->side) CU(cudaStreamSynchronize(h->side));
  if (h->profile) flush_timers(h);
  float ms = 0.f;
  cudaEventElapsedTime(&ms, h->ev0, h->ev1);
  sum->num_behind_camera = (int32_t)h->h_scal[0];

  sum->initial_cost = initial_cost_free + fixed_cost;
  sum->final_cost = minimum_cost + fixed_cost;
  sum->fixed_cost = fixed_cost;
  sum->termination = termination;
  sum->num_iterations = iteration;
  sum->num_successful_steps = n_ok;
  sum->num_unsuccessful_steps = n_bad;
  sum->total_pcg_iterations = total_pcg;
  sum->kernel_launches = h->launch_count;
  sum->solve_ms = ms;
  snprintf(sum->message, sizeof(sum->message), "%s", msg);
  return MPBA_OK;
}

int mpba_solve(mpba_handle* h, const mpba_problem* prob, void* workspace, size_t workspace_bytes,
               const mpba_options* opt_in, mpba_summary* sum) {
  return solve_impl(h, prob, nullptr, workspace, workspace_bytes, opt_in, sum);
}

int mpba_solve_sharded(mpba_handle* h, const mpba_problem* prob, const mpba_shard* shard, void* workspace,
                       size_t workspace_bytes, const mpba_options* opt_in, mpba_summary* sum) {
  if (!h) return MPBA_E_INVALID;
  if (!shard) return fail(h, MPBA_E_INVALID, "shard is NULL");
  return solve_impl(h, prob, shard, workspace, workspace_bytes, opt_in, sum);
}

}  // extern "C"
