// ba_kernels.cuh -- sm_100a kernels of the bundle-adjustment hot path (declarations + launch args).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/mpba.h"
#include "ba_math.cuh"

namespace mpba {

// scalar slots (device doubles) one LM iteration produces
enum Scal {
  S_COST_FREE = 0,   // sum rho over blocks with a free parameter block (not halved)
  S_COST_FIXED = 1,  // sum rho over blocks Ceres removes from the program
  S_GMAX_C = 2,      // max |gc|
  S_GMAX_P = 3,      // max |gp|
  S_MCC = 4,         // model cost change
  S_STEP2_P = 5,     // |X - X_cand|^2 over active points
  S_XNORM2_P = 6,    // |X|^2 over active points
  S_STEP2_C = 7,     // ambient |q - q_cand|^2 + |t - t_cand|^2 over active image blocks
  S_XNORM2_C = 8,
  S_PCG_ITERS = 9,
  S_FILTERED = 9,     // mpba_filter_tracks only (no PCG runs in that call): observations removed
  S_PCG_FLAG = 10,   // 0 converged, 1 iteration cap, 2 breakdown (p'Sp <= 0 or non-finite)
  S_PCG_RES = 11,    // final sqrt(r'M^-1 r / b'M^-1 b)
  S_BAD_POINTS = 12, // number of point blocks whose damped Hpp was not positive definite
  S_PCG_CYC_SPMV = 13, S_PCG_CYC_REDUCE = 14, S_PCG_CYC_UPDATE = 15,  // SM cycles CTA 0 spent per phase
  S_PCG_CYC_DETAIL = 16,  // 7 slots: cluster PCG cycles of CTA 0: sync A, halo, products, sync B, row sums, sync C, update
  S_MCC_C = 23,      // image part of the model cost change (k_cam_update; replicated in a sharded solve)
  S_COUNT = 24
};

constexpr int MPBA_MAX_PRM = 12;
constexpr int IMG_PART = 32;  // doubles per image-chunk partial: 21 Hcc + 6 gc + cost_free + cost_fixed (+pad)
constexpr int PT_PART = 4;    // doubles per track-CTA partial
constexpr int FAC_STRIDE = 16; // doubles per Schur factor record (128 B per observation)

struct ProblemView {
  int C, K, P, N, B, U, num_chunks;
  // owned ranges of a sharded solve (mpba_shard): tracks of points outside [p0, p1) and image-order chunks of
  // images outside [i0, i1) live on other ranks.  Unsharded: [0, P) and [0, C).
  int p0, p1, i0, i1;
  // track-order warp packing (mpba_pack_tracks): warp w owns the whole tracks tw_first[w] .. tw_first[w+1]
  int num_track_warps;
  const int32_t* tw_first;
  // built on the device once per call (launch_track_setup, workspace): point of every track-order slot, and per
  // CTA of the track-order passes {first track, tracks, first slot, observations}
  const int32_t* pm_pt;
  const int4* cta_desc;
  const int4* warp_desc;  // the same per warp group
  const int32_t* img_camera;
  const uint8_t* img_free_mask;
  const double* img_shift_scale;
  const int32_t* cam_model;
  const double* cam_params;
  const uint8_t* pt_const;
  const int32_t* pt_ptr;
  const int32_t* pm_img;
  const double* pm_xy;
  const double* pm_depth;
  const int32_t* img_ptr;
  const int32_t* cm_pt;
  const double* cm_xy;
  const double* cm_depth;
  const double* pm_plane;  // [N][4] normal-prior rows (extension, mpba.h) or nullptr; selects the <PLANE> kernels
  const double* cm_plane;
  const int32_t* chunk_img;
  const int32_t* chunk_begin;
  const int32_t* chunk_end;
  const int32_t* img_chunk_ptr;
  const int32_t* s_row_ptr;
  const int32_t* s_col;
  const int32_t* s_diag;
  const int32_t* s_transpose;
  const int32_t* s_row;
  const int32_t* u_block;
  const int32_t* u_item_ptr;
  const int32_t* item_a;
  const int32_t* item_b;
  LossSpec loss;
};

// dev != nullptr (device-controlled LM loop, lm_loop.cu): {radius, 1 / radius} are read from dev[LD_RADIUS],
// dev[LD_INV_RADIUS] when the kernel starts, and "compute_scaling < 0" means dev[LD_FIRST] != 0
struct LmDiag {
  double radius, lo, hi, inv_radius;
  const double* dev;
};
inline LmDiag make_lm(double radius, double lo, double hi) { return LmDiag{radius, lo, hi, 1.0 / radius, nullptr}; }

// ---- device-controlled LM loop (lm_loop.cu) ---------------------------------------------------
enum LmDevSlot {  // doubles in device memory: the LM state between two iterations of the graph's WHILE body
  LD_RADIUS = 0, LD_INV_RADIUS, LD_FIRST, LD_DECREASE, LD_X_COST, LD_MIN_COST, LD_FIXED_COST, LD_GMAX_C, LD_NEED_GMAX,
  LD_ITERATION, LD_N_OK, LD_N_BAD, LD_INVALID_RUN, LD_TOTAL_PCG, LD_TERMINATION, LD_MSG, LD_ACCEPT, LD_NROWS,
  LD_COUNT = 24
};
enum LmMsg { LM_MSG_NONE = 0, LM_MSG_GRADIENT, LM_MSG_PARAMETER, LM_MSG_FUNCTION, LM_MSG_INVALID, LM_MSG_MAX_ITERATIONS, LM_MSG_MIN_RADIUS };
constexpr int LM_DEVICE_MAX_ITERATIONS = 254;  // rows reserved in the workspace: initial evaluation + one per iteration
struct LmControl {
  double function_tolerance, gradient_tolerance, parameter_tolerance, min_relative_decrease;
  double max_trust_region_radius, min_trust_region_radius;
  int max_num_iterations, max_consecutive_invalid;
};
struct LmAcceptCopy {
  const double* src[4];
  double* dst[4];
  size_t n[4];
};
struct LmTail {  // the controller as the last step of the fused image tail kernel (launch_img_tail)
  double* ld;
  mpba_iteration* rows;
  LmControl ctl;
  cudaGraphConditionalHandle handle;
};
void launch_lm_init(double* ld, mpba_iteration* rows, double radius, double x_cost, double fixed_cost, double gmax_c,
                    double* scal, cudaStream_t st);
void launch_lm_control(double* ld, mpba_iteration* rows, double* scal, const LmControl& o,
                       cudaGraphConditionalHandle handle, int use_handle, cudaStream_t st);
void launch_lm_accept(const double* ld, const LmAcceptCopy& cp, cudaStream_t st);
const char* lm_message(int code);

// ---- image order: Hcc, gc, cost ------------------------------------------------------------
void launch_img_chunks(const ProblemView& pv, const double* pose, const double* X, double* chunk_part,
                       cudaStream_t st);
void launch_img_finalize(const ProblemView& pv, const double* chunk_part, double* Hc, double* gc, double* img_part,
                         cudaStream_t st);
// cost / gradient-max scalars and (compute_scaling) the Jacobi column scales of the image blocks; runs after the
// image blocks are complete (in a sharded solve: after their all-reduce)
void launch_img_scalars(const ProblemView& pv, const double* img_part, const double* Hc, const double* gc,
                        double* c_img, int compute_scaling, int jacobi_scaling, double* scal, cudaStream_t st);

// Small bundles (<= IMG_TAIL_MAX_IMAGES images, single GPU): launch_img_finalize + launch_img_scalars as ONE CTA, same
// sums in the same order; tail != nullptr appends the LM decision (lm_control.cuh) and sets the WHILE condition.
constexpr int IMG_TAIL_MAX_IMAGES = 128;
// pts != nullptr: the reduction of the step's point scalars (launch_pt_scalars_step's work) and of the camera-step
// slices is done by the same CTA (bundles whose track-order passes have at most IMG_TAIL_MAX_PT_ROWS CTAs)
struct PtScalarsArgs {
  const double* pt_part;
  int nblk;  // < 0: nothing to reduce
  int with_gmax;
  const double* cam_stage;
  int cam_rows;
};
constexpr int IMG_TAIL_MAX_PT_ROWS = 8192;
void launch_img_tail(const ProblemView& pv, const double* chunk_part, double* Hc, double* gc, double* img_part,
                     double* c_img, int compute_scaling, int jacobi_scaling, double* scal, const LmTail* tail,
                     const PtScalarsArgs* pts, cudaStream_t st);

// ---- track order: Hpp, gp, point-block Cholesky, Schur factors ------------------------------------
void launch_points(const ProblemView& pv, const double* pose, const double* X, LmDiag lm, double* c_pt,
                   int compute_scaling, int jacobi_scaling, double* Hpp, double* gp, double* Hinv, double* Linv,
                   double* fac, double* pt_part, double* scal, int write_factors, cudaStream_t st);
int points_grid(int num_owned_points);  // k_evaluate: one thread per track
int track_warp_grid(int num_track_warps);
void launch_track_setup(const ProblemView& pv, int32_t* pm_pt, int4* cta_desc, int4* warp_desc, cudaStream_t st);  // k_points / k_back_substitute: CTAs (= rows of pt_part)
void launch_pt_scalars_gmax(const double* pt_part, int nblk, double* scal, cudaStream_t st);

// ---- reduced camera system ----------------------------------------------------------------------
// partial != 0 (sharded solve): S / rhs receive only this rank's -sum over its own tracks; after the all-reduce
// launch_schur_finish adds Hcc + D, gc and the identity rows of fixed tangent coordinates.
void launch_schur_blocks(const ProblemView& pv, const double* fac, const double* Hc, const double* gc,
                         const double* c_img, LmDiag lm, double* S, double* rhs, int partial, double* Minv,
                         cudaStream_t st);  // Minv != nullptr (and !partial): block-Jacobi inverses written by the diagonal CTAs
void launch_schur_finish(const ProblemView& pv, const double* Hc, const double* gc, const double* c_img, LmDiag lm,
                         double* S, double* rhs, double* Minv, cudaStream_t st);
// scal_all [world][S_COUNT] (all-gathered per-rank scalar blocks) -> scal: sums / max in rank order
void launch_combine_scalars(const double* scal_all, int world, double* scal, cudaStream_t st);
// rows [world][S_COUNT]: this rank's scalar block into row `rank`, zeros elsewhere (gathered by a SUM all-reduce)
void launch_stage_scalars(const double* scal, double* rows, int rank, int world, cudaStream_t st);
struct PcgPlan {
  int grid, rows_per_cta, use_smem, max_local_blocks;
  int halo_cap, halo_words;  // halo staging of the grid PCG's product (0: off)
  size_t smem_bytes;
};
PcgPlan pcg_plan(int device, int C, int max_row_blocks);
// two-level preconditioner of the grid kernel: G groups of consecutive CTAs (G == 0: block-Jacobi only)
struct PcgGridCoarse {
  int G = 0;
  const int32_t* grp_row_ptr = nullptr;  // [G+1], device
  const double* Zc = nullptr;
  const double* Ainv = nullptr;
  double* tpart = nullptr;  // [grid][8]
};
// returns cudaError from the cooperative launch
cudaError_t launch_pcg(const ProblemView& pv, const PcgPlan& pl, const double* S, const double* Minv, const double* rhs,
                       double* y, double* vec_z, double* vec_p0, double* vec_p1, double* partial, unsigned* bar,
                       double* scal, double tol, int max_iter, const PcgGridCoarse& co, cudaStream_t st);

// cluster-resident PCG (pcg_cluster.cu): one thread-block cluster of `cl` CTAs, upper blocks of S in
// registers, exchanges through distributed shared memory.  The plan arrays are built on the host from
// the block-CSR structure (ba_solver.cu: build_pcg_cluster_plan) and live in the workspace.
#ifndef MPBA_PCGC_THREADS
#define MPBA_PCGC_THREADS 512  // 128 registers: 88 B of spills outside the iteration loop; measured at cfg3 (r2): 696 us per PCG against 773 us with 384 threads, whose surplus blocks of S sat in shared memory
#endif
constexpr int PCGC_THREADS = MPBA_PCGC_THREADS;
// shared-memory carve of k_pcg_cluster as 32-bit offsets from the dynamic base (doubles; the last three in
// ints): kept in the kernel argument block = constant bank, so they cost no registers next to the 6x6
// block of S every thread holds
struct PcgClusterCarve {
  int dep, x, r, u, p, s, w, minv, sdiag, halo, extra, parts, red;
  int zloc, ainv, rcall, ec, glrp;  // (glrp: int view, local row pointers of the owned groups)
  int zs, zr;  // Z^T s and Z^T r of ALL groups, kept by recurrence on every CTA (pcg_cluster.cu)
  int sc;      // loop-carried scalars (gamma_0, tol^2 gamma_0, 1 / gamma_prev, 1 / alpha_prev, last gamma)
  int stail;   // [2][threads] double2: entries 32..35 of every thread's register block of S
  // two-level preconditioner: Z rows of the owned images, owned rows of A_c^-1, Z^T r, e_c
  int rp;     // row pointers of the owned rows, relative to the first owned full block
  int hdesc;  // halo source: owner CTA << 20 | local row at the owner
  int xmeta;  // spilled blocks: rowl, halo, depA, tcta, tidx
};
struct PcgClusterArgs {
  int C;
  PcgClusterCarve m;
  int max_rows, max_full, max_halo, max_extra;  // per-CTA maxima: uniform shared-memory layout across the cluster
  // coarse space (coarse.cu): G groups of consecutive images, gpc per CTA (G = cl * gpc), nc = 7 G; G == 0: off
  int G, gpc, nc;
  const int32_t* grp_row_ptr;   // [G+1] first image of every group
  const double* Zc;             // [C][6][7] gauge modes
  const double* Ainv;           // [nc][nc] (Z^T S Z)^-1
  const int32_t* row_ptr;       // [C+1] block-CSR row pointers of S (problem)
  const int32_t* s_diag;        // [C]   block index of the diagonal block of every row (problem)
  const int32_t* cta_row_ptr;   // [cl+1] rows owned by CTA k
  const int32_t* cta_blk_ptr;   // [cl+1] upper blocks of CTA k in the ub_* arrays
  const int32_t* cta_halo_ptr;  // [cl+1]
  const int32_t* ub_sblk;       // block index in S
  const int32_t* ub_rowl;       // row of the block, local to its CTA
  const int32_t* ub_halo;       // halo slot holding u of the block's column
  const int32_t* ub_depA;       // deposit slot (local) of S_ij u_j
  const int32_t* ub_tcta;       // CTA owning row j
  const int32_t* ub_tidx;       // deposit slot of S_ij^T u_i at that CTA
  const int32_t* halo_cta;      // owner CTA of the halo row
  const int32_t* halo_rowl;     // its local row index there
  const double* S;
  const double* Minv;
  const double* rhs;
  double* x;
  double* scal;
  double tol;
  int max_iter;
};
size_t pcg_cluster_carve(PcgClusterArgs& a);  // fills a.m from the maxima, returns the dynamic shared-memory bytes
int pcg_cluster_max_smem(int device);
bool pcg_cluster_fits(int device, int cl, size_t smem_bytes);
cudaError_t launch_pcg_cluster(const PcgClusterArgs& a, int cl, size_t smem_bytes, cudaStream_t st);

// coarse space of the two-level preconditioner (coarse.cu)
constexpr int COARSE_MAX_GROUPS = 64;  // 7 * 64 = 448 modes: <= 18 groups one CTA, <= 32 a four-CTA cluster, beyond a sixteen-CTA cluster (coarse.cu)
constexpr int COARSE_MAX_DIM = 7 * COARSE_MAX_GROUPS;
constexpr int COARSE_SPLIT = 8;  // partial coarse matrices of k_coarse_assemble, added in order by k_coarse_invert  // the dense inverse is formed by one CTA in shared memory
struct CoarsePlan {
  int G;
  const int32_t* cb_ptr;   // [G*G+1] fine blocks of every coarse block (gi, gj)
  const int32_t* cb_fine;  // [B] block indices of S, grouped by coarse block, CSR order inside
};
cudaError_t launch_coarse_setup(const CoarsePlan& cp, int C, const double* pose, const uint8_t* free_mask,
                                const int32_t* s_row, const int32_t* s_col, const double* S, double* Zc, double* Ac,
                                double* Ainv, cudaStream_t st, cudaEvent_t after_assemble = nullptr);
// the gauge modes alone (Z of every image at the given poses)
cudaError_t launch_coarse_modes(int C, const double* pose, const uint8_t* free_mask, double* Zc, cudaStream_t st);

// ---- hash of the plan region (ba_solver.cu: upload_plan): order-independent sum of mixed (value, index) words
MPBA_HD unsigned long long plan_hash_term(int32_t x, unsigned long long i) {
  unsigned long long z = (((unsigned long long)(uint32_t)x << 32) | (i & 0xffffffffull)) + 0x9E3779B97F4A7C15ull * (i + 1);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
void launch_plan_hash(const int32_t* region, size_t n, unsigned long long* out, cudaStream_t st);

// ---- dense direct solve of a small reduced camera system (dense_solve.cu) ----------------------------
constexpr int DENSE_SOLVE_MAX_IMAGES = 21;  // 6 * 21 unknowns + the right-hand side = 127 columns of the register tile
bool dense_solve_fits(int num_images);
// pose_cand != nullptr: the camera step (launch_cam_update's work) runs as the epilogue of the same kernel
cudaError_t launch_dense_solve(int C, const int32_t* row_ptr, const int32_t* col, const double* S, const double* rhs,
                               double* y, double* scal, const ProblemView* pv, const double* pose, const double* Hc,
                               const double* gc, const double* c_img, LmDiag lm, double* pose_cand, cudaStream_t st);

// ---- step ---------------------------------------------------------------------------------------
// q (+) -y, t - y, step norms, and the image part of the model cost change (g'y + y'Dy) / 2
void launch_cam_update(const ProblemView& pv, const double* pose, const double* y, const double* Hc, const double* gc,
                       const double* c_img, LmDiag lm, double* pose_cand, double* scal, cudaStream_t st);
// cam != nullptr: the camera step (launch_cam_update's work) is launched first as CTAs of 128 images each
// (k_cam_update_slices); their slice sums land in cam->stage ([cam_step_rows(C)][3]) and are added by
// launch_pt_scalars_step
struct CamStepArgs {
  const double *pose, *Hc, *gc, *c_img;
  double* pose_cand;
  double* stage;
};
int cam_step_rows(int num_images);
void launch_back_substitute(const ProblemView& pv, const double* X, const double* y, const double* Hpp,
                            const double* gp, const double* Linv, const double* c_pt, const double* fac,
                            int use_factors, LmDiag lm, double* X_cand, double* pt_part, const CamStepArgs* cam,
                            cudaStream_t st);
// stage: PT_SCALARS_STAGE doubles of workspace, zero before the first launch (slice partials + the arrival ticket)
constexpr int PT_SCALARS_STAGE = 4 * 64 + 2;
void launch_pt_scalars_step(const double* pt_part, int nblk, double* scal, int with_gmax, double* stage,
                            const double* cam_stage, int cam_rows, cudaStream_t st);

// ---- misc ---------------------------------------------------------------------------------------
void launch_evaluate(const ProblemView& pv, const double* pose, const double* X, double* r_reproj, double* r_depth,
                     cudaStream_t st);
// post-BA track filter (COLMAP FilterPoints3D rule, ba_math.cuh: filter_track); scal[S_FILTERED] += removed observations
void launch_filter_tracks(const ProblemView& pv, const double* pose, const double* X, double max_err2, double min_angle,
                          const uint8_t* pt_select, uint8_t* obs_keep, uint8_t* pt_keep, double* pt_error, double* scal,
                          cudaStream_t st);
// out[0] += observations of this rank's tracks whose point is at z < epsilon in the camera frame
void launch_count_behind(const ProblemView& pv, const double* pose, const double* X, double* out, cudaStream_t st);
// unit quaternions (COLMAP normalises before adding to the problem) + exp(log-scale) into slot 7
void launch_normalize_poses(int C, const double* in, const double* shift_scale, double* out, cudaStream_t st);
void launch_flag_points(int P, const uint8_t* pt_const, double* X, cudaStream_t st);
void launch_expand_hcc(int C, const double* Hc, double* Hcc36, cudaStream_t st);

}  // namespace mpba
