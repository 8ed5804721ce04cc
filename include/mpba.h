/* mpba.h -- C ABI of the B200-native bundle-adjustment engine (libmpba.so).
 *
 * The reference reaches its solver through pybind11 objects, not a C ABI:
 *   mpsfm/sfm/mapper/bundle_adjustment.py:104      pycolmap.create_default_bundle_adjuster
 *   mpsfm/sfm/mapper/bundle_adjustment.py:163-176  pycolmap.create_depth_bundle_adjuster (fork only)
 *   mpsfm/sfm/mapper/bundle_adjustment.py:114-122  pyceres constant blocks / manifolds
 *   mpsfm/sfm/mapper/bundle_adjustment.py:285-293  pyceres.solve (SPARSE_SCHUR, Ceres defaults)
 * This header is the boundary a maintainer binds instead (ctypes stub in INTEGRATION.md).
 * One `mpba_problem` is the array form of the Ceres problem those calls assemble.
 *
 * Conventions
 *   - every pointer inside mpba_problem / workspace is a DEVICE pointer owned by the caller
 *     (PyTorch tensors in this repo); the library never allocates device memory,
 *   - all floating point data is IEEE double, all indices int32,
 *   - every call returns 0 on success or a negative MPBA_E_* code; the message is available from
 *     mpba_last_error(); nothing throws across the boundary,
 *   - slot 7 of an img_pose record and slot 3 of a pt_xyz record are scratch: a solve leaves exp(log-scale) of
 *     the image's depth prior / the point's constant flag there,
 *   - a handle is bound to one device and one stream and is not thread-safe,
 *   - there is no CPU fallback: without a CUDA device mpba_create fails with MPBA_E_CUDA.
 */
#ifndef MPBA_H_
#define MPBA_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MPBA_VERSION 100

enum {
  MPBA_OK = 0,
  MPBA_E_INVALID = -1,   /* bad argument / inconsistent problem */
  MPBA_E_CUDA = -2,      /* CUDA runtime error (message has the cudaError string) */
  MPBA_E_WORKSPACE = -3, /* workspace too small */
  MPBA_E_NUMERIC = -4,   /* non-finite value met where the algorithm cannot continue */
  MPBA_E_NCCL = -5
};

/* COLMAP camera model ids (pycolmap.CameraModelId; upstream colmap/sensor/models.h) */
enum {
  MPBA_SIMPLE_PINHOLE = 0, MPBA_PINHOLE = 1, MPBA_SIMPLE_RADIAL = 2, MPBA_RADIAL = 3, MPBA_OPENCV = 4,
  MPBA_OPENCV_FISHEYE = 5, MPBA_FULL_OPENCV = 6, MPBA_FOV = 7, MPBA_SIMPLE_RADIAL_FISHEYE = 8,
  MPBA_RADIAL_FISHEYE = 9, MPBA_THIN_PRISM_FISHEYE = 10
};

/* Ceres loss functions reachable from bundle_adjustment.py:44-48,:100 (+ Huber, north star) */
enum { MPBA_LOSS_TRIVIAL = 0, MPBA_LOSS_SOFT_L1 = 1, MPBA_LOSS_CAUCHY = 2, MPBA_LOSS_HUBER = 3 };

/* Ceres TerminationType */
enum { MPBA_CONVERGENCE = 0, MPBA_NO_CONVERGENCE = 1, MPBA_FAILURE = 2 };

#define MPBA_MAX_CAMERA_PARAMS 12

/* One bundle-adjustment problem.  C images (poses), K cameras (intrinsics, constant:
 * bundle_adjustment.py:92-94), P points3D, N observations.
 *
 * Observations are given twice, in the two orders the two reduction keys need:
 *   track order  ("pm"): grouped by point3D, pt_ptr[p]..pt_ptr[p+1]; inside a track sorted by image,
 *   image order  ("cm"): grouped by image,   img_ptr[i]..img_ptr[i+1].
 * Both orders hold the same N observations; *_src map each slot back to the caller's
 * (natural) observation index, and is what makes the per-residual outputs comparable bit for bit.
 * A depth prior is attached to an observation by depth weight w > 0 (w == 0: no prior row).
 */
typedef struct mpba_problem {
  int32_t num_images, num_cameras, num_points, num_obs;

  /* images: pose[i] = {qx,qy,qz,qw,tx,ty,tz,unused}  (Eigen quaternion order, cam_from_world) */
  double* img_pose;             /* [C][8]  in/out */
  const int32_t* img_camera;    /* [C]     row of cam_model / cam_params */
  const uint8_t* img_free_mask; /* [C]     bit j: tangent coordinate j free (0-2 rot, 3-5 trans) */
  const double* img_shift_scale;/* [C][2]  depth prior (shift b, log-scale s), constant blocks */
  const int32_t* cam_model;     /* [K] */
  const double* cam_params;     /* [K][12] */

  /* points3D: {x,y,z,unused} */
  double* pt_xyz;               /* [P][4]  in/out */
  const uint8_t* pt_const;      /* [P]     1 = constant parameter block */

  /* track order */
  const int32_t* pt_ptr;        /* [P+1] */
  const int32_t* pm_img;        /* [N] */
  const double* pm_xy;          /* [N][2] */
  const double* pm_depth;       /* [3][N]  rows d, w, a  (prior depth, ScaledLoss weight, loss scale) */
  const int32_t* pm_src;        /* [N]     natural observation index */

  /* image order */
  const int32_t* img_ptr;       /* [C+1] */
  const int32_t* cm_pt;         /* [N] */
  const double* cm_xy;          /* [N][2] */
  const double* cm_depth;       /* [3][N] */
  /* image-order work chunks (at most MPBA_CHUNK observations of one image each) */
  int32_t num_chunks;
  const int32_t* chunk_img;     /* [num_chunks] */
  const int32_t* chunk_begin;   /* [num_chunks] first slot (image order) */
  const int32_t* chunk_end;     /* [num_chunks] */
  const int32_t* img_chunk_ptr; /* [C+1] chunks of image i = img_chunk_ptr[i]..img_chunk_ptr[i+1] */

  /* reduced camera system: block-CSR over co-visible image pairs, full (both triangles),
   * columns sorted inside a row; 6x6 row-major blocks */
  int32_t num_blocks;
  const int32_t* s_row_ptr;     /* [C+1] */
  const int32_t* s_col;         /* [B] */
  const int32_t* s_diag;        /* [C]   block index of (i,i) */
  const int32_t* s_transpose;   /* [B]   block index of (j,i) for block (i,j) */
  const int32_t* s_row;         /* [B]   block row of every block */
  /* Schur work lists: one unit per upper block (row <= col).  Unit u forms block u_block[u] from the
   * items u_item_ptr[u]..u_item_ptr[u+1]; an item is a pair of TRACK-ORDER slots (a <= b) of one track
   * whose images are (row, col) of the block; items of a unit are sorted by track.  Tracks of
   * constant points contribute no items. */
  int32_t num_upper;
  const int32_t* u_block;       /* [U] */
  const int32_t* u_item_ptr;    /* [U+1] */
  const int32_t* item_a;        /* [u_item_ptr[U]] */
  const int32_t* item_b;        /* [u_item_ptr[U]] */

  /* losses: ScaledLoss(loss(scale), weight) shared by all reprojection blocks
   * (bundle_adjustment.py:98-102); per-observation scale/weight for depth (:160-161) */
  int32_t reproj_loss, depth_loss;
  double reproj_loss_scale, reproj_loss_weight;

  /* MPBA_FLAG_*; STRUCTURE_ONLY: every image pose is constant (refine_3d_points,
   * bundle_adjustment.py:276-283) -- the normal equations are block diagonal in the points and
   * the reduced camera system is skipped */
  int32_t flags;
  int32_t s_max_row_blocks;     /* max blocks in one row of S (0 = unknown): sizes the PCG's shared-memory staging */

  /* track-order warp packing (mpba_pack_tracks): the track-order passes run one thread per observation and one
   * warp per group of WHOLE consecutive tracks, tracks track_warp_first[w] .. track_warp_first[w+1] - 1, with at
   * most 32 observations and 32 tracks per group (a longer track is a group of its own).  Covers the owned
   * tracks: track_warp_first[0] = 0 (sharded: pt_begin), track_warp_first[num_track_warps] = num_points (pt_end). */
  int32_t num_track_warps;
  int32_t reserved;
  const int32_t* track_warp_first; /* [num_track_warps + 1] */

  /* EXTENSION without a counterpart in the reference (its bundle adjustment has no residual for monocular normals,
   * mpsfm/sfm/mapper/bundle_adjustment.py:60-100; defined by oracle/residuals.py):
   * point-to-plane rows  r = m . p_cam - c  with  m = sqrt(w_n) n,  c = m . (prior surface point),  loss
   * ScaledLoss(depth_loss(plane_loss_scale), 1).  One record {m_x, m_y, m_z, c} per observation in both orders
   * (m = 0: no row); both NULL: the problem has no such rows and nothing changes.  Sharded solves: the rank's slots,
 * like pm_xy / cm_xy. */
  const double* pm_plane;       /* [N][4] track order, or NULL */
  const double* cm_plane;       /* [N][4] image order, or NULL */
  double plane_loss_scale;      /* > 0 when the arrays are given */
} mpba_problem;

#define MPBA_FLAG_STRUCTURE_ONLY 1

#define MPBA_CHUNK 1024
/* Image-order chunk length the host layout uses for a bundle of num_obs observations: MPBA_CHUNK where the bundle
 * fills the device, halved down to 128 while fewer than ~4 chunks per SM would result -- k_img_chunks is one CTA per
 * chunk, and the mapper's local bundles (10^4 observations) otherwise run on a dozen CTAs. */
int32_t mpba_chunk_size(int64_t num_obs);

/* Ceres Solver::Options subset; defaults (mpba_default_options) are Ceres 2.1's, which is what the
 * reference runs with (bundle_adjustment.py:285-293 sets only solver type / threads / verbosity). */
typedef struct mpba_options {
  int32_t max_num_iterations;              /* 50 */
  int32_t max_num_consecutive_invalid_steps; /* 5 */
  int32_t jacobi_scaling;                  /* 1 */
  int32_t pcg_max_iterations;              /* 2000 */
  double initial_trust_region_radius;      /* 1e4 */
  double max_trust_region_radius;          /* 1e16 */
  double min_trust_region_radius;          /* 1e-32 */
  double min_relative_decrease;            /* 1e-3 */
  double min_lm_diagonal;                  /* 1e-6 */
  double max_lm_diagonal;                  /* 1e32 */
  double function_tolerance;               /* 1e-6 */
  double gradient_tolerance;               /* 1e-10 */
  double parameter_tolerance;              /* 1e-8 */
  double pcg_tolerance;                    /* 1e-10: stop when sqrt(r'M^-1 r / b'M^-1 b) <= tol.  The step minimises
                                              the quadratic model, so a relative step error eps moves the cost by
                                              O(eps^2): 1e-10 leaves the LM trajectory on the exact-solve one to ~1e-13 */
  int32_t profile;                         /* 1: CUDA-event time every kernel launch (see mpba_kernel_times) */
  int32_t verbose;
  int32_t pcg_kernel;                      /* 0: automatic (up to 21 images: dense direct solve by one CTA, no CG at all
                                              -- iterations report pcg_iterations = 0; cluster-resident PCG when S fits
                                              one 16-CTA cluster, grid-wide otherwise), 1: force grid-wide PCG,
                                              2: cluster PCG when it fits (never the dense solve), 3: as 0 */
  int32_t pcg_coarse;                      /* 0: two-level preconditioner (block-Jacobi + gauge-mode coarse space over
                                              groups of images) with the cluster kernel, 1: block-Jacobi only */
  int32_t behind_camera;                   /* point at z < epsilon in the camera frame.  0 (default): COLMAP >= 3.11's
                                              ReprojErrorCostFunctor -- ImgFromCam fails and the block contributes a ZERO
                                              residual and Jacobian; such observations are counted in
                                              mpba_summary.num_behind_camera.  1: strict -- the projection is evaluated as
                                              written (inf / nan) and a non-finite initial cost returns MPBA_E_NUMERIC */
  int32_t coarse_lag;                      /* 1 (default): the dense inverse of the coarse matrix is a preconditioner --
                                              the PCG of linear solve i uses the one built from S of solve i - 1 and
                                              solve i's own is built on a side stream, overlapping the kernels that
                                              follow; 0: rebuilt on the solve stream before every PCG */
  int32_t lm_loop;                         /* where the LM controller (accept / reject, trust region, termination) runs.
                                              0 (default): on the device for bundles below 600 images -- the whole LM
                                              loop is ONE CUDA-graph launch (a WHILE node over the iteration's kernels,
                                              csrc/lm_loop.cu), one synchronisation per solve -- and on the host above;
                                              1: host loop always; 2: device loop wherever it is legal (single GPU,
                                              cluster-resident PCG, profile = verbose = 0, max_num_iterations <= 254).
                                              Both controllers produce the same iterates bit for bit */
  int32_t reserved0;
} mpba_options;

/* ceres::IterationSummary subset, one row per LM iteration (row 0 = initial evaluation) */
typedef struct mpba_iteration {
  double cost;               /* includes fixed_cost */
  double cost_change;
  double gradient_max_norm;
  double step_norm;
  double relative_decrease;
  double trust_region_radius;
  int32_t step_is_valid, step_is_successful;
  int32_t pcg_iterations;
  int32_t reserved;
} mpba_iteration;

typedef struct mpba_summary {
  double initial_cost, final_cost, fixed_cost;
  int32_t termination;       /* MPBA_CONVERGENCE / NO_CONVERGENCE / FAILURE */
  int32_t num_iterations;    /* rows filled in `iterations` minus 1 = LM iterations run */
  int32_t num_successful_steps, num_unsuccessful_steps;
  int64_t total_pcg_iterations;
  int64_t kernel_launches;   /* kernels this call launched */
  double solve_ms;           /* device time of the whole call (CUDA events) */
  mpba_iteration* iterations;/* HOST buffer with capacity rows, caller-owned; may be NULL */
  int32_t capacity;
  int32_t num_behind_camera; /* observations at z < epsilon in their camera frame at the returned state */
  char message[128];
} mpba_summary;

/* names of the kernels for mpba_kernel_times, in slot order */
#define MPBA_NUM_KERNELS 14

typedef struct mpba_handle mpba_handle;

int mpba_version(void);

/* lifetime */
int mpba_create(mpba_handle** out, int device, void* cuda_stream /* cudaStream_t or NULL */);
void mpba_destroy(mpba_handle* h);
const char* mpba_last_error(const mpba_handle* h);
void mpba_default_options(mpba_options* opt);

/* HOST helper: greedy packing of the tracks pt_begin .. pt_end - 1 (HOST array pt_ptr, global indices) into
 * warp groups as mpba_problem.track_warp_first wants them.  first must hold pt_end - pt_begin + 1 entries;
 * *num_warps receives the number of groups. */
int mpba_pack_tracks(const int32_t* pt_ptr, int32_t pt_begin, int32_t pt_end, int32_t* first, int32_t* num_warps);

/* bytes of device workspace mpba_* calls need for this problem */
int mpba_workspace_bytes(const mpba_problem* prob, size_t* bytes);

/* Residuals at the current state, in TRACK order: reproj [N][2], depth [N] (0 where no prior);
 * cost[0] = 1/2 sum rho over blocks with a free parameter block, cost[1] = fixed cost.
 * Output pointers are device pointers; NULL skips that output.  cost is a HOST pointer. */
int mpba_evaluate(mpba_handle* h, const mpba_problem* prob, void* workspace, size_t workspace_bytes,
                  double* r_reproj, double* r_depth, double* cost);

/* Post-BA track filter at the current state.  Replaces pycolmap's
 * ObservationManager.filter_points3D(max_reproj_error, min_tri_angle, point3D_ids) as the mapper calls it after
 * every bundle (mpsfm/sfm/mapper/base.py:700-711 filter_bundle, :772-797 filter_local_points3D), restating COLMAP's
 * rule per track: observations with reprojection error > max_reproj_error [pixels] go (a point behind the camera has
 * infinite error); a track left with < 2 observations loses its point; a point whose kept observations have no pair
 * of projection centres subtending >= min_tri_angle_deg at the point goes too.
 * pt_select [P] (1 = filter this point; NULL = all) -- unselected points are reported kept and not evaluated.
 * Device outputs: obs_keep [N] in TRACK order, pt_keep [P], pt_error [P] = mean kept reprojection error (-1 for
 * removed / unselected points).  num_filtered (HOST, may be NULL) = observations removed, counted as COLMAP does. */
int mpba_filter_tracks(mpba_handle* h, const mpba_problem* prob, void* workspace, size_t workspace_bytes,
                       double max_reproj_error, double min_tri_angle_deg, const uint8_t* pt_select, uint8_t* obs_keep,
                       uint8_t* pt_keep, double* pt_error, int64_t* num_filtered);

/* Normal-equation blocks at the current state (unscaled, undamped), device outputs:
 * Hcc [C][36], gc [C][6], Hpp [P][6] (xx,xy,xz,yy,yz,zz), gp [P][3]; cost as above. */
int mpba_linearize(mpba_handle* h, const mpba_problem* prob, void* workspace, size_t workspace_bytes,
                   double* Hcc, double* gc, double* Hpp, double* gp, double* cost);

/* 3x3 covariance of every variable point conditioned on the poses, (Hpp)^-1 of the undamped, unscaled
 * normal equations at the current state: cov [P][6] (xx,xy,xz,yy,yz,zz), device pointer; zeros for constant
 * points and for blocks that are not positive definite.  Replaces pycolmap.estimate_ba_covariance with
 * BACovarianceOptionsParams.POINTS (mpsfm/sfm/mapper/bundle_adjustment.py:244-261). */
int mpba_point_covariances(mpba_handle* h, const mpba_problem* prob, void* workspace, size_t workspace_bytes,
                           double* cov);

/* Reduced camera system for LM damping lambda_j = clamp(c_j^2 H_jj)/ (radius c_j^2) with Jacobi
 * scaling c taken at the current state: S [B][36], rhs [C][6], and its PCG solution y [C][6]
 * (any may be NULL).  pcg_iterations is a HOST pointer. */
int mpba_reduced_system(mpba_handle* h, const mpba_problem* prob, void* workspace, size_t workspace_bytes,
                        const mpba_options* opt, double radius, double* S, double* rhs, double* y,
                        int32_t* pcg_iterations);

/* Trust-region Levenberg-Marquardt (Ceres 2.1 TrustRegionMinimizer semantics); updates img_pose
 * and pt_xyz in place to the best state visited.
 * Error convention of the reference (bundle_adjustment.py:266, :283, :285-293): pyceres.solve never raises, a
 * failed solve is only visible in the summary.  Likewise here: a non-finite initial cost (e.g. a depth prior on a
 * point with z <= 0, or nan inputs) returns MPBA_OK with termination = MPBA_FAILURE and leaves img_pose / pt_xyz
 * untouched, as Ceres does when the initial residual evaluation fails.  (opt->behind_camera = 1 restores the loud
 * MPBA_E_NUMERIC.) */
int mpba_solve(mpba_handle* h, const mpba_problem* prob, void* workspace, size_t workspace_bytes,
               const mpba_options* opt, mpba_summary* summary);

/* ---- sharded solve: one process per GPU, one problem across `world` GPUs of a box ---------------------------
 * BASELINE.json north star: "observations are sharded by camera across the GPUs ... the partial reduced-camera-
 * system blocks ... are combined with NCCL allreduce over NVLink".  The path has two reduction keys, so every rank
 * holds two shards of the observations ("Multi-GPU"):
 *   image order: the chunks of its OWN images  [img_begin, img_end)  -> Hcc, gc, cost of those images,
 *   track order: the tracks of its OWN points  [pt_begin, pt_end)    -> Hpp, gp, Schur factors, back-substitution.
 * In the rank's mpba_problem every per-image / per-point array is still indexed globally (num_images, num_points
 * are the global counts; poses, points and the structure of the reduced camera system are replicated), img_ptr /
 * img_chunk_ptr / pt_ptr are global-length with EMPTY ranges for what the rank does not own, and the
 * observation arrays, chunk lists and Schur item lists hold the rank's own entries only (local slot numbers).
 * pt_begin = rank * pt_stride (equal counts for the all-gather of the candidate points; the padded point
 * buffers live in the workspace, pt_xyz itself holds num_points entries).
 * Collectives per LM iteration, on the handle's stream: all-reduce of the image blocks (27 C doubles) and of the
 * partial reduced camera system (36 B + 6 C), all-gather of the candidate points (4 pt_stride per rank) and of
 * one block of 24 scalars.  The reduced system is then solved redundantly by every rank (bitwise identical). */
typedef struct mpba_shard {
  int32_t rank, world;
  int32_t pt_begin, pt_end, pt_stride;
  int32_t img_begin, img_end;
  int32_t reserved;
} mpba_shard;

#define MPBA_UNIQUE_ID_BYTES 128
/* rank 0 creates the id (ncclGetUniqueId), the caller broadcasts it (torch.distributed / MPI / files), every rank
 * calls mpba_comm_init (ncclCommInitRank).  NCCL is taken from the process (the libnccl.so.2 PyTorch already
 * loaded) or dlopen'ed from nccl_library (NULL: default search path).  HOST pointers. */
int mpba_comm_unique_id(void* id, const char* nccl_library);
int mpba_comm_init(mpba_handle* h, int world, int rank, const void* id, const char* nccl_library);
int mpba_workspace_bytes_sharded(const mpba_problem* prob, const mpba_shard* shard, size_t* bytes);
/* mpba_solve over the shards; every rank calls it with its own problem / shard and gets the full solution
 * (img_pose, pt_xyz) and the same summary. */
int mpba_solve_sharded(mpba_handle* h, const mpba_problem* prob, const mpba_shard* shard, void* workspace,
                       size_t workspace_bytes, const mpba_options* opt, mpba_summary* summary);

/* ---- problem assembly on the device (the step before the solve in the reference) --------------------------------
 * Batched PriorUtils._data_at_kps (mpsfm/sfm/scene/image/mixins/priorutils.py:49-62; called for depth, variance
 * and validity maps at bundle_adjustment.py:132-135, 215): out[k] = bilinear sample (fp64, align_corners = True,
 * zero padding) of map map_of[k] at kps[k] * (sx, sy) of that map; valid_at_kps is `out == 1` on a 0/1 map.
 * Same operation order as torch.nn.functional.grid_sample, no fused multiply-adds.  maps: the H x W row-major
 * maps back to back; map_offset[m] in doubles; map_hw[m] = {H, W}; map_scale[m] = {sx, sy}.  DEVICE pointers;
 * runs on cuda_stream (NULL: default stream) and returns after it completed. */
int mpba_sample_maps(mpba_handle* h, int32_t n, const double* kps, const int32_t* map_of, int32_t num_maps,
                     const int64_t* map_offset, const int32_t* map_hw, const double* map_scale, const double* maps,
                     double* out, void* cuda_stream);

/* ---- host side of the boundary: the index arrays of mpba_problem from the caller's natural observation list ----
 * What COLMAP's AddImageToProblem + Ceres' block ordering / Schur structure analysis do for the reference
 * (bundle_adjustment.py:104, :285-293) is this call: both observation orders (stable counting sorts: track order by
 * (point, image), image order by (image, point)), the image-order chunks, the block-CSR structure of the reduced
 * camera system over co-visible image pairs, the Schur work lists (items of a unit sorted by track) and the warp
 * groups of mpba_pack_tracks.  HOST pointers in, HOST arrays out (owned by the returned object, valid until
 * mpba_host_layout_destroy); multi-threaded (MPBA_HOST_THREADS, default: all hardware threads).
 * cm_pm[i] = track-order slot of image-order slot i (mpba_gather_observations' permutation). */
typedef struct mpba_host_layout mpba_host_layout;
typedef struct mpba_host_layout_view {
  int32_t num_images, num_points, num_obs;
  int32_t num_chunks, num_blocks, num_upper, num_track_warps;
  int64_t num_items;
  const int32_t *pt_ptr, *pm_img, *pm_src, *img_ptr, *cm_pt, *cm_src, *cm_pm;
  const int32_t *chunk_img, *chunk_begin, *chunk_end, *img_chunk_ptr;
  const int32_t *s_row_ptr, *s_col, *s_diag, *s_transpose, *s_row, *u_block, *u_item_ptr, *item_a, *item_b;
  const int32_t *track_warp_first;
} mpba_host_layout_view;
int mpba_host_layout_create(int32_t num_images, int32_t num_points, int32_t num_obs, const int32_t* obs_img,
                            const int32_t* obs_pt, const uint8_t* pt_const, int32_t structure_only,
                            mpba_host_layout** out, mpba_host_layout_view* view);
void mpba_host_layout_destroy(mpba_host_layout* layout);
/* value arrays of one observation order: out_xy[k] = obs_xy[src[k]], out_depth[r][k] = {d, w, a}[r][src[k]]
 * (src = pm_src or cm_src).  HOST pointers. */
int mpba_host_gather(int32_t num_obs, const int32_t* src, const double* obs_xy, const double* d, const double* w,
                     const double* a, double* out_xy, double* out_depth);

/* The problem holds every observation twice (track order pm_*, image order cm_*).  A caller whose arrays live on the
 * host uploads one order plus a 4-byte permutation and derives the other on the device: dst_xy[i] =
 * src_xy[perm[i]], dst_depth[r][i] = src_depth[r][perm[i]] (r = 0..2, the [3][n] layout of pm_depth / cm_depth).
 * 44 instead of 80 bytes per observation over PCIe.  DEVICE pointers; ASYNCHRONOUS on cuda_stream. */
int mpba_gather_observations(mpba_handle* h, int32_t n, const int32_t* perm, const double* src_xy,
                             const double* src_depth, double* dst_xy, double* dst_depth, void* cuda_stream);

/* Per-observation depth-prior terms of a whole bundle in one launch: the arithmetic of the per-image Python loops at
 * bundle_adjustment.py:136-161 (__build_problem), :239-241 (__build_shiftscale_problem) and :311-330
 * (update_truncation_multiplier).  For sample k of image img_of[k]:
 *   z        = row3[img] . (xyz[k], 1)      row3 = third row of cam_from_world.matrix() (utils/geometry.py:13-19)
 *   mask     = d > 0  [and 1/f < d/z < f if scale_filter_factor = f > 0]  [and |clip(log d,1e-6) - clip(log z,1e-6)|
 *              / sqrt(v) < 3 if gross_outliers]                                          (:139-147)
 *   w        = d*d * (1 / max(v, 1e-6)),   a = loss_multiplier * sqrt(v) / d              (:156-161)
 *   logratio = log(max(z / d, 1e-6))                                                      (:241)
 *   whitened = (log d - log z) / max(sqrt(v) / d, 1e-6)                                   (:326-330)
 * Every output pointer may be NULL (not computed).  w and a are bit-identical to numpy's (IEEE mul / div / sqrt in
 * the reference's order); z, logratio and whitened agree to the last ulps (BLAS matmul order, libm log).
 * DEVICE pointers; runs on cuda_stream and returns after it completed. */
int mpba_depth_prior_terms(mpba_handle* h, int32_t n, const int32_t* img_of, int32_t num_images, const double* row3,
                           const double* xyz, const double* depth, const double* variance, double scale_filter_factor,
                           int32_t gross_outliers, double loss_multiplier, double* out_z, uint8_t* out_mask,
                           double* out_w, double* out_a, double* out_logratio, double* out_whitened,
                           void* cuda_stream);

/* Exact segmented median (numpy.median semantics: mean of the two middle order statistics, NaN if the segment
 * holds a NaN or is empty) and, if out_mad != NULL, the median absolute deviation median(|x - median|) of every
 * segment: optimize_prior_shiftscale's per-image np.median (bundle_adjustment.py:239-241) with one segment per
 * image, fit_robust_gaussian_mad (:10-15) with one segment over all registered images.  Radix select on the
 * order-preserving 64-bit key, 8 passes of one byte, both middle ranks at once; bit-exact against numpy.
 * seg_ptr: int32[num_segments + 1], values[seg_ptr[s] .. seg_ptr[s+1]) is segment s.  DEVICE pointers; the
 * workspace (mpba_segmented_median_workspace_bytes, 256-byte aligned) is caller-owned. */
int mpba_segmented_median_workspace_bytes(int32_t n, int32_t num_segments, size_t* bytes);
int mpba_segmented_median(mpba_handle* h, int32_t n, const double* values, int32_t num_segments,
                          const int32_t* seg_ptr, void* workspace, size_t workspace_bytes, double* out_median,
                          double* out_mad, void* cuda_stream);

/* The same select WITHOUT a caller workspace and without any allocation (no cudaMalloc / cudaFree, no allocating
 * library routine): its scratch is a file-scope `static __device__` array of the library.  No counterpart in the
 * reference: added for callers that cannot allocate (CUDA graph capture, tight
 * serving loops, harnesses that forbid cudaMalloc).  Limits: n <= MPBA_MEDIAN_STATIC_MAX_VALUES and num_segments <=
 * MPBA_MEDIAN_STATIC_MAX_SEGMENTS, beyond them MPBA_E_WORKSPACE with a message (mpba_last_error when h != NULL; h may
 * be NULL).  The array (about 2.6 MB) exists on every GPU that loads the library until the process exits; its device
 * address is looked up once per device (the device CURRENT at the call) with cudaGetSymbolAddress.  Every call shares
 * the array, which keeps the previous call's contents (what must start at zero is cleared by the call): calls that
 * overlap on the GPU -- two streams, two host threads -- race on it and must be serialised by the caller.  Same
 * results, bit for bit, and same completion contract as mpba_segmented_median. */
#define MPBA_MEDIAN_STATIC_MAX_VALUES 262144
#define MPBA_MEDIAN_STATIC_MAX_SEGMENTS 256
int mpba_segmented_median_static(mpba_handle* h, int32_t n, const double* values, int32_t num_segments,
                                 const int32_t* seg_ptr, double* out_median, double* out_mad, void* cuda_stream);


/* ---- depth-map integration with normal priors ------------------------------------------------
 * The step right before every bundle adjustment in the reference (mpsfm/sfm/mapper/base.py:637-644 integrate_bundle
 * -> mpsfm/sfm/scene/image/integration.py:133-137 Image.integrate -> :383-520 Integration._integrate): per image, an
 * IRLS loop on the log-depth map z over the H x W grid,
 *     min  sum wu+ (A1 z + nx)^2 + wu- (A2 z + nx)^2 + wv+ (A3 z + ny)^2 + wv- (A4 z + ny)^2
 *          + lambda1 prec_d (z_prior - z)^2 + lambda2 prec_s (log z_sparse - z)^2                      (:139-165)
 * with one-sided difference operators A1..A4 scaled by the normals (:631-680), sigmoid discontinuity weights
 * (:358-376), the normal equations of every step (:167-234, :445-453) solved by Jacobi-preconditioned CG with the
 * reference's settings (warm start, rtol = cg_tol, scipy/cupy `cg` semantics, :455-466), and the reference's
 * termination rules and per-image state (integrated, energy_old, cached wu / wv).  Replaces the cupy csr_matrix +
 * cupyx cg library calls by ONE kernel launch for a whole batch of images: one thread-block cluster per image, the
 * five-point operator applied matrix-free, cluster-wide dot products through distributed shared memory.
 * Every pointer in a job is a DEVICE pointer; the job / result arrays themselves are HOST arrays. */
#define MPBA_INT_MAX_IRLS 32
typedef struct mpba_integration_conf {
  int32_t max_iter;                 /* 10    IRLS steps (<= MPBA_INT_MAX_IRLS) */
  int32_t cg_max_iter;              /* 5000 */
  double tol;                       /* 5e-2  relative energy change */
  double step_size;                 /* 1 */
  double cg_tol;                    /* 1e-3  rtol of cg */
  double lambda1, lambda2;          /* 1, 1 */
  double k;                         /* 1     sigmoid slope */
  double large_number;              /* 1e6   normal variance of invalid pixels */
  double depth_magnitude_multiplier, normals_magnitude_multiplier; /* 1, 1 */
  double scale_filter_factor;       /* 1.5; 0 = scale filter off (conf.scale_filter = False) */
} mpba_integration_conf;            /* Image.default_conf, mpsfm/sfm/scene/image/base.py:36-53 */

typedef struct mpba_integration_job {
  int32_t H, W;                     /* grid of the prior maps (camera.nshape) */
  int32_t num_sparse;               /* triangulated points projected into this image (:95-120) */
  int32_t integrated;               /* in: the image was integrated before (IntVars.integrated): wu / wv / energy_old hold its state */
  double K[4];                      /* [K11 sy, K00 sx, K12 sy, K02 sx] as _prepare_integration_variables builds it (:123-129) */
  const double* depth_prior;        /* [H][W] depth.data_prior */
  const double* uncertainty;        /* [H][W] depth.uncertainty (variance) */
  const uint8_t* valid;             /* [H][W] depth.valid */
  const double* normals;            /* [H][W][3] normals.data */
  const double* normals_uncertainty;/* [H][W][3][3] normals.uncertainty (only the diagonal is read, :276-278) */
  const int32_t* sparse_xy;         /* [num_sparse][2] integer pixel (x, y) = (kps * (sx, sy) + 0.5).astype(int), inside the grid */
  const double* sparse_depth;       /* [num_sparse] depth3d */
  const double* sparse_zvar;        /* [num_sparse] zvars3d */
  double* depth;                    /* [H][W] in: depth.data (start), out: refined depth when result.updated */
  double* wu;                       /* [H*W] in (integrated) / out: cached discontinuity weights */
  double* wv;
  double energy_old;                /* in: IntVars.energy_old (read when integrated) */
  double* variance;                 /* mpba_integration_variances only: [H*W] out, y = Hessian^-1 1 (see below) */
} mpba_integration_job;

typedef struct mpba_integration_result {
  int32_t updated;                  /* _integrate's return value: depth was replaced */
  int32_t integrated;               /* IntVars.integrated after the call */
  int32_t irls_steps;               /* IRLS steps run */
  int32_t cg_iterations;            /* total */
  double energy_old;                /* IntVars.energy_old after the call */
  double energy[MPBA_INT_MAX_IRLS + 1]; /* energy before the loop and after every step */
  int32_t cg_steps[MPBA_INT_MAX_IRLS];  /* CG iterations of every step */
} mpba_integration_result;

void mpba_integration_default_conf(mpba_integration_conf* conf);
int mpba_integration_workspace_bytes(int32_t num_jobs, const mpba_integration_job* jobs, size_t* bytes);
int mpba_integrate_depth(mpba_handle* h, int32_t num_jobs, const mpba_integration_job* jobs,
                         const mpba_integration_conf* conf, void* workspace, size_t workspace_bytes,
                         mpba_integration_result* results);
/* Variance propagation (integration.py:522-616: calculate_hessian + IntegrationUncertainty.solve).  The reference
 * factorises the Hessian of the integration energy at the current depth (fresh weights; sparse term unfiltered and only
 * when !ignore_depths) with cholespy and, for every query pixel, solves H x = e_pixel and takes x.sum() -- the COLUMN
 * SUM of H^-1.  H is symmetric, so all those numbers are entries of ONE vector y = H^-1 1: job.variance receives y
 * (PCG to relative residual var_tol), the caller reads y[round(y_px) * W + round(x_px)] and multiplies by
 * depth_prior_at_kps^2 (:612-615).  result.cg_iterations = PCG iterations. */
int mpba_integration_variances(mpba_handle* h, int32_t num_jobs, const mpba_integration_job* jobs,
                               const mpba_integration_conf* conf, int32_t ignore_depths, double var_tol, void* workspace,
                               size_t workspace_bytes, mpba_integration_result* results);

/* After a call with opt->profile: per-kernel summed device ms and launch counts (HOST arrays of
 * MPBA_NUM_KERNELS); names from mpba_kernel_name(slot). */
int mpba_kernel_times(const mpba_handle* h, double* ms, int64_t* launches);
const char* mpba_kernel_name(int slot);

#ifdef __cplusplus
}
#endif
#endif /* MPBA_H_ */
