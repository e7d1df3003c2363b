/* sba.h — C-ABI of the B200-native stereo bundle adjuster (libsba.so).
 *
 * Drop-in boundary for the reference's BundleAdjuster (reference: src/bundle_adjuster.h:13-25):
 * everything that BundleAdjuster::Optimize hands to Ceres (src/bundle_adjuster.cpp:59-117) goes
 * through sba_solve() instead; the pack / unpack around it (src/bundle_adjuster.cpp:41-110,121-132)
 * is sba_pack_window() / sba_unpack_window().  Plain pointers and sizes only; no C++/torch types.
 *
 * Conventions (identical to the reference):
 *   poses   f64 [Nc*6]  = [angle-axis w (3), translation t (3)] of T_cw (camera-from-world)
 *                          (src/math_utils.h:12-25, src/bundle_adjuster.cpp:66-69)
 *   points  f64 [Np*3]
 *   obs_uv  f32 [Nobs*4] = (u_l, v_l, u_r, v_r)          (src/observation.h:11-15)
 *   p_l,p_r f64 [12] row-major 3x4 projection matrices    (src/reprojection_error.h:65-66)
 *   objective  1/2 sum_k huber(|| r_k ||^2), r_k the 4-vector of src/reprojection_error.h:35-38
 */
#ifndef SBA_H_
#define SBA_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct sba_handle_s* sba_handle;

typedef enum {
  SBA_OK = 0,
  SBA_NO_CONVERGENCE = 1,   /* iteration / time limit hit (results still written, like the reference) */
  SBA_NUMERICAL_FAILURE = 2,
  SBA_INVALID_ARGUMENT = 3,
  SBA_CUDA_ERROR = 4,
  SBA_NCCL_ERROR = 5
} sba_status;

/* termination reason per window */
enum {
  SBA_RUNNING = 0,
  SBA_TERM_PARAMETER_TOLERANCE = 1,
  SBA_TERM_FUNCTION_TOLERANCE = 2,
  SBA_TERM_GRADIENT_TOLERANCE = 3,
  SBA_TERM_MAX_ITERATIONS = 4,
  SBA_TERM_MAX_TIME = 5,
  SBA_TERM_MIN_RADIUS = 6,
  SBA_TERM_INVALID_STEPS = 7
};

/* replaces the ceres::Solver::Options the reference fills in its ctor
 * (src/bundle_adjuster.cpp:14-36) + the constants of src/params.h:34-47. */
typedef struct sba_options {
  int32_t max_iterations;        /* 50    src/params.h:34 */
  double  max_time_s;            /* 1.0   src/params.h:41 ; <= 0: unlimited (the :40 alternative) */
  double  huber_delta;           /* 1.0   src/bundle_adjuster.cpp:100 */
  double  point_lower_bound;     /* -1e4  src/params.h:44 */
  double  point_upper_bound;     /* +1e4  src/params.h:47 */
  double  parameter_tolerance;   /* 1e-8  Ceres default */
  double  function_tolerance;    /* 1e-16 src/bundle_adjuster.cpp:36 */
  double  gradient_tolerance;    /* 1e-16 src/bundle_adjuster.cpp:35 */
  int32_t residual_dim;          /* 4 = reference (src/reprojection_error.h:58); 3 = (u_l,v_l,u_r) only */
  double  initial_radius;        /* 1e4   Ceres default initial_trust_region_radius */
  int32_t max_pcg_iterations;    /* 500   PCG on the reduced camera system (replaces CHOLMOD, :26-29) */
  double  pcg_tolerance;         /* 1e-12 relative preconditioned-residual tolerance (tight: stands in for the exact solve) */
  int32_t verbose;               /* 0     (reference prints progress + FullReport, :15,:118) */
  int32_t linear_solver;         /* 0     reduced camera system: 0 auto = the reference's exact factorisation (:26-29): dense block
                                  *       Cholesky for windows of <= 64 free cameras, block-sparse Cholesky (nested dissection)
                                  *       above; PCG only if the pattern is too dense to factor.  1 PCG.  2 Cholesky or an error */
} sba_options;

typedef struct sba_summary {
  double  initial_cost;          /* summed over windows */
  double  final_cost;
  int32_t iterations;            /* LM iterations run (max over windows), excluding iteration 0 */
  int32_t successful_steps;      /* summed over windows */
  int32_t linearizations;
  int32_t pcg_iterations;        /* total */
  int32_t termination;           /* reason of window 0 (SBA_TERM_*) */
  int32_t num_windows_converged;
  double  time_total_s;          /* wall, whole call */
  double  time_structure_s;      /* one-off index / pattern build (host) */
  double  time_h2d_s, time_d2h_s;
  double  time_solve_s;          /* LM loop, wall */
  double  ms_linearize;          /* device time (CUDA events), summed over the solve */
  double  ms_schur;
  double  ms_pcg;
  double  ms_update;
} sba_summary;

void sba_default_options(sba_options* opt);
const char* sba_status_string(int status);
const char* sba_last_error(void);

/* ---- handle ------------------------------------------------------------------------------- */
/* One handle = one CUDA device = one host thread at a time.  No global state: device memory comes from a
 * stream-ordered pool owned by the handle. */
int sba_create(int device, sba_handle* out);
void sba_destroy(sba_handle h);

/* Problem structure (constant across LM iterations and across repeated solves of one window).
 *   obs_cam[k], obs_pt[k]  : camera / point index of residual block k, in the reference's scan order
 *                            (src/bundle_adjuster.cpp:62-102); duplicated (cam,pt) pairs allowed.
 *   cam_fixed[c]           : 1 = held constant.  NULL => first camera of every window
 *                            (src/bundle_adjuster.cpp:113).
 *   num_windows / win_*    : independent problems solved in one call, each with its own LM state.
 *                            win_cam_off / win_pt_off [num_windows+1]: cameras and points of window w
 *                            are the contiguous index ranges [off[w], off[w+1]).  NULL => one window.
 * Builds: observations sorted by point and by camera, pair list + block pattern of the reduced
 * camera matrix.  Returns SBA_INVALID_ARGUMENT on out-of-range indices. */
int sba_set_structure(sba_handle h, int32_t num_cameras, int32_t num_points, int64_t num_observations,
                      const int32_t* obs_cam, const int32_t* obs_pt, const uint8_t* cam_fixed,
                      int32_t num_windows, const int32_t* win_cam_off, const int32_t* win_pt_off);

/* Multi-GPU: this rank owns points [pt_begin, pt_end) (and all their observations); cameras are
 * replicated.  Must be called before sba_set_structure; needs sba_comm_init. */
int sba_set_point_shard(sba_handle h, int32_t pt_begin, int32_t pt_end);

/* Upload state + measurements (host pointers). */
int sba_set_state(sba_handle h, const double* poses, const double* points);
int sba_set_measurements(sba_handle h, const float* obs_uv, const double* p_l, const double* p_r);
int sba_get_state(sba_handle h, double* poses, double* points);

/* Run the LM loop on device-resident state. */
int sba_run(sba_handle h, const sba_options* opt, sba_summary* summary);

/* One-shot: structure + upload + LM + download.  poses / points are updated in place — also when the
 * result is SBA_NO_CONVERGENCE or SBA_NUMERICAL_FAILURE (the last ACCEPTED state), because the reference
 * never looks at Ceres' termination type and always writes back (:116-132); untouched on
 * SBA_INVALID_ARGUMENT / SBA_CUDA_ERROR / SBA_NCCL_ERROR. */
int sba_solve(sba_handle h, int32_t num_cameras, int32_t num_points, int64_t num_observations,
              double* poses, double* points, const int32_t* obs_cam, const int32_t* obs_pt,
              const float* obs_uv, const double* p_l, const double* p_r, const uint8_t* cam_fixed,
              int32_t num_windows, const int32_t* win_cam_off, const int32_t* win_pt_off,
              const sba_options* opt, sba_summary* summary);

/* ---- stage-level entry points (parity tests + bench) ------------------------------------------ */
/* K0+K1+K2: residuals, analytic Jacobians, Huber, U/V/W/g accumulation at the current state. */
int sba_linearize(sba_handle h, const sba_options* opt);
/* K3: damped Schur complement for trust-region radius `radius` (all windows). */
int sba_schur(sba_handle h, const sba_options* opt, double radius);
/* K4: PCG on the reduced system; returns iterations used. */
int sba_pcg(sba_handle h, const sba_options* opt, int32_t* iterations);
/* repeat (linearize + schur) `repeats` times; device ms per stage averaged (CUDA events). */
int sba_time_linearize_schur(sba_handle h, const sba_options* opt, double radius, int32_t repeats,
                             double* ms_linearize, double* ms_schur);

/* Time one stage in isolation (bench.py's roofline leg): `repeats` launches, each timed with CUDA
 * events on the handle's stream; flush_l2 != 0 overwrites a 256 MiB scratch buffer between launches
 * so every launch starts from a cold L2.  Stages: 0 k_lin_chunk (K0+K1+K2) 1 k_assemble_schur 2 k_schur_chunk (K3)
 * 3 whole linearisation 4 whole Schur step 5 k_backsub 6 k_trial_points 7 one multi-launch PCG iteration
 * 8 linearisation + Schur (one bench step; on a single GPU the step is replayed as one CUDA graph) 9 k_assemble_cams 10 k_lin_finish
 * 11 k_fused_chunk (K1-K3 of every chunk in one kernel: what a bench step and the LM loop run from the second linearisation on).  Needs a prior sba_linearize (and
 * sba_schur for stages 1, 5-7).  *ms = mean device milliseconds per launch. */
int sba_time_stage(sba_handle h, const sba_options* opt, double radius, int32_t stage, int32_t repeats,
                   int32_t flush_l2, double* ms);

/* FP64 FMA peak of the handle's device in TFLOP/s (register-only DFMA chains; best of 5 launches): the second
 * roofline the fp64 kernels are reported against. */
int sba_measure_fp64_peak(sba_handle h, double* tflops);

/* Introspection, all host output buffers, any may be NULL.  Orders are the caller's own:
 *   r [Nobs*4] corrected residuals in input observation order; W [Nobs*18] (6x3 row-major) likewise;
 *   U [Nc*36] full symmetric 6x6; V [Np*9]; gc [Nc*6]; gp [Np*3]; cost [num_windows]. */
int sba_get_linearization(sba_handle h, double* r, double* U, double* V, double* W, double* gc,
                          double* gp, double* cost);
/* Block pattern of the reduced camera matrix over FREE cameras (full: both triangles), sorted by
 * (row, col).  Returns the block count; fills rows/cols up to `capacity`. */
int64_t sba_get_schur_pattern(sba_handle h, int32_t* rows, int32_t* cols, int64_t capacity);
/* Values of those blocks (36 doubles each, row-major), rhs [6*Nfree], and PCG solution [6*Nfree]. */
int sba_get_schur_values(sba_handle h, double* blocks, double* rhs, double* solution);
/* sorted views: obs index (input order) at each position of the point-sorted / camera-sorted lists */
int sba_get_sorted_views(sba_handle h, int32_t* by_point, int32_t* by_camera);
/* Symbolic phase of the exact block-sparse Cholesky used when a window has more than 64 free cameras (the global BA
 * calls, src/slam.cpp:156; stands in for CHOLMOD's analysis, src/bundle_adjuster.cpp:26-29).  info = {block columns n,
 * blocks in L, block products, levels, chains}; call once with NULL arrays for the sizes.  perm [n] new -> old
 * free-camera index; col_ptr [n+1] / row_idx [nnzL] pattern of L (diagonal first); init_src [nnzL] block of
 * sba_get_schur_pattern's UPPER blocks that lands there (b, ~b = transposed, INT32_MIN = fill); upd_ptr [nnzL+1] /
 * upd_a, upd_b [nupd] products L_a L_b^T each block receives; lvl_ptr [levels+1] / lvl_cols, lvl_len [chains]: level sets
 * of the elimination tree's chains (columns lvl_cols[c] .. lvl_cols[c] + lvl_len[c] - 1, each the only child of the next).
 * SBA_INVALID_ARGUMENT when the structure has no sparse factorisation. */
int sba_get_factor_symbolic(sba_handle h, int64_t info[5], int32_t* perm, int32_t* col_ptr, int32_t* row_idx,
                            int32_t* init_src, int64_t* upd_ptr, int32_t* upd_a, int32_t* upd_b, int32_t* lvl_ptr,
                            int32_t* lvl_cols, int32_t* lvl_len);
int32_t sba_num_free_cameras(sba_handle h);
/* number of kernels launched by this handle so far (bench.py's gpu_launches) */
int64_t sba_launch_count(sba_handle h);
/* which kernel runs K1-K3 of the chunks after the first linearisation of the current structure (diagnostics: bench.py names
 * its dominant kernel with it): 0 the two-kernel path (k_lin_chunk + k_schur_chunk), 1 k_fused_chunk (256 threads),
 * 2 k_fused_chunk (128 threads), 3 k_fused_persist */
int32_t sba_fused_kernel_variant(sba_handle h);

/* ---- multi-GPU (NCCL over NVLink) -------------------------------------------------------------- */
/* id: 128-byte ncclUniqueId made by rank 0 with sba_comm_unique_id and broadcast by the caller. */
int sba_comm_unique_id(uint8_t id[128]);
int sba_comm_init(sba_handle h, int32_t rank, int32_t world_size, const uint8_t id[128]);

/* ---- device-resident map ------------------------------------------------------- */
/* The reference's BundleAdjuster holds references to the SLAM map (m_cam_frames / m_ldm_points,
 * src/bundle_adjuster.h:21-22) and is called once per frame on (n-1, n) and every refine_interval frames on a sliding
 * window (src/slam.cpp:122-129).  With these entry points the map lives inside the handle between calls: map points and
 * stereo pixels stay in HBM (only a new frame's observations and points are uploaded, once), the integer side (which
 * point an observation sees) and the poses are mirrored on the host, where the reference's index build (:57-102) and
 * its f32 pose (de)parameterisation (src/math_utils.h:12-41) run.  Results of sba_map_optimize are bit-identical to
 * sba_pack_window + sba_solve + sba_unpack_window on the same map.  Single GPU. */
int sba_map_reset(sba_handle h);
/* VisualOdometer::Update (src/visual_odometer.cpp:335-444): a new frame with its pose T_wc (f32 4x4 row-major), its
 * observations (global point id + u_l v_l u_r v_r) and the points it creates (ids continue the map's numbering). */
int sba_map_append_frame(sba_handle h, const float* T_wc16, int64_t nobs, const uint32_t* obs_point_id, const float* obs_uv,
                         int32_t num_new_points, const float* new_points_xyz);
int64_t sba_map_num_frames(sba_handle h);
int64_t sba_map_num_points(sba_handle h);
/* BundleAdjuster::Optimize(start, end) on the resident map (src/bundle_adjuster.cpp:39-133): half-open window, first
 * camera fixed, every window camera and point written back (poses through PoseToMatrix / inverse / Normalize as the
 * Frame::GlobalPose setter does, src/camera_frame.h:32-49; points narrowed to f32, src/map_point.h:30-35). */
int sba_map_optimize(sba_handle h, int32_t start, int32_t end, const double* p_l, const double* p_r, const sba_options* opt,
                     sba_summary* summary);
int sba_map_get_poses(sba_handle h, int32_t first, int32_t count, float* T_wc16);
int sba_map_set_pose(sba_handle h, int32_t frame, const float* T_wc16);          /* Frame::GlobalPose(pose) by other callers */
int sba_map_get_points(sba_handle h, int64_t first, int64_t count, float* xyz);
int sba_map_set_points(sba_handle h, int64_t first, int64_t count, const float* xyz);
/* ids of the points the last sba_map_optimize wrote back (returns their number) and its phase times in ms:
 * {index build + poses, structure, state on device, solve, unpack}. */
int64_t sba_map_last_window(sba_handle h, uint32_t* point_ids, int64_t capacity, double* ms5);

/* ---- pose-graph optimisation ---------------------------------------------------- */
/* What PoseGraphOptimizer::Optimize hands to g2o (src/pose_graph_optimizer.cpp:34-96): SE3 vertices = the frames'
 * world-from-camera poses, edges (i, j) with measurement Z = X_i^-1 X_j (odometry :143-168: Frame::RelativePose; loop
 * closures :192-211), information diag(0.01, 0.01, 0.01, 1, 1, 1) (translation first, :23-26), Huber kernel (:162),
 * vertex 0 fixed (:123-126), 10 Levenberg iterations (:66).  Poses are 4x4 row-major f64 (g2o works in double on the
 * Frames' float poses), updated in place.  fixed NULL = vertex 0; information6 NULL = the reference's diagonal. */
typedef struct sba_pgo_summary {
  double  initial_chi2, final_chi2, final_lambda;
  int32_t iterations, trials;    /* Levenberg iterations run; linear solves tried (g2o retries with a larger lambda) */
  int32_t termination;           /* 0 iteration limit, 1 g2o's Terminate (10 failed trials / zero gain / lambda overflow) */
  double  time_total_s, time_symbolic_s;
} sba_pgo_summary;
int sba_pgo_optimize(sba_handle h, int32_t num_vertices, double* T_wc, int32_t num_edges, const int32_t* edge_i, const int32_t* edge_j,
                     const double* Z, const uint8_t* fixed, const double* information6, double huber_delta, int32_t iterations,
                     sba_pgo_summary* summary);

/* ---- N4: stereo RANSAC rigid alignment ---------------------------------------------------------- */
/* Replaces RANSAC::Solver<ObservationPair, Eigen::Matrix4f>::Solve over StereoTransModel
 * (ransac_solver/include/ransac_solver.h:51-166; ransac_solver/include/stereo_reprojection.h:28-134), which
 * VisualOdometer::CalcTransformation (src/visual_odometer.cpp:93-103) calls once per frame and the loop closer once
 * per candidate (src/pose_graph_optimizer.cpp:189).  Options = RANSAC::Option (ransac_solver.h:13-19) with the values
 * src/params.h:18-27 sets, plus the model's pixel threshold (stereo_reprojection.h:126). */
typedef struct sba_ransac_options {
  int32_t max_iterations;        /* 100 */
  int32_t final_model_fitting;   /* 0 */
  int32_t early_termination;     /* 0 */
  float   concensus_ratio;       /* 0.8 (the reference's spelling) */
  float   inlier_threshold;      /* 3.0 px on sqrt(du_l^2 + dv_l^2 + du_r^2 + dv_r^2) */
} sba_ransac_options;
void sba_ransac_default_options(sba_ransac_options* o);
/* num_problems independent alignments in one call; problem b owns elements [offsets[b], offsets[b+1]).
 *   obs_first / obs_second [n*4] f32 (u_l, v_l, u_r, v_r) = ObservationPair::first / ::second: the model maps the
 *     points triangulated from `first` onto those from `second`
 *   P_l / P_r [12] f32 row-major 3x4 projection matrices (StereoTransModel::SetCameraParams)
 *   samples [num_problems * max_iterations * 3] i32 element indices local to the problem: the three draws
 *     `rand() % size` of every iteration (ransac_solver.h:84-90), made by the caller so that its generator — the
 *     reference's rand() included — reproduces; NULL = a counter-based draw from `seed`
 * Outputs: T [num_problems*16] f32 row-major (Solve's return value), labels [n] u8 and losses [n] f32 (its `labels`
 * / `losses` out-parameters); optional best_iteration [num_problems] and counts [num_problems*max_iterations]
 * (inlier count of every hypothesis; -1 = not run because of early termination).
 * Element weights are not an argument: the reference's models ignore them.  A problem with no elements gets T = 0.
 * CalcTransformation's "fewer than 10 pairs -> false" guard belongs to the caller. */
int sba_ransac_stereo(sba_handle h, int32_t num_problems, const int64_t* offsets, const float* obs_first, const float* obs_second,
                      const float* P_l, const float* P_r, const sba_ransac_options* options, const int32_t* samples, uint64_t seed,
                      float* T, uint8_t* labels, float* losses, int32_t* best_iteration, int32_t* counts);
/* triangulated points of the last sba_ransac_stereo call's inputs are not kept; this evaluates cv::triangulatePoints
 * + division by w (stereo_reprojection.h:79-92) for n stereo observations: xyz [n*3] f32 */
int sba_triangulate_stereo(sba_handle h, int64_t n, const float* obs, const float* P_l, const float* P_r, float* xyz);

/* ---- allocation-free variant of the stereo RANSAC alignment -------------------------------------- */
/* NO COUNTERPART IN THE REFERENCE: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same computation and
 * same kernels as sba_ransac_stereo, but:
 *   - inputs and outputs are DEVICE pointers supplied by the caller (d_offsets [num_problems+1] i64, d_obs_first /
 *     d_obs_second [num_elements*4] f32 16-byte aligned, d_samples [num_problems*max_iterations*3] i32 or NULL = draw
 *     from `seed`; d_T [num_problems*16] f32, d_labels [num_elements] u8, d_losses [num_elements] f32, optional
 *     d_best_iteration [num_problems] i32 and d_counts [num_problems*max_iterations] i32); P_l / P_r / options are
 *     small HOST arguments that travel as kernel parameters; sample indices are not validated (out-of-range -> 0);
 *   - the scratch (triangulated points of both observation sets, the models of all hypotheses, and the counts /
 *     chosen iteration when the caller passes NULL for them) lives in ONE file-scope device array of fixed size
 *     (about 3 MiB, held by every GPU that loads the library until exit); problems beyond
 *     SBA_NOALLOC_MAX_ELEMENTS elements in total or SBA_NOALLOC_MAX_HYPOTHESES = num_problems * max_iterations
 *     hypotheses are rejected with SBA_INVALID_ARGUMENT;
 *   - the call allocates nothing on the device or in pinned memory (no cudaMalloc* / cudaFree* / cudaHostAlloc, no
 *     memory pool); it enqueues two memsets and three kernels on `stream` (a cudaStream_t; NULL = the handle's own
 *     stream, and then the call also waits for completion) and otherwise returns without synchronising.
 * Every call on a GPU shares the array and finds the previous call's contents in it (what must start at zero is
 * cleared by the call itself): calls that overlap on the same GPU — different streams, different handles or
 * processes' threads — RACE.  Serialise them. */
#define SBA_NOALLOC_MAX_ELEMENTS   65536
#define SBA_NOALLOC_MAX_HYPOTHESES 16384
int sba_ransac_stereo_noalloc(sba_handle h, int32_t num_problems, int64_t num_elements, const int64_t* d_offsets,
                              const float* d_obs_first, const float* d_obs_second, const float* P_l, const float* P_r,
                              const sba_ransac_options* options, const int32_t* d_samples, uint64_t seed, float* d_T,
                              uint8_t* d_labels, float* d_losses, int32_t* d_best_iteration, int32_t* d_counts, void* stream);

/* ---- host side of BundleAdjuster::Optimize ------------------------------------------------------ */
/* Pack window [start,end) of a map (reference: src/bundle_adjuster.cpp:41-110).
 *   T_wc [N*16] f32 row-major world-from-camera (Frame::GlobalPose), map_points [M*3] f32,
 *   frame_off [N+1] offsets into obs_point_id / obs_uv_in (per-frame observation lists, stored flat).
 * Outputs (caller-allocated; capacities: poses 6*(end-start), the rest sized by the window's
 * observation count): point numbering = order of first appearance, bit-exact with the reference.
 * Returns number of points via *num_points. */
int sba_pack_window(const float* T_wc, const float* map_points, const int64_t* frame_off,
                    const uint32_t* obs_point_id, const float* obs_uv_in, int32_t start, int32_t end,
                    double* poses, double* points, uint32_t* point_ids, int32_t* obs_cam,
                    int32_t* obs_pt, float* obs_uv, int32_t* num_points);
/* Write back (reference: src/bundle_adjuster.cpp:121-132 + Frame::GlobalPose setter,
 * src/camera_frame.h:32-49): T_wc rows [start,end) and the window's map points are overwritten. */
int sba_unpack_window(float* T_wc, float* map_points, int32_t start, int32_t end, const double* poses,
                      const double* points, const uint32_t* point_ids, int32_t num_points);
/* MatrixToPose / PoseToMatrix (src/math_utils.h:12-41), float arithmetic like the reference. */
void sba_matrix_to_pose(const float* T16, double* pose6);
void sba_pose_to_matrix(const double* pose6, float* T16);
/* float 4x4 helpers the Frame::GlobalPose setter needs (src/camera_frame.h:32-49): general inverse
 * (Matrix4f::inverse), Normalize (src/math_utils.h:6-10), product out = A*B.  Row-major. */
int sba_invert_pose(const float* T16, float* out16);
void sba_normalize_pose(float* T16);
void sba_compose_pose(const float* A16, const float* B16, float* out16);

#ifdef __cplusplus
}
#endif
#endif /* SBA_H_ */
