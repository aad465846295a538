/* acino_b200 — C ABI of the B200-native AcinoSet TRI + SBA hot path.
 *
 * The reference (DenDen047/AcinoSet) has no FFI: its seam is Python function injection
 * (src/lib/app.py:123-138 passes triangulate_func / project_func into src/lib/sba.py and
 * src/lib/utils.py). Each entry point below names the reference function whose arithmetic it
 * replaces; acinoset_b200/{calib,utils,sba}.py bind them with ctypes behind the reference's own
 * signatures. Plain pointers and sizes only; every function returns 0 on success or a negative
 * ACINO_E* code (message via acino_last_error()); nothing throws or exits.
 *
 * Camera block ("cam24"): 24 doubles per camera
 *   [0] fx [1] fy [2] cx [3] cy [4] alpha=K01/K00 [5..8] k1..k4 [9..17] R row-major (world->camera)
 *   [18..20] t [21..23] reserved (0)            (scene arrays of src/lib/utils.py:55-74)
 * Dense observation store: point p = (frame, keypoint); slot (p, c) holds camera c's pixel:
 *   uv   double[P][C][2]      obs_mask uint8[P]  (bit c set <=> slot (p,c) is an observation)
 * All arithmetic is fp64. Pointers named *_dev are device pointers on the current CUDA device.
 */
#ifndef ACINO_B200_H
#define ACINO_B200_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define ACINO_CAM_STRIDE 24
#define ACINO_MAX_CAMS 8

#define ACINO_OK 0
#define ACINO_EINVAL (-1)   /* bad argument */
#define ACINO_ECUDA (-2)    /* CUDA runtime error (no device, launch failure, OOM) */
#define ACINO_ESTATE (-3)   /* call order violated (e.g. solve before set_observations) */
#define ACINO_ENCCL (-4)    /* NCCL unavailable or failed */

const char* acino_last_error(void);
int acino_version(void);
int acino_device_count(void);           /* <0 on error, 0 when no CUDA device is visible */
int acino_set_device(int device);

/* device memory helpers so that a ctypes host can keep inputs resident in HBM */
int acino_dev_alloc(void** out_dev, int64_t bytes);
int acino_dev_free(void* dev);
int acino_h2d(void* dst_dev, const void* src_host, int64_t bytes);
int acino_d2h(void* dst_host, const void* src_dev, int64_t bytes);
int acino_sync(void);
/* The same copies on a stream of the caller's (no counterpart in the reference: plumbing). With page-locked host memory
 * they return at once, so a host can upload the next scene while the solver works on the current one
 * (dense.ScenePipeline.run_many); with pageable memory they behave like acino_h2d / acino_d2h. */
int acino_stream_create(void** out_stream);
int acino_stream_destroy(void* stream);
int acino_stream_sync(void* stream);
int acino_h2d_async(void* dst_dev, const void* src_host, int64_t bytes, void* stream);
int acino_d2h_async(void* dst_host, const void* src_dev, int64_t bytes, void* stream);

/* K1 — project_points_fisheye, src/lib/calib.py:132-136 (cv::fisheye::projectPoints).
 * X double[N][3] -> uv double[N][2]. Host pointers. */
int acino_project(const double* X, int64_t N, const double* cam24, double* uv);
int acino_project_dev(const double* X_dev, int64_t N, const double* cam24_host, double* uv_dev, void* stream);
/* K1 for all cameras in one launch — the projection loop of save_3d_cheetah_as_2d, src/lib/utils.py:255-263.
 * X double[N][3] -> uv double[n_cams][N][2]. width > 0 applies the reference's range rule (a pixel outside
 * [0,width] x [0,height] becomes NaN, utils.py:262-263); width <= 0 leaves pixels as projected. Host pointers. */
int acino_project_all(const double* X, int64_t N, const double* cams24, int n_cams, double width, double height, double* uv);

/* K1 — cost_func_points_only / cost_func_points_extrinsics, src/lib/sba.py:142-153, on observation
 * lists: f[2*o], f[2*o+1] = proj(X[point_idx[o]], cam cam_idx[o]) - points_2d[o]. Host pointers. */
int acino_residuals_list(const double* X, int64_t n_points, const int32_t* point_idx, const int32_t* cam_idx,
                         const double* points_2d, int64_t n_obs, const double* cams24, int n_cams, double* f);

/* Allocation-free residual vector + robust cost of an observation list. This entry point has NO counterpart in the
 * reference; it was added for callers that cannot allocate (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid cudaMalloc). It calls no cudaMalloc / cudaFree / allocating
 * library routine: the per-block partial sums of the cost live in a file-scope device array of the library (128.25 KB,
 * held by every GPU that loads the library until the process exits), and the cameras travel in the launch parameters.
 *   f_dev[2*o], f_dev[2*o+1] = proj(X[point_idx[o]], cam cam_idx[o]) - points_2d[o]   (as acino_residuals_list; may be NULL)
 *   cost_dev[0] = 0.5 f_scale^2 sum log1p((f / f_scale)^2) over all 2 n_obs residuals (scipy's Cauchy cost, deterministic sum)
 * Limits: n_obs <= 4,194,304 and 1 <= n_cams <= 8 per call (larger sizes are rejected with ACINO_EINVAL and a message).
 * All pointers are DEVICE pointers on the current device except cams24_host. The work is enqueued on `stream` and not
 * waited for. An observation whose point or camera index is out of range gives NaN residuals and a NaN cost. The first
 * call on a device looks up the scratch array's device address (cudaGetSymbolAddress): make one call before a graph
 * capture. All calls share the scratch array, which keeps the last call's contents (the entry point clears what must
 * start at zero): calls that OVERLAP on one GPU race on it - issue them on one stream or serialise them. */
int acino_residual_cost_noalloc(const double* X_dev, int64_t n_points, const int32_t* point_idx_dev,
                                const int32_t* cam_idx_dev, const double* points_2d_dev, int64_t n_obs,
                                const double* cams24_host, int n_cams, double f_scale, double* f_dev, double* cost_dev,
                                void* stream);

/* The analytic Jacobian of that residual (the reference has none: scipy differentiates cost_func_points_extrinsics,
 * src/lib/sba.py:142-146, by forward differences), for inspection and tests: per observation J double[2][9] =
 * d(u, v) / d(w, tau, X), (w, tau) the world-frame twist of the camera (R <- R Exp(w), t <- t + R tau) the solver
 * steps in, X the point. Host pointers. */
int acino_jacobian_list(const double* X, int64_t n_points, const int32_t* point_idx, const int32_t* cam_idx, int64_t n_obs,
                        const double* cams24, int n_cams, double* J);

/* N3 — the residual vector of adjust_extrinsics_manual_points, src/lib/calib.py:224-256, for n_batch candidate
 * corrections at once (the current (rvec, t) and scipy's forward differences). img_pts double[n_points][n_cams][2];
 * cameras in correct_cams take (r R^T, t - r t_corr); every pair (a, b) of cam_pairs triangulates all points (K2), every
 * camera reprojects them (K1), errors pass the reference's redescending loss (src/lib/misc.py:329-343, a<b<c)
 * elementwise. f double[n_batch][n_cams][n_pairs][n_points][2] = the reference's flatten order. Host pointers. */
int acino_manual_residuals(const double* img_pts, int n_points, const double* cams24, int n_cams, const int32_t* correct_cams,
                           int n_correct, const int32_t* cam_pairs, int n_pairs, const double* params, int n_batch,
                           double loss_a, double loss_b, double loss_c, double* f);

/* N3 — threshold search of initialize_marker_3d, src/lib/app.py:44-57: crit[p] = max over ring pairs of
 * min(lik[p][c], lik[p][c+1]) = the supremum of thresholds t for which `likelihood > t` still triangulates point p
 * (-inf when no pair exists; NaN likelihoods never pass). lik double[P][n_cams]. Host pointers. */
int acino_ring_critical(const double* lik, int64_t P, int n_cams, double* crit);

/* N4 — the measurement of the EKF update, src/core/ekf.py:66-96,245-251: the skeleton's markers for the pose x
 * (get_3d_marker_coords, src/lib/misc.py:144-327; mode 0 default / 1 head / 2 upper_body / 3 head_stabilize,
 * n_pose = 29 / 6 / 11 / 11) projected into every camera, h double[n_cams*n_markers*2] in the reference's order
 * [camera][marker][x,y], and the forward-difference Jacobian of numerical_jacobian (eps as given, 1e-3 in the
 * reference) H double[n_cams*n_markers*2][n_states], columns >= n_pose zero. coords (optional) receives the
 * markers' 3-D positions double[n_markers][3]. Host pointers. */
int acino_ekf_measurement(const double* x, int n_pose, int mode, const double* cams24, int n_cams, double eps, int n_states,
                          double* h, double* H, double* coords);

/* cv::Rodrigues(rvec)->R for n vectors (src/lib/sba.py:29), the device routine the extrinsics solver uses. */
int acino_rodrigues(const double* rvecs, int n, double* R_out);

/* K2 — triangulate_points_fisheye, src/lib/calib.py:120-129: fisheye undistort (Newton, <=10 steps,
 * eps 1e-8) of both pixels, 4x4 homogeneous DLT with P=[R|t], dehomogenise. uv1, uv2 double[M][2]. */
int acino_triangulate_pair(const double* uv1, const double* uv2, int64_t M, const double* cam24_a,
                           const double* cam24_b, double* xyz);

/* K2 — get_pairwise_3d_points_from_df, src/lib/utils.py:319-349, fused with the caller's
 * `likelihood > dlc_thresh` (src/core/tri.py:41, strict >): ring pairs (c,(c+1)%C), unweighted mean of
 * the pair estimates; a point exists iff some ring pair has both views. Outputs xyz double[P][3]
 * (NaN when absent), exists uint8[P], and obs_mask uint8[P] = thresholded slots of existing points (0 for
 * absent points): exactly the observations the inner merge at src/lib/sba.py:299 hands to SBA. */
int acino_triangulate_ring(const double* uv, const double* lik, double thresh, int64_t P, int C,
                           const double* cams24, double* xyz, uint8_t* obs_mask, uint8_t* exists);
int acino_triangulate_ring_dev(const double* uv_dev, const double* lik_dev, double thresh, int64_t P, int C,
                               const double* cams24_host, double* xyz_dev, uint8_t* obs_mask_dev,
                               uint8_t* exists_dev, void* stream);

/* ---- SBA solver handle: bundle_adjust_points_only (src/lib/sba.py:181-195) and
 * bundle_adjust_points_and_extrinsics (src/lib/sba.py:158-178) ---- */
typedef struct acino_sba acino_sba_t;

#define ACINO_MODE_POINTS_ONLY 0
#define ACINO_MODE_POINTS_AND_EXTRINSICS 1

typedef struct {
  int32_t mode;            /* ACINO_MODE_* */
  int32_t max_iter;        /* LM iterations (per point in points-only mode) */
  int32_t inner_iters;     /* points-only: LM iterations fused into one kernel launch (>=1) */
  int32_t force_iters;     /* !=0: never early-out on convergence (fixed work per iteration; bench) */
  double f_scale;          /* Cauchy scale C: cost = 0.5*C^2*sum(log1p((f/C)^2)), src/lib/sba.py:181 */
  double gtol;             /* stop when max|g| <= gtol */
  double xtol;             /* stop when an accepted step has max|dx| <= xtol*(max|x|+xtol) */
  double lambda0;          /* initial Marquardt damping */
  double ftol;             /* points+extrinsics: stop when an accepted step reduces the cost by <= ftol*cost */
  double hess_floor;       /* points+extrinsics: floor of the Gauss-Newton curvature weight relative to rho'
                              (rho' + 2 rho'' f^2 = rho' (1-z)/(1+z) is negative beyond |f| > C): 1.0 = plain IRLS weights */
  double point_refine_iters; /* points+extrinsics: from the second linearisation on, every point takes up to this many LM
                                iterations of its own with the cameras fixed, inside the linearisation kernel (each point has
                                its own damping, kept across iterations) */
  double point_refine_skip;  /* the refinement leaves a point alone when its model predicts a cost decrease
                              * <= point_refine_skip * max(cost of the point, f_scale^2/2); 0 = refine every point */
  double polish_tol;         /* points+extrinsics, acino_sba_solve: once an accepted step gains less than polish_tol
                              * (relative), and again before ftol may end the solve, the points whose own refinement ran
                              * out of steps are solved to convergence with the cameras fixed (96 LM iterations each);
                              * 0 = never */
  double lambda_dec;         /* points+extrinsics: the smallest factor by which an accepted step may shrink the damping
                              * (lambda <- lambda * max(lambda_dec, 1 - (2 rho - 1)^3)). Nielsen's rule is 1/3; the default
                              * 0.2 reaches the same stopping rule in about 10 % fewer iterations at a final cost a few
                              * 1e-6 (relative) higher (profiles/r02_lambda_dec_sweep.txt); 0 = default */
} acino_sba_opts;

typedef struct {
  int64_t n_points, n_obs;
  int64_t n_unconverged;   /* points (points-only) / 0-1 (extrinsics) not converged at exit */
  int32_t iterations;      /* LM iterations executed (sweeps) */
  int32_t launches;        /* CUDA kernel launches issued by the solve */
  double cost_initial, cost_final;   /* scipy-convention robust cost */
  double solve_ms;         /* device time of the solve (CUDA events) */
  double status;           /* points+extrinsics exit reason: 1 gtol, 2 ftol, 3 xtol, 4 stalled, 0 max_iter (7: peer exchange timed out, the solve returns an error) */
  double n_accept;         /* accepted LM steps */
  double lambda_final;     /* Marquardt damping at exit */
  double gmax;             /* max |gradient| at the last linearisation */
  double init_ms;          /* points+extrinsics: of solve_ms, the once-per-solve set-up (points sorted by observation mask,
                              observations in that layout, tiles; K3's visiting order). Host clock around the synchronised section */
  double polish_ms;        /* points+extrinsics: of solve_ms, the polishing launches (K3 to convergence, cameras fixed) */
  double syncs;            /* host synchronisations inside the solve with nothing queued behind them (the GPU idles while the
                            * host decides): set-up, polishes, the end. The per-iteration read of the LM state is made one
                            * iteration late, with the next iteration already queued, and is not counted */
} acino_sba_stats;

void acino_sba_default_opts(acino_sba_opts* o, int mode);
acino_sba_t* acino_sba_create(int n_cams, int64_t n_points);      /* NULL on error */
void acino_sba_destroy(acino_sba_t* h);
int acino_sba_set_cameras(acino_sba_t* h, const double* cams24);                    /* host */
int acino_sba_set_observations(acino_sba_t* h, const double* uv, const uint8_t* obs_mask);      /* host, copied */
int acino_sba_set_observations_dev(acino_sba_t* h, const double* uv_dev, const uint8_t* obs_mask_dev); /* borrowed */
int acino_sba_set_points(acino_sba_t* h, const double* xyz);                        /* host */
int acino_sba_set_points_dev(acino_sba_t* h, const double* xyz_dev);                /* d2d copy */
int acino_sba_get_points(acino_sba_t* h, double* xyz);
/* the same into a device buffer of the caller's (a copy the caller can download on a stream of its own while the handle
 * goes on to the next solve: dense.ScenePipeline.run_many) */
int acino_sba_get_points_dev(acino_sba_t* h, double* xyz_dev);
int acino_sba_get_cameras(acino_sba_t* h, double* cams24);
int acino_sba_get_rvec_t(acino_sba_t* h, double* rvecs, double* tvecs);             /* extrinsics mode parameters */
int acino_sba_set_rvec_t(acino_sba_t* h, const double* rvecs, const double* tvecs);
/* dense residuals f double[P][C][2] at the current parameters (0 in empty slots) */
int acino_sba_residuals(acino_sba_t* h, double* f);
int acino_sba_solve(acino_sba_t* h, const acino_sba_opts* opts, acino_sba_stats* stats);
/* one LM iteration over the whole scene on the handle's stream, no host sync (bench / roofline) */
int acino_sba_iterate(acino_sba_t* h, const acino_sba_opts* opts);
int acino_sba_reset_state(acino_sba_t* h);
void* acino_sba_stream(acino_sba_t* h);

/* per-kernel device times of acino_sba_iterate in points+extrinsics mode (CUDA events on the handle's stream):
 * enable, run <=256 iterations, read the averages in ms (build = K4a linearisation + K4b Schur accumulation,
 * backsub = K6). profile_read_kernels: ms4 = [K4a, K4b, fold + exchange + K5, K6]. */
int acino_sba_profile(acino_sba_t* h, int enable);
int acino_sba_profile_read(acino_sba_t* h, double* build_ms, double* backsub_ms, int* n_iterations);
int acino_sba_profile_read_kernels(acino_sba_t* h, double* ms4, int* n_iterations);
/* LM state after iterate / solve: out[0]=cost, [1]=lambda, [2]=gmax, [3]=iterations, [4]=accepted, [5]=status,
 * [6]=trial cost, [7]=predicted reduction, [8]=gain ratio, [9]=max |dx|, [10]=max |x|, [11]=nu, [12]=relative decrease of
 * the last accepted step, [13]=rejected, [14]=Cholesky failures, [15]=cost at the first linearisation */
int acino_sba_lm_state(acino_sba_t* h, double* out16);
/* The iteration log of the latest points+extrinsics acino_sba_solve - what scipy prints with verbose=2
 * (src/lib/sba.py:170: least_squares(..., verbose=2)): rows of 6 doubles [iteration, cost, cost reduction, max |step|,
 * max |gradient| (first-order optimality), accepted]; row 0 is the starting point. n_rows receives the number of rows
 * the log holds, at most cap_rows are copied. */
int acino_sba_history(acino_sba_t* h, double* rows6, int cap_rows, int* n_rows);

/* multi-GPU (frame-sharded) — one handle per rank (one process per GPU); NCCL is dlopen'ed from the process
 * (libnccl.so.2). Every rank calls solve / iterate collectively; the reduced camera system (6C x 6C + gradient + cost)
 * is all-reduced once per LM iteration and the trial scalars once more. comm_init also maps a small exchange buffer of
 * every peer (CUDA IPC over NVLink): when that succeeds on all ranks the two all-reduces are done by the kernels
 * themselves (each rank stores its partial sums into every peer's buffer, the consuming kernel waits for all ranks and
 * adds the rows in rank order; csrc/acino_sba_xchg.cuh), otherwise by one grouped NCCL launch each. ACINO_P2P=0 in the
 * environment forces the NCCL path. acino_sba_comm_mode: 0 single rank, 1 NCCL all-reduce, 2 peer-memory exchange. */
int acino_nccl_unique_id(void* id128);
int acino_sba_comm_init(acino_sba_t* h, const void* id128, int rank, int nranks);
int acino_sba_comm_mode(acino_sba_t* h);

/* fp64 FMA throughput of the device (dependent-chain-free DFMA microbenchmark), TFLOP/s: the second roofline
 * denominator bench.py reports next to the measured HBM bandwidth. */
int acino_fp64_peak(double* tflops);

#ifdef __cplusplus
}
#endif
#endif
