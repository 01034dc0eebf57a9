/* mcc.h -- C ABI of the B200-native multicam_calibration solve path (libmcc.so).
 *
 * The reference has no FFI layer: its boundary for this path is the C++ class
 * multicam_calibration::Calibrator (/root/reference/include/multicam_calibration/calibrator.h:17-49), which
 * hands the problem to Ceres (/root/reference/src/calibrator.cpp:285-313).  Every entry point below cites the
 * Calibrator member (or Ceres call site) it replaces.  include/multicam_calibration/calibrator.h is the C++
 * facade with the reference's own method names on top of this ABI; INTEGRATION.md shows the binding.
 *
 * Conventions: plain pointers and sizes only; every input is copied before the call returns; matrices are
 * row-major; 4x4 transforms are 16 doubles row-major; one caller thread per handle (the reference's nodelet is
 * single-threaded, launch/calibration.launch:5-6); no exception crosses the ABI; every function returns an
 * mcc_status (0 = ok, <0 = error, text via mcc_last_error).  All arithmetic on the device is fp64.
 * There is NO CPU fallback: without a CUDA device the compute entry points return MCC_ERR_NO_DEVICE.
 */
#ifndef MCC_H_
#define MCC_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct mcc_handle mcc_handle;
typedef int mcc_status;

enum {
  MCC_OK = 0,
  MCC_ERR_NO_DATA = -1,    /* runCalibration's "no data to run on!" (calibrator.cpp:286-289) */
  MCC_ERR_BAD_ARG = -2,    /* addPoints' "number of cams != width of array!" (calibrator.cpp:236-240) and friends */
  MCC_ERR_BAD_MODEL = -3,  /* FrameResidual's "unknown distortion model" (calibrator.cpp:131-133) */
  MCC_ERR_CUDA = -4,
  MCC_ERR_NCCL = -5,
  MCC_ERR_NUMERICAL = -6,
  MCC_ERR_NO_DEVICE = -7,
  MCC_ERR_TOO_LARGE = -8
};

enum { MCC_MODEL_EQUIDISTANT = 0, MCC_MODEL_RADTAN = 1 };
#define MCC_MAX_CAMERAS 32
#define MCC_MAX_DIST 8
#define MCC_MAX_TRACE 256
#define MCC_UNIQUE_ID_BYTES 128

/* One camera = CalibrationData (calibration_data.h:20-36) restricted to what the solve reads. */
typedef struct mcc_camera {
  int model;               /* MCC_MODEL_*  (CameraIntrinsics::distortion_model) */
  int n_intr;              /* must be 4: fx, fy, cx, cy (camera_intrinsics.h:16) */
  int n_dist;              /* 0..8 distortion coefficients */
  const double* intr;      /* [n_intr] */
  const double* dist;      /* [n_dist] */
  double T_cn_cnm1[16];    /* ignored for camera 0 */
  int fix_intr, fix_extr, active;
} mcc_camera;

/* ceres::Solver::Options as the reference sets them (calibrator.cpp:300-305) plus the Ceres defaults that
 * are in force; mcc_default_options() fills exactly those. */
typedef struct mcc_options {
  int max_num_iterations;            /* 200 */
  double function_tolerance;         /* 1e-12 */
  double parameter_tolerance;        /* 1e-12 */
  double gradient_tolerance;         /* 1e-10 */
  double initial_trust_region_radius;/* 1e4 */
  double max_trust_region_radius;    /* 1e16 */
  double min_trust_region_radius;    /* 1e-32 */
  double min_relative_decrease;      /* 1e-3 */
  double min_lm_diagonal;            /* 1e-6 */
  double max_lm_diagonal;            /* 1e32 */
  int max_num_consecutive_invalid_steps; /* 5 */
  int jacobi_scaling;                /* 1 */
  int use_loss;                      /* 1: HuberLoss(1.0) per frame block (calibrator.cpp:506) */
  int verbose;                       /* 1: print Ceres-style progress rows (minimizer_progress_to_stdout) */
  int keep_params;                   /* 0: params_.clear(); initializeVariables() first, as runCalibration does;
                                        1: start from the parameters currently held by the handle */
} mcc_options;

/* ceres::Solver::Summary, the parts FullReport() prints (calibrator.cpp:312) + per-iteration rows. */
typedef struct mcc_summary {
  double initial_cost, final_cost;
  int num_iterations;      /* "Minimizer iterations": logged rows, row 0 included */
  int num_successful, num_unsuccessful;
  int termination;         /* 0 CONVERGENCE, 1 NO_CONVERGENCE, 2 FAILURE */
  int why;                 /* 0 gradient, 1 parameter, 2 function tolerance, 3 max iterations, 4 invalid steps,
                              5 min radius, 6 linear solver, 7 no data */
  double why_value;
  int num_params, num_reduced_params;
  long long num_residuals;
  double t_upload, t_jacobian, t_linear, t_residual, t_total; /* host wall clock, seconds */
  int n_jacobian_evals, n_residual_evals, n_linear_solves;
  long long n_kernel_launches; /* kernels of this library launched by this call */
  long long h2d_bytes, d2h_bytes;
  int trace_len;
  double tr_cost[MCC_MAX_TRACE], tr_cost_change[MCC_MAX_TRACE], tr_gradient_max[MCC_MAX_TRACE],
      tr_step_norm[MCC_MAX_TRACE], tr_ratio[MCC_MAX_TRACE], tr_radius[MCC_MAX_TRACE], tr_iter_time[MCC_MAX_TRACE];
} mcc_summary;

const char* mcc_version(void);
void mcc_default_options(mcc_options* o);
int mcc_summary_size(void);

/* new Calibrator (calibration_nodelet.cpp:55) / destructor */
mcc_status mcc_create(mcc_handle** out);
void mcc_destroy(mcc_handle* h);
const char* mcc_last_error(const mcc_handle* h);

/* CUDA device ordinal this handle computes on (default: the current device at first use). */
mcc_status mcc_set_device(mcc_handle* h, int device);

/* Several GPUs behind ONE handle, one process (no reference counterpart: the reference is single-device; its caller,
 * calibration_nodelet.cpp:196-208 / 642, is one process that calls addPoints per frame and then runCalibration, and with this
 * call it keeps doing exactly that).  Call it before frames are added.  The handle then deals the frames out over the
 * devices (whole frames, levelled by observation count; page-locked arrays given to mcc_add_frames go to the devices over
 * their own PCIe links at the same time), runs every compute call on one host thread per device, and sums the reduced
 * camera system over the devices through peer windows (cudaDeviceEnablePeerAccess; devices without peer access to each
 * other are refused at the first computation with MCC_ERR_NCCL).  Results, parameter layout, residual and mask order are
 * those of a single-device handle.  ndev == 1 is mcc_set_device; at most 8 devices.  Not combinable with mcc_comm_init. */
mcc_status mcc_set_devices(mcc_handle* h, int ndev, const int* device_ids);

/* Calibrator::setCameras (calibrator.cpp:225-227) */
mcc_status mcc_set_cameras(mcc_handle* h, int ncam, const mcc_camera* cams);
/* Calibrator::setFixAllIntrinsics (calibrator.h:22) */
mcc_status mcc_set_fix_all_intrinsics(mcc_handle* h, int fix);
/* the per-camera flags the set_parameter service changes (calibration_nodelet.cpp:257-285) */
mcc_status mcc_set_camera_flags(mcc_handle* h, int cam, int fix_intr, int fix_extr, int active);

/* Calibrator::addPoints (calibrator.cpp:229-249): one frame; world_xyz[k] -> 3*npts[k] doubles, image_uv[k] ->
 * 2*npts[k] doubles, T_c0_w = cam0 pose guess.  ncam must equal the camera count (else MCC_ERR_BAD_ARG). */
mcc_status mcc_add_frame(mcc_handle* h, int ncam, const int* npts, const double* const* world_xyz,
                         const double* const* image_uv, const double* T_c0_w);
/* The same for many frames at once, flat: npts[nframes*ncam]; world_xyz/image_uv hold all observations
 * frame-major, camera-minor, in insertion order; poses[nframes*16]. */
mcc_status mcc_add_frames(mcc_handle* h, int nframes, int ncam, const int* npts, const double* world_xyz,
                          const double* image_uv, const double* poses);
mcc_status mcc_clear_frames(mcc_handle* h);
long long mcc_num_frames(const mcc_handle* h);
long long mcc_num_observations(const mcc_handle* h);

/* Calibrator::initializeVariables (calibrator.cpp:251-283): (re)pack params from cameras + pose guesses. */
mcc_status mcc_init_params(mcc_handle* h);
/* params_ (calibrator.h:47), layout of calibrator.cpp:36-46: [intr,dist per cam | rvec,t of T_cn_c0, n>=1 |
 * rvec,t of the cam0 pose per frame]. */
long long mcc_num_params(const mcc_handle* h);
mcc_status mcc_get_params(const mcc_handle* h, double* out, size_t n);
mcc_status mcc_set_params(mcc_handle* h, const double* in, size_t n);

/* Stage observations, masks and parameters on the device now (otherwise mcc_run does it on first use). */
mcc_status mcc_upload(mcc_handle* h);

/* Calibrator::runCalibration (calibrator.cpp:285-313) = ceres::Solve on the frame-block problem of
 * setupOptimizationProblem (calibrator.cpp:466-523).  opts NULL = the reference's options. */
mcc_status mcc_run(mcc_handle* h, const mcc_options* opts, mcc_summary* out);

/* Calibrator::getCalibrationResults (calibrator.cpp:415-463), one camera: intr[4], dist[n_dist],
 * T_cn_cnm1[16] (identity for camera 0). */
mcc_status mcc_get_camera(const mcc_handle* h, int k, double* intr, double* dist, double* T_cn_cnm1);

/* Calibrator::testCalibration (calibrator.cpp:316-351): raw residuals (no loss) of every valid point of every
 * active camera, frame -> camera -> point order, (x,y) interleaved; n_out = capacity of out_xy in doubles. */
long long mcc_num_residuals(const mcc_handle* h);
mcc_status mcc_eval_residuals(mcc_handle* h, double* out_xy, size_t n_out);
/* Huber cost 1/2 sum_f rho(s_f) at the current parameters (what Ceres reports as "cost"). */
mcc_status mcc_eval_cost(mcc_handle* h, int use_loss, double* cost);

/* Calibrator::removeOutliers (calibrator.cpp:372-413): masks points with |r|^2 > max_err^2. */
mcc_status mcc_remove_outliers(mcc_handle* h, double max_err, long long* ndropped);
mcc_status mcc_get_mask(const mcc_handle* h, uint8_t* out, size_t n);
mcc_status mcc_set_mask(mcc_handle* h, const uint8_t* in, size_t n);

/* Test hook: the normal-equation blocks ceres would form at the current parameters (Huber-corrected, before
 * Jacobi scaling): cost, Hcc[C*C], gc[C] over the reduced camera columns, and per frame WITH residuals
 * Hff[36], gf[6], Hfc[6*C]; frame_ids receives their frame numbers.  Any output may be NULL. */
int mcc_num_reduced_cam_cols(const mcc_handle* h);
long long mcc_num_active_frames(const mcc_handle* h);
mcc_status mcc_eval_normal_blocks(mcc_handle* h, int use_loss, double* cost, double* Hcc, double* gc, double* Hff,
                                  double* gf, double* Hfc, int* frame_ids);

/* Multi-GPU: one process per GPU, frames sharded by the caller (each rank adds only its frames); the reduced
 * camera system and the scalars are all-reduced over NCCL.  Rank 0 makes an id, the host layer broadcasts it
 * (torch.distributed in bench.py), every rank calls mcc_comm_init.  No reference counterpart: the reference is
 * single-process (calibrator.cpp:303 only sets Ceres' thread count). */
mcc_status mcc_comm_unique_id(unsigned char id[MCC_UNIQUE_ID_BYTES]);
mcc_status mcc_comm_init(mcc_handle* h, int nranks, int rank, const unsigned char id[MCC_UNIQUE_ID_BYTES]);
/* How the per-iteration sums travel between the ranks of this handle: 0 = single rank, 1 = the library's own one-shot
 * all-reduce over NVLink peer windows, 2 = NCCL.  Valid after the first mcc_upload / mcc_run.  An exchange over peer windows
 * waits at most MCC_PEER_TIMEOUT_S seconds (environment, default 60) for another rank; after that the call fails with
 * MCC_ERR_NCCL "peer exchange timed out" (NCCL would wait forever).  A rank whose frame shard is empty still has to make
 * the same mcc_* calls as the others. */
int mcc_comm_transport(const mcc_handle* h);

/* Measurement hooks for bench.py (no reference counterpart: Ceres only prints wall-clock phase times,
 * calibrator.cpp:312).  mcc_timer_start/stop bracket a region with CUDA events on the handle's stream (the
 * stream every kernel of this library is launched on).  With kernel timing enabled every launch is bracketed
 * by an event pair and accumulated per kernel class; mcc_get_kernel_times returns ms and launch counts for
 * mcc_num_kernel_classes() classes named by mcc_kernel_class_name().  mcc_measure_fp64_peak times a
 * dependent-chain DFMA kernel: the measured FP64 roofline denominator of this GPU. */
mcc_status mcc_timer_start(mcc_handle* h);
mcc_status mcc_timer_stop(mcc_handle* h, double* ms);
mcc_status mcc_set_kernel_timing(mcc_handle* h, int enabled);
int mcc_num_kernel_classes(void);
const char* mcc_kernel_class_name(int c);
mcc_status mcc_get_kernel_times(mcc_handle* h, double* ms, long long* counts, int n);
mcc_status mcc_measure_fp64_peak(int device, double* tflops);

/* Test hook (host only, no device needed; no reference counterpart): the lane -> (unit, row slice) map that the tile
 * kernel's contraction uses for a camera with n_dist distortion coefficients when rows are `rowstride` doubles apart
 * (15 or 21).  lane_u / lane_q: 32 entries each, -1 for idle lanes; partner: 3 x 8 partner block columns of the units
 * (-1: none).  Returns the number of row slices, or a negative mcc_status. */
int mcc_debug_lane_map(int n_dist, int rowstride, signed char* lane_u, signed char* lane_q, signed char* partner, int* n_units);

/* No counterpart in the reference.  Sum of squared reprojection residuals of n observations of ONE camera under
 * ONE rigid transform -- the arithmetic of the residual-only pass (FrameResidual with T = double) -- for callers
 * that cannot allocate device memory at the point of the call: a per-frame quality monitor inside a live image
 * callback, where cudaMalloc's device-wide synchronisation and unbounded latency stall the other streams of the
 * process, or a process whose device memory is already claimed up front by another allocator (a perception stack's
 * arena, torch's caching allocator) so that a cudaMalloc can simply fail.  This call performs NO cudaMalloc /
 * cudaFree and calls no allocating library routine; it needs no handle.
 *   Inputs, all DEVICE pointers owned by the caller: d_intr = [fx, fy, cx, cy, dist[0..n_dist)], d_Rt = 12 doubles
 *   (R row-major, then t; X_cam = R X_world + t, i.e. R = R_c R_f and t = R_c t_f + t_c of calibrator.cpp:124-125),
 *   SoA observations d_X, d_Y, d_Z, d_U, d_V [n], d_mask [n] or NULL (all valid).
 *   Output: d_sum_sq[0] = sum over valid points of |projection - (u,v)|^2 (device); cost without loss = half of it.
 *   stream: a cudaStream_t (NULL = default stream); the call only enqueues two kernels and returns.
 * Scratch: one partial sum per 128 observations in a file-scope `static __device__` array of
 * MCC_NOALLOC_MAX_OBS / 128 doubles (256 KiB, held by every GPU that loads the library until the process exits).
 * n > MCC_NOALLOC_MAX_OBS is rejected with MCC_ERR_TOO_LARGE (message in errbuf, which may be NULL).
 * The FIRST call on each device looks up the array's device address (cudaGetSymbolAddress, that device current);
 * make that first call outside any stream capture.  Under CUDA's lazy module loading that first call (like the first
 * launch of any kernel of this library) also makes the driver load the library's device code and the 256 KiB array,
 * which takes device memory and may synchronise the context: make one warm-up call per device at start-up, before the
 * remaining memory is handed to another allocator and before the latency-critical section begins.  The looked-up address
 * is cached per device ordinal for the life of the process and is not valid across cudaDeviceReset.
 * All calls on a device share the one array, which keeps the
 * previous call's partial sums: every slot a call reads it has first overwritten itself, so nothing needs
 * clearing, but two calls whose kernels overlap on the GPU (different streams, or several host threads) race on
 * it -- serialise them on one stream. */
#define MCC_NOALLOC_MAX_OBS 4194304
mcc_status mcc_reprojection_cost_noalloc(int model, int n_dist, const double* d_intr, const double* d_Rt, size_t n,
                                         const double* d_X, const double* d_Y, const double* d_Z, const double* d_U,
                                         const double* d_V, const uint8_t* d_mask, double* d_sum_sq, void* stream,
                                         char* errbuf, size_t errbuf_len);

/* get_init_pose::get_init_pose (/root/reference/src/get_init_pose.cpp:81-140): initial pose T_cam_world of ONE
 * camera view from the homography between the board plane (world x, y; z ignored) and the undistorted image points,
 * decomposed as in decomposeHomography (:46-75).  Host code (an O(N) one-off before the solve).  Needs >= 4 points
 * (MCC_ERR_NO_DATA otherwise); MCC_ERR_NUMERICAL for degenerate configurations. */
mcc_status mcc_init_pose(int model, const double* intr, int n_dist, const double* dist, int npts, const double* world_xyz,
                         const double* image_uv, double* T_cam_world);

/* The same for many views at once: the views are independent (one per (frame, camera) that saw the board -- the nodelet
 * computes one per camera per frame, calibration_nodelet.cpp:510-543), so they are spread over host threads.  A 10 000-frame,
 * 8-camera corners file is 80 000 views, ~0.1 ms each: ~9 s on one core (measured here: 8 000 views 0.91 s on one thread,
 * 0.33 s on four).  T_cam_world: n_views x 16 doubles, row-major; a
 * view that fails gets the zero matrix (what get_init_pose returns, get_init_pose.cpp:88) and its status in status[v] (may
 * be NULL).  n_threads <= 0: one per hardware thread.  Returns MCC_OK unless an argument is invalid: per-view failures are not
 * an error of the call.  Results are bit-identical to n_views calls of mcc_init_pose, whatever the thread count. */
typedef struct mcc_view {
  int model, n_dist, npts;
  const double* intr;        /* fx fy cx cy */
  const double* dist;        /* n_dist coefficients */
  const double* world_xyz;   /* npts x 3 */
  const double* image_uv;    /* npts x 2 */
} mcc_view;
mcc_status mcc_init_pose_batch(int n_views, const mcc_view* views, double* T_cam_world, mcc_status* status, int n_threads);

/* Transform helpers used by the facade: Eigen AngleAxisd(R) route of transform_to_rvec_tvec
 * (calibrator.cpp:157-170) and utils::rotation_matrix (utils.h:29-39). */
void mcc_transform_to_rvec_tvec(const double T[16], double out6[6]);
void mcc_rvec_tvec_to_transform(const double in6[6], double T[16]);

#ifdef __cplusplus
}
#endif
#endif /* MCC_H_ */
