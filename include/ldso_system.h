/* ldso_system.h — C ABI of the C++ orchestrator: FullSystem::addActiveFrame as ONE call per camera frame.
 *
 * Replaces ldso::FullSystem (reference include/frontend/FullSystem.h:70 `addActiveFrame`, src/frontend/FullSystem.cc):
 * the host control flow of the reference — trackNewCoarse with its 83 motion hypotheses (:218-468), the keyframe
 * decision (:152-176), makeKeyFrame (:485-699) with flagFramesForMarginalization (:763-842), activatePointsMT
 * (:1197-1337), optimize (:844-1007), removeOutliers (:1644-1665), flagPointsForRemoval (:1356-1424),
 * EnergyFunctional::marginalizePointsF / marginalizeFrame (EnergyFunctional.cc:109-262), makeNewTraces (:1426-1482) and
 * CoarseInitializer bootstrap (:95-128, initializeFromInitializer :1484-1642) — stays C++ (include/ldso_system.hpp) and
 * drives the CUDA kernels through the C ABI of include/ldso_gpu.h.  Vision-only (setting_vi_enable = false),
 * linearizeOperation = true (tracking and mapping on the caller's thread: examples/run_dso_tum_mono.cc:73,348),
 * loop closing off.
 *
 * Conventions as in ldso_gpu.h: int return, 0 = OK, < 0 = failure (message: ldso_gpu_system_last_error), plain
 * pointers and sizes, poses as double[12] = row-major R(9) then t(3).
 */
#ifndef LDSO_SYSTEM_H_
#define LDSO_SYSTEM_H_

#include "ldso_gpu.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct ldso_gpu_system ldso_gpu_system;

/* The FullSystem-level subset of the setting_* globals (src/Setting.cc; defaults = preset 0 of
 * examples/run_dso_tum_mono.cc:65-83 with mode=1: no photometric calibration, affine a,b optimised without prior). */
typedef struct ldso_system_config {
    ldso_gpu_config gpu;                 /* hot-path settings and capacities of the underlying context */
    float desired_immature_density;      /* setting_desiredImmatureDensity = 1500 (:29) */
    float desired_point_density;         /* setting_desiredPointDensity = 2000 (:30) */
    float min_points_remaining;          /* setting_minPointsRemaining = 0.05 (:31) */
    float max_log_aff_fac_in_window;     /* setting_maxLogAffFacInWindow = 0.7 (:32) */
    int min_frames;                      /* setting_minFrames = 5 (:33) */
    int max_frames;                      /* setting_maxFrames = 7 (:34) */
    int min_frame_age;                   /* setting_minFrameAge = 1 (:35) */
    int max_opt_iterations;              /* setting_maxOptIterations = 6 (:36) */
    int min_opt_iterations;              /* setting_minOptIterations = 1 (:37) */
    float th_opt_iterations;             /* setting_thOptIterations = 1.2 (:38) */
    float max_shift_weight_t;            /* setting_maxShiftWeightT = 0.04 (:11) */
    float max_shift_weight_r;            /* setting_maxShiftWeightR = 0 (:12) */
    float max_shift_weight_rt;           /* setting_maxShiftWeightRT = 0.02 (:13) */
    float kf_global_weight;              /* setting_kfGlobalWeight = 1 (:14) */
    float max_affine_weight;             /* setting_maxAffineWeight = 2 (:15) */
    float idepth_fix_prior_marg_fac;     /* setting_idepthFixPriorMargFac = 600^2 (:17) */
    float initial_rot_prior;             /* setting_initialRotPrior = 1e11 (:18) */
    float initial_trans_prior;           /* setting_initialTransPrior = 1e10 (:19) */
    float initial_aff_b_prior;           /* setting_initialAffBPrior = 1e14 (:20) */
    float initial_aff_a_prior;           /* setting_initialAffAPrior = 1e14 (:21) */
    float initial_calib_hessian;         /* setting_initialCalibHessian = 5e9 (:22) */
    float min_idepth_h_marg;             /* setting_minIdepthH_marg = 50 (:26) */
    float marg_weight_fac;               /* setting_margWeightFac = 0.25 (:45) */
    float min_trace_quality;             /* setting_minTraceQuality = 3 (:51) */
    float re_track_threshold;            /* setting_reTrackThreshold = 1.5 (:53) */
    int min_good_active_res_for_marg;    /* setting_minGoodActiveResForMarg = 3 (:54) */
    int min_good_res_for_marg;           /* setting_minGoodResForMarg = 4 (:55) */
    float initializer_slack_factor;      /* benchmark_initializerSlackFactor = 1 (:100) */
    int point_selection;                 /* setting_pointSelection: 0 = DSO PixelSelector, 1 = LDSO corner detector (N3) */
    /* EnergyFunctional::nPoints steers the activation density (FullSystem.cc:1199-1215).  The fork only ever DECREMENTS
     * it (EnergyFunctional.cc:200-201; the insertPoint of upstream DSO is gone), so it is <= 0, currentMinActDist falls to 0
     * and every activatable candidate is activated.  0 = as written in the fork, 1 = count the active points (upstream
     * DSO's behaviour: the window settles at desired_point_density points, BASELINE.json configs[1]). */
    int npoints_counter_mode;
    int reserved[8];
} ldso_system_config;

void ldso_gpu_system_default_config(ldso_system_config* cfg);
int ldso_gpu_system_create(const ldso_system_config* cfg, ldso_gpu_system** out);
void ldso_gpu_system_destroy(ldso_gpu_system* sys);
const char* ldso_gpu_system_last_error(const ldso_gpu_system* sys);
/* the underlying context (timers, profile, launch / byte counters of ldso_gpu.h) */
ldso_gpu_ctx* ldso_gpu_system_ctx(ldso_gpu_system* sys);

/* status_out slots of ldso_gpu_system_add_active_frame */
#define LDSO_SYS_ST_INITIALIZED 0   /* FullSystem::initialized after the call */
#define LDSO_SYS_ST_KEYFRAME 1      /* the frame became a keyframe */
#define LDSO_SYS_ST_LOST 2          /* FullSystem::isLost */
#define LDSO_SYS_ST_INIT_FAILED 3   /* FullSystem::initFailed */
#define LDSO_SYS_ST_WINDOW 4        /* keyframes in the window after the call */
#define LDSO_SYS_ST_POINTS 5        /* active points */
#define LDSO_SYS_ST_RESIDUALS 6     /* residuals of the active points */
#define LDSO_SYS_ST_IMMATURE 7      /* immature points */
#define LDSO_SYS_ST_TRIES 8         /* motion hypotheses tried by trackNewCoarse */
#define LDSO_SYS_ST_FRAME_ID 9      /* Frame::id given to the frame */
#define LDSO_SYS_ST_N 16

/* what the image pointer of add_active_frame / batch_* is (the `image_is_device` argument; 0 and 1 keep their old meaning) */
#define LDSO_IMAGE_HOST_F32 0     /* w*h floats in host memory: ImageAndExposure::image, what Undistort hands over */
#define LDSO_IMAGE_DEVICE_F32 1   /* the same, already resident in HBM (a DEVICE pointer) */
#define LDSO_IMAGE_HOST_RAW8 2    /* the 8-bit sensor image (w_org*h_org bytes, host): Undistort::undistort<unsigned char> */
#define LDSO_IMAGE_HOST_RAW16 3   /* (src/frontend/Undistort.cc:366-458) runs on the device in front of makeImages; the tables
                                     are those of ldso_gpu_undistort_set on ldso_gpu_system_ctx(sys); a quarter (half) of the bytes
                                     of the float image cross the bus */

/* FullSystem::addActiveFrame(ImageAndExposure*, imu) (src/frontend/FullSystem.cc:78-187).  image: w*h floats (irradiance
 * 0..255, what Undistort hands over); image_is_device: one of LDSO_IMAGE_* (1: `image` is a DEVICE pointer, inputs already
 * resident in HBM; 2 / 3: `image` points to the raw 8 / 16-bit sensor image).
 * pose_out (may be NULL): worldToCam of the frame after the call (Frame::getPose), identity while initialising.
 * status_out (may be NULL): LDSO_SYS_ST_N ints.  Returns 0; 1 when the system is lost (the reference returns at once). */
int ldso_gpu_system_add_active_frame(ldso_gpu_system* sys, const float* image, int image_is_device, double timestamp,
                                     float exposure_time, double pose_out[12], int* status_out);

/* Frame::getPose of every frame with a valid pose, in frame order: ids[n], poses[n][12] (worldToCam, keyframes carry their
 * optimised pose), is_kf[n].  The reference's printResult (src/frontend/FullSystem.cc:2259-2288) writes the keyframes only;
 * a non-keyframe keeps the pose it was tracked at (relative_to_ref = 0) or, with relative_to_ref != 0, follows its
 * tracking reference: camToTrackingRef composed with the reference keyframe's current pose (upstream DSO's
 * FrameShell::trackingRef output).  Returns the number of frames written (at most max_n). */
int ldso_gpu_system_get_trajectory(ldso_gpu_system* sys, int max_n, int* ids, double* poses, int* is_kf, int relative_to_ref);
/* ids (Frame::id) of the keyframes in the window */
int ldso_gpu_system_get_window(ldso_gpu_system* sys, int max_n, int* ids);

/* Counters since creation (host bookkeeping, for tests and the bench log) */
#define LDSO_SYS_CT_FRAMES 0            /* frames fed */
#define LDSO_SYS_CT_KEYFRAMES 1         /* keyframes made (FullSystem::makeKeyFrame calls) */
#define LDSO_SYS_CT_MARG_FRAMES 2       /* frames marginalised (EnergyFunctional::marginalizeFrame calls) */
#define LDSO_SYS_CT_MARG_POINTS 3       /* points marginalised into HM, bM (marginalizePointsF) */
#define LDSO_SYS_CT_LIN_RESIDUALS 4     /* residuals whose linearisation was fixed (fixLinearizationF) */
#define LDSO_SYS_CT_DROPPED_POINTS 5    /* points dropped as OUT / OUTLIER */
#define LDSO_SYS_CT_ACTIVATED 6         /* immature points activated */
#define LDSO_SYS_CT_BA_ITERATIONS 7     /* solveSystemF calls */
#define LDSO_SYS_CT_TRACK_CALLS 8       /* trackNewestCoarse calls */
#define LDSO_SYS_CT_RES_IN_M 9          /* EnergyFunctional::resInM */
#define LDSO_SYS_CT_N 16
int ldso_gpu_system_get_counters(ldso_gpu_system* sys, long long* out /* LDSO_SYS_CT_N */);
/* the marginalisation prior as it stands: HM ((8n+4)^2 row-major), bM (8n+4), n = window size; returns n (pointers may
 * be NULL to query n) */
int ldso_gpu_system_get_prior(ldso_gpu_system* sys, double* HM, double* bM);
/* last keyframe: [rmse returned by optimize, points, residuals before optimize, points marginalised, points dropped,
 * residuals linearised, activated, frames marginalised, currentMinActDist, frameEnergyTH of the new keyframe,
 * BA iterations, |HM|_F] */
int ldso_gpu_system_last_keyframe_log(ldso_gpu_system* sys, double out[12]);
/* the first n values of glibc's rand() after srand(seed), from the library's private restatement of the generator
 * (FullSystem::initializeFromInitializer draws from std::rand; sequences stepped concurrently must not share it) */
int ldso_gpu_system_rand_sequence(unsigned seed, int n, int* out);
/* host wall-clock seconds spent per phase since creation (0 trackNewCoarse incl. makeImages, 1 traceNewCoarse,
 * 2 activatePointsMT, 3 optimize, 4 marginalisation, 5 makeNewTraces + setCoarseTrackingRef, 6 initializer, 7 other;
 * inside 0: 8 the makeImages call, 9 the trackNewestCoarse calls, 10 building the motion hypotheses; 11 unused) */
int ldso_gpu_system_get_phase_seconds(ldso_gpu_system* sys, double out[12]);

/* ---- several sequences per GPU (BASELINE.json configs[4]: 64 independent sequences over 1/2/4/8 GPUs) ----
 * A batch steps n systems together: worker thread t (of n_threads) owns systems t, t + n_threads, ... and feeds each its
 * next frame; the call returns when every system has consumed its frame.  Sequences share nothing (own context, streams,
 * device pools), so their kernels overlap on the GPU.  The systems stay owned by the caller and must outlive the batch;
 * while a batch exists its systems must only be stepped through it. */
typedef struct ldso_gpu_batch ldso_gpu_batch;
int ldso_gpu_batch_create(int n, ldso_gpu_system** systems, int n_threads, ldso_gpu_batch** out);
void ldso_gpu_batch_destroy(ldso_gpu_batch* batch);
/* images[n]: one w*h float image per system (all host or all device pointers); timestamps / exposures (may be NULL: 0 / 1);
 * sync_after != 0: every worker drains its contexts' streams (ldso_gpu_sync: the queued traceNewCoarse too) before the call
 * returns; poses_out [n][12] and status_out [n][LDSO_SYS_ST_N] may be NULL.  Returns 0, or a negative return of a system. */
int ldso_gpu_batch_add_active_frames(ldso_gpu_batch* batch, const float* const* images, int image_is_device,
                                     const double* timestamps, const float* exposures, int sync_after, double* poses_out,
                                     int* status_out);

/* The same for n_frames frames per system in ONE call, without a rendezvous between frames: images[n][n_frames]
 * (system-major; they must stay valid until the call returns), exposure 1, timestamps counting up.  Every worker runs its
 * systems as fibers: where the library would wait for the device it switches to the worker's next system
 * (ldso_gpu_set_wait_hook), and frame k+1 of a system is prepared (copy, Undistort, pyramid) behind the tracker launch of its
 * frame k; a keyframe of one sequence holds back neither the other workers nor the worker's other sequences.
 * poses_out [n][n_frames][12], status_out [n][n_frames][LDSO_SYS_ST_N]. */
int ldso_gpu_batch_run(ldso_gpu_batch* batch, int n_frames, const float* const* images, int image_is_device, int sync_after,
                       double* poses_out, int* status_out);

#ifdef __cplusplus
}
#endif
#endif /* LDSO_SYSTEM_H_ */
