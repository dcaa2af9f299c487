/* ldso_gpu.h — C ABI of libldso_gpu.so: the B200 (sm_100a) replacement for LDSO's photometric
 * front-end + windowed-BA hot path.
 *
 * The reference (n-lalanne/LDSO) has no FFI layer: its boundary is the C++ class surface in
 * namespace ldso / ldso::internal that FullSystem calls.  Every entry point below names the reference
 * method it replaces (file:line relative to the reference tree).  The C++ shim classes in
 * include/ldso_shim.h keep the reference's class and method names and forward here.
 *
 * Conventions
 *   - every function returns int: 0 = OK, <0 = CUDA/driver/usage failure (sticky, see
 *     ldso_gpu_last_error), >0 = algorithmic condition (documented per function);
 *   - plain pointers + sizes only; all pointers are HOST pointers unless the name ends in _dev;
 *   - poses cross the boundary as double[12] = row-major R(9) followed by t(3);
 *   - nothing throws or aborts across the boundary; one ctx per sequence, no global device state.
 */
#ifndef LDSO_GPU_H_
#define LDSO_GPU_H_

#ifdef __cplusplus
extern "C" {
#endif

#define LDSO_MAX_LEVELS 6      /* PYR_LEVELS, include/Settings.h:10 */
#define LDSO_PATTERN_NUM 8     /* patternNum, include/Settings.h:225 */
#define LDSO_NUM_TRACKERS 2    /* coarseTracker + coarseTracker_forNewKF, include/frontend/FullSystem.h */

typedef struct ldso_gpu_ctx ldso_gpu_ctx;

/* Per-sequence configuration: replaces the process-global calib (src/internal/GlobalCalib.cc:7-18)
 * and the hot-path subset of the ~170 setting_* globals (src/Setting.cc). */
typedef struct ldso_gpu_config {
    int w, h;                       /* level-0 image size (after the reference's crop/rectify) */
    float fx, fy, cx, cy;           /* level-0 pinhole intrinsics handed to setGlobalCalib */
    int device;                     /* CUDA device ordinal */
    int n_frame_slots;              /* pyramids resident at once: window (setting_maxFrames) + in flight */
    int max_points;                 /* capacity of the active-point SoA */
    int max_immature;               /* capacity of the immature-point SoA */
    int max_residuals;              /* capacity of the residual SoA */
    /* src/Setting.cc defaults */
    float huber_th;                 /* setting_huberTH = 9 (:76) */
    float coarse_cutoff_th;         /* setting_coarseCutoffTH = 20 (:82) */
    int gamma_weights_pixel_select; /* setting_gammaWeightsPixelSelect = 1 (:68) */
    float affine_opt_mode_a;        /* setting_affineOptModeA (1e12; mode=1 -> 0) */
    float affine_opt_mode_b;        /* setting_affineOptModeB (1e8;  mode=1 -> 0) */
    float outlier_th;               /* setting_outlierTH = 144 */
    float outlier_th_sum_component; /* setting_outlierTHSumComponent = 2500 */
    float overall_energy_th_weight; /* setting_overallEnergyTHWeight = 1 */
    float max_pix_search;           /* setting_maxPixSearch = 0.027 */
    float trace_stepsize;           /* 1.0 */
    int trace_gn_iterations;        /* 3 */
    float trace_gn_threshold;       /* 0.1 */
    float trace_extra_slack_on_th;  /* 1.2 */
    float trace_slack_interval;     /* 1.5 */
    float trace_min_improvement;    /* 2 */
    int min_trace_test_radius;      /* 2 */
    float min_grad_hist_cut;        /* 0.5 */
    float min_grad_hist_add;        /* 7 */
    float grad_downweight_per_level;/* 0.75 */
    int select_direction_distribution; /* true */
    float idepth_fix_prior;         /* setting_idepthFixPrior = 2500 */
    float min_idepth_h_act;         /* 100 */
    int gn_its_on_point_activation; /* 3 */
    float frame_energy_th_n;        /* 0.7 */
    float frame_energy_th_fac_median; /* 1.5 */
    float frame_energy_th_const_weight; /* 0.5 */
    int tracker_solve_mode;         /* 0 = FULL_SYM (upstream DSO), 1 = FORK_LOWER_ONLY */
    int reserved[8];
} ldso_gpu_config;

/* Fill cfg with src/Setting.cc defaults + preset=0 (examples/run_dso_tum_mono.cc:65-83). */
void ldso_gpu_default_config(ldso_gpu_config* cfg);

/* Context life cycle.  create() derives the pyramid level count with the rule of setGlobalCalib
 * (src/internal/GlobalCalib.cc:20-28), allocates every device buffer once (no allocation on the
 * per-frame path), and creates the ctx's streams. */
int ldso_gpu_create(const ldso_gpu_config* cfg, ldso_gpu_ctx** out);
void ldso_gpu_destroy(ldso_gpu_ctx* ctx);
const char* ldso_gpu_last_error(const ldso_gpu_ctx* ctx);
int ldso_gpu_pyr_levels(const ldso_gpu_ctx* ctx);                     /* pyrLevelsUsed */
int ldso_gpu_level_size(const ldso_gpu_ctx* ctx, int lvl, int* w, int* h); /* wG[lvl], hG[lvl] */
int ldso_gpu_sync(ldso_gpu_ctx* ctx);                                 /* drain the ctx's streams */
/* number of kernel launches issued through this ctx since creation (bench.py "gpu_launches") */
long long ldso_gpu_launch_count(const ldso_gpu_ctx* ctx);
/* CUDA-event stopwatch on the ctx's tracking stream (the stream every kernel of this ctx is launched on) */
int ldso_gpu_timer_start(ldso_gpu_ctx* ctx);
int ldso_gpu_timer_stop(ldso_gpu_ctx* ctx, float* ms);
/* Per-kernel CUDA-event profile.  enable(1) zeroes the counters; which = 0: fused tracker kernel (K3+K4).
 * total_ms = summed launch durations, units = items processed (points). */
int ldso_gpu_profile_enable(ldso_gpu_ctx* ctx, int on);
int ldso_gpu_profile_get(ldso_gpu_ctx* ctx, int which, double* total_ms, long long* launches, double* units);

/* CalibHessian::B inverse-response LUT used by getBGradOnly (include/internal/CalibHessian.h:102-111).
 * Default: identity (B[i] = i), i.e. no photometric calibration. */
int ldso_gpu_set_gamma(ldso_gpu_ctx* ctx, const float B[256]);

/* ---- K1: FrameHessian::makeImages (src/internal/FrameHessian.cc:59-115) ------------------------
 * Builds dIp[lvl] = {I, dx, dy} and absSquaredGrad[lvl] for every pyramid level of frame `slot`
 * from a w*h float irradiance image.  Device layout: one float4 {I,dx,dy,absSquaredGrad} per pixel. */
int ldso_gpu_frame_make_images(ldso_gpu_ctx* ctx, int slot, const float* img);
int ldso_gpu_frame_make_images_dev(ldso_gpu_ctx* ctx, int slot, const float* img_dev);
/* Host mirror in the reference's layout.  what = 0: Eigen::Vector3f dIp[lvl] (3 floats / px, AoS);
 * what = 1: float absSquaredGrad[lvl]. */
int ldso_gpu_frame_download(ldso_gpu_ctx* ctx, int slot, int lvl, int what, float* dst);

/* ---- K2: CoarseTracker::makeK (src/frontend/CoarseTracker.cc:312-339) and
 *          CoarseTracker::setCoarseTrackingRef / makeCoarseDepthL0 (:347-538) ----------------------
 * make_k takes the current CalibHessian fxl,fyl,cxl,cyl.  set_ref takes, for every active point whose
 * newest residual is IN and targets the new reference frame (:363-382): centerProjectedTo (u,v,new
 * idepth) and PointHessian::HdiF; plus the reference frame's pyramid slot, exposure and aff_g2l. */
int ldso_gpu_tracker_make_k(ldso_gpu_ctx* ctx, int tracker, float fx, float fy, float cx, float cy);
int ldso_gpu_tracker_set_ref(ldso_gpu_ctx* ctx, int tracker, int ref_slot, int n_points,
                             const float* u, const float* v, const float* idepth, const float* hdi,
                             float ref_exposure, float ref_aff_a, float ref_aff_b);
/* pc_n[lvl] and (optionally, any pointer may be NULL) pc_u/pc_v/pc_idepth/pc_color[lvl] */
int ldso_gpu_tracker_get_pc(ldso_gpu_ctx* ctx, int tracker, int lvl, int* n,
                            float* u, float* v, float* idepth, float* color);

/* ---- K3: CoarseTracker::calcRes (src/frontend/CoarseTracker.cc:540-673) ------------------------
 * out6 = [E, numTermsInE, flowT, 0, flowRT, saturatedRatio].  The warped-point buffers of the
 * reference are never materialised: the same launch also accumulates the K4 system for this pose. */
int ldso_gpu_tracker_calc_res(ldso_gpu_ctx* ctx, int tracker, int new_slot, float new_exposure, int lvl,
                              const double ref_to_new[12], float aff_a, float aff_b, float cutoff_th,
                              float out6[6]);
/* ---- K4: CoarseTracker::calcGSSSE (:675-741) over the points warped by the LAST calc_res ---------
 * H (8x8 row-major) and b already divided by the padded count and scaled (SCALE_XI_TRANS etc). */
int ldso_gpu_tracker_calc_gs(ldso_gpu_ctx* ctx, int tracker, int lvl, const double ref_to_new[12],
                             float aff_a, float aff_b, double H[64], double b[8]);
/* ---- CoarseTracker::trackNewestCoarse (:61-310): the whole coarse-to-fine LM loop ---------------
 * pose_io / aff_io: initial guess in, result out (only written when the return value is 0).
 * Returns 0 = true, 1 = false (the reference's bool).  last_residuals[5], flow[3] as the public members
 * lastResiduals / lastFlowIndicators.  iters_out (may be NULL): [calcRes evaluations, accepted steps]. */
int ldso_gpu_tracker_track(ldso_gpu_ctx* ctx, int tracker, int new_slot, float new_exposure,
                           double pose_io[12], float aff_io[2], int coarsest_lvl,
                           const float min_res_for_abort[5], float last_residuals[5], float flow[3],
                           int iters_out[2]);

#ifdef __cplusplus
}
#endif
#endif /* LDSO_GPU_H_ */
