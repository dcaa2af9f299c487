/* ldso_gpu.h — C ABI of libldso_gpu.so: the B200 (sm_100a) replacement for LDSO's photometric
 * front-end + windowed-BA hot path.
 *
 * The reference (n-lalanne/LDSO) has no FFI layer: its boundary is the C++ class surface in
 * namespace ldso / ldso::internal that FullSystem calls.  Every entry point below names the reference
 * method it replaces (file:line relative to the reference tree).  The C++ shim classes in
 * include/ldso_shim.hpp keep the reference's class and method names and forward here.
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
    int tracker_lm_mode;            /* 0 = the whole coarse-to-fine LM loop resident on the device on ONE thread-block cluster
                                     * (DSMEM fold + cluster barrier, ordinary launch), 1 = host-driven loop, one launch +
                                     * read-back per evaluation, 2 = device-resident on a cooperative grid (grid barrier) */
    int reserved[7];
} ldso_gpu_config;

/* Fill cfg with src/Setting.cc defaults + preset=0 (examples/run_dso_tum_mono.cc:65-83). */
void ldso_gpu_default_config(ldso_gpu_config* cfg);

/* Cooperative waiting (no counterpart in the reference; used by ldso_gpu_batch_*).  While a hook is installed for the CALLING
 * THREAD, every wait inside the library (stream / event synchronisation, the tracker's mailbox poll) polls and calls
 * hook(arg) between polls instead of spinning, so a scheduler can run another sequence on the same thread until the device
 * has caught up.  hook = NULL restores the plain blocking waits.  Thread-local: it affects calls made from this thread only. */
int ldso_gpu_set_wait_hook(void (*hook)(void*), void* arg);

/* Context life cycle.  create() derives the pyramid level count with the rule of setGlobalCalib
 * (src/internal/GlobalCalib.cc:20-28), allocates the frame pyramids, the two trackers, the window (points, residuals,
 * accumulators) and the tracker's reduction workspace, and creates the ctx's streams.  The other subsystems allocate their
 * buffers once, at their first use: selector, immature pool, corner detector (first keyframe), undistorter
 * (ldso_gpu_undistort_set), initializer (ldso_gpu_init_set_first: its per-level point arrays are sized by the first
 * frame's point count, so each set_first allocates them anew).  A tracked frame (makeImages, trackNewestCoarse,
 * traceNewCoarse) and a keyframe of an initialised sequence allocate nothing. */
int ldso_gpu_create(const ldso_gpu_config* cfg, ldso_gpu_ctx** out);
void ldso_gpu_destroy(ldso_gpu_ctx* ctx);
const char* ldso_gpu_last_error(const ldso_gpu_ctx* ctx);
int ldso_gpu_pyr_levels(const ldso_gpu_ctx* ctx);                     /* pyrLevelsUsed */
int ldso_gpu_level_size(const ldso_gpu_ctx* ctx, int lvl, int* w, int* h); /* wG[lvl], hG[lvl] */
int ldso_gpu_sync(ldso_gpu_ctx* ctx);                                 /* drain the ctx's streams */
/* makes the ctx's CUDA device the calling thread's current device (cudaSetDevice(cfg.device)): call once from every host
 * thread that is going to use the ctx (the reference's tracking and mapping threads, a batch worker) */
int ldso_gpu_bind_thread(ldso_gpu_ctx* ctx);
/* number of kernel launches issued through this ctx since creation (bench.py "gpu_launches") */
long long ldso_gpu_launch_count(const ldso_gpu_ctx* ctx);
/* bytes copied host->device / device->host by the entry points of this ctx since creation (bench.py "e2e") */
int ldso_gpu_io_bytes(ldso_gpu_ctx* ctx, long long* h2d, long long* d2h);
/* CUDA-event stopwatch on the ctx's tracking stream (the stream every kernel of this ctx is launched on) */
int ldso_gpu_timer_start(ldso_gpu_ctx* ctx);
int ldso_gpu_timer_stop(ldso_gpu_ctx* ctx, float* ms);
/* Per-kernel CUDA-event profile (events on the launching stream, resolved in profile_get).  enable(1) zeroes the
 * counters.  total_ms = summed launch durations, units = items processed (points / residuals / pixels). */
#define LDSO_PROF_K34 0      /* fused tracker kernel, units = points */
#define LDSO_PROF_K5 1       /* traceOn, units = epipolar search steps (+9 per searched point for the fixed part) */
#define LDSO_PROF_K7 2       /* linearize, units = residuals */
#define LDSO_PROF_K8 3       /* top Hessian accumulation (both modes), units = residuals */
#define LDSO_PROF_K9 4       /* Schur accumulation, units = residuals */
#define LDSO_PROF_K1 5       /* pyramid (both kernels), units = frames */
#define LDSO_PROF_A2 6       /* selector kernels of one makeMaps call, units = level-0 pixels per pass */
#define LDSO_PROF_K6 7       /* activation, units = candidates */
#define LDSO_PROF_K10 8      /* resubstitute, units = points */
#define LDSO_PROF_K2 9       /* makeCoarseDepthL0 kernels, units = points */
#define LDSO_PROF_N2 10      /* distance map, units = level-1 cells */
#define LDSO_PROF_N2SEL 11   /* activation walk, units = pool points */
#define LDSO_PROF_SLOTS 12
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
/* ---- allocation-free variant of ldso_gpu_tracker_calc_res -------------------------------------------
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot call cudaMalloc
 * (code under CUDA-graph capture, tight serving loops, harnesses that forbid it).  Same fused K3+K4 kernel and the same
 * result as ldso_gpu_tracker_calc_res, but the reduction scratch (per-CTA partial sums, arrival counter, result block)
 * is a file-scope `static __device__ __align__(256)` array of 246 KB inside the library (every GPU that loads the
 * library holds it until the process exits) instead of the ctx's workspace, and the call allocates nothing: no
 * cudaMalloc / cudaFree / cudaMallocHost, no allocating library routine.  Inputs are the ctx's resident tracker
 * reference (ldso_gpu_tracker_set_ref) and frame pyramid `new_slot`, as for ldso_gpu_tracker_calc_res; the tracker's
 * calc_gs state is NOT touched.  Limit: at most LDSO_NOALLOC_MAX_POINTS points on the level (pc_n[lvl]); more returns -2
 * with a message in ldso_gpu_last_error.  The ctx's device must be the calling thread's current device.  All calls on a
 * device share the one array, which keeps the previous call's contents (the call clears what must start at zero):
 * calls that overlap on the GPU (two ctxs / two streams at once) race — serialise them.
 * out6 as ldso_gpu_tracker_calc_res; sums52 (may be NULL): the 45 upper-triangle sums of the 9x9 system, E, flowT,
 * flowRT, then numTermsInE, numSaturated, numTermsInWarped, flowNum. */
#define LDSO_NOALLOC_MAX_POINTS 131072
int ldso_gpu_tracker_calc_res_noalloc(ldso_gpu_ctx* ctx, int tracker, int new_slot, float new_exposure, int lvl,
                                      const double ref_to_new[12], float aff_a, float aff_b, float cutoff_th,
                                      float out6[6], double sums52[52]);
/* ---- CoarseTracker::trackNewestCoarse (:61-310): the whole coarse-to-fine LM loop ---------------
 * pose_io / aff_io: initial guess in, result out (only written when the return value is 0).
 * Returns 0 = true, 1 = false (the reference's bool).  last_residuals[5], flow[3] as the public members
 * lastResiduals / lastFlowIndicators.  iters_out (may be NULL): [calcRes evaluations, accepted steps]. */
int ldso_gpu_tracker_track(ldso_gpu_ctx* ctx, int tracker, int new_slot, float new_exposure,
                           double pose_io[12], float aff_io[2], int coarsest_lvl,
                           const float min_res_for_abort[5], float last_residuals[5], float flow[3],
                           int iters_out[2]);
/* The same call in two halves (CoarseTracker::trackNewestCoarse :61-310, no second counterpart in the reference): _begin
 * launches the tracker kernel and returns without waiting (with tracker_lm_mode = 1, the host-driven loop, it only keeps
 * its arguments); _end waits for the result and returns what ldso_gpu_tracker_track returns, pose / affine of the result in
 * pose_out / aff_out (the initial guess when the return value is not 0).  One pair at a time per ctx; no other call on the ctx
 * between the two.  A host thread that steps several sequences begins all of them before it ends the first
 * (ldso_gpu_batch_*), so their trackers run side by side. */
int ldso_gpu_tracker_track_begin(ldso_gpu_ctx* ctx, int tracker, int new_slot, float new_exposure, const double pose[12],
                                 const float aff[2], int coarsest_lvl, const float min_res_for_abort[5]);
int ldso_gpu_tracker_track_end(ldso_gpu_ctx* ctx, double pose_out[12], float aff_out[2], float last_residuals[5], float flow[3],
                               int iters_out[2]);


/* ---- A2: PixelSelector::makeMaps (src/frontend/PixelSelector2.cc:111-169) incl. makeHists (:36-109) and
 * select (:171-316).  map_out: w*h floats in {0,1,2,4} (PixelSelectorStatus as the reference stores it);
 * potential_io: PixelSelector::currentPotential, persistent across keyframes (initially 3); returns the
 * number of selected pixels in *n_selected.  The threshold tables are zero-filled at allocation (the
 * reference leaves them uninitialised off the 32-pixel grid, SURVEY.md trap T3). */
int ldso_gpu_select_make_maps(ldso_gpu_ctx* ctx, int slot, float density, int recursions_left, float th_factor,
                              int* potential_io, float* map_out, int* n_selected);

/* PixelSelector::makeMaps followed by the pixel loop of FullSystem::makeNewTraces (src/frontend/FullSystem.cc:1440-1460):
 * the selected pixels with border <= x < w - border, border <= y < h - border in row-major order as (u, v, type) lists,
 * type = the map value 1 / 2 / 4.  At most max_points are returned. */
int ldso_gpu_select_points(ldso_gpu_ctx* ctx, int slot, float density, int recursions_left, float th_factor,
                           int* potential_io, int border, int max_points, float* u_out, float* v_out, float* type_out,
                           int* n_points);

/* ---- K5 / K6: immature points (src/internal/ImmaturePoint.cc) ---------------------------------------
 * The pool of immature points lives on the device; the host re-sends it after it adds or removes points
 * (once per keyframe) and reads back what it branches on.  ImmaturePointStatus codes as in
 * include/internal/ImmaturePoint.h:31-38 (0 GOOD, 1 OOB, 2 OUTLIER, 3 SKIPPED, 4 BADCONDITION, 5 UNINITIALIZED). */
/* ImmaturePoint::ImmaturePoint (:14-39): color[8], weights[8], gradH (xx,xy,yy) and energyTH of n new points
 * hosted in frame `slot` (energyTH = NaN marks a point whose patch is not finite) */
int ldso_gpu_immature_create(ldso_gpu_ctx* ctx, int slot, int n, const float* u, const float* v, float* color,
                             float* weights, float* grad_h, float* energy_th);
/* upload the pool: host = index into the per-host tables passed to trace / the window frames for activate */
int ldso_gpu_immature_set(ldso_gpu_ctx* ctx, int n, const int* host, const float* u, const float* v, const float* color,
                          const float* weights, const float* grad_h, const float* energy_th, const float* idepth_min,
                          const float* idepth_max, const float* quality, const int* status);
/* FullSystem::traceNewCoarse (src/frontend/FullSystem.cc:1157-1195) = ImmaturePoint::traceOn (:47-317) for every
 * pool point against frame new_slot; KRKi [n_hosts][9], Kt [n_hosts][3], aff [n_hosts][2] are the per-host
 * hostToFrame tables of :1171-1176.  counts_out[6]: number of points per resulting status. */
int ldso_gpu_immature_trace(ldso_gpu_ctx* ctx, int new_slot, int n_hosts, const float* KRKi, const float* Kt,
                            const float* aff, int counts_out[6]);
/* the same trace without waiting for it: the kernel is queued behind the frame's other work and the call returns; the
 * status counts are fetched later with ldso_gpu_immature_trace_counts (which waits if it has to).  Every other pool entry
 * point completes a pending trace first, so results are the same as with ldso_gpu_immature_trace.  The reference consumes
 * trace results only at the next keyframe (activatePointsMT, FullSystem.cc:1203-1320), so the wait leaves the frame's
 * critical path. */
int ldso_gpu_immature_trace_async(ldso_gpu_ctx* ctx, int new_slot, int n_hosts, const float* KRKi, const float* Kt,
                                  const float* aff);
int ldso_gpu_immature_trace_counts(ldso_gpu_ctx* ctx, int counts_out[6]);
/* any pointer may be NULL */
int ldso_gpu_immature_get(ldso_gpu_ctx* ctx, float* idepth_min, float* idepth_max, float* quality, int* status,
                          float* last_trace_uv, float* last_trace_pixel_interval);
/* FullSystem::optimizeImmaturePoint (:1035-1155) for the listed pool points against the current window frames
 * (ldso_gpu_window_set_frames): ok_out[i] = 1 when a PointHessian would be created, idepth_out its inverse depth,
 * res_state_out [n_cand][n_frames] the final ResState per target (-1 for the host frame itself). */
int ldso_gpu_immature_activate(ldso_gpu_ctx* ctx, int n_cand, const int* cand, int min_obs, float* idepth_out,
                               int* ok_out, int* res_state_out);

/* Device-resident pool maintenance: the pool never has to come back to the host to change shape.
 * compact: drop every point hosted by frame index drop_host (-1: none) — the frame that leaves the window
 *          (FullSystem::marginalizeFrame) —, every point whose status is OOB when drop_oob != 0 (FullSystem.cc:1279-1284),
 *          and the listed pool indices (points that were activated or discarded, :1301-1318).  Order is preserved and
 *          host indices above drop_host move down by one.  n_out: new pool size.
 * append:  ImmaturePoint constructor (:14-39) for n pixels of frame `slot` + append to the pool with host index host_idx,
 *          idepth_min 0, idepth_max NaN, quality 10000, status UNINITIALIZED (OOB where the patch is not finite, the
 *          reference deletes such points: FullSystem.cc:1456-1458).  n_pool: new pool size. */
int ldso_gpu_immature_compact(ldso_gpu_ctx* ctx, int drop_host, int drop_oob, int n_remove, const int* remove_idx, int* n_out);
int ldso_gpu_immature_append(ldso_gpu_ctx* ctx, int slot, int host_idx, int n, const float* u, const float* v, int* n_pool);

/* ---- N2: CoarseDistanceMap and the candidate walk of FullSystem::activatePointsMT ----------------------
 * ImmaturePoint::my_type of the pool rows [first, first + n): 1 / 2 / 4, the `type` of ldso_gpu_select_points
 * (include/internal/ImmaturePoint.h, set in FullSystem::makeNewTraces :1449).  Rows made by immature_set /
 * immature_append start with 1. */
int ldso_gpu_immature_set_types(ldso_gpu_ctx* ctx, int first, int n, const float* type);
/* CoarseDistanceMap::makeDistanceMap + growDistBFS (src/frontend/CoarseTracker.cc:795-923) on the level-1 grid
 * (w/2 x h/2).  The n_points active points (host = window index, u, v, idepth_scaled) are projected with
 * KRKi[host] = K[1] R Ki[0] and Kt[host] = K[1] t of host -> newest frame (FullSystem.cc:1236-1238); cells they hit
 * start at 0, every other cell at 1000, and 39 rings grow from them, odd rings over 8 neighbours, even rings over 4,
 * never out of a border cell.  Points with host < 0 are skipped (the newest frame's own points). */
int ldso_gpu_distmap_make(ldso_gpu_ctx* ctx, int n_hosts, const float* KRKi, const float* Kt, int n_points,
                          const int* host, const float* u, const float* v, const float* idepth);
/* CoarseDistanceMap::fwdWarpedIDDistFinal, (w/2)*(h/2) floats */
int ldso_gpu_distmap_get(ldso_gpu_ctx* ctx, float* dist);
/* The walk over the immature points of activatePointsMT (FullSystem.cc:1226-1298) in pool order (= frames, then
 * features) against the current distance map, which every accepted point updates at once
 * (CoarseDistanceMap::addIntoDistFinal, CoarseTracker.cc:925-930) — the walk is order dependent.
 * action_out[n_pool]: 0 stays immature, 1 deleted (never traced successfully, outlier on the last trace, cannot be
 * activated and will be out, or projects outside the newest frame), 2 accepted for activation (toOptimize, pool order).
 * host_skip[h] != 0: frame h is the newest keyframe, its points are not visited; host_flagged[h]:
 * flaggedForMarginalization.  min_trace_quality = setting_minTraceQuality (3).  counts_out = {accepted, deleted}. */
int ldso_gpu_immature_select_activation(ldso_gpu_ctx* ctx, int n_hosts, const float* KRKi, const float* Kt,
                                        const unsigned char* host_skip, const unsigned char* host_flagged,
                                        float current_min_act_dist, float min_trace_quality, unsigned char* action_out,
                                        int counts_out[2]);

/* ---- N1: CoarseInitializer (src/frontend/CoarseInitializer.cc, include/frontend/CoarseInitializer.h) ----------
 * One initializer per ctx.  Point state (struct Pnt, CoarseInitializer.h:23-63) lives on the device per pyramid level.
 * setFirst (:656-737): makeK (:811-837); level 0 = PixelSelector::makeMaps(density 0.03 w h, recursions 1, thFactor 2)
 * with currentPotential reset to 3, levels > 0 = makePixelStatus (include/frontend/PixelSelector2.h:63-278) with
 * densities {0.05, 0.15, 0.5, 1} w h and the file-static sparsityFactor carried from level to level (starts at 5);
 * points are the selected pixels with patternPadding + 1 <= x < w_l - patternPadding - 2 (same for y) in row-major
 * order at (x + 0.1, y + 0.1), idepth = iR = 1; makeNN (:839-906): 10 nearest neighbours per point on its level
 * (the point itself comes first; equal distances resolve to the lower index — nanoflann's tie order is not specified)
 * and the nearest point of the next coarser level to (0.5 u - 0.25, 0.5 v - 0.25) as parent.
 * num_points_out[levels]. */
int ldso_gpu_init_set_first(ldso_gpu_ctx* ctx, int slot, int* num_points_out);
/* trackFrame (:45-183) of the pyramid in `slot` against the first frame: per level, coarse to fine, propagateDown
 * (:606-631), resetPoints (:739-761), then the Levenberg-Marquardt loop over calcResAndGS (:186-342 =
 * calculateAllResiduals :391-518 + calculateAlphaEnergy :376-389 + calculateSC :349-374), doStep (:763-789),
 * calcEC (:520-536), applyStep (:791-809), optReg (:538-567); finally propagateUp (:570-604).
 * exposure_first / exposure_new: ab_exposure of the two frames (:71-73).  Out: thisToNext as 12 doubles (R row-major,
 * then t), thisToNext_aff, snapped, the return value of trackFrame in *ready (snapped && frameID > snappedAt + 5),
 * latest_res[3] = the last level-0 {energy, alphaEnergy, num} of calcResAndGS, iterations per level (may be NULL). */
int ldso_gpu_init_track_frame(ldso_gpu_ctx* ctx, int slot, float exposure_first, float exposure_new, double pose_out[12],
                              float aff_out[2], int* snapped, int* ready, float latest_res[3], int iters_out[LDSO_MAX_LEVELS]);
/* Pnt fields of level lvl (any pointer may be NULL): u, v, idepth, iR, isGood, my_type, lastHessian, energy[0] */
int ldso_gpu_init_get_points(ldso_gpu_ctx* ctx, int lvl, int* n, float* u, float* v, float* idepth, float* iR,
                             unsigned char* is_good, float* my_type, float* last_hessian, float* energy);
/* neighbours [n][10] and parent [n] of level lvl (makeNN) */
int ldso_gpu_init_get_graph(ldso_gpu_ctx* ctx, int lvl, int* neighbours, int* parent);
/* resetPoints(lvl) (:739-761) + one calcResAndGS (:186-342) at level lvl for the given refToNew / affine, on the current point
 * state (as trackFrame starts a level, :84-85): H, Hsc 8x8 and
 * b, bsc as floats, res[3] = {energy, alphaEnergy, num}.  For parity tests of the per-point kernels. */
int ldso_gpu_init_calc_res_and_gs(ldso_gpu_ctx* ctx, int slot, int lvl, const double ref_to_new[12], float aff_a, float aff_b,
                                  float H[64], float b[8], float Hsc[64], float bsc[8], float res[3]);

/* ---- N3: FeatureDetector::DetectCorners (src/frontend/FeatureDetector.cc:34-124), the part before the descriptors ------
 * The image is cut into cells of gridsize = int(sqrtf(w h / n_features) + 0.5); in every cell away from the border
 * (skip = 2 HALF_PATCH_SIZE / gridsize + 1 cells) the pixels whose absSquaredGrad exceeds max(0.5 * cell maximum, 5) are
 * scored with ShiTomasiScore (include/frontend/FeatureDetector.h:50-81, 8x8 box of dIp gradients) and the best
 * floor(nfeatInGrid) + 1 become features, cells in column-major order (gx outer, gy inner), features of a cell by
 * descending score (equal scores keep the candidate order x outer, y inner: std::sort's order among ties is unspecified).
 * is_corner: score > 0.01 * (maximum score over ALL candidates), minus the lower-scored one of every pair of such
 * features closer than 5 px (:95-110).  IC_Angle and the rotated-BRIEF descriptor (:112-123) stay on the host.
 * Returns at most max_out features; *n_out = number written, *n_corners = how many carry is_corner. */
int ldso_gpu_detect_corners(ldso_gpu_ctx* ctx, int slot, int n_features, int max_out, float* x, float* y, float* score,
                            unsigned char* is_corner, int* n_out, int* n_corners);

/* ---- N4: photometric + geometric undistortion in front of makeImages (src/frontend/Undistort.cc) -----------------
 * Tables of an Undistort object, uploaded once: the raw sensor size, the inverse response G (PhotometricUndistorter::G,
 * g_depth = 256 or 65536 entries; NULL = no photometric calibration), vignetteMapInv [w_org*h_org] (NULL = none), and the
 * rectification lookup remapX / remapY [w*h] of the output size of this ctx (Undistort.cc:300-364 builds them on the host
 * from the camera model; NULL = passthrough, which needs w_org == w and h_org == h).  photometric_mode =
 * setting_photometricCalibration (0 off, 1 response only, 2 response + vignette). */
int ldso_gpu_undistort_set(ldso_gpu_ctx* ctx, int w_org, int h_org, int g_depth, const float* G, const float* vignette_inv,
                           const float* remap_x, const float* remap_y, int photometric_mode);
/* Undistort::undistort<T> (:366-458) = PhotometricUndistorter::processFrame<T> (:190-230: factor * raw when there is no
 * calibration or exposure_time <= 0, else G[raw] (* vignetteMapInv)) followed by the bilinear remap (out = 0 where
 * remapX < 0), then FrameHessian::makeImages on the result, all on the device: only the raw 8- or 16-bit image
 * (bytes_per_pixel 1 or 2) crosses the bus.  The benchmark-only noise / blur options (benchmark_varNoise,
 * benchmark_varBlurNoise) are not part of this path. */
int ldso_gpu_frame_make_images_raw(ldso_gpu_ctx* ctx, int slot, const void* raw, int bytes_per_pixel, float exposure_time,
                                   float factor);
/* FrameHessian::makeImages (FrameHessian.cc:59-128; with kind 2 / 3 behind Undistort::undistort<T>, as
 * ldso_gpu_frame_make_images_raw with factor 1) QUEUED without waiting: for a caller that was handed its frames up front
 * (ldso_gpu_batch_run prepares the next frame of a sequence while the current one is being tracked).  kind: 0 host float
 * image, 1 device float image, 2 / 3 host 8 / 16-bit sensor image (the LDSO_IMAGE_* values of include/ldso_system.h).  A
 * page-locked `image` must stay valid until the ctx's tracking stream has been waited for (ldso_gpu_sync, or the end of a
 * tracker call on that slot); a pageable one is consumed before the call returns. */
int ldso_gpu_frame_make_images_async(ldso_gpu_ctx* ctx, int slot, const void* image, int kind, float exposure_time);
/* ImageAndExposure::image of the last ldso_gpu_frame_make_images_raw: w*h floats */
int ldso_gpu_undistort_get_image(ldso_gpu_ctx* ctx, float* image_out);

/* ==== windowed bundle adjustment (K7-K10, H1) ======================================================
 * The window is described by the host (it owns the shared_ptr graph) as index arrays; the device keeps
 * points and residuals as SoA/AoS pools.  Index conventions follow the reference: pair tables are
 * indexed [host + n * target] (EnergyFunctional.cc:523-609), points are in allPoints order (frames ->
 * features, EnergyFunctional.cc:505-519; hence contiguous per host frame), residuals in p->residuals
 * order, flattened point-major with a CSR offset array. */
#define LDSO_MAX_WINDOW 12
#define LDSO_PRECALC_STRIDE 48
/* one FrameFramePrecalc (src/internal/FrameFramePrecalc.cc:6-35) as 48 floats:
 * [0..8] PRE_RTll_0 (row-major) [9..11] PRE_tTll_0 [12..20] PRE_KRKiTll [21..23] PRE_KtTll
 * [24..25] PRE_aff_mode [26] PRE_b0_mode [27..35] PRE_RTll [36..38] PRE_tTll [39] distanceLL [40..47] reserved */
/* ResState (include/internal/Residuals.h:33-35) */
#define LDSO_RES_IN 0
#define LDSO_RES_OOB 1
#define LDSO_RES_OUTLIER 2
/* residual flag bits */
#define LDSO_RF_LINEARIZED 1   /* PointFrameResidual::isLinearized */
#define LDSO_RF_ACTIVE 2       /* isActiveAndIsGoodNEW */
#define LDSO_RF_NEW 4          /* isNew */
/* RawResidualJacobian (include/internal/RawResidualJacobian.h:13-39) as 72 floats:
 * [0..7] resF [8..15] JIdx[0] [16..23] JIdx[1] [24..31] JabF[0] [32..39] JabF[1] [40..45] Jpdxi[0] [46..51] Jpdxi[1]
 * [52..55] Jpdc[0] [56..59] Jpdc[1] [60..61] Jpdd [62..64] JIdx2 (00,01,11) [65..68] JabJIdx (00,01,10,11)
 * [69..71] Jab2 (00,01,11) */
#define LDSO_J_FLOATS 72

/* frames of the window: pyramid slot of each, n*n precalc table, frameEnergyTH per frame, and
 * CalibHessian fxl,fyl,cxl,cyl (setPrecalcValues, src/frontend/FullSystem.cc:1667-1675) */
int ldso_gpu_window_set_frames(ldso_gpu_ctx* ctx, int n_frames, const int* slots, const float* precalc,
                               const float* frame_energy_th, const float calib[4]);
/* active points (PointHessian fields: u,v,idepth(_scaled),idepth_zero(_scaled),color[8],weights[8],priorF;
 * host = frame index) and the CSR offsets of their residual lists (n_points+1 entries) */
int ldso_gpu_window_set_points(ldso_gpu_ctx* ctx, int n_points, const float* u, const float* v, const float* idepth,
                               const float* idepth_zero, const int* host, const float* color, const float* weights,
                               const float* prior, const int* res_begin);
/* residuals (PointFrameResidual): target frame index, state_state, flag bits; optional res_to_zero [n][8] */
int ldso_gpu_window_set_residuals(ldso_gpu_ctx* ctx, int n_res, const int* target, const int* state,
                                  const unsigned char* flags, const float* res_to_zero);
/* EnergyFunctional::setAdjointsF / setDeltaF products (EnergyFunctional.cc:523-609): adHost, adTarget
 * [n*n][64] row-major double, adHTdeltaF [n*n][8], cDeltaF[4], cPrior[4], per-frame prior[8] and delta_prior[8] */
int ldso_gpu_window_set_adjoints(ldso_gpu_ctx* ctx, const double* ad_host, const double* ad_target, const float* ad_ht_delta,
                                 const float c_delta[4], const double c_prior[4], const double* frame_prior,
                                 const double* frame_delta_prior);
/* new point depths after a step (doStepFromBackup, FullSystem.cc:1826-1932) without re-sending the window */
int ldso_gpu_window_set_idepth(ldso_gpu_ctx* ctx, const float* idepth, const float* idepth_zero);

/* ---- K7: FullSystem::linearizeAll -> PointFrameResidual::linearize (src/internal/Residuals.cc:15-217)
 * over every residual that is not linearized; stats = [sum energy, 0, #state_NewState==IN]
 * (FullSystem.cc:1716-1741).  Also runs setNewFrameEnergyTH (:2078-2109) for the newest frame and returns
 * it in new_frame_energy_th.  fix_linearization != 0 additionally does applyRes(true) and the isNew
 * bookkeeping (maxRelBaseline, numGoodResiduals) of linearizeAll_Reductor (:1786-1820). */
int ldso_gpu_window_linearize(ldso_gpu_ctx* ctx, int fix_linearization, double stats[3], float* new_frame_energy_th);
/* FullSystem::applyRes_Reductor(true) (FullSystem.cc:2022-2027; include/internal/Residuals.h:70-88,124-129) */
int ldso_gpu_window_apply_res(ldso_gpu_ctx* ctx);
/* PointFrameResidual::resetOOB for every non-linearized residual (FullSystem.cc:864-867) */
int ldso_gpu_window_reset_oob(ldso_gpu_ctx* ctx);
/* PointFrameResidual::fixLinearizationF (Residuals.cc:219-245) for the residuals whose mask byte is non-zero */
int ldso_gpu_window_fix_linearization(ldso_gpu_ctx* ctx, const unsigned char* mask);

/* ---- K8 + K9: EnergyFunctional::accumulateAF_MT / accumulateLF_MT / accumulateSCF_MT (EnergyFunctional.cc:670-749)
 * = AccumulatedTopHessianSSE::addPoint<0>, <1> + stitchDoubleMT, AccumulatedSCHessianSSE::addPoint + stitch.
 * Each H is (8n+4)^2 row-major double, each b is 8n+4.  counts = [resInA, resInL]. */
int ldso_gpu_window_accumulate(ldso_gpu_ctx* ctx, double* HA, double* bA, double* HL, double* bL, double* Hsc, double* bsc,
                               int counts[2]);
/* ---- H1: the dense part of EnergyFunctional::solveSystemF (EnergyFunctional.cc:310-432), host code.
 * HM,bM: marginalisation prior; delta: getStitchedDeltaF(); N: (8n+4) x n_null nullspace columns (may be NULL)
 * used by orthogonalize() when iteration >= 2.  x_out: 8n+4. */
int ldso_gpu_window_solve(ldso_gpu_ctx* ctx, int iteration, double lambda, const double* HA, const double* bA,
                          const double* HL, const double* bL, const double* Hsc, const double* bsc, const double* HM,
                          const double* bM, const double* delta, const double* N, int n_null, double* x_out);
/* ---- K10: EnergyFunctional::resubstituteF_MT / resubstituteFPt (EnergyFunctional.cc:611-667): per-point step */
int ldso_gpu_window_resubstitute(ldso_gpu_ctx* ctx, const double* x, float lambda, float* step_out);

/* ---- marginalisation (EnergyFunctional.cc:109-262) and the energies the step acceptance test reads (:473-497, :751-806)
 *
 * marginalizePointsF (:205-262): marg[np] != 0 marks the points whose status is MARGINALIZED.  Their prior is multiplied
 * by idepth_fix_prior_marg_fac (setting_idepthFixPriorMargFac = 600^2), AccumulatedTopHessianSSE::addPoint<2>
 * (AccumulatedTopHessian.cc:9-118, resApprox = res_toZeroF) and AccumulatedSCHessianSSE::addPoint(p, false) run over
 * them, both accumulators are stitched without priors and H = M - Msc, b = Mb - Mbsc are returned ((8n+4)^2, 8n+4).
 * The caller keeps HM,bM on the host and adds marg_weight_fac * H, b (setting_margWeightFac = 0.25) — with the default
 * solver mode no orthogonalisation happens here.  n_res = residuals that went in (resInM +=).  As removePoint does,
 * the residuals of those points are inactive afterwards. */
int ldso_gpu_window_marginalize_points(ldso_gpu_ctx* ctx, const unsigned char* marg, float idepth_fix_prior_marg_fac,
                                       double* H, double* b, int* n_res);
/* The marginalisation of a keyframe in one call, on the window that is already resident (the one the last
 * ldso_gpu_window_linearize(fix = 1) ran on; nothing is uploaded again): for the points with point_mask != 0 the inner loop of
 * FullSystem::flagPointsForRemoval (src/frontend/FullSystem.cc:1390-1399: resetOOB, linearize, applyRes, fixLinearizationF of the
 * active residuals) followed by EnergyFunctional::marginalizePointsF (as ldso_gpu_window_marginalize_points).  residual_dead[r]
 * != 0: the host has dropped residual r since the upload (linearizeAll(true)'s removal list).  frame_energy_th[n]: the
 * thresholds as they stand now (setNewFrameEnergyTH ran after the last linearisation).  n_lin: residuals whose linearisation
 * was fixed.  Afterwards the resident window is spent: upload it again before the next optimisation. */
int ldso_gpu_window_marginalize_subset(ldso_gpu_ctx* ctx, const unsigned char* point_mask, const unsigned char* residual_dead,
                                       const float* frame_energy_th, float idepth_fix_prior_marg_fac, double* H, double* b,
                                       int* n_res, int* n_lin);
/* marginalizeFrame (:109-191), the dense part (H1, host code): moves frame idx (0-based window position, n frames) to
 * the end of HM,bM ((8n+4)^2, 8n+4), adds its prior / prior*delta_prior, takes the diagonally scaled Schur complement
 * with util::MatrixInverter::invertPosDef (src/util/MatrixInverter.cc:31-50, SVD form) of the 8x8 block and
 * symmetrises.  HM_out,bM_out: (8(n-1)+4)^2, 8(n-1)+4.  Needs no window state. */
int ldso_gpu_window_marginalize_frame(ldso_gpu_ctx* ctx, int n, int idx, const double* HM, const double* bM,
                                      const double prior[8], const double delta_prior[8], double* HM_out, double* bM_out);
/* calcLEnergyF_MT (:481-497, calcLEnergyPt :751-806) and calcMEnergyF (:473-479).  energy[0] = linearised energy
 * (frame priors + calibration prior + sum over linearised active residuals of (2 res_toZeroF + J delta) . J delta +
 * per-point deltaF^2 priorF), energy[1] = delta . (2 bM + HM delta).  Uses the adjoint / delta tables of
 * ldso_gpu_window_set_adjoints and the idepths of set_points / set_idepth. */
int ldso_gpu_window_calc_energy(ldso_gpu_ctx* ctx, const double* HM, const double* bM, const double* delta, double energy[2]);

/* read-backs (host mirrors of public members other reference code reads directly) */
/* what: 0 J [n][72], 1 res_toZeroF [n][8], 2 JpJdF [n][8], 3 centerProjectedTo [n][3],
 *       4 state_NewEnergy, 5 state_NewEnergyWithOutlier, 6 state_energy  (float each) */
int ldso_gpu_window_get_residual_floats(ldso_gpu_ctx* ctx, int what, float* dst);
/* what: 0 state_state, 1 state_NewState, 2 flag bits */
int ldso_gpu_window_get_residual_ints(ldso_gpu_ctx* ctx, int what, int* dst);
/* what: 0 Hdd_accAF 1 bd_accAF 2 Hcd_accAF[4] 3 Hdd_accLF 4 bd_accLF 5 Hcd_accLF[4] 6 HdiF 7 bdSumF
 *       8 idepth_hessian 9 maxRelBaseline 10 numGoodResiduals (as float) 11 step */
int ldso_gpu_window_get_point_floats(ldso_gpu_ctx* ctx, int what, float* dst);

/* everything FullSystem reads back after linearizeAll(true) in one call (one synchronisation): state_state [n_res]
 * (-> PointHessian::lastResiduals, FullSystem.cc:1745-1751), centerProjectedTo [n_res][3] and HdiF [n_points]
 * (-> CoarseTracker::makeCoarseDepthL0, CoarseTracker.cc:363-382), idepth_hessian, numGoodResiduals (as float) and
 * maxRelBaseline [n_points] (-> flagPointsForRemoval, FullSystem.cc:1377-1417).  Any pointer may be NULL. */
int ldso_gpu_window_get_keyframe_results(ldso_gpu_ctx* ctx, int* state, float* center3, float* hdi, float* idepth_hessian,
                                         float* num_good, float* max_rel_baseline);

#ifdef __cplusplus
}
#endif
#endif /* LDSO_GPU_H_ */
