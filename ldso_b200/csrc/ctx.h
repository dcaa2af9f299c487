// ctx.h — per-sequence device context of libldso_gpu.so (internal; the public face is include/ldso_gpu.h).
//
// One ctx = one sequence: replaces the reference's process-global calibration
// (src/internal/GlobalCalib.cc:7-18) and settings (src/Setting.cc) so that many sequences can live in
// one process (BASELINE.json config 5).  All device memory is allocated once in ldso_gpu_create.
#pragma once
#include <cuda_runtime.h>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/ldso_gpu.h"
#include "../../include/ldso_smallmat.h"
#include "window.h"

#define LDSO_CUDA_OK(ctx, expr)                                                                      \
    do {                                                                                             \
        cudaError_t _e = (expr);                                                                     \
        if (_e != cudaSuccess) {                                                                     \
            (ctx)->set_error(std::string(#expr) + ": " + cudaGetErrorString(_e), __FILE__, __LINE__); \
            return -1;                                                                               \
        }                                                                                            \
    } while (0)

#define LDSO_REQUIRE(ctx, cond)                                                      \
    do {                                                                             \
        if (!(cond)) {                                                               \
            (ctx)->set_error(std::string("bad argument: ") + #cond, __FILE__, __LINE__); \
            return -2;                                                               \
        }                                                                            \
    } while (0)

// include/Settings.h:28-35
#define LDSO_SCALE_XI_ROT 1.0f
#define LDSO_SCALE_XI_TRANS 0.5f
#define LDSO_SCALE_A 10.0f
#define LDSO_SCALE_B 1000.0f

// values reduced by the fused K3+K4 launch (see k34_tracker.cu)
#define TRK_NRED 56          // 45 (9x9 upper) + E + flowT + flowRT + pad ; ints kept apart

struct TrackerDev {
    // makeK (CoarseTracker.cc:312-339), host copies
    int w[LDSO_MAX_LEVELS], h[LDSO_MAX_LEVELS];
    float fx[LDSO_MAX_LEVELS], fy[LDSO_MAX_LEVELS], cx[LDSO_MAX_LEVELS], cy[LDSO_MAX_LEVELS];
    float K[LDSO_MAX_LEVELS][9], Ki[LDSO_MAX_LEVELS][9];
    // device: per-level idepth / weightSums / weightSums_bak maps and the compacted point cloud
    float* idepth[LDSO_MAX_LEVELS];
    float* wsum[LDSO_MAX_LEVELS];
    float* wsum_bak[LDSO_MAX_LEVELS];
    float4* pc[LDSO_MAX_LEVELS];   // {u, v, idepth, color}  (pc_u, pc_v, pc_idepth, pc_color)
    int* row_cnt;                  // per-row valid counts / offsets for the ordered compaction (all levels)
    int* pc_n_dev;                 // [LDSO_MAX_LEVELS]
    int pc_n[LDSO_MAX_LEVELS];     // host mirror (valid after trk_pc_n_sync)
    int* pc_n_pinned = nullptr;    // [LDSO_MAX_LEVELS] pinned: setCoarseTrackingRef's counts arrive here without a wait
    cudaEvent_t ev_pc_n = nullptr; // recorded behind that copy
    bool pc_n_valid = true;
    int row_base[LDSO_MAX_LEVELS + 1];
    int ref_slot = -1;
    float ref_exposure = 1, ref_aff_a = 0, ref_aff_b = 0;
    // state of the last calc_res (for calc_gs)
    int last_lvl = -1, new_slot = -1;
    float new_exposure = 1;
    double last_sums[48];
    int last_counts[4];
    float last_aff_a = 0, last_aff_b = 0;
};

struct ldso_gpu_ctx {
    ldso_gpu_config cfg;
    int levels = 0;
    int wG[LDSO_MAX_LEVELS], hG[LDSO_MAX_LEVELS];
    size_t lvl_off[LDSO_MAX_LEVELS + 1];   // pixel offset of each level inside a frame slot
    cudaStream_t stream = nullptr;          // tracking stream
    cudaStream_t stream_map = nullptr;      // mapping stream
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev_h2d = nullptr, ev_trk = nullptr;
    float4** frame = nullptr;               // [n_frame_slots] device float4 pyramids {I,dx,dy,absSquaredGrad}
    float* img_dev = nullptr;               // staged input image
    float* img_pinned = nullptr;
    float* gammaB_dev = nullptr;            // B[256]
    float* scratch_host = nullptr;          // pinned scratch, >= 16 B * pixels of level 0
    TrackerDev trk[LDSO_NUM_TRACKERS];
    // reduction workspace of the fused tracker launch
    int trk_max_blocks = 0;                 // ceil(w*h / 128): every level-0 pixel could be a tracker point
    float* trk_partials = nullptr;          // [trk_max_blocks][TRK_NRED]
    int* trk_partials_i = nullptr;          // [trk_max_blocks][4]
    unsigned int* trk_counter = nullptr;
    double* trk_result_dev = nullptr;       // [64]
    double* trk_result_pinned = nullptr;    // [64] scratch for small read-backs
    double* trk_mailbox = nullptr;          // [64] mapped pinned: the fused tracker kernel writes its sums + a sequence number here
    double* trk_mailbox_dev = nullptr;      // device alias of trk_mailbox
    unsigned long long trk_seq = 0;
    // ldso_gpu_tracker_track_begin .. _end: the launch in flight and the arguments of the call
    struct TrackPending {
        bool active = false;
        int tracker = 0, new_slot = 0, coarsest = 0, prof_idx = -1, dbg = 0;
        float new_exposure = 1, aff[2] = {0, 0}, min_res[5] = {0, 0, 0, 0, 0};
        double pose[12] = {0}, seq = 0;
    } trk_pending;
    double lm_dbg_clocks[8] = {0};          // cfg.reserved[0] == 3: SM clocks per LM phase (measurement only)
    int trk_coop_max = 0;                   // co-resident CTA budget of the cooperative LM kernel (tracker_lm_mode = 2)
    int trk_cluster = 0;                    // CTAs of the tracker's thread-block cluster (tracker_lm_mode = 0), chosen at first use
    bool k1_attr_set = false, k8_attr_set = false;   // opt-in dynamic shared memory set for this ctx's device
    int sm_count = 148;                     // multiprocessors of the ctx's device
    // calls on one ctx are serialised: the reference's tracking thread (trackMutex) and mapping thread (mapMutex) may both
    // hold a pointer to it, and the two trackers, the pinned scratch and the profile lists are shared state
    std::recursive_mutex mu;
    WindowDev win;
    void* selector = nullptr;               // SelectorDev (a2_selector.cu), created on first use
    void* immature = nullptr;               // ImmatureDev (k56_immature.cu), created on first use
    void* initializer = nullptr;            // InitDev (n1_initializer.cu), created on first use
    void* undistort = nullptr;              // UndistortDev (n4_undistort.cu), created by ldso_gpu_undistort_set
    void* corners = nullptr;                // N3Dev (n3_corners.cu), created on first use
    long long launches = 0;
    long long h2d_bytes = 0, d2h_bytes = 0;   // bytes moved by the entry points (bench.py e2e accounting)
    // optional per-kernel CUDA-event profile (ldso_gpu_profile_*); slots: see LDSO_PROF_* in include/ldso_gpu.h.
    // Event pairs are recorded around launches and resolved lazily in profile_get, so profiling adds no sync.
    int profile_on = 0;
    struct ProfRec { int which; double units; cudaEvent_t e0, e1; };
    std::vector<ProfRec> prof_pending;
    std::vector<cudaEvent_t> prof_free;
    cudaEvent_t prof_cur = nullptr;
    double prof_ms[LDSO_PROF_SLOTS] = {0};
    long long prof_launches[LDSO_PROF_SLOTS] = {0};
    double prof_units[LDSO_PROF_SLOTS] = {0};
    std::string err;

    void set_error(const std::string& what, const char* file, int line) {
        char buf[64];
        snprintf(buf, sizeof(buf), " (%s:%d)", strrchr(file, '/') ? strrchr(file, '/') + 1 : file, line);
        err = what + buf;
    }
};

// first statement of every entry point that takes a ctx: serialise on the ctx and make its device current for the calling
// thread (cudaSetDevice is per thread; a second host thread would otherwise run against device 0)
#define LDSO_ENTER(ctx)                                               \
    std::lock_guard<std::recursive_mutex> ldso_enter_guard((ctx)->mu); \
    do {                                                              \
        int ldso_cur_dev = -1;                                        \
        if (cudaGetDevice(&ldso_cur_dev) != cudaSuccess || ldso_cur_dev != (ctx)->cfg.device) cudaSetDevice((ctx)->cfg.device); \
    } while (0)

// How a host thread waits for the device.  Default: spin (lowest latency, one core per waiting thread).  With
// LDSO_WAIT_YIELD=1 the mailbox polls call sched_yield() and the context is switched to cudaDeviceScheduleYield, so that
// more sequence threads than cores can share a host (ldso_gpu_batch_* with many sequences per GPU).
bool ldso_wait_yield();
// Every wait of the library goes through these two (and ldso_cpu_relax for the mailbox polls).  Without a hook they are
// cudaStreamSynchronize / cudaEventSynchronize.  With a hook installed for the calling thread (ldso_gpu_set_wait_hook: the
// batch runner's cooperative scheduler) they poll and call the hook between polls, so the thread runs another sequence
// instead of spinning.
cudaError_t ldso_stream_sync(cudaStream_t s);
cudaError_t ldso_event_sync(cudaEvent_t e);
bool ldso_wait_block();   // LDSO_WAIT_BLOCK=1: waits sleep in the driver (cudaDeviceScheduleBlockingSync, blocking events): no core is burnt while waiting
void ldso_cpu_relax();

// cudaFuncAttributeMaxDynamicSharedMemorySize of `fn` on the ctx's device is raised to at least `bytes` (never lowered: other
// contexts of the process launch the same kernel with their own sizes).  slot: one per kernel, see api.cu
int ldso_raise_smem(ldso_gpu_ctx* c, const void* fn, size_t bytes, int slot);

// Pinned staging ring of the ctx (window_api.cu): a block for a queued host-to-device copy on the TRACKING stream; the caller may
// return without waiting (the ring is rewound only after a synchronisation of that stream)
void* ldso_staging(ldso_gpu_ctx* c, size_t bytes);

// every host<->device copy of the library goes through here so that the traffic is accounted
inline cudaError_t ldso_copy(ldso_gpu_ctx* c, void* dst, const void* src, size_t bytes, cudaMemcpyKind kind, cudaStream_t s) {
    if (kind == cudaMemcpyHostToDevice) c->h2d_bytes += (long long)bytes;
    else if (kind == cudaMemcpyDeviceToHost) c->d2h_bytes += (long long)bytes;
    return cudaMemcpyAsync(dst, src, bytes, kind, s);
}

// kernels (one translation unit each)
void selector_free(ldso_gpu_ctx* c);
void prof_begin(ldso_gpu_ctx* c, cudaStream_t s);
int prof_end(ldso_gpu_ctx* c, cudaStream_t s, int which, double units);   // index into prof_pending or -1
void prof_drain(ldso_gpu_ctx* c);
int immature_trace_fence(ldso_gpu_ctx* c, int slot);   // before a frame slot is rewritten on the tracking stream
int immature_trace_finalize(ldso_gpu_ctx* c);   // completes a trace launched without waiting (k56_immature.cu)
void immature_free(ldso_gpu_ctx* c);
void initializer_free(ldso_gpu_ctx* c);
void undistort_free(ldso_gpu_ctx* c);
void corners_free(ldso_gpu_ctx* c);
int k1_make_images(ldso_gpu_ctx* c, int slot, const float* img_dev);
int k2_set_ref(ldso_gpu_ctx* c, int tracker, int ref_slot, int n, const float* u, const float* v, const float* id,
               const float* hdi);
struct TrackEvalOut {
    double sums[48];   // 45 H entries (9x9 upper, row-major), E, flowT, flowRT
    int counts[4];     // numTermsInE, numSaturated, numTermsInWarped, flowNum
};
// the host mirror of the tracker's point counts: waits for the copy queued by k2_set_ref if it has not been read yet
int trk_pc_n_sync(ldso_gpu_ctx* c, TrackerDev& T);
int k34_track_all_launch(ldso_gpu_ctx* c, int tracker, int coarsest_lvl, const int* maxIt, const float* min_res_for_abort,
                         const ldso_sm::SE3& cur, float aff_a, float aff_b);
int k34_track_all_wait(ldso_gpu_ctx* c, ldso_sm::SE3& cur, float& aff_a, float& aff_b, float last_residuals[5], float flow[3], int* status,
                       int* nEval, int* nAccept);
int k34_eval(ldso_gpu_ctx* c, int tracker, int lvl, const ldso_sm::SE3& refToNew, float aff_a, float aff_b, float cutoffTH,
             TrackEvalOut* out);
