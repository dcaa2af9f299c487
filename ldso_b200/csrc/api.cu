// api.cu — context life cycle and the C-ABI entry points of the photometric front-end
// (frame pyramids, coarse tracker).  See include/ldso_gpu.h for the reference methods each one replaces.
#include "ctx.h"

#include <cuda.h>
#include <atomic>

#include <immintrin.h>
#include <sched.h>

#include <cmath>
#include <new>
#include <vector>

using ldso_sm::SE3;

extern "C" void ldso_gpu_default_config(ldso_gpu_config* c) {
    // src/Setting.cc:8-182 defaults, preset 0 (examples/run_dso_tum_mono.cc:65-83)
    memset(c, 0, sizeof(*c));
    c->w = 640; c->h = 480;
    c->fx = 384; c->fy = 384; c->cx = 319.5f; c->cy = 239.5f;
    c->device = 0;
    c->n_frame_slots = 10;
    c->max_points = 4096; c->max_immature = 16384; c->max_residuals = 32768;
    c->huber_th = 9; c->coarse_cutoff_th = 20; c->gamma_weights_pixel_select = 1;
    c->affine_opt_mode_a = 1e12f; c->affine_opt_mode_b = 1e8f;
    c->outlier_th = 144; c->outlier_th_sum_component = 2500; c->overall_energy_th_weight = 1;
    c->max_pix_search = 0.027f; c->trace_stepsize = 1.0f; c->trace_gn_iterations = 3;
    c->trace_gn_threshold = 0.1f; c->trace_extra_slack_on_th = 1.2f; c->trace_slack_interval = 1.5f;
    c->trace_min_improvement = 2; c->min_trace_test_radius = 2;
    c->min_grad_hist_cut = 0.5f; c->min_grad_hist_add = 7; c->grad_downweight_per_level = 0.75f;
    c->select_direction_distribution = 1;
    c->idepth_fix_prior = 2500; c->min_idepth_h_act = 100; c->gn_its_on_point_activation = 3;
    c->frame_energy_th_n = 0.7f; c->frame_energy_th_fac_median = 1.5f; c->frame_energy_th_const_weight = 0.5f;
    c->tracker_solve_mode = 0;
}

static int create_impl(ldso_gpu_ctx* c) {
    const ldso_gpu_config& cfg = c->cfg;
    LDSO_REQUIRE(c, cfg.w > 8 && cfg.h > 8 && cfg.n_frame_slots > 0);
    int ndev = 0;
    LDSO_CUDA_OK(c, cudaGetDeviceCount(&ndev));
    LDSO_REQUIRE(c, cfg.device >= 0 && cfg.device < ndev);
    LDSO_CUDA_OK(c, cudaSetDevice(cfg.device));
    LDSO_CUDA_OK(c, cudaDeviceGetAttribute(&c->sm_count, cudaDevAttrMultiProcessorCount, cfg.device));
    // pyramid level rule of setGlobalCalib (src/internal/GlobalCalib.cc:20-28)
    int wl = cfg.w, hl = cfg.h, lv = 1;
    while (wl % 2 == 0 && hl % 2 == 0 && wl * hl > 5000 && lv < LDSO_MAX_LEVELS) { wl /= 2; hl /= 2; lv++; }
    c->levels = lv;
    size_t off = 0;
    for (int l = 0; l < lv; l++) {
        c->wG[l] = cfg.w >> l;
        c->hG[l] = cfg.h >> l;
        c->lvl_off[l] = off;
        off += (size_t)c->wG[l] * c->hG[l];
    }
    c->lvl_off[lv] = off;
    const size_t npx0 = (size_t)cfg.w * cfg.h;
    LDSO_CUDA_OK(c, cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    LDSO_CUDA_OK(c, cudaStreamCreateWithFlags(&c->stream_map, cudaStreamNonBlocking));
    LDSO_CUDA_OK(c, cudaEventCreate(&c->ev0));
    LDSO_CUDA_OK(c, cudaEventCreate(&c->ev1));
    LDSO_CUDA_OK(c, cudaEventCreateWithFlags(&c->ev_h2d, cudaEventDisableTiming | (ldso_wait_block() ? cudaEventBlockingSync : 0)));
    LDSO_CUDA_OK(c, cudaEventCreateWithFlags(&c->ev_trk, cudaEventDisableTiming | (ldso_wait_block() ? cudaEventBlockingSync : 0)));
    c->frame = new float4*[cfg.n_frame_slots]();
    for (int s = 0; s < cfg.n_frame_slots; s++) LDSO_CUDA_OK(c, cudaMalloc(&c->frame[s], off * sizeof(float4)));
    LDSO_CUDA_OK(c, cudaMalloc(&c->img_dev, npx0 * sizeof(float)));
    LDSO_CUDA_OK(c, cudaMallocHost(&c->img_pinned, npx0 * sizeof(float)));
    LDSO_CUDA_OK(c, cudaMallocHost(&c->scratch_host, npx0 * sizeof(float4)));
    LDSO_CUDA_OK(c, cudaMalloc(&c->gammaB_dev, 256 * sizeof(float)));
    float B[256];
    for (int i = 0; i < 256; i++) B[i] = (float)i;
    LDSO_CUDA_OK(c, cudaMemcpy(c->gammaB_dev, B, sizeof(B), cudaMemcpyHostToDevice));
    // trackers (CoarseTracker::CoarseTracker, src/frontend/CoarseTracker.cc:30-59)
    for (int t = 0; t < LDSO_NUM_TRACKERS; t++) {
        TrackerDev& T = c->trk[t];
        T.row_base[0] = 0;
        for (int l = 0; l < lv; l++) {
            size_t n = (size_t)c->wG[l] * c->hG[l];
            LDSO_CUDA_OK(c, cudaMalloc(&T.idepth[l], n * sizeof(float)));
            LDSO_CUDA_OK(c, cudaMalloc(&T.wsum[l], n * sizeof(float)));
            LDSO_CUDA_OK(c, cudaMalloc(&T.wsum_bak[l], n * sizeof(float)));
            LDSO_CUDA_OK(c, cudaMalloc(&T.pc[l], n * sizeof(float4)));
            T.pc_n[l] = 0;
            int rows = c->hG[l] - 4;
            T.row_base[l + 1] = T.row_base[l] + (rows > 0 ? rows : 0);
        }
        LDSO_CUDA_OK(c, cudaMalloc(&T.row_cnt, (T.row_base[lv] + 1) * sizeof(int)));
        LDSO_CUDA_OK(c, cudaMalloc(&T.pc_n_dev, LDSO_MAX_LEVELS * sizeof(int)));
        LDSO_CUDA_OK(c, cudaMemset(T.pc_n_dev, 0, LDSO_MAX_LEVELS * sizeof(int)));
        LDSO_CUDA_OK(c, cudaMallocHost((void**)&T.pc_n_pinned, LDSO_MAX_LEVELS * sizeof(int)));
        memset(T.pc_n_pinned, 0, LDSO_MAX_LEVELS * sizeof(int));
        LDSO_CUDA_OK(c, cudaEventCreateWithFlags(&T.ev_pc_n, cudaEventDisableTiming));
        int rc = ldso_gpu_tracker_make_k(c, t, cfg.fx, cfg.fy, cfg.cx, cfg.cy);
        if (rc) return rc;
    }
    c->trk_max_blocks = (int)((npx0 + 127) / 128);
    LDSO_CUDA_OK(c, cudaMalloc(&c->trk_partials, (size_t)c->trk_max_blocks * TRK_NRED * sizeof(float)));
    LDSO_CUDA_OK(c, cudaMalloc(&c->trk_partials_i, (size_t)c->trk_max_blocks * 4 * sizeof(int)));
    LDSO_CUDA_OK(c, cudaMalloc(&c->trk_counter, sizeof(unsigned int)));
    LDSO_CUDA_OK(c, cudaMemset(c->trk_counter, 0, sizeof(unsigned int)));
    LDSO_CUDA_OK(c, cudaMalloc(&c->trk_result_dev, 64 * sizeof(double)));
    LDSO_CUDA_OK(c, cudaMallocHost(&c->trk_result_pinned, 64 * sizeof(double)));
    LDSO_CUDA_OK(c, cudaHostAlloc(&c->trk_mailbox, 64 * sizeof(double), cudaHostAllocMapped));
    LDSO_CUDA_OK(c, cudaHostGetDevicePointer((void**)&c->trk_mailbox_dev, c->trk_mailbox, 0));
    memset(c->trk_mailbox, 0, 64 * sizeof(double));
    if (win_alloc(c)) return -1;
    LDSO_CUDA_OK(c, cudaDeviceSynchronize());
    return 0;
}

int ldso_raise_smem(ldso_gpu_ctx* c, const void* fn, size_t bytes, int slot) {
    constexpr int MAXDEV = 64, NSLOT = 4;
    static size_t have[MAXDEV][NSLOT] = {};
    static std::mutex mu;
    const int dev = c->cfg.device;
    LDSO_REQUIRE(c, dev >= 0 && dev < MAXDEV && slot >= 0 && slot < NSLOT);
    std::lock_guard<std::mutex> lk(mu);
    if (bytes <= have[dev][slot]) return 0;
    LDSO_CUDA_OK(c, cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    have[dev][slot] = bytes;
    return 0;
}

bool ldso_wait_yield() {
    static const bool y = [] { const char* e = getenv("LDSO_WAIT_YIELD"); return e && atoi(e) != 0; }();
    return y;
}
bool ldso_wait_block() {
    static const bool b = [] { const char* e = getenv("LDSO_WAIT_BLOCK"); return e && atoi(e) != 0; }();
    return b;
}
namespace {
thread_local void (*t_wait_hook)(void*) = nullptr;
thread_local void* t_wait_arg = nullptr;
// cudaErrorNotReady from a query is not a failure: make sure it does not linger as the thread's "last error"
inline void forget_not_ready() {
    if (cudaPeekAtLastError() == cudaErrorNotReady) (void)cudaGetLastError();
}
}  // namespace
extern "C" int ldso_gpu_set_wait_hook(void (*hook)(void*), void* arg) {
    t_wait_hook = hook; t_wait_arg = arg;
    return 0;
}
void ldso_cpu_relax() {
    if (t_wait_hook) t_wait_hook(t_wait_arg);
    else if (ldso_wait_yield()) sched_yield();
    else _mm_pause();
}
namespace {
// Hooked stream wait without driver calls in the poll loop: a stream memory operation (cuStreamWriteValue32, obtained through
// the runtime so that the library does not link against libcuda) writes a sequence number into mapped pinned memory behind
// the work queued so far, and the host polls that word.  One 4 KB pinned block per waiting thread, one word per stream.
typedef CUresult (*StreamWrite32Fn)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);
StreamWrite32Fn stream_write32() {
    static const StreamWrite32Fn fn = [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (getenv("LDSO_WAIT_QUERY")) return (StreamWrite32Fn) nullptr;     // measurement knob: poll with cudaStreamQuery instead
        if (cudaGetDriverEntryPoint("cuStreamWriteValue32", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) {
            (void)cudaGetLastError();
            return (StreamWrite32Fn) nullptr;
        }
        return (StreamWrite32Fn)p;
    }();
    return fn;
}
struct WaitSlots {
    volatile unsigned* host = nullptr;
    unsigned* dev = nullptr;
    int n = 0;
    cudaStream_t stream[1024];
    unsigned seq[1024];
    bool failed = false;
};
thread_local WaitSlots t_slots;
int wait_slot(cudaStream_t s) {
    WaitSlots& W = t_slots;
    for (int i = 0; i < W.n; i++) if (W.stream[i] == s) return i;
    if (W.failed || W.n >= 1024) return -1;
    if (!W.host) {
        void* h = nullptr; void* d = nullptr;
        if (cudaHostAlloc(&h, 4096, cudaHostAllocMapped | cudaHostAllocPortable) != cudaSuccess || cudaHostGetDevicePointer(&d, h, 0) != cudaSuccess) {
            (void)cudaGetLastError();
            W.failed = true;
            return -1;
        }
        memset(h, 0, 4096);
        W.host = (volatile unsigned*)h; W.dev = (unsigned*)d;
    }
    W.stream[W.n] = s; W.seq[W.n] = 0;
    return W.n++;
}
}  // namespace
cudaError_t ldso_stream_sync(cudaStream_t s) {
    if (!t_wait_hook) return cudaStreamSynchronize(s);
    const StreamWrite32Fn wr = stream_write32();
    const int slot = wr ? wait_slot(s) : -1;
    if (slot >= 0) {
        WaitSlots& W = t_slots;
        const unsigned want = ++W.seq[slot];
        if (wr((CUstream)s, (CUdeviceptr)(W.dev + slot), want, 0) == CUDA_SUCCESS) {
            for (unsigned spin = 1; W.host[slot] != want; spin++) {
                t_wait_hook(t_wait_arg);
                if ((spin & 0xfff) == 0) {                 // now and then make sure the stream is still alive
                    const cudaError_t e = cudaStreamQuery(s);
                    if (e != cudaErrorNotReady && e != cudaSuccess) return e;
                    if (e == cudaErrorNotReady) forget_not_ready();
                }
            }
            std::atomic_thread_fence(std::memory_order_acquire);
            return cudaSuccess;
        }
    }
    for (bool waited = false;; waited = true) {
        const cudaError_t e = cudaStreamQuery(s);
        if (e != cudaErrorNotReady) { if (waited) forget_not_ready(); return e; }
        t_wait_hook(t_wait_arg);
    }
}
cudaError_t ldso_event_sync(cudaEvent_t ev) {
    if (!t_wait_hook) return cudaEventSynchronize(ev);
    for (bool waited = false;; waited = true) {
        const cudaError_t e = cudaEventQuery(ev);
        if (e != cudaErrorNotReady) { if (waited) forget_not_ready(); return e; }
        t_wait_hook(t_wait_arg);
    }
}

extern "C" int ldso_gpu_create(const ldso_gpu_config* cfg, ldso_gpu_ctx** out) {
    if (!cfg || !out) return -2;
    ldso_gpu_ctx* c = new (std::nothrow) ldso_gpu_ctx();
    if (!c) return -3;
    c->cfg = *cfg;
    for (int t = 0; t < LDSO_NUM_TRACKERS; t++) {
        memset(c->trk[t].idepth, 0, sizeof(c->trk[t].idepth));
        memset(c->trk[t].wsum, 0, sizeof(c->trk[t].wsum));
        memset(c->trk[t].wsum_bak, 0, sizeof(c->trk[t].wsum_bak));
        memset(c->trk[t].pc, 0, sizeof(c->trk[t].pc));
        c->trk[t].row_cnt = nullptr;
        c->trk[t].pc_n_dev = nullptr;
    }
    if (ldso_wait_block()) { cudaSetDevice(cfg->device); cudaSetDeviceFlags(cudaDeviceScheduleBlockingSync); cudaGetLastError(); }
    else if (ldso_wait_yield()) { cudaSetDevice(cfg->device); cudaSetDeviceFlags(cudaDeviceScheduleYield); cudaGetLastError(); }
    int rc = create_impl(c);
    *out = c;   // returned even on failure so that the caller can read last_error, then destroy
    return rc;
}

extern "C" void ldso_gpu_destroy(ldso_gpu_ctx* c) {
    if (!c) return;
    int caller_dev = -1;
    cudaGetDevice(&caller_dev);              // the caller's current device is restored below
    cudaSetDevice(c->cfg.device);
    cudaDeviceSynchronize();
    if (c->frame) {
        for (int s = 0; s < c->cfg.n_frame_slots; s++) cudaFree(c->frame[s]);
        delete[] c->frame;
    }
    for (int t = 0; t < LDSO_NUM_TRACKERS; t++) {
        for (int l = 0; l < LDSO_MAX_LEVELS; l++) {
            cudaFree(c->trk[t].idepth[l]); cudaFree(c->trk[t].wsum[l]); cudaFree(c->trk[t].wsum_bak[l]); cudaFree(c->trk[t].pc[l]);
        }
        cudaFree(c->trk[t].row_cnt); cudaFree(c->trk[t].pc_n_dev);
        if (c->trk[t].pc_n_pinned) cudaFreeHost(c->trk[t].pc_n_pinned);
        if (c->trk[t].ev_pc_n) cudaEventDestroy(c->trk[t].ev_pc_n);
    }
    win_free(c);
    selector_free(c);
    immature_free(c);
    initializer_free(c);
    undistort_free(c);
    corners_free(c);
    cudaFree(c->img_dev); cudaFreeHost(c->img_pinned); cudaFreeHost(c->scratch_host); cudaFree(c->gammaB_dev);
    cudaFree(c->trk_partials); cudaFree(c->trk_partials_i); cudaFree(c->trk_counter); cudaFree(c->trk_result_dev);
    cudaFreeHost(c->trk_result_pinned);
    cudaFreeHost(c->trk_mailbox);
    if (c->ev0) cudaEventDestroy(c->ev0);
    if (c->ev1) cudaEventDestroy(c->ev1);
    if (c->ev_h2d) cudaEventDestroy(c->ev_h2d);
    if (c->ev_trk) cudaEventDestroy(c->ev_trk);
    prof_drain(c);
    for (cudaEvent_t e : c->prof_free) cudaEventDestroy(e);
    if (c->stream) cudaStreamDestroy(c->stream);
    if (c->stream_map) cudaStreamDestroy(c->stream_map);
    const int own_dev = c->cfg.device;
    delete c;
    if (caller_dev >= 0 && caller_dev != own_dev) cudaSetDevice(caller_dev);
}

extern "C" const char* ldso_gpu_last_error(const ldso_gpu_ctx* c) { return c ? c->err.c_str() : "null ctx"; }
extern "C" int ldso_gpu_pyr_levels(const ldso_gpu_ctx* c) { return c->levels; }
extern "C" int ldso_gpu_level_size(const ldso_gpu_ctx* c, int lvl, int* w, int* h) {
    if (lvl < 0 || lvl >= c->levels) return -2;
    *w = c->wG[lvl]; *h = c->hG[lvl];
    return 0;
}
extern "C" int ldso_gpu_sync(ldso_gpu_ctx* c) {
    LDSO_ENTER(c);
    LDSO_CUDA_OK(c, ldso_stream_sync(c->stream));
    LDSO_CUDA_OK(c, ldso_stream_sync(c->stream_map));
    return 0;
}
extern "C" int ldso_gpu_bind_thread(ldso_gpu_ctx* c) {
    LDSO_ENTER(c);
    LDSO_CUDA_OK(c, cudaSetDevice(c->cfg.device));
    return 0;
}
extern "C" long long ldso_gpu_launch_count(const ldso_gpu_ctx* c) { return c->launches; }

extern "C" int ldso_gpu_io_bytes(ldso_gpu_ctx* c, long long* h2d, long long* d2h) {
    LDSO_ENTER(c);
    *h2d = c->h2d_bytes; *d2h = c->d2h_bytes;
    return 0;
}
extern "C" int ldso_gpu_timer_start(ldso_gpu_ctx* c) {
    LDSO_ENTER(c);
    LDSO_CUDA_OK(c, cudaEventRecord(c->ev0, c->stream));
    return 0;
}
extern "C" int ldso_gpu_timer_stop(ldso_gpu_ctx* c, float* ms) {
    LDSO_ENTER(c);
    // the window closes over BOTH streams: work queued on the mapping stream (the asynchronous traceNewCoarse) is part of
    // the frame, so the tracking stream waits for it before the stop event is recorded
    LDSO_CUDA_OK(c, cudaEventRecord(c->ev_h2d, c->stream_map));
    LDSO_CUDA_OK(c, cudaStreamWaitEvent(c->stream, c->ev_h2d, 0));
    LDSO_CUDA_OK(c, cudaEventRecord(c->ev1, c->stream));
    LDSO_CUDA_OK(c, ldso_event_sync(c->ev1));
    LDSO_CUDA_OK(c, cudaEventElapsedTime(ms, c->ev0, c->ev1));
    return 0;
}

static cudaEvent_t prof_event(ldso_gpu_ctx* c) {
    if (!c->prof_free.empty()) { cudaEvent_t e = c->prof_free.back(); c->prof_free.pop_back(); return e; }
    cudaEvent_t e = nullptr;
    cudaEventCreate(&e);
    return e;
}
void prof_begin(ldso_gpu_ctx* c, cudaStream_t s) {
    if (!c->profile_on) return;
    c->prof_cur = prof_event(c);
    cudaEventRecord(c->prof_cur, s);
}
int prof_end(ldso_gpu_ctx* c, cudaStream_t s, int which, double units) {
    if (!c->profile_on || !c->prof_cur) return -1;
    cudaEvent_t e1 = prof_event(c);
    cudaEventRecord(e1, s);
    c->prof_pending.push_back({which, units, c->prof_cur, e1});
    c->prof_cur = nullptr;
    return (int)c->prof_pending.size() - 1;
}
void prof_drain(ldso_gpu_ctx* c) {
    for (auto& r : c->prof_pending) {
        float ms = 0;
        ldso_event_sync(r.e1);
        if (cudaEventElapsedTime(&ms, r.e0, r.e1) == cudaSuccess) {
            c->prof_ms[r.which] += ms; c->prof_launches[r.which]++; c->prof_units[r.which] += r.units;
        }
        c->prof_free.push_back(r.e0); c->prof_free.push_back(r.e1);
    }
    c->prof_pending.clear();
}
extern "C" int ldso_gpu_profile_enable(ldso_gpu_ctx* c, int on) {
    LDSO_ENTER(c);
    immature_trace_finalize(c);
    prof_drain(c);
    c->profile_on = on;
    if (on) for (int i = 0; i < LDSO_PROF_SLOTS; i++) { c->prof_ms[i] = 0; c->prof_launches[i] = 0; c->prof_units[i] = 0; }
    return 0;
}
extern "C" int ldso_gpu_profile_get(ldso_gpu_ctx* c, int which, double* total_ms, long long* launches, double* units) {
    LDSO_ENTER(c);
    LDSO_REQUIRE(c, which >= 0 && which < LDSO_PROF_SLOTS);
    immature_trace_finalize(c);
    prof_drain(c);
    *total_ms = c->prof_ms[which]; *launches = c->prof_launches[which]; *units = c->prof_units[which];
    if (c->cfg.reserved[0] == 3 && which == LDSO_PROF_K34) {
        fprintf(stderr, "lm phase clocks: eval %.0f barrier %.0f fold %.0f decide %.0f solve %.0f candidate %.0f\n", c->lm_dbg_clocks[0],
                c->lm_dbg_clocks[1], c->lm_dbg_clocks[2], c->lm_dbg_clocks[3], c->lm_dbg_clocks[4], c->lm_dbg_clocks[5]);
        for (int i = 0; i < 8; i++) c->lm_dbg_clocks[i] = 0;
    }
    return 0;
}

extern "C" int ldso_gpu_set_gamma(ldso_gpu_ctx* c, const float B[256]) {
    LDSO_ENTER(c);
    LDSO_CUDA_OK(c, ldso_copy(c, c->gammaB_dev, B, 256 * sizeof(float), cudaMemcpyHostToDevice, c->stream));
    LDSO_CUDA_OK(c, ldso_stream_sync(c->stream));
    return 0;
}

// ---- K1 ------------------------------------------------------------------------------------------
extern "C" int ldso_gpu_frame_make_images(ldso_gpu_ctx* c, int slot, const float* img) {
    LDSO_ENTER(c);
    LDSO_REQUIRE(c, slot >= 0 && slot < c->cfg.n_frame_slots && img);
    const size_t bytes = (size_t)c->cfg.w * c->cfg.h * sizeof(float);
    // the caller deletes its ImageAndExposure right after addActiveFrame (run_dso_tum_mono.cc:405-429):
    // consume the image synchronously into pinned staging, then go asynchronous.
    cudaPointerAttributes attr;
    if (cudaPointerGetAttributes(&attr, img) == cudaSuccess && attr.type == cudaMemoryTypeHost) {
        // the caller's image already lives in page-locked memory (cudaMallocHost / cudaHostRegister, e.g. the shim's
        // ImageAndExposure allocator): DMA straight out of it, queue the pyramid behind the copy, and give the buffer
        // back as soon as the copy — not the kernel — has finished
        LDSO_CUDA_OK(c, ldso_copy(c, c->img_dev, img, bytes, cudaMemcpyHostToDevice, c->stream));
        LDSO_CUDA_OK(c, cudaEventRecord(c->ev_h2d, c->stream));
        int rc = k1_make_images(c, slot, c->img_dev);
        LDSO_CUDA_OK(c, ldso_event_sync(c->ev_h2d));
        return rc;
    }
    (void)cudaGetLastError();                            // an unregistered pointer is reported as an error on some drivers
    LDSO_CUDA_OK(c, ldso_stream_sync(c->stream));   // previous H2D out of img_pinned has drained
    memcpy(c->img_pinned, img, bytes);
    LDSO_CUDA_OK(c, ldso_copy(c, c->img_dev, c->img_pinned, bytes, cudaMemcpyHostToDevice, c->stream));
    return k1_make_images(c, slot, c->img_dev);
}
// makeImages of a frame whose image stays valid until the ctx's tracking stream has been waited for (a batch that was handed
// all its frames up front, ldso_gpu_batch_run): the copy and the pyramid are queued and the call returns at once.  kind: an
// LDSO_IMAGE_* value (include/ldso_system.h).  A pageable host pointer cannot be copied asynchronously: it takes the waiting path.
extern "C" int ldso_gpu_frame_make_images_raw(ldso_gpu_ctx* c, int slot, const void* raw, int bytes_per_pixel, float exposure_time,
                                              float factor);
int n4_make_images_raw_async(ldso_gpu_ctx* c, int slot, const void* raw, int bytes_per_pixel, float exposure_time, float factor);
extern "C" int ldso_gpu_frame_make_images_async(ldso_gpu_ctx* c, int slot, const void* image, int kind, float exposure_time) {
    LDSO_ENTER(c);
    LDSO_REQUIRE(c, slot >= 0 && slot < c->cfg.n_frame_slots && image && kind >= 0 && kind <= 3);
    if (kind == 1) return k1_make_images(c, slot, (const float*)image);
    cudaPointerAttributes attr;
    const bool pinned = cudaPointerGetAttributes(&attr, image) == cudaSuccess && attr.type == cudaMemoryTypeHost;
    if (!pinned) {
        (void)cudaGetLastError();
        if (kind == 0) return ldso_gpu_frame_make_images(c, slot, (const float*)image);
        return ldso_gpu_frame_make_images_raw(c, slot, image, kind == 2 ? 1 : 2, exposure_time, 1.0f);
    }
    if (kind == 0) {
        const size_t bytes = (size_t)c->cfg.w * c->cfg.h * sizeof(float);
        LDSO_CUDA_OK(c, ldso_copy(c, c->img_dev, image, bytes, cudaMemcpyHostToDevice, c->stream));
        return k1_make_images(c, slot, c->img_dev);
    }
    return n4_make_images_raw_async(c, slot, image, kind == 2 ? 1 : 2, exposure_time, 1.0f);
}

extern "C" int ldso_gpu_frame_make_images_dev(ldso_gpu_ctx* c, int slot, const float* img_dev) {
    LDSO_ENTER(c);
    LDSO_REQUIRE(c, slot >= 0 && slot < c->cfg.n_frame_slots && img_dev);
    return k1_make_images(c, slot, img_dev);
}

extern "C" int ldso_gpu_frame_download(ldso_gpu_ctx* c, int slot, int lvl, int what, float* dst) {
    LDSO_ENTER(c);
    LDSO_REQUIRE(c, slot >= 0 && slot < c->cfg.n_frame_slots && lvl >= 0 && lvl < c->levels && dst);
    const size_t n = (size_t)c->wG[lvl] * c->hG[lvl];
    float4* h = (float4*)c->scratch_host;
    LDSO_CUDA_OK(c, ldso_copy(c, h, c->frame[slot] + c->lvl_off[lvl], n * sizeof(float4), cudaMemcpyDeviceToHost, c->stream));
    LDSO_CUDA_OK(c, ldso_stream_sync(c->stream));
    if (what == 0) {
        for (size_t i = 0; i < n; i++) { dst[3 * i] = h[i].x; dst[3 * i + 1] = h[i].y; dst[3 * i + 2] = h[i].z; }
    } else {
        for (size_t i = 0; i < n; i++) dst[i] = h[i].w;
    }
    return 0;
}

// ---- K2 ------------------------------------------------------------------------------------------
extern "C" int ldso_gpu_tracker_make_k(ldso_gpu_ctx* c, int tracker, float fx, float fy, float cx, float cy) {
    LDSO_ENTER(c);
    LDSO_REQUIRE(c, tracker >= 0 && tracker < LDSO_NUM_TRACKERS);
    TrackerDev& T = c->trk[tracker];
    // CoarseTracker::makeK (src/frontend/CoarseTracker.cc:312-339): float state, double intermediates
    T.w[0] = c->wG[0]; T.h[0] = c->hG[0];
    T.fx[0] = fx; T.fy[0] = fy; T.cx[0] = cx; T.cy[0] = cy;
    for (int l = 1; l < c->levels; l++) {
        T.w[l] = T.w[0] >> l;
        T.h[l] = T.h[0] >> l;
        T.fx[l] = (float)(T.fx[l - 1] * 0.5);
        T.fy[l] = (float)(T.fy[l - 1] * 0.5);
        T.cx[l] = (float)((T.cx[0] + 0.5) / ((int)1 << l) - 0.5);
        T.cy[l] = (float)((T.cy[0] + 0.5) / ((int)1 << l) - 0.5);
    }
    for (int l = 0; l < c->levels; l++) {
        float* K = T.K[l];
        K[0] = T.fx[l]; K[1] = 0; K[2] = T.cx[l];
        K[3] = 0; K[4] = T.fy[l]; K[5] = T.cy[l];
        K[6] = 0; K[7] = 0; K[8] = 1;
        ldso_sm::mat3_inv<float>(K, T.Ki[l]);
    }
    return 0;
}

extern "C" int ldso_gpu_tracker_set_ref(ldso_gpu_ctx* c, int tracker, int ref_slot, int n_points, const float* u, const float* v,
                                        const float* idepth, const float* hdi, float ref_exposure, float ref_aff_a,
                                        float ref_aff_b) {
    LDSO_ENTER(c);
    LDSO_REQUIRE(c, tracker >= 0 && tracker < LDSO_NUM_TRACKERS);
    LDSO_REQUIRE(c, ref_slot >= 0 && ref_slot < c->cfg.n_frame_slots);
    LDSO_REQUIRE(c, n_points >= 0 && (size_t)n_points * 4 <= (size_t)c->cfg.w * c->cfg.h);
    LDSO_REQUIRE(c, n_points == 0 || (u && v && idepth && hdi));
    TrackerDev& T = c->trk[tracker];
    T.ref_exposure = ref_exposure; T.ref_aff_a = ref_aff_a; T.ref_aff_b = ref_aff_b;
    return k2_set_ref(c, tracker, ref_slot, n_points, u, v, idepth, hdi);
}

extern "C" int ldso_gpu_tracker_get_pc(ldso_gpu_ctx* c, int tracker, int lvl, int* n, float* u, float* v, float* idepth,
                                       float* color) {
    LDSO_ENTER(c);
    LDSO_REQUIRE(c, tracker >= 0 && tracker < LDSO_NUM_TRACKERS && lvl >= 0 && lvl < c->levels && n);
    TrackerDev& T = c->trk[tracker];
    if (trk_pc_n_sync(c, T)) return -1;
    int m = T.pc_n[lvl];
    *n = m;
    if (m == 0 || !(u || v || idepth || color)) return 0;
    float4* h = (float4*)c->scratch_host;
    LDSO_CUDA_OK(c, ldso_copy(c, h, T.pc[lvl], (size_t)m * sizeof(float4), cudaMemcpyDeviceToHost, c->stream));
    LDSO_CUDA_OK(c, ldso_stream_sync(c->stream));
    for (int i = 0; i < m; i++) {
        if (u) u[i] = h[i].x;
        if (v) v[i] = h[i].y;
        if (idepth) idepth[i] = h[i].z;
        if (color) color[i] = h[i].w;
    }
    return 0;
}

// ---- K3 / K4 -------------------------------------------------------------------------------------
// CoarseTracker.cc:664-672
static void fill_vec6(const TrackEvalOut& e, double rs[6]) {
    rs[0] = (float)e.sums[45];
    rs[1] = e.counts[0];
    float num = (float)e.counts[3];
    rs[2] = (float)e.sums[46] / (num + 0.1);
    rs[3] = 0;
    rs[4] = (float)e.sums[47] / (num + 0.1);
    rs[5] = e.counts[1] / (float)e.counts[0];
}
// CoarseTracker.cc:725-740.  n is the warped count padded to a multiple of 4 (trap T7).
static void fill_Hb(const TrackEvalOut& e, double H[64], double b[8]) {
    int n = (e.counts[2] + 3) & ~3;
    double invn = (double)(1.0f / n);
    static const double sc[8] = {LDSO_SCALE_XI_TRANS, LDSO_SCALE_XI_TRANS, LDSO_SCALE_XI_TRANS, LDSO_SCALE_XI_ROT,
                                 LDSO_SCALE_XI_ROT,   LDSO_SCALE_XI_ROT,   LDSO_SCALE_A,        LDSO_SCALE_B};
    int k = 0;
    for (int r = 0; r < 9; r++)
        for (int cc = r; cc < 9; cc++, k++) {
            double v = e.sums[k] * invn;
            if (r < 8 && cc < 8) H[r * 8 + cc] = H[cc * 8 + r] = v * sc[r] * sc[cc];
            else if (r < 8) b[r] = v * sc[r];
        }
}

extern "C" int ldso_gpu_tracker_calc_res(ldso_gpu_ctx* c, int tracker, int new_slot, float new_exposure, int lvl,
                                         const double ref_to_new[12], float aff_a, float aff_b, float cutoff_th, float out6[6]) {
    LDSO_ENTER(c);
    LDSO_REQUIRE(c, tracker >= 0 && tracker < LDSO_NUM_TRACKERS && lvl >= 0 && lvl < c->levels);
    LDSO_REQUIRE(c, new_slot >= 0 && new_slot < c->cfg.n_frame_slots && ref_to_new && out6);
    TrackerDev& T = c->trk[tracker];
    LDSO_REQUIRE(c, T.ref_slot >= 0);
    T.new_slot = new_slot; T.new_exposure = new_exposure;
    TrackEvalOut e;
    int rc = k34_eval(c, tracker, lvl, SE3::fromRt(ref_to_new), aff_a, aff_b, cutoff_th, &e);
    if (rc) return rc;
    memcpy(T.last_sums, e.sums, sizeof(e.sums));
    memcpy(T.last_counts, e.counts, sizeof(e.counts));
    T.last_lvl = lvl; T.last_aff_a = aff_a; T.last_aff_b = aff_b;
    double rs[6];
    fill_vec6(e, rs);
    for (int i = 0; i < 6; i++) out6[i] = (float)rs[i];
    return 0;
}

extern "C" int ldso_gpu_tracker_calc_gs(ldso_gpu_ctx* c, int tracker, int lvl, const double ref_to_new[12], float aff_a,
                                        float aff_b, double H[64], double b[8]) {
    LDSO_ENTER(c);
    LDSO_REQUIRE(c, tracker >= 0 && tracker < LDSO_NUM_TRACKERS && H && b);
    TrackerDev& T = c->trk[tracker];
    // the reference accumulates over the buffers left by the LAST calcRes; the fused launch already did
    LDSO_REQUIRE(c, T.last_lvl == lvl && T.last_aff_a == aff_a && T.last_aff_b == aff_b);
    (void)ref_to_new;
    TrackEvalOut e;
    memcpy(e.sums, T.last_sums, sizeof(e.sums));
    memcpy(e.counts, T.last_counts, sizeof(e.counts));
    fill_Hb(e, H, b);
    return 0;
}

// ---- CoarseTracker::trackNewestCoarse (src/frontend/CoarseTracker.cc:61-310), VI terms identically 0 ----
namespace {

// LM step on the 8-dof system: Hl = H with diag*(1+lambda), inc = Hl^-1 (-b)  (CoarseTracker.cc:130-184)
void solve_step(const double* H, const double* b, float lambda, int mode, float optA, float optB, double* inc) {
    double Hl[64], rhs[8];
    for (int i = 0; i < 8; i++) {
        for (int j = 0; j < 8; j++) Hl[i * 8 + j] = (mode == 1 && j < i) ? 0.0 : H[i * 8 + j];   // trap T4
        Hl[i * 8 + i] *= (1 + lambda);
        rhs[i] = -b[i];
    }
    auto lead = [&](int k, const double* M, const double* r, double* x) {
        double S[64];
        for (int i = 0; i < k; i++)
            for (int j = 0; j < k; j++) S[i * k + j] = M[i * 8 + j];
        ldso_sm::ldlt_solve(k, S, r, x);
    };
    const bool fixA = optA < 0, fixB = optB < 0;
    lead(8, Hl, rhs, inc);
    if (fixA && fixB) { lead(6, Hl, rhs, inc); inc[6] = inc[7] = 0; }
    else if (!fixA && fixB) { lead(7, Hl, rhs, inc); inc[7] = 0; }
    else if (fixA && !fixB) {
        double Hs[64], rs[8], xs[8];
        memcpy(Hs, Hl, sizeof(Hs));
        for (int i = 0; i < 8; i++) Hs[i * 8 + 6] = Hs[i * 8 + 7];
        for (int j = 0; j < 8; j++) Hs[6 * 8 + j] = Hs[7 * 8 + j];
        for (int i = 0; i < 8; i++) rs[i] = -b[i];
        rs[6] = -b[7];
        lead(7, Hs, rs, xs);
        for (int i = 0; i < 6; i++) inc[i] = xs[i];
        inc[6] = 0; inc[7] = xs[6];
    }
}

}  // namespace

// CoarseTracker::trackNewestCoarse in two halves (phase 1 = begin: launch without waiting, 2 = end: wait and finish,
// 0 = both).  With the host-driven LM loop (tracker_lm_mode = 1) nothing can be launched ahead: begin only keeps its arguments.
static int track_impl(ldso_gpu_ctx* c, int phase, int tracker, int new_slot, float new_exposure, double pose_io[12], float aff_io[2],
                      int coarsest_lvl, const float min_res_for_abort[5], float last_residuals[5], float flow[3], int iters_out[2]) {
    LDSO_REQUIRE(c, tracker >= 0 && tracker < LDSO_NUM_TRACKERS);
    LDSO_REQUIRE(c, new_slot >= 0 && new_slot < c->cfg.n_frame_slots && pose_io && aff_io && min_res_for_abort);
    LDSO_REQUIRE(c, coarsest_lvl >= 0 && coarsest_lvl < 5 && coarsest_lvl < c->levels);   // CoarseTracker.cc:65
    TrackerDev& T = c->trk[tracker];
    LDSO_REQUIRE(c, T.ref_slot >= 0);
    static const int maxIterations[5] = {10, 20, 50, 50, 50};
    const float extrapLimit = 0.001f;
    const float optA = c->cfg.affine_opt_mode_a, optB = c->cfg.affine_opt_mode_b;
    SE3 cur = SE3::fromRt(pose_io);
    float cur_a = aff_io[0], cur_b = aff_io[1];
    const bool device_lm = c->cfg.tracker_lm_mode != 1;
    if (phase != 2) {
        T.new_slot = new_slot; T.new_exposure = new_exposure;
        T.last_lvl = -1;
        if (device_lm) {
            int rc = k34_track_all_launch(c, tracker, coarsest_lvl, maxIterations, min_res_for_abort, cur, cur_a, cur_b);
            if (rc) return rc;
        }
        if (phase == 1) return 0;
    }
    for (int i = 0; i < 5; i++) last_residuals[i] = NAN;
    for (int i = 0; i < 3; i++) flow[i] = 1000;
    bool repeated = false;
    int nEval = 0, nAccept = 0, rc;
    TrackEvalOut ev;
    if (device_lm) {
        // the level loop with its LM iterations runs inside one kernel (thread-block cluster, or cooperative grid in mode 2)
        int status = 0;
        if ((rc = k34_track_all_wait(c, cur, cur_a, cur_b, last_residuals, flow, &status, &nEval, &nAccept))) return rc;
        if (status) {
            if (iters_out) { iters_out[0] = nEval; iters_out[1] = nAccept; }
            return 1;
        }
    }
    for (int lvl = coarsest_lvl; lvl >= 0 && !device_lm; lvl--) {
        double H[64], b[8], resOld[6], resNew[6];
        float cutoffRepeat = 1;
        {
            if ((rc = k34_eval(c, tracker, lvl, cur, cur_a, cur_b, c->cfg.coarse_cutoff_th * cutoffRepeat, &ev))) return rc;
            nEval++;
            fill_vec6(ev, resOld);
            while (resOld[5] > 0.6 && cutoffRepeat < 50) {
                cutoffRepeat *= 2;
                if ((rc = k34_eval(c, tracker, lvl, cur, cur_a, cur_b, c->cfg.coarse_cutoff_th * cutoffRepeat, &ev))) return rc;
                nEval++;
                fill_vec6(ev, resOld);
            }
            fill_Hb(ev, H, b);
            float lambda = 0.01f;
            for (int it = 0; it < maxIterations[lvl]; it++) {
                double inc[8], incS[8];
                solve_step(H, b, lambda, c->cfg.tracker_solve_mode, optA, optB, inc);
                float extrap = 1;
                if (lambda < extrapLimit) extrap = sqrtf(sqrt(extrapLimit / lambda));
                double sq = 0;
                for (int i = 0; i < 8; i++) {
                    inc[i] *= extrap;
                    incS[i] = inc[i] * (i < 3 ? LDSO_SCALE_XI_TRANS : i < 6 ? LDSO_SCALE_XI_ROT : i == 6 ? LDSO_SCALE_A : LDSO_SCALE_B);
                    sq += incS[i] * incS[i];
                }
                if (!std::isfinite(sq))
                    for (int i = 0; i < 8; i++) inc[i] = incS[i] = 0;
                SE3 cand = SE3::exp(incS) * cur;
                float cand_a = cur_a, cand_b = cur_b;
                cand_a += incS[6];
                cand_b += incS[7];
                if ((rc = k34_eval(c, tracker, lvl, cand, cand_a, cand_b, c->cfg.coarse_cutoff_th * cutoffRepeat, &ev))) return rc;
                nEval++;
                fill_vec6(ev, resNew);
                if ((resNew[0] / resNew[1]) < (resOld[0] / resOld[1])) {
                    fill_Hb(ev, H, b);   // the fused launch already accumulated the system at the accepted pose
                    memcpy(resOld, resNew, sizeof(resOld));
                    cur = cand; cur_a = cand_a; cur_b = cand_b;
                    lambda *= 0.5;
                    nAccept++;
                } else {
                    lambda *= 4.0;
                    if (lambda < extrapLimit) lambda = extrapLimit;
                }
                double nrm = 0;
                for (int i = 0; i < 8; i++) nrm += inc[i] * inc[i];
                if (!(std::sqrt(nrm) > 1e-3)) break;
            }
        }
        last_residuals[lvl] = sqrtf((float)(resOld[0] / resOld[1]));
        flow[0] = (float)resOld[2]; flow[1] = (float)resOld[3]; flow[2] = (float)resOld[4];
        if (last_residuals[lvl] > 1.5 * min_res_for_abort[lvl]) {
            if (iters_out) { iters_out[0] = nEval; iters_out[1] = nAccept; }
            return 1;
        }
        if (cutoffRepeat > 1 && !repeated) { lvl++; repeated = true; }
    }
    if (iters_out) { iters_out[0] = nEval; iters_out[1] = nAccept; }
    if ((optA != 0 && fabsf(cur_a) > 1.2) || (optB != 0 && fabsf(cur_b) > 200)) return 1;
    float relA, relB;
    {
        float eF = T.ref_exposure, eT = new_exposure;
        if (eF == 0 || eT == 0) eT = eF = 1;
        relA = exp(cur_a - T.ref_aff_a) * eT / eF;   // include/AffLight.h:27-35
        relB = cur_b - relA * T.ref_aff_b;
    }
    if ((optA == 0 && fabsf(logf(relA)) > 1.5) || (optB == 0 && fabsf(relB) > 200)) return 1;
    if (optA < 0) cur_a = 0;
    if (optB < 0) cur_b = 0;
    cur.toRt(pose_io);
    aff_io[0] = cur_a; aff_io[1] = cur_b;
    return 0;
}

extern "C" int ldso_gpu_tracker_track(ldso_gpu_ctx* c, int tracker, int new_slot, float new_exposure, double pose_io[12],
                                      float aff_io[2], int coarsest_lvl, const float min_res_for_abort[5],
                                      float last_residuals[5], float flow[3], int iters_out[2]) {
    LDSO_ENTER(c);
    LDSO_REQUIRE(c, last_residuals && flow);
    LDSO_REQUIRE(c, !c->trk_pending.active);
    return track_impl(c, 0, tracker, new_slot, new_exposure, pose_io, aff_io, coarsest_lvl, min_res_for_abort, last_residuals, flow, iters_out);
}

extern "C" int ldso_gpu_tracker_track_begin(ldso_gpu_ctx* c, int tracker, int new_slot, float new_exposure, const double pose[12],
                                            const float aff[2], int coarsest_lvl, const float min_res_for_abort[5]) {
    LDSO_ENTER(c);
    LDSO_REQUIRE(c, pose && aff && min_res_for_abort);
    LDSO_REQUIRE(c, !c->trk_pending.active);
    auto& P = c->trk_pending;
    P.tracker = tracker; P.new_slot = new_slot; P.new_exposure = new_exposure; P.coarsest = coarsest_lvl;
    for (int i = 0; i < 12; i++) P.pose[i] = pose[i];
    for (int i = 0; i < 2; i++) P.aff[i] = aff[i];
    for (int i = 0; i < 5; i++) P.min_res[i] = min_res_for_abort[i];
    const int rc = track_impl(c, 1, tracker, new_slot, new_exposure, P.pose, P.aff, coarsest_lvl, P.min_res, nullptr, nullptr, nullptr);
    if (rc == 0) P.active = true;
    return rc;
}

extern "C" int ldso_gpu_tracker_track_end(ldso_gpu_ctx* c, double pose_out[12], float aff_out[2], float last_residuals[5], float flow[3],
                                          int iters_out[2]) {
    LDSO_ENTER(c);
    LDSO_REQUIRE(c, pose_out && aff_out && last_residuals && flow);
    auto& P = c->trk_pending;
    LDSO_REQUIRE(c, P.active);
    P.active = false;
    const int rc = track_impl(c, 2, P.tracker, P.new_slot, P.new_exposure, P.pose, P.aff, P.coarsest, P.min_res, last_residuals, flow, iters_out);
    for (int i = 0; i < 12; i++) pose_out[i] = P.pose[i];
    for (int i = 0; i < 2; i++) aff_out[i] = P.aff[i];
    return rc;
}
