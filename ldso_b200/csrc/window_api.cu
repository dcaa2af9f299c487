// window_api.cu — C-ABI entry points of the windowed bundle adjustment (ldso_gpu_window_*): window upload,
// launch of K7-K10, and the small dense fp64 host algebra (H1) the reference keeps on the host too:
// stitching the accumulators with the adjoints (AccumulatedTopHessian.cc:193-255,
// AccumulatedSCHessian.cc:53-119) and the dense part of EnergyFunctional::solveSystemF (:310-432).
#include "ctx.h"

#include <cstdlib>

#include <algorithm>
#include <cmath>

#define CPARS 4

void* ldso_staging(ldso_gpu_ctx* c, size_t bytes) {
    WindowDev& W = c->win;
    const size_t need = (bytes + 255) & ~(size_t)255;
    if (W.pinned_off + need > W.pinned_bytes) {
        if (W.pinned_off > 0) { ldso_stream_sync(c->stream); W.pinned_off = 0; }
        if (need > W.pinned_bytes) {
            if (W.pinned) cudaFreeHost(W.pinned);
            W.pinned = nullptr;
            size_t nb = std::max(need, (size_t)4 << 20);
            if (cudaMallocHost(&W.pinned, nb) != cudaSuccess) { W.pinned_bytes = 0; return nullptr; }
            W.pinned_bytes = nb;
        }
    }
    void* p = (char*)W.pinned + W.pinned_off;
    W.pinned_off += need;
    return p;
}


namespace {

template <class T> int dalloc(ldso_gpu_ctx* c, T** p, size_t n) {
    LDSO_CUDA_OK(c, cudaMalloc((void**)p, n * sizeof(T)));
    return 0;
}

// Pinned staging ring.  Uploads are queued: the host buffer is copied into the ring at the cursor and the H2D copy
// is issued without waiting (the caller may reuse its buffer at once; kernels on the same stream are ordered behind
// the copy).  The ring is only rewound after a stream synchronisation, so a window upload costs one sync at most.
inline void* staging(ldso_gpu_ctx* c, size_t bytes) { return ldso_staging(c, bytes); }

int upload(ldso_gpu_ctx* c, void* dst, const void* src, size_t bytes) {
    if (bytes == 0) return 0;
    void* st = staging(c, bytes);
    if (!st) { c->set_error("pinned staging allocation failed", __FILE__, __LINE__); return -1; }
    memcpy(st, src, bytes);
    LDSO_CUDA_OK(c, ldso_copy(c, dst, st, bytes, cudaMemcpyHostToDevice, c->stream));
    return 0;
}

int download(ldso_gpu_ctx* c, void* dst, const void* src, size_t bytes) {
    if (bytes == 0) return 0;
    void* st = staging(c, bytes);
    if (!st) { c->set_error("pinned staging allocation failed", __FILE__, __LINE__); return -1; }
    LDSO_CUDA_OK(c, ldso_copy(c, st, src, bytes, cudaMemcpyDeviceToHost, c->stream));
    LDSO_CUDA_OK(c, ldso_stream_sync(c->stream));
    memcpy(dst, st, bytes);
    c->win.pinned_off = 0;   // everything queued before has completed
    return 0;
}

}  // namespace

int win_alloc(ldso_gpu_ctx* c) {
    WindowDev& W = c->win;
    W.cap_p = std::max(c->cfg.max_points, 16);
    W.cap_r = std::max(c->cfg.max_residuals, 16);
    const size_t P = W.cap_p, R = W.cap_r, M = LDSO_MAX_WINDOW;
    int rc = 0;
    rc |= dalloc(c, &W.precalc, M * M * LDSO_PRECALC_STRIDE);
    rc |= dalloc(c, &W.pool_p, (P + 8) * 24);
    rc |= dalloc(c, &W.pt_out, (WIN_PT_ROWS + 2) * P);          // + max_rel_baseline, num_good: one memset per upload
    W.max_rel_baseline = W.pt_out + (size_t)WIN_PT_ROWS * P; W.num_good = W.max_rel_baseline + P;
    rc |= dalloc(c, &W.pool_ri, (R + 8) * 5);
    rc |= dalloc(c, &W.pool_rz, (R + 8) * 103);
    rc |= dalloc(c, &W.order, R); rc |= dalloc(c, &W.chunk, 3 * (R / 1 + 2 * M * M + 8));
    rc |= dalloc(c, &W.sc_blocks, 3 * (P + M + 8));
    rc |= dalloc(c, &W.ridx, P * M); rc |= dalloc(c, &W.pmask, P);
    rc |= dalloc(c, &W.adHTdelta, M * M * 8); rc |= dalloc(c, &W.xAd, M * M * 8 + 8); rc |= dalloc(c, &W.stats, 4);
    rc |= dalloc(c, &W.acc_top, 2 * M * M * WIN_TOP_STRIDE); rc |= dalloc(c, &W.acc_top_cnt, 2 * M * M);
    W.acc_sc_size = M * M * M * 64 + M * M * 40 + 20;
    rc |= dalloc(c, &W.acc_sc, W.acc_sc_size);
    rc |= dalloc(c, &W.adH_dev, M * M * 64); rc |= dalloc(c, &W.adT_dev, M * M * 64); rc |= dalloc(c, &W.prior_dev, 8 + 16 * M);
    rc |= dalloc(c, &W.stitched, 3 * ((8 * M + 4) * (8 * M + 4) + (8 * M + 4)));
    return rc ? -1 : 0;
}

void win_free(ldso_gpu_ctx* c) {
    WindowDev& W = c->win;
    void* ptrs[] = {W.precalc, W.pool_p, W.pt_out, W.pool_ri, W.pool_rz, W.order, W.chunk, W.sc_blocks, W.ridx, W.pmask, W.adHTdelta, W.xAd,
                    W.stats, W.acc_top, W.acc_top_cnt, W.acc_sc, W.adH_dev, W.adT_dev, W.prior_dev, W.stitched};
    for (void* p : ptrs) cudaFree(p);
    if (W.pinned) cudaFreeHost(W.pinned);
}

// The per-upload arrays are carved from their pools with strides rounded to the upload's size, in the order the host stages
// them, so that a window upload is ONE host-to-device copy (+ one memset) per group instead of one per array.
static size_t carve_points(WindowDev& W, int np) {
    const size_t q = ((size_t)np + 4) & ~(size_t)3;             // >= np + 1 (res_begin), multiple of 4 floats (16 B)
    float* b = W.pool_p;
    W.u = b; W.v = b + q; W.idepth = b + 2 * q; W.idepth_zero = b + 3 * q; W.prior = b + 4 * q;
    W.host = (int*)(b + 5 * q); W.res_begin = (int*)(b + 6 * q); W.color = b + 7 * q; W.weights = b + 15 * q;
    return q;
}
static size_t carve_residuals(WindowDev& W, int nr) {
    const size_t r = ((size_t)nr + 7) & ~(size_t)7;             // multiple of 8: the 288-byte J records stay 32-byte aligned
    int* bi = W.pool_ri;
    W.r_target = bi; W.r_point = bi + r; W.state = bi + 2 * r; W.flags = bi + 3 * r; W.new_state = bi + 4 * r;
    float* z = W.pool_rz;
    W.energy = z; W.new_energy = z + r; W.new_energy_wo = z + 2 * r; W.rtz = z + 3 * r; W.JpJdF = z + 11 * r; W.center = z + 19 * r;
    W.res_pt = z + 23 * r; W.J = z + 31 * r;
    return r;
}

// ---- window description --------------------------------------------------------------------------
extern "C" int ldso_gpu_window_set_frames(ldso_gpu_ctx* c, int n, const int* slots, const float* precalc, const float* eth,
                                          const float calib[4]) {
    LDSO_ENTER(c);
    LDSO_REQUIRE(c, n >= 1 && n <= LDSO_MAX_WINDOW && slots && precalc && eth && calib);
    WindowDev& W = c->win;
    for (int i = 0; i < n; i++) {
        LDSO_REQUIRE(c, slots[i] >= 0 && slots[i] < c->cfg.n_frame_slots);
        W.slots[i] = slots[i];
        W.frameEnergyTH[i] = eth[i];
    }
    W.n = n;
    W.fxl = calib[0]; W.fyl = calib[1]; W.cxl = calib[2]; W.cyl = calib[3];
    W.fxli = 1.0f / W.fxl; W.fyli = 1.0f / W.fyl;   // CalibHessian::value_scaledi (include/internal/CalibHessian.h:93-99)
    return upload(c, W.precalc, precalc, sizeof(float) * n * n * LDSO_PRECALC_STRIDE);
}

extern "C" int ldso_gpu_window_set_points(ldso_gpu_ctx* c, int np, const float* u, const float* v, const float* idepth,
                                          const float* idepth_zero, const int* host, const float* color, const float* weights,
                                          const float* prior, const int* res_begin) {
    LDSO_ENTER(c);
    WindowDev& W = c->win;
    LDSO_REQUIRE(c, np >= 0 && np <= W.cap_p && W.n > 0);
    LDSO_REQUIRE(c, np == 0 || (u && v && idepth && idepth_zero && host && color && weights && prior && res_begin));
    for (int i = 0; i < np; i++) LDSO_REQUIRE(c, host[i] >= 0 && host[i] < W.n && (i == 0 || host[i] >= host[i - 1]));
    W.np = np;
    W.h_host.assign(host, host + np);
    W.h_res_begin.assign(res_begin, res_begin + np + 1);
    {
        const size_t q = carve_points(W, np);
        const size_t bytes = sizeof(float) * 23 * q;
        float* st = (float*)staging(c, bytes);
        LDSO_REQUIRE(c, st != nullptr);
        memcpy(st, u, sizeof(float) * np); memcpy(st + q, v, sizeof(float) * np); memcpy(st + 2 * q, idepth, sizeof(float) * np);
        memcpy(st + 3 * q, idepth_zero, sizeof(float) * np); memcpy(st + 4 * q, prior, sizeof(float) * np);
        memcpy(st + 5 * q, host, sizeof(int) * np); memcpy(st + 6 * q, res_begin, sizeof(int) * (np + 1));
        memcpy(st + 7 * q, color, sizeof(float) * 8 * np); memcpy(st + 15 * q, weights, sizeof(float) * 8 * np);
        LDSO_CUDA_OK(c, ldso_copy(c, W.pool_p, st, bytes, cudaMemcpyHostToDevice, c->stream));
    }
    LDSO_CUDA_OK(c, cudaMemsetAsync(W.pt_out, 0, sizeof(float) * (WIN_PT_ROWS + 2) * W.cap_p, c->stream));
    // K9 CTA table: consecutive points of one host per CTA — 64 (one staging tile) for a single window, more tiles per CTA
    // once that still leaves two CTAs per SM (batched windows): the cross-group fold and the ~3 k fp64 atomics of the
    // flush are paid per CTA (64x window on B200: 150 us at 64 points per CTA, 95 us at 448)
    int per = 64;
    const int slots = 2 * c->sm_count - LDSO_MAX_WINDOW;                    // every host frame may add one short block
    if (np > 64 * slots) per = std::min(1024, ((np + slots - 1) / slots + 63) / 64 * 64);   // one wave: no tail
    if (const char* e = getenv("LDSO_K9_BLOCK")) per = atoi(e);            // tuning knob of profiles/kernel_bench.py
    std::vector<int> blk;
    int p = 0;
    while (p < np) {
        int h = host[p], e = p;
        while (e < np && host[e] == h && e - p < per) e++;
        blk.push_back(h); blk.push_back(p); blk.push_back(e);
        p = e;
    }
    W.n_sc_blocks = (int)blk.size() / 3;
    return upload(c, W.sc_blocks, blk.data(), sizeof(int) * blk.size());
}

// (re)build the K8 work list: residual ids sorted by key = mode * n^2 + host + n * target, cut into chunks of <= 128 (one window) to <= 1024 (batched) ids
// With `only` (a point mask) the list holds just those points' residuals, all under the mode-0 keys (addPoint<2>).
static int build_order(ldso_gpu_ctx* c, const unsigned char* only = nullptr) {
    WindowDev& W = c->win;
    const int nn = W.n * W.n, nr = W.nr;
    std::vector<int> cnt(2 * nn + 1, 0), key(nr);
    for (int r = 0; r < nr; r++) {
        if (only && !only[W.h_point[r]]) { key[r] = -1; continue; }
        key[r] = (!only && W.h_lin[r] ? nn : 0) + W.h_host[W.h_point[r]] + W.n * W.h_target[r];
        cnt[key[r] + 1]++;
    }
    for (int k = 0; k < 2 * nn; k++) cnt[k + 1] += cnt[k];
    std::vector<int> order(std::max(nr, 1)), pos(cnt.begin(), cnt.end() - 1), chunk;
    for (int r = 0; r < nr; r++) if (key[r] >= 0) order[pos[key[r]]++] = r;
    // chunk length: 128 (thread per residual) while the list is a wave or two of CTAs (one window: ~100 chunks); beyond that
    // k8_top_stream_kernel with chunks of 384..1024 ids, sized to keep the list ~8 chunks per SM deep (measured on B200:
    // 96 k residuals 22.9 us at 128 vs 16.5 us at 384; 769 k residuals flat from 384 to 1024)
    const int total = cnt[2 * nn];
    int clen = 128;
    if (total > 128 * 2 * c->sm_count) clen = std::min(1024, std::max(384, total / (8 * c->sm_count) / 128 * 128));
    if (const char* e = getenv("LDSO_K8_CHUNK")) clen = atoi(e);           // tuning knob of profiles/kernel_bench.py
    W.chunk_len = clen;
    for (int k = 0; k < 2 * nn; k++)
        for (int s = cnt[k]; s < cnt[k + 1]; s += clen) {
            chunk.push_back(k); chunk.push_back(s); chunk.push_back(std::min(clen, cnt[k + 1] - s));
        }
    W.n_chunks = (int)chunk.size() / 3;
    int rc = upload(c, W.order, order.data(), sizeof(int) * cnt[2 * nn]);
    rc |= upload(c, W.chunk, chunk.data(), sizeof(int) * chunk.size());
    return rc ? -1 : 0;
}

extern "C" int ldso_gpu_window_set_residuals(ldso_gpu_ctx* c, int nr, const int* target, const int* state,
                                             const unsigned char* flags, const float* res_to_zero) {
    LDSO_ENTER(c);
    WindowDev& W = c->win;
    LDSO_REQUIRE(c, nr >= 0 && nr <= W.cap_r && (int)W.h_res_begin.size() == W.np + 1 && W.h_res_begin[W.np] == nr);
    LDSO_REQUIRE(c, nr == 0 || (target && state && flags));
    W.nr = nr;
    W.h_target.assign(target, target + nr);
    W.h_point.resize(nr);
    W.h_lin.resize(nr);
    std::vector<int> fl(nr);
    for (int p = 0; p < W.np; p++)
        for (int r = W.h_res_begin[p]; r < W.h_res_begin[p + 1]; r++) W.h_point[r] = p;
    for (int r = 0; r < nr; r++) {
        LDSO_REQUIRE(c, target[r] >= 0 && target[r] < W.n);
        fl[r] = flags[r];
        W.h_lin[r] = (flags[r] & LDSO_RF_LINEARIZED) ? 1 : 0;
    }
    int rc = 0;
    {
        const size_t r = carve_residuals(W, nr);
        const size_t bytes = sizeof(int) * 5 * r;
        int* st = (int*)staging(c, std::max<size_t>(bytes, 16));
        LDSO_REQUIRE(c, st != nullptr);
        if (nr) {
            memcpy(st, target, sizeof(int) * nr); memcpy(st + r, W.h_point.data(), sizeof(int) * nr); memcpy(st + 2 * r, state, sizeof(int) * nr);
            memcpy(st + 3 * r, fl.data(), sizeof(int) * nr);
            for (int i = 0; i < nr; i++) st[4 * r + i] = LDSO_RES_OUTLIER;
            LDSO_CUDA_OK(c, ldso_copy(c, W.pool_ri, st, bytes, cudaMemcpyHostToDevice, c->stream));
            LDSO_CUDA_OK(c, cudaMemsetAsync(W.pool_rz, 0, sizeof(float) * 103 * r, c->stream));
        }
    }
    if (res_to_zero) rc |= upload(c, W.rtz, res_to_zero, sizeof(float) * 8 * (size_t)nr);
    if (rc) return -1;
    return build_order(c);
}

extern "C" int ldso_gpu_window_set_adjoints(ldso_gpu_ctx* c, const double* adH, const double* adT, const float* adHTdelta,
                                            const float cDelta[4], const double cPrior[4], const double* fprior,
                                            const double* fdelta_prior) {
    LDSO_ENTER(c);
    WindowDev& W = c->win;
    LDSO_REQUIRE(c, W.n > 0 && adH && adT && adHTdelta && cDelta && cPrior && fprior && fdelta_prior);
    const int nn = W.n * W.n;
    // the adjoints change only with the linearisation point (setAdjointsF): inside optimize's iterations the same matrices come
    // again with new deltas, and their 2 x 25 KB are not sent a second time
    const bool same_ad = (int)W.adHost.size() == 64 * nn && (int)W.adTarget.size() == 64 * nn && W.ad_dev_valid &&
                         memcmp(W.adHost.data(), adH, sizeof(double) * 64 * nn) == 0 && memcmp(W.adTarget.data(), adT, sizeof(double) * 64 * nn) == 0;
    if (!same_ad) {
        W.adHost.assign(adH, adH + 64 * nn); W.adTarget.assign(adT, adT + 64 * nn);
        W.adHostF.resize(64 * nn); W.adTargetF.resize(64 * nn);
        for (int i = 0; i < 64 * nn; i++) { W.adHostF[i] = (float)adH[i]; W.adTargetF[i] = (float)adT[i]; }
    }
    for (int i = 0; i < 4; i++) { W.cDeltaF[i] = cDelta[i]; W.cPrior[i] = cPrior[i]; }
    W.fprior.assign(fprior, fprior + 8 * W.n); W.fdelta_prior.assign(fdelta_prior, fdelta_prior + 8 * W.n);
    std::vector<double> pr(8 + 16 * (size_t)W.n);
    for (int i = 0; i < 4; i++) { pr[i] = cPrior[i]; pr[4 + i] = (double)cDelta[i]; }
    for (int i = 0; i < 8 * W.n; i++) { pr[8 + i] = fprior[i]; pr[8 + 8 * W.n + i] = fdelta_prior[i]; }
    int rc = 0;
    if (!same_ad) {
        rc |= upload(c, W.adH_dev, adH, sizeof(double) * 64 * nn);
        rc |= upload(c, W.adT_dev, adT, sizeof(double) * 64 * nn);
        W.ad_dev_valid = rc == 0;
    }
    rc |= upload(c, W.prior_dev, pr.data(), sizeof(double) * pr.size());
    rc |= upload(c, W.adHTdelta, adHTdelta, sizeof(float) * 8 * nn);
    return rc ? -1 : 0;
}

extern "C" int ldso_gpu_window_set_idepth(ldso_gpu_ctx* c, const float* idepth, const float* idepth_zero) {
    LDSO_ENTER(c);
    WindowDev& W = c->win;
    LDSO_REQUIRE(c, idepth && idepth_zero);
    if (W.np == 0) return 0;
    const size_t q = (size_t)(W.idepth_zero - W.idepth);        // the two arrays are neighbours in the point pool (carve_points)
    float* st = (float*)staging(c, sizeof(float) * 2 * q);
    LDSO_REQUIRE(c, st != nullptr);
    memcpy(st, idepth, sizeof(float) * W.np); memcpy(st + q, idepth_zero, sizeof(float) * W.np);
    LDSO_CUDA_OK(c, ldso_copy(c, W.idepth, st, sizeof(float) * (q + W.np), cudaMemcpyHostToDevice, c->stream));
    return 0;
}

// ---- K7 ------------------------------------------------------------------------------------------
extern "C" int ldso_gpu_window_linearize(ldso_gpu_ctx* c, int fix, double stats[3], float* new_th) {
    LDSO_ENTER(c);
    WindowDev& W = c->win;
    LDSO_REQUIRE(c, W.n > 0 && stats);
    int rc = k7_linearize(c, fix);
    if (rc) return rc;
    // read back [sum E, #IN] and the per-residual energies the frame threshold is drawn from
    const size_t bytes = 4 * sizeof(double) + sizeof(float) * (size_t)W.nr;
    char* st = (char*)staging(c, bytes);
    LDSO_REQUIRE(c, st != nullptr);
    LDSO_CUDA_OK(c, ldso_copy(c, st, W.stats, 4 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    if (W.nr) LDSO_CUDA_OK(c, ldso_copy(c, st + 4 * sizeof(double), W.new_energy_wo, sizeof(float) * W.nr, cudaMemcpyDeviceToHost, c->stream));
    LDSO_CUDA_OK(c, ldso_stream_sync(c->stream));
    W.pinned_off = 0;
    const double* ds = (const double*)st;
    stats[0] = ds[0]; stats[1] = 0; stats[2] = ds[1];
    // FullSystem::setNewFrameEnergyTH (src/frontend/FullSystem.cc:2078-2109): 70th percentile over the residuals
    // into the newest frame — a selection over ~2000 floats whose result the host needs anyway
    const float* e = (const float*)(st + 4 * sizeof(double));
    std::vector<float> all;
    const int newest = W.n - 1;
    for (int r = 0; r < W.nr; r++)
        if (!W.h_lin[r] && W.h_target[r] == newest && e[r] >= 0) all.push_back(e[r]);
    float th;
    if (all.empty()) {
        th = 12 * 12 * LDSO_PATTERN_NUM;
    } else {
        int nth = (int)(c->cfg.frame_energy_th_n * all.size());
        std::nth_element(all.begin(), all.begin() + nth, all.end());
        th = sqrtf(all[nth]) * c->cfg.frame_energy_th_fac_median;
        th = 26.0f * c->cfg.frame_energy_th_const_weight + th * (1 - c->cfg.frame_energy_th_const_weight);
        th = th * th;
        th *= c->cfg.overall_energy_th_weight * c->cfg.overall_energy_th_weight;
    }
    W.frameEnergyTH[newest] = th;
    if (new_th) *new_th = th;
    return 0;
}

extern "C" int ldso_gpu_window_apply_res(ldso_gpu_ctx* c) {
    LDSO_ENTER(c); return k7_apply_res(c); }
extern "C" int ldso_gpu_window_reset_oob(ldso_gpu_ctx* c) {
    LDSO_ENTER(c); return k7_reset_oob(c); }

extern "C" int ldso_gpu_window_fix_linearization(ldso_gpu_ctx* c, const unsigned char* mask) {
    LDSO_ENTER(c);
    WindowDev& W = c->win;
    LDSO_REQUIRE(c, mask || W.nr == 0);
    if (W.nr == 0) return 0;
    // the mask is staged in res_pt (scratch at this point)
    unsigned char* dmask = reinterpret_cast<unsigned char*>(W.res_pt);   // res_pt is rewritten by the next K8 launch
    int rc = upload(c, dmask, mask, W.nr);
    if (rc) return rc;
    rc = k7_fix_linearization(c, dmask);
    if (rc) return rc;
    for (int r = 0; r < W.nr; r++)
        if (mask[r]) W.h_lin[r] = 1;
    LDSO_CUDA_OK(c, ldso_stream_sync(c->stream));
    return build_order(c);
}

// ---- K8 + K9 + stitching (h1_stitch.cu) --------------------------------------------------------------
extern "C" int ldso_gpu_window_accumulate(ldso_gpu_ctx* c, double* HA, double* bA, double* HL, double* bL, double* Hsc,
                                          double* bsc, int counts[2]) {
    LDSO_ENTER(c);
    WindowDev& W = c->win;
    LDSO_REQUIRE(c, W.n > 0 && HA && bA && HL && bL && Hsc && bsc);
    LDSO_REQUIRE(c, (int)W.adHost.size() == 64 * W.n * W.n);
    int rc = k8_accumulate_top(c);
    if (rc) return rc;
    rc = k9_accumulate_sc(c);
    if (rc) return rc;
    rc = h1_stitch(c);
    if (rc) return rc;
    const int n = W.n, nn = n * n, D = 8 * n + CPARS;
    const size_t sysz = (size_t)D * D + D;
    const size_t bytes = sizeof(double) * 3 * sysz + sizeof(int) * 2 * nn;
    char* st = (char*)staging(c, bytes);
    LDSO_REQUIRE(c, st != nullptr);
    LDSO_CUDA_OK(c, ldso_copy(c, st, W.stitched, sizeof(double) * 3 * sysz, cudaMemcpyDeviceToHost, c->stream));
    LDSO_CUDA_OK(c, ldso_copy(c, st + sizeof(double) * 3 * sysz, W.acc_top_cnt, sizeof(int) * 2 * nn, cudaMemcpyDeviceToHost, c->stream));
    LDSO_CUDA_OK(c, ldso_stream_sync(c->stream));
    W.pinned_off = 0;
    const double* sys = (const double*)st;
    double* Hs[3] = {HA, HL, Hsc};
    double* bs[3] = {bA, bL, bsc};
    for (int q = 0; q < 3; q++) {
        memcpy(Hs[q], sys + q * sysz, sizeof(double) * D * D);
        memcpy(bs[q], sys + q * sysz + (size_t)D * D, sizeof(double) * D);
    }
    if (counts) {
        const int* cnt = (const int*)(st + sizeof(double) * 3 * sysz);
        counts[0] = counts[1] = 0;
        for (int k = 0; k < nn; k++) { counts[0] += cnt[k]; counts[1] += cnt[nn + k]; }
    }
    return 0;
}

// ---- H1: dense solve ------------------------------------------------------------------------------
namespace {

// x -= P x with P the symmetrised projector onto span(N) built from N's pseudo-inverse
// (EnergyFunctional::orthogonalize, EnergyFunctional.cc:809-841; only the `b` form is used by solveSystemF)
void project_out_nullspace(int D, const double* Nin, int k, double* x) {
    std::vector<double> N((size_t)D * k), U((size_t)D * k), S(k), V((size_t)k * k), Npi((size_t)D * k);
    for (int j = 0; j < k; j++) {
        double s = 0;
        for (int i = 0; i < D; i++) s += Nin[(size_t)i * k + j] * Nin[(size_t)i * k + j];
        s = std::sqrt(s);
        for (int i = 0; i < D; i++) N[(size_t)i * k + j] = Nin[(size_t)i * k + j] / s;
    }
    ldso_sm::jacobi_svd(D, k, N.data(), U.data(), S.data(), V.data());
    double smax = 0;
    for (int i = 0; i < k; i++) smax = std::max(smax, S[i]);
    for (int i = 0; i < k; i++) S[i] = S[i] > 1e-5 * smax ? 1.0 / S[i] : 0.0;   // setting_solverModeDelta (src/Setting.cc:24)
    for (int i = 0; i < D; i++)
        for (int j = 0; j < k; j++) {
            double s = 0;
            for (int l = 0; l < k; l++) s += U[(size_t)i * k + l] * S[l] * V[(size_t)j * k + l];
            Npi[(size_t)i * k + j] = s;
        }
    std::vector<double> px(k), nx(k);
    for (int j = 0; j < k; j++) {
        double a = 0, b = 0;
        for (int i = 0; i < D; i++) { a += Npi[(size_t)i * k + j] * x[i]; b += N[(size_t)i * k + j] * x[i]; }
        px[j] = a; nx[j] = b;
    }
    for (int i = 0; i < D; i++) {
        double s = 0;
        for (int j = 0; j < k; j++) s += N[(size_t)i * k + j] * px[j] + Npi[(size_t)i * k + j] * nx[j];
        x[i] -= 0.5 * s;
    }
}

}  // namespace

extern "C" int ldso_gpu_window_solve(ldso_gpu_ctx* c, int iteration, double lambda, const double* HA, const double* bA,
                                     const double* HL, const double* bL, const double* Hsc, const double* bsc, const double* HM,
                                     const double* bM, const double* delta, const double* N, int n_null, double* x_out) {
    LDSO_ENTER(c);
    WindowDev& W = c->win;
    LDSO_REQUIRE(c, W.n > 0 && HA && bA && HL && bL && Hsc && bsc && HM && bM && delta && x_out);
    const int D = 8 * W.n + CPARS;
    // setting_solverMode = SOLVER_FIX_LAMBDA | SOLVER_ORTHOGONALIZE_X_LATER (src/Setting.cc:23): lambda is pinned
    (void)lambda;
    const double lam = 1e-5;
    const double scw = (double)(1.0f / (1 + lam));
    std::vector<double> Hf((size_t)D * D), rhs(D), sdiag(D), y(D);
    for (int i = 0; i < D; i++) {
        double bm = bM[i];
        for (int j = 0; j < D; j++) bm += HM[(size_t)i * D + j] * delta[j];   // bM_top = bM + HM * delta (:310)
        rhs[i] = bL[i] + bm + bA[i] - bsc[i] / (1 + lam);
        for (int j = i; j < D; j++) {   // only the upper triangle is defined (:347, :378), mirrored by selfadjointView
            double v = HL[(size_t)i * D + j] + HM[(size_t)i * D + j] + HA[(size_t)i * D + j];
            if (j == i) v *= (1 + lam);
            v -= Hsc[(size_t)i * D + j] * scw;
            Hf[(size_t)i * D + j] = Hf[(size_t)j * D + i] = v;
        }
    }
    for (int i = 0; i < D; i++) sdiag[i] = 1.0 / std::sqrt(Hf[(size_t)i * D + i] + 10);   // :415
    for (int i = 0; i < D; i++) {
        rhs[i] *= sdiag[i];
        for (int j = 0; j < D; j++) Hf[(size_t)i * D + j] *= sdiag[i] * sdiag[j];
    }
    ldso_sm::ldlt_solve(D, Hf.data(), rhs.data(), y.data());
    for (int i = 0; i < D; i++) x_out[i] = sdiag[i] * y[i];
    if (iteration >= 2 && N && n_null > 0) project_out_nullspace(D, N, n_null, x_out);
    return 0;
}

// ---- K10 -----------------------------------------------------------------------------------------
extern "C" int ldso_gpu_window_resubstitute(ldso_gpu_ctx* c, const double* x, float lambda, float* step_out) {
    LDSO_ENTER(c);
    WindowDev& W = c->win;
    LDSO_REQUIRE(c, W.n > 0 && x && (int)W.adHostF.size() == 64 * W.n * W.n);
    const int n = W.n;
    // xAd[h*n + t] = x_h^T adHostF[h + n t] + x_t^T adTargetF[h + n t]   (EnergyFunctional.cc:618-627), float
    std::vector<float> tab((size_t)n * n * 8 + 4);
    for (int h = 0; h < n; h++)
        for (int t = 0; t < n; t++) {
            const float* AH = &W.adHostF[64 * (h + n * t)];
            const float* AT = &W.adTargetF[64 * (h + n * t)];
            for (int j = 0; j < 8; j++) {
                float a = 0, b = 0;
                for (int i = 0; i < 8; i++) {
                    a += (float)x[CPARS + 8 * h + i] * AH[i * 8 + j];
                    b += (float)x[CPARS + 8 * t + i] * AT[i * 8 + j];
                }
                tab[(size_t)(n * h + t) * 8 + j] = a + b;
            }
        }
    for (int i = 0; i < 4; i++) tab[(size_t)n * n * 8 + i] = (float)x[i];
    int rc = upload(c, W.xAd, tab.data(), sizeof(float) * tab.size());
    if (rc) return rc;
    rc = k10_resubstitute(c, lambda);
    if (rc) return rc;
    if (step_out) return download(c, step_out, W.pt_out + (size_t)PT_STEP * W.cap_p, sizeof(float) * W.np);
    return 0;
}

// ---- read-backs -----------------------------------------------------------------------------------
extern "C" int ldso_gpu_window_get_residual_floats(ldso_gpu_ctx* c, int what, float* dst) {
    LDSO_ENTER(c);
    WindowDev& W = c->win;
    LDSO_REQUIRE(c, dst && what >= 0 && what <= 6);
    const size_t nr = W.nr;
    switch (what) {
        case 0: return download(c, dst, W.J, sizeof(float) * nr * LDSO_J_FLOATS);
        case 1: return download(c, dst, W.rtz, sizeof(float) * nr * 8);
        case 2: return download(c, dst, W.JpJdF, sizeof(float) * nr * 8);
        case 3: {
            std::vector<float> t(nr * 4);
            int rc = download(c, t.data(), W.center, sizeof(float) * nr * 4);
            for (size_t r = 0; r < nr; r++) { dst[3 * r] = t[4 * r]; dst[3 * r + 1] = t[4 * r + 1]; dst[3 * r + 2] = t[4 * r + 2]; }
            return rc;
        }
        case 4: return download(c, dst, W.new_energy, sizeof(float) * nr);
        case 5: return download(c, dst, W.new_energy_wo, sizeof(float) * nr);
        default: return download(c, dst, W.energy, sizeof(float) * nr);
    }
}

extern "C" int ldso_gpu_window_get_residual_ints(ldso_gpu_ctx* c, int what, int* dst) {
    LDSO_ENTER(c);
    WindowDev& W = c->win;
    LDSO_REQUIRE(c, dst && what >= 0 && what <= 2);
    const int* src = what == 0 ? W.state : what == 1 ? W.new_state : W.flags;
    return download(c, dst, src, sizeof(int) * (size_t)W.nr);
}

extern "C" int ldso_gpu_window_get_point_floats(ldso_gpu_ctx* c, int what, float* dst) {
    LDSO_ENTER(c);
    WindowDev& W = c->win;
    LDSO_REQUIRE(c, dst && what >= 0 && what <= 11);
    const size_t np = W.np, cap = W.cap_p;
    auto row = [&](int r, float* out) { return download(c, out, W.pt_out + (size_t)r * cap, sizeof(float) * np); };
    auto rows4 = [&](int r0) {
        std::vector<float> t(np);
        for (int i = 0; i < 4; i++) {
            int rc = row(r0 + i, t.data());
            if (rc) return rc;
            for (size_t p = 0; p < np; p++) dst[4 * p + i] = t[p];
        }
        return 0;
    };
    switch (what) {
        case 0: return row(PT_HDD_A, dst);
        case 1: return row(PT_BD_A, dst);
        case 2: return rows4(PT_HCD_A);
        case 3: return row(PT_HDD_L, dst);
        case 4: return row(PT_BD_L, dst);
        case 5: return rows4(PT_HCD_L);
        case 6: return row(PT_HDI, dst);
        case 7: return row(PT_BDSUM, dst);
        case 8: return row(PT_IDH, dst);
        case 9: return download(c, dst, W.max_rel_baseline, sizeof(float) * np);
        case 10: return download(c, dst, W.num_good, sizeof(float) * np);
        default: return row(PT_STEP, dst);
    }
}

// what FullSystem reads after linearizeAll(true) (states -> lastResiduals, centerProjectedTo + HdiF -> makeCoarseDepthL0,
// idepth_hessian / numGoodResiduals -> flagPointsForRemoval): ONE staged device->host copy sequence and one sync
extern "C" int ldso_gpu_window_get_keyframe_results(ldso_gpu_ctx* c, int* state, float* center3, float* hdi, float* idepth_hessian,
                                                    float* num_good, float* max_rel_baseline) {
    LDSO_ENTER(c);
    WindowDev& W = c->win;
    const size_t nr = W.nr, np = W.np, cap = W.cap_p;
    const size_t bytes = nr * (sizeof(int) + 4 * sizeof(float)) + 4 * np * sizeof(float) + 64;
    char* st = (char*)staging(c, bytes);
    LDSO_REQUIRE(c, st != nullptr);
    int* h_state = (int*)st;
    float* h_center = (float*)(h_state + nr);
    float* h_pt = h_center + 4 * nr;
    if (nr) {
        LDSO_CUDA_OK(c, ldso_copy(c, h_state, W.state, sizeof(int) * nr, cudaMemcpyDeviceToHost, c->stream));
        LDSO_CUDA_OK(c, ldso_copy(c, h_center, W.center, sizeof(float) * 4 * nr, cudaMemcpyDeviceToHost, c->stream));
    }
    if (np) {
        LDSO_CUDA_OK(c, ldso_copy(c, h_pt, W.pt_out + (size_t)PT_HDI * cap, sizeof(float) * np, cudaMemcpyDeviceToHost, c->stream));
        LDSO_CUDA_OK(c, ldso_copy(c, h_pt + np, W.pt_out + (size_t)PT_IDH * cap, sizeof(float) * np, cudaMemcpyDeviceToHost, c->stream));
        LDSO_CUDA_OK(c, ldso_copy(c, h_pt + 2 * np, W.num_good, sizeof(float) * np, cudaMemcpyDeviceToHost, c->stream));
        LDSO_CUDA_OK(c, ldso_copy(c, h_pt + 3 * np, W.max_rel_baseline, sizeof(float) * np, cudaMemcpyDeviceToHost, c->stream));
    }
    LDSO_CUDA_OK(c, ldso_stream_sync(c->stream));
    W.pinned_off = 0;
    if (state) memcpy(state, h_state, sizeof(int) * nr);
    if (center3) for (size_t r = 0; r < nr; r++) { center3[3 * r] = h_center[4 * r]; center3[3 * r + 1] = h_center[4 * r + 1]; center3[3 * r + 2] = h_center[4 * r + 2]; }
    if (hdi) memcpy(hdi, h_pt, sizeof(float) * np);
    if (idepth_hessian) memcpy(idepth_hessian, h_pt + np, sizeof(float) * np);
    if (num_good) memcpy(num_good, h_pt + 2 * np, sizeof(float) * np);
    if (max_rel_baseline) memcpy(max_rel_baseline, h_pt + 3 * np, sizeof(float) * np);
    return 0;
}

// ---- marginalisation + energies (EnergyFunctional.cc:109-262, :473-497) -----------------------------
namespace {

// EnergyFunctional::removePoint for the masked points: their residuals leave the active set
__global__ void win_deactivate_kernel(int nr, const int* __restrict__ r_point, const unsigned char* __restrict__ mask,
                                      int* __restrict__ flags) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r < nr && mask[r_point[r]]) flags[r] &= ~LDSO_RF_ACTIVE;
}

// util::MatrixInverter::invertPosDef, SVD form (src/util/MatrixInverter.cc:31-50); only the upper triangle of the
// result is defined there, the caller reads it through selfadjointView<Upper>
void invert_pos_def8(const double* M, double* res) {
    const int n = 8;
    double SVecI[8], Ms[64], U[64], S[8], V[64];
    for (int i = 0; i < n; i++) SVecI[i] = 1.0 / std::sqrt(std::fabs(M[i * n + i]) + 10);
    for (int i = 0; i < n; i++)
        for (int j = i; j < n; j++) Ms[i * n + j] = Ms[j * n + i] = SVecI[i] * M[i * n + j] * SVecI[j];
    ldso_sm::jacobi_svd(n, n, Ms, U, S, V);
    for (int i = 0; i < n; i++) S[i] = S[i] <= 0 ? 0.0 : 1.0 / S[i];
    for (int i = 0; i < n; i++)
        for (int j = i; j < n; j++) {
            double s = 0;
            for (int l = 0; l < n; l++) s += V[i * n + l] * S[l] * U[j * n + l];
            res[i * n + j] = res[j * n + i] = SVecI[i] * s * SVecI[j];
        }
}

}  // namespace

extern "C" int ldso_gpu_window_marginalize_points(ldso_gpu_ctx* c, const unsigned char* marg, float fac, double* H, double* b,
                                                  int* n_res) {
    LDSO_ENTER(c);
    WindowDev& W = c->win;
    LDSO_REQUIRE(c, W.n > 0 && H && b && (marg || W.np == 0));
    LDSO_REQUIRE(c, (int)W.adHost.size() == 64 * W.n * W.n);
    const int n = W.n, nn = n * n, D = 8 * n + CPARS;
    int rc = upload(c, W.pmask, marg, W.np);
    if (rc) return rc;
    rc = build_order(c, marg);
    if (rc) return rc;
    rc = k8_accumulate_top(c, 1);
    if (rc) return rc;
    rc = k9_accumulate_sc(c, W.pmask, fac);
    if (rc) return rc;
    rc = h1_stitch(c);
    if (rc) return rc;
    if (W.nr > 0) {
        win_deactivate_kernel<<<(W.nr + 255) / 256, 256, 0, c->stream>>>(W.nr, W.r_point, W.pmask, W.flags);
        c->launches++;
    }
    const size_t sysz = (size_t)D * D + D;
    const size_t bytes = sizeof(double) * 3 * sysz + sizeof(int) * nn;
    char* st = (char*)staging(c, bytes);
    LDSO_REQUIRE(c, st != nullptr);
    LDSO_CUDA_OK(c, ldso_copy(c, st, W.stitched, sizeof(double) * 3 * sysz, cudaMemcpyDeviceToHost, c->stream));
    LDSO_CUDA_OK(c, ldso_copy(c, st + sizeof(double) * 3 * sysz, W.acc_top_cnt, sizeof(int) * nn, cudaMemcpyDeviceToHost, c->stream));
    LDSO_CUDA_OK(c, ldso_stream_sync(c->stream));
    W.pinned_off = 0;
    const double *M = (const double*)st, *Msc = M + 2 * sysz;
    for (size_t i = 0; i < (size_t)D * D; i++) H[i] = M[i] - Msc[i];
    for (int i = 0; i < D; i++) b[i] = M[(size_t)D * D + i] - Msc[(size_t)D * D + i];
    if (n_res) {
        const int* cnt = (const int*)(st + sizeof(double) * 3 * sysz);
        *n_res = 0;
        for (int k = 0; k < nn; k++) *n_res += cnt[k];
    }
    return build_order(c);   // back to the A/L work list of the remaining residuals (inactive ones are skipped by flag)
}

// FullSystem::flagPointsForRemoval's inner loop (FullSystem.cc:1390-1399) + EnergyFunctional::marginalizePointsF (:205-262) for a
// subset of the points of the window that is ALREADY resident (the one optimize just ran on): nothing is re-uploaded.
// The reference re-linearises every remaining residual of such a point (resetOOB, linearize, applyRes) before it fixes the
// linearisation.  State, tables and inverse depths have not moved since the last linearizeAll(true), so that pass reproduces the
// Jacobians bit for bit; the one input that did change is the newest frame's energy threshold (setNewFrameEnergyTH ran after the
// last linearisation), which can turn a residual IN -> OUTLIER.  The kernel therefore re-judges state_NewEnergyWithOutlier
// against max(frameEnergyTH[host], frameEnergyTH[target]), applies the result (isActive) and fixes the linearisation of the
// active ones; residuals the host has dropped since (dead != 0) stay out.
namespace {
__global__ void win_rejudge_fix_kernel(int nr, int n, const unsigned char* __restrict__ pmask, const unsigned char* __restrict__ dead,
                                       const int* __restrict__ r_point, const int* __restrict__ r_target, const int* __restrict__ host,
                                       const float* __restrict__ e_wo, const float* __restrict__ eth, int* __restrict__ state,
                                       int* __restrict__ new_state, int* __restrict__ flags, const float* __restrict__ J,
                                       const float* __restrict__ adHTdelta, float4 cDelta, const float* __restrict__ idepth,
                                       const float* __restrict__ idepth_zero, float* __restrict__ rtz, int* __restrict__ n_lin) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= nr) return;
    const int pt = r_point[r];
    if (!pmask[pt]) return;
    int fl = flags[r] & ~LDSO_RF_LINEARIZED;
    if (dead[r]) { flags[r] = fl & ~LDSO_RF_ACTIVE; return; }
    const int h = host[pt], t = r_target[r];
    const float th = fmaxf(eth[h], eth[t]);
    const int ns = e_wo[r] > th ? LDSO_RES_OUTLIER : LDSO_RES_IN;   // every remaining residual was IN with wJI2_sum >= 2
    new_state[r] = ns;
    state[r] = ns;                                                  // applyRes(true) after resetOOB (state IN): state = NewState
    if (ns != LDSO_RES_IN) { flags[r] = fl & ~LDSO_RF_ACTIVE; return; }
    fl |= LDSO_RF_ACTIVE;
    const float* j = J + (size_t)r * LDSO_J_FLOATS;
    const float* dp = adHTdelta + 8 * (h + n * t);
    const float deltaF = idepth[pt] - idepth_zero[pt];
    float dx = 0, dy = 0;
#pragma unroll
    for (int i = 0; i < 6; i++) { dx += j[40 + i] * dp[i]; dy += j[46 + i] * dp[i]; }
    const float cd[4] = {cDelta.x, cDelta.y, cDelta.z, cDelta.w};
    float cx = 0, cy = 0;
#pragma unroll
    for (int i = 0; i < 4; i++) { cx += j[52 + i] * cd[i]; cy += j[56 + i] * cd[i]; }
    const float Jx = dx + cx + j[60] * deltaF, Jy = dy + cy + j[61] * deltaF;
#pragma unroll
    for (int i = 0; i < 8; i++) {   // fixLinearizationF (Residuals.cc:219-245), as k7_fix_lin_kernel
        float v = j[i];
        v = v - j[8 + i] * Jx;
        v = v - j[16 + i] * Jy;
        v = v - j[24 + i] * dp[6];
        v = v - j[32 + i] * dp[7];
        rtz[8 * (size_t)r + i] = v;
    }
    flags[r] = fl | LDSO_RF_LINEARIZED;
    atomicAdd(n_lin, 1);
}
}  // namespace

extern "C" int ldso_gpu_window_marginalize_subset(ldso_gpu_ctx* c, const unsigned char* point_mask, const unsigned char* residual_dead,
                                                  const float* frame_energy_th, float fac, double* H, double* b, int* n_res,
                                                  int* n_lin) {
    LDSO_ENTER(c);
    WindowDev& W = c->win;
    LDSO_REQUIRE(c, W.n > 0 && H && b && frame_energy_th && (point_mask || W.np == 0) && (residual_dead || W.nr == 0));
    LDSO_REQUIRE(c, (int)W.adHost.size() == 64 * W.n * W.n);
    const int n = W.n, nn = n * n, D = 8 * n + CPARS;
    for (int i = 0; i < n; i++) W.frameEnergyTH[i] = frame_energy_th[i];
    // small uploads ride in scratch that the accumulate path rewrites anyway: thresholds + lin counter in xAd, dead mask in res_pt
    int rc = upload(c, W.pmask, point_mask, W.np);
    unsigned char* ddead = reinterpret_cast<unsigned char*>(W.res_pt);
    rc |= upload(c, ddead, residual_dead, W.nr);
    float tab[LDSO_MAX_WINDOW + 1];
    for (int i = 0; i < n; i++) tab[i] = frame_energy_th[i];
    tab[LDSO_MAX_WINDOW] = 0.f;
    float* deth = W.xAd;
    int* dlin = reinterpret_cast<int*>(W.xAd + LDSO_MAX_WINDOW);
    rc |= upload(c, deth, tab, sizeof(tab));
    if (rc) return -1;
    if (W.nr > 0) {
        float4 cd = make_float4(W.cDeltaF[0], W.cDeltaF[1], W.cDeltaF[2], W.cDeltaF[3]);
        win_rejudge_fix_kernel<<<(W.nr + 255) / 256, 256, 0, c->stream>>>(W.nr, n, W.pmask, ddead, W.r_point, W.r_target, W.host,
                                                                         W.new_energy_wo, deth, W.state, W.new_state, W.flags, W.J,
                                                                         W.adHTdelta, cd, W.idepth, W.idepth_zero, W.rtz, dlin);
        c->launches++;
        LDSO_CUDA_OK(c, cudaGetLastError());
    }
    int lin_host = 0;
    char* stl = (char*)staging(c, 64);
    LDSO_REQUIRE(c, stl != nullptr);
    LDSO_CUDA_OK(c, ldso_copy(c, stl, dlin, sizeof(int), cudaMemcpyDeviceToHost, c->stream));   // before K8 reuses the scratch? xAd is K10's
    rc = build_order(c, point_mask);
    if (rc) return rc;
    rc = k8_accumulate_top(c, 1);
    if (rc) return rc;
    rc = k9_accumulate_sc(c, W.pmask, fac);
    if (rc) return rc;
    rc = h1_stitch(c);
    if (rc) return rc;
    if (W.nr > 0) {
        win_deactivate_kernel<<<(W.nr + 255) / 256, 256, 0, c->stream>>>(W.nr, W.r_point, W.pmask, W.flags);
        c->launches++;
    }
    const size_t sysz = (size_t)D * D + D;
    const size_t bytes = sizeof(double) * 3 * sysz + sizeof(int) * nn;
    char* st = (char*)staging(c, bytes);
    LDSO_REQUIRE(c, st != nullptr);
    LDSO_CUDA_OK(c, ldso_copy(c, st, W.stitched, sizeof(double) * 3 * sysz, cudaMemcpyDeviceToHost, c->stream));
    LDSO_CUDA_OK(c, ldso_copy(c, st + sizeof(double) * 3 * sysz, W.acc_top_cnt, sizeof(int) * nn, cudaMemcpyDeviceToHost, c->stream));
    LDSO_CUDA_OK(c, ldso_stream_sync(c->stream));
    lin_host = *(const int*)stl;
    W.pinned_off = 0;
    const double *M = (const double*)st, *Msc = M + 2 * sysz;
    for (size_t i = 0; i < (size_t)D * D; i++) H[i] = M[i] - Msc[i];
    for (int i = 0; i < D; i++) b[i] = M[(size_t)D * D + i] - Msc[(size_t)D * D + i];
    if (n_res) {
        const int* cnt = (const int*)(st + sizeof(double) * 3 * sysz);
        *n_res = 0;
        for (int k = 0; k < nn; k++) *n_res += cnt[k];
    }
    if (n_lin) *n_lin = lin_host;
    return 0;   // the window is not reusable for another optimisation: the next keyframe uploads it afresh
}

extern "C" int ldso_gpu_window_marginalize_frame(ldso_gpu_ctx* c, int n, int idx, const double* HMin, const double* bMin,
                                                 const double prior[8], const double delta_prior[8], double* HM_out,
                                                 double* bM_out) {
    LDSO_ENTER(c);
    LDSO_REQUIRE(c, n >= 1 && n <= LDSO_MAX_WINDOW && idx >= 0 && idx < n && HMin && bMin && prior && delta_prior && HM_out && bM_out);
    const int odim = 8 * n + CPARS, ndim = odim - 8;
    // permutation that moves the frame's 8 rows/cols to the end, keeping the order of the rest (:127-146)
    std::vector<int> perm(odim);
    {
        int k = 0;
        const int io = 8 * idx + CPARS;
        for (int i = 0; i < odim; i++) if (i < io || i >= io + 8) perm[k++] = i;
        for (int i = 0; i < 8; i++) perm[k++] = io + i;
    }
    std::vector<double> HM((size_t)odim * odim), bM(odim), SVec(odim), SVecI(odim);
    for (int i = 0; i < odim; i++) {
        bM[i] = bMin[perm[i]];
        for (int j = 0; j < odim; j++) HM[(size_t)i * odim + j] = HMin[(size_t)perm[i] * odim + perm[j]];
    }
    for (int i = 0; i < 8; i++) {   // the frame's own prior goes in here instead of into the active system (:150-151)
        HM[(size_t)(ndim + i) * odim + ndim + i] += prior[i];
        bM[ndim + i] += prior[i] * delta_prior[i];
    }
    for (int i = 0; i < odim; i++) { SVec[i] = std::sqrt(std::fabs(HM[(size_t)i * odim + i]) + 10); SVecI[i] = 1.0 / SVec[i]; }
    for (int i = 0; i < odim; i++) {
        bM[i] *= SVecI[i];
        for (int j = 0; j < odim; j++) HM[(size_t)i * odim + j] *= SVecI[i] * SVecI[j];
    }
    double hp[64], hpi[64];
    for (int i = 0; i < 8; i++)
        for (int j = 0; j < 8; j++) hp[i * 8 + j] = HM[(size_t)(ndim + i) * odim + ndim + j];
    invert_pos_def8(hp, hpi);
    // bli = bottomLeft^T * hpi ; top-left -= bli * bottomLeft ; b_head -= bli * b_tail   (:166-169)
    std::vector<double> bli((size_t)ndim * 8);
    for (int i = 0; i < ndim; i++)
        for (int j = 0; j < 8; j++) {
            double s = 0;
            for (int l = 0; l < 8; l++) s += HM[(size_t)(ndim + l) * odim + i] * hpi[l * 8 + j];
            bli[(size_t)i * 8 + j] = s;
        }
    std::vector<double> Hn((size_t)ndim * ndim);
    for (int i = 0; i < ndim; i++) {
        double sb = 0;
        for (int l = 0; l < 8; l++) sb += bli[(size_t)i * 8 + l] * bM[ndim + l];
        bM_out[i] = (bM[i] - sb) * SVec[i];
        for (int j = 0; j < ndim; j++) {
            double s = 0;
            for (int l = 0; l < 8; l++) s += bli[(size_t)i * 8 + l] * HM[(size_t)(ndim + l) * odim + j];
            Hn[(size_t)i * ndim + j] = (HM[(size_t)i * odim + j] - s) * SVec[i] * SVec[j];
        }
    }
    for (int i = 0; i < ndim; i++)
        for (int j = 0; j < ndim; j++) HM_out[(size_t)i * ndim + j] = 0.5 * (Hn[(size_t)i * ndim + j] + Hn[(size_t)j * ndim + i]);
    return 0;
}

extern "C" int ldso_gpu_window_calc_energy(ldso_gpu_ctx* c, const double* HM, const double* bM, const double* delta,
                                           double energy[2]) {
    LDSO_ENTER(c);
    WindowDev& W = c->win;
    LDSO_REQUIRE(c, W.n > 0 && HM && bM && delta && energy && (int)W.fprior.size() == 8 * W.n);
    const int D = 8 * W.n + CPARS;
    int rc = k8_linearized_energy(c, W.stats + 2);
    if (rc) return rc;
    double eL = 0;
    rc = download(c, &eL, W.stats + 2, sizeof(double));
    if (rc) return rc;
    double E = 0;
    for (int i = 0; i < 8 * W.n; i++) E += W.fdelta_prior[i] * W.fprior[i] * W.fdelta_prior[i];
    float ec = 0;
    for (int i = 0; i < 4; i++) ec += W.cDeltaF[i] * (float)W.cPrior[i] * W.cDeltaF[i];   // cDeltaF.cwiseProduct(cPriorF).dot(cDeltaF)
    energy[0] = E + ec + eL;
    double em = 0;   // calcMEnergyF: delta . (2 bM + HM delta)
    for (int i = 0; i < D; i++) {
        double s = 2 * bM[i];
        for (int j = 0; j < D; j++) s += HM[(size_t)i * D + j] * delta[j];
        em += delta[i] * s;
    }
    energy[1] = em;
    return 0;
}
