// oracle/orc_frontend.cpp — TEST INFRASTRUCTURE (see orc_common.h).  PARITY UNPINNED.
// CPU restatement of: setGlobalCalib, FrameHessian::makeImages, CoarseTracker::{makeK,
// makeCoarseDepthL0, calcRes, calcGSSSE, trackNewestCoarse}, Accumulator9.
#include "orc_common.h"

#include <map>
#include <mutex>

#include <algorithm>

using namespace orc;
using ldso_sm::SE3;

// =================================================================================================
// config defaults: src/Setting.cc:8-182 + preset 0 (examples/run_dso_tum_mono.cc:65-83)
extern "C" void orc_default_config(ldso_gpu_config* c) {
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

// src/internal/GlobalCalib.cc:20-28 (level rule) and :43-72 (sizes)
static int derive_levels(int w, int h, int* wG, int* hG) {
    int wlvl = w, hlvl = h, lv = 1;
    while (wlvl % 2 == 0 && hlvl % 2 == 0 && wlvl * hlvl > 5000 && lv < LDSO_MAX_LEVELS) {
        wlvl /= 2; hlvl /= 2; lv++;
    }
    for (int l = 0; l < lv; l++) { wG[l] = w >> l; hG[l] = h >> l; }
    return lv;
}

static void tracker_alloc(orc_ctx* c, Tracker& t) {
    // CoarseTracker::CoarseTracker (src/frontend/CoarseTracker.cc:30-59)
    int ww = c->wG[0], hh = c->hG[0];
    for (int l = 0; l < c->levels; l++) {
        int n = (ww >> l) * (hh >> l);
        t.idepth[l].assign(n, 0); t.weightSums[l].assign(n, 0); t.weightSums_bak[l].assign(n, 0);
        t.pc_u[l].assign(n, 0); t.pc_v[l].assign(n, 0); t.pc_idepth[l].assign(n, 0); t.pc_color[l].assign(n, 0);
        t.pc_n[l] = 0;
    }
    size_t n = (size_t)ww * hh + 8;
    t.buf_store.assign(8 * n + 8, 0);
    float* base = (float*)((((uintptr_t)t.buf_store.data()) + 15) & ~(uintptr_t)15);
    t.buf_warped_idepth = base; t.buf_warped_u = base + n; t.buf_warped_v = base + 2 * n; t.buf_warped_dx = base + 3 * n;
    t.buf_warped_dy = base + 4 * n; t.buf_warped_residual = base + 5 * n; t.buf_warped_weight = base + 6 * n;
    t.buf_warped_refColor = base + 7 * n;
    t.w[0] = t.h[0] = 0;
}

extern "C" int orc_tracker_make_k(orc_ctx* c, int tracker, float fx, float fy, float cx, float cy);

extern "C" int orc_create(const ldso_gpu_config* cfg, orc_ctx** out) {
    orc_ctx* c = new orc_ctx();
    c->cfg = *cfg;
    c->levels = derive_levels(cfg->w, cfg->h, c->wG, c->hG);
    for (int i = 0; i < 256; i++) c->B[i] = (float)i;
    if (const char* e = getenv("ORC_THREADS")) c->threads = std::max(1, atoi(e));
    c->frames.resize(cfg->n_frame_slots);
    for (auto& f : c->frames)
        for (int l = 0; l < c->levels; l++) {
            f.dIp[l].assign((size_t)3 * c->wG[l] * c->hG[l], 0);
            f.absSquaredGrad[l].assign((size_t)c->wG[l] * c->hG[l], 0);
        }
    for (int t = 0; t < LDSO_NUM_TRACKERS; t++) {
        tracker_alloc(c, c->trk[t]);
        orc_tracker_make_k(c, t, cfg->fx, cfg->fy, cfg->cx, cfg->cy);
    }
    *out = c;
    return 0;
}
extern "C" void orc_destroy(orc_ctx* c) { delete c; }
extern "C" int orc_io_bytes(orc_ctx*, long long* h2d, long long* d2h) { *h2d = 0; *d2h = 0; return 0; }   // no device
extern "C" const char* orc_last_error(const orc_ctx* c) { return c->err.c_str(); }
extern "C" int orc_pyr_levels(const orc_ctx* c) { return c->levels; }
extern "C" int orc_level_size(const orc_ctx* c, int lvl, int* w, int* h) {
    if (lvl < 0 || lvl >= c->levels) return -1;
    *w = c->wG[lvl]; *h = c->hG[lvl];
    return 0;
}
extern "C" int orc_sync(orc_ctx*) { return 0; }
extern "C" int orc_bind_thread(orc_ctx*) { return 0; }   // no device on the CPU
extern "C" long long orc_launch_count(const orc_ctx*) { return 0; }
extern "C" int orc_set_gamma(orc_ctx* c, const float B[256]) { memcpy(c->B, B, sizeof(float) * 256); return 0; }

// =================================================================================================
// K1 — FrameHessian::makeImages (src/internal/FrameHessian.cc:59-115)
static inline float getBGradOnly(const float* B, float color) {   // include/internal/CalibHessian.h:102-111
    int c = color + 0.5f;
    if (c < 5) c = 5;
    if (c > 250) c = 250;
    return B[c + 1] - B[c];
}

extern "C" int orc_frame_make_images(orc_ctx* c, int slot, const float* color) {
    if (slot < 0 || slot >= (int)c->frames.size()) return -1;
    Frame& f = c->frames[slot];
    for (int i = 0; i < c->levels; i++) {   // :61-66 new[] + memset
        std::fill(f.dIp[i].begin(), f.dIp[i].end(), 0.f);
        std::fill(f.absSquaredGrad[i].begin(), f.absSquaredGrad[i].end(), 0.f);
    }
    int w = c->wG[0], h = c->hG[0];
    float* dI = f.dIp[0].data();
    for (int i = 0; i < w * h; i++) dI[3 * i] = color[i];   // :72-74
    for (int lvl = 0; lvl < c->levels; lvl++) {
        int wl = c->wG[lvl], hl = c->hG[lvl];
        float* dI_l = f.dIp[lvl].data();
        float* dabs_l = f.absSquaredGrad[lvl].data();
        if (lvl > 0) {   // :80-93
            int wlm1 = c->wG[lvl - 1];
            const float* dI_lm = f.dIp[lvl - 1].data();
            for (int y = 0; y < hl; y++)
                for (int x = 0; x < wl; x++) {
                    dI_l[3 * (x + y * wl)] = 0.25f * (dI_lm[3 * (2 * x + 2 * y * wlm1)] + dI_lm[3 * (2 * x + 1 + 2 * y * wlm1)] +
                                                      dI_lm[3 * (2 * x + 2 * y * wlm1 + wlm1)] +
                                                      dI_lm[3 * (2 * x + 1 + 2 * y * wlm1 + wlm1)]);
                }
        }
        for (int idx = wl; idx < wl * (hl - 1); idx++) {   // :95-114
            float dx = 0.5f * (dI_l[3 * (idx + 1)] - dI_l[3 * (idx - 1)]);
            float dy = 0.5f * (dI_l[3 * (idx + wl)] - dI_l[3 * (idx - wl)]);
            if (std::isnan(dx) || std::fabs(dx) > 255.0) dx = 0;
            if (std::isnan(dy) || std::fabs(dy) > 255.0) dy = 0;
            dI_l[3 * idx + 1] = dx;
            dI_l[3 * idx + 2] = dy;
            dabs_l[idx] = dx * dx + dy * dy;
            if (c->cfg.gamma_weights_pixel_select == 1) {
                float gw = getBGradOnly(c->B, dI_l[3 * idx]);
                dabs_l[idx] *= gw * gw;
            }
        }
    }
    f.valid = true;
    return 0;
}
extern "C" int orc_frame_make_images_dev(orc_ctx* c, int slot, const float* img) { return orc_frame_make_images(c, slot, img); }

extern "C" int orc_frame_download(orc_ctx* c, int slot, int lvl, int what, float* dst) {
    if (slot < 0 || slot >= (int)c->frames.size() || lvl < 0 || lvl >= c->levels) return -1;
    Frame& f = c->frames[slot];
    if (what == 0) memcpy(dst, f.dIp[lvl].data(), f.dIp[lvl].size() * sizeof(float));
    else memcpy(dst, f.absSquaredGrad[lvl].data(), f.absSquaredGrad[lvl].size() * sizeof(float));
    return 0;
}

// =================================================================================================
// Accumulator9 (include/internal/OptimizationBackend/MatrixAccumulators.h:1104-1643)
void Accumulator9::initialize() {   // :1112-1119
    memset(H, 0, sizeof(H));
    memset(SSEData, 0, sizeof(SSEData)); memset(SSEData1k, 0, sizeof(SSEData1k)); memset(SSEData1m, 0, sizeof(SSEData1m));
    num = 0; numIn1 = numIn1k = numIn1m = 0;
}
void Accumulator9::shiftUp(bool force) {   // :1624-1642
    if (numIn1 > 1000 || force) {
        for (int i = 0; i < 45; i++)
            _mm_store_ps(SSEData1k + 4 * i, _mm_add_ps(_mm_load_ps(SSEData + 4 * i), _mm_load_ps(SSEData1k + 4 * i)));
        numIn1k += numIn1; numIn1 = 0;
        memset(SSEData, 0, sizeof(SSEData));
    }
    if (numIn1k > 1000 || force) {
        for (int i = 0; i < 45; i++)
            _mm_store_ps(SSEData1m + 4 * i, _mm_add_ps(_mm_load_ps(SSEData1k + 4 * i), _mm_load_ps(SSEData1m + 4 * i)));
        numIn1m += numIn1k; numIn1k = 0;
        memset(SSEData1k, 0, sizeof(SSEData1k));
    }
}
void Accumulator9::finish() {   // :1121-1135
    memset(H, 0, sizeof(H));
    shiftUp(true);
    int idx = 0;
    for (int r = 0; r < 9; r++)
        for (int c = r; c < 9; c++) {
            float d = SSEData1m[idx + 0] + SSEData1m[idx + 1] + SSEData1m[idx + 2] + SSEData1m[idx + 3];
            H[r][c] = H[c][r] = d;
            idx += 4;
        }
}
void Accumulator9::updateSSE_eighted(const __m128* J, __m128 w) {   // :1250-1370 (45 unrolled stores)
    float* pt = SSEData;
    for (int r = 0; r < 9; r++) {
        __m128 Jw = _mm_mul_ps(J[r], w);
        for (int c = r; c < 9; c++) {
            _mm_store_ps(pt, _mm_add_ps(_mm_load_ps(pt), _mm_mul_ps(Jw, J[c])));
            pt += 4;
        }
    }
    num += 4;
    numIn1++;
    shiftUp(false);
}
void Accumulator9::updateSSE(const __m128* J) {   // :1138-1248
    float* pt = SSEData;
    for (int r = 0; r < 9; r++)
        for (int c = r; c < 9; c++) {
            _mm_store_ps(pt, _mm_add_ps(_mm_load_ps(pt), _mm_mul_ps(J[r], J[c])));
            pt += 4;
        }
    num += 4;
    numIn1++;
    shiftUp(false);
}
void Accumulator9::updateSingle(const float* J, int off) {   // :1373-1490 (lane `off` of each 4-group)
    float* pt = SSEData + off;
    for (int r = 0; r < 9; r++)
        for (int c = r; c < 9; c++) { *pt += J[r] * J[c]; pt += 4; }
    num++;
    numIn1++;
    shiftUp(false);
}
void Accumulator9::updateSingleWeighted(const float* Jin, float w, int off) {   // :1493-1615
    float J[9];
    for (int i = 0; i < 9; i++) J[i] = Jin[i];
    float* pt = SSEData + off;
    for (int r = 0; r < 9; r++) {
        float Jw = J[r] * w;
        for (int c = r; c < 9; c++) { *pt += Jw * J[c]; pt += 4; }
    }
    num++;
    numIn1++;
    shiftUp(false);
}

// =================================================================================================
// K2 — CoarseTracker::makeK (src/frontend/CoarseTracker.cc:312-339)
extern "C" int orc_tracker_make_k(orc_ctx* c, int tracker, float fx, float fy, float cx, float cy) {
    if (tracker < 0 || tracker >= LDSO_NUM_TRACKERS) return -1;
    Tracker& t = c->trk[tracker];
    t.w[0] = c->wG[0]; t.h[0] = c->hG[0];
    t.fx[0] = fx; t.fy[0] = fy; t.cx[0] = cx; t.cy[0] = cy;
    for (int level = 1; level < c->levels; ++level) {
        t.w[level] = t.w[0] >> level;
        t.h[level] = t.h[0] >> level;
        t.fx[level] = t.fx[level - 1] * 0.5;
        t.fy[level] = t.fy[level - 1] * 0.5;
        t.cx[level] = (t.cx[0] + 0.5) / ((int)1 << level) - 0.5;
        t.cy[level] = (t.cy[0] + 0.5) / ((int)1 << level) - 0.5;
    }
    for (int level = 0; level < c->levels; ++level) {
        float* K = t.K[level];
        K[0] = t.fx[level]; K[1] = 0; K[2] = t.cx[level];
        K[3] = 0; K[4] = t.fy[level]; K[5] = t.cy[level];
        K[6] = 0; K[7] = 0; K[8] = 1;
        ldso_sm::mat3_inv<float>(K, t.Ki[level]);
    }
    return 0;
}

// K2 — CoarseTracker::setCoarseTrackingRef / makeCoarseDepthL0 (src/frontend/CoarseTracker.cc:347-538)
// The walk over frames/features/points (:363-382) is host bookkeeping: the caller hands over
// centerProjectedTo and HdiF of every qualifying point in that iteration order.
extern "C" int orc_tracker_set_ref(orc_ctx* c, int tracker, int ref_slot, int n_points, const float* pu, const float* pv,
                                   const float* pid, const float* phdi, float ref_exposure, float ref_aff_a, float ref_aff_b) {
    if (tracker < 0 || tracker >= LDSO_NUM_TRACKERS) return -1;
    Tracker& t = c->trk[tracker];
    t.ref_slot = ref_slot; t.ref_exposure = ref_exposure; t.ref_aff_a = ref_aff_a; t.ref_aff_b = ref_aff_b;
    const int* w = t.w; const int* h = t.h;
    float** idepth = nullptr; (void)idepth;
    memset(t.idepth[0].data(), 0, sizeof(float) * w[0] * h[0]);
    memset(t.weightSums[0].data(), 0, sizeof(float) * w[0] * h[0]);
    for (int k = 0; k < n_points; k++) {   // :363-382
        int u = pu[k] + 0.5f;
        int v = pv[k] + 0.5f;
        float new_idepth = pid[k];
        float weight = sqrtf(1e-3 / (phdi[k] + 1e-12));
        t.idepth[0][u + w[0] * v] += new_idepth * weight;
        t.weightSums[0][u + w[0] * v] += weight;
    }
    for (int lvl = 1; lvl < c->levels; lvl++) {   // :384-409
        int lvlm1 = lvl - 1;
        int wl = w[lvl], hl = h[lvl], wlm1 = w[lvlm1];
        float* idepth_l = t.idepth[lvl].data(); float* weightSums_l = t.weightSums[lvl].data();
        float* idepth_lm = t.idepth[lvlm1].data(); float* weightSums_lm = t.weightSums[lvlm1].data();
        for (int y = 0; y < hl; y++)
            for (int x = 0; x < wl; x++) {
                int bidx = 2 * x + 2 * y * wlm1;
                idepth_l[x + y * wl] = idepth_lm[bidx] + idepth_lm[bidx + 1] + idepth_lm[bidx + wlm1] + idepth_lm[bidx + wlm1 + 1];
                weightSums_l[x + y * wl] =
                    weightSums_lm[bidx] + weightSums_lm[bidx + 1] + weightSums_lm[bidx + wlm1] + weightSums_lm[bidx + wlm1 + 1];
            }
    }
    for (int lvl = 0; lvl < c->levels; lvl++) {   // :412-454 (lvl<2: diagonal) and :458-495 (lvl>=2: axis)
        int wh = w[lvl] * h[lvl] - w[lvl];
        int wl = w[lvl];
        float* weightSumsl = t.weightSums[lvl].data();
        float* weightSumsl_bak = t.weightSums_bak[lvl].data();
        memcpy(weightSumsl_bak, weightSumsl, w[lvl] * h[lvl] * sizeof(float));
        float* idepthl = t.idepth[lvl].data();
        int o0, o1, o2, o3;
        if (lvl < 2) { o0 = 1 + wl; o1 = -1 - wl; o2 = wl - 1; o3 = -wl + 1; }
        else { o0 = 1; o1 = -1; o2 = wl; o3 = -wl; }
        for (int i = w[lvl]; i < wh; i++) {
            if (weightSumsl_bak[i] <= 0) {
                float sum = 0, num = 0, numn = 0;
                // the reference reads one element before / after its arrays for the first cell of row 1 and the last cell of
                // row h-2 (:437-466; cells that never reach the point cloud): a neighbour outside the plane counts as weight 0
                const int nAll = w[lvl] * h[lvl];
                auto wgt = [&](int j) { return (j >= 0 && j < nAll) ? weightSumsl_bak[j] : 0.f; };
                if (wgt(i + o0) > 0) { sum += idepthl[i + o0]; num += weightSumsl_bak[i + o0]; numn++; }
                if (wgt(i + o1) > 0) { sum += idepthl[i + o1]; num += weightSumsl_bak[i + o1]; numn++; }
                if (wgt(i + o2) > 0) { sum += idepthl[i + o2]; num += weightSumsl_bak[i + o2]; numn++; }
                if (wgt(i + o3) > 0) { sum += idepthl[i + o3]; num += weightSumsl_bak[i + o3]; numn++; }
                if (numn > 0) { idepthl[i] = sum / numn; weightSumsl[i] = num / numn; }
            }
        }
    }
    const Frame& ref = c->frames[ref_slot];
    for (int lvl = 0; lvl < c->levels; lvl++) {   // :498-537
        float* weightSumsl = t.weightSums[lvl].data();
        float* idepthl = t.idepth[lvl].data();
        const float* dIRefl = ref.dIp[lvl].data();
        int wl = w[lvl], hl = h[lvl];
        int lpc_n = 0;
        float* lpc_u = t.pc_u[lvl].data(); float* lpc_v = t.pc_v[lvl].data();
        float* lpc_idepth = t.pc_idepth[lvl].data(); float* lpc_color = t.pc_color[lvl].data();
        for (int y = 2; y < hl - 2; y++)
            for (int x = 2; x < wl - 2; x++) {
                int i = x + y * wl;
                if (weightSumsl[i] > 0) {
                    idepthl[i] /= weightSumsl[i];
                    lpc_u[lpc_n] = x; lpc_v[lpc_n] = y;
                    lpc_idepth[lpc_n] = idepthl[i];
                    lpc_color[lpc_n] = dIRefl[3 * i];
                    if (!std::isfinite(lpc_color[lpc_n]) || !(idepthl[i] > 0)) {
                        idepthl[i] = -1;
                        continue;
                    }
                    lpc_n++;
                } else
                    idepthl[i] = -1;
                weightSumsl[i] = 1;
            }
        t.pc_n[lvl] = lpc_n;
    }
    return 0;
}

extern "C" int orc_tracker_get_pc(orc_ctx* c, int tracker, int lvl, int* n, float* u, float* v, float* idepth, float* color) {
    Tracker& t = c->trk[tracker];
    int m = t.pc_n[lvl];
    *n = m;
    if (u) memcpy(u, t.pc_u[lvl].data(), m * sizeof(float));
    if (v) memcpy(v, t.pc_v[lvl].data(), m * sizeof(float));
    if (idepth) memcpy(idepth, t.pc_idepth[lvl].data(), m * sizeof(float));
    if (color) memcpy(color, t.pc_color[lvl].data(), m * sizeof(float));
    return 0;
}

// =================================================================================================
// include/AffLight.h:27-35
static inline void fromToVecExposure(float exposureF, float exposureT, float g2F_a, float g2F_b, float g2T_a, float g2T_b,
                                     double* out) {
    if (exposureF == 0 || exposureT == 0) exposureT = exposureF = 1;
    float a = exp(g2T_a - g2F_a) * exposureT / exposureF;
    float b = g2T_b - a * g2F_b;
    out[0] = a; out[1] = b;
}

// include/internal/GlobalFuncs.h:89-103
static inline void getInterpolatedElement33(const float* mat, float x, float y, int width, float* out) {
    int ix = (int)x, iy = (int)y;
    float dx = x - ix, dy = y - iy, dxdy = dx * dy;
    const float* bp = mat + 3 * (ix + iy * width);
    for (int k = 0; k < 3; k++)
        out[k] = dxdy * bp[3 * (1 + width) + k] + (dy - dxdy) * bp[3 * width + k] + (dx - dxdy) * bp[3 + k] +
                 (1 - dx - dy + dxdy) * bp[k];
}

// K3 — CoarseTracker::calcRes (src/frontend/CoarseTracker.cc:540-673)
static void calcRes(orc_ctx* c, Tracker& T, int lvl, const SE3& refToNew, float aff_a, float aff_b, float cutoffTH, double* rs) {
    float E = 0;
    int numTermsInE = 0, numTermsInWarped = 0, numSaturated = 0;
    int wl = T.w[lvl], hl = T.h[lvl];
    const float* dINewl = c->frames[T.new_slot].dIp[lvl].data();
    float fxl = T.fx[lvl], fyl = T.fy[lvl], cxl = T.cx[lvl], cyl = T.cy[lvl];
    double Rd[9];
    refToNew.rotationMatrix(Rd);
    float Rf[9], RKi[9];
    for (int i = 0; i < 9; i++) Rf[i] = (float)Rd[i];
    ldso_sm::mat3_mul<float>(Rf, T.Ki[lvl], RKi);
    float t[3] = {(float)refToNew.t[0], (float)refToNew.t[1], (float)refToNew.t[2]};
    double affd[2];
    fromToVecExposure(T.ref_exposure, T.new_exposure, T.ref_aff_a, T.ref_aff_b, aff_a, aff_b, affd);
    float affLL[2] = {(float)affd[0], (float)affd[1]};
    float sumSquaredShiftT = 0, sumSquaredShiftRT = 0, sumSquaredShiftNum = 0;
    const float huberTH = c->cfg.huber_th;
    float maxEnergy = 2 * huberTH * cutoffTH - huberTH * huberTH;
    int nl = T.pc_n[lvl];
    const float* lpc_u = T.pc_u[lvl].data(); const float* lpc_v = T.pc_v[lvl].data();
    const float* lpc_idepth = T.pc_idepth[lvl].data(); const float* lpc_color = T.pc_color[lvl].data();
    const float* Ki = T.Ki[lvl];
    for (int i = 0; i < nl; i++) {
        float id = lpc_idepth[i], x = lpc_u[i], y = lpc_v[i];
        float pt[3] = {RKi[0] * x + RKi[1] * y + RKi[2] * 1 + t[0] * id, RKi[3] * x + RKi[4] * y + RKi[5] * 1 + t[1] * id,
                       RKi[6] * x + RKi[7] * y + RKi[8] * 1 + t[2] * id};
        float u = pt[0] / pt[2], v = pt[1] / pt[2];
        float Ku = fxl * u + cxl, Kv = fyl * v + cyl;
        float new_idepth = id / pt[2];
        if (lvl == 0 && i % 32 == 0) {   // :587-617
            float k[3] = {Ki[0] * x + Ki[1] * y + Ki[2] * 1, Ki[3] * x + Ki[4] * y + Ki[5] * 1, Ki[6] * x + Ki[7] * y + Ki[8] * 1};
            float ptT[3] = {k[0] + t[0] * id, k[1] + t[1] * id, k[2] + t[2] * id};
            float KuT = fxl * (ptT[0] / ptT[2]) + cxl, KvT = fyl * (ptT[1] / ptT[2]) + cyl;
            float ptT2[3] = {k[0] - t[0] * id, k[1] - t[1] * id, k[2] - t[2] * id};
            float KuT2 = fxl * (ptT2[0] / ptT2[2]) + cxl, KvT2 = fyl * (ptT2[1] / ptT2[2]) + cyl;
            float r[3] = {RKi[0] * x + RKi[1] * y + RKi[2] * 1, RKi[3] * x + RKi[4] * y + RKi[5] * 1, RKi[6] * x + RKi[7] * y + RKi[8] * 1};
            float pt3[3] = {r[0] - t[0] * id, r[1] - t[1] * id, r[2] - t[2] * id};
            float Ku3 = fxl * (pt3[0] / pt3[2]) + cxl, Kv3 = fyl * (pt3[1] / pt3[2]) + cyl;
            sumSquaredShiftT += (KuT - x) * (KuT - x) + (KvT - y) * (KvT - y);
            sumSquaredShiftT += (KuT2 - x) * (KuT2 - x) + (KvT2 - y) * (KvT2 - y);
            sumSquaredShiftRT += (Ku - x) * (Ku - x) + (Kv - y) * (Kv - y);
            sumSquaredShiftRT += (Ku3 - x) * (Ku3 - x) + (Kv3 - y) * (Kv3 - y);
            sumSquaredShiftNum += 2;
        }
        if (!(Ku > 2 && Kv > 2 && Ku < wl - 3 && Kv < hl - 3 && new_idepth > 0)) continue;
        float refColor = lpc_color[i];
        float hitColor[3];
        getInterpolatedElement33(dINewl, Ku, Kv, wl, hitColor);
        if (!std::isfinite((float)hitColor[0])) continue;
        float residual = hitColor[0] - (float)(affLL[0] * refColor + affLL[1]);
        float hw = fabs(residual) < huberTH ? 1 : huberTH / fabs(residual);
        if (fabs(residual) > cutoffTH) {
            E += maxEnergy; numTermsInE++; numSaturated++;
        } else {
            E += hw * residual * residual * (2 - hw);
            numTermsInE++;
            T.buf_warped_idepth[numTermsInWarped] = new_idepth;
            T.buf_warped_u[numTermsInWarped] = u;
            T.buf_warped_v[numTermsInWarped] = v;
            T.buf_warped_dx[numTermsInWarped] = hitColor[1];
            T.buf_warped_dy[numTermsInWarped] = hitColor[2];
            T.buf_warped_residual[numTermsInWarped] = residual;
            T.buf_warped_weight[numTermsInWarped] = hw;
            T.buf_warped_refColor[numTermsInWarped] = lpc_color[i];
            numTermsInWarped++;
        }
    }
    while (numTermsInWarped % 4 != 0) {   // :651-661
        T.buf_warped_idepth[numTermsInWarped] = 0; T.buf_warped_u[numTermsInWarped] = 0; T.buf_warped_v[numTermsInWarped] = 0;
        T.buf_warped_dx[numTermsInWarped] = 0; T.buf_warped_dy[numTermsInWarped] = 0; T.buf_warped_residual[numTermsInWarped] = 0;
        T.buf_warped_weight[numTermsInWarped] = 0; T.buf_warped_refColor[numTermsInWarped] = 0;
        numTermsInWarped++;
    }
    T.buf_warped_n = numTermsInWarped;
    rs[0] = E; rs[1] = numTermsInE;
    rs[2] = sumSquaredShiftT / (sumSquaredShiftNum + 0.1);
    rs[3] = 0;
    rs[4] = sumSquaredShiftRT / (sumSquaredShiftNum + 0.1);
    rs[5] = numSaturated / (float)numTermsInE;
}

// K4 — CoarseTracker::calcGSSSE (src/frontend/CoarseTracker.cc:675-741)
static void calcGSSSE(orc_ctx* c, Tracker& T, int lvl, double* H_out, double* b_out, float aff_a, float aff_b) {
    Accumulator9& acc = T.acc;
    acc.initialize();
    __m128 fxl = _mm_set1_ps(T.fx[lvl]), fyl = _mm_set1_ps(T.fy[lvl]);
    __m128 b0 = _mm_set1_ps(T.ref_aff_b);
    double affd[2];
    fromToVecExposure(T.ref_exposure, T.new_exposure, T.ref_aff_a, T.ref_aff_b, aff_a, aff_b, affd);
    __m128 a = _mm_set1_ps((float)affd[0]);
    __m128 one = _mm_set1_ps(1), minusOne = _mm_set1_ps(-1), zero = _mm_set1_ps(0);
    int n = T.buf_warped_n;
    for (int i = 0; i < n; i += 4) {
        __m128 dx = _mm_mul_ps(_mm_load_ps(T.buf_warped_dx + i), fxl);
        __m128 dy = _mm_mul_ps(_mm_load_ps(T.buf_warped_dy + i), fyl);
        __m128 u = _mm_load_ps(T.buf_warped_u + i);
        __m128 v = _mm_load_ps(T.buf_warped_v + i);
        __m128 id = _mm_load_ps(T.buf_warped_idepth + i);
        __m128 J[9];
        J[0] = _mm_mul_ps(id, dx);
        J[1] = _mm_mul_ps(id, dy);
        J[2] = _mm_sub_ps(zero, _mm_mul_ps(id, _mm_add_ps(_mm_mul_ps(u, dx), _mm_mul_ps(v, dy))));
        J[3] = _mm_sub_ps(zero, _mm_add_ps(_mm_mul_ps(_mm_mul_ps(u, v), dx), _mm_mul_ps(dy, _mm_add_ps(one, _mm_mul_ps(v, v)))));
        J[4] = _mm_add_ps(_mm_mul_ps(_mm_mul_ps(u, v), dy), _mm_mul_ps(dx, _mm_add_ps(one, _mm_mul_ps(u, u))));
        J[5] = _mm_sub_ps(_mm_mul_ps(u, dy), _mm_mul_ps(v, dx));
        J[6] = _mm_mul_ps(a, _mm_sub_ps(b0, _mm_load_ps(T.buf_warped_refColor + i)));
        J[7] = minusOne;
        J[8] = _mm_load_ps(T.buf_warped_residual + i);
        acc.updateSSE_eighted(J, _mm_load_ps(T.buf_warped_weight + i));
    }
    acc.finish();
    float invn = 1.0f / n;
    for (int r = 0; r < 8; r++) {
        for (int cc = 0; cc < 8; cc++) H_out[r * 8 + cc] = (double)acc.H[r][cc] * invn;
        b_out[r] = (double)acc.H[r][8] * invn;
    }
    const double sc[8] = {SCALE_XI_TRANS, SCALE_XI_TRANS, SCALE_XI_TRANS, SCALE_XI_ROT, SCALE_XI_ROT, SCALE_XI_ROT, SCALE_A, SCALE_B};
    for (int r = 0; r < 8; r++) {
        for (int cc = 0; cc < 8; cc++) H_out[r * 8 + cc] *= sc[r] * sc[cc];
        b_out[r] *= sc[r];
    }
    (void)c;
}

extern "C" int orc_tracker_calc_res(orc_ctx* c, int tracker, int new_slot, float new_exposure, int lvl, const double ref_to_new[12],
                                    float aff_a, float aff_b, float cutoff_th, float out6[6]) {
    Tracker& T = c->trk[tracker];
    T.new_slot = new_slot; T.new_exposure = new_exposure;
    double rs[6];
    calcRes(c, T, lvl, SE3::fromRt(ref_to_new), aff_a, aff_b, cutoff_th, rs);
    for (int i = 0; i < 6; i++) out6[i] = (float)rs[i];
    return 0;
}
extern "C" int orc_tracker_calc_gs(orc_ctx* c, int tracker, int lvl, const double ref_to_new[12], float aff_a, float aff_b,
                                   double H[64], double b[8]) {
    (void)ref_to_new;
    calcGSSSE(c, c->trk[tracker], lvl, H, b, aff_a, aff_b);
    return 0;
}

// 8x8 (or leading k x k) LM solve of CoarseTracker.cc:130-184 with the inertial terms identically zero
// (SURVEY.md §3.2).  solve_mode 0: Hl symmetric (upstream DSO); 1: as the fork wrote it — only the
// upper triangle is filled and Eigen's LDLT reads the lower one, i.e. diag(H)(1+lambda) (trap T4).
static void lm_solve(const double* H, const double* b, float lambda, int solve_mode, float modeA, float modeB, double* inc) {
    double Hl[64];
    for (int i = 0; i < 8; i++)
        for (int j = 0; j < 8; j++) {
            double v = H[i * 8 + j];
            if (solve_mode == 1 && j < i) v = 0;   // lower triangle never written
            Hl[i * 8 + j] = v;
        }
    for (int i = 0; i < 8; i++) Hl[i * 8 + i] *= (1 + lambda);
    double nb[8];
    for (int i = 0; i < 8; i++) nb[i] = -b[i];
    auto solve_k = [&](int k, const double* Hm, const double* rhs, double* out) {
        double Hs[64];
        for (int i = 0; i < k; i++)
            for (int j = 0; j < k; j++) Hs[i * k + j] = Hm[i * 8 + j];
        ldso_sm::ldlt_solve(k, Hs, rhs, out);
    };
    solve_k(8, Hl, nb, inc);
    if (modeA < 0 && modeB < 0) { solve_k(6, Hl, nb, inc); inc[6] = inc[7] = 0; }
    if (!(modeA < 0) && modeB < 0) { solve_k(7, Hl, nb, inc); inc[7] = 0; }
    if (modeA < 0 && !(modeB < 0)) {
        double Hs[64], bs[8];
        memcpy(Hs, Hl, sizeof(Hs));
        for (int i = 0; i < 8; i++) bs[i] = b[i];
        for (int i = 0; i < 8; i++) Hs[i * 8 + 6] = Hs[i * 8 + 7];
        for (int j = 0; j < 8; j++) Hs[6 * 8 + j] = Hs[7 * 8 + j];
        bs[6] = bs[7];
        double nbs[8], is[8];
        for (int i = 0; i < 8; i++) nbs[i] = -bs[i];
        solve_k(7, Hs, nbs, is);
        for (int i = 0; i < 8; i++) inc[i] = 0;
        for (int i = 0; i < 6; i++) inc[i] = is[i];
        inc[6] = 0; inc[7] = is[6];
    }
}

extern "C" int orc_tracker_track(orc_ctx* c, int tracker, int new_slot, float new_exposure, double pose_io[12], float aff_io[2],
                                 int coarsestLvl, const float minResForAbort[5], float lastResiduals[5], float lastFlow[3],
                                 int iters_out[2]);
// the two halves of the C ABI (ldso_gpu_tracker_track_begin / _end): on the CPU nothing runs ahead, begin keeps the arguments
namespace {
struct TrackPending { int tracker, new_slot, coarsest; float new_exposure, aff[2], min_res[5]; double pose[12]; };
std::map<orc_ctx*, TrackPending> g_track_pending;
std::mutex g_track_pending_mu;
}
extern "C" int orc_tracker_track_begin(orc_ctx* c, int tracker, int new_slot, float new_exposure, const double pose[12], const float aff[2],
                                       int coarsestLvl, const float minResForAbort[5]) {
    TrackPending P;
    P.tracker = tracker; P.new_slot = new_slot; P.coarsest = coarsestLvl; P.new_exposure = new_exposure;
    for (int i = 0; i < 12; i++) P.pose[i] = pose[i];
    for (int i = 0; i < 2; i++) P.aff[i] = aff[i];
    for (int i = 0; i < 5; i++) P.min_res[i] = minResForAbort[i];
    std::lock_guard<std::mutex> lk(g_track_pending_mu);
    g_track_pending[c] = P;
    return 0;
}
extern "C" int orc_tracker_track_end(orc_ctx* c, double pose_out[12], float aff_out[2], float lastResiduals[5], float lastFlow[3],
                                     int iters_out[2]) {
    TrackPending P;
    {
        std::lock_guard<std::mutex> lk(g_track_pending_mu);
        auto it = g_track_pending.find(c);
        if (it == g_track_pending.end()) return -2;
        P = it->second;
        g_track_pending.erase(it);
    }
    const int rc = orc_tracker_track(c, P.tracker, P.new_slot, P.new_exposure, P.pose, P.aff, P.coarsest, P.min_res, lastResiduals, lastFlow, iters_out);
    for (int i = 0; i < 12; i++) pose_out[i] = P.pose[i];
    for (int i = 0; i < 2; i++) aff_out[i] = P.aff[i];
    return rc;
}

// CoarseTracker::trackNewestCoarse (src/frontend/CoarseTracker.cc:61-310), setting_vi_enable=false
extern "C" int orc_tracker_track(orc_ctx* c, int tracker, int new_slot, float new_exposure, double pose_io[12], float aff_io[2],
                                 int coarsestLvl, const float minResForAbort[5], float lastResiduals[5], float lastFlow[3],
                                 int iters_out[2]) {
    Tracker& T = c->trk[tracker];
    T.new_slot = new_slot; T.new_exposure = new_exposure;
    if (!(coarsestLvl < 5 && coarsestLvl < c->levels)) return -1;
    for (int i = 0; i < 5; i++) lastResiduals[i] = NAN;
    for (int i = 0; i < 3; i++) lastFlow[i] = 1000;
    int maxIterations[] = {10, 20, 50, 50, 50};
    float lambdaExtrapolationLimit = 0.001;
    SE3 refToNew_current = SE3::fromRt(pose_io);
    float affc_a = aff_io[0], affc_b = aff_io[1];
    bool haveRepeated = false;
    int nRes = 0, nAcc = 0;
    const float modeA = c->cfg.affine_opt_mode_a, modeB = c->cfg.affine_opt_mode_b;
    const float cutoff0 = c->cfg.coarse_cutoff_th;
    for (int lvl = coarsestLvl; lvl >= 0; lvl--) {
        double H[64], b[8];
        float levelCutoffRepeat = 1;
        double resOld[6];
        calcRes(c, T, lvl, refToNew_current, affc_a, affc_b, cutoff0 * levelCutoffRepeat, resOld); nRes++;
        while (resOld[5] > 0.6 && levelCutoffRepeat < 50) {
            levelCutoffRepeat *= 2;
            calcRes(c, T, lvl, refToNew_current, affc_a, affc_b, cutoff0 * levelCutoffRepeat, resOld); nRes++;
        }
        float lambda = 0.01;
        calcGSSSE(c, T, lvl, H, b, affc_a, affc_b);
        for (int iteration = 0; iteration < maxIterations[lvl]; iteration++) {
            double inc[8];
            lm_solve(H, b, lambda, c->cfg.tracker_solve_mode, modeA, modeB, inc);
            float extrapFac = 1;
            if (lambda < lambdaExtrapolationLimit) extrapFac = sqrtf(sqrt(lambdaExtrapolationLimit / lambda));
            for (int i = 0; i < 8; i++) inc[i] *= extrapFac;
            double incScaled[8];
            for (int i = 0; i < 3; i++) incScaled[i] = inc[i] * SCALE_XI_TRANS;
            for (int i = 3; i < 6; i++) incScaled[i] = inc[i] * SCALE_XI_ROT;
            incScaled[6] = inc[6] * SCALE_A;
            incScaled[7] = inc[7] * SCALE_B;
            double sq = 0;
            for (int i = 0; i < 8; i++) sq += incScaled[i] * incScaled[i];
            if (!std::isfinite(sq)) {
                for (int i = 0; i < 8; i++) incScaled[i] = inc[i] = 0;
            }
            SE3 refToNew_new = SE3::exp(incScaled) * refToNew_current;
            float affn_a = affc_a, affn_b = affc_b;
            affn_a += incScaled[6];
            affn_b += incScaled[7];
            double resNew[6];
            calcRes(c, T, lvl, refToNew_new, affn_a, affn_b, cutoff0 * levelCutoffRepeat, resNew); nRes++;
            bool accept = (resNew[0] / resNew[1]) < (resOld[0] / resOld[1]);
            if (accept) {
                calcGSSSE(c, T, lvl, H, b, affn_a, affn_b);
                memcpy(resOld, resNew, sizeof(resOld));
                affc_a = affn_a; affc_b = affn_b;
                refToNew_current = refToNew_new;
                lambda *= 0.5;   // setting_coarse_tracker_lambda_decrease (src/Setting.cc:136-137)
                nAcc++;
            } else {
                lambda *= 4.0;   // setting_coarse_tracker_lambda_increase
                if (lambda < lambdaExtrapolationLimit) lambda = lambdaExtrapolationLimit;
            }
            double nrm = 0;
            for (int i = 0; i < 8; i++) nrm += inc[i] * inc[i];
            if (!(std::sqrt(nrm) > 1e-3)) break;
        }
        lastResiduals[lvl] = sqrtf((float)(resOld[0] / resOld[1]));
        lastFlow[0] = resOld[2]; lastFlow[1] = resOld[3]; lastFlow[2] = resOld[4];
        if (lastResiduals[lvl] > 1.5 * minResForAbort[lvl]) {
            if (iters_out) { iters_out[0] = nRes; iters_out[1] = nAcc; }
            return 1;
        }
        if (levelCutoffRepeat > 1 && !haveRepeated) { lvl++; haveRepeated = true; }
    }
    if (iters_out) { iters_out[0] = nRes; iters_out[1] = nAcc; }
    float out_a = affc_a, out_b = affc_b;
    if ((modeA != 0 && (fabsf(out_a) > 1.2)) || (modeB != 0 && (fabsf(out_b) > 200))) return 1;
    double rel[2];
    fromToVecExposure(T.ref_exposure, T.new_exposure, T.ref_aff_a, T.ref_aff_b, out_a, out_b, rel);
    float relAff0 = (float)rel[0], relAff1 = (float)rel[1];
    if ((modeA == 0 && (fabsf(logf(relAff0)) > 1.5)) || (modeB == 0 && (fabsf(relAff1) > 200))) return 1;
    // NB: the reference has already assigned lastToNew_out / aff_g2l_out before these checks
    // (CoarseTracker.cc:285-286); callers ignore them on `false`, and so does this ABI.
    if (modeA < 0) out_a = 0;
    if (modeB < 0) out_b = 0;
    refToNew_current.toRt(pose_io);
    aff_io[0] = out_a; aff_io[1] = out_b;
    return 0;
}

// =================================================================================================
// N4 — Undistort::undistort<T> (src/frontend/Undistort.cc:366-458) with PhotometricUndistorter::processFrame<T>
// (:190-230), in the reference's two passes: the whole raw image through the response / vignette first, then the remap.
namespace {
struct UndistortState {
    int w_org = 0, h_org = 0, g_depth = 0, mode = 0;
    bool has_G = false, has_vig = false, has_remap = false;
    std::vector<float> G, vig, remap_x, remap_y, image;
};
UndistortState* undist(orc_ctx* c) {
    if (!c->undistort) c->undistort = std::make_shared<UndistortState>();
    return (UndistortState*)c->undistort.get();
}
}  // namespace

extern "C" int orc_undistort_set(orc_ctx* c, int w_org, int h_org, int g_depth, const float* G, const float* vignette_inv,
                                 const float* remap_x, const float* remap_y, int photometric_mode) {
    if (w_org <= 0 || h_org <= 0 || ((remap_x == nullptr) != (remap_y == nullptr))) return -2;
    if (!remap_x && (w_org != c->cfg.w || h_org != c->cfg.h)) return -2;
    c->undistort = std::make_shared<UndistortState>();
    UndistortState* U = undist(c);
    U->w_org = w_org; U->h_org = h_org; U->g_depth = g_depth; U->mode = photometric_mode;
    const size_t no = (size_t)w_org * h_org, n = (size_t)c->cfg.w * c->cfg.h;
    if (G) { U->G.assign(G, G + g_depth); U->has_G = true; }
    if (vignette_inv) { U->vig.assign(vignette_inv, vignette_inv + no); U->has_vig = true; }
    if (remap_x) { U->remap_x.assign(remap_x, remap_x + n); U->remap_y.assign(remap_y, remap_y + n); U->has_remap = true; }
    return 0;
}

extern "C" int orc_frame_make_images_raw(orc_ctx* c, int slot, const void* raw, int bytes_per_pixel, float exposure_time, float factor) {
    auto U = undist(c);
    if (U->w_org == 0 || !raw || (bytes_per_pixel != 1 && bytes_per_pixel != 2)) return -2;
    const int wOrg = U->w_org, hOrg = U->h_org, w = c->cfg.w, h = c->cfg.h;
    const int wh = wOrg * hOrg;
    std::vector<float> data(wh);
    auto in = [&](int i) -> unsigned { return bytes_per_pixel == 1 ? ((const unsigned char*)raw)[i] : ((const unsigned short*)raw)[i]; };
    if (!U->has_G || exposure_time <= 0 || U->mode == 0) {                       // processFrame :197-204
        for (int i = 0; i < wh; i++) data[i] = factor * (float)in(i);
    } else {
        for (int i = 0; i < wh; i++) data[i] = U->G[in(i)];
        if (U->mode == 2 && U->has_vig)
            for (int i = 0; i < wh; i++) data[i] *= U->vig[i];
    }
    U->image.assign((size_t)w * h, 0.f);
    if (U->has_remap) {
        for (int idx = w * h - 1; idx >= 0; idx--) {                             // undistort :400-447
            float xx = U->remap_x[idx], yy = U->remap_y[idx];
            if (xx < 0) { U->image[idx] = 0; continue; }
            int xxi = xx, yyi = yy;
            xx -= xxi; yy -= yyi;
            float xxyy = xx * yy;
            const float* src = data.data() + xxi + yyi * wOrg;
            U->image[idx] = xxyy * src[1 + wOrg] + (yy - xxyy) * src[wOrg] + (xx - xxyy) * src[1] + (1 - xx - yy + xxyy) * src[0];
        }
    } else {
        memcpy(U->image.data(), data.data(), sizeof(float) * (size_t)w * h);
    }
    return orc_frame_make_images(c, slot, U->image.data());
}

extern "C" int orc_undistort_get_image(orc_ctx* c, float* image_out) {
    auto U = undist(c);
    if (U->image.empty()) return -2;
    memcpy(image_out, U->image.data(), sizeof(float) * U->image.size());
    return 0;
}

// ldso_gpu_frame_make_images_async: on the CPU nothing is queued, the image is consumed at once
extern "C" int orc_frame_make_images(orc_ctx* c, int slot, const float* img);
extern "C" int orc_frame_make_images_async(orc_ctx* c, int slot, const void* image, int kind, float exposure_time) {
    if (kind == 0 || kind == 1) return orc_frame_make_images(c, slot, (const float*)image);
    return orc_frame_make_images_raw(c, slot, image, kind == 2 ? 1 : 2, exposure_time, 1.0f);
}

// ldso_gpu_set_wait_hook: the CPU path never waits for a device
extern "C" int orc_set_wait_hook(void (*)(void*), void*) { return 0; }
