// ldso_system.hpp — the host side of LDSO's per-frame pipeline in C++, over the C ABI of include/ldso_gpu.h.
//
// ldso::FullSystemT<Api> restates the control flow of the reference's ldso::FullSystem (src/frontend/FullSystem.cc) and
// of the host half of ldso::internal::EnergyFunctional (src/internal/OptimizationBackend/EnergyFunctional.cc) with the
// reference's member and type names (FrameHessian, PointHessian, PointFrameResidual, FrameFramePrecalc, CalibHessian,
// AffLight).  Every loop over pixels, points or residuals is a call into the ABI (CUDA kernels in libldso_gpu.so); what
// stays here is what is sequential by nature: motion hypotheses, keyframe decision, window bookkeeping, FEJ state
// algebra, the (8n+4)^2 marginalisation prior.  `Api` binds the entry points by symbol prefix, so the SAME control flow
// is compiled once into libldso_gpu.so (ldso_gpu_system_*, the product) and once into the CPU oracle (orc_system_*, test
// infrastructure): trajectories of the two can be compared frame by frame.
//
// Scope: vision only (setting_vi_enable = false: every inertial term vanishes), linearizeOperation =
// true, loop closing off, setting_solverMode = SOLVER_FIX_LAMBDA | SOLVER_ORTHOGONALIZE_X_LATER (Setting.cc:23),
// setting_forceAceptStep = true (:73).  Fork behaviour is kept as written where it differs from upstream DSO
// (DESIGN.md section 2): SE(3)-composed state differences (T9), PointFrameResidual::isNew never
// cleared, EnergyFunctional::nPoints never incremented (npoints_counter_mode), FrameHessian::aff_g2l() of an untracked
// frame being zero in the keyframe test.
#pragma once
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdlib>
#include <cstdio>
#include <cstring>
#include <condition_variable>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include <ucontext.h>

#include "ldso_gpu.h"
#include "ldso_smallmat.h"
#include "ldso_system.h"

// ---- binding of the C ABI by symbol prefix ---------------------------------------------------------------------------
#ifdef LDSO_API_TIMING   // measurement build: seconds spent inside the C ABI, so that (wall - this) is the orchestrator's own host work
#include <chrono>
namespace ldso { inline double& api_seconds() { static thread_local double s = 0; return s; }
struct ApiTimer { std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
                  ~ApiTimer() { api_seconds() += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count(); } }; }
#define LDSO_API_FN(P, name) \
    template <class... A> static inline auto name(A... a) -> decltype(P##name(a...)) { ldso::ApiTimer ldso_api_timer; return P##name(a...); }
#else
#define LDSO_API_FN(P, name) \
    template <class... A> static inline auto name(A... a) -> decltype(P##name(a...)) { return P##name(a...); }
#endif
#define LDSO_DEFINE_API(NAME, P, CTX)                                                                                   \
    struct NAME {                                                                                                       \
        typedef CTX ctx_t;                                                                                              \
        LDSO_API_FN(P, default_config) LDSO_API_FN(P, create) LDSO_API_FN(P, destroy) LDSO_API_FN(P, last_error)         \
        LDSO_API_FN(P, pyr_levels) LDSO_API_FN(P, level_size) LDSO_API_FN(P, sync) LDSO_API_FN(P, bind_thread) LDSO_API_FN(P, set_wait_hook)                                    \
        LDSO_API_FN(P, frame_make_images) LDSO_API_FN(P, frame_make_images_dev) LDSO_API_FN(P, frame_make_images_raw)    \
        LDSO_API_FN(P, frame_make_images_async)                                                                         \
        LDSO_API_FN(P, tracker_make_k) LDSO_API_FN(P, tracker_set_ref) LDSO_API_FN(P, tracker_track)                     \
        LDSO_API_FN(P, tracker_track_begin) LDSO_API_FN(P, tracker_track_end)                                            \
        LDSO_API_FN(P, select_points) LDSO_API_FN(P, detect_corners)                                                     \
        LDSO_API_FN(P, immature_create) LDSO_API_FN(P, immature_append) LDSO_API_FN(P, immature_set_types)               \
        LDSO_API_FN(P, immature_compact) LDSO_API_FN(P, immature_trace_async) LDSO_API_FN(P, immature_trace_counts)      \
        LDSO_API_FN(P, immature_activate) LDSO_API_FN(P, distmap_make) LDSO_API_FN(P, immature_select_activation)        \
        LDSO_API_FN(P, init_set_first) LDSO_API_FN(P, init_track_frame) LDSO_API_FN(P, init_get_points)                  \
        LDSO_API_FN(P, window_set_frames) LDSO_API_FN(P, window_set_points) LDSO_API_FN(P, window_set_residuals)         \
        LDSO_API_FN(P, window_set_adjoints) LDSO_API_FN(P, window_set_idepth) LDSO_API_FN(P, window_linearize)           \
        LDSO_API_FN(P, window_apply_res) LDSO_API_FN(P, window_fix_linearization) LDSO_API_FN(P, window_accumulate)      \
        LDSO_API_FN(P, window_solve) LDSO_API_FN(P, window_resubstitute) LDSO_API_FN(P, window_marginalize_points)       \
        LDSO_API_FN(P, window_marginalize_frame) LDSO_API_FN(P, window_get_residual_floats)                              \
        LDSO_API_FN(P, window_get_residual_ints) LDSO_API_FN(P, window_get_point_floats)                                 \
        LDSO_API_FN(P, window_get_keyframe_results) LDSO_API_FN(P, window_marginalize_subset)                            \
    };
// optional fine-grained host timers of the orchestrator (LDSO_SYS_PROF=1, FullSystemT::prof): label -> seconds
#ifdef LDSO_API_TIMING
#define LDSO_T(label, expr) do { const double _t0 = now(), _a0 = ldso::api_seconds(); expr; prof.add(label, now() - _t0, ldso::api_seconds() - _a0); } while (0)
#else
#define LDSO_T(label, expr) do { const double _t0 = now(); expr; prof.add(label, now() - _t0); } while (0)
#endif
// declarations of the same entry points under another prefix (test builds: the CPU oracle mirrors the ABI as orc_*)
#define LDSO_DECLARE_API(P, CTX)                                                                                         \
    extern "C" {                                                                                                        \
    void P##default_config(ldso_gpu_config*);                                                                            \
    int P##create(const ldso_gpu_config*, CTX**);                                                                        \
    void P##destroy(CTX*);                                                                                               \
    const char* P##last_error(const CTX*);                                                                               \
    int P##pyr_levels(const CTX*);                                                                                       \
    int P##level_size(const CTX*, int, int*, int*);                                                                      \
    int P##sync(CTX*);                                                                                                   \
    int P##bind_thread(CTX*);                                                                                            \
    int P##set_wait_hook(void (*)(void*), void*);                                                                        \
    int P##frame_make_images(CTX*, int, const float*);                                                                   \
    int P##frame_make_images_dev(CTX*, int, const float*);                                                               \
    int P##frame_make_images_raw(CTX*, int, const void*, int, float, float);                                             \
    int P##frame_make_images_async(CTX*, int, const void*, int, float);                                                  \
    int P##tracker_make_k(CTX*, int, float, float, float, float);                                                        \
    int P##tracker_set_ref(CTX*, int, int, int, const float*, const float*, const float*, const float*, float, float, float); \
    int P##tracker_track(CTX*, int, int, float, double*, float*, int, const float*, float*, float*, int*);               \
    int P##tracker_track_begin(CTX*, int, int, float, const double*, const float*, int, const float*);                   \
    int P##tracker_track_end(CTX*, double*, float*, float*, float*, int*);                                               \
    int P##select_points(CTX*, int, float, int, float, int*, int, int, float*, float*, float*, int*);                    \
    int P##detect_corners(CTX*, int, int, int, float*, float*, float*, unsigned char*, int*, int*);                      \
    int P##immature_create(CTX*, int, int, const float*, const float*, float*, float*, float*, float*);                  \
    int P##immature_append(CTX*, int, int, int, const float*, const float*, int*);                                       \
    int P##immature_set_types(CTX*, int, int, const float*);                                                             \
    int P##immature_compact(CTX*, int, int, int, const int*, int*);                                                      \
    int P##immature_trace_async(CTX*, int, int, const float*, const float*, const float*);                               \
    int P##immature_trace_counts(CTX*, int*);                                                                            \
    int P##immature_activate(CTX*, int, const int*, int, float*, int*, int*);                                            \
    int P##distmap_make(CTX*, int, const float*, const float*, int, const int*, const float*, const float*, const float*); \
    int P##immature_select_activation(CTX*, int, const float*, const float*, const unsigned char*, const unsigned char*, \
                                      float, float, unsigned char*, int*);                                               \
    int P##init_set_first(CTX*, int, int*);                                                                              \
    int P##init_track_frame(CTX*, int, float, float, double*, float*, int*, int*, float*, int*);                         \
    int P##init_get_points(CTX*, int, int*, float*, float*, float*, float*, unsigned char*, float*, float*, float*);     \
    int P##window_set_frames(CTX*, int, const int*, const float*, const float*, const float*);                           \
    int P##window_set_points(CTX*, int, const float*, const float*, const float*, const float*, const int*, const float*, \
                             const float*, const float*, const int*);                                                    \
    int P##window_set_residuals(CTX*, int, const int*, const int*, const unsigned char*, const float*);                  \
    int P##window_set_adjoints(CTX*, const double*, const double*, const float*, const float*, const double*, const double*, \
                               const double*);                                                                           \
    int P##window_set_idepth(CTX*, const float*, const float*);                                                          \
    int P##window_linearize(CTX*, int, double*, float*);                                                                 \
    int P##window_apply_res(CTX*);                                                                                       \
    int P##window_fix_linearization(CTX*, const unsigned char*);                                                         \
    int P##window_accumulate(CTX*, double*, double*, double*, double*, double*, double*, int*);                          \
    int P##window_solve(CTX*, int, double, const double*, const double*, const double*, const double*, const double*,     \
                        const double*, const double*, const double*, const double*, const double*, int, double*);        \
    int P##window_resubstitute(CTX*, const double*, float, float*);                                                      \
    int P##window_marginalize_points(CTX*, const unsigned char*, float, double*, double*, int*);                         \
    int P##window_marginalize_frame(CTX*, int, int, const double*, const double*, const double*, const double*, double*, double*); \
    int P##window_get_residual_floats(CTX*, int, float*);                                                                \
    int P##window_get_residual_ints(CTX*, int, int*);                                                                    \
    int P##window_get_point_floats(CTX*, int, float*);                                                                   \
    int P##window_get_keyframe_results(CTX*, int*, float*, float*, float*, float*, float*);                              \
    int P##window_marginalize_subset(CTX*, const unsigned char*, const unsigned char*, const float*, float, double*, double*, int*, \
                                     int*);                                                                              \
    }

namespace ldso {

using SE3 = ldso_sm::SE3;

// include/Settings.h:28-52
constexpr float SCALE_IDEPTH = 1.0f, SCALE_XI_ROT = 1.0f, SCALE_XI_TRANS = 0.5f, SCALE_F = 50.0f, SCALE_C = 50.0f;
constexpr float SCALE_A = 10.0f, SCALE_B = 1000.0f;
constexpr float SCALE_XI_ROT_INVERSE = 1.0f / SCALE_XI_ROT, SCALE_XI_TRANS_INVERSE = 1.0f / SCALE_XI_TRANS;
constexpr float SCALE_F_INVERSE = 1.0f / SCALE_F, SCALE_C_INVERSE = 1.0f / SCALE_C;
constexpr float SCALE_A_INVERSE = 1.0f / SCALE_A, SCALE_B_INVERSE = 1.0f / SCALE_B;
constexpr int CPARS = 4;
constexpr int patternNum = 8, patternPadding = 2;

// glibc's srand() / rand() (TYPE_3 additive feedback generator, r[i] = r[i-3] + r[i-31], output >> 1) with private state.
// FullSystem::initializeFromInitializer sub-selects the initializer's points with std::srand(549812361) / rand()
// (FullSystem.cc:1486, :1516); the process-global generator would interleave between sequences stepped concurrently.
struct GlibcRand {
    explicit GlibcRand(unsigned seed) {
        int w[31];
        w[0] = seed ? (int)seed : 1;
        for (int i = 1; i < 31; i++) {
            long long hi = w[i - 1] / 127773, lo = w[i - 1] % 127773, v = 16807 * lo - 2836 * hi;
            if (v < 0) v += 2147483647;
            w[i] = (int)v;
        }
        for (int i = 0; i < 31; i++) ring[i] = (unsigned)w[i];
        fi = 3; ri = 0;
        for (int i = 0; i < 310; i++) next();
    }
    int next() {
        ring[fi] += ring[ri];
        const int out = (int)(ring[fi] >> 1);
        if (++fi >= 31) fi = 0;
        if (++ri >= 31) ri = 0;
        return out;
    }
private:
    unsigned ring[31];
    int fi, ri;
};

struct AffLight {   // include/AffLight.h
    float a = 0, b = 0;
    AffLight() {}
    AffLight(float a_, float b_) : a(a_), b(b_) {}
    static void fromToVecExposure(float exposureF, float exposureT, AffLight g2F, AffLight g2T, double out[2]) {   // :27-35
        if (exposureF == 0 || exposureT == 0) exposureT = exposureF = 1;
        float a = std::exp(g2T.a - g2F.a) * exposureT / exposureF;
        float b = g2T.b - a * g2F.b;
        out[0] = a;
        out[1] = b;
    }
};

enum ResState { IN = 0, OOB = 1, OUTLIER = 2 };   // include/internal/Residuals.h:33-35
enum class PointStatus { ACTIVE = 0, OUTLIER, OUT, MARGINALIZED };   // include/Point.h

namespace internal {

struct FrameHessian;
struct PointHessian;

// include/internal/CalibHessian.h (value, scaled value, zero point; the gamma LUT lives in the ctx: ldso_gpu_set_gamma)
struct CalibHessian {
    double value[4], value_zero[4], value_scaled[4], value_minus_value_zero[4], value_backup[4], step[4];
    float value_scaledf[4], value_scaledi[4];
    void init(float fx, float fy, float cx, float cy) {
        double v[4] = {fx, fy, cx, cy};
        for (int i = 0; i < 4; i++) value_zero[i] = 0;
        setValueScaled(v);
        for (int i = 0; i < 4; i++) { value_zero[i] = value[i]; value_minus_value_zero[i] = 0; step[i] = 0; value_backup[i] = value[i]; }
    }
    float fxl() const { return value_scaledf[0]; }
    float fyl() const { return value_scaledf[1]; }
    float cxl() const { return value_scaledf[2]; }
    float cyl() const { return value_scaledf[3]; }
    void fill_inverse() {
        for (int i = 0; i < 4; i++) value_scaledf[i] = (float)value_scaled[i];
        value_scaledi[0] = 1.0f / value_scaledf[0];
        value_scaledi[1] = 1.0f / value_scaledf[1];
        value_scaledi[2] = -value_scaledf[2] / value_scaledf[0];
        value_scaledi[3] = -value_scaledf[3] / value_scaledf[1];
    }
    void setValue(const double* v) {   // :69-83
        for (int i = 0; i < 4; i++) value[i] = v[i];
        value_scaled[0] = SCALE_F * value[0]; value_scaled[1] = SCALE_F * value[1];
        value_scaled[2] = SCALE_C * value[2]; value_scaled[3] = SCALE_C * value[3];
        fill_inverse();
        for (int i = 0; i < 4; i++) value_minus_value_zero[i] = value[i] - value_zero[i];
    }
    void setValueScaled(const double* vs) {   // :85-98
        for (int i = 0; i < 4; i++) value_scaled[i] = vs[i];
        value[0] = SCALE_F_INVERSE * value_scaled[0]; value[1] = SCALE_F_INVERSE * value_scaled[1];
        value[2] = SCALE_C_INVERSE * value_scaled[2]; value[3] = SCALE_C_INVERSE * value_scaled[3];
        for (int i = 0; i < 4; i++) value_minus_value_zero[i] = value[i] - value_zero[i];
        fill_inverse();
    }
};

// include/internal/FrameFramePrecalc.h: one (host, target) pair, float
struct FrameFramePrecalc {
    float PRE_RTll_0[9], PRE_tTll_0[3], PRE_KRKiTll[9], PRE_RKiTll[9], PRE_KtTll[3], PRE_RTll[9], PRE_tTll[3];
    float PRE_aff_mode[2], PRE_b0_mode, distanceLL;
    void Set(const FrameHessian& host, const FrameHessian& target, const CalibHessian& HCalib);   // FrameFramePrecalc.cc:6-35
    void pack(float* o) const {   // the 48-float record of include/ldso_gpu.h (LDSO_PRECALC_STRIDE)
        for (int i = 0; i < LDSO_PRECALC_STRIDE; i++) o[i] = 0;
        for (int i = 0; i < 9; i++) { o[i] = PRE_RTll_0[i]; o[12 + i] = PRE_KRKiTll[i]; o[27 + i] = PRE_RTll[i]; }
        for (int i = 0; i < 3; i++) { o[9 + i] = PRE_tTll_0[i]; o[21 + i] = PRE_KtTll[i]; o[36 + i] = PRE_tTll[i]; }
        o[24] = PRE_aff_mode[0]; o[25] = PRE_aff_mode[1]; o[26] = PRE_b0_mode; o[39] = distanceLL;
    }
};

// include/internal/Residuals.h:42-131.  J (RawResidualJacobian), res_toZeroF and JpJdF stay on the device
// (ldso_gpu_window_get_residual_floats mirrors them when other host code wants them).
struct PointFrameResidual {
    FrameHessian* host = nullptr;
    FrameHessian* target = nullptr;
    ResState state_state = ResState::OUTLIER;
    double state_energy = 0;
    ResState state_NewState = ResState::OUTLIER;
    double state_NewEnergy = 0;
    bool isNew = true;
    bool isLinearized = false;
    bool isActiveAndIsGoodNEW = false;
    float centerProjectedTo[3] = {0, 0, 0};
    int hostIDX = 0, targetIDX = 0;
    bool isActive() const { return isActiveAndIsGoodNEW; }
    void resetOOB() {   // :63-67
        state_NewEnergy = state_energy = 0;
        state_NewState = ResState::OUTLIER;
        state_state = ResState::IN;
    }
};

// include/internal/PointHessian.h:19-132
struct PointHessian {
    FrameHessian* host = nullptr;
    int feat_idx = 0;                // position among the host frame's features (Frame::features order = allPoints order)
    int upload_idx = -1;             // row of the point in the device-resident window (set by the last upload)
    PointStatus status = PointStatus::ACTIVE;   // Point::status
    float u = 0, v = 0, my_type = 0;
    bool hasDepthPrior = false;
    float idepth_scaled = 0, idepth_zero_scaled = 0, idepth_zero = 0, idepth = 0, step = 0, idepth_backup = 0;
    float nullspaces_scale = 0, idepth_hessian = 0, maxRelBaseline = 0;
    int numGoodResiduals = 0;
    std::vector<PointFrameResidual> residuals;
    // lastResiduals[k].first is identified by its target frame (one residual per (point, target)); nullptr = none
    FrameHessian* lastResiduals_first[2] = {nullptr, nullptr};
    ResState lastResiduals_second[2] = {ResState::OOB, ResState::OOB};
    float color[patternNum], weights[patternNum];
    float priorF = 0, deltaF = 0, HdiF = 0;
    void setIdepth(float id) { idepth = id; idepth_scaled = SCALE_IDEPTH * id; }
    void setIdepthScaled(float ids) { idepth = (1.0f / SCALE_IDEPTH) * ids; idepth_scaled = ids; }
    void setIdepthZero(float id) {
        idepth_zero = id;
        idepth_zero_scaled = SCALE_IDEPTH * id;
        nullspaces_scale = (float)(-(id * 1.001 - id / 1.001) * 500);
    }
    void takeData(float idepthFixPrior) {   // :109-114
        priorF = hasDepthPrior ? idepthFixPrior * SCALE_IDEPTH * SCALE_IDEPTH : 0;
        deltaF = idepth - idepth_zero;
    }
    PointFrameResidual* find(const FrameHessian* target) {
        for (auto& r : residuals) if (r.target == target) return &r;
        return nullptr;
    }
    // :53-73
    bool isOOB(const std::vector<FrameHessian*>& toMarg, int minGoodActiveResForMarg, int minGoodResForMarg) const {
        int visInToMarg = 0;
        for (const auto& r : residuals) {
            if (r.state_state != ResState::IN) continue;
            for (const FrameHessian* k : toMarg) if (r.target == k) visInToMarg++;
        }
        if ((int)residuals.size() >= minGoodActiveResForMarg && numGoodResiduals > minGoodResForMarg + 10 &&
            (int)residuals.size() - visInToMarg < minGoodActiveResForMarg) return true;
        if (lastResiduals_second[0] == ResState::OOB) return true;
        if (residuals.size() < 2) return false;
        if (lastResiduals_second[0] == ResState::OUTLIER && lastResiduals_second[1] == ResState::OUTLIER) return true;
        return false;
    }
    bool isInlierNew(int minGoodActiveResForMarg, int minGoodResForMarg) const {   // :75-78
        return (int)residuals.size() >= minGoodActiveResForMarg && numGoodResiduals >= minGoodResForMarg;
    }
};

// an immature feature of a keyframe: the ImmaturePoint itself lives in the device pool (K5 / K6), the host keeps the
// static fields a PointHessian is built from (PointHessian::PointHessian(ImmaturePoint), src/internal/PointHessian.cc:9-23)
struct ImmatureFeature {
    float u, v, my_type;
    int feat_idx;
    float color[patternNum], weights[patternNum];
};

// include/internal/FrameHessian.h + the shell fields of include/Frame.h that the path reads
struct FrameHessian {
    // ---- Frame (shell) ----
    int id = 0;                 // Frame::id: running number over all frames
    int kfId = -1;              // Frame::kfId = FrameHessian::frameID
    double timeStamp = 0;
    SE3 Tcw;                    // Frame::getPose(): worldToCam
    AffLight shell_aff_g2l;     // Frame::aff_g2l
    bool poseValid = true;
    FrameHessian* trackingRef = nullptr;   // the keyframe this frame was tracked against and the relative pose found then
    SE3 camToTrackingRef;                   // (upstream DSO's FrameShell::trackingRef / camToTrackingRef)
    // ---- FrameHessian ----
    int frameID = 0, idx = 0, slot = -1;
    float ab_exposure = 0, frameEnergyTH = 8 * 8 * patternNum;
    bool flaggedForMarginalization = false;
    double nullspaces_pose[6][6] = {{0}}, nullspaces_scale[6] = {0}, nullspaces_affine[4][2] = {{0}};   // [row][col]
    SE3 worldToCam_evalPT, PRE_worldToCam, PRE_camToWorld;
    double state[10] = {0}, step[10] = {0}, state_backup[10] = {0}, state_zero[10] = {0}, state_scaled[10] = {0};
    double prior[8] = {0}, delta_prior[8] = {0}, delta[8] = {0};
    std::vector<FrameFramePrecalc> targetPrecalc;
    // ---- Frame::features, split by status ----
    std::vector<ImmatureFeature> immature;                   // FeatureStatus::IMMATURE, in feature order = pool order
    std::vector<std::unique_ptr<PointHessian>> points;       // VALID features whose point is ACTIVE, in feature order
    int n_out = 0;                                           // every other feature (OUTLIER, or point no longer ACTIVE)
    int n_features = 0;

    AffLight aff_g2l() const { return AffLight((float)state_scaled[6], (float)state_scaled[7]); }
    AffLight aff_g2l_0() const { return AffLight((float)(state_zero[6] * SCALE_A), (float)(state_zero[7] * SCALE_B)); }
    void refreshPRE() {
        PRE_worldToCam = SE3::exp(state_scaled) * worldToCam_evalPT;   // w2c_leftEps() = state_scaled.head<6>()
        PRE_camToWorld = PRE_worldToCam.inverse();
    }
    void setState(const double* s) {   // FrameHessian.h:87-101
        for (int i = 0; i < 10; i++) state[i] = s[i];
        for (int i = 0; i < 3; i++) { state_scaled[i] = SCALE_XI_TRANS * state[i]; state_scaled[3 + i] = SCALE_XI_ROT * state[3 + i]; }
        state_scaled[6] = SCALE_A * state[6]; state_scaled[7] = SCALE_B * state[7];
        state_scaled[8] = SCALE_A * state[8]; state_scaled[9] = SCALE_B * state[9];
        refreshPRE();
    }
    void setStateScaled(const double* ss) {   // :103-117
        for (int i = 0; i < 10; i++) state_scaled[i] = ss[i];
        for (int i = 0; i < 3; i++) { state[i] = SCALE_XI_TRANS_INVERSE * state_scaled[i]; state[3 + i] = SCALE_XI_ROT_INVERSE * state_scaled[3 + i]; }
        state[6] = SCALE_A_INVERSE * state_scaled[6]; state[7] = SCALE_B_INVERSE * state_scaled[7];
        state[8] = SCALE_A_INVERSE * state_scaled[8]; state[9] = SCALE_B_INVERSE * state_scaled[9];
        refreshPRE();
    }
    void setStateZero(const double* sz) {   // src/internal/FrameHessian.cc:27-57
        for (int i = 0; i < 10; i++) state_zero[i] = sz[i];
        const SE3 evalInv = worldToCam_evalPT.inverse();
        for (int i = 0; i < 6; i++) {
            double eps[6] = {0, 0, 0, 0, 0, 0}, meps[6] = {0, 0, 0, 0, 0, 0}, lp[6], lm[6];
            eps[i] = 1e-3; meps[i] = -1e-3;
            SE3 P = (worldToCam_evalPT * SE3::exp(eps)) * evalInv, M = (worldToCam_evalPT * SE3::exp(meps)) * evalInv;
            P.log(lp); M.log(lm);
            for (int k = 0; k < 6; k++) nullspaces_pose[k][i] = (lp[k] - lm[k]) / (2e-3);
        }
        SE3 P = worldToCam_evalPT, M = worldToCam_evalPT;
        for (int k = 0; k < 3; k++) { P.t[k] *= 1.00001; M.t[k] /= 1.00001; }
        P = P * evalInv; M = M * evalInv;
        double lp[6], lm[6];
        P.log(lp); M.log(lm);
        for (int k = 0; k < 6; k++) nullspaces_scale[k] = (lp[k] - lm[k]) / (2e-3);
        for (int r = 0; r < 4; r++) nullspaces_affine[r][0] = nullspaces_affine[r][1] = 0;
        nullspaces_affine[0][0] = 1; nullspaces_affine[1][0] = 0;
        nullspaces_affine[0][1] = 0; nullspaces_affine[1][1] = std::exp(aff_g2l_0().a) * ab_exposure;   // expf(float) * float
    }
    void setEvalPT(const SE3& w2c_evalPT, const double* s) {   // FrameHessian.h:119-124
        worldToCam_evalPT = w2c_evalPT;
        setState(s);
        setStateZero(s);
    }
    void setEvalPT_scaled(const SE3& w2c_evalPT, const AffLight& aff) {   // :127-134
        double init[10] = {0, 0, 0, 0, 0, 0, aff.a, aff.b, 0, 0};
        worldToCam_evalPT = w2c_evalPT;
        setStateScaled(init);
        setStateZero(state);
    }
    void get_state_minus_stateZero(double* res) const {   // :59-64 (fork: SE(3)-composed, SURVEY.md trap T9)
        for (int i = 0; i < 10; i++) res[i] = state[i] - state_zero[i];
        double nz[6];
        for (int i = 0; i < 6; i++) nz[i] = -state_zero[i];
        (SE3::exp(nz) * SE3::exp(state)).log(res);
    }
    void get_state_minus_statePriorZero(double* res) const {   // :66-71 (getPriorZero() = 0)
        for (int i = 0; i < 10; i++) res[i] = state[i];
        double nz[6] = {0, 0, 0, 0, 0, 0};
        (SE3::exp(nz) * SE3::exp(state)).log(res);
    }
};

inline void FrameFramePrecalc::Set(const FrameHessian& host, const FrameHessian& target, const CalibHessian& HCalib) {
    double R[9];
    SE3 leftToLeft_0 = target.worldToCam_evalPT * host.worldToCam_evalPT.inverse();
    leftToLeft_0.rotationMatrix(R);
    for (int i = 0; i < 9; i++) PRE_RTll_0[i] = (float)R[i];
    for (int i = 0; i < 3; i++) PRE_tTll_0[i] = (float)leftToLeft_0.t[i];
    SE3 leftToLeft = target.PRE_worldToCam * host.PRE_camToWorld;
    leftToLeft.rotationMatrix(R);
    for (int i = 0; i < 9; i++) PRE_RTll[i] = (float)R[i];
    for (int i = 0; i < 3; i++) PRE_tTll[i] = (float)leftToLeft.t[i];
    distanceLL = (float)std::sqrt(leftToLeft.t[0] * leftToLeft.t[0] + leftToLeft.t[1] * leftToLeft.t[1] + leftToLeft.t[2] * leftToLeft.t[2]);
    float K[9] = {HCalib.fxl(), 0, HCalib.cxl(), 0, HCalib.fyl(), HCalib.cyl(), 0, 0, 1}, Ki[9], KR[9];
    ldso_sm::mat3_inv<float>(K, Ki);
    ldso_sm::mat3_mul<float>(K, PRE_RTll, KR);
    ldso_sm::mat3_mul<float>(KR, Ki, PRE_KRKiTll);
    ldso_sm::mat3_mul<float>(PRE_RTll, Ki, PRE_RKiTll);
    ldso_sm::mat3_vec<float>(K, PRE_tTll, PRE_KtTll);
    double ab[2];
    AffLight::fromToVecExposure(host.ab_exposure, target.ab_exposure, host.aff_g2l(), target.aff_g2l(), ab);
    PRE_aff_mode[0] = (float)ab[0]; PRE_aff_mode[1] = (float)ab[1];
    PRE_b0_mode = host.aff_g2l_0().b;
}

}  // namespace internal

// =========================================================================================================================
template <class Api> class FullSystemT {
public:
    typedef typename Api::ctx_t ctx_t;
    typedef internal::FrameHessian FrameHessian;
    typedef internal::PointHessian PointHessian;
    typedef internal::PointFrameResidual PointFrameResidual;

    ldso_system_config cfg;
    ctx_t* ctx = nullptr;
    std::string err;

    // ---- FullSystem members (include/frontend/FullSystem.h) ----
    bool initialized = false, isLost = false, initFailed = false;
    std::vector<std::unique_ptr<FrameHessian>> allFrameHistory;   // Frame + FrameHessian of every frame fed
    std::vector<FrameHessian*> frames;                            // the sliding window, oldest first
    internal::CalibHessian Hcalib;
    float currentMinActDist = 2;
    double lastCoarseRMSE[5] = {100, 100, 100, 100, 100};
    int numKFs = 0;                                               // globalMap->NumFrames()
    // the two CoarseTrackers are device slots 0 / 1 of the ctx; the host keeps what FullSystem reads from them
    struct TrackerHost {
        int dev = 0, refFrameID = -1;
        FrameHessian* lastRef = nullptr;
        AffLight lastRef_aff_g2l;
        float firstCoarseRMSE = -1;
        float lastResiduals[5] = {0, 0, 0, 0, 0}, lastFlowIndicators[3] = {0, 0, 0};
    } trackerA, trackerB;
    TrackerHost *coarseTracker = &trackerA, *coarseTracker_forNewKF = &trackerB;
    // CoarseInitializer
    int initializerFrameID = -1;
    FrameHessian* initializerFirstFrame = nullptr;
    // PixelSelector
    int currentPotential = 3;
    // ---- EnergyFunctional members ----
    std::vector<double> HM, bM;                                   // (8n+4)^2 row-major, 8n+4
    int nPoints = 0, nResiduals = 0, resInA = 0, resInM = 0;
    std::vector<double> adHost, adTarget, lastX;                  // [h + n t][64]
    std::vector<float> adHostF, adTargetF, adHTdeltaF;
    float cDeltaF[4] = {0, 0, 0, 0};
    double cPrior[4] = {0, 0, 0, 0};
    std::vector<PointHessian*> allPoints;
    // ---- bookkeeping ----
    std::vector<int> free_slots;
    long long counters[LDSO_SYS_CT_N] = {0};
    double kf_log[12] = {0};
    double phase_s[12] = {0};   // [0..7] see ldso_gpu_system_get_phase_seconds; [8] makeImages call, [9] trackNewestCoarse calls, [10] hypotheses
    int last_tries = 0;
    int n_pool = 0;                                               // size of the device-resident immature pool

    explicit FullSystemT(const ldso_system_config& c) : cfg(c) {}
    ~FullSystemT() {
        if (ctx) Api::destroy(ctx);
    }
    FullSystemT(const FullSystemT&) = delete;
    FullSystemT& operator=(const FullSystemT&) = delete;

    int create() {
        if (cfg.max_frames + 1 > LDSO_MAX_WINDOW) return fail("max_frames + 1 exceeds LDSO_MAX_WINDOW");
        cfg.gpu.n_frame_slots = std::max(cfg.gpu.n_frame_slots, cfg.max_frames + 4);
        int rc = Api::create(&cfg.gpu, &ctx);
        if (rc != 0) {
            err = ctx ? Api::last_error(ctx) : "create failed";
            if (ctx) Api::destroy(ctx);
            ctx = nullptr;
            return rc < 0 ? rc : -1;
        }
        for (int s = cfg.gpu.n_frame_slots - 1; s >= 0; s--) free_slots.push_back(s);
        Hcalib.init(cfg.gpu.fx, cfg.gpu.fy, cfg.gpu.cx, cfg.gpu.cy);
        trackerA.dev = 0; trackerB.dev = 1;
        HM.assign((size_t)CPARS * CPARS, 0.0);
        bM.assign(CPARS, 0.0);
        Api::level_size(ctx, 0, &wG0, &hG0);
        levels = Api::pyr_levels(ctx);
        return 0;
    }

    // ---- FullSystem::addActiveFrame (src/frontend/FullSystem.cc:78-187) ----------------------------------------------
    // image_kind: LDSO_IMAGE_HOST_F32 / _DEVICE_F32 (ImageAndExposure::image as Undistort hands it over) or LDSO_IMAGE_HOST_RAW8 /
    // _RAW16: the sensor image, through Undistort::undistort<T> on the device in front of makeImages (run_dso_tum_mono.cc:405-429)
    int addActiveFrame(const float* image, int image_kind, double timestamp, float exposure_time, double pose_out[12],
                       int* status_out) {
        const int rc = addActiveFrameBegin(image, image_kind, timestamp, exposure_time, pose_out, status_out);
        if (rc < 0) return rc;
        return addActiveFrameEnd();
    }

    // The same call in two halves, so that ONE host thread can keep several sequences busy (BatchRunner): Begin consumes the
    // image (makeImages) and, once the system is initialised, builds the motion hypotheses and LAUNCHES the tracker on the
    // first one without waiting; End waits for it and does everything else.  Between the two halves the thread is free to
    // begin the frames of its other sequences, whose kernels then run beside this tracker.
    struct PendingFrame {
        bool active = false, lost = false, tracking = false, launched = false;
        FrameHessian* fh = nullptr;
        double t0 = 0;
        double* pose_out = nullptr;
        int* status_out = nullptr;
        // trackNewCoarse between its halves
        FrameHessian* lastF = nullptr;
        AffLight aff_last_2_l{0, 0};
        std::vector<SE3> tries;
        float minRes0[5];
    } pend;

    // The NEXT frame's pyramid, built one frame ahead (BatchRunner, which holds all the frames of a call): queued behind the
    // current frame's tracker on the tracking stream, it runs while the host finishes the current frame, and the next
    // addActiveFrameBegin finds its image already consumed.  The image must stay valid until that frame has been tracked.
    struct Prefetched { bool valid = false; const float* image = nullptr; int slot = -1; } pre;
    int prefetchImage(const float* image, int image_kind, float exposure_time) {
        if (pre.valid || isLost || !initialized || free_slots.size() < 3) return 0;
        const int slot = free_slots.back();
        if (Api::frame_make_images_async(ctx, slot, (const void*)image, image_kind, exposure_time) < 0) return fail_api();
        free_slots.pop_back();
        pre.valid = true; pre.image = image; pre.slot = slot;
        return 0;
    }

    int addActiveFrameBegin(const float* image, int image_kind, double timestamp, float exposure_time, double pose_out[12],
                            int* status_out) {
        if (pend.active) return fail("addActiveFrameBegin: the previous frame was not finished (addActiveFrameEnd)");
        pend = PendingFrame();
        pend.active = true; pend.pose_out = pose_out; pend.status_out = status_out;
        if (isLost) { pend.lost = true; return 0; }
        const double t0 = now();
        pend.t0 = t0;
        allFrameHistory.emplace_back(new FrameHessian());
        FrameHessian* fh = allFrameHistory.back().get();
        pend.fh = fh;
        fh->id = (int)allFrameHistory.size() - 1;
        fh->timeStamp = timestamp;
        fh->ab_exposure = exposure_time;
        if (pre.valid && pre.image != image) {   // a frame was prepared that is not the one that came: give its slot back
            free_slots.push_back(pre.slot);
            pre.valid = false;
        }
        if (!pre.valid && free_slots.empty()) { pend.active = false; return fail("no free frame slot"); }
        counters[LDSO_SYS_CT_FRAMES]++;
        // fh->makeImages(image, Hcalib) (:93)
        int rc;
        if (pre.valid) {
            fh->slot = pre.slot;
            pre.valid = false;
            rc = 0;
        } else {
            fh->slot = free_slots.back();
            free_slots.pop_back();
            rc = 1;
        }
        if (rc == 0) {}
        else if (image_kind == LDSO_IMAGE_DEVICE_F32) rc = Api::frame_make_images_dev(ctx, fh->slot, image);
        else if (image_kind == LDSO_IMAGE_HOST_F32) rc = Api::frame_make_images(ctx, fh->slot, image);
        else if (image_kind == LDSO_IMAGE_HOST_RAW8 || image_kind == LDSO_IMAGE_HOST_RAW16)
            rc = Api::frame_make_images_raw(ctx, fh->slot, (const void*)image, image_kind == LDSO_IMAGE_HOST_RAW8 ? 1 : 2, exposure_time, 1.0f);
        else { pend.active = false; return fail("unknown image kind"); }
        if (rc < 0) { pend.active = false; return fail_api(); }
        phase_s[8] += now() - t0;
        if (initialized) {
            // :130-138  swap tracking reference?
            if (coarseTracker_forNewKF->refFrameID > coarseTracker->refFrameID) std::swap(coarseTracker, coarseTracker_forNewKF);
            pend.tracking = true;
            LDSO_T("frame trackNewCoarseBegin", rc = trackNewCoarseBegin(fh));
            if (rc < 0) { pend.active = false; return rc; }
        }
        phase_s[0] += now() - t0;
        return 0;
    }

    int addActiveFrameEnd() {
        if (!pend.active) return fail("addActiveFrameEnd without addActiveFrameBegin");
        pend.active = false;
        double* pose_out = pend.pose_out;
        int* status_out = pend.status_out;
        if (pend.lost) { fill_status(status_out, nullptr, false); return 1; }
        FrameHessian* fh = pend.fh;
        const double t0 = now();
        bool madeKF = false;
        int rc;
        if (!pend.tracking) {
            // :95-128
            int ready = 0;
            if (initializerFrameID < 0) {
                std::vector<int> np(LDSO_MAX_LEVELS, 0);
                if (Api::init_set_first(ctx, fh->slot, np.data()) < 0) return fail_api();
                initializerFirstFrame = fh;
                initializerFrameID = 0;
                phase_s[6] += now() - t0;
            } else {
                double pose[12];
                float aff[2], latest[3];
                int snapped = 0, iters[LDSO_MAX_LEVELS];
                if (Api::init_track_frame(ctx, fh->slot, initializerFirstFrame->ab_exposure, fh->ab_exposure, pose, aff, &snapped,
                                          &ready, latest, iters) < 0)
                    return fail_api();
                initializerFrameID++;
                phase_s[6] += now() - t0;
                if (ready) {
                    rc = initializeFromInitializer(fh, SE3::fromRt(pose));
                    if (rc < 0) return rc;
                    rc = deliverTrackedFrame(fh, true);
                    if (rc < 0) return rc;
                    madeKF = true;
                } else {
                    fh->poseValid = false;   // frame->ReleaseAll(): the pyramid is not needed any more
                    release_slot(fh);
                }
            }
        } else {
            double tres[4];
            LDSO_T("frame trackNewCoarseEnd", rc = trackNewCoarseEnd(fh, tres));
            if (rc < 0) return rc;
            phase_s[0] += now() - t0;
            if (!std::isfinite(tres[0]) || !std::isfinite(tres[1]) || !std::isfinite(tres[2]) || !std::isfinite(tres[3])) {
                isLost = true;   // :144-150
                fill_status(status_out, fh, false);
                if (pose_out) fh->Tcw.toRt(pose_out);
                return 0;
            }
            // :152-176, setting_keyframesPerSecond = 0
            bool needToMakeKF = false;
            {
                double refToFh[2];
                AffLight::fromToVecExposure(coarseTracker->lastRef->ab_exposure, fh->ab_exposure, coarseTracker->lastRef_aff_g2l,
                                            fh->aff_g2l(), refToFh);
                float b = cfg.kf_global_weight * cfg.max_shift_weight_t * sqrtf((float)tres[1]) +
                          cfg.kf_global_weight * cfg.max_shift_weight_r * sqrtf((float)tres[2]) +
                          cfg.kf_global_weight * cfg.max_shift_weight_rt * sqrtf((float)tres[3]) +
                          cfg.kf_global_weight * cfg.max_affine_weight * std::fabs(logf((float)refToFh[0]));
                bool b1 = b > 1;
                bool b2 = 2 * coarseTracker->firstCoarseRMSE < tres[0];
                needToMakeKF = allFrameHistory.size() == 1 || b1 || b2;
            }
            LDSO_T("frame deliverTrackedFrame (KF or trace)", rc = deliverTrackedFrame(fh, needToMakeKF));
            if (rc < 0) return rc;
            madeKF = needToMakeKF;
        }
        if (pose_out) fh->Tcw.toRt(pose_out);
        fill_status(status_out, fh, madeKF);
        return 0;
    }

    // relative_to_ref = 0: Frame::getPose as the reference keeps it (a non-keyframe stays where it was tracked);
    // 1: non-keyframes follow their tracking reference (camToTrackingRef composed with the reference's CURRENT pose, as
    // upstream DSO's printResult does), so the whole trajectory is consistent with the bundle-adjusted keyframes
    int getTrajectory(int max_n, int* ids, double* poses, int* is_kf, int relative_to_ref) const {
        int k = 0;
        for (const auto& f : allFrameHistory) {
            if (!f->poseValid || k >= max_n) continue;
            if (!initialized && f.get() != initializerFirstFrame) continue;
            if (ids) ids[k] = f->id;
            SE3 Tcw = f->Tcw;
            if (relative_to_ref && f->kfId < 0 && f->trackingRef) Tcw = (f->trackingRef->Tcw.inverse() * f->camToTrackingRef).inverse();
            if (poses) Tcw.toRt(poses + 12 * (size_t)k);
            if (is_kf) is_kf[k] = f->kfId >= 0 ? 1 : 0;
            k++;
        }
        return k;
    }

    int countResiduals() const {
        int n = 0;
        for (const FrameHessian* f : frames) for (const auto& p : f->points) n += (int)p->residuals.size();
        return n;
    }
    int countPoints() const {
        int n = 0;
        for (const FrameHessian* f : frames) n += (int)f->points.size();
        return n;
    }
    int countImmature() const {
        int n = 0;
        for (const FrameHessian* f : frames) n += (int)f->immature.size();
        return n;
    }

private:
    int wG0 = 0, hG0 = 0, levels = 0;

    // optional fine-grained host timers (LDSO_SYS_PROF=1): label -> seconds, printed by the destructor
    struct Prof {
        std::vector<std::pair<std::string, std::pair<double, double>>> acc;   // label -> (wall seconds, seconds inside the C ABI)
        bool on = getenv("LDSO_SYS_PROF") != nullptr;
        void add(const char* label, double dt, double api = 0) {
            if (!on) return;
            for (auto& a : acc) if (a.first == label) { a.second.first += dt; a.second.second += api; return; }
            acc.emplace_back(label, std::make_pair(dt, api));
        }
        ~Prof() {
            for (auto& a : acc)
                fprintf(stderr, "  [sys prof] %-40s %.3f ms (host outside the C ABI %.3f ms)\n", a.first.c_str(), 1e3 * a.second.first,
                        1e3 * (a.second.first - a.second.second));
        }
    } prof;

    static double now() {
        return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
    }
    int fail(const char* what) { err = what; return -2; }
    int fail_api() { err = Api::last_error(ctx); return -1; }
    void release_slot(FrameHessian* fh) {
        // oldest-first reuse: a trace queued on the mapping stream may still read the pyramid of the frame just released
        if (fh->slot >= 0) free_slots.insert(free_slots.begin(), fh->slot);
        fh->slot = -1;
    }
    void fill_status(int* st, const FrameHessian* fh, bool kf) const {
        if (!st) return;
        for (int i = 0; i < LDSO_SYS_ST_N; i++) st[i] = 0;
        st[LDSO_SYS_ST_INITIALIZED] = initialized; st[LDSO_SYS_ST_KEYFRAME] = kf; st[LDSO_SYS_ST_LOST] = isLost;
        st[LDSO_SYS_ST_INIT_FAILED] = initFailed; st[LDSO_SYS_ST_WINDOW] = (int)frames.size();
        st[LDSO_SYS_ST_POINTS] = countPoints(); st[LDSO_SYS_ST_RESIDUALS] = kf ? countResiduals() : 0;
        st[LDSO_SYS_ST_IMMATURE] = countImmature(); st[LDSO_SYS_ST_TRIES] = last_tries;
        st[LDSO_SYS_ST_FRAME_ID] = fh ? fh->id : -1;
    }

    // ---- FullSystem::deliverTrackedFrame (:189-216), linearizeOperation = true ----
    int deliverTrackedFrame(FrameHessian* fh, bool needKF) { return needKF ? makeKeyFrame(fh) : makeNonKeyFrame(fh); }

    // ---- FullSystem::trackNewCoarse (:218-468) ----------------------------------------------------------------------
    static SE3 quatRot(double w, double x, double y, double z) {   // SE3(Eigen::Quaterniond(w,x,y,z), 0): Sophus normalises
        SE3 r;
        r.qw = w; r.qx = x; r.qy = y; r.qz = z;
        r.normalize();
        return r;
    }
    // first half: the motion hypotheses (:225-380) and the launch of trackNewestCoarse on the first one
    int trackNewCoarseBegin(FrameHessian* fh) {
        const double t_h0 = now();
        FrameHessian* lastF = coarseTracker->lastRef;
        AffLight aff_last_2_l(0, 0);
        std::vector<SE3>& lastF_2_fh_tries = pend.tries;
        lastF_2_fh_tries.clear();
        if (allFrameHistory.size() == 2) {
            lastF_2_fh_tries.push_back(SE3());
        } else {
            FrameHessian* slast = allFrameHistory[allFrameHistory.size() - 2].get();
            FrameHessian* sprelast = allFrameHistory[allFrameHistory.size() - 3].get();
            SE3 slast_2_sprelast = sprelast->Tcw * slast->Tcw.inverse();
            SE3 lastF_2_slast = slast->Tcw * lastF->Tcw.inverse();
            aff_last_2_l = slast->shell_aff_g2l;
            SE3 fh_2_slast = slast_2_sprelast;   // assumed to be the same as fh_2_slast
            const SE3 fi = fh_2_slast.inverse();
            lastF_2_fh_tries.push_back(fi * lastF_2_slast);        // constant motion
            lastF_2_fh_tries.push_back(fi * fi * lastF_2_slast);   // double motion (frame skipped)
            {
                double l[6];
                fh_2_slast.log(l);
                for (int i = 0; i < 6; i++) l[i] *= 0.5;
                lastF_2_fh_tries.push_back(SE3::exp(l).inverse() * lastF_2_slast);   // half motion
            }
            lastF_2_fh_tries.push_back(lastF_2_slast);             // zero motion
            lastF_2_fh_tries.push_back(SE3());                     // zero motion from the keyframe
            // 26 small rotations x 3 magnitudes (:287-367); the float loop 0.02, 0.03, 0.04 as the reference runs it
            static const int sgn[26][3] = {{1, 0, 0},   {0, 1, 0},   {0, 0, 1},   {-1, 0, 0},  {0, -1, 0},   {0, 0, -1},  {1, 1, 0},
                                           {0, 1, 1},   {1, 0, 1},   {-1, 1, 0},  {0, -1, 1},  {-1, 0, 1},   {1, -1, 0},  {0, 1, -1},
                                           {1, 0, -1},  {-1, -1, 0}, {0, -1, -1}, {-1, 0, -1}, {-1, -1, -1}, {-1, -1, 1}, {-1, 1, -1},
                                           {-1, 1, 1},  {1, -1, -1}, {1, -1, 1},  {1, 1, -1},  {1, 1, 1}};
            const SE3 base = fi * lastF_2_slast;
            for (float rotDelta = 0.02; rotDelta < 0.05; rotDelta += 0.01)
                for (int k = 0; k < 26; k++)
                    lastF_2_fh_tries.push_back(base * quatRot(1, sgn[k][0] * rotDelta, sgn[k][1] * rotDelta, sgn[k][2] * rotDelta));
            if (!slast->poseValid || !sprelast->poseValid || !lastF->poseValid) {
                lastF_2_fh_tries.clear();
                lastF_2_fh_tries.push_back(SE3());
            }
        }
        phase_s[10] += now() - t_h0;
        pend.lastF = lastF;
        pend.aff_last_2_l = aff_last_2_l;
        {
            double pose[12];
            float aff[2] = {aff_last_2_l.a, aff_last_2_l.b};
            lastF_2_fh_tries[0].toRt(pose);
            for (int k = 0; k < 5; k++) pend.minRes0[k] = INFINITY;
            const double t_t0 = now();
            if (Api::tracker_track_begin(ctx, coarseTracker->dev, fh->slot, fh->ab_exposure, pose, aff, levels - 1, pend.minRes0) < 0) return fail_api();
            phase_s[9] += now() - t_t0;
            pend.launched = true;
        }
        return 0;
    }
    // second half: the result of the first hypothesis, further hypotheses while tracking is not good enough (:382-468)
    int trackNewCoarseEnd(FrameHessian* fh, double tres[4]) {
        FrameHessian* lastF = pend.lastF;
        const AffLight aff_last_2_l = pend.aff_last_2_l;
        const std::vector<SE3>& lastF_2_fh_tries = pend.tries;
        double flowVecs[3] = {100, 100, 100};
        SE3 lastF_2_fh;
        AffLight aff_g2l(0, 0);
        double achievedRes[5];
        for (int i = 0; i < 5; i++) achievedRes[i] = INFINITY;
        bool haveOneGood = false;
        int tryIterations = 0;
        for (size_t i = 0; i < lastF_2_fh_tries.size(); i++) {
            AffLight aff_g2l_this = aff_last_2_l;
            SE3 lastF_2_fh_this = lastF_2_fh_tries[i];
            // coarseTracker->trackNewestCoarse(fh, pose, aff, pyrLevelsUsed - 1, achievedRes) (:396-399)
            double pose[12];
            float aff[2] = {aff_g2l_this.a, aff_g2l_this.b}, minRes[5], lastRes[5], flow[3];
            int iters[2];
            lastF_2_fh_this.toRt(pose);
            for (int k = 0; k < 5; k++) minRes[k] = (float)achievedRes[k];
            const double t_t0 = now();
            int rc;
            if (i == 0 && pend.launched) rc = Api::tracker_track_end(ctx, pose, aff, lastRes, flow, iters);   // launched by trackNewCoarseBegin
            else rc = Api::tracker_track(ctx, coarseTracker->dev, fh->slot, fh->ab_exposure, pose, aff, levels - 1, minRes, lastRes, flow, iters);
            phase_s[9] += now() - t_t0;
            if (rc < 0) return fail_api();
            counters[LDSO_SYS_CT_TRACK_CALLS]++;
            const bool trackingIsGood = rc == 0;
            if (trackingIsGood) { lastF_2_fh_this = SE3::fromRt(pose); aff_g2l_this = AffLight(aff[0], aff[1]); }
            for (int k = 0; k < 5; k++) coarseTracker->lastResiduals[k] = lastRes[k];
            for (int k = 0; k < 3; k++) coarseTracker->lastFlowIndicators[k] = flow[k];
            tryIterations++;
            if (trackingIsGood && std::isfinite(lastRes[0]) && !(lastRes[0] >= achievedRes[0])) {
                for (int k = 0; k < 3; k++) flowVecs[k] = flow[k];
                aff_g2l = aff_g2l_this;
                lastF_2_fh = lastF_2_fh_this;
                haveOneGood = true;
            }
            if (haveOneGood)
                for (int k = 0; k < 5; k++)
                    if (!std::isfinite((float)achievedRes[k]) || achievedRes[k] > lastRes[k]) achievedRes[k] = lastRes[k];
            if (haveOneGood && achievedRes[0] < lastCoarseRMSE[0] * cfg.re_track_threshold) break;
        }
        last_tries = tryIterations;
        if (!haveOneGood) {   // :429-435
            flowVecs[0] = flowVecs[1] = flowVecs[2] = 0;
            aff_g2l = aff_last_2_l;
            lastF_2_fh = lastF_2_fh_tries[0];
        }
        for (int k = 0; k < 5; k++) lastCoarseRMSE[k] = achievedRes[k];
        SE3 camToWorld = lastF->Tcw.inverse() * lastF_2_fh.inverse();
        fh->Tcw = camToWorld.inverse();
        fh->trackingRef = lastF;
        fh->camToTrackingRef = lastF_2_fh.inverse();
        fh->shell_aff_g2l = aff_g2l;
        if (coarseTracker->firstCoarseRMSE < 0) coarseTracker->firstCoarseRMSE = (float)achievedRes[0];
        tres[0] = achievedRes[0]; tres[1] = flowVecs[0]; tres[2] = flowVecs[1]; tres[3] = flowVecs[2];
        return 0;
    }

    // ---- FullSystem::traceNewCoarse (:1157-1195) ----------------------------------------------------------------------
    int traceNewCoarse(FrameHessian* fh) {
        const double t0 = now();
        const int n = (int)frames.size();
        if (n == 0 || n_pool == 0) return 0;
        float K[9] = {Hcalib.fxl(), 0, Hcalib.cxl(), 0, Hcalib.fyl(), Hcalib.cyl(), 0, 0, 1}, Ki[9];
        ldso_sm::mat3_inv<float>(K, Ki);
        std::vector<float> KRKi(9 * (size_t)n), Kt(3 * (size_t)n), aff(2 * (size_t)n);
        for (int h = 0; h < n; h++) {
            const FrameHessian* host = frames[h];
            SE3 hostToNew = fh->PRE_worldToCam * host->PRE_camToWorld;
            double Rd[9];
            float R[9], KR[9], t[3];
            hostToNew.rotationMatrix(Rd);
            for (int i = 0; i < 9; i++) R[i] = (float)Rd[i];
            for (int i = 0; i < 3; i++) t[i] = (float)hostToNew.t[i];
            ldso_sm::mat3_mul<float>(K, R, KR);
            ldso_sm::mat3_mul<float>(KR, Ki, &KRKi[9 * (size_t)h]);
            ldso_sm::mat3_vec<float>(K, t, &Kt[3 * (size_t)h]);
            double ab[2];
            AffLight::fromToVecExposure(host->ab_exposure, fh->ab_exposure, host->aff_g2l(), fh->aff_g2l(), ab);
            aff[2 * (size_t)h] = (float)ab[0]; aff[2 * (size_t)h + 1] = (float)ab[1];
        }
        // queued behind the frame's pyramid on the mapping stream; every later pool call completes it first
        if (Api::immature_trace_async(ctx, fh->slot, n, KRKi.data(), Kt.data(), aff.data()) < 0) return fail_api();
        phase_s[1] += now() - t0;
        return 0;
    }

    // ---- FullSystem::makeNonKeyFrame (:701-716) ----
    int makeNonKeyFrame(FrameHessian* fh) {
        fh->setEvalPT_scaled(fh->Tcw, fh->shell_aff_g2l);
        int rc = traceNewCoarse(fh);
        release_slot(fh);   // fh->frame->ReleaseAll()
        return rc;
    }

    // ---- EnergyFunctional::setAdjointsF (EnergyFunctional.cc:551-609) ----
    void setAdjointsF() {
        const int n = (int)frames.size();
        adHost.assign(64 * (size_t)n * n, 0.0);
        adTarget.assign(64 * (size_t)n * n, 0.0);
        const double rowScale[8] = {SCALE_XI_TRANS, SCALE_XI_TRANS, SCALE_XI_TRANS, SCALE_XI_ROT, SCALE_XI_ROT, SCALE_XI_ROT, SCALE_A, SCALE_B};
        for (int h = 0; h < n; h++)
            for (int t = 0; t < n; t++) {
                const FrameHessian *host = frames[h], *target = frames[t];
                SE3 hostToTarget = target->worldToCam_evalPT * host->worldToCam_evalPT.inverse();
                double Ad[36];
                hostToTarget.Adj(Ad);
                double* AH = &adHost[64 * (size_t)(h + t * n)];
                double* AT = &adTarget[64 * (size_t)(h + t * n)];
                for (int i = 0; i < 8; i++) AH[i * 8 + i] = AT[i * 8 + i] = 1;
                for (int i = 0; i < 6; i++)
                    for (int j = 0; j < 6; j++) AH[i * 8 + j] = -Ad[j * 6 + i];   // -Adj^T
                double ab[2];
                AffLight::fromToVecExposure(host->ab_exposure, target->ab_exposure, host->aff_g2l_0(), target->aff_g2l_0(), ab);
                const float affLL0 = (float)ab[0];
                AT[6 * 8 + 6] = -affLL0; AH[6 * 8 + 6] = affLL0; AT[7 * 8 + 7] = -1; AH[7 * 8 + 7] = affLL0;
                for (int i = 0; i < 8; i++)
                    for (int j = 0; j < 8; j++) { AH[i * 8 + j] *= rowScale[i]; AT[i * 8 + j] *= rowScale[i]; }
            }
        for (int i = 0; i < 4; i++) cPrior[i] = cfg.initial_calib_hessian;
        adHostF.resize(adHost.size());
        adTargetF.resize(adTarget.size());
        for (size_t i = 0; i < adHost.size(); i++) { adHostF[i] = (float)adHost[i]; adTargetF[i] = (float)adTarget[i]; }
    }

    // ---- EnergyFunctional::setDeltaF (:523-549) ----
    void setDeltaF() {
        const int n = (int)frames.size();
        adHTdeltaF.assign(8 * (size_t)n * n, 0.f);
        std::vector<float> d(8 * (size_t)n);
        for (int h = 0; h < n; h++) {
            double r[10];
            frames[h]->get_state_minus_stateZero(r);
            for (int i = 0; i < 8; i++) d[8 * (size_t)h + i] = (float)r[i];
        }
        for (int h = 0; h < n; h++)
            for (int t = 0; t < n; t++) {
                const int idx = h + t * n;
                const float *AH = &adHostF[64 * (size_t)idx], *AT = &adTargetF[64 * (size_t)idx];
                for (int j = 0; j < 8; j++) {
                    float a = 0, b = 0;
                    for (int i = 0; i < 8; i++) { a += d[8 * (size_t)h + i] * AH[i * 8 + j]; b += d[8 * (size_t)t + i] * AT[i * 8 + j]; }
                    adHTdeltaF[8 * (size_t)idx + j] = a + b;
                }
            }
        for (int i = 0; i < 4; i++) cDeltaF[i] = (float)Hcalib.value_minus_value_zero[i];
        for (FrameHessian* f : frames) {
            double r[10];
            f->get_state_minus_stateZero(r);
            for (int i = 0; i < 8; i++) f->delta[i] = r[i];
            f->get_state_minus_statePriorZero(r);
            for (int i = 0; i < 8; i++) f->delta_prior[i] = r[i];
            for (auto& p : f->points) p->deltaF = p->idepth - p->idepth_zero;
        }
    }

    // ---- FullSystem::setPrecalcValues (:1667-1675) ----
    void setPrecalcValues() {
        const int n = (int)frames.size();
        for (FrameHessian* fr : frames) {
            fr->targetPrecalc.resize(n);
            for (int i = 0; i < n; i++) fr->targetPrecalc[i].Set(*fr, *frames[i], Hcalib);
        }
        setDeltaF();
    }

    // FrameHessian::getPrior (include/internal/FrameHessian.h:142-173) / takeData (src/internal/FrameHessian.cc:130-134)
    void frameTakeData(FrameHessian* fh) {
        double p[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        if (fh->id == 0) {
            for (int i = 0; i < 3; i++) { p[i] = cfg.initial_trans_prior; p[3 + i] = cfg.initial_rot_prior; }
            p[6] = cfg.initial_aff_a_prior; p[7] = cfg.initial_aff_b_prior;
        } else {
            p[6] = cfg.gpu.affine_opt_mode_a < 0 ? cfg.initial_aff_a_prior : cfg.gpu.affine_opt_mode_a;
            p[7] = cfg.gpu.affine_opt_mode_b < 0 ? cfg.initial_aff_b_prior : cfg.gpu.affine_opt_mode_b;
        }
        double r[10];
        for (int i = 0; i < 8; i++) fh->prior[i] = p[i];
        fh->get_state_minus_stateZero(r);
        for (int i = 0; i < 8; i++) fh->delta[i] = r[i];
        fh->get_state_minus_statePriorZero(r);
        for (int i = 0; i < 8; i++) fh->delta_prior[i] = r[i];
    }

    // ---- EnergyFunctional::insertFrame (:51-100) ----
    void insertFrame(FrameHessian* fh) {
        frameTakeData(fh);
        const int n = (int)frames.size();   // fh is already in `frames`
        const int D = 8 * n + CPARS, Do = D - 8;
        std::vector<double> H2((size_t)D * D, 0.0), b2(D, 0.0);
        for (int i = 0; i < Do; i++) {
            b2[i] = bM[i];
            for (int j = 0; j < Do; j++) H2[(size_t)i * D + j] = HM[(size_t)i * Do + j];
        }
        HM.swap(H2);
        bM.swap(b2);
        setAdjointsF();
        makeIDX();
    }

    // ---- EnergyFunctional::makeIDX (:499-521) ----
    void makeIDX() {
        for (size_t i = 0; i < frames.size(); i++) frames[i]->idx = (int)i;
        allPoints.clear();
        for (FrameHessian* f : frames)
            for (auto& p : f->points) {
                allPoints.push_back(p.get());
                for (auto& r : p->residuals) { r.hostIDX = r.host->idx; r.targetIDX = r.target->idx; }
            }
    }

    // EnergyFunctional::removePoint (:193-203): the fork never sets alreadyRemoved, nPoints only ever goes down
    void removePoint(PointHessian* ph) {
        nResiduals -= (int)ph->residuals.size();
        ph->residuals.clear();
        nPoints--;
    }
    // EnergyFunctional::dropResidual (:102-107) + deleteOut (include/internal/GlobalFuncs.h:24-34): swap with the last
    void dropResidual(PointHessian* p, size_t i) {
        p->residuals[i] = p->residuals.back();
        p->residuals.pop_back();
        nResiduals--;
    }
    // erase the points of `f` whose status is no longer ACTIVE (they stay in Frame::features in the reference, where every
    // loop skips them; here they only keep being counted as "out" by flagFramesForMarginalization)
    void sweepPoints(FrameHessian* f) {
        size_t k = 0;
        for (size_t i = 0; i < f->points.size(); i++) {
            if (f->points[i]->status == PointStatus::ACTIVE) {
                if (k != i) f->points[k] = std::move(f->points[i]);
                k++;
            } else {
                f->n_out++;
            }
        }
        f->points.resize(k);
    }
    int activePointCountForDensity() const { return cfg.npoints_counter_mode == 1 ? countPoints() : nPoints; }

    // ---- FullSystem::initializeFromInitializer (:1484-1642) -----------------------------------------------------------
    int initializeFromInitializer(FrameHessian* newFrame, const SE3& thisToNext) {
        const double t0 = now();
        GlibcRand rng(549812361);   // std::srand(549812361) (:1486), private to this sequence
        FrameHessian* firstFrame = initializerFirstFrame;
        firstFrame->idx = (int)frames.size();
        frames.push_back(firstFrame);
        firstFrame->frameID = firstFrame->kfId = numKFs;
        // the reference calls ef->insertFrame / setPrecalcValues here with the frame's default state; the evaluation point
        // is set below (:1620) and both tables are rebuilt by makeKeyFrame before anything reads them
        firstFrame->setEvalPT_scaled(firstFrame->Tcw, firstFrame->shell_aff_g2l);
        insertFrame(firstFrame);
        setPrecalcValues();
        int n0 = 0;
        if (Api::init_get_points(ctx, 0, &n0, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr) < 0) return fail_api();
        std::vector<float> pu(n0), pv(n0), piR(n0), ptype(n0);
        if (Api::init_get_points(ctx, 0, &n0, pu.data(), pv.data(), nullptr, piR.data(), nullptr, ptype.data(), nullptr, nullptr) < 0)
            return fail_api();
        float sumID = 1e-5, numID = 1e-5;
        for (int i = 0; i < n0; i++) { sumID += piR[i]; numID++; }
        const float rescaleFactor = 1 / (sumID / numID);
        const float keepPercentage = cfg.desired_point_density / n0;
        std::vector<float> su, sv, sid, st;
        for (int i = 0; i < n0; i++) {
            if (rng.next() / (float)2147483647 > keepPercentage) continue;   // rand() / (float)RAND_MAX (:1516)
            su.push_back(pu[i] + 0.5f); sv.push_back(pv[i] + 0.5f);
            sid.push_back(piR[i] * rescaleFactor); st.push_back(ptype[i]);
        }
        const int ns = (int)su.size();
        std::vector<float> col(8 * (size_t)ns), wgt(8 * (size_t)ns), gh(3 * (size_t)ns), eth(ns);
        if (ns && Api::immature_create(ctx, firstFrame->slot, ns, su.data(), sv.data(), col.data(), wgt.data(), gh.data(), eth.data()) < 0)
            return fail_api();
        for (int i = 0; i < ns; i++) {
            if (!std::isfinite(eth[i])) continue;   // :1528-1531 (PointHessian::energyTH is the ImmaturePoint's)
            std::unique_ptr<PointHessian> ph(new PointHessian());
            ph->host = firstFrame;
            ph->feat_idx = firstFrame->n_features++;
            ph->u = su[i]; ph->v = sv[i]; ph->my_type = st[i];
            memcpy(ph->color, &col[8 * (size_t)i], sizeof(ph->color));
            memcpy(ph->weights, &wgt[8 * (size_t)i], sizeof(ph->weights));
            ph->setIdepthScaled(sid[i]);
            ph->setIdepthZero(ph->idepth);
            ph->hasDepthPrior = true;
            ph->status = PointStatus::ACTIVE;
            ph->takeData(cfg.gpu.idepth_fix_prior);
            firstFrame->points.push_back(std::move(ph));
        }
        SE3 firstToNew = thisToNext;
        for (int k = 0; k < 3; k++) firstToNew.t[k] /= rescaleFactor;
        // :1617-1634 (vision only): first frame at the origin, new frame at firstToNew
        firstFrame->Tcw = SE3();
        firstFrame->setEvalPT_scaled(firstFrame->Tcw, firstFrame->shell_aff_g2l);
        newFrame->Tcw = firstToNew;
        newFrame->setEvalPT_scaled(newFrame->Tcw, newFrame->shell_aff_g2l);
        initialized = true;
        numKFs++;   // globalMap->AddKeyFrame(fr)
        phase_s[6] += now() - t0;
        return 0;
    }

    // ---- FullSystem::flagFramesForMarginalization (:763-842) -------------------------------------------------------------
    void flagFramesForMarginalization() {
        const int n = (int)frames.size();
        if (cfg.min_frame_age > cfg.max_frames) {
            for (int i = cfg.max_frames; i < n; i++) frames[i - cfg.max_frames]->flaggedForMarginalization = true;
            return;
        }
        int flagged = 0;
        for (int i = 0; i < n; i++) {
            FrameHessian* fh = frames[i];
            const int in = (int)fh->immature.size() + (int)fh->points.size(), out = fh->n_out;
            double refToFh[2];
            AffLight::fromToVecExposure(frames.back()->ab_exposure, fh->ab_exposure, frames.back()->aff_g2l(), fh->aff_g2l(), refToFh);
            if (coarseTracker->lastRef && fh->kfId != coarseTracker->lastRef->kfId) {
                if ((in < cfg.min_points_remaining * (in + out) || std::fabs(logf((float)refToFh[0])) > cfg.max_log_aff_fac_in_window) &&
                    n - flagged > cfg.min_frames) {
                    fh->flaggedForMarginalization = true;
                    flagged++;
                }
            }
        }
        if (n - flagged >= cfg.max_frames) {
            double smallestScore = 1;
            FrameHessian* toMarginalize = nullptr;
            FrameHessian* latest = frames.back();
            for (FrameHessian* fr : frames) {
                if (fr->frameID > latest->frameID - cfg.min_frame_age || fr->frameID == 0 || fr == coarseTracker->lastRef) continue;
                double distScore = 0;
                for (size_t t = 0; t < fr->targetPrecalc.size(); t++) {
                    const FrameHessian* target = frames[t];
                    if (target->frameID > latest->frameID - cfg.min_frame_age + 1 || target == fr) continue;
                    distScore += 1 / (1e-5 + fr->targetPrecalc[t].distanceLL);
                }
                distScore *= -sqrtf(fr->targetPrecalc.back().distanceLL);
                if (distScore < smallestScore) { smallestScore = distScore; toMarginalize = fr; }
            }
            if (toMarginalize) toMarginalize->flaggedForMarginalization = true;
        }
    }

    // ---- upload of (a subset of) the window to the device pools ------------------------------------------------------
    struct Upload {
        std::vector<PointHessian*> pts;
        std::vector<int> res_begin;
        std::vector<FrameHessian*> res_target;   // target of every uploaded residual, in upload order
        std::vector<unsigned char> dead;         // residuals dropped on the host since the upload
        int nr = 0;
        bool resident = false;                   // the device window still holds exactly this upload (+ its linearisation)
    } up;
    std::vector<float> precalcBuf, ethBuf;
    std::vector<double> fpriorBuf, fdeltaPriorBuf;

    int uploadFrames() {   // FrameFramePrecalc tables + thresholds + CalibHessian (setPrecalcValues' device half)
        const int n = (int)frames.size();
        precalcBuf.resize((size_t)n * n * LDSO_PRECALC_STRIDE);
        ethBuf.resize(n);
        std::vector<int> slots(n);
        for (int h = 0; h < n; h++) {
            slots[h] = frames[h]->slot;
            ethBuf[h] = frames[h]->frameEnergyTH;
            for (int t = 0; t < n; t++) frames[h]->targetPrecalc[t].pack(&precalcBuf[(size_t)(h + n * t) * LDSO_PRECALC_STRIDE]);
        }
        const float calib[4] = {Hcalib.fxl(), Hcalib.fyl(), Hcalib.cxl(), Hcalib.cyl()};
        if (Api::window_set_frames(ctx, n, slots.data(), precalcBuf.data(), ethBuf.data(), calib) < 0) return fail_api();
        return 0;
    }
    int uploadAdjoints() {   // setAdjointsF / setDeltaF products + priors
        const int n = (int)frames.size();
        fpriorBuf.resize(8 * (size_t)n);
        fdeltaPriorBuf.resize(8 * (size_t)n);
        for (int h = 0; h < n; h++)
            for (int i = 0; i < 8; i++) { fpriorBuf[8 * (size_t)h + i] = frames[h]->prior[i]; fdeltaPriorBuf[8 * (size_t)h + i] = frames[h]->delta_prior[i]; }
        if (Api::window_set_adjoints(ctx, adHost.data(), adTarget.data(), adHTdeltaF.data(), cDeltaF, cPrior, fpriorBuf.data(),
                                     fdeltaPriorBuf.data()) < 0)
            return fail_api();
        return 0;
    }
    int uploadPoints(const std::vector<PointHessian*>& pts) {
        const int np = (int)pts.size();
        up.pts = pts;
        up.resident = false;
        up.res_begin.assign(np + 1, 0);
        std::vector<float> u(np), v(np), id(np), idz(np), col(8 * (size_t)np), wgt(8 * (size_t)np), prior(np);
        std::vector<int> host(np);
        for (int i = 0; i < np; i++) {
            const PointHessian* p = pts[i];
            pts[i]->upload_idx = i;
            u[i] = p->u; v[i] = p->v; id[i] = p->idepth_scaled; idz[i] = p->idepth_zero_scaled; host[i] = p->host->idx; prior[i] = p->priorF;
            memcpy(&col[8 * (size_t)i], p->color, sizeof(p->color));
            memcpy(&wgt[8 * (size_t)i], p->weights, sizeof(p->weights));
            up.res_begin[i + 1] = up.res_begin[i] + (int)p->residuals.size();
        }
        const int nr = up.res_begin[np];
        up.nr = nr;
        if (np > cfg.gpu.max_points || nr > cfg.gpu.max_residuals) return fail("window exceeds max_points / max_residuals of the context");
        std::vector<int> target(nr), state(nr);
        std::vector<unsigned char> flags(nr);
        up.res_target.resize(nr);
        up.dead.assign(nr, 0);
        for (int i = 0, r = 0; i < np; i++)
            for (const auto& res : pts[i]->residuals) {
                up.res_target[r] = res.target;
                target[r] = res.target->idx;
                state[r] = (int)res.state_state;
                flags[r] = (unsigned char)((res.isLinearized ? LDSO_RF_LINEARIZED : 0) | (res.isActiveAndIsGoodNEW ? LDSO_RF_ACTIVE : 0) |
                                           (res.isNew ? LDSO_RF_NEW : 0));
                r++;
            }
        if (Api::window_set_points(ctx, np, u.data(), v.data(), id.data(), idz.data(), host.data(), col.data(), wgt.data(), prior.data(),
                                   up.res_begin.data()) < 0)
            return fail_api();
        if (Api::window_set_residuals(ctx, nr, target.data(), state.data(), flags.data(), nullptr) < 0) return fail_api();
        return 0;
    }

    // ---- FullSystem::linearizeAll (:1716-1769) ------------------------------------------------------------------------------
    // device: linearizeAll_Reductor over every non-linearised residual + setNewFrameEnergyTH; host (fixLinearization only):
    // state_state -> lastResiduals[].second, removal of the residuals that are not active any more
    int linearizeAll(bool fixLinearization, double out[3]) {
        float th = 0;
        if (Api::window_linearize(ctx, fixLinearization ? 1 : 0, out, &th) < 0) return fail_api();
        frames.back()->frameEnergyTH = th;
        if (!fixLinearization) return 0;
        const int np = (int)up.pts.size(), nr = up.nr;
        std::vector<int> st(std::max(nr, 1));
        std::vector<float> cen(3 * (size_t)std::max(nr, 1)), hdi(std::max(np, 1)), idh(std::max(np, 1)), ngood(std::max(np, 1)),
            mrb(std::max(np, 1));
        if (Api::window_get_keyframe_results(ctx, st.data(), cen.data(), hdi.data(), idh.data(), ngood.data(), mrb.data()) < 0) return fail_api();
        for (int i = 0; i < np; i++) {
            PointHessian* p = up.pts[i];
            p->HdiF = hdi[i];
            p->idepth_hessian = idh[i];
            p->numGoodResiduals += (int)ngood[i];
            if (mrb[i] > p->maxRelBaseline) p->maxRelBaseline = mrb[i];
            int r = up.res_begin[i];
            for (auto& res : p->residuals) {
                // applyRes(true) (include/internal/Residuals.h:70-88) ran on the device: active <=> the state it left is IN
                res.state_state = (ResState)st[r];
                res.isActiveAndIsGoodNEW = res.state_state == ResState::IN;
                for (int k = 0; k < 3; k++) res.centerProjectedTo[k] = cen[3 * (size_t)r + k];
                if (p->lastResiduals_first[0] == res.target) p->lastResiduals_second[0] = res.state_state;
                else if (p->lastResiduals_first[1] == res.target) p->lastResiduals_second[1] = res.state_state;
                r++;
            }
            // toRemove (:1749-1764): in residual order, each removal swaps the last residual in (deleteOut)
            std::vector<FrameHessian*> rm;
            {
                int rr = up.res_begin[i];
                for (const auto& res : p->residuals) { if (!res.isActive()) { rm.push_back(res.target); up.dead[rr] = 1; } rr++; }
            }
            for (FrameHessian* t : rm) {
                if (p->lastResiduals_first[0] == t) p->lastResiduals_first[0] = nullptr;
                else if (p->lastResiduals_first[1] == t) p->lastResiduals_first[1] = nullptr;
                for (size_t k = 0; k < p->residuals.size(); k++)
                    if (p->residuals[k].target == t) { dropResidual(p, k); break; }
            }
        }
        return 0;
    }

    // ---- FullSystem::getNullspaces (:2028-2076): pose (6) and scale (1) columns, the ones orthogonalize() uses ----------
    void getNullspaces(std::vector<double>& N) const {
        const int n = (int)frames.size(), D = CPARS + 8 * n, K = 7;
        N.assign((size_t)D * K, 0.0);
        for (const FrameHessian* fh : frames) {
            const int o = CPARS + 8 * fh->idx;
            for (int i = 0; i < 6; i++)
                for (int k = 0; k < 6; k++)
                    N[(size_t)(o + k) * K + i] = fh->nullspaces_pose[k][i] * (k < 3 ? SCALE_XI_TRANS_INVERSE : SCALE_XI_ROT_INVERSE);
            for (int k = 0; k < 6; k++) N[(size_t)(o + k) * K + 6] = fh->nullspaces_scale[k] * (k < 3 ? SCALE_XI_TRANS_INVERSE : SCALE_XI_ROT_INVERSE);
        }
    }

    // ---- FullSystem::optimize (:844-1007) -----------------------------------------------------------------------------------
    int optimize(int mnumOptIts, float* rmse_out) {
        *rmse_out = 0;
        if (frames.size() < 2) return 0;
        if (frames.size() < 3) mnumOptIts = 20;
        if (frames.size() < 4) mnumOptIts = 15;
        const int n = (int)frames.size(), D = CPARS + 8 * n;
        // activeResiduals = every residual of every active point (none is linearised outside flagPointsForRemoval), resetOOB
        for (PointHessian* p : allPoints)
            for (auto& r : p->residuals) r.resetOOB();
        int rc;
        LDSO_T("opt uploadFrames", rc = uploadFrames()); if (rc < 0) return rc;
        LDSO_T("opt uploadPoints", rc = uploadPoints(allPoints)); if (rc < 0) return rc;
        LDSO_T("opt uploadAdjoints", rc = uploadAdjoints()); if (rc < 0) return rc;
        kf_log[1] = (double)allPoints.size();
        kf_log[2] = (double)up.nr;
        double lastEnergy[3], newEnergy[3];
        LDSO_T("opt linearize first", rc = linearizeAll(false, lastEnergy)); if (rc < 0) return rc;
        if (Api::window_apply_res(ctx) < 0) return fail_api();
        double lambda = 1e-1;
        std::vector<double> HA((size_t)D * D), HL((size_t)D * D), Hsc((size_t)D * D), bA(D), bL(D), bsc(D), x(D), delta(D), N;
        std::vector<float> pstep(std::max<size_t>(allPoints.size(), 1)), idv(std::max<size_t>(allPoints.size(), 1));
        int its = 0;
        for (int iteration = 0; iteration < mnumOptIts; iteration++) {
            // backupState (:1934-1989, no momentum)
            for (int i = 0; i < 4; i++) Hcalib.value_backup[i] = Hcalib.value[i];
            for (FrameHessian* fh : frames) {
                for (int i = 0; i < 10; i++) fh->state_backup[i] = fh->state[i];
                for (auto& p : fh->points) p->idepth_backup = p->idepth;
            }
            // solveSystem (:1707-1714) -> EnergyFunctional::solveSystemF (EnergyFunctional.cc:280-471)
            getNullspaces(N);
            int counts[2] = {0, 0};
            LDSO_T("opt accumulate", rc = Api::window_accumulate(ctx, HA.data(), bA.data(), HL.data(), bL.data(), Hsc.data(), bsc.data(), counts)); if (rc < 0) return fail_api();
            resInA = counts[0];
            for (int i = 0; i < 4; i++) delta[i] = (double)cDeltaF[i];   // getStitchedDeltaF (EnergyFunctional.h:192-198)
            for (int h = 0; h < n; h++)
                for (int i = 0; i < 8; i++) delta[CPARS + 8 * h + i] = frames[h]->delta[i];
            LDSO_T("opt solve", rc = Api::window_solve(ctx, iteration, lambda, HA.data(), bA.data(), HL.data(), bL.data(), Hsc.data(), bsc.data(), HM.data(),
                                  bM.data(), delta.data(), N.data(), 7, x.data()));
            if (rc < 0) return fail_api();
            lastX = x;
            counters[LDSO_SYS_CT_BA_ITERATIONS]++;
            its++;
            // resubstituteF_MT (:611-637): currentLambda is the pinned 1e-5 of SOLVER_FIX_LAMBDA
            LDSO_T("opt resubstitute", rc = Api::window_resubstitute(ctx, x.data(), 1e-5f, pstep.data())); if (rc < 0) return fail_api();
            for (int i = 0; i < 4; i++) Hcalib.step[i] = -x[i];
            for (FrameHessian* h : frames) {
                for (int i = 0; i < 8; i++) h->step[i] = -x[CPARS + 8 * h->idx + i];
                h->step[8] = h->step[9] = 0;
            }
            for (size_t i = 0; i < allPoints.size(); i++) allPoints[i]->step = pstep[i];
            double xn = 0;
            for (int i = 0; i < D; i++) xn += x[i] * x[i];
            if (std::isnan(std::sqrt(xn))) { isLost = true; *rmse_out = 3.4e38f; return 0; }   // :907-911
            bool canbreak = false;
            LDSO_T("opt doStepFromBackup", canbreak = doStepFromBackup());
            // device half of setPrecalcValues + the new inverse depths
            for (size_t i = 0; i < allPoints.size(); i++) idv[i] = allPoints[i]->idepth_scaled;
            LDSO_T("opt it uploads", rc = uploadFrames(); if (rc >= 0) rc = uploadAdjoints(); if (rc >= 0 && !allPoints.empty() && Api::window_set_idepth(ctx, idv.data(), idv.data()) < 0) rc = fail_api());
            if (rc < 0) return rc;
            LDSO_T("opt it linearize", rc = linearizeAll(false, newEnergy)); if (rc < 0) return rc;
            // setting_forceAceptStep (:935): always accepted
            if (Api::window_apply_res(ctx) < 0) return fail_api();
            for (int i = 0; i < 3; i++) lastEnergy[i] = newEnergy[i];
            lambda *= 0.25;
            if (canbreak && iteration >= cfg.min_opt_iterations) break;
        }
        kf_log[10] = its;
        // new FEJ zero point for the newest frame (:972-983)
        double newStateZero[10] = {0};
        FrameHessian* fh = frames.back();
        newStateZero[6] = fh->state[6]; newStateZero[7] = fh->state[7];
        fh->setEvalPT(fh->PRE_worldToCam, newStateZero);
        LDSO_T("opt final host tables", setAdjointsF(); setPrecalcValues());
        LDSO_T("opt final uploads", rc = uploadFrames(); if (rc >= 0) rc = uploadAdjoints()); if (rc < 0) return rc;
        LDSO_T("opt final linearize(true)+readback", rc = linearizeAll(true, lastEnergy)); if (rc < 0) return rc;
        up.resident = true;
        if (!std::isfinite(lastEnergy[0]) || !std::isfinite(lastEnergy[1]) || !std::isfinite(lastEnergy[2])) isLost = true;
        for (FrameHessian* fr : frames) {   // :995-1004
            fr->Tcw = fr->PRE_camToWorld.inverse();
            fr->shell_aff_g2l = fr->aff_g2l();
        }
        *rmse_out = sqrtf((float)(lastEnergy[0] / (patternNum * resInA)));
        return 0;
    }

    // ---- FullSystem::doStepFromBackup (:1826-1932), no momentum, all step factors 1 ----
    bool doStepFromBackup() {
        float sumA = 0, sumB = 0, sumT = 0, sumR = 0, sumID = 0, numID = 0, sumNID = 0;
        double nv[4];
        for (int i = 0; i < 4; i++) nv[i] = Hcalib.value_backup[i] + 1.0f * Hcalib.step[i];
        Hcalib.setValue(nv);
        for (FrameHessian* fh : frames) {
            double st[10], s6[6];
            for (int i = 0; i < 10; i++) st[i] = fh->state_backup[i] + 1.0 * fh->step[i];
            for (int i = 0; i < 6; i++) s6[i] = 1.0 * fh->step[i];
            (SE3::exp(s6) * SE3::exp(fh->state_backup)).log(st);   // fork: composed on SE(3) (SURVEY.md trap T9)
            fh->setState(st);
            sumA += (float)(fh->step[6] * fh->step[6]);
            sumB += (float)(fh->step[7] * fh->step[7]);
            sumT += (float)(fh->step[0] * fh->step[0] + fh->step[1] * fh->step[1] + fh->step[2] * fh->step[2]);
            sumR += (float)(fh->step[3] * fh->step[3] + fh->step[4] * fh->step[4] + fh->step[5] * fh->step[5]);
            for (auto& ph : fh->points) {
                ph->setIdepth(ph->idepth_backup + 1.0f * ph->step);
                sumID += ph->step * ph->step;
                sumNID += fabsf(ph->idepth_backup);
                numID++;
                ph->setIdepthZero(ph->idepth_backup + 1.0f * ph->step);
            }
        }
        const float nf = (float)frames.size();
        sumA /= nf; sumB /= nf; sumR /= nf; sumT /= nf; sumID /= numID; sumNID /= numID;
        setPrecalcValues();
        return sqrtf(sumA) < 0.0005 * cfg.th_opt_iterations && sqrtf(sumB) < 0.00005 * cfg.th_opt_iterations &&
               sqrtf(sumR) < 0.00005 * cfg.th_opt_iterations && sqrtf(sumT) * sumNID < 0.00005 * cfg.th_opt_iterations;
    }

    // ---- FullSystem::activatePointsMT (:1197-1337) ---------------------------------------------------------------------
    void makeKLevel1(float K1[9], float Ki0[9]) const {   // CoarseDistanceMap::makeK = CoarseTracker::makeK (CoarseTracker.cc:312-339)
        const float fx0 = Hcalib.fxl(), fy0 = Hcalib.fyl(), cx0 = Hcalib.cxl(), cy0 = Hcalib.cyl();
        const float fx1 = fx0 * 0.5f, fy1 = fy0 * 0.5f, cx1 = (cx0 + 0.5f) / ((int)1 << 1) - 0.5f, cy1 = (cy0 + 0.5f) / ((int)1 << 1) - 0.5f;
        const float K0[9] = {fx0, 0, cx0, 0, fy0, cy0, 0, 0, 1};
        const float k1[9] = {fx1, 0, cx1, 0, fy1, cy1, 0, 0, 1};
        for (int i = 0; i < 9; i++) K1[i] = k1[i];
        ldso_sm::mat3_inv<float>(K0, Ki0);
    }
    int activatePointsMT() {
        const double t0 = now();
        const int np_ctrl = activePointCountForDensity();
        const float dens = cfg.desired_point_density;
        if (np_ctrl < dens * 0.66) currentMinActDist -= 0.8;
        if (np_ctrl < dens * 0.8) currentMinActDist -= 0.5;
        else if (np_ctrl < dens * 0.9) currentMinActDist -= 0.2;
        else if (np_ctrl < dens) currentMinActDist -= 0.1;
        if (np_ctrl > dens * 1.5) currentMinActDist += 0.8;
        if (np_ctrl > dens * 1.3) currentMinActDist += 0.5;
        if (np_ctrl > dens * 1.15) currentMinActDist += 0.2;
        if (np_ctrl > dens) currentMinActDist += 0.1;
        if (currentMinActDist < 0) currentMinActDist = 0;
        if (currentMinActDist > 4) currentMinActDist = 4;
        kf_log[8] = currentMinActDist;
        kf_log[6] = 0;
        if (n_pool == 0) { phase_s[2] += now() - t0; return 0; }
        const int n = (int)frames.size();
        FrameHessian* newestFr = frames.back();
        float K1[9], Ki0[9];
        makeKLevel1(K1, Ki0);
        std::vector<float> KRKi(9 * (size_t)n), Kt(3 * (size_t)n);
        std::vector<unsigned char> skip(n, 0), flagged(n, 0);
        for (int h = 0; h < n; h++) {
            const FrameHessian* host = frames[h];
            SE3 fhToNew = newestFr->PRE_worldToCam * host->PRE_camToWorld;
            double Rd[9];
            float R[9], KR[9], t[3];
            fhToNew.rotationMatrix(Rd);
            for (int i = 0; i < 9; i++) R[i] = (float)Rd[i];
            for (int i = 0; i < 3; i++) t[i] = (float)fhToNew.t[i];
            ldso_sm::mat3_mul<float>(K1, R, KR);
            ldso_sm::mat3_mul<float>(KR, Ki0, &KRKi[9 * (size_t)h]);
            ldso_sm::mat3_vec<float>(K1, t, &Kt[3 * (size_t)h]);
            skip[h] = host == newestFr;
            flagged[h] = host->flaggedForMarginalization;
        }
        // coarseDistanceMap->makeDistanceMap (CoarseTracker.cc:795-830): every active point of the other frames
        std::vector<int> ph_host;
        std::vector<float> pu, pv, pid;
        for (const FrameHessian* f : frames) {
            if (f == newestFr) continue;
            for (const auto& p : f->points) { ph_host.push_back(f->idx); pu.push_back(p->u); pv.push_back(p->v); pid.push_back(p->idepth_scaled); }
        }
        int rcp = 0;
        LDSO_T("act distmap_make", rcp = Api::distmap_make(ctx, n, KRKi.data(), Kt.data(), (int)pu.size(), ph_host.data(), pu.data(), pv.data(), pid.data()));
        if (rcp < 0) return fail_api();
        // the ordered candidate walk (:1233-1298)
        std::vector<unsigned char> action(std::max(n_pool, 1), 0);
        int cnt[2] = {0, 0};
        LDSO_T("act select_activation", rcp = Api::immature_select_activation(ctx, n, KRKi.data(), Kt.data(), skip.data(), flagged.data(), currentMinActDist,
                                            cfg.min_trace_quality, action.data(), cnt));
        if (rcp < 0) return fail_api();
        std::vector<int> cand, gone;
        for (int i = 0; i < n_pool; i++) {
            if (action[i] == 2) cand.push_back(i);
            if (action[i] != 0) gone.push_back(i);
        }
        // optimizeImmaturePoint for every candidate (:1304-1311), against the window's FrameFramePrecalc tables
        std::vector<float> idp(std::max<size_t>(cand.size(), 1));
        std::vector<int> ok(std::max<size_t>(cand.size(), 1)), rst(std::max<size_t>(cand.size(), 1) * n);
        if (!cand.empty()) {
            int rc = uploadFrames();
            if (rc < 0) return rc;
            LDSO_T("act immature_activate", rcp = Api::immature_activate(ctx, (int)cand.size(), cand.data(), 1, idp.data(), ok.data(), rst.data())); if (rcp < 0) return fail_api();
        }
        // walk the pool in order (frames, then features) and apply the outcome (:1313-1336)
        int pi = 0, ci = 0, n_act = 0;
        for (FrameHessian* f : frames) {
            std::vector<internal::ImmatureFeature> keep;
            std::vector<std::unique_ptr<PointHessian>> born;
            for (size_t k = 0; k < f->immature.size(); k++, pi++) {
                const internal::ImmatureFeature& im = f->immature[k];
                if (action[pi] == 0) { keep.push_back(im); continue; }
                if (action[pi] == 1) { f->n_out++; continue; }
                const int c = ci++;
                if (!ok[c]) { f->n_out++; continue; }
                std::unique_ptr<PointHessian> p(new PointHessian());
                p->host = f; p->feat_idx = im.feat_idx; p->u = im.u; p->v = im.v; p->my_type = im.my_type;
                memcpy(p->color, im.color, sizeof(p->color));
                memcpy(p->weights, im.weights, sizeof(p->weights));
                p->setIdepth(idp[c]);   // the PointHessian constructor's value is overwritten at once (:1121-1122)
                p->setIdepthZero(idp[c]);
                p->status = PointStatus::ACTIVE;
                for (int t = 0; t < n; t++) {
                    if (frames[t] == f || rst[(size_t)c * n + t] != (int)ResState::IN) continue;
                    PointFrameResidual r;
                    r.host = f; r.target = frames[t];
                    r.resetOOB();   // constructor (Residuals.h:49-56), then state_NewEnergy = state_energy = 0, OUTLIER / IN
                    p->residuals.push_back(r);
                    if (frames[t] == frames.back()) { p->lastResiduals_first[0] = frames[t]; p->lastResiduals_second[0] = ResState::IN; }
                    else if (n >= 2 && frames[t] == frames[n - 2]) { p->lastResiduals_first[1] = frames[t]; p->lastResiduals_second[1] = ResState::IN; }
                }
                p->takeData(cfg.gpu.idepth_fix_prior);
                nResiduals += (int)p->residuals.size();   // ef->insertResidual
                born.push_back(std::move(p));
                n_act++;
            }
            f->immature.swap(keep);
            if (!born.empty()) {   // merge by feature position: Frame::features order is the allPoints order
                std::vector<std::unique_ptr<PointHessian>> merged;
                merged.reserve(f->points.size() + born.size());
                size_t a = 0, b = 0;
                while (a < f->points.size() || b < born.size()) {
                    if (b >= born.size() || (a < f->points.size() && f->points[a]->feat_idx < born[b]->feat_idx)) merged.push_back(std::move(f->points[a++]));
                    else merged.push_back(std::move(born[b++]));
                }
                f->points.swap(merged);
            }
        }
        if (!gone.empty()) {
            int np2 = 0;
            LDSO_T("act compact", rcp = Api::immature_compact(ctx, -1, 0, (int)gone.size(), gone.data(), &np2)); if (rcp < 0) return fail_api();
            n_pool = np2;
        }
        counters[LDSO_SYS_CT_ACTIVATED] += n_act;
        kf_log[6] = n_act;
        phase_s[2] += now() - t0;
        return 0;
    }

    // ---- FullSystem::removeOutliers (:1644-1665) ----
    void removeOutliers() {
        int dropped = 0;
        for (FrameHessian* fr : frames) {
            for (auto& ph : fr->points)
                if (ph->residuals.empty()) { ph->status = PointStatus::OUTLIER; dropped++; }
        }
        dropPointsF();
        counters[LDSO_SYS_CT_DROPPED_POINTS] += dropped;
        kf_log[4] += dropped;
    }
    // ---- EnergyFunctional::dropPointsF (:264-278) ----
    void dropPointsF() {
        for (FrameHessian* f : frames) {
            for (auto& p : f->points)
                if (p->status == PointStatus::OUTLIER || p->status == PointStatus::OUT) removePoint(p.get());
            // MARGINALIZED points are still needed by marginalizePointsF: only the dropped ones leave here
            size_t k = 0;
            for (size_t i = 0; i < f->points.size(); i++) {
                if (f->points[i]->status == PointStatus::OUTLIER || f->points[i]->status == PointStatus::OUT) { f->n_out++; continue; }
                if (k != i) f->points[k] = std::move(f->points[i]);
                k++;
            }
            f->points.resize(k);
        }
        makeIDXKeepMarginalized();
    }
    void makeIDXKeepMarginalized() {
        for (size_t i = 0; i < frames.size(); i++) frames[i]->idx = (int)i;
        allPoints.clear();
        for (FrameHessian* f : frames)
            for (auto& p : f->points)
                if (p->status == PointStatus::ACTIVE) allPoints.push_back(p.get());
    }

    // ---- FullSystem::flagPointsForRemoval (:1356-1424) + EnergyFunctional::marginalizePointsF (:205-262) -----------------
    // The points to marginalise are re-linearised at the current state, their linearisation is fixed and they are folded into
    // HM, bM.  On the device: a window holding just those points (same frames, same tables) goes through K7 (linearize +
    // applyRes), fixLinearizationF for the active residuals, addPoint<2> + the Schur accumulator, stitched to (8n+4)^2.
    int flagAndMarginalizePoints() {
        std::vector<FrameHessian*> fhsToMargPoints;
        for (FrameHessian* f : frames) if (f->flaggedForMarginalization) fhsToMargPoints.push_back(f);
        std::vector<PointHessian*> toMarg;
        int flag_out = 0;
        for (FrameHessian* fr : frames)
            for (auto& php : fr->points) {
                PointHessian* ph = php.get();
                if (ph->idepth_scaled < 0 || ph->residuals.empty()) {
                    ph->status = PointStatus::OUTLIER;
                    flag_out++;
                } else if (ph->isOOB(fhsToMargPoints, cfg.min_good_active_res_for_marg, cfg.min_good_res_for_marg) || fr->flaggedForMarginalization) {
                    if (ph->isInlierNew(cfg.min_good_active_res_for_marg, cfg.min_good_res_for_marg)) {
                        if (ph->idepth_hessian > cfg.min_idepth_h_marg) { ph->status = PointStatus::MARGINALIZED; toMarg.push_back(ph); }
                        else { ph->status = PointStatus::OUT; flag_out++; }
                    } else {
                        ph->status = PointStatus::OUT;
                        flag_out++;
                    }
                }
            }
        // (the re-linearisation of the OUT points of :1390-1399 has no effect: they are dropped right away)
        dropPointsF();   // :626
        counters[LDSO_SYS_CT_DROPPED_POINTS] += flag_out;
        kf_log[4] += flag_out;
        kf_log[3] = (double)toMarg.size();
        kf_log[5] = 0;
        const int n = (int)frames.size(), D = CPARS + 8 * n;
        if (!toMarg.empty()) {
            int rc;
            std::vector<double> H((size_t)D * D), b(D);
            int nres = 0, nlin = 0;
            if (up.resident && !getenv("LDSO_SYS_NO_SUBSET")) {
                // the window optimize ran on is still resident: re-judge, fix and marginalise in place (one call, no upload)
                std::vector<unsigned char> pm(up.pts.size(), 0);
                for (PointHessian* p : toMarg) pm[p->upload_idx] = 1;
                std::vector<float> eth(n);
                for (int i = 0; i < n; i++) eth[i] = frames[i]->frameEnergyTH;
                LDSO_T("marg subset", rc = Api::window_marginalize_subset(ctx, pm.data(), up.dead.data(), eth.data(), cfg.idepth_fix_prior_marg_fac,
                                                                          H.data(), b.data(), &nres, &nlin));
                if (rc < 0) return fail_api();
                up.resident = false;
            } else {
                // general path: a window holding just those points goes through K7 (linearize + applyRes), fixLinearizationF,
                // addPoint<2> + the Schur accumulator
                for (PointHessian* p : toMarg)
                    for (auto& r : p->residuals) { r.resetOOB(); r.isLinearized = false; }
                if ((rc = uploadFrames()) < 0 || (rc = uploadPoints(toMarg)) < 0 || (rc = uploadAdjoints()) < 0) return rc;
                double stats[3];
                float th_keep = frames.back()->frameEnergyTH, th = 0;
                if (Api::window_linearize(ctx, 0, stats, &th) < 0) return fail_api();   // r->linearize (:1392)
                frames.back()->frameEnergyTH = th_keep;                                 // no setNewFrameEnergyTH on this path
                if (Api::window_apply_res(ctx) < 0) return fail_api();                  // r->applyRes(true) (:1394)
                std::vector<int> st(std::max(up.nr, 1));
                if (up.nr && Api::window_get_residual_ints(ctx, 0, st.data()) < 0) return fail_api();
                std::vector<unsigned char> mask(std::max(up.nr, 1), 0), pm(toMarg.size(), 1);
                for (size_t i = 0, r = 0; i < toMarg.size(); i++)
                    for (auto& res : toMarg[i]->residuals) {
                        if ((ResState)st[r] == ResState::IN) { mask[r] = 1; nlin++; }   // fixLinearizationF (:1396)
                        r++;
                    }
                if (up.nr && Api::window_fix_linearization(ctx, mask.data()) < 0) return fail_api();
                if (Api::window_marginalize_points(ctx, pm.data(), cfg.idepth_fix_prior_marg_fac, H.data(), b.data(), &nres) < 0) return fail_api();
            }
            counters[LDSO_SYS_CT_LIN_RESIDUALS] += nlin;
            kf_log[5] = nlin;
            // marginalizePointsF: priorF *= setting_idepthFixPriorMargFac, addPoint<2>, SC addPoint, stitch, H = M - Msc
            resInM += nres;
            counters[LDSO_SYS_CT_RES_IN_M] = resInM;
            for (size_t i = 0; i < (size_t)D * D; i++) HM[i] += cfg.marg_weight_fac * H[i];
            for (int i = 0; i < D; i++) bM[i] += cfg.marg_weight_fac * b[i];
            for (PointHessian* p : toMarg) { p->priorF *= cfg.idepth_fix_prior_marg_fac; removePoint(p); }
            counters[LDSO_SYS_CT_MARG_POINTS] += (long long)toMarg.size();
        }
        // the marginalised points leave the frames' active lists (they remain "out" features)
        for (FrameHessian* f : frames) sweepPoints(f);
        makeIDX();
        return 0;
    }

    // ---- FullSystem::makeNewTraces (:1426-1482) ---------------------------------------------------------------------------
    int makeNewTraces(FrameHessian* newFrame) {
        const int cap = std::max(16, (wG0 * hG0) / 4);
        std::vector<float> u(cap), v(cap), typ(cap);
        int n = 0;
        if (cfg.point_selection == 1) {
            std::vector<float> score(cap);
            std::vector<unsigned char> corner(cap);
            int nc = 0;
            if (Api::detect_corners(ctx, newFrame->slot, (int)cfg.desired_immature_density, cap, u.data(), v.data(), score.data(), corner.data(),
                                    &n, &nc) < 0)
                return fail_api();
            for (int i = 0; i < n; i++) typ[i] = 1;
        } else {
            // pixelSelector->makeMaps(newFrame, selectionMap, setting_desiredImmatureDensity) and the pixel loop: the selector
            // yields 4 <= x < w - 5, 4 <= y <= h - 4 (PixelSelector2.cc:218), the loop keeps 3 <= x < w - 4, 3 <= y < h - 4
            if (Api::select_points(ctx, newFrame->slot, cfg.desired_immature_density, 1, 1.0f, &currentPotential, patternPadding + 2, cap,
                                   u.data(), v.data(), typ.data(), &n) < 0)
                return fail_api();
        }
        std::vector<float> col(8 * (size_t)std::max(n, 1)), wgt(8 * (size_t)std::max(n, 1)), gh(3 * (size_t)std::max(n, 1)), eth(std::max(n, 1));
        if (n && Api::immature_create(ctx, newFrame->slot, n, u.data(), v.data(), col.data(), wgt.data(), gh.data(), eth.data()) < 0)
            return fail_api();
        std::vector<float> ku, kv, kt;
        for (int i = 0; i < n; i++) {
            if (!std::isfinite(eth[i])) continue;   // :1456-1459
            internal::ImmatureFeature im;
            im.u = u[i]; im.v = v[i]; im.my_type = typ[i]; im.feat_idx = newFrame->n_features++;
            memcpy(im.color, &col[8 * (size_t)i], sizeof(im.color));
            memcpy(im.weights, &wgt[8 * (size_t)i], sizeof(im.weights));
            newFrame->immature.push_back(im);
            ku.push_back(u[i]); kv.push_back(v[i]); kt.push_back(typ[i]);
        }
        if (!ku.empty()) {
            const int first = n_pool;
            int np2 = 0;
            if (Api::immature_append(ctx, newFrame->slot, newFrame->idx, (int)ku.size(), ku.data(), kv.data(), &np2) < 0) return fail_api();
            n_pool = np2;
            if (Api::immature_set_types(ctx, first, (int)kt.size(), kt.data()) < 0) return fail_api();
        }
        return 0;
    }

    // ---- CoarseTracker::makeK + setCoarseTrackingRef (CoarseTracker.cc:312-355), host half of makeCoarseDepthL0 (:357-382) --
    int setCoarseTrackingRef(TrackerHost* trk) {
        FrameHessian* lastRef = frames.back();
        if (Api::tracker_make_k(ctx, trk->dev, Hcalib.fxl(), Hcalib.fyl(), Hcalib.cxl(), Hcalib.cyl()) < 0) return fail_api();
        std::vector<float> u, v, id, hdi;
        for (const FrameHessian* fh : frames)
            for (const auto& ph : fh->points) {
                if (ph->lastResiduals_first[0] == nullptr || ph->lastResiduals_second[0] != ResState::IN) continue;
                const PointFrameResidual* r = ph->find(ph->lastResiduals_first[0]);
                if (!r) continue;
                u.push_back(r->centerProjectedTo[0]); v.push_back(r->centerProjectedTo[1]); id.push_back(r->centerProjectedTo[2]);
                hdi.push_back(ph->HdiF);
            }
        const AffLight aff = lastRef->aff_g2l();
        if (Api::tracker_set_ref(ctx, trk->dev, lastRef->slot, (int)u.size(), u.data(), v.data(), id.data(), hdi.data(), lastRef->ab_exposure,
                                 aff.a, aff.b) < 0)
            return fail_api();
        trk->lastRef = lastRef;
        trk->refFrameID = lastRef->id;
        trk->lastRef_aff_g2l = aff;
        trk->firstCoarseRMSE = -1;
        return 0;
    }

    // ---- FullSystem::marginalizeFrame (:718-761) + EnergyFunctional::marginalizeFrame (EnergyFunctional.cc:109-191) --------
    int marginalizeFrame(FrameHessian* frame) {
        const int n = (int)frames.size(), D = CPARS + 8 * n, Dn = D - 8;
        std::vector<double> H2((size_t)Dn * Dn), b2(Dn);
        if (Api::window_marginalize_frame(ctx, n, frame->idx, HM.data(), bM.data(), frame->prior, frame->delta_prior, H2.data(), b2.data()) < 0)
            return fail_api();
        HM.swap(H2);
        bM.swap(b2);
        // drop all observations of existing points in that frame (:724-748)
        for (FrameHessian* fr : frames) {
            if (fr == frame) continue;
            for (auto& ph : fr->points) {
                size_t nres = ph->residuals.size();
                for (size_t i = 0; i < nres; i++) {
                    if (ph->residuals[i].target == frame) {
                        if (ph->lastResiduals_first[0] == frame) ph->lastResiduals_first[0] = nullptr;
                        else if (ph->lastResiduals_first[1] == frame) ph->lastResiduals_first[1] = nullptr;
                        dropResidual(ph.get(), i);
                        i--;
                        nres--;
                    }
                }
            }
        }
        // the frame's immature points leave the device pool, host indices above it move down
        if (n_pool > 0) {
            int np2 = 0;
            if (Api::immature_compact(ctx, frame->idx, 0, 0, nullptr, &np2) < 0) return fail_api();
            n_pool = np2;
        }
        frame->immature.clear();
        frame->points.clear();
        release_slot(frame);   // frame->ReleaseAll()
        frames.erase(frames.begin() + frame->idx);   // deleteOutOrder
        for (size_t i = 0; i < frames.size(); i++) frames[i]->idx = (int)i;
        makeIDX();
        setPrecalcValues();
        setAdjointsF();
        counters[LDSO_SYS_CT_MARG_FRAMES]++;
        kf_log[7] += 1;
        return 0;
    }

    // ---- FullSystem::makeKeyFrame (:485-699) ---------------------------------------------------------------------------------
    int makeKeyFrame(FrameHessian* fh) {
        int rc;
        double t0 = now();
        for (int i = 0; i < 12; i++) kf_log[i] = 0;
        fh->setEvalPT_scaled(fh->Tcw, fh->shell_aff_g2l);
        if ((rc = traceNewCoarse(fh)) < 0) return rc;
        flagFramesForMarginalization();
        fh->idx = (int)frames.size();
        frames.push_back(fh);
        fh->kfId = fh->frameID = numKFs;
        insertFrame(fh);
        setPrecalcValues();
        // new residuals of the old points into the new frame (:546-571)
        for (FrameHessian* fht : frames) {
            if (fht == fh) continue;
            for (auto& ph : fht->points) {
                PointFrameResidual r;
                r.host = fht; r.target = fh;
                r.resetOOB();
                ph->residuals.push_back(r);
                nResiduals++;   // ef->insertResidual
                ph->lastResiduals_first[1] = ph->lastResiduals_first[0]; ph->lastResiduals_second[1] = ph->lastResiduals_second[0];
                ph->lastResiduals_first[0] = fh; ph->lastResiduals_second[0] = ResState::IN;
            }
        }
        phase_s[7] += now() - t0;
        LDSO_T("KF activatePointsMT", rc = activatePointsMT()); if (rc < 0) return rc;
        makeIDX();
        t0 = now();
        fh->frameEnergyTH = frames.back()->frameEnergyTH;
        float rmse = 0;
        LDSO_T("KF optimize", rc = optimize(cfg.max_opt_iterations, &rmse)); if (rc < 0) return rc;
        kf_log[0] = rmse;
        kf_log[9] = fh->frameEnergyTH;
        phase_s[3] += now() - t0;
        // initialisation failure tests (:584-601); numKFs counts the new frame only after AddKeyFrame below
        if (numKFs <= 4) {
            if (numKFs == 2 && rmse > 20 * cfg.initializer_slack_factor) initFailed = true;
            if (numKFs == 3 && rmse > 13 * cfg.initializer_slack_factor) initFailed = true;
            if (numKFs == 4 && rmse > 9 * cfg.initializer_slack_factor) initFailed = true;
        }
        if (isLost) return 0;
        t0 = now();
        LDSO_T("KF removeOutliers", removeOutliers());
        phase_s[4] += now() - t0;
        t0 = now();
        LDSO_T("KF setCoarseTrackingRef", rc = setCoarseTrackingRef(coarseTracker_forNewKF)); if (rc < 0) return rc;
        phase_s[5] += now() - t0;
        t0 = now();
        LDSO_T("KF flagAndMarginalizePoints", rc = flagAndMarginalizePoints()); if (rc < 0) return rc;
        phase_s[4] += now() - t0;
        t0 = now();
        LDSO_T("KF makeNewTraces", rc = makeNewTraces(fh)); if (rc < 0) return rc;
        phase_s[5] += now() - t0;
        t0 = now();
        for (size_t i = 0; i < frames.size(); i++)
            if (frames[i]->flaggedForMarginalization) {
                LDSO_T("KF marginalizeFrame", rc = marginalizeFrame(frames[i])); if (rc < 0) return rc;
                i = 0;   // as the reference: restart the scan (after its i++)
            }
        double hn = 0;
        for (double v : HM) hn += v * v;
        kf_log[11] = std::sqrt(hn);
        phase_s[4] += now() - t0;
        numKFs++;   // globalMap->AddKeyFrame(fh->frame)
        counters[LDSO_SYS_CT_KEYFRAMES]++;
        return 0;
    }
};

}  // namespace ldso

// ---- several sequences stepped together (BASELINE.json configs[4]: batches of independent sequences per GPU) -----------
// "one host thread per GPU ... steps its contexts".  Worker t owns systems t, t + T, ...; a step hands
// every system its next frame and returns when all of them are done.  Nothing is shared between sequences: each has its
// own context, streams and device pools, so the kernels of different sequences overlap on the GPU.
namespace ldso {
template <class Sys, class Api> class BatchRunner {
public:
    struct Job {
        const float* const* images = nullptr;   // n_frames image pointers of this system, fed in order
        int n_frames = 1;
        int is_device = 0;                      // LDSO_IMAGE_* kind of the pointers
        double timestamp = 0;                   // of the first frame; later frames count up by one
        float exposure = 1;
        double* pose_out = nullptr;             // [n_frames][12]
        int* status_out = nullptr;              // [n_frames][LDSO_SYS_ST_N]
        int sync_after = 0;
        int rc = 0;
    };
    std::vector<Sys*> sys;
    std::vector<Job> jobs;
    bool prefetch = getenv("LDSO_NO_PREFETCH") == nullptr;   // measurement knob: LDSO_NO_PREFETCH=1 prepares no frame ahead
    bool fibers = getenv("LDSO_NO_FIBERS") == nullptr;       // measurement knob: LDSO_NO_FIBERS=1 runs a worker's sequences one after the other
    BatchRunner(int n, Sys** systems, int n_threads) : sys(systems, systems + n), jobs(n) {
        const int T = std::max(1, std::min(n_threads, n));
        for (int t = 0; t < T; t++) workers.emplace_back([this, t, T] { loop(t, T); });
    }
    ~BatchRunner() {
        {
            std::lock_guard<std::mutex> lk(mu);
            quit = true;
            generation++;
        }
        cv.notify_all();
        for (auto& w : workers) w.join();
    }
    void step() {
        {
            std::lock_guard<std::mutex> lk(mu);
            pending = (int)workers.size();
            generation++;
        }
        cv.notify_all();
        std::unique_lock<std::mutex> lk(mu);
        done.wait(lk, [this] { return pending == 0; });
    }

private:
    std::vector<std::thread> workers;
    std::mutex mu;
    std::condition_variable cv, done;
    long long generation = 0;
    int pending = 0;
    bool quit = false;
    // ---- fibers: one per sequence of a worker -------------------------------------------------------------------------
    static constexpr size_t kFiberStack = 1 << 20;
    struct Fiber {
        ucontext_t uc;
        std::vector<char> stack;
        BatchRunner* runner = nullptr;
        size_t job = 0;
        bool done = false;
    };
    struct Worker {
        ucontext_t sched;
        std::vector<Fiber> fib;
        Fiber* cur = nullptr;
        unsigned long long resumed_at = 0;      // time-stamp counter when the current fiber was resumed
        int idle_polls = 0;                     // consecutive resumes that only polled (no host work between resume and yield)
    };
    static Fiber*& current_fiber() { static thread_local Fiber* f = nullptr; return f; }
    static void fiber_entry() {
        Fiber* f = current_fiber();
        f->runner->run_job(f->job);
        f->done = true;                                    // returning resumes uc_link = the scheduler
    }
    static void fiber_yield(void* arg) {                   // the library's wait hook
        Worker* w = (Worker*)arg;
        if (!w->cur) return;
        // a fiber that yields again within ~1 us of being resumed has only looked at its flag
        if (__builtin_ia32_rdtsc() - w->resumed_at < 3000) w->idle_polls++;
        else w->idle_polls = 0;
        swapcontext(&w->cur->uc, &w->sched);
    }
    // all frames of one sequence: begin (makeImages + the tracker launch, nothing waits), prepare the next frame behind it, end
    void run_job(size_t i) {
        Job& j = jobs[i];
        for (int f = 0; f < j.n_frames && j.rc >= 0; f++) {
            j.rc = sys[i]->impl.addActiveFrameBegin(j.images[f], j.is_device, j.timestamp + f, j.exposure,
                                                    j.pose_out ? j.pose_out + 12 * (size_t)f : nullptr,
                                                    j.status_out ? j.status_out + LDSO_SYS_ST_N * (size_t)f : nullptr);
            // the next frame of this call is at hand: its copy and pyramid queue up behind the tracker just launched
            if (j.rc >= 0 && prefetch && f + 1 < j.n_frames && sys[i]->impl.prefetchImage(j.images[f + 1], j.is_device, j.exposure) < 0)
                j.rc = -1;
            if (j.rc >= 0) j.rc = sys[i]->impl.addActiveFrameEnd();
        }
        if (j.sync_after && j.rc >= 0 && Api::sync(sys[i]->impl.ctx) < 0) j.rc = -1;   // drains the queued trace too
    }
    void loop(int t, int T) {
        long long seen = 0;
        bool bound = false;
        Worker w;                                   // the fibers' stacks live as long as the worker: allocated once
        for (;;) {
            {
                std::unique_lock<std::mutex> lk(mu);
                cv.wait(lk, [&] { return generation != seen; });
                seen = generation;
                if (quit) return;
            }
            // Every own sequence runs its frames on a FIBER of this thread (its own stack, ucontext): wherever the library
            // would wait for the device (tracker mailbox, stream / event synchronisation: ~30 round trips per keyframe) the
            // wait hook switches to the scheduler below, which resumes the next sequence.  The thread therefore never spins
            // while one of its sequences has host work to do, a keyframe of one sequence holds back neither the other
            // workers nor the worker's other sequences, and the kernels of all of them overlap on the GPU.
            std::vector<size_t> own;
            for (size_t i = t; i < sys.size(); i += T) {
                if (!bound && sys[i]->impl.ctx) Api::bind_thread(sys[i]->impl.ctx);   // the device of the ctx becomes this thread's
                jobs[i].rc = 0;
                own.push_back(i);
            }
            if (own.size() <= 1 || !fibers) {
                for (size_t i : own) run_job(i);
            } else {
                if (w.fib.size() != own.size()) w.fib = std::vector<Fiber>(own.size());   // (a worker's share never changes)
                w.idle_polls = 0;
                for (size_t k = 0; k < own.size(); k++) {
                    Fiber& f = w.fib[k];
                    f.runner = this; f.job = own[k]; f.done = false;
                    if (f.stack.size() != kFiberStack) f.stack.resize(kFiberStack);
                    getcontext(&f.uc);
                    f.uc.uc_stack.ss_sp = f.stack.data();
                    f.uc.uc_stack.ss_size = f.stack.size();
                    f.uc.uc_link = &w.sched;
                    makecontext(&f.uc, (void (*)())&BatchRunner::fiber_entry, 0);
                }
                Api::set_wait_hook(&BatchRunner::fiber_yield, &w);
                for (size_t left = own.size(); left > 0;) {
                    for (size_t k = 0; k < own.size(); k++) {
                        Fiber& f = w.fib[k];
                        if (f.done) continue;
                        w.cur = &f;
                        current_fiber() = &f;
                        w.resumed_at = __builtin_ia32_rdtsc();
                        swapcontext(&w.sched, &f.uc);      // runs the sequence until its next wait (or its end: uc_link)
                        w.cur = nullptr;
                        if (f.done) { left--; w.idle_polls = 0; }
                        // every live sequence of this worker is waiting for the device: ease off for ~1 us instead of hammering the
                        // flags (the core's other hardware thread and the memory system belong to workers that have work)
                        if (w.idle_polls >= (int)left && left > 0) {
                            for (int i = 0; i < 32; i++) __builtin_ia32_pause();
                            w.idle_polls = 0;
                        }
                    }
                }
                Api::set_wait_hook(nullptr, nullptr);
            }
            bound = true;
            {
                std::lock_guard<std::mutex> lk(mu);
                if (--pending == 0) done.notify_all();
            }
        }
    }
};
}  // namespace ldso

#ifdef LDSO_API_TIMING
#define LDSO_API_TIMING_EXPORT out[11] = ldso::api_seconds();
#else
#define LDSO_API_TIMING_EXPORT
#endif
// ---- the extern "C" face (include/ldso_system.h) for a symbol prefix ----------------------------------------------------
#define LDSO_DEFINE_SYSTEM_EXPORTS(P, API)                                                                                 \
    struct P##system {                                                                                                    \
        ldso::FullSystemT<API> impl;                                                                                      \
        explicit P##system(const ldso_system_config& c) : impl(c) {}                                                      \
    };                                                                                                                    \
    extern "C" void P##system_default_config(ldso_system_config* c) {                                                     \
        memset(c, 0, sizeof(*c));                                                                                         \
        P##default_config(&c->gpu);                                                                                       \
        /* mode=1 of examples/run_dso_tum_mono.cc:283-288: no photometric calibration, a and b free without prior */      \
        c->gpu.affine_opt_mode_a = 0; c->gpu.affine_opt_mode_b = 0;                                                       \
        c->gpu.max_points = 12288; c->gpu.max_residuals = 98304; c->gpu.max_immature = 16384; c->gpu.n_frame_slots = 12;  \
        c->desired_immature_density = 1500; c->desired_point_density = 2000; c->min_points_remaining = 0.05f;             \
        c->max_log_aff_fac_in_window = 0.7f; c->min_frames = 5; c->max_frames = 7; c->min_frame_age = 1;                  \
        c->max_opt_iterations = 6; c->min_opt_iterations = 1; c->th_opt_iterations = 1.2f;                                \
        c->max_shift_weight_t = 0.04f; c->max_shift_weight_r = 0.0f; c->max_shift_weight_rt = 0.02f;                      \
        c->kf_global_weight = 1.0f; c->max_affine_weight = 2.0f; c->idepth_fix_prior_marg_fac = 600 * 600;                \
        c->initial_rot_prior = 1e11f; c->initial_trans_prior = 1e10f; c->initial_aff_b_prior = 1e14f;                     \
        c->initial_aff_a_prior = 1e14f; c->initial_calib_hessian = 5e9f; c->min_idepth_h_marg = 50;                       \
        c->marg_weight_fac = 0.5f * 0.5f; c->min_trace_quality = 3; c->re_track_threshold = 1.5f;                         \
        c->min_good_active_res_for_marg = 3; c->min_good_res_for_marg = 4; c->initializer_slack_factor = 1;               \
        c->point_selection = 0; c->npoints_counter_mode = 0;                                                              \
    }                                                                                                                     \
    extern "C" int P##system_create(const ldso_system_config* cfg, P##system** out) {                                     \
        if (!cfg || !out) return -2;                                                                                      \
        P##system* s = new (std::nothrow) P##system(*cfg);                                                                \
        if (!s) return -3;                                                                                                \
        *out = s; /* returned even on failure so that the caller can read last_error, then destroy */                     \
        return s->impl.create();                                                                                          \
    }                                                                                                                     \
    extern "C" void P##system_destroy(P##system* s) { delete s; }                                                         \
    extern "C" const char* P##system_last_error(const P##system* s) { return s ? s->impl.err.c_str() : "null system"; }   \
    extern "C" API::ctx_t* P##system_ctx(P##system* s) { return s->impl.ctx; }                                            \
    extern "C" int P##system_add_active_frame(P##system* s, const float* image, int image_is_device, double timestamp,    \
                                              float exposure_time, double pose_out[12], int* status_out) {                \
        if (!s || !image) return -2;                                                                                      \
        return s->impl.addActiveFrame(image, image_is_device, timestamp, exposure_time, pose_out, status_out);            \
    }                                                                                                                     \
    extern "C" int P##system_get_trajectory(P##system* s, int max_n, int* ids, double* poses, int* is_kf,                 \
                                            int relative_to_ref) {                                                        \
        return s->impl.getTrajectory(max_n, ids, poses, is_kf, relative_to_ref);                                          \
    }                                                                                                                     \
    extern "C" int P##system_get_window(P##system* s, int max_n, int* ids) {                                              \
        int k = 0;                                                                                                        \
        for (auto* f : s->impl.frames) if (k < max_n) ids[k++] = f->id;                                                   \
        return k;                                                                                                         \
    }                                                                                                                     \
    extern "C" int P##system_get_counters(P##system* s, long long* out) {                                                 \
        for (int i = 0; i < LDSO_SYS_CT_N; i++) out[i] = s->impl.counters[i];                                             \
        return 0;                                                                                                         \
    }                                                                                                                     \
    extern "C" int P##system_get_prior(P##system* s, double* HM, double* bM) {                                            \
        if (HM) memcpy(HM, s->impl.HM.data(), sizeof(double) * s->impl.HM.size());                                        \
        if (bM) memcpy(bM, s->impl.bM.data(), sizeof(double) * s->impl.bM.size());                                        \
        return (int)s->impl.frames.size();                                                                                \
    }                                                                                                                     \
    extern "C" int P##system_last_keyframe_log(P##system* s, double out[12]) {                                            \
        for (int i = 0; i < 12; i++) out[i] = s->impl.kf_log[i];                                                          \
        return 0;                                                                                                         \
    }                                                                                                                     \
    extern "C" int P##system_rand_sequence(unsigned seed, int n, int* out) { /* the private glibc rand() restatement */   \
        ldso::GlibcRand g(seed);                                                                                          \
        for (int i = 0; i < n; i++) out[i] = g.next();                                                                    \
        return 0;                                                                                                         \
    }                                                                                                                     \
    extern "C" int P##system_get_phase_seconds(P##system* s, double out[12]) {                                            \
        for (int i = 0; i < 12; i++) out[i] = s->impl.phase_s[i];                                                         \
        LDSO_API_TIMING_EXPORT                                                                                            \
        return 0;                                                                                                         \
    }                                                                                                                     \
    struct P##batch {                                                                                                     \
        ldso::BatchRunner<P##system, API> run;                                                                            \
        P##batch(int n, P##system** s, int t) : run(n, s, t) {}                                                           \
    };                                                                                                                    \
    extern "C" int P##batch_create(int n, P##system** systems, int n_threads, P##batch** out) {                           \
        if (n <= 0 || !systems || !out) return -2;                                                                        \
        *out = new (std::nothrow) P##batch(n, systems, n_threads);                                                        \
        return *out ? 0 : -3;                                                                                             \
    }                                                                                                                     \
    extern "C" void P##batch_destroy(P##batch* b) { delete b; }                                                           \
    extern "C" int P##batch_add_active_frames(P##batch* b, const float* const* images, int image_is_device,              \
                                              const double* timestamps, const float* exposures, int sync_after,           \
                                              double* poses_out, int* status_out) {                                       \
        if (!b || !images) return -2;                                                                                     \
        const int n = (int)b->run.sys.size();                                                                             \
        for (int i = 0; i < n; i++) {                                                                                     \
            auto& j = b->run.jobs[i];                                                                                     \
            j.images = images + i; j.n_frames = 1; j.is_device = image_is_device;                                         \
            j.timestamp = timestamps ? timestamps[i] : 0;                                                                 \
            j.exposure = exposures ? exposures[i] : 1.0f; j.sync_after = sync_after;                                      \
            j.pose_out = poses_out ? poses_out + 12 * (size_t)i : nullptr;                                                \
            j.status_out = status_out ? status_out + LDSO_SYS_ST_N * (size_t)i : nullptr;                                 \
        }                                                                                                                 \
        b->run.step();                                                                                                    \
        int worst = 0;                                                                                                    \
        for (int i = 0; i < n; i++) if (b->run.jobs[i].rc < 0) worst = b->run.jobs[i].rc;                                 \
        return worst;                                                                                                     \
    }                                                                                                                     \
    extern "C" int P##batch_run(P##batch* b, int n_frames, const float* const* images, int image_is_device,              \
                                int sync_after, double* poses_out, int* status_out) {                                     \
        if (!b || !images || n_frames < 1) return -2;                                                                     \
        const int n = (int)b->run.sys.size();                                                                             \
        for (int i = 0; i < n; i++) {                                                                                     \
            auto& j = b->run.jobs[i];                                                                                     \
            j.images = images + (size_t)i * n_frames; j.n_frames = n_frames; j.is_device = image_is_device;               \
            j.timestamp = (double)b->run.sys[i]->impl.allFrameHistory.size(); j.exposure = 1.0f;                          \
            j.sync_after = sync_after;                                                                                    \
            j.pose_out = poses_out ? poses_out + 12 * (size_t)i * n_frames : nullptr;                                     \
            j.status_out = status_out ? status_out + LDSO_SYS_ST_N * (size_t)i * n_frames : nullptr;                      \
        }                                                                                                                 \
        b->run.step();                                                                                                    \
        int worst = 0;                                                                                                    \
        for (int i = 0; i < n; i++) if (b->run.jobs[i].rc < 0) worst = b->run.jobs[i].rc;                                 \
        return worst;                                                                                                     \
    }
