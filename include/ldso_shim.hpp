// ldso_shim.hpp — C++ facade over the C ABI (include/ldso_gpu.h) with the reference's class and member names.
//
// The reference (n-lalanne/LDSO) has no plugin layer: FullSystem calls ldso::CoarseTracker, ldso::PixelSelector,
// ldso::internal::FrameHessian / ImmaturePoint / EnergyFunctional directly.  This header shows — and lets C++ code use
// — the same surface on top of libldso_gpu.so without Eigen / Sophus: poses are ldso_sm::SE3 (unit quaternion +
// translation, the representation Sophus uses), affine brightness is AffLight {a, b}.  Each method names the reference
// member it stands for; INTEGRATION.md lists the corresponding edits inside the reference itself.
//
// Error behaviour: a negative return of the C ABI (CUDA failure / bad argument) throws std::runtime_error with
// ldso_gpu_last_error(); algorithmic outcomes keep the reference's conventions (trackNewestCoarse returns bool, …).
//
// The symbol prefix is a macro so that the test-suite can compile the SAME facade against the CPU oracle
// (-DLDSO_SHIM_PREFIX=orc_ -DLDSO_SHIM_CTX=orc_ctx); product code never defines it.
#pragma once
#include <stdexcept>
#include <string>
#include <vector>

#include "ldso_gpu.h"
#include "ldso_smallmat.h"

#ifndef LDSO_SHIM_PREFIX
#define LDSO_SHIM_PREFIX ldso_gpu_
#define LDSO_SHIM_CTX ldso_gpu_ctx
#else
struct LDSO_SHIM_CTX;
#endif
#define LDSO_SHIM_CAT2(a, b) a##b
#define LDSO_SHIM_CAT(a, b) LDSO_SHIM_CAT2(a, b)
#define LDSO_API(name) LDSO_SHIM_CAT(LDSO_SHIM_PREFIX, name)

#ifdef LDSO_SHIM_DECLARE_FOREIGN
// declarations of the mirrored entry points when the prefix is not ldso_gpu_ (test builds only)
extern "C" {
void LDSO_API(default_config)(ldso_gpu_config*);
int LDSO_API(create)(const ldso_gpu_config*, LDSO_SHIM_CTX**);
void LDSO_API(destroy)(LDSO_SHIM_CTX*);
const char* LDSO_API(last_error)(const LDSO_SHIM_CTX*);
int LDSO_API(pyr_levels)(const LDSO_SHIM_CTX*);
int LDSO_API(frame_make_images)(LDSO_SHIM_CTX*, int, const float*);
int LDSO_API(frame_download)(LDSO_SHIM_CTX*, int, int, int, float*);
int LDSO_API(tracker_make_k)(LDSO_SHIM_CTX*, int, float, float, float, float);
int LDSO_API(tracker_set_ref)(LDSO_SHIM_CTX*, int, int, int, const float*, const float*, const float*, const float*, float, float,
                              float);
int LDSO_API(tracker_track)(LDSO_SHIM_CTX*, int, int, float, double*, float*, int, const float*, float*, float*, int*);
int LDSO_API(select_make_maps)(LDSO_SHIM_CTX*, int, float, int, float, int*, float*, int*);
int LDSO_API(select_points)(LDSO_SHIM_CTX*, int, float, int, float, int*, int, int, float*, float*, float*, int*);
int LDSO_API(immature_set)(LDSO_SHIM_CTX*, int, const int*, const float*, const float*, const float*, const float*, const float*,
                           const float*, const float*, const float*, const float*, const int*);
int LDSO_API(immature_append)(LDSO_SHIM_CTX*, int, int, int, const float*, const float*, int*);
int LDSO_API(immature_compact)(LDSO_SHIM_CTX*, int, int, int, const int*, int*);
int LDSO_API(immature_trace)(LDSO_SHIM_CTX*, int, int, const float*, const float*, const float*, int*);
int LDSO_API(immature_trace_async)(LDSO_SHIM_CTX*, int, int, const float*, const float*, const float*);
int LDSO_API(immature_trace_counts)(LDSO_SHIM_CTX*, int*);
int LDSO_API(immature_get)(LDSO_SHIM_CTX*, float*, float*, float*, int*, float*, float*);
int LDSO_API(immature_activate)(LDSO_SHIM_CTX*, int, const int*, int, float*, int*, int*);
int LDSO_API(distmap_make)(LDSO_SHIM_CTX*, int, const float*, const float*, int, const int*, const float*, const float*, const float*);
int LDSO_API(distmap_get)(LDSO_SHIM_CTX*, float*);
int LDSO_API(immature_select_activation)(LDSO_SHIM_CTX*, int, const float*, const float*, const unsigned char*, const unsigned char*,
                                         float, float, unsigned char*, int*);
int LDSO_API(init_set_first)(LDSO_SHIM_CTX*, int, int*);
int LDSO_API(init_track_frame)(LDSO_SHIM_CTX*, int, float, float, double*, float*, int*, int*, float*, int*);
int LDSO_API(init_get_points)(LDSO_SHIM_CTX*, int, int*, float*, float*, float*, float*, unsigned char*, float*, float*, float*);
int LDSO_API(window_set_frames)(LDSO_SHIM_CTX*, int, const int*, const float*, const float*, const float*);
int LDSO_API(window_set_points)(LDSO_SHIM_CTX*, int, const float*, const float*, const float*, const float*, const int*,
                                const float*, const float*, const float*, const int*);
int LDSO_API(window_set_residuals)(LDSO_SHIM_CTX*, int, const int*, const int*, const unsigned char*, const float*);
int LDSO_API(window_set_adjoints)(LDSO_SHIM_CTX*, const double*, const double*, const float*, const float*, const double*,
                                  const double*, const double*);
int LDSO_API(window_set_idepth)(LDSO_SHIM_CTX*, const float*, const float*);
int LDSO_API(window_linearize)(LDSO_SHIM_CTX*, int, double*, float*);
int LDSO_API(window_apply_res)(LDSO_SHIM_CTX*);
int LDSO_API(window_accumulate)(LDSO_SHIM_CTX*, double*, double*, double*, double*, double*, double*, int*);
int LDSO_API(window_solve)(LDSO_SHIM_CTX*, int, double, const double*, const double*, const double*, const double*, const double*,
                           const double*, const double*, const double*, const double*, const double*, int, double*);
int LDSO_API(window_resubstitute)(LDSO_SHIM_CTX*, const double*, float, float*);
int LDSO_API(window_marginalize_points)(LDSO_SHIM_CTX*, const unsigned char*, float, double*, double*, int*);
int LDSO_API(window_marginalize_frame)(LDSO_SHIM_CTX*, int, int, const double*, const double*, const double*, const double*, double*, double*);
int LDSO_API(window_calc_energy)(LDSO_SHIM_CTX*, const double*, const double*, const double*, double*);
}
#endif

namespace ldso_b200 {

using SE3 = ldso_sm::SE3;

struct AffLight {                       // include/AffLight.h
    float a = 0, b = 0;
};

enum class ImmaturePointStatus { IPS_GOOD = 0, IPS_OOB, IPS_OUTLIER, IPS_SKIPPED, IPS_BADCONDITION, IPS_UNINITIALIZED };

// One sequence = one context (replaces the process-global calibration and settings, GlobalCalib.cc / Setting.cc).
class Context {
public:
    ldso_gpu_config cfg;
    LDSO_SHIM_CTX* h = nullptr;
    explicit Context(int w, int hgt, float fx, float fy, float cx, float cy) {
        LDSO_API(default_config)(&cfg);
        cfg.w = w; cfg.h = hgt; cfg.fx = fx; cfg.fy = fy; cfg.cx = cx; cfg.cy = cy;
    }
    void create() {
        if (LDSO_API(create)(&cfg, &h) != 0) {
            std::string m = h ? LDSO_API(last_error)(h) : "create failed";
            if (h) LDSO_API(destroy)(h);
            h = nullptr;
            throw std::runtime_error(m);
        }
    }
    ~Context() { if (h) LDSO_API(destroy)(h); }
    Context(const Context&) = delete;
    Context& operator=(const Context&) = delete;
    int check(int rc) const {
        if (rc < 0) throw std::runtime_error(LDSO_API(last_error)(h));
        return rc;
    }
    int pyrLevelsUsed() const { return LDSO_API(pyr_levels)(h); }
};

// ldso::internal::FrameHessian (include/internal/FrameHessian.h): the pyramid lives in a device slot
class FrameHessian {
public:
    int slot = -1;
    float ab_exposure = 1;
    AffLight aff_g2l;
    // void FrameHessian::makeImages(float* color, const shared_ptr<CalibHessian>&)  (src/internal/FrameHessian.cc:59)
    void makeImages(Context& c, int slot_, const float* color) {
        slot = slot_;
        c.check(LDSO_API(frame_make_images)(c.h, slot, color));
    }
    // host mirror of dIp[lvl] (3 floats / px) or absSquaredGrad[lvl], for code that still reads them
    void download(Context& c, int lvl, bool absSquaredGrad, float* dst) const {
        c.check(LDSO_API(frame_download)(c.h, slot, lvl, absSquaredGrad ? 1 : 0, dst));
    }
};

// ldso::PixelSelector (include/frontend/PixelSelector2.h)
class PixelSelector {
public:
    int currentPotential = 3;
    // int makeMaps(fh, float* map_out, float density, int recursionsLeft = 1, bool plot = false, float thFactor = 1)
    int makeMaps(Context& c, const FrameHessian& fh, float* map_out, float density, int recursionsLeft = 1, bool /*plot*/ = false,
                 float thFactor = 1) {
        int n = 0;
        c.check(LDSO_API(select_make_maps)(c.h, fh.slot, density, recursionsLeft, thFactor, &currentPotential, map_out, &n));
        return n;
    }
    // makeMaps + the pixel loop of FullSystem::makeNewTraces (FullSystem.cc:1440-1460)
    int selectPoints(Context& c, const FrameHessian& fh, float density, int border, std::vector<float>& u, std::vector<float>& v,
                     std::vector<float>& type, int recursionsLeft = 1, float thFactor = 1) {
        const int cap = c.cfg.w * c.cfg.h / 2;
        u.resize(cap); v.resize(cap); type.resize(cap);
        int n = 0;
        c.check(LDSO_API(select_points)(c.h, fh.slot, density, recursionsLeft, thFactor, &currentPotential, border, cap, u.data(),
                                        v.data(), type.data(), &n));
        u.resize(n); v.resize(n); type.resize(n);
        return n;
    }
};

// ldso::CoarseTracker (include/frontend/CoarseTracker.h:41-64)
class CoarseTracker {
public:
    int id;                                    // 0: coarseTracker, 1: coarseTracker_forNewKF
    double lastResiduals[5] = {0, 0, 0, 0, 0};
    double lastFlowIndicators[3] = {0, 0, 0};
    double firstCoarseRMSE = -1;
    int refFrameID = -1;
    const FrameHessian* lastRef = nullptr;
    AffLight lastRef_aff_g2l;
    explicit CoarseTracker(int id_ = 0) : id(id_) {}
    // void makeK(shared_ptr<CalibHessian>)  (CoarseTracker.cc:312)
    void makeK(Context& c, float fx, float fy, float cx, float cy) { c.check(LDSO_API(tracker_make_k)(c.h, id, fx, fy, cx, cy)); }
    // void setCoarseTrackingRef(vector<shared_ptr<FrameHessian>>&) + makeCoarseDepthL0 (:347-538): the caller gathers, for every
    // active residual targeting the reference frame, centerProjectedTo (u, v, new idepth) and the point's HdiF (:366-383)
    void setCoarseTrackingRef(Context& c, const FrameHessian& ref, int frameID, const std::vector<float>& u, const std::vector<float>& v,
                              const std::vector<float>& idepth, const std::vector<float>& HdiF) {
        lastRef = &ref; refFrameID = frameID; lastRef_aff_g2l = ref.aff_g2l; firstCoarseRMSE = -1;
        c.check(LDSO_API(tracker_set_ref)(c.h, id, ref.slot, (int)u.size(), u.data(), v.data(), idepth.data(), HdiF.data(),
                                          ref.ab_exposure, ref.aff_g2l.a, ref.aff_g2l.b));
    }
    // bool trackNewestCoarse(newFrameHessian, SE3& lastToNew_out, AffLight& aff_g2l_out, int coarsestLvl, Vec5 minResForAbort)
    bool trackNewestCoarse(Context& c, const FrameHessian& newFrame, SE3& lastToNew_out, AffLight& aff_g2l_out, int coarsestLvl,
                           const double minResForAbort[5]) {
        double Rt[12];
        lastToNew_out.toRt(Rt);
        float aff[2] = {aff_g2l_out.a, aff_g2l_out.b}, minRes[5], lastRes[5], flow[3];
        int iters[2] = {0, 0};
        for (int i = 0; i < 5; i++) minRes[i] = (float)minResForAbort[i];
        const int rc = c.check(LDSO_API(tracker_track)(c.h, id, newFrame.slot, newFrame.ab_exposure, Rt, aff, coarsestLvl, minRes,
                                                      lastRes, flow, iters));
        for (int i = 0; i < 5; i++) lastResiduals[i] = lastRes[i];
        for (int i = 0; i < 3; i++) lastFlowIndicators[i] = flow[i];
        if (rc != 0) return false;                            // the reference's `return false` paths (:286-309)
        if (firstCoarseRMSE < 0) firstCoarseRMSE = lastResiduals[0];
        lastToNew_out = SE3::fromRt(Rt);
        aff_g2l_out.a = aff[0]; aff_g2l_out.b = aff[1];
        return true;
    }
};

struct ImmaturePoolSize { int size = 0; };   // base of ImmaturePool (declared early for CoarseDistanceMap)

// ldso::CoarseInitializer (include/frontend/CoarseInitializer.h:66-196)
class CoarseInitializer {
public:
    struct Pnt { float u, v, idepth, iR, my_type, lastHessian, energy; bool isGood; };   // the fields other code reads (:23-63)
    int frameID = -1;
    int numPoints[LDSO_MAX_LEVELS] = {0};
    SE3 thisToNext;
    AffLight thisToNext_aff;
    const FrameHessian* firstFrame = nullptr;
    bool snapped = false;
    // void setFirst(shared_ptr<CalibHessian>, shared_ptr<FrameHessian>)  (CoarseInitializer.cc:656-737)
    void setFirst(Context& c, const FrameHessian& first) {
        firstFrame = &first;
        c.check(LDSO_API(init_set_first)(c.h, first.slot, numPoints));
        thisToNext = SE3(); thisToNext_aff = AffLight(); snapped = false; frameID = 0;
    }
    // bool trackFrame(shared_ptr<FrameHessian>)  (:45-183)
    bool trackFrame(Context& c, const FrameHessian& newFrame) {
        double Rt[12];
        float aff[2], res[3];
        int sn = 0, ready = 0, iters[LDSO_MAX_LEVELS];
        c.check(LDSO_API(init_track_frame)(c.h, newFrame.slot, firstFrame->ab_exposure, newFrame.ab_exposure, Rt, aff, &sn, &ready, res,
                                           iters));
        thisToNext = SE3::fromRt(Rt);
        thisToNext_aff.a = aff[0]; thisToNext_aff.b = aff[1];
        snapped = sn != 0;
        frameID++;
        return ready != 0;
    }
    // points[lvl] as FullSystem::initializeFromInitializer reads them (FullSystem.cc:1500-1544)
    std::vector<Pnt> points(Context& c, int lvl) {
        int n = numPoints[lvl];
        std::vector<float> u(n + 1), v(n + 1), id(n + 1), iR(n + 1), ty(n + 1), lh(n + 1), en(n + 1);
        std::vector<unsigned char> good(n + 1);
        c.check(LDSO_API(init_get_points)(c.h, lvl, &n, u.data(), v.data(), id.data(), iR.data(), good.data(), ty.data(), lh.data(),
                                          en.data()));
        std::vector<Pnt> out(n);
        for (int i = 0; i < n; i++) out[i] = Pnt{u[i], v[i], id[i], iR[i], ty[i], lh[i], en[i], good[i] != 0};
        return out;
    }
};

// ldso::CoarseDistanceMap (include/frontend/CoarseTracker.h:95-135) + the candidate walk of FullSystem::activatePointsMT
class CoarseDistanceMap {
public:
    // void makeDistanceMap(frameHessians, frame)  (CoarseTracker.cc:795-835): host[i] = window index of point i's host frame (-1: the
    // newest frame's own points), KRKi / Kt = K[1] R Ki[0], K[1] t of host -> newest frame (FullSystem.cc:1236-1238)
    void makeDistanceMap(Context& c, int nHosts, const float* KRKi, const float* Kt, const std::vector<int>& host,
                         const std::vector<float>& u, const std::vector<float>& v, const std::vector<float>& idepth) {
        c.check(LDSO_API(distmap_make)(c.h, nHosts, KRKi, Kt, (int)host.size(), host.data(), u.data(), v.data(), idepth.data()));
    }
    // float* fwdWarpedIDDistFinal, (w/2)*(h/2)
    void fwdWarpedIDDistFinal(Context& c, std::vector<float>& out, int w1, int h1) {
        out.resize((size_t)w1 * h1);
        c.check(LDSO_API(distmap_get)(c.h, out.data()));
    }
    // FullSystem.cc:1226-1298 over the immature pool, addIntoDistFinal (CoarseTracker.cc:925-930) for every accepted point;
    // action[i]: 0 keep, 1 delete, 2 optimise
    void selectForActivation(Context& c, const ImmaturePoolSize& pool, int nHosts, const float* KRKi, const float* Kt,
                             const unsigned char* hostIsNewest, const unsigned char* hostFlagged, float currentMinActDist,
                             float minTraceQuality, std::vector<unsigned char>& action, int counts[2]) {
        action.resize(pool.size > 0 ? pool.size : 1);
        c.check(LDSO_API(immature_select_activation)(c.h, nHosts, KRKi, Kt, hostIsNewest, hostFlagged, currentMinActDist, minTraceQuality,
                                                     action.data(), counts));
        action.resize(pool.size);
    }
};

// The immature points of the window as one device-resident pool (ldso::internal::ImmaturePoint, batched)
class ImmaturePool : public ImmaturePoolSize {
public:
    void clear(Context& c) {
        int dummy_i = 0;
        float dummy_f = 0;
        c.check(LDSO_API(immature_set)(c.h, 0, &dummy_i, &dummy_f, &dummy_f, &dummy_f, &dummy_f, &dummy_f, &dummy_f, &dummy_f, &dummy_f,
                                       &dummy_f, &dummy_i));
        size = 0;
    }
    // new ImmaturePoint(...) for the selected pixels of a keyframe (ImmaturePoint.cc:14-39, FullSystem.cc:1448-1460)
    int add(Context& c, const FrameHessian& host, int hostIdx, const std::vector<float>& u, const std::vector<float>& v) {
        c.check(LDSO_API(immature_append)(c.h, host.slot, hostIdx, (int)u.size(), u.data(), v.data(), &size));
        return size;
    }
    // FullSystem::traceNewCoarse (FullSystem.cc:1157-1195): ImmaturePoint::traceOn for every pool point; tables per host
    // as built at :1171-1176; returns the per-status counts
    void traceOn(Context& c, const FrameHessian& frame, int nHosts, const float* KRKi, const float* Kt, const float* affine,
                 int counts[6]) {
        c.check(LDSO_API(immature_trace)(c.h, frame.slot, nHosts, KRKi, Kt, affine, counts));
    }
    // the same trace queued without waiting (nothing reads its results before the next keyframe's activatePointsMT; any
    // other pool call completes it); traceCounts waits if it still runs
    void traceOnAsync(Context& c, const FrameHessian& frame, int nHosts, const float* KRKi, const float* Kt, const float* affine) {
        c.check(LDSO_API(immature_trace_async)(c.h, frame.slot, nHosts, KRKi, Kt, affine));
    }
    void traceCounts(Context& c, int counts[6]) { c.check(LDSO_API(immature_trace_counts)(c.h, counts)); }
    // deletion of OOB / activated points and of the points of a marginalised host frame
    int remove(Context& c, int dropHostIdx, bool dropOOB, const std::vector<int>& idx) {
        c.check(LDSO_API(immature_compact)(c.h, dropHostIdx, dropOOB ? 1 : 0, (int)idx.size(), idx.data(), &size));
        return size;
    }
    void sync(Context& c, std::vector<float>& idepth_min, std::vector<float>& idepth_max, std::vector<float>& quality,
              std::vector<int>& status) {
        idepth_min.resize(size); idepth_max.resize(size); quality.resize(size); status.resize(size);
        if (size) c.check(LDSO_API(immature_get)(c.h, idepth_min.data(), idepth_max.data(), quality.data(), status.data(), nullptr, nullptr));
    }
    // FullSystem::optimizeImmaturePoint for a batch of candidates (FullSystem.cc:1035-1155)
    void optimizeImmaturePoints(Context& c, const std::vector<int>& cand, int minObs, int nFrames, std::vector<float>& idepth,
                                std::vector<int>& ok, std::vector<int>& resState) {
        idepth.resize(cand.size()); ok.resize(cand.size()); resState.resize(cand.size() * (size_t)nFrames);
        if (!cand.empty())
            c.check(LDSO_API(immature_activate)(c.h, (int)cand.size(), cand.data(), minObs, idepth.data(), ok.data(), resState.data()));
    }
};

// ldso::internal::EnergyFunctional + the linearizeAll half of FullSystem (EnergyFunctional.h:63-186)
class EnergyFunctional {
public:
    int nFrames = 0, nPoints = 0, nResiduals = 0;
    std::vector<double> HA, bA, HL, bL, Hsc, bsc, lastX;
    int resInA = 0, resInL = 0;
    int dim() const { return 8 * nFrames + 4; }
    // makeIDX + setPrecalcValues: the flattened window (see include/ldso_gpu.h for the index conventions)
    void setFrames(Context& c, const std::vector<int>& slots, const float* precalc, const float* frameEnergyTH, const float calib[4]) {
        nFrames = (int)slots.size();
        c.check(LDSO_API(window_set_frames)(c.h, nFrames, slots.data(), precalc, frameEnergyTH, calib));
    }
    void setPoints(Context& c, int n, const float* u, const float* v, const float* idepth, const float* idepth_zero, const int* host,
                   const float* color, const float* weights, const float* prior, const int* res_begin) {
        nPoints = n;
        c.check(LDSO_API(window_set_points)(c.h, n, u, v, idepth, idepth_zero, host, color, weights, prior, res_begin));
    }
    void setResiduals(Context& c, int n, const int* target, const int* state, const unsigned char* flags, const float* res_toZeroF) {
        nResiduals = n;
        c.check(LDSO_API(window_set_residuals)(c.h, n, target, state, flags, res_toZeroF));
    }
    // setAdjointsF / setDeltaF (EnergyFunctional.cc:523-609)
    void setAdjointsF(Context& c, const double* adHost, const double* adTarget, const float* adHTdeltaF, const float cDeltaF[4],
                      const double cPrior[4], const double* framePrior, const double* frameDeltaPrior) {
        c.check(LDSO_API(window_set_adjoints)(c.h, adHost, adTarget, adHTdeltaF, cDeltaF, cPrior, framePrior, frameDeltaPrior));
    }
    // Vec3 FullSystem::linearizeAll(bool fixLinearization) (FullSystem.cc:1716-1769) + applyRes_Reductor
    void linearizeAll(Context& c, bool fixLinearization, double stats[3], float* newFrameEnergyTH = nullptr) {
        c.check(LDSO_API(window_linearize)(c.h, fixLinearization ? 1 : 0, stats, newFrameEnergyTH));
    }
    void applyRes(Context& c) { c.check(LDSO_API(window_apply_res)(c.h)); }
    // accumulateAF_MT / accumulateLF_MT / accumulateSCF_MT (EnergyFunctional.cc:670-749)
    void accumulate(Context& c) {
        const size_t D = dim();
        HA.assign(D * D, 0); HL.assign(D * D, 0); Hsc.assign(D * D, 0); bA.assign(D, 0); bL.assign(D, 0); bsc.assign(D, 0);
        int cnt[2];
        c.check(LDSO_API(window_accumulate)(c.h, HA.data(), bA.data(), HL.data(), bL.data(), Hsc.data(), bsc.data(), cnt));
        resInA = cnt[0]; resInL = cnt[1];
    }
    // void solveSystemF(int iteration, double lambda, HCalib, HInertial) (EnergyFunctional.cc:280-471); HM/bM: marginalisation
    // prior, delta: getStitchedDeltaF(), N: nullspace columns (may be null)
    void solveSystemF(Context& c, int iteration, double lambda, const double* HM, const double* bM, const double* delta,
                      const double* N = nullptr, int nNull = 0) {
        accumulate(c);
        lastX.assign(dim(), 0);
        c.check(LDSO_API(window_solve)(c.h, iteration, lambda, HA.data(), bA.data(), HL.data(), bL.data(), Hsc.data(), bsc.data(), HM,
                                       bM, delta, N, nNull, lastX.data()));
    }
    // resubstituteF_MT (EnergyFunctional.cc:611-667): per-point step
    void resubstituteF_MT(Context& c, float lambda, std::vector<float>& step) {
        step.resize(nPoints);
        c.check(LDSO_API(window_resubstitute)(c.h, lastX.data(), lambda, step.data()));
    }
    void setIdepth(Context& c, const float* idepth, const float* idepth_zero) { c.check(LDSO_API(window_set_idepth)(c.h, idepth, idepth_zero)); }
    // void marginalizePointsF() (EnergyFunctional.cc:205-262): HM += margWeightFac * (M - Msc), bM likewise; HM,bM are the caller's
    // dim() x dim() / dim() prior (the reference keeps them as members); marg[nPoints] marks Point::PointStatus::MARGINALIZED
    int resInM = 0;
    void marginalizePointsF(Context& c, const unsigned char* marg, std::vector<double>& HM, std::vector<double>& bM,
                            float idepthFixPriorMargFac = 600.f * 600.f, double margWeightFac = 0.5 * 0.5) {
        const size_t D = dim();
        std::vector<double> H(D * D), b(D);
        int nres = 0;
        c.check(LDSO_API(window_marginalize_points)(c.h, marg, idepthFixPriorMargFac, H.data(), b.data(), &nres));
        resInM += nres;
        for (size_t i = 0; i < D * D; i++) HM[i] += margWeightFac * H[i];
        for (size_t i = 0; i < D; i++) bM[i] += margWeightFac * b[i];
    }
    // void marginalizeFrame(shared_ptr<FrameHessian> fh) (:109-191): fh->idx = idx, fh->prior / fh->delta_prior; shrinks HM,bM
    void marginalizeFrame(Context& c, int idx, const double prior[8], const double deltaPrior[8], std::vector<double>& HM,
                          std::vector<double>& bM) {
        const size_t Dn = dim() - 8;
        std::vector<double> Ho(Dn * Dn), bo(Dn);
        c.check(LDSO_API(window_marginalize_frame)(c.h, nFrames, idx, HM.data(), bM.data(), prior, deltaPrior, Ho.data(), bo.data()));
        HM.swap(Ho); bM.swap(bo);
        nFrames--;
    }
    // double calcLEnergyF_MT() / double calcMEnergyF() (:473-497)
    void calcEnergies(Context& c, const double* HM, const double* bM, const double* delta, double& LEnergy, double& MEnergy) {
        double e[2];
        c.check(LDSO_API(window_calc_energy)(c.h, HM, bM, delta, e));
        LEnergy = e[0]; MEnergy = e[1];
    }
};

}  // namespace ldso_b200
