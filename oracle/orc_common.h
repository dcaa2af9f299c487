// oracle/orc_common.h — TEST INFRASTRUCTURE, not product code.
//
// CPU restatement of the LDSO hot path (reference: n-lalanne/LDSO, CPU/SSE only).  It exists to
// (a) check the CUDA path in tests/ and __graft_entry__.smoke(), and (b) be timed as the CPU baseline
// by bench.py.  Nothing under ldso_b200/ may include, link or call it.
//
// PARITY UNPINNED: the reference ships no tests, golden vectors or fixtures for this path
//  and cannot be compiled here (Eigen/glog/OpenCV/DBoW3 absent).  Correctness of
// this restatement is argued by line-by-line correspondence (every function cites the reference
// file:line it follows), numpy/scipy cross-checks of the small dense algebra, and end-to-end
// behaviour on synthetic sequences with known ground truth.
//
// Every exported function has the same signature as its ldso_gpu_* counterpart in
// include/ldso_gpu.h with the prefix orc_ and ctx type orc_ctx, so the parity tests drive both
// libraries through one harness.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <string>
#include <vector>
#include <xmmintrin.h>

#include "../include/ldso_gpu.h"
#include "../include/ldso_smallmat.h"

namespace orc {

// include/Settings.h:28-35
constexpr float SCALE_IDEPTH = 1.0f, SCALE_XI_ROT = 1.0f, SCALE_XI_TRANS = 0.5f, SCALE_F = 50.0f, SCALE_C = 50.0f,
                SCALE_W = 1.0f, SCALE_A = 10.0f, SCALE_B = 1000.0f;

struct Frame {
    // FrameHessian::dIp / absSquaredGrad (include/internal/FrameHessian.h)
    std::vector<float> dIp[LDSO_MAX_LEVELS];   // 3 floats per pixel {I,dx,dy}
    std::vector<float> absSquaredGrad[LDSO_MAX_LEVELS];
    bool valid = false;
};

// include/internal/OptimizationBackend/MatrixAccumulators.h:1104-1643
struct Accumulator9 {
    float H[9][9];
    size_t num;
    alignas(16) float SSEData[4 * 45], SSEData1k[4 * 45], SSEData1m[4 * 45];
    float numIn1, numIn1k, numIn1m;
    void initialize();
    void finish();
    void updateSSE_eighted(const __m128* J, __m128 w);   // J[0..8]
    void updateSSE(const __m128* J);
    void updateSingle(const float* J, int off = 0);
    void updateSingleWeighted(const float* J, float w, int off = 0);
    void shiftUp(bool force);
};

struct Tracker {
    // CoarseTracker members (include/frontend/CoarseTracker.h)
    int w[LDSO_MAX_LEVELS], h[LDSO_MAX_LEVELS];
    float fx[LDSO_MAX_LEVELS], fy[LDSO_MAX_LEVELS], cx[LDSO_MAX_LEVELS], cy[LDSO_MAX_LEVELS];
    float K[LDSO_MAX_LEVELS][9], Ki[LDSO_MAX_LEVELS][9];
    std::vector<float> idepth[LDSO_MAX_LEVELS], weightSums[LDSO_MAX_LEVELS], weightSums_bak[LDSO_MAX_LEVELS];
    std::vector<float> pc_u[LDSO_MAX_LEVELS], pc_v[LDSO_MAX_LEVELS], pc_idepth[LDSO_MAX_LEVELS], pc_color[LDSO_MAX_LEVELS];
    int pc_n[LDSO_MAX_LEVELS];
    float *buf_warped_idepth, *buf_warped_u, *buf_warped_v, *buf_warped_dx, *buf_warped_dy, *buf_warped_residual,
        *buf_warped_weight, *buf_warped_refColor;
    std::vector<float> buf_store;
    int buf_warped_n = 0;
    int ref_slot = -1;
    float ref_exposure = 1, ref_aff_a = 0, ref_aff_b = 0;   // lastRef->ab_exposure, lastRef_aff_g2l
    int new_slot = -1;
    float new_exposure = 1;
    Accumulator9 acc;
};


// ---- windowed BA state (host structs of the reference flattened; see include/ldso_gpu.h) ----
// include/internal/OptimizationBackend/MatrixAccumulators.h:749-1102
struct AccumulatorApprox {
    float H[13][13];
    size_t num;
    alignas(16) float Data[60], Data1k[60], Data1m[60];
    alignas(16) float TopRight_Data[32], TopRight_Data1k[32], TopRight_Data1m[32];
    alignas(16) float BotRight_Data[8], BotRight_Data1k[8], BotRight_Data1m[8];
    float numIn1, numIn1k, numIn1m;
    void initialize();
    void finish();
    void update(const float* x4, const float* x6, const float* y4, const float* y6, float a, float b, float c);
    void updateTopRight(const float* x4, const float* x6, const float* y4, const float* y6, float TR00, float TR10, float TR01,
                        float TR11, float TR02, float TR12);
    void updateBotRight(float a00, float a01, float a02, float a11, float a12, float a22);
    void shiftUp(bool force);
};
// MatrixAccumulators.h:20-64 (AccumulatorXX<i,j>) and :150-198 (AccumulatorX<i>), sizes fixed at run time
struct AccumulatorMat {
    int rows = 0, cols = 0;
    std::vector<float> A, A1k, A1m;
    size_t num = 0;
    float numIn1 = 0, numIn1k = 0, numIn1m = 0;
    void initialize(int r, int c);
    void update(const float* L, const float* R, float w);   // A += w * L * R^T   (cols==1, R==nullptr: A += w*L)
    void finish();
    void shiftUp(bool force);
};

struct Window {
    int n = 0;
    int slots[LDSO_MAX_WINDOW];
    float precalc[LDSO_MAX_WINDOW * LDSO_MAX_WINDOW][LDSO_PRECALC_STRIDE];
    float frameEnergyTH[LDSO_MAX_WINDOW];
    float fxl, fyl, cxl, cyl, fxli, fyli, cxli, cyli;
    // points
    int np = 0;
    std::vector<float> u, v, idepth, idepth_zero, prior, color, weights, deltaF;
    std::vector<int> host, res_begin;
    std::vector<float> Hdd_accAF, bd_accAF, Hcd_accAF, Hdd_accLF, bd_accLF, Hcd_accLF, HdiF, bdSumF, idepth_hessian,
        maxRelBaseline, numGoodResiduals, step;
    // residuals
    int nr = 0;
    std::vector<int> r_point, r_target, state_state, state_NewState, flags;
    std::vector<double> state_energy, state_NewEnergy, state_NewEnergyWithOutlier;
    std::vector<float> J, res_toZeroF, JpJdF, center;
    // adjoints / deltas
    std::vector<double> adHost, adTarget, fprior, fdelta_prior;
    std::vector<float> adHostF, adTargetF, adHTdeltaF;
    float cDeltaF[4] = {0, 0, 0, 0};
    double cPrior[4] = {0, 0, 0, 0};
};

}  // namespace orc

struct orc_ctx {
    ldso_gpu_config cfg;
    int levels;
    int wG[LDSO_MAX_LEVELS], hG[LDSO_MAX_LEVELS];
    float B[256];
    std::vector<orc::Frame> frames;
    orc::Tracker trk[LDSO_NUM_TRACKERS];
    orc::Window win;
    std::shared_ptr<void> selector, immature, initializer, undistort;   // per-ctx state of orc_selector.cpp / orc_immature.cpp / orc_initializer.cpp (die with the ctx)
    std::string err;
    long long launches = 0;
    int threads = 1;   // worker threads of the timing build (env ORC_THREADS at create; the reference uses NUM_THREADS = 6)
};

#ifdef _OPENMP
#include <omp.h>
inline int orc_threads(const orc_ctx* c) { return c->threads > 1 ? c->threads : 1; }
inline int orc_thread_id() { return omp_get_thread_num(); }
#else
inline int orc_threads(const orc_ctx*) { return 1; }
inline int orc_thread_id() { return 0; }
#endif
