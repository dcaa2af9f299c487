// oracle/orc_backend.cpp — TEST INFRASTRUCTURE (see orc_common.h).  PARITY UNPINNED.
// CPU restatement of the windowed-BA hot path: PointFrameResidual::linearize / applyRes / takeData /
// fixLinearizationF, setNewFrameEnergyTH, AccumulatedTopHessianSSE::addPoint<0|1> + stitch,
// AccumulatedSCHessianSSE::addPoint + stitch, AccumulatorApprox / AccumulatorXX / AccumulatorX,
// EnergyFunctional::resubstituteFPt and the dense part of solveSystemF.
#include "orc_common.h"

#include <algorithm>

using namespace orc;

static const int patternP[8][2] = {{0, -2}, {-1, -1}, {1, -1}, {-2, 0}, {0, 0}, {2, 0}, {-1, 1}, {0, 2}};   // src/Setting.cc:275
static const float SETTING_INITIAL_CALIB_HESSIAN = 5e9f;
#define CPARS 4

// ================================================================================================
// window description
extern "C" int orc_window_set_frames(orc_ctx* c, int n, const int* slots, const float* precalc, const float* eth,
                                     const float calib[4]) {
    Window& W = c->win;
    if (n < 1 || n > LDSO_MAX_WINDOW) return -2;
    W.n = n;
    for (int i = 0; i < n; i++) { W.slots[i] = slots[i]; W.frameEnergyTH[i] = eth[i]; }
    memcpy(W.precalc, precalc, sizeof(float) * n * n * LDSO_PRECALC_STRIDE);
    W.fxl = calib[0]; W.fyl = calib[1]; W.cxl = calib[2]; W.cyl = calib[3];
    // CalibHessian::setValueScaled (include/internal/CalibHessian.h:93-99)
    W.fxli = 1.0f / W.fxl; W.fyli = 1.0f / W.fyl; W.cxli = -W.cxl / W.fxl; W.cyli = -W.cyl / W.fyl;
    return 0;
}

extern "C" int orc_window_set_points(orc_ctx* c, int np, const float* u, const float* v, const float* idepth,
                                     const float* idepth_zero, const int* host, const float* color, const float* weights,
                                     const float* prior, const int* res_begin) {
    Window& W = c->win;
    W.np = np;
    W.u.assign(u, u + np); W.v.assign(v, v + np); W.idepth.assign(idepth, idepth + np);
    W.idepth_zero.assign(idepth_zero, idepth_zero + np); W.host.assign(host, host + np);
    W.color.assign(color, color + 8 * (size_t)np); W.weights.assign(weights, weights + 8 * (size_t)np);
    W.prior.assign(prior, prior + np); W.res_begin.assign(res_begin, res_begin + np + 1);
    W.deltaF.resize(np);
    for (int i = 0; i < np; i++) W.deltaF[i] = W.idepth[i] - W.idepth_zero[i];   // PointHessian::takeData / setDeltaF
    for (auto* a : {&W.Hdd_accAF, &W.bd_accAF, &W.Hdd_accLF, &W.bd_accLF, &W.HdiF, &W.bdSumF, &W.idepth_hessian,
                    &W.maxRelBaseline, &W.numGoodResiduals, &W.step})
        a->assign(np, 0.f);
    W.Hcd_accAF.assign(4 * (size_t)np, 0.f); W.Hcd_accLF.assign(4 * (size_t)np, 0.f);
    return 0;
}

extern "C" int orc_window_set_idepth(orc_ctx* c, const float* idepth, const float* idepth_zero) {
    Window& W = c->win;
    for (int i = 0; i < W.np; i++) {
        W.idepth[i] = idepth[i]; W.idepth_zero[i] = idepth_zero[i];
        W.deltaF[i] = idepth[i] - idepth_zero[i];
    }
    return 0;
}

extern "C" int orc_window_set_residuals(orc_ctx* c, int nr, const int* target, const int* state, const unsigned char* flags,
                                        const float* res_to_zero) {
    Window& W = c->win;
    if (W.res_begin.empty() || W.res_begin[W.np] != nr) return -2;
    W.nr = nr;
    W.r_target.assign(target, target + nr);
    W.state_state.assign(state, state + nr);
    W.state_NewState.assign(nr, LDSO_RES_OUTLIER);
    W.flags.resize(nr);
    for (int i = 0; i < nr; i++) W.flags[i] = flags[i];
    W.r_point.resize(nr);
    for (int p = 0; p < W.np; p++)
        for (int r = W.res_begin[p]; r < W.res_begin[p + 1]; r++) W.r_point[r] = p;
    W.state_energy.assign(nr, 0); W.state_NewEnergy.assign(nr, 0); W.state_NewEnergyWithOutlier.assign(nr, 0);
    W.J.assign((size_t)nr * LDSO_J_FLOATS, 0.f);
    W.res_toZeroF.assign((size_t)nr * 8, 0.f);
    if (res_to_zero) memcpy(W.res_toZeroF.data(), res_to_zero, sizeof(float) * 8 * nr);
    W.JpJdF.assign((size_t)nr * 8, 0.f);
    W.center.assign((size_t)nr * 3, 0.f);
    return 0;
}

extern "C" int orc_window_set_adjoints(orc_ctx* c, const double* adH, const double* adT, const float* adHTdelta,
                                       const float cDelta[4], const double cPrior[4], const double* fprior,
                                       const double* fdelta_prior) {
    Window& W = c->win;
    int nn = W.n * W.n;
    W.adHost.assign(adH, adH + 64 * nn); W.adTarget.assign(adT, adT + 64 * nn);
    W.adHostF.resize(64 * nn); W.adTargetF.resize(64 * nn);
    for (int i = 0; i < 64 * nn; i++) { W.adHostF[i] = (float)adH[i]; W.adTargetF[i] = (float)adT[i]; }
    W.adHTdeltaF.assign(adHTdelta, adHTdelta + 8 * nn);
    for (int i = 0; i < 4; i++) { W.cDeltaF[i] = cDelta[i]; W.cPrior[i] = cPrior[i]; }
    W.fprior.assign(fprior, fprior + 8 * W.n); W.fdelta_prior.assign(fdelta_prior, fdelta_prior + 8 * W.n);
    return 0;
}

// ================================================================================================
// accumulators
void AccumulatorApprox::initialize() {   // MatrixAccumulators.h:756-769
    memset(Data, 0, sizeof(Data)); memset(Data1k, 0, sizeof(Data1k)); memset(Data1m, 0, sizeof(Data1m));
    memset(TopRight_Data, 0, sizeof(TopRight_Data)); memset(TopRight_Data1k, 0, sizeof(TopRight_Data1k));
    memset(TopRight_Data1m, 0, sizeof(TopRight_Data1m));
    memset(BotRight_Data, 0, sizeof(BotRight_Data)); memset(BotRight_Data1k, 0, sizeof(BotRight_Data1k));
    memset(BotRight_Data1m, 0, sizeof(BotRight_Data1m));
    num = 0; numIn1 = numIn1k = numIn1m = 0;
}
void AccumulatorApprox::shiftUp(bool force) {   // :1063-1100
    if (numIn1 > 1000 || force) {
        for (int i = 0; i < 60; i++) Data1k[i] = Data[i] + Data1k[i];
        for (int i = 0; i < 32; i++) TopRight_Data1k[i] = TopRight_Data[i] + TopRight_Data1k[i];
        for (int i = 0; i < 8; i++) BotRight_Data1k[i] = BotRight_Data[i] + BotRight_Data1k[i];
        numIn1k += numIn1; numIn1 = 0;
        memset(Data, 0, sizeof(Data)); memset(TopRight_Data, 0, sizeof(TopRight_Data)); memset(BotRight_Data, 0, sizeof(BotRight_Data));
    }
    if (numIn1k > 1000 || force) {
        for (int i = 0; i < 60; i++) Data1m[i] = Data1k[i] + Data1m[i];
        for (int i = 0; i < 32; i++) TopRight_Data1m[i] = TopRight_Data1k[i] + TopRight_Data1m[i];
        for (int i = 0; i < 8; i++) BotRight_Data1m[i] = BotRight_Data1k[i] + BotRight_Data1m[i];
        numIn1m += numIn1k; numIn1k = 0;
        memset(Data1k, 0, sizeof(Data1k)); memset(TopRight_Data1k, 0, sizeof(TopRight_Data1k));
        memset(BotRight_Data1k, 0, sizeof(BotRight_Data1k));
    }
}
void AccumulatorApprox::finish() {   // :771-800
    memset(H, 0, sizeof(H));
    shiftUp(true);
    int idx = 0;
    for (int r = 0; r < 10; r++)
        for (int cc = r; cc < 10; cc++) { H[r][cc] = H[cc][r] = Data1m[idx]; idx++; }
    idx = 0;
    for (int r = 0; r < 10; r++)
        for (int cc = 0; cc < 3; cc++) { H[r][cc + 10] = H[cc + 10][r] = TopRight_Data1m[idx]; idx++; }
    H[10][10] = BotRight_Data1m[0];
    H[10][11] = H[11][10] = BotRight_Data1m[1];
    H[10][12] = H[12][10] = BotRight_Data1m[2];
    H[11][11] = BotRight_Data1m[3];
    H[11][12] = H[12][11] = BotRight_Data1m[4];
    H[12][12] = BotRight_Data1m[5];
    num = numIn1 + numIn1k + numIn1m;
}
void AccumulatorApprox::update(const float* x4, const float* x6, const float* y4, const float* y6, float a, float b, float c) {
    // :893-960 — Data[idx] += a*x[c]*x[r] + c*y[c]*y[r] + b*(x[c]*y[r] + y[c]*x[r]) over r<=c of the 10-vector [x4|x6]
    float x[10], y[10];
    for (int i = 0; i < 4; i++) { x[i] = x4[i]; y[i] = y4[i]; }
    for (int i = 0; i < 6; i++) { x[4 + i] = x6[i]; y[4 + i] = y6[i]; }
    int idx = 0;
    for (int r = 0; r < 10; r++)
        for (int cc = r; cc < 10; cc++) {
            Data[idx] += a * x[cc] * x[r] + c * y[cc] * y[r] + b * (x[cc] * y[r] + y[cc] * x[r]);
            idx++;
        }
    num++;
    numIn1++;
    shiftUp(false);
}
void AccumulatorApprox::updateTopRight(const float* x4, const float* x6, const float* y4, const float* y6, float TR00, float TR10,
                                       float TR01, float TR11, float TR02, float TR12) {   // :962-1002
    float x[10], y[10];
    for (int i = 0; i < 4; i++) { x[i] = x4[i]; y[i] = y4[i]; }
    for (int i = 0; i < 6; i++) { x[4 + i] = x6[i]; y[4 + i] = y6[i]; }
    for (int r = 0; r < 10; r++) {
        TopRight_Data[3 * r + 0] += x[r] * TR00 + y[r] * TR10;
        TopRight_Data[3 * r + 1] += x[r] * TR01 + y[r] * TR11;
        TopRight_Data[3 * r + 2] += x[r] * TR02 + y[r] * TR12;
    }
}
void AccumulatorApprox::updateBotRight(float a00, float a01, float a02, float a11, float a12, float a22) {   // :1004-1017
    BotRight_Data[0] += a00; BotRight_Data[1] += a01; BotRight_Data[2] += a02;
    BotRight_Data[3] += a11; BotRight_Data[4] += a12; BotRight_Data[5] += a22;
}

void AccumulatorMat::initialize(int r, int c) {
    rows = r; cols = c;
    A.assign(r * c, 0.f); A1k.assign(r * c, 0.f); A1m.assign(r * c, 0.f);
    num = 0; numIn1 = numIn1k = numIn1m = 0;
}
void AccumulatorMat::shiftUp(bool force) {
    if (numIn1 > 1000 || force) {
        for (size_t i = 0; i < A.size(); i++) { A1k[i] += A[i]; A[i] = 0; }
        numIn1k += numIn1; numIn1 = 0;
    }
    if (numIn1k > 1000 || force) {
        for (size_t i = 0; i < A.size(); i++) { A1m[i] += A1k[i]; A1k[i] = 0; }
        numIn1m += numIn1k; numIn1k = 0;
    }
}
void AccumulatorMat::update(const float* L, const float* R, float w) {
    // Eigen evaluates `w * L * R.transpose()` as (w*L) * R^T
    if (R) {
        for (int i = 0; i < rows; i++) {
            float wl = w * L[i];
            for (int j = 0; j < cols; j++) A[i * cols + j] += wl * R[j];
        }
    } else {
        for (int i = 0; i < rows; i++) A[i] += w * L[i];
    }
    numIn1++;
    shiftUp(false);
}
void AccumulatorMat::finish() {
    shiftUp(true);
    num = numIn1 + numIn1k + numIn1m;
}

// ================================================================================================
// K7 — PointFrameResidual::linearize (src/internal/Residuals.cc:15-217)
static inline void interp33(const float* mat, float x, float y, int width, float* out) {   // GlobalFuncs.h:89-103
    int ix = (int)x, iy = (int)y;
    float dx = x - ix, dy = y - iy, dxdy = dx * dy;
    const float* bp = mat + 3 * (ix + iy * width);
    for (int k = 0; k < 3; k++)
        out[k] = dxdy * bp[3 * (1 + width) + k] + (dy - dxdy) * bp[3 * width + k] + (dx - dxdy) * bp[3 + k] +
                 (1 - dx - dy + dxdy) * bp[k];
}

static double linearize_one(orc_ctx* c, int r) {
    Window& W = c->win;
    W.state_NewEnergyWithOutlier[r] = -1;
    if (W.state_state[r] == LDSO_RES_OOB) { W.state_NewState[r] = LDSO_RES_OOB; return W.state_energy[r]; }
    const int p = W.r_point[r], h = W.host[p], t = W.r_target[r];
    const float* pc = W.precalc[h + W.n * t];
    const float* R0 = pc;        // PRE_RTll_0
    const float* t0 = pc + 9;    // PRE_tTll_0
    const float* KRKi = pc + 12; // PRE_KRKiTll
    const float* Kt = pc + 21;   // PRE_KtTll
    const float affLL0 = pc[24], affLL1 = pc[25], b0 = pc[26];
    const float* color = &W.color[8 * p];
    const float* weights = &W.weights[8 * p];
    const float* dIl = c->frames[W.slots[t]].dIp[0].data();
    const float wM3G = c->wG[0] - 3, hM3G = c->hG[0] - 3;
    float* J = &W.J[(size_t)r * LDSO_J_FLOATS];
    float energyLeft = 0;
    float d_xi_x[6], d_xi_y[6], d_C_x[4], d_C_y[4], d_d_x, d_d_y;
    {
        // projectPoint (ResidualProjections.h:57-84) at the FEJ point
        float KliP[3] = {(W.u[p] + 0 - W.cxl) * W.fxli, (W.v[p] + 0 - W.cyl) * W.fyli, 1};
        const float idz = W.idepth_zero[p];
        float ptp[3] = {R0[0] * KliP[0] + R0[1] * KliP[1] + R0[2] * KliP[2] + t0[0] * idz,
                        R0[3] * KliP[0] + R0[4] * KliP[1] + R0[5] * KliP[2] + t0[1] * idz,
                        R0[6] * KliP[0] + R0[7] * KliP[1] + R0[8] * KliP[2] + t0[2] * idz};
        float drescale = 1.0f / ptp[2];
        float new_idepth = idz * drescale;
        if (!(drescale > 0)) { W.state_NewState[r] = LDSO_RES_OOB; return W.state_energy[r]; }
        float u = ptp[0] * drescale, v = ptp[1] * drescale;
        float Ku = u * W.fxl + W.cxl, Kv = v * W.fyl + W.cyl;
        if (!(Ku > 1.1f && Kv > 1.1f && Ku < wM3G && Kv < hM3G)) { W.state_NewState[r] = LDSO_RES_OOB; return W.state_energy[r]; }
        W.center[3 * r] = Ku; W.center[3 * r + 1] = Kv; W.center[3 * r + 2] = new_idepth;
        d_d_x = drescale * (t0[0] - t0[2] * u) * SCALE_IDEPTH * W.fxl;
        d_d_y = drescale * (t0[1] - t0[2] * v) * SCALE_IDEPTH * W.fyl;
        d_C_x[2] = drescale * (R0[6] * u - R0[0]);
        d_C_x[3] = W.fxl * drescale * (R0[7] * u - R0[1]) * W.fyli;
        d_C_x[0] = KliP[0] * d_C_x[2];
        d_C_x[1] = KliP[1] * d_C_x[3];
        d_C_y[2] = W.fyl * drescale * (R0[6] * v - R0[3]) * W.fxli;
        d_C_y[3] = drescale * (R0[7] * v - R0[4]);
        d_C_y[0] = KliP[0] * d_C_y[2];
        d_C_y[1] = KliP[1] * d_C_y[3];
        d_C_x[0] = (d_C_x[0] + u) * SCALE_F;
        d_C_x[1] *= SCALE_F;
        d_C_x[2] = (d_C_x[2] + 1) * SCALE_C;
        d_C_x[3] *= SCALE_C;
        d_C_y[0] *= SCALE_F;
        d_C_y[1] = (d_C_y[1] + v) * SCALE_F;
        d_C_y[2] *= SCALE_C;
        d_C_y[3] = (d_C_y[3] + 1) * SCALE_C;
        d_xi_x[0] = new_idepth * W.fxl;
        d_xi_x[1] = 0;
        d_xi_x[2] = -new_idepth * u * W.fxl;
        d_xi_x[3] = -u * v * W.fxl;
        d_xi_x[4] = (1 + u * u) * W.fxl;
        d_xi_x[5] = -v * W.fxl;
        d_xi_y[0] = 0;
        d_xi_y[1] = new_idepth * W.fyl;
        d_xi_y[2] = -new_idepth * v * W.fyl;
        d_xi_y[3] = -(1 + v * v) * W.fyl;
        d_xi_y[4] = u * v * W.fyl;
        d_xi_y[5] = u * W.fyl;
    }
    for (int i = 0; i < 6; i++) { J[40 + i] = d_xi_x[i]; J[46 + i] = d_xi_y[i]; }
    for (int i = 0; i < 4; i++) { J[52 + i] = d_C_x[i]; J[56 + i] = d_C_y[i]; }
    J[60] = d_d_x; J[61] = d_d_y;
    float JIdxJIdx_00 = 0, JIdxJIdx_11 = 0, JIdxJIdx_10 = 0;
    float JabJIdx_00 = 0, JabJIdx_01 = 0, JabJIdx_10 = 0, JabJIdx_11 = 0;
    float JabJab_00 = 0, JabJab_01 = 0, JabJab_11 = 0;
    float wJI2_sum = 0;
    const float huberTH = c->cfg.huber_th, thsc = c->cfg.outlier_th_sum_component;
    for (int idx = 0; idx < 8; idx++) {
        // projectPoint (ResidualProjections.h:24-33) at the current state
        float up = W.u[p] + patternP[idx][0], vp = W.v[p] + patternP[idx][1];
        float ptp[3] = {KRKi[0] * up + KRKi[1] * vp + KRKi[2] * 1 + Kt[0] * W.idepth[p],
                        KRKi[3] * up + KRKi[4] * vp + KRKi[5] * 1 + Kt[1] * W.idepth[p],
                        KRKi[6] * up + KRKi[7] * vp + KRKi[8] * 1 + Kt[2] * W.idepth[p]};
        float Ku = ptp[0] / ptp[2], Kv = ptp[1] / ptp[2];
        if (!(Ku > 1.1f && Kv > 1.1f && Ku < wM3G && Kv < hM3G)) { W.state_NewState[r] = LDSO_RES_OOB; return W.state_energy[r]; }
        float hitColor[3];
        interp33(dIl, Ku, Kv, c->wG[0], hitColor);
        float residual = hitColor[0] - (float)(affLL0 * color[idx] + affLL1);
        float drdA = (color[idx] - b0);
        if (!std::isfinite((float)hitColor[0])) { W.state_NewState[r] = LDSO_RES_OOB; return W.state_energy[r]; }
        float w = sqrtf(thsc / (thsc + (hitColor[1] * hitColor[1] + hitColor[2] * hitColor[2])));
        w = 0.5f * (w + weights[idx]);
        float hw = fabsf(residual) < huberTH ? 1 : huberTH / fabsf(residual);
        energyLeft += w * w * hw * residual * residual * (2 - hw);
        {
            if (hw < 1) hw = sqrtf(hw);
            hw = hw * w;
            hitColor[1] *= hw;
            hitColor[2] *= hw;
            J[idx] = residual * hw;
            J[8 + idx] = hitColor[1];
            J[16 + idx] = hitColor[2];
            J[24 + idx] = drdA * hw;
            J[32 + idx] = hw;
            JIdxJIdx_00 += hitColor[1] * hitColor[1];
            JIdxJIdx_11 += hitColor[2] * hitColor[2];
            JIdxJIdx_10 += hitColor[1] * hitColor[2];
            JabJIdx_00 += drdA * hw * hitColor[1];
            JabJIdx_01 += drdA * hw * hitColor[2];
            JabJIdx_10 += hw * hitColor[1];
            JabJIdx_11 += hw * hitColor[2];
            JabJab_00 += drdA * drdA * hw * hw;
            JabJab_01 += drdA * hw * hw;
            JabJab_11 += hw * hw;
            wJI2_sum += hw * hw * (hitColor[1] * hitColor[1] + hitColor[2] * hitColor[2]);
            if (c->cfg.affine_opt_mode_a < 0) J[24 + idx] = 0;
            if (c->cfg.affine_opt_mode_b < 0) J[32 + idx] = 0;
        }
    }
    J[62] = JIdxJIdx_00; J[63] = JIdxJIdx_10; J[64] = JIdxJIdx_11;
    J[65] = JabJIdx_00; J[66] = JabJIdx_01; J[67] = JabJIdx_10; J[68] = JabJIdx_11;
    J[69] = JabJab_00; J[70] = JabJab_01; J[71] = JabJab_11;
    W.state_NewEnergyWithOutlier[r] = energyLeft;
    float th = std::max<float>(W.frameEnergyTH[h], W.frameEnergyTH[t]);
    if (energyLeft > th || wJI2_sum < 2) {
        energyLeft = th;
        W.state_NewState[r] = LDSO_RES_OUTLIER;
    } else {
        W.state_NewState[r] = LDSO_RES_IN;
    }
    W.state_NewEnergy[r] = energyLeft;
    return energyLeft;
}

// PointFrameResidual::takeData (include/internal/Residuals.h:124-129)
static void take_data(Window& W, int r) {
    const float* J = &W.J[(size_t)r * LDSO_J_FLOATS];
    float* o = &W.JpJdF[8 * (size_t)r];
    // JIdx2 * Jpdd, JIdx2 = [[00, 10],[10, 11]]
    float a0 = J[62] * J[60] + J[63] * J[61];
    float a1 = J[63] * J[60] + J[64] * J[61];
    for (int i = 0; i < 6; i++) o[i] = J[40 + i] * a0 + J[46 + i] * a1;
    o[6] = J[65] * J[60] + J[66] * J[61];
    o[7] = J[67] * J[60] + J[68] * J[61];
}

// PointFrameResidual::applyRes (include/internal/Residuals.h:70-88)
static void apply_res(Window& W, int r, bool copyJacobians) {
    if (copyJacobians) {
        if (W.state_state[r] == LDSO_RES_OOB) return;
        if (W.state_NewState[r] == LDSO_RES_IN) {
            W.flags[r] |= LDSO_RF_ACTIVE;
            take_data(W, r);
        } else {
            W.flags[r] &= ~LDSO_RF_ACTIVE;
        }
    }
    W.state_state[r] = W.state_NewState[r];
    W.state_energy[r] = W.state_NewEnergy[r];
}

// FullSystem::setNewFrameEnergyTH (src/frontend/FullSystem.cc:2078-2109)
static float new_frame_energy_th(orc_ctx* c) {
    Window& W = c->win;
    std::vector<float> all;
    const int newest = W.n - 1;
    for (int r = 0; r < W.nr; r++)
        if (!(W.flags[r] & LDSO_RF_LINEARIZED) && W.state_NewEnergyWithOutlier[r] >= 0 && W.r_target[r] == newest)
            all.push_back((float)W.state_NewEnergyWithOutlier[r]);
    if (all.empty()) return 12 * 12 * 8;
    int nthIdx = c->cfg.frame_energy_th_n * all.size();
    std::nth_element(all.begin(), all.begin() + nthIdx, all.end());
    float nthElement = sqrtf(all[nthIdx]);
    float th = nthElement * c->cfg.frame_energy_th_fac_median;
    th = 26.0f * c->cfg.frame_energy_th_const_weight + th * (1 - c->cfg.frame_energy_th_const_weight);
    th = th * th;
    th *= c->cfg.overall_energy_th_weight * c->cfg.overall_energy_th_weight;
    return th;
}

// FullSystem::linearizeAll + linearizeAll_Reductor (src/frontend/FullSystem.cc:1716-1823), single thread
extern "C" int orc_window_linearize(orc_ctx* c, int fix, double stats[3], float* new_th) {
    Window& W = c->win;
    double E = 0, num = 0;
    const int T = orc_threads(c);
    // FullSystem::linearizeAll_Reductor splits activeResiduals over the worker threads (FullSystem.cc:1716-1824); here the
    // split is by point so that the per-point bookkeeping of the fix pass stays race-free
#pragma omp parallel for num_threads(T) if (T > 1) schedule(static) reduction(+ : E, num)
    for (int p = 0; p < W.np; p++)
    for (int r = W.res_begin[p]; r < W.res_begin[p + 1]; r++) {
        if (W.flags[r] & LDSO_RF_LINEARIZED) continue;   // activeResiduals = !isLinearized (FullSystem.cc:864)
        E += linearize_one(c, r);
        if (W.state_NewState[r] == LDSO_RES_IN) num += 1;
        if (fix) {
            apply_res(W, r, true);
            if ((W.flags[r] & LDSO_RF_ACTIVE) && (W.flags[r] & LDSO_RF_NEW)) {
                int p = W.r_point[r];
                const float* pc = W.precalc[W.host[p] + W.n * W.r_target[r]];
                const float* KRKi = pc + 12; const float* Kt = pc + 21;
                float inf[3] = {KRKi[0] * W.u[p] + KRKi[1] * W.v[p] + KRKi[2] * 1, KRKi[3] * W.u[p] + KRKi[4] * W.v[p] + KRKi[5] * 1,
                                KRKi[6] * W.u[p] + KRKi[7] * W.v[p] + KRKi[8] * 1};
                float ptp[3] = {inf[0] + Kt[0] * W.idepth[p], inf[1] + Kt[1] * W.idepth[p], inf[2] + Kt[2] * W.idepth[p]};
                float d0 = inf[0] / inf[2] - ptp[0] / ptp[2], d1 = inf[1] / inf[2] - ptp[1] / ptp[2];
                float relBS = 0.01 * std::sqrt(d0 * d0 + d1 * d1);
                if (relBS > W.maxRelBaseline[p]) W.maxRelBaseline[p] = relBS;
                W.numGoodResiduals[p] += 1;
            }
        }
    }
    float th = new_frame_energy_th(c);
    W.frameEnergyTH[W.n - 1] = th;
    if (new_th) *new_th = th;
    stats[0] = E; stats[1] = 0; stats[2] = num;
    return 0;
}

extern "C" int orc_window_apply_res(orc_ctx* c) {
    Window& W = c->win;
    for (int r = 0; r < W.nr; r++)
        if (!(W.flags[r] & LDSO_RF_LINEARIZED)) apply_res(W, r, true);
    return 0;
}

extern "C" int orc_window_reset_oob(orc_ctx* c) {   // include/internal/Residuals.h:63-67
    Window& W = c->win;
    for (int r = 0; r < W.nr; r++)
        if (!(W.flags[r] & LDSO_RF_LINEARIZED)) {
            W.state_NewEnergy[r] = W.state_energy[r] = 0;
            W.state_NewState[r] = LDSO_RES_OUTLIER;
            W.state_state[r] = LDSO_RES_IN;
        }
    return 0;
}

// PointFrameResidual::fixLinearizationF (src/internal/Residuals.cc:219-245)
extern "C" int orc_window_fix_linearization(orc_ctx* c, const unsigned char* mask) {
    Window& W = c->win;
    for (int r = 0; r < W.nr; r++) {
        if (!mask[r]) continue;
        const int p = W.r_point[r];
        const float* J = &W.J[(size_t)r * LDSO_J_FLOATS];
        const float* dp = &W.adHTdeltaF[8 * (W.host[p] + W.n * W.r_target[r])];
        float dx = 0, dy = 0, cx = 0, cy = 0;
        for (int i = 0; i < 6; i++) { dx += J[40 + i] * dp[i]; dy += J[46 + i] * dp[i]; }
        for (int i = 0; i < 4; i++) { cx += J[52 + i] * W.cDeltaF[i]; cy += J[56 + i] * W.cDeltaF[i]; }
        float Jp_delta_x = dx + cx + J[60] * W.deltaF[p];
        float Jp_delta_y = dy + cy + J[61] * W.deltaF[p];
        for (int i = 0; i < 8; i++) {
            float rtz = J[i];
            rtz = rtz - J[8 + i] * Jp_delta_x;
            rtz = rtz - J[16 + i] * Jp_delta_y;
            rtz = rtz - J[24 + i] * dp[6];
            rtz = rtz - J[32 + i] * dp[7];
            W.res_toZeroF[8 * (size_t)r + i] = rtz;
        }
        W.flags[r] |= LDSO_RF_LINEARIZED;
    }
    return 0;
}

// ================================================================================================
// K8 — AccumulatedTopHessianSSE::addPoint<mode> (src/internal/OptimizationBackend/AccumulatedTopHessian.cc:9-118)
static void top_add_point(orc_ctx* c, int mode, int p, std::vector<AccumulatorApprox>& acc, int& nres) {
    Window& W = c->win;
    const float* dc = W.cDeltaF;
    float dd = W.deltaF[p];
    float bd_acc = 0, Hdd_acc = 0, Hcd_acc[4] = {0, 0, 0, 0};
    for (int r = W.res_begin[p]; r < W.res_begin[p + 1]; r++) {
        const bool lin = W.flags[r] & LDSO_RF_LINEARIZED, act = W.flags[r] & LDSO_RF_ACTIVE;
        if (mode == 0 && (lin || !act)) continue;
        if (mode == 1 && (!lin || !act)) continue;
        if (mode == 2 && !act) continue;
        const float* J = &W.J[(size_t)r * LDSO_J_FLOATS];
        int htIDX = W.host[p] + W.r_target[r] * W.n;
        const float* dp = &W.adHTdeltaF[8 * htIDX];
        float resApprox[8];
        if (mode == 0) for (int i = 0; i < 8; i++) resApprox[i] = J[i];
        if (mode == 2) for (int i = 0; i < 8; i++) resApprox[i] = W.res_toZeroF[8 * (size_t)r + i];
        if (mode == 1) {
            float dx = 0, dy = 0, cx = 0, cy = 0;
            for (int i = 0; i < 6; i++) { dx += J[40 + i] * dp[i]; dy += J[46 + i] * dp[i]; }
            for (int i = 0; i < 4; i++) { cx += J[52 + i] * dc[i]; cy += J[56 + i] * dc[i]; }
            float Jp_delta_x = dx + cx + J[60] * dd, Jp_delta_y = dy + cy + J[61] * dd;
            for (int i = 0; i < 8; i++) {
                float rtz = W.res_toZeroF[8 * (size_t)r + i];
                rtz = rtz + J[8 + i] * Jp_delta_x;
                rtz = rtz + J[16 + i] * Jp_delta_y;
                rtz = rtz + J[24 + i] * dp[6];
                rtz = rtz + J[32 + i] * dp[7];
                resApprox[i] = rtz;
            }
        }
        float JI_r[2] = {0, 0}, Jab_r[2] = {0, 0}, rr = 0;
        for (int i = 0; i < 8; i++) {
            JI_r[0] += resApprox[i] * J[8 + i];
            JI_r[1] += resApprox[i] * J[16 + i];
            Jab_r[0] += resApprox[i] * J[24 + i];
            Jab_r[1] += resApprox[i] * J[32 + i];
            rr += resApprox[i] * resApprox[i];
        }
        AccumulatorApprox& A = acc[htIDX];
        A.update(J + 52, J + 40, J + 56, J + 46, J[62], J[63], J[64]);
        A.updateBotRight(J[69], J[70], Jab_r[0], J[71], Jab_r[1], rr);
        A.updateTopRight(J + 52, J + 40, J + 56, J + 46, J[65], J[66], J[67], J[68], JI_r[0], JI_r[1]);
        float Ji2_Jpdd[2] = {J[62] * J[60] + J[63] * J[61], J[63] * J[60] + J[64] * J[61]};
        bd_acc += JI_r[0] * J[60] + JI_r[1] * J[61];
        Hdd_acc += Ji2_Jpdd[0] * J[60] + Ji2_Jpdd[1] * J[61];
        for (int i = 0; i < 4; i++) Hcd_acc[i] += J[52 + i] * Ji2_Jpdd[0] + J[56 + i] * Ji2_Jpdd[1];
        nres++;
    }
    if (mode == 0) {
        W.Hdd_accAF[p] = Hdd_acc; W.bd_accAF[p] = bd_acc;
        for (int i = 0; i < 4; i++) W.Hcd_accAF[4 * p + i] = Hcd_acc[i];
    }
    if (mode == 1 || mode == 2) {
        W.Hdd_accLF[p] = Hdd_acc; W.bd_accLF[p] = bd_acc;
        for (int i = 0; i < 4; i++) W.Hcd_accLF[4 * p + i] = Hcd_acc[i];
    }
    if (mode == 2) {
        W.Hdd_accAF[p] = 0; W.bd_accAF[p] = 0;
        for (int i = 0; i < 4; i++) W.Hcd_accAF[4 * p + i] = 0;
    }
}

// small dense helpers (row-major double)
static void mm(const double* A, const double* B, double* C, int n, int k, int m, bool tB = false) {   // C[n x m] = A[n x k] * B
    for (int i = 0; i < n; i++)
        for (int j = 0; j < m; j++) {
            double s = 0;
            for (int l = 0; l < k; l++) s += A[i * k + l] * (tB ? B[j * k + l] : B[l * m + j]);
            C[i * m + j] = s;
        }
}

// AccumulatedTopHessianSSE::stitchDoubleInternal + stitchDoubleMT tail (AccumulatedTopHessian.cc:193-255, .h:64-105)
static void top_stitch(orc_ctx* c, std::vector<AccumulatorApprox>& acc, bool usePrior, double* H, double* b) {
    Window& W = c->win;
    const int n = W.n, D = 8 * n + CPARS;
    std::fill(H, H + D * D, 0.0);
    std::fill(b, b + D, 0.0);
    auto Hb = [&](int r, int cc) -> double& { return H[r * D + cc]; };
    for (int k = 0; k < n * n; k++) {
        int h = k % n, t = k / n;
        int hIdx = CPARS + h * 8, tIdx = CPARS + t * 8;
        acc[k].finish();
        double accH[13][13];
        for (int i = 0; i < 13; i++)
            for (int j = 0; j < 13; j++) accH[i][j] = acc[k].num == 0 ? 0.0 : (double)acc[k].H[i][j];
        const double* AH = &W.adHost[64 * k];
        const double* AT = &W.adTarget[64 * k];
        double M88[64], M8C[32], v8[8], T1[64], T2[64];
        for (int i = 0; i < 8; i++) {
            for (int j = 0; j < 8; j++) M88[i * 8 + j] = accH[CPARS + i][CPARS + j];
            for (int j = 0; j < 4; j++) M8C[i * 4 + j] = accH[CPARS + i][j];
            v8[i] = accH[CPARS + i][8 + CPARS];
        }
        mm(AH, M88, T1, 8, 8, 8);
        mm(T1, AH, T2, 8, 8, 8, true);
        for (int i = 0; i < 8; i++) for (int j = 0; j < 8; j++) Hb(hIdx + i, hIdx + j) += T2[i * 8 + j];
        mm(T1, AT, T2, 8, 8, 8, true);
        for (int i = 0; i < 8; i++) for (int j = 0; j < 8; j++) Hb(hIdx + i, tIdx + j) += T2[i * 8 + j];
        mm(AT, M88, T1, 8, 8, 8);
        mm(T1, AT, T2, 8, 8, 8, true);
        for (int i = 0; i < 8; i++) for (int j = 0; j < 8; j++) Hb(tIdx + i, tIdx + j) += T2[i * 8 + j];
        double T3[32];
        mm(AH, M8C, T3, 8, 8, 4);
        for (int i = 0; i < 8; i++) for (int j = 0; j < 4; j++) Hb(hIdx + i, j) += T3[i * 4 + j];
        mm(AT, M8C, T3, 8, 8, 4);
        for (int i = 0; i < 8; i++) for (int j = 0; j < 4; j++) Hb(tIdx + i, j) += T3[i * 4 + j];
        for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) Hb(i, j) += accH[i][j];
        double w8[8];
        mm(AH, v8, w8, 8, 8, 1);
        for (int i = 0; i < 8; i++) b[hIdx + i] += w8[i];
        mm(AT, v8, w8, 8, 8, 1);
        for (int i = 0; i < 8; i++) b[tIdx + i] += w8[i];
        for (int i = 0; i < 4; i++) b[i] += accH[i][8 + CPARS];
    }
    if (usePrior) {
        for (int i = 0; i < 4; i++) { Hb(i, i) += W.cPrior[i]; b[i] += W.cPrior[i] * (double)W.cDeltaF[i]; }
        for (int h = 0; h < n; h++)
            for (int i = 0; i < 8; i++) {
                Hb(CPARS + h * 8 + i, CPARS + h * 8 + i) += W.fprior[8 * h + i];
                b[CPARS + h * 8 + i] += W.fprior[8 * h + i] * W.fdelta_prior[8 * h + i];
            }
    }
    for (int h = 0; h < n; h++) {   // make diagonal by copying over parts (.h:93-103)
        int hIdx = CPARS + h * 8;
        for (int i = 0; i < 4; i++) for (int j = 0; j < 8; j++) Hb(i, hIdx + j) = Hb(hIdx + j, i);
        for (int t = h + 1; t < n; t++) {
            int tIdx = CPARS + t * 8;
            for (int i = 0; i < 8; i++) for (int j = 0; j < 8; j++) Hb(hIdx + i, tIdx + j) += Hb(tIdx + j, hIdx + i);
            for (int i = 0; i < 8; i++) for (int j = 0; j < 8; j++) Hb(tIdx + i, hIdx + j) = Hb(hIdx + j, tIdx + i);
        }
    }
}

// K9 — AccumulatedSCHessianSSE::addPoint (AccumulatedSCHessian.cc:9-51)
struct SCAcc {
    std::vector<AccumulatorMat> accE, accEB, accD;
    AccumulatorMat accHcc, accbc;
};
static void sc_add_point(orc_ctx* c, int p, bool shiftPriorToZero, SCAcc& S) {
    Window& W = c->win;
    const int n = W.n;
    int ngoodres = 0;
    for (int r = W.res_begin[p]; r < W.res_begin[p + 1]; r++) if (W.flags[r] & LDSO_RF_ACTIVE) ngoodres++;
    if (ngoodres == 0) { W.HdiF[p] = 0; W.bdSumF[p] = 0; W.idepth_hessian[p] = 0; W.maxRelBaseline[p] = 0; return; }
    float H = W.Hdd_accAF[p] + W.Hdd_accLF[p] + W.prior[p];
    if (H < 1e-10) H = 1e-10;
    W.idepth_hessian[p] = H;
    W.HdiF[p] = 1.0 / H;
    W.bdSumF[p] = W.bd_accAF[p] + W.bd_accLF[p];
    if (shiftPriorToZero) W.bdSumF[p] += W.prior[p] * W.deltaF[p];
    float Hcd[4];
    for (int i = 0; i < 4; i++) Hcd[i] = W.Hcd_accAF[4 * p + i] + W.Hcd_accLF[4 * p + i];
    S.accHcc.update(Hcd, Hcd, W.HdiF[p]);
    S.accbc.update(Hcd, nullptr, W.bdSumF[p] * W.HdiF[p]);
    const int nFrames2 = n * n;
    for (int r1 = W.res_begin[p]; r1 < W.res_begin[p + 1]; r1++) {
        if (!(W.flags[r1] & LDSO_RF_ACTIVE)) continue;
        int r1ht = W.host[p] + W.r_target[r1] * n;
        for (int r2 = W.res_begin[p]; r2 < W.res_begin[p + 1]; r2++) {
            if (!(W.flags[r2] & LDSO_RF_ACTIVE)) continue;
            S.accD[r1ht + W.r_target[r2] * nFrames2].update(&W.JpJdF[8 * (size_t)r1], &W.JpJdF[8 * (size_t)r2], W.HdiF[p]);
        }
        S.accE[r1ht].update(&W.JpJdF[8 * (size_t)r1], Hcd, W.HdiF[p]);
        S.accEB[r1ht].update(&W.JpJdF[8 * (size_t)r1], nullptr, W.HdiF[p] * W.bdSumF[p]);
    }
}

// AccumulatedSCHessianSSE::stitchDoubleInternal + MT tail (AccumulatedSCHessian.cc:53-119, .h:64-100)
static void sc_stitch(orc_ctx* c, SCAcc& S, double* H, double* b) {
    Window& W = c->win;
    const int nf = W.n, D = 8 * nf + CPARS, nframes2 = nf * nf;
    std::fill(H, H + D * D, 0.0);
    std::fill(b, b + D, 0.0);
    auto Hb = [&](int r, int cc) -> double& { return H[r * D + cc]; };
    for (int k = 0; k < nframes2; k++) {
        int i = k % nf, j = k / nf;
        int iIdx = CPARS + i * 8, jIdx = CPARS + j * 8, ijIdx = i + nf * j;
        S.accE[ijIdx].finish(); S.accEB[ijIdx].finish();
        double Hpc[32], bp[8];
        for (int q = 0; q < 32; q++) Hpc[q] = S.accE[ijIdx].A1m[q];
        for (int q = 0; q < 8; q++) bp[q] = S.accEB[ijIdx].A1m[q];
        const double* AHij = &W.adHost[64 * ijIdx];
        const double* ATij = &W.adTarget[64 * ijIdx];
        double T3[32], w8[8];
        mm(AHij, Hpc, T3, 8, 8, 4);
        for (int a = 0; a < 8; a++) for (int q = 0; q < 4; q++) Hb(iIdx + a, q) += T3[a * 4 + q];
        mm(ATij, Hpc, T3, 8, 8, 4);
        for (int a = 0; a < 8; a++) for (int q = 0; q < 4; q++) Hb(jIdx + a, q) += T3[a * 4 + q];
        mm(AHij, bp, w8, 8, 8, 1);
        for (int a = 0; a < 8; a++) b[iIdx + a] += w8[a];
        mm(ATij, bp, w8, 8, 8, 1);
        for (int a = 0; a < 8; a++) b[jIdx + a] += w8[a];
        for (int kk = 0; kk < nf; kk++) {
            int kIdx = CPARS + kk * 8, ijkIdx = ijIdx + kk * nframes2, ikIdx = i + nf * kk;
            S.accD[ijkIdx].finish();
            if (S.accD[ijkIdx].num == 0) continue;
            double Dm[64], T1[64], T2[64];
            for (int q = 0; q < 64; q++) Dm[q] = S.accD[ijkIdx].A1m[q];
            const double* AHik = &W.adHost[64 * ikIdx];
            const double* ATik = &W.adTarget[64 * ikIdx];
            mm(AHij, Dm, T1, 8, 8, 8);
            mm(T1, AHik, T2, 8, 8, 8, true);
            for (int a = 0; a < 8; a++) for (int q = 0; q < 8; q++) Hb(iIdx + a, iIdx + q) += T2[a * 8 + q];
            mm(T1, ATik, T2, 8, 8, 8, true);
            for (int a = 0; a < 8; a++) for (int q = 0; q < 8; q++) Hb(iIdx + a, kIdx + q) += T2[a * 8 + q];
            mm(ATij, Dm, T1, 8, 8, 8);
            mm(T1, ATik, T2, 8, 8, 8, true);
            for (int a = 0; a < 8; a++) for (int q = 0; q < 8; q++) Hb(jIdx + a, kIdx + q) += T2[a * 8 + q];
            mm(T1, AHik, T2, 8, 8, 8, true);
            for (int a = 0; a < 8; a++) for (int q = 0; q < 8; q++) Hb(jIdx + a, iIdx + q) += T2[a * 8 + q];
        }
    }
    S.accHcc.finish(); S.accbc.finish();
    for (int a = 0; a < 4; a++) {
        for (int q = 0; q < 4; q++) Hb(a, q) += S.accHcc.A1m[a * 4 + q];
        b[a] += S.accbc.A1m[a];
    }
    // copy transposed parts (AccumulatedSCHessian.h:91-99)
    for (int h = 0; h < nf; h++) {
        int hIdx = CPARS + h * 8;
        for (int a = 0; a < 4; a++) for (int q = 0; q < 8; q++) Hb(a, hIdx + q) = Hb(hIdx + q, a);
    }
}

// EnergyFunctional::accumulateAF_MT / LF_MT / SCF_MT, MT=false paths (EnergyFunctional.cc:670-749)
// Per-thread accumulators merged before stitching, as the reference's accSSE_top_A[tid] / accE[tid] are
// (AccumulatedTopHessian.cc:193-213, AccumulatedSCHessian.cc:53-75).  Only the timing build is threaded.
static void merge_into(AccumulatorApprox& d, AccumulatorApprox& s) {
    d.finish(); s.finish();
    for (int i = 0; i < 60; i++) d.Data1m[i] += s.Data1m[i];
    for (int i = 0; i < 32; i++) d.TopRight_Data1m[i] += s.TopRight_Data1m[i];
    for (int i = 0; i < 8; i++) d.BotRight_Data1m[i] += s.BotRight_Data1m[i];
    d.numIn1m += s.num;
}
static void merge_into(AccumulatorMat& d, AccumulatorMat& s) {
    d.finish(); s.finish();
    for (size_t i = 0; i < d.A1m.size(); i++) d.A1m[i] += s.A1m[i];
    d.numIn1m += s.num;
}

extern "C" int orc_window_accumulate(orc_ctx* c, double* HA, double* bA, double* HL, double* bL, double* Hsc, double* bsc,
                                     int counts[2]) {
    Window& W = c->win;
    const int n = W.n;
    const int T = orc_threads(c);
    std::vector<std::vector<AccumulatorApprox>> accA(T, std::vector<AccumulatorApprox>(n * n)), accL = accA;
    std::vector<SCAcc> S(T);
    std::vector<int> nA(T, 0), nL(T, 0);
#pragma omp parallel num_threads(T) if (T > 1)
    {
        const int tid = orc_thread_id();
        for (auto& a : accA[tid]) a.initialize();
        for (auto& a : accL[tid]) a.initialize();
        SCAcc& St = S[tid];
        St.accE.resize(n * n); St.accEB.resize(n * n); St.accD.resize(n * n * n);
        for (auto& a : St.accE) a.initialize(8, 4);
        for (auto& a : St.accEB) a.initialize(8, 1);
        for (auto& a : St.accD) a.initialize(8, 8);
        St.accHcc.initialize(4, 4); St.accbc.initialize(4, 1);
        const int p0 = (int)((long long)W.np * tid / T), p1 = (int)((long long)W.np * (tid + 1) / T);
        for (int p = p0; p < p1; p++) top_add_point(c, 0, p, accA[tid], nA[tid]);
        for (int p = p0; p < p1; p++) top_add_point(c, 1, p, accL[tid], nL[tid]);
        for (int p = p0; p < p1; p++) sc_add_point(c, p, true, St);
    }
    int nresA = nA[0], nresL = nL[0];
    for (int t = 1; t < T; t++) {
        for (int k = 0; k < n * n; k++) { merge_into(accA[0][k], accA[t][k]); merge_into(accL[0][k], accL[t][k]); }
        for (int k = 0; k < n * n; k++) { merge_into(S[0].accE[k], S[t].accE[k]); merge_into(S[0].accEB[k], S[t].accEB[k]); }
        for (int k = 0; k < n * n * n; k++) merge_into(S[0].accD[k], S[t].accD[k]);
        merge_into(S[0].accHcc, S[t].accHcc); merge_into(S[0].accbc, S[t].accbc);
        nresA += nA[t]; nresL += nL[t];
    }
    top_stitch(c, accA[0], false, HA, bA);
    top_stitch(c, accL[0], true, HL, bL);
    sc_stitch(c, S[0], Hsc, bsc);
    if (counts) { counts[0] = nresA; counts[1] = nresL; }
    return 0;
}

// EnergyFunctional::orthogonalize (EnergyFunctional.cc:809-841), b-only form used by solveSystemF
static void orthogonalize_x(int D, const double* Nin, int k, double* x) {
    std::vector<double> N(D * k), U(D * k), S(k), V(k * k);
    for (int j = 0; j < k; j++) {
        double nrm = 0;
        for (int i = 0; i < D; i++) nrm += Nin[i * k + j] * Nin[i * k + j];
        nrm = std::sqrt(nrm);
        for (int i = 0; i < D; i++) N[i * k + j] = Nin[i * k + j] / nrm;
    }
    ldso_sm::jacobi_svd(D, k, N.data(), U.data(), S.data(), V.data());
    double maxSv = 0;
    for (int i = 0; i < k; i++) maxSv = std::max(maxSv, S[i]);
    for (int i = 0; i < k; i++) S[i] = (S[i] > 0.00001 * maxSv) ? 1.0 / S[i] : 0;   // setting_solverModeDelta
    std::vector<double> Npi(D * k, 0.0);   // U S V^T
    for (int i = 0; i < D; i++)
        for (int j = 0; j < k; j++) {
            double s = 0;
            for (int l = 0; l < k; l++) s += U[i * k + l] * S[l] * V[j * k + l];
            Npi[i * k + j] = s;
        }
    // NNpiTS = 0.5 (N Npi^T + Npi N^T);  x -= NNpiTS x
    std::vector<double> a(k), bb(k);
    for (int j = 0; j < k; j++) {
        double s1 = 0, s2 = 0;
        for (int i = 0; i < D; i++) { s1 += Npi[i * k + j] * x[i]; s2 += N[i * k + j] * x[i]; }
        a[j] = s1; bb[j] = s2;
    }
    for (int i = 0; i < D; i++) {
        double s = 0;
        for (int j = 0; j < k; j++) s += 0.5 * (N[i * k + j] * a[j] + Npi[i * k + j] * bb[j]);
        x[i] -= s;
    }
}

// dense part of EnergyFunctional::solveSystemF (EnergyFunctional.cc:310-432), setting_vi_enable=false,
// setting_solverMode = SOLVER_FIX_LAMBDA | SOLVER_ORTHOGONALIZE_X_LATER (src/Setting.cc:23)
extern "C" int orc_window_solve(orc_ctx* c, int iteration, double lambda, const double* HA, const double* bA, const double* HL,
                                const double* bL, const double* Hsc, const double* bsc, const double* HM, const double* bM,
                                const double* delta, const double* N, int n_null, double* x) {
    const int D = 8 * c->win.n + CPARS;
    lambda = 1e-5;   // SOLVER_FIX_LAMBDA (:283)
    std::vector<double> Hf(D * D), bf(D), bMtop(D);
    for (int i = 0; i < D; i++) {
        double s = bM[i];
        for (int j = 0; j < D; j++) s += HM[i * D + j] * delta[j];
        bMtop[i] = s;
    }
    // upper triangle = HL + HM + HA, diag *= (1+lambda), upper -= Hsc * (1.0f/(1+lambda)); then selfadjointView<Upper>
    const double isc = (double)(1.0f / (1 + lambda));
    for (int i = 0; i < D; i++) {
        bf[i] = bL[i] + bMtop[i] + bA[i] - bsc[i] / (1 + lambda);
        for (int j = i; j < D; j++) {
            double v = HL[i * D + j] + HM[i * D + j] + HA[i * D + j];
            if (i == j) v *= (1 + lambda);
            v -= Hsc[i * D + j] * isc;
            Hf[i * D + j] = Hf[j * D + i] = v;
        }
    }
    std::vector<double> SVecI(D), Hs(D * D), bs(D), y(D);
    for (int i = 0; i < D; i++) SVecI[i] = 1.0 / std::sqrt(Hf[i * D + i] + 10);
    for (int i = 0; i < D; i++) {
        bs[i] = SVecI[i] * bf[i];
        for (int j = 0; j < D; j++) Hs[i * D + j] = SVecI[i] * Hf[i * D + j] * SVecI[j];
    }
    ldso_sm::ldlt_solve(D, Hs.data(), bs.data(), y.data());
    for (int i = 0; i < D; i++) x[i] = SVecI[i] * y[i];
    if (iteration >= 2 && N && n_null > 0) orthogonalize_x(D, N, n_null, x);
    return 0;
}

// K10 — EnergyFunctional::resubstituteF_MT / resubstituteFPt (EnergyFunctional.cc:611-667)
extern "C" int orc_window_resubstitute(orc_ctx* c, const double* x, float lambda, float* step_out) {
    Window& W = c->win;
    const int n = W.n, D = 8 * n + CPARS;
    std::vector<float> xF(D);
    for (int i = 0; i < D; i++) xF[i] = (float)x[i];
    std::vector<float> xAd(8 * n * n);
    for (int h = 0; h < n; h++)
        for (int t = 0; t < n; t++)
            for (int j = 0; j < 8; j++) {
                float s1 = 0, s2 = 0;
                for (int i = 0; i < 8; i++) {
                    s1 += xF[CPARS + 8 * h + i] * W.adHostF[64 * (h + n * t) + i * 8 + j];
                    s2 += xF[CPARS + 8 * t + i] * W.adTargetF[64 * (h + n * t) + i * 8 + j];
                }
                xAd[8 * (n * h + t) + j] = s1 + s2;
            }
    for (int p = 0; p < W.np; p++) {
        int ngoodres = 0;
        for (int r = W.res_begin[p]; r < W.res_begin[p + 1]; r++) if (W.flags[r] & LDSO_RF_ACTIVE) ngoodres++;
        if (ngoodres == 0) { W.step[p] = 0; continue; }
        float b = W.bdSumF[p];
        float d = 0;
        for (int i = 0; i < 4; i++) d += xF[i] * (W.Hcd_accAF[4 * p + i] + W.Hcd_accLF[4 * p + i]);
        b -= d;
        for (int r = W.res_begin[p]; r < W.res_begin[p + 1]; r++) {
            if (!(W.flags[r] & LDSO_RF_ACTIVE)) continue;
            const float* xa = &xAd[8 * (W.host[p] * n + W.r_target[r])];
            float s = 0;
            for (int i = 0; i < 8; i++) s += xa[i] * W.JpJdF[8 * (size_t)r + i];
            b -= s;
        }
        if (!std::isfinite(b)) break;   // the reference returns from the whole loop (:660-662)
        W.step[p] = -b * W.HdiF[p] / (1 + (double)lambda);   // currentLambda is a double in the reference
    }
    if (step_out) memcpy(step_out, W.step.data(), sizeof(float) * W.np);
    return 0;
}

// ================================================================================================
// read-backs
extern "C" int orc_window_get_residual_floats(orc_ctx* c, int what, float* dst) {
    Window& W = c->win;
    switch (what) {
        case 0: memcpy(dst, W.J.data(), sizeof(float) * W.J.size()); break;
        case 1: memcpy(dst, W.res_toZeroF.data(), sizeof(float) * W.res_toZeroF.size()); break;
        case 2: memcpy(dst, W.JpJdF.data(), sizeof(float) * W.JpJdF.size()); break;
        case 3: memcpy(dst, W.center.data(), sizeof(float) * W.center.size()); break;
        case 4: for (int i = 0; i < W.nr; i++) dst[i] = (float)W.state_NewEnergy[i]; break;
        case 5: for (int i = 0; i < W.nr; i++) dst[i] = (float)W.state_NewEnergyWithOutlier[i]; break;
        case 6: for (int i = 0; i < W.nr; i++) dst[i] = (float)W.state_energy[i]; break;
        default: return -2;
    }
    return 0;
}
extern "C" int orc_window_get_residual_ints(orc_ctx* c, int what, int* dst) {
    Window& W = c->win;
    const std::vector<int>& s = what == 0 ? W.state_state : what == 1 ? W.state_NewState : W.flags;
    memcpy(dst, s.data(), sizeof(int) * W.nr);
    return 0;
}
extern "C" int orc_window_get_point_floats(orc_ctx* c, int what, float* dst) {
    Window& W = c->win;
    const std::vector<float>* s[] = {&W.Hdd_accAF, &W.bd_accAF, &W.Hcd_accAF, &W.Hdd_accLF, &W.bd_accLF, &W.Hcd_accLF, &W.HdiF,
                                     &W.bdSumF, &W.idepth_hessian, &W.maxRelBaseline, &W.numGoodResiduals, &W.step};
    if (what < 0 || what > 11) return -2;
    memcpy(dst, s[what]->data(), sizeof(float) * s[what]->size());
    return 0;
}

// everything FullSystem reads after linearizeAll(true) in one call (the GPU library answers it with one device->host copy)
extern "C" int orc_window_get_keyframe_results(orc_ctx* c, int* state, float* center3, float* hdi, float* idepth_hessian,
                                               float* num_good, float* max_rel_baseline) {
    int rc = 0;
    if (state) rc |= orc_window_get_residual_ints(c, 0, state);
    if (center3) rc |= orc_window_get_residual_floats(c, 3, center3);
    if (hdi) rc |= orc_window_get_point_floats(c, 6, hdi);
    if (idepth_hessian) rc |= orc_window_get_point_floats(c, 8, idepth_hessian);
    if (num_good) rc |= orc_window_get_point_floats(c, 10, num_good);
    if (max_rel_baseline) rc |= orc_window_get_point_floats(c, 9, max_rel_baseline);
    return rc;
}

// ================================================================================================
// EnergyFunctional::marginalizePointsF (EnergyFunctional.cc:205-262), single-threaded as in the reference
extern "C" int orc_window_marginalize_points(orc_ctx* c, const unsigned char* marg, float fac, double* H, double* b, int* n_res) {
    Window& W = c->win;
    const int n = W.n, D = 8 * n + CPARS;
    std::vector<AccumulatorApprox> acc(n * n);
    for (auto& a : acc) a.initialize();
    SCAcc S;
    S.accE.resize(n * n); S.accEB.resize(n * n); S.accD.resize(n * n * n);
    for (auto& a : S.accE) a.initialize(8, 4);
    for (auto& a : S.accEB) a.initialize(8, 1);
    for (auto& a : S.accD) a.initialize(8, 8);
    S.accHcc.initialize(4, 4); S.accbc.initialize(4, 1);
    int nres = 0;
    for (int p = 0; p < W.np; p++) {
        if (!marg[p]) continue;
        W.prior[p] *= fac;                       // p->priorF *= setting_idepthFixPriorMargFac (:217)
        top_add_point(c, 2, p, acc, nres);       // :235
        sc_add_point(c, p, false, S);            // :236
        for (int r = W.res_begin[p]; r < W.res_begin[p + 1]; r++) W.flags[r] &= ~LDSO_RF_ACTIVE;   // removePoint (:193-203)
    }
    std::vector<double> M((size_t)D * D), Mb(D), Msc((size_t)D * D), Mbsc(D);
    top_stitch(c, acc, false, M.data(), Mb.data());   // stitchDouble(M, Mb, this, false, false) (:242)
    sc_stitch(c, S, Msc.data(), Mbsc.data());
    for (size_t i = 0; i < (size_t)D * D; i++) H[i] = M[i] - Msc[i];
    for (int i = 0; i < D; i++) b[i] = Mb[i] - Mbsc[i];
    if (n_res) *n_res = nres;
    return 0;
}

// FullSystem::flagPointsForRemoval's inner loop (src/frontend/FullSystem.cc:1390-1399) followed by
// EnergyFunctional::marginalizePointsF (:205-262) for the masked points of the resident window, literally: resetOOB,
// linearize, isLinearized = false, applyRes(true), fixLinearizationF for the active ones, then addPoint<2> + the Schur accumulator.
// residual_dead marks the residuals FullSystem has dropped since the upload (linearizeAll(true)'s toRemove list).
extern "C" int orc_window_marginalize_subset(orc_ctx* c, const unsigned char* point_mask, const unsigned char* residual_dead,
                                             const float* frame_energy_th, float fac, double* H, double* b, int* n_res, int* n_lin) {
    Window& W = c->win;
    for (int i = 0; i < W.n; i++) W.frameEnergyTH[i] = frame_energy_th[i];
    std::vector<unsigned char> fix(std::max(W.nr, 1), 0);
    int nlin = 0;
    for (int p = 0; p < W.np; p++) {
        if (!point_mask[p]) continue;
        for (int r = W.res_begin[p]; r < W.res_begin[p + 1]; r++) {
            W.flags[r] &= ~LDSO_RF_LINEARIZED;
            if (residual_dead[r]) { W.flags[r] &= ~LDSO_RF_ACTIVE; continue; }
            W.state_NewEnergy[r] = W.state_energy[r] = 0;      // resetOOB
            W.state_NewState[r] = LDSO_RES_OUTLIER;
            W.state_state[r] = LDSO_RES_IN;
            linearize_one(c, r);
            apply_res(W, r, true);
            if (W.flags[r] & LDSO_RF_ACTIVE) { fix[r] = 1; nlin++; }
        }
    }
    orc_window_fix_linearization(c, fix.data());
    if (n_lin) *n_lin = nlin;
    return orc_window_marginalize_points(c, point_mask, fac, H, b, n_res);
}

// util::MatrixInverter::invertPosDef(M, fast = false) (src/util/MatrixInverter.cc:31-50)
static void invert_pos_def(int n, const double* M, double* res) {
    std::vector<double> sv(n), Ms((size_t)n * n), U((size_t)n * n), S(n), V((size_t)n * n);
    for (int i = 0; i < n; i++) sv[i] = 1.0 / std::sqrt(std::fabs(M[i * n + i]) + 10);
    for (int i = 0; i < n; i++)
        for (int j = 0; j < n; j++) {
            const double m = i <= j ? M[i * n + j] : M[j * n + i];   // selfadjointView<Upper>
            Ms[i * n + j] = sv[i] * m * sv[j];
        }
    ldso_sm::jacobi_svd(n, n, Ms.data(), U.data(), S.data(), V.data());
    for (int i = 0; i < n; i++) S[i] = S[i] <= 0 ? 0 : 1 / S[i];
    for (int i = 0; i < n; i++)
        for (int j = 0; j < n; j++) {
            double s = 0;
            for (int l = 0; l < n; l++) s += V[i * n + l] * S[l] * U[j * n + l];
            res[i * n + j] = sv[i] * s * sv[j];
        }
    for (int i = 0; i < n; i++)
        for (int j = 0; j < i; j++) res[i * n + j] = res[j * n + i];   // read back through selfadjointView<Upper>
}

// EnergyFunctional::marginalizeFrame (EnergyFunctional.cc:109-191): the moves are done block by block as there
extern "C" int orc_window_marginalize_frame(orc_ctx*, int nFrames, int idx, const double* HMin, const double* bMin,
                                            const double prior[8], const double delta_prior[8], double* HM_out, double* bM_out) {
    const int ndim = nFrames * 8 + CPARS - 8, odim = nFrames * 8 + CPARS;
    std::vector<double> HM(HMin, HMin + (size_t)odim * odim), bM(bMin, bMin + odim);
    auto at = [&](std::vector<double>& A, int r, int cc) -> double& { return A[(size_t)r * odim + cc]; };
    if (idx != nFrames - 1) {
        const int io = idx * 8 + CPARS, ntail = 8 * (nFrames - idx - 1);
        std::vector<double> bTmp(bM.begin() + io, bM.begin() + io + 8), tail(bM.end() - ntail, bM.end());
        for (int i = 0; i < ntail; i++) bM[io + i] = tail[i];
        for (int i = 0; i < 8; i++) bM[odim - 8 + i] = bTmp[i];
        std::vector<double> T = HM;                                   // columns
        for (int r = 0; r < odim; r++) {
            for (int i = 0; i < ntail; i++) at(HM, r, io + i) = at(T, r, odim - ntail + i);
            for (int i = 0; i < 8; i++) at(HM, r, odim - 8 + i) = at(T, r, io + i);
        }
        T = HM;                                                       // rows
        for (int cc = 0; cc < odim; cc++) {
            for (int i = 0; i < ntail; i++) at(HM, io + i, cc) = at(T, odim - ntail + i, cc);
            for (int i = 0; i < 8; i++) at(HM, odim - 8 + i, cc) = at(T, io + i, cc);
        }
    }
    for (int i = 0; i < 8; i++) {
        at(HM, ndim + i, ndim + i) += prior[i];
        bM[ndim + i] += prior[i] * delta_prior[i];
    }
    std::vector<double> SVec(odim), SVecI(odim);
    for (int i = 0; i < odim; i++) { SVec[i] = std::sqrt(std::fabs(at(HM, i, i)) + 10); SVecI[i] = 1 / SVec[i]; }
    std::vector<double> HS((size_t)odim * odim), bS(odim);
    for (int i = 0; i < odim; i++) {
        bS[i] = SVecI[i] * bM[i];
        for (int j = 0; j < odim; j++) at(HS, i, j) = SVecI[i] * at(HM, i, j) * SVecI[j];
    }
    double hp[64], hpi[64];
    for (int i = 0; i < 8; i++) for (int j = 0; j < 8; j++) hp[i * 8 + j] = at(HS, ndim + i, ndim + j);
    invert_pos_def(8, hp, hpi);
    std::vector<double> BL((size_t)8 * ndim), BLt((size_t)ndim * 8), bli((size_t)ndim * 8), prod((size_t)ndim * ndim), pb(ndim);
    for (int i = 0; i < 8; i++) for (int j = 0; j < ndim; j++) { BL[(size_t)i * ndim + j] = at(HS, ndim + i, j); BLt[(size_t)j * 8 + i] = BL[(size_t)i * ndim + j]; }
    mm(BLt.data(), hpi, bli.data(), ndim, 8, 8);
    mm(bli.data(), BL.data(), prod.data(), ndim, 8, ndim);
    mm(bli.data(), &bS[ndim], pb.data(), ndim, 8, 1);
    for (int i = 0; i < ndim; i++) {
        bS[i] -= pb[i];
        for (int j = 0; j < ndim; j++) at(HS, i, j) -= prod[(size_t)i * ndim + j];
    }
    for (int i = 0; i < odim; i++) {
        bS[i] *= SVec[i];
        for (int j = 0; j < odim; j++) at(HS, i, j) *= SVec[i] * SVec[j];
    }
    for (int i = 0; i < ndim; i++) {
        bM_out[i] = bS[i];
        for (int j = 0; j < ndim; j++) HM_out[(size_t)i * ndim + j] = 0.5 * (at(HS, i, j) + at(HS, j, i));
    }
    return 0;
}

// EnergyFunctional::calcLEnergyF_MT / calcLEnergyPt (:481-497, :751-806) with Accumulator11
// (MatrixAccumulators.h:68-142), and calcMEnergyF (:473-479)
namespace {
struct Acc11 {
    float A = 0, d[4] = {0, 0, 0, 0}, d1k[4] = {0, 0, 0, 0}, d1m[4] = {0, 0, 0, 0}, n1 = 0, n1k = 0, n1m = 0;
    void shiftUp(bool force) {
        if (n1 > 1000 || force) { for (int i = 0; i < 4; i++) { d1k[i] = d[i] + d1k[i]; d[i] = 0; } n1k += n1; n1 = 0; }
        if (n1k > 1000 || force) { for (int i = 0; i < 4; i++) { d1m[i] = d1k[i] + d1m[i]; d1k[i] = 0; } n1m += n1k; n1k = 0; }
    }
    void sse_noshift(const float* v) { for (int i = 0; i < 4; i++) d[i] += v[i]; n1++; }
    void single(float v) { d[0] += v; n1++; shiftUp(false); }
    void finish() { shiftUp(true); A = d1m[0] + d1m[1] + d1m[2] + d1m[3]; }
};
}  // namespace

extern "C" int orc_window_calc_energy(orc_ctx* c, const double* HM, const double* bM, const double* delta, double energy[2]) {
    Window& W = c->win;
    const int n = W.n, D = 8 * n + CPARS;
    double E = 0;
    for (int i = 0; i < 8 * n; i++) E += W.fdelta_prior[i] * W.fprior[i] * W.fdelta_prior[i];
    float ec = 0;
    for (int i = 0; i < 4; i++) ec += W.cDeltaF[i] * (float)W.cPrior[i] * W.cDeltaF[i];
    E += ec;
    Acc11 A;
    const float* dc = W.cDeltaF;
    for (int p = 0; p < W.np; p++) {
        const float dd = W.deltaF[p];
        for (int r = W.res_begin[p]; r < W.res_begin[p + 1]; r++) {
            if (!(W.flags[r] & LDSO_RF_LINEARIZED) || !(W.flags[r] & LDSO_RF_ACTIVE)) continue;
            const float* J = &W.J[(size_t)r * LDSO_J_FLOATS];
            const float* dp = &W.adHTdeltaF[8 * (W.host[p] + n * W.r_target[r])];
            float dx = 0, dy = 0, cx = 0, cy = 0;
            for (int i = 0; i < 6; i++) { dx += J[40 + i] * dp[i]; dy += J[46 + i] * dp[i]; }
            for (int i = 0; i < 4; i++) { cx += J[52 + i] * dc[i]; cy += J[56 + i] * dc[i]; }
            const float Jx = dx + cx + J[60] * dd, Jy = dy + cy + J[61] * dd;
            for (int i = 0; i + 3 < 8; i += 4) {
                float v[4];
                for (int l = 0; l < 4; l++) {
                    float Jd = J[8 + i + l] * Jx;
                    Jd = Jd + J[16 + i + l] * Jy;
                    Jd = Jd + J[24 + i + l] * dp[6];
                    Jd = Jd + J[32 + i + l] * dp[7];
                    float r0 = W.res_toZeroF[8 * (size_t)r + i + l];
                    r0 = r0 + r0;
                    r0 = r0 + Jd;
                    v[l] = Jd * r0;
                }
                A.sse_noshift(v);
            }
        }
        A.single(dd * dd * W.prior[p]);
    }
    A.finish();
    energy[0] = E + A.A;
    double em = 0;
    for (int i = 0; i < D; i++) {
        double s = 2 * bM[i];
        for (int j = 0; j < D; j++) s += HM[(size_t)i * D + j] * delta[j];
        em += delta[i] * s;
    }
    energy[1] = em;
    return 0;
}
