// oracle/orc_initializer.cpp — TEST INFRASTRUCTURE, not product code (see orc_common.h).
//
// CPU restatement of CoarseInitializer (reference src/frontend/CoarseInitializer.cc, include/frontend/CoarseInitializer.h)
// and makePixelStatus / gridMaxSelection (include/frontend/PixelSelector2.h:60-278): SURVEY.md section 8f row N1.
// PARITY UNPINNED like the rest of the oracle.  Two documented departures from a literal transcription:
//   * the 10-NN / parent graph comes from include/ldso_knn.h (exact search, ties to the lower index) because nanoflann's
//     order among equidistant points is an implementation detail of its kd-tree;
//   * the 6x6 / 8x8 solve of trackFrame (:104-111, Eigen float LDLT) goes through ldso_sm::ldlt_solve in double.
#include "orc_common.h"
#include "../include/ldso_knn.h"

#include <algorithm>

using ldso_sm::SE3;
using orc::Accumulator9;

extern "C" int orc_select_make_maps(orc_ctx* c, int slot, float density, int recursions_left, float th_factor, int* potential_io,
                                    float* map_out, int* n_selected);

namespace {

const int patternP[8][2] = {{0, -2}, {-1, -1}, {1, -1}, {-2, 0}, {0, 0}, {2, 0}, {-1, 1}, {0, 2}};   // src/Setting.cc:275
constexpr int patternNum = 8, patternPadding = 2;

struct Pnt {   // CoarseInitializer.h:23-63
    float u, v, idepth;
    bool isGood;
    float energy[2];
    bool isGood_new;
    float idepth_new, energy_new[2], iR, iRSumNum, lastHessian, lastHessian_new, maxstep;
    int parent;
    float parentDist;
    int neighbours[10];
    float neighboursDist[10];
    float my_type, outlierTH;
};

// MatrixAccumulators.h:68-142
struct Accumulator11 {
    float A, S[4], S1k[4], S1m[4], numIn1, numIn1k, numIn1m;
    size_t num;
    void initialize() { A = 0; for (int i = 0; i < 4; i++) S[i] = S1k[i] = S1m[i] = 0; num = 0; numIn1 = numIn1k = numIn1m = 0; }
    void shiftUp(bool force) {
        if (numIn1 > 1000 || force) { for (int i = 0; i < 4; i++) { S1k[i] = S[i] + S1k[i]; S[i] = 0; } numIn1k += numIn1; numIn1 = 0; }
        if (numIn1k > 1000 || force) { for (int i = 0; i < 4; i++) { S1m[i] = S1k[i] + S1m[i]; S1k[i] = 0; } numIn1m += numIn1k; numIn1k = 0; }
    }
    void updateSingle(float v) { S[0] += v; num++; numIn1++; shiftUp(false); }
    void finish() { shiftUp(true); A = S1m[0] + S1m[1] + S1m[2] + S1m[3]; }
};

struct Init {
    int levels = 0;
    int w[LDSO_MAX_LEVELS], h[LDSO_MAX_LEVELS];
    double fx[LDSO_MAX_LEVELS], fy[LDSO_MAX_LEVELS], cx[LDSO_MAX_LEVELS], cy[LDSO_MAX_LEVELS];
    double Ki[LDSO_MAX_LEVELS][9];
    std::vector<Pnt> points[LDSO_MAX_LEVELS];
    std::vector<float> Jb, Jb_new;   // Vec10f per pixel slot (sized for level 0)
    int first_slot = -1, new_slot = -1;
    SE3 thisToNext;
    double aff_a = 0, aff_b = 0;     // thisToNext_aff
    bool snapped = false, fixAffine = true;
    int frameID = -1, snappedAt = 0;
    float alphaK, alphaW, regWeight, couplingWeight;
    Accumulator9 acc9, acc9SC;
    Accumulator11 accE, accEAlpha;
    int sparsityFactor = 5;          // file-static of PixelSelector2.h
};

Init* get(orc_ctx* c) {
    if (!c->initializer) c->initializer = std::make_shared<Init>();
    return (Init*)c->initializer.get();
}

// PixelSelector2.h:63-148 / :151-236 (the template and the run-time pot variant are the same loop)
int gridMaxSelection(const float* grads, unsigned char* map_out, int w, int h, int pot, float THFac) {
    memset(map_out, 0, (size_t)w * h);
    int numGood = 0;
    for (int y = 1; y < h - pot; y += pot)
        for (int x = 1; x < w - pot; x += pot) {
            int bestXXID = -1, bestYYID = -1, bestXYID = -1, bestYXID = -1;
            float bestXX = 0, bestYY = 0, bestXY = 0, bestYX = 0;
            const float* grads0 = grads + 3 * (x + y * w);
            for (int dx = 0; dx < pot; dx++)
                for (int dy = 0; dy < pot; dy++) {
                    const int idx = dx + dy * w;
                    const float* g = grads0 + 3 * idx;
                    const float sqgd = g[1] * g[1] + g[2] * g[2];
                    const float TH = THFac * 10.0f * 0.75f;   // minUseGrad_pixsel = 10
                    if (sqgd > TH * TH) {
                        const float agx = fabsf(g[1]);
                        if (agx > bestXX) { bestXX = agx; bestXXID = idx; }
                        const float agy = fabsf(g[2]);
                        if (agy > bestYY) { bestYY = agy; bestYYID = idx; }
                        const float gxpy = fabsf(g[1] - g[2]);
                        if (gxpy > bestXY) { bestXY = gxpy; bestXYID = idx; }
                        const float gxmy = fabsf(g[1] + g[2]);
                        if (gxmy > bestYX) { bestYX = gxmy; bestYXID = idx; }
                    }
                }
            unsigned char* map0 = map_out + x + y * w;
            const int ids[4] = {bestXXID, bestYYID, bestXYID, bestYXID};
            for (int k = 0; k < 4; k++)
                if (ids[k] >= 0) {
                    if (!map0[ids[k]]) numGood++;
                    map0[ids[k]] = 1;
                }
        }
    return numGood;
}

// PixelSelector2.h:239-277
int makePixelStatus(Init* I, const float* grads, unsigned char* map, int w, int h, float desiredDensity, int recsLeft = 5, float THFac = 1) {
    if (I->sparsityFactor < 1) I->sparsityFactor = 1;
    const int numGoodPoints = gridMaxSelection(grads, map, w, h, I->sparsityFactor, THFac);
    const float quotia = numGoodPoints / (float)(desiredDensity);
    int newSparsity = (I->sparsityFactor * sqrtf(quotia)) + 0.7f;
    if (newSparsity < 1) newSparsity = 1;
    const float oldTHFac = THFac;
    if (newSparsity == 1 && I->sparsityFactor == 1) THFac = 0.5;
    if ((abs(newSparsity - I->sparsityFactor) < 1 && THFac == oldTHFac) || (quotia > 0.8 && 1.0f / quotia > 0.8) || recsLeft == 0) {
        I->sparsityFactor = newSparsity;
        return numGoodPoints;
    }
    I->sparsityFactor = newSparsity;
    return makePixelStatus(I, grads, map, w, h, desiredDensity, recsLeft - 1, THFac);
}

// CoarseInitializer::makeK (:811-837)
void makeK(orc_ctx* c, Init* I) {
    I->levels = c->levels;
    I->w[0] = c->wG[0]; I->h[0] = c->hG[0];
    I->fx[0] = c->cfg.fx; I->fy[0] = c->cfg.fy; I->cx[0] = c->cfg.cx; I->cy[0] = c->cfg.cy;
    for (int l = 1; l < I->levels; l++) {
        I->w[l] = I->w[0] >> l; I->h[l] = I->h[0] >> l;
        I->fx[l] = I->fx[l - 1] * 0.5; I->fy[l] = I->fy[l - 1] * 0.5;
        I->cx[l] = (I->cx[0] + 0.5) / ((int)1 << l) - 0.5;
        I->cy[l] = (I->cy[0] + 0.5) / ((int)1 << l) - 0.5;
    }
    for (int l = 0; l < I->levels; l++) {
        const double K[9] = {I->fx[l], 0, I->cx[l], 0, I->fy[l], I->cy[l], 0, 0, 1};
        ldso_sm::mat3_inv(K, I->Ki[l]);
    }
}

// CoarseInitializer::makeNN (:839-906) on include/ldso_knn.h
void makeNN(Init* I) {
    const float NNDistFactor = 0.05f;
    std::vector<float> us[LDSO_MAX_LEVELS], vs[LDSO_MAX_LEVELS];
    std::vector<ldso_knn::Grid> grids(I->levels);
    for (int l = 0; l < I->levels; l++) {
        for (const Pnt& p : I->points[l]) { us[l].push_back(p.u); vs[l].push_back(p.v); }
        grids[l].build((int)us[l].size(), us[l].data(), vs[l].data(), I->w[l], I->h[l], 8.f);
    }
    for (int l = 0; l < I->levels; l++)
        for (Pnt& p : I->points[l]) {
            int ret_index[10];
            float ret_dist[10];
            const int found = grids[l].query(p.u, p.v, 10, ret_index, ret_dist);
            float sumDF = 0;
            for (int k = 0; k < 10; k++) {
                if (k >= found) { p.neighbours[k] = -1; p.neighboursDist[k] = 0; continue; }
                p.neighbours[k] = ret_index[k];
                const float df = expf(-ret_dist[k] * NNDistFactor);
                sumDF += df;
                p.neighboursDist[k] = df;
            }
            for (int k = 0; k < 10; k++) p.neighboursDist[k] *= 10 / sumDF;
            if (l < I->levels - 1 && !I->points[l + 1].empty()) {
                int pi; float pd;
                grids[l + 1].query(p.u * 0.5f - 0.25f, p.v * 0.5f - 0.25f, 1, &pi, &pd);
                p.parent = pi;
                p.parentDist = expf(-pd * NNDistFactor);
            } else {
                p.parent = -1; p.parentDist = -1;
            }
        }
}

}  // namespace

// ---- setFirst (:656-737) ---------------------------------------------------------------------------------------
extern "C" int orc_init_set_first(orc_ctx* c, int slot, int* num_points_out) {
    Init* I = get(c);
    makeK(c, I);
    I->first_slot = slot;
    const int w0 = I->w[0], h0 = I->h[0];
    std::vector<float> statusMap((size_t)w0 * h0, 0.f);
    std::vector<unsigned char> statusMapB((size_t)w0 * h0, 0);
    const float densities[] = {0.03f, 0.05f, 0.15f, 0.5f, 1.f};
    I->Jb.assign((size_t)10 * w0 * h0, 0.f);
    I->Jb_new.assign((size_t)10 * w0 * h0, 0.f);
    for (int lvl = 0; lvl < I->levels; lvl++) {
        const int wl = I->w[lvl], hl = I->h[lvl];
        const float* dI = c->frames[slot].dIp[lvl].data();
        if (lvl == 0) {
            int potential = 3, nsel = 0;   // sel.currentPotential = 3 (:672)
            int rc = orc_select_make_maps(c, slot, densities[lvl] * w0 * h0, 1, 2.f, &potential, statusMap.data(), &nsel);
            if (rc) return rc;
        } else {
            makePixelStatus(I, dI, statusMapB.data(), wl, hl, densities[lvl] * w0 * h0);
        }
        std::vector<Pnt>& pl = I->points[lvl];
        pl.clear();
        for (int y = patternPadding + 1; y < hl - patternPadding - 2; y++)
            for (int x = patternPadding + 1; x < wl - patternPadding - 2; x++) {
                if ((lvl != 0 && statusMapB[x + y * wl]) || (lvl == 0 && statusMap[x + y * wl] != 0)) {
                    Pnt p;
                    memset(&p, 0, sizeof(p));
                    p.u = x + 0.1; p.v = y + 0.1;
                    p.idepth = 1; p.iR = 1; p.isGood = true;
                    p.energy[0] = p.energy[1] = 0;
                    p.lastHessian = 0; p.lastHessian_new = 0;
                    p.my_type = (lvl != 0) ? 1 : statusMap[x + y * wl];
                    p.outlierTH = patternNum * c->cfg.outlier_th;
                    pl.push_back(p);
                }
            }
        num_points_out[lvl] = (int)pl.size();
    }
    makeNN(I);
    I->thisToNext = SE3();
    I->aff_a = I->aff_b = 0;
    I->snapped = false;
    I->frameID = I->snappedAt = 0;
    return 0;
}

namespace {

inline void interp33(const float* mat, float x, float y, int width, float* out) {   // GlobalFuncs.h:89-103
    int ix = (int)x, iy = (int)y;
    float dx = x - ix, dy = y - iy, dxdy = dx * dy;
    const float* bp = mat + 3 * (ix + iy * width);
    for (int k = 0; k < 3; k++)
        out[k] = dxdy * bp[3 * (1 + width) + k] + (dy - dxdy) * bp[3 * width + k] + (dx - dxdy) * bp[3 + k] +
                 (1 - dx - dy + dxdy) * bp[k];
}
inline float interp31(const float* mat, float x, float y, int width) {   // GlobalFuncs.h:145-159
    int ix = (int)x, iy = (int)y;
    float dx = x - ix, dy = y - iy, dxdy = dx * dy;
    const float* bp = mat + 3 * (ix + iy * width);
    return dxdy * bp[3 * (1 + width)] + (dy - dxdy) * bp[3 * width] + (dx - dxdy) * bp[3] + (1 - dx - dy + dxdy) * bp[0];
}

// calculateAllResiduals (:391-518)
void calculateAllResiduals(orc_ctx* c, Init* I, int lvl, const float* RKi, const float* t, float aff0, float aff1) {
    std::vector<Pnt>& pts = I->points[lvl];
    const int wl = I->w[lvl], hl = I->h[lvl];
    const float fxl = I->fx[lvl], fyl = I->fy[lvl], cxl = I->cx[lvl], cyl = I->cy[lvl];
    const float* colorRef = c->frames[I->first_slot].dIp[lvl].data();
    const float* colorNew = c->frames[I->new_slot].dIp[lvl].data();
    const float huberTH = c->cfg.huber_th;
    for (size_t i = 0; i < pts.size(); i++) {
        Pnt* point = &pts[i];
        point->maxstep = 1e10;
        if (!point->isGood) {
            I->accE.updateSingle(point->energy[0]);
            point->energy_new[0] = point->energy[0]; point->energy_new[1] = point->energy[1];
            point->isGood_new = false;
            continue;
        }
        alignas(16) float dp[9][8];   // dp0..dp7, r
        float dd[8];
        float* Jb = &I->Jb_new[10 * i];
        for (int k = 0; k < 10; k++) Jb[k] = 0;
        bool isGood = true;
        float energy = 0;
        for (int idx = 0; idx < patternNum; idx++) {
            const int dx = patternP[idx][0], dy = patternP[idx][1];
            const float px = point->u + dx, py = point->v + dy;
            float pt[3];
            for (int r = 0; r < 3; r++) pt[r] = (RKi[3 * r] * px + RKi[3 * r + 1] * py + RKi[3 * r + 2] * 1.f) + t[r] * point->idepth_new;
            const float u = pt[0] / pt[2], v = pt[1] / pt[2];
            const float Ku = fxl * u + cxl, Kv = fyl * v + cyl;
            const float new_idepth = point->idepth_new / pt[2];
            if (!(Ku > 1 && Kv > 1 && Ku < wl - 2 && Kv < hl - 2 && new_idepth > 0)) { isGood = false; break; }
            float hitColor[3];
            interp33(colorNew, Ku, Kv, wl, hitColor);
            const float rlR = interp31(colorRef, px, py, wl);
            if (!std::isfinite(rlR) || !std::isfinite(hitColor[0])) { isGood = false; break; }
            const float residual = hitColor[0] - aff0 * rlR - aff1;
            float hw = fabsf(residual) < huberTH ? 1 : huberTH / fabsf(residual);
            energy += hw * residual * residual * (2 - hw);
            const float dxdd = (t[0] - t[2] * u) / pt[2];
            const float dydd = (t[1] - t[2] * v) / pt[2];
            if (hw < 1) hw = sqrtf(hw);
            const float dxInterp = hw * hitColor[1] * fxl;
            const float dyInterp = hw * hitColor[2] * fyl;
            dp[0][idx] = new_idepth * dxInterp;
            dp[1][idx] = new_idepth * dyInterp;
            dp[2][idx] = -new_idepth * (u * dxInterp + v * dyInterp);
            dp[3][idx] = -u * v * dxInterp - (1 + v * v) * dyInterp;
            dp[4][idx] = (1 + u * u) * dxInterp + u * v * dyInterp;
            dp[5][idx] = -v * dxInterp + u * dyInterp;
            dp[6][idx] = -hw * aff0 * rlR;
            dp[7][idx] = -hw * 1;
            dd[idx] = dxInterp * dxdd + dyInterp * dydd;
            dp[8][idx] = hw * residual;
            const float a = dxdd * fxl, b = dydd * fyl;
            const float maxstep = 1.0f / sqrtf(a * a + b * b);
            if (maxstep < point->maxstep) point->maxstep = maxstep;
            for (int k = 0; k < 8; k++) Jb[k] += dp[k][idx] * dd[idx];
            Jb[8] += dp[8][idx] * dd[idx];
            Jb[9] += dd[idx] * dd[idx];
        }
        if (!isGood || energy > point->outlierTH * 20) {
            I->accE.updateSingle(point->energy[0]);
            point->isGood_new = false;
            point->energy_new[0] = point->energy[0]; point->energy_new[1] = point->energy[1];
            continue;
        }
        I->accE.updateSingle(energy);
        point->isGood_new = true;
        point->energy_new[0] = energy;
        for (int k = 0; k + 3 < patternNum; k += 4) {
            __m128 J[9];
            for (int r = 0; r < 9; r++) J[r] = _mm_load_ps(&dp[r][k]);
            I->acc9.updateSSE(J);
        }
    }
}

// calculateAlphaEnergy (:376-389)
void calculateAlphaEnergy(Init* I, int lvl) {
    for (Pnt& p : I->points[lvl]) {
        if (!p.isGood_new) {
            I->accEAlpha.updateSingle(p.energy[1]);
        } else {
            p.energy_new[1] = (p.idepth_new - 1) * (p.idepth_new - 1);
            I->accEAlpha.updateSingle(p.energy_new[1]);
        }
    }
}

// calculateSC (:349-374)
void calculateSC(Init* I, int lvl, float alphaOpt) {
    std::vector<Pnt>& pts = I->points[lvl];
    for (size_t i = 0; i < pts.size(); i++) {
        Pnt* point = &pts[i];
        if (!point->isGood_new) continue;
        float* Jb = &I->Jb_new[10 * i];
        point->lastHessian_new = Jb[9];
        Jb[8] += alphaOpt * (point->idepth_new - 1);
        Jb[9] += alphaOpt;
        if (alphaOpt == 0) {
            Jb[8] += I->couplingWeight * (point->idepth_new - point->iR);
            Jb[9] += I->couplingWeight;
        }
        Jb[9] = 1 / (1 + Jb[9]);
        I->acc9SC.updateSingleWeighted(Jb, Jb[9]);
    }
}

// calcResAndGS (:186-342); H, Hsc 8x8 row-major, returns {energy, alphaEnergy, num}
void calcResAndGS(orc_ctx* c, Init* I, int lvl, float* H_out, float* b_out, float* Hsc_out, float* bsc_out, const SE3& refToNew,
                  double aff_a, double aff_b, float res[3]) {
    double R[9];
    refToNew.rotationMatrix(R);
    double RKid[9];
    ldso_sm::mat3_mul(R, I->Ki[lvl], RKid);
    float RKi[9], t[3];
    for (int k = 0; k < 9; k++) RKi[k] = (float)RKid[k];
    for (int k = 0; k < 3; k++) t[k] = (float)refToNew.t[k];
    const float aff0 = (float)exp(aff_a), aff1 = (float)aff_b;
    I->acc9.initialize(); I->acc9SC.initialize(); I->accE.initialize(); I->accEAlpha.initialize();
    const int npts = (int)I->points[lvl].size();
    calculateAllResiduals(c, I, lvl, RKi, t, aff0, aff1);
    I->accE.finish();
    I->acc9.finish();
    calculateAlphaEnergy(I, lvl);
    I->accEAlpha.finish();
    float alphaEnergy = (refToNew.t[0] * refToNew.t[0] + refToNew.t[1] * refToNew.t[1] + refToNew.t[2] * refToNew.t[2]) * npts;
    alphaEnergy += I->accEAlpha.A;
    alphaEnergy *= I->alphaW;
    float alphaOpt;
    if (alphaEnergy > I->alphaK * npts) {
        alphaOpt = 0;
        alphaEnergy = I->alphaK * npts;
    } else {
        alphaOpt = I->alphaW;
    }
    calculateSC(I, lvl, alphaOpt);
    I->acc9SC.finish();
    for (int r = 0; r < 8; r++) {
        for (int q = 0; q < 8; q++) { H_out[8 * r + q] = I->acc9.H[r][q]; Hsc_out[8 * r + q] = I->acc9SC.H[r][q]; }
        b_out[r] = I->acc9.H[r][8];
        bsc_out[r] = I->acc9SC.H[r][8];
    }
    H_out[0] += alphaOpt * npts; H_out[9] += alphaOpt * npts; H_out[18] += alphaOpt * npts;
    double lg[6];
    refToNew.log(lg);
    for (int k = 0; k < 3; k++) b_out[k] += (float)lg[k] * alphaOpt * npts;
    res[0] = I->accE.A; res[1] = alphaEnergy; res[2] = (float)I->accE.num;
}

}  // namespace

namespace {

// calcEC (:520-536): {old, new} regularisation energy; AccumulatorX<2>::updateNoWeight is a plain float sum with the
// 1k / 1M shift-up tiers (MatrixAccumulators.h:150-198)
void calcEC(Init* I, int lvl, float out[3]) {
    const int npts = (int)I->points[lvl].size();
    if (!I->snapped) { out[0] = 0; out[1] = 0; out[2] = (float)npts; return; }
    float A[2] = {0, 0}, A1k[2] = {0, 0}, A1m[2] = {0, 0};
    float numIn1 = 0, numIn1k = 0, numIn1m = 0;
    auto shiftUp = [&](bool force) {
        if (numIn1 > 1000 || force) { for (int k = 0; k < 2; k++) { A1k[k] += A[k]; A[k] = 0; } numIn1k += numIn1; numIn1 = 0; }
        if (numIn1k > 1000 || force) { for (int k = 0; k < 2; k++) { A1m[k] += A1k[k]; A1k[k] = 0; } numIn1m += numIn1k; numIn1k = 0; }
    };
    for (const Pnt& p : I->points[lvl]) {
        if (!p.isGood_new) continue;
        const float rOld = (p.idepth - p.iR), rNew = (p.idepth_new - p.iR);
        A[0] += rOld * rOld; A[1] += rNew * rNew;
        numIn1++;
        shiftUp(false);
    }
    shiftUp(true);
    out[0] = I->couplingWeight * A1m[0]; out[1] = I->couplingWeight * A1m[1]; out[2] = numIn1m;
}

// optReg (:538-567): in-place and in index order, so later points see the regularised iR of earlier ones
void optReg(Init* I, int lvl) {
    std::vector<Pnt>& ptsl = I->points[lvl];
    if (!I->snapped) {
        for (Pnt& p : ptsl) p.iR = 1;
        return;
    }
    for (Pnt& point : ptsl) {
        if (!point.isGood) continue;
        float idnn[10];
        int nnn = 0;
        for (int j = 0; j < 10; j++) {
            if (point.neighbours[j] == -1) continue;
            const Pnt& other = ptsl[point.neighbours[j]];
            if (!other.isGood) continue;
            idnn[nnn++] = other.iR;
        }
        if (nnn > 2) {
            std::nth_element(idnn, idnn + nnn / 2, idnn + nnn);
            point.iR = (1 - I->regWeight) * point.idepth + I->regWeight * idnn[nnn / 2];
        }
    }
}

// propagateUp (:570-604)
void propagateUp(Init* I, int srcLvl) {
    std::vector<Pnt>& ptss = I->points[srcLvl];
    std::vector<Pnt>& ptst = I->points[srcLvl + 1];
    for (Pnt& parent : ptst) { parent.iR = 0; parent.iRSumNum = 0; }
    for (Pnt& point : ptss) {
        if (!point.isGood || point.parent < 0) continue;
        Pnt& parent = ptst[point.parent];
        parent.iR += point.iR * point.lastHessian;
        parent.iRSumNum += point.lastHessian;
    }
    for (Pnt& parent : ptst)
        if (parent.iRSumNum > 0) {
            parent.idepth = parent.iR = (parent.iR / parent.iRSumNum);
            parent.isGood = true;
        }
    optReg(I, srcLvl + 1);
}

// propagateDown (:606-631)
void propagateDown(Init* I, int srcLvl) {
    std::vector<Pnt>& ptss = I->points[srcLvl];
    std::vector<Pnt>& ptst = I->points[srcLvl - 1];
    for (Pnt& point : ptst) {
        if (point.parent < 0) continue;
        const Pnt& parent = ptss[point.parent];
        if (!parent.isGood || parent.lastHessian < 0.1) continue;
        if (!point.isGood) {
            point.iR = point.idepth = point.idepth_new = parent.iR;
            point.isGood = true;
            point.lastHessian = 0;
        } else {
            const float newiR = (point.iR * point.lastHessian * 2 + parent.iR * parent.lastHessian) /
                                (point.lastHessian * 2 + parent.lastHessian);
            point.iR = point.idepth = point.idepth_new = newiR;
        }
    }
    optReg(I, srcLvl - 1);
}

// resetPoints (:739-761)
void resetPoints(Init* I, int lvl) {
    std::vector<Pnt>& pts = I->points[lvl];
    for (Pnt& p : pts) {
        p.energy[0] = p.energy[1] = 0;
        p.idepth_new = p.idepth;
        if (lvl == I->levels - 1 && !p.isGood) {
            float snd = 0, sn = 0;
            for (int n = 0; n < 10; n++) {
                if (p.neighbours[n] == -1 || !pts[p.neighbours[n]].isGood) continue;
                snd += pts[p.neighbours[n]].iR;
                sn += 1;
            }
            if (sn > 0) {
                p.isGood = true;
                p.iR = p.idepth = p.idepth_new = snd / sn;
            }
        }
    }
}

// doStep (:763-789)
void doStep(Init* I, int lvl, float lambda, const float inc[8]) {
    const float maxPixelStep = 0.25f, idMaxStep = 1e10f;
    std::vector<Pnt>& pts = I->points[lvl];
    for (size_t i = 0; i < pts.size(); i++) {
        if (!pts[i].isGood) continue;
        const float* Jb = &I->Jb[10 * i];
        float dot = 0;
        for (int k = 0; k < 8; k++) dot += Jb[k] * inc[k];
        const float b = Jb[8] + dot;
        float step = -b * Jb[9] / (1 + lambda);
        float maxstep = maxPixelStep * pts[i].maxstep;
        if (maxstep > idMaxStep) maxstep = idMaxStep;
        if (step > maxstep) step = maxstep;
        if (step < -maxstep) step = -maxstep;
        float newIdepth = pts[i].idepth + step;
        if (newIdepth < 1e-3f) newIdepth = 1e-3f;
        if (newIdepth > 50) newIdepth = 50;
        pts[i].idepth_new = newIdepth;
    }
}

// applyStep (:791-809)
void applyStep(Init* I, int lvl) {
    for (Pnt& p : I->points[lvl]) {
        if (!p.isGood) {
            p.idepth = p.idepth_new = p.iR;
            continue;
        }
        p.energy[0] = p.energy_new[0]; p.energy[1] = p.energy_new[1];
        p.isGood = p.isGood_new;
        p.idepth = p.idepth_new;
        p.lastHessian = p.lastHessian_new;
    }
    std::swap(I->Jb, I->Jb_new);
}

}  // namespace

// ---- trackFrame (:45-183) --------------------------------------------------------------------------------------
extern "C" int orc_init_track_frame(orc_ctx* c, int slot, float exposure_first, float exposure_new, double pose_out[12],
                                    float aff_out[2], int* snapped_out, int* ready, float latest_res[3], int* iters_out) {
    Init* I = get(c);
    if (I->first_slot < 0) return -2;
    I->new_slot = slot;
    const int maxIterations[] = {5, 5, 10, 30, 50};
    I->alphaK = 2.5f * 2.5f;
    I->alphaW = 150 * 150;
    I->regWeight = 0.8f;
    I->couplingWeight = 1;
    const int L = I->levels;
    if (!I->snapped) {
        I->thisToNext.t[0] = I->thisToNext.t[1] = I->thisToNext.t[2] = 0;
        for (int lvl = 0; lvl < L; lvl++)
            for (Pnt& p : I->points[lvl]) { p.iR = 1; p.idepth_new = 1; p.lastHessian = 0; }
    }
    SE3 refToNew_current = I->thisToNext;
    double aff_cur_a = I->aff_a, aff_cur_b = I->aff_b;
    if (exposure_first > 0 && exposure_new > 0) { aff_cur_a = logf(exposure_new / exposure_first); aff_cur_b = 0; }
    // wM (:29-32): SCALE_XI_ROT on the first three entries, SCALE_XI_TRANS on the next three, as the reference has it
    const float wM[8] = {orc::SCALE_XI_ROT, orc::SCALE_XI_ROT, orc::SCALE_XI_ROT, orc::SCALE_XI_TRANS, orc::SCALE_XI_TRANS,
                         orc::SCALE_XI_TRANS, orc::SCALE_A, orc::SCALE_B};
    float latestRes[3] = {0, 0, 0};
    for (int lvl = L - 1; lvl >= 0; lvl--) {
        if (lvl < L - 1) propagateDown(I, lvl + 1);
        float H[64], b[8], Hsc[64], bsc[8], resOld[3];
        resetPoints(I, lvl);
        calcResAndGS(c, I, lvl, H, b, Hsc, bsc, refToNew_current, aff_cur_a, aff_cur_b, resOld);
        applyStep(I, lvl);
        float lambda = 0.1f;
        const float eps = 1e-4f;
        int fails = 0, iteration = 0;
        while (true) {
            float Hl[64], bl[8];
            for (int k = 0; k < 64; k++) Hl[k] = H[k];
            for (int k = 0; k < 8; k++) Hl[9 * k] *= (1 + lambda);
            for (int k = 0; k < 64; k++) Hl[k] -= Hsc[k] * (1 / (1 + lambda));
            for (int k = 0; k < 8; k++) bl[k] = b[k] - bsc[k] * (1 / (1 + lambda));
            const float sc = 0.01f / (I->w[lvl] * I->h[lvl]);
            for (int r = 0; r < 8; r++) {
                for (int q = 0; q < 8; q++) Hl[8 * r + q] = wM[r] * Hl[8 * r + q] * wM[q] * sc;
                bl[r] = wM[r] * bl[r] * sc;
            }
            float inc[8];
            const int nd = I->fixAffine ? 6 : 8;
            {
                double Hd[64], bd[8], x[8];
                for (int r = 0; r < nd; r++) { for (int q = 0; q < nd; q++) Hd[nd * r + q] = Hl[8 * r + q]; bd[r] = bl[r]; }
                ldso_sm::ldlt_solve(nd, Hd, bd, x);
                for (int k = 0; k < 8; k++) inc[k] = k < nd ? -(wM[k] * (float)x[k]) : 0.f;
            }
            double incd[6];
            for (int k = 0; k < 6; k++) incd[k] = inc[k];
            SE3 refToNew_new = SE3::exp(incd) * refToNew_current;
            const double aff_new_a = aff_cur_a + inc[6], aff_new_b = aff_cur_b + inc[7];
            doStep(I, lvl, lambda, inc);
            float H_new[64], b_new[8], Hsc_new[64], bsc_new[8], resNew[3], regEnergy[3];
            calcResAndGS(c, I, lvl, H_new, b_new, Hsc_new, bsc_new, refToNew_new, aff_new_a, aff_new_b, resNew);
            calcEC(I, lvl, regEnergy);
            const float eTotalNew = (resNew[0] + resNew[1] + regEnergy[1]);
            const float eTotalOld = (resOld[0] + resOld[1] + regEnergy[0]);
            const bool accept = eTotalOld > eTotalNew;
            if (accept) {
                if (resNew[1] == I->alphaK * (float)I->points[lvl].size()) I->snapped = true;
                memcpy(H, H_new, sizeof(H)); memcpy(b, b_new, sizeof(b)); memcpy(Hsc, Hsc_new, sizeof(H)); memcpy(bsc, bsc_new, sizeof(b));
                memcpy(resOld, resNew, sizeof(resOld));
                aff_cur_a = aff_new_a; aff_cur_b = aff_new_b;
                refToNew_current = refToNew_new;
                applyStep(I, lvl);
                optReg(I, lvl);
                lambda *= 0.5f;
                fails = 0;
                if (lambda < 0.0001f) lambda = 0.0001f;
            } else {
                fails++;
                lambda *= 4;
                if (lambda > 10000) lambda = 10000;
            }
            float n2 = 0;
            for (int k = 0; k < 8; k++) n2 += inc[k] * inc[k];
            const bool quitOpt = !(sqrtf(n2) > eps) || iteration >= maxIterations[lvl] || fails >= 2;
            if (quitOpt) break;
            iteration++;
        }
        if (iters_out) iters_out[lvl] = iteration;
        memcpy(latestRes, resOld, sizeof(latestRes));
    }
    I->thisToNext = refToNew_current;
    I->aff_a = aff_cur_a; I->aff_b = aff_cur_b;
    for (int i = 0; i < L - 1; i++) propagateUp(I, i);
    I->frameID++;
    if (!I->snapped) I->snappedAt = 0;
    if (I->snapped && I->snappedAt == 0) I->snappedAt = I->frameID;
    I->thisToNext.toRt(pose_out);
    aff_out[0] = (float)I->aff_a; aff_out[1] = (float)I->aff_b;
    *snapped_out = I->snapped ? 1 : 0;
    *ready = (I->snapped && I->frameID > I->snappedAt + 5) ? 1 : 0;
    if (latest_res) memcpy(latest_res, latestRes, sizeof(latestRes));
    return 0;
}

extern "C" int orc_init_get_points(orc_ctx* c, int lvl, int* n, float* u, float* v, float* idepth, float* iR, unsigned char* is_good,
                                   float* my_type, float* last_hessian, float* energy) {
    Init* I = get(c);
    if (lvl < 0 || lvl >= I->levels) return -2;
    const std::vector<Pnt>& p = I->points[lvl];
    *n = (int)p.size();
    for (size_t i = 0; i < p.size(); i++) {
        if (u) u[i] = p[i].u;
        if (v) v[i] = p[i].v;
        if (idepth) idepth[i] = p[i].idepth;
        if (iR) iR[i] = p[i].iR;
        if (is_good) is_good[i] = p[i].isGood ? 1 : 0;
        if (my_type) my_type[i] = p[i].my_type;
        if (last_hessian) last_hessian[i] = p[i].lastHessian;
        if (energy) energy[i] = p[i].energy[0];
    }
    return 0;
}

extern "C" int orc_init_get_graph(orc_ctx* c, int lvl, int* neighbours, int* parent) {
    Init* I = get(c);
    if (lvl < 0 || lvl >= I->levels) return -2;
    const std::vector<Pnt>& p = I->points[lvl];
    for (size_t i = 0; i < p.size(); i++) {
        for (int k = 0; k < 10; k++) neighbours[10 * i + k] = p[i].neighbours[k];
        parent[i] = p[i].parent;
    }
    return 0;
}

extern "C" int orc_init_calc_res_and_gs(orc_ctx* c, int slot, int lvl, const double ref_to_new[12], float aff_a, float aff_b,
                                        float H[64], float b[8], float Hsc[64], float bsc[8], float res[3]) {
    Init* I = get(c);
    if (I->first_slot < 0 || lvl < 0 || lvl >= I->levels) return -2;
    I->new_slot = slot;
    I->alphaK = 2.5f * 2.5f; I->alphaW = 150 * 150; I->regWeight = 0.8f; I->couplingWeight = 1;
    resetPoints(I, lvl);                     // trackFrame's resetPoints(lvl) (:84): idepth_new = idepth (setFirst leaves it unset)
    calcResAndGS(c, I, lvl, H, b, Hsc, bsc, SE3::fromRt(ref_to_new), aff_a, aff_b, res);
    return 0;
}
