// oracle/orc_immature.cpp — TEST INFRASTRUCTURE (see orc_common.h).  PARITY UNPINNED.
// CPU restatement of ImmaturePoint (src/internal/ImmaturePoint.cc): constructor (:14-39), traceOn (:47-317),
// linearizeResidual (:319-389); FullSystem::traceNewCoarse (src/frontend/FullSystem.cc:1157-1195) and
// FullSystem::optimizeImmaturePoint (:1035-1155).
#include "orc_common.h"


using namespace orc;

namespace {

const int patternP[8][2] = {{0, -2}, {-1, -1}, {1, -1}, {-2, 0}, {0, 0}, {2, 0}, {-1, 1}, {0, 2}};

struct Pool {
    int n = 0;
    std::vector<int> host, status;
    std::vector<float> u, v, color, weights, gradH, energyTH, idepth_min, idepth_max, quality, lastUV, lastInterval, type;
    // CoarseDistanceMap (include/frontend/CoarseTracker.h:120-160): fwdWarpedIDDistFinal and the two BFS lists
    std::vector<float> dist;
    std::vector<int> bfs1, bfs2;
    int trace_counts[6] = {0, 0, 0, 0, 0, 0};   // last trace (orc_immature_trace_async / _counts)
};
Pool& pool_of(orc_ctx* c) {
    if (!c->immature) c->immature = std::make_shared<Pool>();
    return *(Pool*)c->immature.get();
}

// include/internal/GlobalFuncs.h:185-207
void interp33BiLin(const float* mat, float x, float y, int width, float* out) {
    int ix = (int)x, iy = (int)y;
    const float* bp = mat + 3 * (ix + iy * width);
    float tl = bp[0], tr = bp[3], bl = bp[3 * width], br = bp[3 * width + 3];
    float dx = x - ix, dy = y - iy;
    float topInt = dx * tr + (1 - dx) * tl;
    float botInt = dx * br + (1 - dx) * bl;
    float leftInt = dy * bl + (1 - dy) * tl;
    float rightInt = dy * br + (1 - dy) * tr;
    out[0] = dx * rightInt + (1 - dx) * leftInt;
    out[1] = rightInt - leftInt;
    out[2] = botInt - topInt;
}
void interp33(const float* mat, float x, float y, int width, float* out) {   // GlobalFuncs.h:89-103
    int ix = (int)x, iy = (int)y;
    float dx = x - ix, dy = y - iy, dxdy = dx * dy;
    const float* bp = mat + 3 * (ix + iy * width);
    for (int k = 0; k < 3; k++)
        out[k] = dxdy * bp[3 * (1 + width) + k] + (dy - dxdy) * bp[3 * width + k] + (dx - dxdy) * bp[3 + k] +
                 (1 - dx - dy + dxdy) * bp[k];
}
float interp31(const float* mat, float x, float y, int width) {   // GlobalFuncs.h:145-159
    int ix = (int)x, iy = (int)y;
    float dx = x - ix, dy = y - iy, dxdy = dx * dy;
    const float* bp = mat + 3 * (ix + iy * width);
    return dxdy * bp[3 * (1 + width)] + (dy - dxdy) * bp[3 * width] + (dx - dxdy) * bp[3] + (1 - dx - dy + dxdy) * bp[0];
}

}  // namespace

// ImmaturePoint::ImmaturePoint (ImmaturePoint.cc:14-39)
extern "C" int orc_immature_create(orc_ctx* c, int slot, int n, const float* pu, const float* pv, float* color, float* weights,
                                   float* gradH, float* energyTH) {
    const float* dI = c->frames[slot].dIp[0].data();
    const int w = c->wG[0];
    const float thsc = c->cfg.outlier_th_sum_component;
    for (int k = 0; k < n; k++) {
        float g00 = 0, g01 = 0, g11 = 0;
        bool bad = false;
        for (int idx = 0; idx < 8; idx++) color[8 * k + idx] = weights[8 * k + idx] = 0;   // defined where the ctor returns early
        for (int idx = 0; idx < 8; idx++) {
            float ptc[3];
            interp33BiLin(dI, pu[k] + patternP[idx][0], pv[k] + patternP[idx][1], w, ptc);
            color[8 * k + idx] = ptc[0];
            if (!std::isfinite(ptc[0])) { bad = true; break; }
            g00 += ptc[1] * ptc[1]; g01 += ptc[1] * ptc[2]; g11 += ptc[2] * ptc[2];
            weights[8 * k + idx] = sqrtf(thsc / (thsc + (ptc[1] * ptc[1] + ptc[2] * ptc[2])));
        }
        gradH[3 * k] = g00; gradH[3 * k + 1] = g01; gradH[3 * k + 2] = g11;
        if (bad) { energyTH[k] = NAN; continue; }
        float e = 8 * c->cfg.outlier_th;
        e *= c->cfg.overall_energy_th_weight * c->cfg.overall_energy_th_weight;
        energyTH[k] = e;
    }
    return 0;
}

extern "C" int orc_immature_set(orc_ctx* c, int n, const int* host, const float* u, const float* v, const float* color,
                                const float* weights, const float* gradH, const float* energyTH, const float* idmin,
                                const float* idmax, const float* quality, const int* status) {
    Pool& P = pool_of(c);
    P.n = n;
    P.host.assign(host, host + n); P.status.assign(status, status + n);
    P.u.assign(u, u + n); P.v.assign(v, v + n); P.color.assign(color, color + 8 * n); P.weights.assign(weights, weights + 8 * n);
    P.gradH.assign(gradH, gradH + 3 * n); P.energyTH.assign(energyTH, energyTH + n);
    P.idepth_min.assign(idmin, idmin + n); P.idepth_max.assign(idmax, idmax + n); P.quality.assign(quality, quality + n);
    P.lastUV.assign(2 * n, 0.f); P.lastInterval.assign(n, 0.f); P.type.assign(n, 1.f);
    return 0;
}

// ImmaturePoint::traceOn (ImmaturePoint.cc:47-317); enum values as ImmaturePointStatus
static int trace_on(orc_ctx* c, Pool& P, int k, const float* dI, const float* KRKi, const float* Kt, const float* aff) {
    enum { GOOD = 0, OOB, OUTLIER, SKIPPED, BADCONDITION, UNINIT };
    const int wG = c->wG[0], hG = c->hG[0];
    int& last = P.status[k];
    float& idepth_min = P.idepth_min[k];
    float& idepth_max = P.idepth_max[k];
    auto fail = [&](int st) { P.lastUV[2 * k] = -1; P.lastUV[2 * k + 1] = -1; P.lastInterval[k] = 0; return last = st; };
    if (last == OOB) return last;
    const ldso_gpu_config& s = c->cfg;
    float maxPixSearch = (wG + hG) * s.max_pix_search;
    const float uu = P.u[k], vv = P.v[k];
    float pr[3] = {KRKi[0] * uu + KRKi[1] * vv + KRKi[2] * 1, KRKi[3] * uu + KRKi[4] * vv + KRKi[5] * 1,
                   KRKi[6] * uu + KRKi[7] * vv + KRKi[8] * 1};
    float ptpMin[3] = {pr[0] + Kt[0] * idepth_min, pr[1] + Kt[1] * idepth_min, pr[2] + Kt[2] * idepth_min};
    float uMin = ptpMin[0] / ptpMin[2], vMin = ptpMin[1] / ptpMin[2];
    if (!(uMin > 4 && vMin > 4 && uMin < wG - 5 && vMin < hG - 5)) return fail(OOB);
    float dist, uMax, vMax, ptpMax[3];
    if (std::isfinite(idepth_max)) {
        for (int i = 0; i < 3; i++) ptpMax[i] = pr[i] + Kt[i] * idepth_max;
        uMax = ptpMax[0] / ptpMax[2]; vMax = ptpMax[1] / ptpMax[2];
        if (!(uMax > 4 && vMax > 4 && uMax < wG - 5 && vMax < hG - 5)) return fail(OOB);
        dist = (uMin - uMax) * (uMin - uMax) + (vMin - vMax) * (vMin - vMax);
        dist = sqrtf(dist);
        if (dist < s.trace_slack_interval) {
            P.lastUV[2 * k] = (uMax + uMin) * 0.5; P.lastUV[2 * k + 1] = (vMax + vMin) * 0.5;
            P.lastInterval[k] = dist;
            return last = SKIPPED;
        }
    } else {
        dist = maxPixSearch;
        for (int i = 0; i < 3; i++) ptpMax[i] = pr[i] + Kt[i] * 0.01f;
        uMax = ptpMax[0] / ptpMax[2]; vMax = ptpMax[1] / ptpMax[2];
        float dx = uMax - uMin, dy = vMax - vMin;
        float d = 1.0f / sqrtf(dx * dx + dy * dy);
        uMax = uMin + dist * dx * d;
        vMax = vMin + dist * dy * d;
        if (!(uMax > 4 && vMax > 4 && uMax < wG - 5 && vMax < hG - 5)) return fail(OOB);
    }
    if (!(idepth_min < 0 || (ptpMin[2] > 0.75 && ptpMin[2] < 1.5))) return fail(OOB);
    float dx = s.trace_stepsize * (uMax - uMin), dy = s.trace_stepsize * (vMax - vMin);
    const float g00 = P.gradH[3 * k], g01 = P.gradH[3 * k + 1], g11 = P.gradH[3 * k + 2];
    // v^T G v with Eigen's evaluation order: (v^T G) first, then dot with v
    float a = (dx * g00 + dy * g01) * dx + (dx * g01 + dy * g11) * dy;
    float b = (dy * g00 + (-dx) * g01) * dy + (dy * g01 + (-dx) * g11) * (-dx);
    float errorInPixel = 0.2f + 0.2f * (a + b) / a;
    if (errorInPixel * s.trace_min_improvement > dist && std::isfinite(idepth_max)) {
        P.lastUV[2 * k] = (uMax + uMin) * 0.5; P.lastUV[2 * k + 1] = (vMax + vMin) * 0.5;
        P.lastInterval[k] = dist;
        return last = BADCONDITION;
    }
    if (errorInPixel > 10) errorInPixel = 10;
    dx /= dist; dy /= dist;
    if (dist > maxPixSearch) {
        uMax = uMin + maxPixSearch * dx;
        vMax = vMin + maxPixSearch * dy;
        dist = maxPixSearch;
    }
    int numSteps = 1.9999f + dist / s.trace_stepsize;
    float randShift = uMin * 1000 - floorf(uMin * 1000);
    float ptx = uMin - randShift * dx, pty = vMin - randShift * dy;
    float rot[8][2];
    for (int idx = 0; idx < 8; idx++) {
        rot[idx][0] = KRKi[0] * patternP[idx][0] + KRKi[1] * patternP[idx][1];
        rot[idx][1] = KRKi[3] * patternP[idx][0] + KRKi[4] * patternP[idx][1];
    }
    if (!std::isfinite(dx) || !std::isfinite(dy)) return fail(OOB);
    const float* color = &P.color[8 * k];
    const float* weights = &P.weights[8 * k];
    float errors[100];
    float bestU = 0, bestV = 0, bestEnergy = 1e10;
    int bestIdx = -1;
    if (numSteps >= 100) numSteps = 99;
    for (int i = 0; i < numSteps; i++) {
        float energy = 0;
        for (int idx = 0; idx < 8; idx++) {
            float hitColor = interp31(dI, (float)(ptx + rot[idx][0]), (float)(pty + rot[idx][1]), wG);
            if (!std::isfinite(hitColor)) { energy += 1e5; continue; }
            float residual = hitColor - (float)(aff[0] * color[idx] + aff[1]);
            float hw = fabs(residual) < s.huber_th ? 1 : s.huber_th / fabs(residual);
            energy += hw * residual * residual * (2 - hw);
        }
        errors[i] = energy;
        if (energy < bestEnergy) { bestU = ptx; bestV = pty; bestEnergy = energy; bestIdx = i; }
        ptx += dx; pty += dy;
    }
    float secondBest = 1e10;
    for (int i = 0; i < numSteps; i++)
        if ((i < bestIdx - s.min_trace_test_radius || i > bestIdx + s.min_trace_test_radius) && errors[i] < secondBest)
            secondBest = errors[i];
    float newQuality = secondBest / bestEnergy;
    if (newQuality < P.quality[k] || numSteps > 10) P.quality[k] = newQuality;
    float uBak = bestU, vBak = bestV, gnstepsize = 1, stepBack = 0;
    if (s.trace_gn_iterations > 0) bestEnergy = 1e5;
    for (int it = 0; it < s.trace_gn_iterations; it++) {
        float H = 1, bb = 0, energy = 0;
        for (int idx = 0; idx < 8; idx++) {
            float hit[3];
            interp33(dI, (float)(bestU + rot[idx][0]), (float)(bestV + rot[idx][1]), wG, hit);
            if (!std::isfinite((float)hit[0])) { energy += 1e5; continue; }
            float residual = hit[0] - (aff[0] * color[idx] + aff[1]);
            float dResdDist = dx * hit[1] + dy * hit[2];
            float hw = fabs(residual) < s.huber_th ? 1 : s.huber_th / fabs(residual);
            H += hw * dResdDist * dResdDist;
            bb += hw * residual * dResdDist;
            energy += weights[idx] * weights[idx] * hw * residual * residual * (2 - hw);
        }
        if (energy > bestEnergy) {
            stepBack *= 0.5;
            bestU = uBak + stepBack * dx;
            bestV = vBak + stepBack * dy;
        } else {
            float step = -gnstepsize * bb / H;
            if (step < -0.5) step = -0.5;
            else if (step > 0.5) step = 0.5;
            if (!std::isfinite(step)) step = 0;
            uBak = bestU; vBak = bestV; stepBack = step;
            bestU += step * dx; bestV += step * dy;
            bestEnergy = energy;
        }
        if (fabsf(stepBack) < s.trace_gn_threshold) break;
    }
    if (!(bestEnergy < P.energyTH[k] * s.trace_extra_slack_on_th)) {
        P.lastInterval[k] = 0; P.lastUV[2 * k] = -1; P.lastUV[2 * k + 1] = -1;
        if (last == OUTLIER) return last = OOB;
        return last = OUTLIER;
    }
    if (dx * dx > dy * dy) {
        idepth_min = (pr[2] * (bestU - errorInPixel * dx) - pr[0]) / (Kt[0] - Kt[2] * (bestU - errorInPixel * dx));
        idepth_max = (pr[2] * (bestU + errorInPixel * dx) - pr[0]) / (Kt[0] - Kt[2] * (bestU + errorInPixel * dx));
    } else {
        idepth_min = (pr[2] * (bestV - errorInPixel * dy) - pr[1]) / (Kt[1] - Kt[2] * (bestV - errorInPixel * dy));
        idepth_max = (pr[2] * (bestV + errorInPixel * dy) - pr[1]) / (Kt[1] - Kt[2] * (bestV + errorInPixel * dy));
    }
    if (idepth_min > idepth_max) std::swap<float>(idepth_min, idepth_max);
    if (!std::isfinite(idepth_min) || !std::isfinite(idepth_max) || (idepth_max < 0)) {
        P.lastInterval[k] = 0; P.lastUV[2 * k] = -1; P.lastUV[2 * k + 1] = -1;
        return last = OUTLIER;
    }
    P.lastInterval[k] = 2 * errorInPixel;
    P.lastUV[2 * k] = bestU; P.lastUV[2 * k + 1] = bestV;
    return last = GOOD;
}

extern "C" int orc_immature_trace(orc_ctx* c, int new_slot, int n_hosts, const float* KRKi, const float* Kt, const float* aff,
                                  int counts[6]) {
    Pool& P = pool_of(c);
    const float* dI = c->frames[new_slot].dIp[0].data();
    for (int i = 0; i < 6; i++) counts[i] = 0;
    for (int k = 0; k < P.n; k++) {
        int h = P.host[k];
        if (h < 0 || h >= n_hosts) return -2;
        int st = trace_on(c, P, k, dI, KRKi + 9 * h, Kt + 3 * h, aff + 2 * h);
        counts[st]++;
    }
    return 0;
}

// the asynchronous pair of the product ABI: the oracle traces at once and keeps the counts
extern "C" int orc_immature_trace_async(orc_ctx* c, int new_slot, int n_hosts, const float* KRKi, const float* Kt, const float* aff) {
    Pool& P = pool_of(c);
    return orc_immature_trace(c, new_slot, n_hosts, KRKi, Kt, aff, P.trace_counts);
}

extern "C" int orc_immature_trace_counts(orc_ctx* c, int counts[6]) {
    Pool& P = pool_of(c);
    for (int i = 0; i < 6; i++) counts[i] = P.trace_counts[i];
    return 0;
}

extern "C" int orc_immature_get(orc_ctx* c, float* idmin, float* idmax, float* quality, int* status, float* uv, float* interval) {
    Pool& P = pool_of(c);
    if (idmin) memcpy(idmin, P.idepth_min.data(), sizeof(float) * P.n);
    if (idmax) memcpy(idmax, P.idepth_max.data(), sizeof(float) * P.n);
    if (quality) memcpy(quality, P.quality.data(), sizeof(float) * P.n);
    if (status) memcpy(status, P.status.data(), sizeof(int) * P.n);
    if (uv) memcpy(uv, P.lastUV.data(), sizeof(float) * 2 * P.n);
    if (interval) memcpy(interval, P.lastInterval.data(), sizeof(float) * P.n);
    return 0;
}

// ImmaturePoint::linearizeResidual (ImmaturePoint.cc:319-389)
struct TmpRes { int state_state, state_NewState; double state_energy, state_NewEnergy; int target; };
static double linearize_residual(orc_ctx* c, Pool& P, int k, float outlierTHSlack, TmpRes& tr, float& Hdd, float& bd, float idepth) {
    Window& W = c->win;
    if (tr.state_state == LDSO_RES_OOB) { tr.state_NewState = LDSO_RES_OOB; return tr.state_energy; }
    const float* pc = W.precalc[P.host[k] + W.n * tr.target];
    const float* R = pc + 27; const float* t = pc + 36;
    const float aff0 = pc[24], aff1 = pc[25];
    const float* dIl = c->frames[W.slots[tr.target]].dIp[0].data();
    const float wM3G = c->wG[0] - 3, hM3G = c->hG[0] - 3;
    float energyLeft = 0;
    for (int idx = 0; idx < 8; idx++) {
        int dx = patternP[idx][0], dy = patternP[idx][1];
        float KliP[3] = {(P.u[k] + dx - W.cxl) * W.fxli, (P.v[k] + dy - W.cyl) * W.fyli, 1};
        float ptp[3] = {R[0] * KliP[0] + R[1] * KliP[1] + R[2] * KliP[2] + t[0] * idepth,
                        R[3] * KliP[0] + R[4] * KliP[1] + R[5] * KliP[2] + t[1] * idepth,
                        R[6] * KliP[0] + R[7] * KliP[1] + R[8] * KliP[2] + t[2] * idepth};
        float drescale = 1.0f / ptp[2];
        if (!(drescale > 0)) { tr.state_NewState = LDSO_RES_OOB; return tr.state_energy; }
        float u = ptp[0] * drescale, v = ptp[1] * drescale;
        float Ku = u * W.fxl + W.cxl, Kv = v * W.fyl + W.cyl;
        if (!(Ku > 1.1f && Kv > 1.1f && Ku < wM3G && Kv < hM3G)) { tr.state_NewState = LDSO_RES_OOB; return tr.state_energy; }
        float hit[3];
        interp33(dIl, Ku, Kv, c->wG[0], hit);
        if (!std::isfinite((float)hit[0])) { tr.state_NewState = LDSO_RES_OOB; return tr.state_energy; }
        float residual = hit[0] - (aff0 * P.color[8 * k + idx] + aff1);
        float hw = fabsf(residual) < c->cfg.huber_th ? 1 : c->cfg.huber_th / fabsf(residual);
        const float wgt = P.weights[8 * k + idx];
        energyLeft += wgt * wgt * hw * residual * residual * (2 - hw);
        float dxInterp = hit[1] * W.fxl, dyInterp = hit[2] * W.fyl;
        float d_idepth = (dxInterp * drescale * (t[0] - t[2] * u) + dyInterp * drescale * (t[1] - t[2] * v)) * SCALE_IDEPTH;
        hw *= wgt * wgt;
        Hdd += (hw * d_idepth) * d_idepth;
        bd += (hw * residual) * d_idepth;
    }
    if (energyLeft > P.energyTH[k] * outlierTHSlack) {
        energyLeft = P.energyTH[k] * outlierTHSlack;
        tr.state_NewState = LDSO_RES_OUTLIER;
    } else {
        tr.state_NewState = LDSO_RES_IN;
    }
    tr.state_NewEnergy = energyLeft;
    return energyLeft;
}

// FullSystem::optimizeImmaturePoint (src/frontend/FullSystem.cc:1035-1155)
extern "C" int orc_immature_activate(orc_ctx* c, int n_cand, const int* cand, int min_obs, float* idepth_out, int* ok_out,
                                     int* res_state_out) {
    Pool& P = pool_of(c);
    Window& W = c->win;
    const int T = orc_threads(c);   // activatePointsMT_Reductor (FullSystem.cc:1339-1354)
#pragma omp parallel for num_threads(T) if (T > 1) schedule(static)
    for (int q = 0; q < n_cand; q++) {
        const int k = cand[q];
        ok_out[q] = 0; idepth_out[q] = 0;
        for (int f = 0; f < W.n; f++) res_state_out[q * W.n + f] = -1;
        TmpRes res[LDSO_MAX_WINDOW];
        int nres = 0;
        for (int f = 0; f < W.n; f++)
            if (f != P.host[k]) { res[nres] = {LDSO_RES_IN, LDSO_RES_OUTLIER, 0, 0, f}; nres++; }
        float lastEnergy = 0, lastHdd = 0, lastbd = 0;
        float currentIdepth = (P.idepth_max[k] + P.idepth_min[k]) * 0.5f;
        for (int i = 0; i < nres; i++) {
            lastEnergy += linearize_residual(c, P, k, 1000, res[i], lastHdd, lastbd, currentIdepth);
            res[i].state_state = res[i].state_NewState;
            res[i].state_energy = res[i].state_NewEnergy;
        }
        if (!std::isfinite(lastEnergy) || lastHdd < c->cfg.min_idepth_h_act) continue;
        float lambda = 0.1;
        bool dead = false;
        for (int iteration = 0; iteration < c->cfg.gn_its_on_point_activation; iteration++) {
            float H = lastHdd;
            H *= 1 + lambda;
            float step = (1.0 / H) * lastbd;
            float newIdepth = currentIdepth - step;
            float newHdd = 0, newbd = 0, newEnergy = 0;
            for (int i = 0; i < nres; i++) newEnergy += linearize_residual(c, P, k, 1, res[i], newHdd, newbd, newIdepth);
            if (!std::isfinite(lastEnergy) || newHdd < c->cfg.min_idepth_h_act) { dead = true; break; }
            if (newEnergy < lastEnergy) {
                currentIdepth = newIdepth; lastHdd = newHdd; lastbd = newbd; lastEnergy = newEnergy;
                for (int i = 0; i < nres; i++) { res[i].state_state = res[i].state_NewState; res[i].state_energy = res[i].state_NewEnergy; }
                lambda *= 0.5;
            } else {
                lambda *= 5;
            }
            if (fabsf(step) < 0.0001 * currentIdepth) break;
        }
        if (dead || !std::isfinite(currentIdepth)) continue;
        int numGoodRes = 0;
        for (int i = 0; i < nres; i++) if (res[i].state_state == LDSO_RES_IN) numGoodRes++;
        for (int i = 0; i < nres; i++) res_state_out[q * W.n + res[i].target] = res[i].state_state;
        if (numGoodRes < min_obs) continue;
        ok_out[q] = 1;
        idepth_out[q] = currentIdepth;
    }
    return 0;
}

// device-resident pool maintenance of the GPU library, restated on the host vectors
extern "C" int orc_immature_compact(orc_ctx* c, int drop_host, int drop_oob, int n_remove, const int* remove_idx, int* n_out) {
    Pool& P = pool_of(c);
    std::vector<char> rm(P.n, 0);
    for (int i = 0; i < n_remove; i++) {
        if (remove_idx[i] < 0 || remove_idx[i] >= P.n) return -2;
        rm[remove_idx[i]] = 1;
    }
    int o = 0;
    for (int i = 0; i < P.n; i++) {
        if (rm[i] || (drop_host >= 0 && P.host[i] == drop_host) || (drop_oob && P.status[i] == 1)) continue;
        P.host[o] = (drop_host >= 0 && P.host[i] > drop_host) ? P.host[i] - 1 : P.host[i];
        P.status[o] = P.status[i]; P.u[o] = P.u[i]; P.v[o] = P.v[i]; P.energyTH[o] = P.energyTH[i];
        P.idepth_min[o] = P.idepth_min[i]; P.idepth_max[o] = P.idepth_max[i]; P.quality[o] = P.quality[i];
        P.type[o] = P.type[i];
        P.lastInterval[o] = P.lastInterval[i]; P.lastUV[2 * o] = P.lastUV[2 * i]; P.lastUV[2 * o + 1] = P.lastUV[2 * i + 1];
        for (int k = 0; k < 3; k++) P.gradH[3 * o + k] = P.gradH[3 * i + k];
        for (int k = 0; k < 8; k++) { P.color[8 * o + k] = P.color[8 * i + k]; P.weights[8 * o + k] = P.weights[8 * i + k]; }
        o++;
    }
    P.n = o;
    *n_out = o;
    return 0;
}

extern "C" int orc_immature_append(orc_ctx* c, int slot, int host_idx, int n, const float* u, const float* v, int* n_pool) {
    Pool& P = pool_of(c);
    std::vector<float> col(8 * (size_t)n), wgt(8 * (size_t)n), gh(3 * (size_t)n), eth(n);
    if (n > 0) orc_immature_create(c, slot, n, u, v, col.data(), wgt.data(), gh.data(), eth.data());
    const size_t N = (size_t)P.n + n;
    P.host.resize(N); P.status.resize(N); P.u.resize(N); P.v.resize(N); P.energyTH.resize(N); P.idepth_min.resize(N);
    P.idepth_max.resize(N); P.quality.resize(N); P.lastInterval.resize(N); P.lastUV.resize(2 * N); P.gradH.resize(3 * N);
    P.color.resize(8 * N); P.weights.resize(8 * N); P.type.resize(N);
    for (int i = 0; i < n; i++) {
        const int o = P.n + i;
        P.type[o] = 1;
        P.host[o] = host_idx; P.u[o] = u[i]; P.v[o] = v[i]; P.energyTH[o] = eth[i];
        P.status[o] = std::isfinite(eth[i]) ? 5 : 1;
        P.idepth_min[o] = 0; P.idepth_max[o] = NAN; P.quality[o] = 10000; P.lastInterval[o] = 0;
        P.lastUV[2 * o] = P.lastUV[2 * o + 1] = 0;
        for (int k = 0; k < 3; k++) P.gradH[3 * o + k] = gh[3 * i + k];
        for (int k = 0; k < 8; k++) { P.color[8 * o + k] = col[8 * i + k]; P.weights[8 * o + k] = wgt[8 * i + k]; }
    }
    P.n = (int)N;
    *n_pool = P.n;
    return 0;
}

// ================================================================================================
// N2 — CoarseDistanceMap (src/frontend/CoarseTracker.cc:795-930) and the candidate walk of
// FullSystem::activatePointsMT (src/frontend/FullSystem.cc:1226-1298)
extern "C" int orc_immature_set_types(orc_ctx* c, int first, int n, const float* type) {
    Pool& P = pool_of(c);
    if (first < 0 || n < 0 || first + n > P.n) return -2;
    for (int i = 0; i < n; i++) P.type[first + i] = type[i];
    return 0;
}

namespace {
// CoarseDistanceMap::growDistBFS (:828-923); the lists hold packed (x, y)
void grow_dist_bfs(orc_ctx* c, Pool& P, int bfsNum) {
    const int w1 = c->wG[1], h1 = c->hG[1];
    float* D = P.dist.data();
    for (int k = 1; k < 40; k++) {
        int bfsNum2 = bfsNum;
        std::swap(P.bfs1, P.bfs2);
        bfsNum = 0;
        auto visit = [&](int x, int y) {
            float& d = D[x + y * w1];
            if (d > k) { d = (float)k; P.bfs1[2 * bfsNum] = x; P.bfs1[2 * bfsNum + 1] = y; bfsNum++; }
        };
        for (int i = 0; i < bfsNum2; i++) {
            int x = P.bfs2[2 * i], y = P.bfs2[2 * i + 1];
            if (x == 0 || y == 0 || x == w1 - 1 || y == h1 - 1) continue;
            visit(x + 1, y); visit(x - 1, y); visit(x, y + 1); visit(x, y - 1);
            if (k % 2 != 0) { visit(x + 1, y + 1); visit(x - 1, y + 1); visit(x - 1, y - 1); visit(x + 1, y - 1); }
        }
    }
}
void dist_alloc(orc_ctx* c, Pool& P) {
    const size_t wh1 = (size_t)c->wG[1] * c->hG[1];
    if (P.dist.size() != wh1) { P.dist.assign(wh1, 1000.f); P.bfs1.assign(2 * wh1, 0); P.bfs2.assign(2 * wh1, 0); }
}
}  // namespace

extern "C" int orc_distmap_make(orc_ctx* c, int n_hosts, const float* KRKi, const float* Kt, int n_points, const int* host,
                                const float* u, const float* v, const float* idepth) {
    Pool& P = pool_of(c);
    dist_alloc(c, P);
    const int w1 = c->wG[1], h1 = c->hG[1];
    for (auto& d : P.dist) d = 1000;
    int numItems = 0;
    for (int i = 0; i < n_points; i++) {
        const int h = host[i];
        if (h < 0) continue;
        if (h >= n_hosts) return -2;
        const float *K = KRKi + 9 * h, *t = Kt + 3 * h;
        float ptp[3];
        for (int r = 0; r < 3; r++) ptp[r] = (K[3 * r] * u[i] + K[3 * r + 1] * v[i] + K[3 * r + 2] * 1) + t[r] * idepth[i];
        int x = ptp[0] / ptp[2] + 0.5f;
        int y = ptp[1] / ptp[2] + 0.5f;
        if (!(x > 0 && y > 0 && x < w1 && y < h1)) continue;
        P.dist[x + w1 * y] = 0;
        P.bfs1[2 * numItems] = x; P.bfs1[2 * numItems + 1] = y;
        numItems++;
        if (2 * (size_t)(numItems + 1) > P.bfs1.size()) { P.bfs1.resize(2 * P.bfs1.size()); P.bfs2.resize(P.bfs1.size()); }
    }
    grow_dist_bfs(c, P, numItems);
    return 0;
}

extern "C" int orc_distmap_get(orc_ctx* c, float* dist) {
    Pool& P = pool_of(c);
    dist_alloc(c, P);
    std::copy(P.dist.begin(), P.dist.end(), dist);
    return 0;
}

extern "C" int orc_immature_select_activation(orc_ctx* c, int n_hosts, const float* KRKi, const float* Kt,
                                              const unsigned char* host_skip, const unsigned char* host_flagged,
                                              float currentMinActDist, float minTraceQuality, unsigned char* action_out,
                                              int counts_out[2]) {
    Pool& P = pool_of(c);
    dist_alloc(c, P);
    const int w1 = c->wG[1], h1 = c->hG[1];
    counts_out[0] = counts_out[1] = 0;
    for (int i = 0; i < P.n; i++) {
        const int h = P.host[i];
        if (h < 0 || h >= n_hosts) return -2;
        action_out[i] = 0;
        if (host_skip[h]) continue;                                           // host == newestFr (:1231)
        const int st = P.status[i];
        if (!std::isfinite(P.idepth_max[i]) || st == 2) { action_out[i] = 1; counts_out[1]++; continue; }   // :1247-1251
        const bool canActivate = (st == 0 || st == 3 || st == 4 || st == 1) && P.lastInterval[i] < 8 &&
                                 P.quality[i] > minTraceQuality && (P.idepth_max[i] + P.idepth_min[i]) > 0;       // :1253-1259
        if (!canActivate) {
            if (host_flagged[h] || st == 1) { action_out[i] = 1; counts_out[1]++; }                           // :1261-1268
            continue;
        }
        const float *K = KRKi + 9 * h, *t = Kt + 3 * h;
        const float idp = 0.5f * (P.idepth_max[i] + P.idepth_min[i]);
        float ptp[3];
        for (int r = 0; r < 3; r++) ptp[r] = (K[3 * r] * P.u[i] + K[3 * r + 1] * P.v[i] + K[3 * r + 2] * 1) + t[r] * idp;
        const float px = ptp[0] / ptp[2], py = ptp[1] / ptp[2];
        int x = px + 0.5f;
        int y = py + 0.5f;
        if (x > 0 && y > 0 && x < w1 && y < h1) {
            // NOTE: the reference adds the fractional part of ptp[0] (not of ptp[0] / ptp[2]) — :1279-1280 reads
            // `ptp[0] - floorf((float)(ptp[0]))` on the homogeneous coordinate; kept
            float dist = P.dist[x + w1 * y] + (ptp[0] - floorf((float)(ptp[0])));
            if (dist >= currentMinActDist * P.type[i]) {
                P.bfs1[0] = x; P.bfs1[1] = y;                                // addIntoDistFinal (:925-930)
                P.dist[x + w1 * y] = 0;
                grow_dist_bfs(c, P, 1);
                action_out[i] = 2; counts_out[0]++;
            }
        } else {
            action_out[i] = 1; counts_out[1]++;
        }
    }
    return 0;
}
