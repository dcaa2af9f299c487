// k56_immature.cu — K5: ImmaturePoint::traceOn (reference src/internal/ImmaturePoint.cc:47-317) driven by
// FullSystem::traceNewCoarse (src/frontend/FullSystem.cc:1157-1195); K6: the ImmaturePoint constructor (:14-39),
// ImmaturePoint::linearizeResidual (:319-389) and FullSystem::optimizeImmaturePoint (FullSystem.cc:1035-1155).
//
// The pool of immature points is a device SoA.  Eight lanes cooperate on one point, one per pixel of the
// 8-pixel pattern; sums over the pattern are all-gathered in pattern order (seq8) so that every energy,
// Hessian and step is bit-identical to a sequential strict-IEEE evaluation of the reference loops and no
// status / accept decision can flip on summation order.  The epipolar search keeps its per-step energies in
// shared memory (100 floats per point) for the second-best scan of :213-219.
#include "immature.h"

#include <algorithm>

namespace {

constexpr int IM_THREADS = 128;              // 16 points per CTA
constexpr int IM_PTS = IM_THREADS / 8;


__device__ __forceinline__ float seq8(float v, int base) {
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < 8; j++) s += __shfl_sync(0xffffffffu, v, base + j);
    return s;
}
__device__ __forceinline__ float seq8_from(float s, float v, int base) {
#pragma unroll
    for (int j = 0; j < 8; j++) s += __shfl_sync(0xffffffffu, v, base + j);
    return s;
}
__device__ __forceinline__ void pattern(int k, int& pdx, int& pdy) {
    pdx = (k == 1 || k == 6) ? -1 : (k == 2) ? 1 : (k == 3) ? -2 : (k == 5) ? 2 : 0;
    pdy = (k == 0) ? -2 : (k == 1 || k == 2) ? -1 : (k == 6) ? 1 : (k == 7) ? 2 : 0;
}
// getInterpolatedElement33 (include/internal/GlobalFuncs.h:89-103) on the float4 texel layout
__device__ __forceinline__ float3 interp33(const float4* img, float x, float y, int w) {
    const int ix = (int)x, iy = (int)y;
    const float dx = x - ix, dy = y - iy, dxdy = dx * dy;
    const float4* bp = img + ix + iy * w;
    const float4 a = __ldg(bp + 1 + w), b = __ldg(bp + w), c = __ldg(bp + 1), d = __ldg(bp);
    const float w0 = dxdy, w1 = dy - dxdy, w2 = dx - dxdy, w3 = 1 - dx - dy + dxdy;
    return make_float3(w0 * a.x + w1 * b.x + w2 * c.x + w3 * d.x, w0 * a.y + w1 * b.y + w2 * c.y + w3 * d.y,
                       w0 * a.z + w1 * b.z + w2 * c.z + w3 * d.z);
}
__device__ __forceinline__ float interp31(const float4* img, float x, float y, int w) {   // GlobalFuncs.h:145-159
    const int ix = (int)x, iy = (int)y;
    const float dx = x - ix, dy = y - iy, dxdy = dx * dy;
    const float4* bp = img + ix + iy * w;
    return dxdy * __ldg(&bp[1 + w].x) + (dy - dxdy) * __ldg(&bp[w].x) + (dx - dxdy) * __ldg(&bp[1].x) +
           (1 - dx - dy + dxdy) * __ldg(&bp[0].x);
}

// ---- K6a: constructor ----------------------------------------------------------------------------------
__global__ void __launch_bounds__(IM_THREADS) k6_create_kernel(const float4* img, int w, int n, float thsc, float energyTH,
                                                               float* cr) {
    const int gid = blockIdx.x * IM_THREADS + threadIdx.x;
    const int pt = gid >> 3, k = gid & 7;
    const int lane = threadIdx.x & 31, base = lane & ~7;
    const bool live = pt < n;
    float col = 0, gx = 0, gy = 0, wgt = 0;
    if (live) {
        int pdx, pdy;
        pattern(k, pdx, pdy);
        const float x = cr[pt] + pdx, y = cr[n + pt] + pdy;
        // getInterpolatedElement33BiLin (GlobalFuncs.h:185-207)
        const int ix = (int)x, iy = (int)y;
        const float4* bp = img + ix + iy * w;
        const float tl = __ldg(&bp[0].x), tr = __ldg(&bp[1].x), bl = __ldg(&bp[w].x), br = __ldg(&bp[w + 1].x);
        const float dx = x - ix, dy = y - iy;
        const float topInt = dx * tr + (1 - dx) * tl, botInt = dx * br + (1 - dx) * bl;
        const float leftInt = dy * bl + (1 - dy) * tl, rightInt = dy * br + (1 - dy) * tr;
        col = dx * rightInt + (1 - dx) * leftInt;
        gx = rightInt - leftInt;
        gy = botInt - topInt;
        wgt = sqrtf(thsc / (thsc + (gx * gx + gy * gy)));
    }
    // the reference stops at the first non-finite colour: later pattern pixels contribute nothing
    const unsigned bal = (__ballot_sync(0xffffffffu, live && !isfinite(col)) >> base) & 0xffu;
    const int first_bad = bal ? __ffs(bal) - 1 : 8;
    const bool counted = k < first_bad;
    const float g00 = seq8(counted ? gx * gx : 0.f, base), g01 = seq8(counted ? gx * gy : 0.f, base),
                g11 = seq8(counted ? gy * gy : 0.f, base);
    if (!live) return;
    float* out = cr + 2 * (size_t)n;
    out[(size_t)pt * 20 + k] = (k <= first_bad) ? col : 0.f;
    out[(size_t)pt * 20 + 8 + k] = counted ? wgt : 0.f;
    if (k == 0) {
        out[(size_t)pt * 20 + 16] = g00; out[(size_t)pt * 20 + 17] = g01; out[(size_t)pt * 20 + 18] = g11;
        out[(size_t)pt * 20 + 19] = bal ? __int_as_float(0x7fc00000) : energyTH;
    }
}

// ---- K5: traceOn -----------------------------------------------------------------------------------------
struct K5Params {
    const float4* img;
    int n, w, h, gn_its, min_radius;
    float max_pix_search, slack_interval, stepsize, min_improvement, huberTH, gn_threshold, extra_slack;
    const int* host;
    int* status;
    const float *u, *v, *color, *weights, *gradH, *energyTH;
    float *idmin, *idmax, *quality, *lastUV, *interval;
    int* counts;         // [8] of this launch (zero on entry)
    int* counts_next;    // [8] of the next launch: zeroed here (the launches are ordered on one stream or by events)
    float tab[16 * LDSO_MAX_WINDOW];   // per host: KRKi[9] Kt[3] aff[2] pad[2] — rides in the kernel parameters, no copy
};

__global__ void __launch_bounds__(IM_THREADS) k5_trace_kernel(const K5Params p) {
    enum { GOOD = 0, OOB, OUTLIER, SKIPPED, BADCONDITION, UNINIT };
    __shared__ float errors[IM_PTS][100];
    __shared__ int cnt[8];   // [0..5] status histogram, [6] search steps, [7] points that reached the search
    if (threadIdx.x < 8) cnt[threadIdx.x] = 0;
    if (blockIdx.x == 0 && threadIdx.x < 8) p.counts_next[threadIdx.x] = 0;
    __syncthreads();
    const int gid = blockIdx.x * IM_THREADS + threadIdx.x;
    const int pt = gid >> 3, k = gid & 7;
    const int lane = threadIdx.x & 31, base = lane & ~7;
    const int slot = threadIdx.x >> 3;
    const bool live = pt < p.n;
    const int wG = p.w, hG = p.h;

    // every lane of the group runs the same scalar control flow; `done` parks the group until the end
    int result = -1;
    float out_u = -1, out_v = -1, out_int = 0;
    bool write_trace = false, write_depth = false;
    float idepth_min = 0, idepth_max = 0, quality = 0;
    int last = OOB;
    float uMin = 0, vMin = 0, uMax = 0, vMax = 0, dist = 0, dx = 0, dy = 0, errorInPixel = 0;
    float pr0 = 0, pr1 = 0, pr2 = 0, Kt0 = 0, Kt1 = 0, Kt2 = 0, aff0 = 0, aff1 = 0, R00 = 0, R01 = 0, R10 = 0, R11 = 0;
    const float maxPixSearch = (wG + hG) * p.max_pix_search;
    if (live) {
        last = p.status[pt];
        if (last == OOB) result = OOB;   // :53 (nothing is written)
    } else {
        result = -2;
    }
    if (result == -1) {
        write_trace = true;
        const float* T = p.tab + 16 * p.host[pt];
        Kt0 = T[9]; Kt1 = T[10]; Kt2 = T[11]; aff0 = T[12]; aff1 = T[13];
        R00 = T[0]; R01 = T[1]; R10 = T[3]; R11 = T[4];
        idepth_min = p.idmin[pt]; idepth_max = p.idmax[pt]; quality = p.quality[pt];
        const float uu = p.u[pt], vv = p.v[pt];
        pr0 = T[0] * uu + T[1] * vv + T[2] * 1; pr1 = T[3] * uu + T[4] * vv + T[5] * 1; pr2 = T[6] * uu + T[7] * vv + T[8] * 1;
        const float m0 = pr0 + Kt0 * idepth_min, m1 = pr1 + Kt1 * idepth_min, m2 = pr2 + Kt2 * idepth_min;
        uMin = m0 / m2; vMin = m1 / m2;
        if (!(uMin > 4 && vMin > 4 && uMin < wG - 5 && vMin < hG - 5)) result = OOB;
        if (result == -1) {
            if (isfinite(idepth_max)) {
                const float x0 = pr0 + Kt0 * idepth_max, x1 = pr1 + Kt1 * idepth_max, x2 = pr2 + Kt2 * idepth_max;
                uMax = x0 / x2; vMax = x1 / x2;
                if (!(uMax > 4 && vMax > 4 && uMax < wG - 5 && vMax < hG - 5)) result = OOB;
                if (result == -1) {
                    dist = (uMin - uMax) * (uMin - uMax) + (vMin - vMax) * (vMin - vMax);
                    dist = sqrtf(dist);
                    if (dist < p.slack_interval) {
                        out_u = (uMax + uMin) * 0.5f; out_v = (vMax + vMin) * 0.5f; out_int = dist;
                        result = SKIPPED;
                    }
                }
            } else {
                dist = maxPixSearch;
                const float x0 = pr0 + Kt0 * 0.01f, x1 = pr1 + Kt1 * 0.01f, x2 = pr2 + Kt2 * 0.01f;
                uMax = x0 / x2; vMax = x1 / x2;
                const float ddx = uMax - uMin, ddy = vMax - vMin;
                const float d = 1.0f / sqrtf(ddx * ddx + ddy * ddy);
                uMax = uMin + dist * ddx * d;
                vMax = vMin + dist * ddy * d;
                if (!(uMax > 4 && vMax > 4 && uMax < wG - 5 && vMax < hG - 5)) result = OOB;
            }
        }
        if (result == -1 && !(idepth_min < 0 || (m2 > 0.75f && m2 < 1.5f))) result = OOB;
        if (result == -1) {
            dx = p.stepsize * (uMax - uMin); dy = p.stepsize * (vMax - vMin);
            const float g00 = p.gradH[3 * pt], g01 = p.gradH[3 * pt + 1], g11 = p.gradH[3 * pt + 2];
            const float a = (dx * g00 + dy * g01) * dx + (dx * g01 + dy * g11) * dy;
            const float b = (dy * g00 + (-dx) * g01) * dy + (dy * g01 + (-dx) * g11) * (-dx);
            errorInPixel = 0.2f + 0.2f * (a + b) / a;
            if (errorInPixel * p.min_improvement > dist && isfinite(idepth_max)) {
                out_u = (uMax + uMin) * 0.5f; out_v = (vMax + vMin) * 0.5f; out_int = dist;
                result = BADCONDITION;
            }
        }
    }
    int numSteps = 0;
    float ptx = 0, pty = 0, rx = 0, ry = 0, mycolor = 0, myweight = 0;
    if (result == -1) {
        if (errorInPixel > 10) errorInPixel = 10;
        dx /= dist; dy /= dist;
        if (dist > maxPixSearch) {
            uMax = uMin + maxPixSearch * dx; vMax = vMin + maxPixSearch * dy;
            dist = maxPixSearch;
        }
        numSteps = (int)(1.9999f + dist / p.stepsize);
        const float randShift = uMin * 1000 - floorf(uMin * 1000);
        ptx = uMin - randShift * dx; pty = vMin - randShift * dy;
        int pdx, pdy;
        pattern(k, pdx, pdy);
        rx = R00 * pdx + R01 * pdy; ry = R10 * pdx + R11 * pdy;
        if (!isfinite(dx) || !isfinite(dy)) result = OOB;
        mycolor = p.color[8 * pt + k]; myweight = p.weights[8 * pt + k];
    }
    // ---- discrete search (:183-210): lane k evaluates pattern pixel k of every step ----
    float bestU = 0, bestV = 0, bestEnergy = 1e10f;
    int bestIdx = -1;
    if (numSteps >= 100) numSteps = 99;
    const bool search = result == -1;
    const unsigned grp = 0xffu << base;
    // groups of one warp may have different step counts: loop to the warp maximum so the shuffles stay converged
    int warpSteps = search ? numSteps : 0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) warpSteps = max(warpSteps, __shfl_xor_sync(0xffffffffu, warpSteps, o));
    const float refc = (float)(aff0 * mycolor + aff1);
    for (int i = 0; i < warpSteps; i++) {
        float term = 0.f;
        const bool act = search && i < numSteps;
        if (act) {
            const float hitColor = interp31(p.img, ptx + rx, pty + ry, wG);
            if (!isfinite(hitColor)) {
                term = 1e5f;
            } else {
                const float residual = hitColor - refc;
                const float hw = fabsf(residual) < p.huberTH ? 1 : p.huberTH / fabsf(residual);
                term = hw * residual * residual * (2 - hw);
            }
        }
        const float energy = seq8(term, base);
        if (act) {
            if (k == 0) errors[slot][i] = energy;
            if (energy < bestEnergy) { bestU = ptx; bestV = pty; bestEnergy = energy; bestIdx = i; }
            ptx += dx; pty += dy;
        }
    }
    __syncwarp();
    if (search) {
        // :213-219 — the eight lanes split the scan; min is order-independent
        float secondBest = 1e10f;
        for (int i = k; i < numSteps; i += 8)
            if ((i < bestIdx - p.min_radius || i > bestIdx + p.min_radius) && errors[slot][i] < secondBest) secondBest = errors[slot][i];
#pragma unroll
        for (int o = 4; o > 0; o >>= 1) secondBest = fminf(secondBest, __shfl_xor_sync(grp, secondBest, o));
        const float newQuality = secondBest / bestEnergy;
        if (newQuality < quality || numSteps > 10) quality = newQuality;
    }
    // ---- GN refinement (:222-268) ----
    float uBak = bestU, vBak = bestV, stepBack = 0;
    if (p.gn_its > 0) bestEnergy = 1e5f;
    bool gn_live = search;
    for (int it = 0; it < p.gn_its; it++) {
        if (!__any_sync(0xffffffffu, gn_live)) break;
        float tH = 0, tb = 0, te = 0;
        if (gn_live) {
            const float3 hit = interp33(p.img, bestU + rx, bestV + ry, wG);
            if (!isfinite(hit.x)) {
                te = 1e5f;
            } else {
                const float residual = hit.x - (aff0 * mycolor + aff1);
                const float dResdDist = dx * hit.y + dy * hit.z;
                const float hw = fabsf(residual) < p.huberTH ? 1 : p.huberTH / fabsf(residual);
                tH = hw * dResdDist * dResdDist;
                tb = hw * residual * dResdDist;
                te = myweight * myweight * hw * residual * residual * (2 - hw);
            }
        }
        const float H = seq8_from(1.f, tH, base), b = seq8(tb, base), energy = seq8(te, base);
        if (gn_live) {
            if (energy > bestEnergy) {
                stepBack *= 0.5f;
                bestU = uBak + stepBack * dx; bestV = vBak + stepBack * dy;
            } else {
                float step = -1.f * b / H;
                if (step < -0.5f) step = -0.5f;
                else if (step > 0.5f) step = 0.5f;
                if (!isfinite(step)) step = 0;
                uBak = bestU; vBak = bestV; stepBack = step;
                bestU += step * dx; bestV += step * dy;
                bestEnergy = energy;
            }
            if (fabsf(stepBack) < p.gn_threshold) gn_live = false;
        }
    }
    if (search) {
        if (!(bestEnergy < p.energyTH[pt] * p.extra_slack)) {
            result = (last == OUTLIER) ? OOB : OUTLIER;
        } else {
            if (dx * dx > dy * dy) {
                idepth_min = (pr2 * (bestU - errorInPixel * dx) - pr0) / (Kt0 - Kt2 * (bestU - errorInPixel * dx));
                idepth_max = (pr2 * (bestU + errorInPixel * dx) - pr0) / (Kt0 - Kt2 * (bestU + errorInPixel * dx));
            } else {
                idepth_min = (pr2 * (bestV - errorInPixel * dy) - pr1) / (Kt1 - Kt2 * (bestV - errorInPixel * dy));
                idepth_max = (pr2 * (bestV + errorInPixel * dy) - pr1) / (Kt1 - Kt2 * (bestV + errorInPixel * dy));
            }
            if (idepth_min > idepth_max) { const float t = idepth_min; idepth_min = idepth_max; idepth_max = t; }
            write_depth = true;
            if (!isfinite(idepth_min) || !isfinite(idepth_max) || (idepth_max < 0)) {
                result = OUTLIER;
            } else {
                out_int = 2 * errorInPixel; out_u = bestU; out_v = bestV;
                result = GOOD;
            }
        }
    }
    if (live && k == 0) {
        if (write_trace) {
            p.status[pt] = result;
            p.lastUV[2 * pt] = out_u; p.lastUV[2 * pt + 1] = out_v; p.interval[pt] = out_int;
            p.quality[pt] = quality;
            if (write_depth) { p.idmin[pt] = idepth_min; p.idmax[pt] = idepth_max; }
        }
        atomicAdd(&cnt[result], 1);
        if (numSteps > 0) { atomicAdd(&cnt[6], numSteps); atomicAdd(&cnt[7], 1); }
    }
    __syncthreads();
    if (threadIdx.x < 8 && cnt[threadIdx.x]) atomicAdd(&p.counts[threadIdx.x], cnt[threadIdx.x]);
}

// ---- K6b: activation (optimizeImmaturePoint) ----------------------------------------------------------------
struct K6Params {
    const float4* frame[LDSO_MAX_WINDOW];
    int n, n_cand, w, h, gn_its, min_obs;
    float fxl, fyl, cxl, cyl, fxli, fyli, huberTH, min_idepth_h;
    const float* precalc;
    const int *cand, *host;
    const float *u, *v, *color, *weights, *energyTH, *idmin, *idmax;
    float* idepth_out;
    int *ok_out, *state_out;
};

struct LinOut { float ret; int new_state; float new_energy; };

// ImmaturePoint::linearizeResidual (:319-389) for pattern pixel k of one group; all 32 lanes must call it.
__device__ __forceinline__ LinOut k6_lin_res(const K6Params& p, bool act, int st_in, float st_energy, const float* pc,
                                             const float4* img, float pu, float pv, float idepth, int k, int base, float col,
                                             float wgt, float eth_slack, float& Hdd, float& bd) {
    LinOut o;
    o.ret = st_energy; o.new_state = LDSO_RES_OOB; o.new_energy = 0.f;
    const bool run = act && st_in != LDSO_RES_OOB;
    bool bad = false;
    float tE = 0, tH = 0, tb = 0;
    if (run) {
        int pdx, pdy;
        pattern(k, pdx, pdy);
        const float K0 = (pu + pdx - p.cxl) * p.fxli, K1 = (pv + pdy - p.cyl) * p.fyli;
        const float x0 = pc[27] * K0 + pc[28] * K1 + pc[29] * 1 + pc[36] * idepth;
        const float x1 = pc[30] * K0 + pc[31] * K1 + pc[32] * 1 + pc[37] * idepth;
        const float x2 = pc[33] * K0 + pc[34] * K1 + pc[35] * 1 + pc[38] * idepth;
        const float drescale = 1.0f / x2;
        if (!(drescale > 0)) {
            bad = true;
        } else {
            const float u = x0 * drescale, v = x1 * drescale;
            const float Ku = u * p.fxl + p.cxl, Kv = v * p.fyl + p.cyl;
            if (!(Ku > 1.1f && Kv > 1.1f && Ku < (float)(p.w - 3) && Kv < (float)(p.h - 3))) {
                bad = true;
            } else {
                const float3 hit = interp33(img, Ku, Kv, p.w);
                if (!isfinite(hit.x)) {
                    bad = true;
                } else {
                    const float residual = hit.x - (pc[24] * col + pc[25]);
                    float hw = fabsf(residual) < p.huberTH ? 1 : p.huberTH / fabsf(residual);
                    tE = wgt * wgt * hw * residual * residual * (2 - hw);
                    const float dxInterp = hit.y * p.fxl, dyInterp = hit.z * p.fyl;
                    const float d_idepth =
                        (dxInterp * drescale * (pc[36] - pc[38] * u) + dyInterp * drescale * (pc[37] - pc[38] * v)) * 1.0f;
                    hw *= wgt * wgt;
                    tH = (hw * d_idepth) * d_idepth;
                    tb = (hw * residual) * d_idepth;
                }
            }
        }
    }
    // the reference returns at the first failing pattern pixel: earlier pixels have already been added to Hdd / bd
    const unsigned bal = (__ballot_sync(0xffffffffu, bad) >> base) & 0xffu;
    const int first_bad = bal ? __ffs(bal) - 1 : 8;
    const bool counted = run && k < first_bad;
    const float newH = seq8_from(Hdd, counted ? tH : 0.f, base);
    const float newb = seq8_from(bd, counted ? tb : 0.f, base);
    float energyLeft = seq8(counted ? tE : 0.f, base);
    if (run) {
        Hdd = newH; bd = newb;
        if (!bal) {
            if (energyLeft > eth_slack) { energyLeft = eth_slack; o.new_state = LDSO_RES_OUTLIER; }
            else o.new_state = LDSO_RES_IN;
            o.new_energy = energyLeft;
            o.ret = energyLeft;
        }
    }
    return o;
}

__global__ void __launch_bounds__(IM_THREADS) k6_activate_kernel(const K6Params p) {
    const int gid = blockIdx.x * IM_THREADS + threadIdx.x;
    const int q = gid >> 3, k = gid & 7;
    const int lane = threadIdx.x & 31, base = lane & ~7;
    const bool live = q < p.n_cand;
    int pt = 0, hst = -1;
    float pu = 0, pv = 0, col = 0, wgt = 0, eth = 0, currentIdepth = 0;
    if (live) {
        pt = p.cand[q]; hst = p.host[pt];
        pu = p.u[pt]; pv = p.v[pt]; col = p.color[8 * pt + k]; wgt = p.weights[8 * pt + k]; eth = p.energyTH[pt];
        currentIdepth = (p.idmax[pt] + p.idmin[pt]) * 0.5f;
    }
    int st[LDSO_MAX_WINDOW], nst[LDSO_MAX_WINDOW];
    float en[LDSO_MAX_WINDOW], nen[LDSO_MAX_WINDOW];
#pragma unroll
    for (int f = 0; f < LDSO_MAX_WINDOW; f++) { st[f] = LDSO_RES_IN; nst[f] = LDSO_RES_OUTLIER; en[f] = 0; nen[f] = 0; }
    float lastEnergy = 0, lastHdd = 0, lastbd = 0;
#pragma unroll 1
    for (int f = 0; f < p.n; f++) {
        const bool act = live && f != hst;
        const float* pc = p.precalc + (size_t)((act ? hst : 0) + p.n * f) * LDSO_PRECALC_STRIDE;
        const LinOut o = k6_lin_res(p, act, st[f], en[f], pc, p.frame[f], pu, pv, currentIdepth, k, base, col, wgt, eth * 1000,
                                    lastHdd, lastbd);
        if (act) { lastEnergy += o.ret; st[f] = o.new_state; en[f] = o.new_energy; nst[f] = o.new_state; nen[f] = o.new_energy; }
    }
    bool alive = live, ok = true;
    if (live && (!isfinite(lastEnergy) || lastHdd < p.min_idepth_h)) { alive = false; ok = false; }
    float lambda = 0.1f;
    for (int it = 0; it < p.gn_its; it++) {
        if (!__any_sync(0xffffffffu, alive)) break;
        float H = lastHdd;
        H *= 1 + lambda;
        const float step = (float)((1.0 / (double)H) * (double)lastbd);
        const float newIdepth = currentIdepth - step;
        float newHdd = 0, newbd = 0, newEnergy = 0;
#pragma unroll 1
        for (int f = 0; f < p.n; f++) {
            const bool act = alive && f != hst;
            const float* pc = p.precalc + (size_t)((act ? hst : 0) + p.n * f) * LDSO_PRECALC_STRIDE;
            const LinOut o = k6_lin_res(p, act, st[f], en[f], pc, p.frame[f], pu, pv, newIdepth, k, base, col, wgt, eth * 1, newHdd,
                                        newbd);
            if (act) { newEnergy += o.ret; nst[f] = o.new_state; if (o.new_state != LDSO_RES_OOB) nen[f] = o.new_energy; }
        }
        if (alive) {
            if (!isfinite(lastEnergy) || newHdd < p.min_idepth_h) {
                alive = false; ok = false;
            } else {
                if (newEnergy < lastEnergy) {
                    currentIdepth = newIdepth; lastHdd = newHdd; lastbd = newbd; lastEnergy = newEnergy;
#pragma unroll
                    for (int f = 0; f < LDSO_MAX_WINDOW; f++) { st[f] = nst[f]; en[f] = nen[f]; }
                    lambda *= 0.5f;
                } else {
                    lambda *= 5;
                }
                if ((double)fabsf(step) < 0.0001 * (double)currentIdepth) alive = false;
            }
        }
    }
    if (!live || k != 0) return;
    int numGood = 0;
    const bool fin = ok && isfinite(currentIdepth);
    for (int f = 0; f < p.n; f++) {
        const bool is_res = f != hst;
        if (is_res && st[f] == LDSO_RES_IN) numGood++;
        p.state_out[(size_t)q * p.n + f] = (fin && is_res) ? st[f] : -1;
    }
    const bool good = fin && numGood >= p.min_obs;
    p.ok_out[q] = good ? 1 : 0;
    p.idepth_out[q] = good ? currentIdepth : 0.f;
}

// ---- device-resident pool maintenance --------------------------------------------------------------------------

__global__ void im_flag_kernel(int n, const int* __restrict__ host, const int* __restrict__ status, int drop_host, int drop_oob,
                               int* __restrict__ keep) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    // keep[] was cleared to 0 and the explicit removals were marked -1 before this launch
    const bool rm = keep[i] < 0 || (drop_host >= 0 && host[i] == drop_host) || (drop_oob && status[i] == 1);
    keep[i] = rm ? 0 : 1;
}
__global__ void im_fill_kernel(int n, float v, float* __restrict__ dst) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) dst[i] = v;
}
__global__ void im_mark_kernel(int n_remove, const int* __restrict__ idx, int* __restrict__ keep) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_remove) keep[idx[i]] = -1;
}
// exclusive scan of keep[0..n) into pos[0..n], pos[n] = total; one CTA
__global__ void __launch_bounds__(1024) im_scan_kernel(int n, const int* __restrict__ keep, int* __restrict__ pos) {
    __shared__ int s[1024];
    __shared__ int carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int st = 0; st < n; st += 1024) {
        const int i = st + threadIdx.x;
        const int v = i < n ? keep[i] : 0;
        s[threadIdx.x] = v;
        __syncthreads();
        for (int o = 1; o < 1024; o <<= 1) {
            const int t = threadIdx.x >= o ? s[threadIdx.x - o] : 0;
            __syncthreads();
            s[threadIdx.x] += t;
            __syncthreads();
        }
        if (i < n) pos[i] = carry + s[threadIdx.x] - v;
        __syncthreads();
        if (threadIdx.x == 1023) carry += s[1023];
        __syncthreads();
    }
    if (threadIdx.x == 0) pos[n] = carry;
}
__global__ void im_scatter_kernel(int n, const int* __restrict__ keep, const int* __restrict__ pos, int drop_host, const PoolPtrs a,
                                  const PoolPtrs b) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n || !keep[i]) return;
    const int o = pos[i];
    const int h = a.host[i];
    b.host[o] = (drop_host >= 0 && h > drop_host) ? h - 1 : h;
    b.status[o] = a.status[i];
    b.u[o] = a.u[i]; b.v[o] = a.v[i]; b.energyTH[o] = a.energyTH[i];
    b.idmin[o] = a.idmin[i]; b.idmax[o] = a.idmax[i]; b.quality[o] = a.quality[i]; b.interval[o] = a.interval[i];
    b.type[o] = a.type[i];
    b.lastUV[2 * o] = a.lastUV[2 * i]; b.lastUV[2 * o + 1] = a.lastUV[2 * i + 1];
    b.gradH[3 * o] = a.gradH[3 * i]; b.gradH[3 * o + 1] = a.gradH[3 * i + 1]; b.gradH[3 * o + 2] = a.gradH[3 * i + 2];
    const float4* ca = reinterpret_cast<const float4*>(a.color + 8 * (size_t)i);
    const float4* wa = reinterpret_cast<const float4*>(a.weights + 8 * (size_t)i);
    float4* cb = reinterpret_cast<float4*>(b.color + 8 * (size_t)o);
    float4* wb = reinterpret_cast<float4*>(b.weights + 8 * (size_t)o);
    cb[0] = ca[0]; cb[1] = ca[1]; wb[0] = wa[0]; wb[1] = wa[1];
}
// new points of one keyframe: constructor results (cr scratch) -> pool rows [base, base + n)
__global__ void im_append_kernel(int n, int base, int host_idx, const float* __restrict__ cr, const PoolPtrs a) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float* out = cr + 2 * (size_t)n + 20 * (size_t)i;
    const int o = base + i;
    a.host[o] = host_idx;
    a.u[o] = cr[i]; a.v[o] = cr[n + i];
    for (int k = 0; k < 8; k++) { a.color[8 * (size_t)o + k] = out[k]; a.weights[8 * (size_t)o + k] = out[8 + k]; }
    a.gradH[3 * o] = out[16]; a.gradH[3 * o + 1] = out[17]; a.gradH[3 * o + 2] = out[18];
    const float eth = out[19];
    a.energyTH[o] = eth;
    a.status[o] = isfinite(eth) ? 5 : 1;     // a non-finite patch never becomes a point (FullSystem.cc:1456-1458): OOB
    a.idmin[o] = 0.f; a.idmax[o] = __int_as_float(0x7fc00000); a.quality[o] = 10000.f;
    a.lastUV[2 * o] = 0.f; a.lastUV[2 * o + 1] = 0.f; a.interval[o] = 0.f;
    a.type[o] = 1.f;
}

// ---- host side ------------------------------------------------------------------------------------------------
}  // namespace

// a trace launched by ldso_gpu_immature_trace_async is completed (stream sync, counts, profile units) before anything
// else touches the pool or its pinned staging buffer
int immature_trace_finalize(ldso_gpu_ctx* c) {
    ImmatureDev* P = (ImmatureDev*)c->immature;
    if (!P || !P->trace_pending) return 0;
    P->trace_pending = false;
    LDSO_CUDA_OK(c, ldso_stream_sync(P->trace_stream));
    P->trace_counts_fresh = false;
    if (P->trace_cnt_valid) {
        for (int i = 0; i < 6; i++) P->trace_counts[i] = P->trace_cnt[i];
        P->trace_counts_fresh = true;
        // profile unit of K5 = one epipolar search step (8 bilinear samples); the fixed per-point part counts as 9 steps
        if (P->trace_prof_idx >= 0 && P->trace_prof_idx < (int)c->prof_pending.size()) c->prof_pending[P->trace_prof_idx].units = (double)P->trace_cnt[6] + 9.0 * P->trace_cnt[7];
    }
    P->trace_prof_idx = -1;
    return 0;
}

// a frame slot is about to be rewritten on the tracking stream: if the pending trace (mapping stream) still reads it, the
// tracking stream waits for it on the device
int immature_trace_fence(ldso_gpu_ctx* c, int slot) {
    ImmatureDev* P = (ImmatureDev*)c->immature;
    if (P && P->trace_pending && P->trace_slot == slot && P->trace_stream != c->stream) LDSO_CUDA_OK(c, cudaStreamWaitEvent(c->stream, P->ev_done, 0));
    return 0;
}

ImmatureDev* immature_pool(ldso_gpu_ctx* c) {
    if (c->immature) {
        immature_trace_finalize(c);
        return (ImmatureDev*)c->immature;
    }
    ImmatureDev* P = new ImmatureDev();
    P->cap = c->cfg.max_immature > 16 ? c->cfg.max_immature : 16;
    const size_t N = P->cap;
    bool ok = true;
    auto da = [&](void** p, size_t bytes) { ok = ok && cudaMalloc(p, bytes) == cudaSuccess; };
    da((void**)&P->host, 4 * N); da((void**)&P->status, 4 * N); da((void**)&P->u, 4 * N); da((void**)&P->v, 4 * N);
    da((void**)&P->color, 32 * N); da((void**)&P->weights, 32 * N); da((void**)&P->gradH, 12 * N); da((void**)&P->energyTH, 4 * N);
    da((void**)&P->idmin, 4 * N); da((void**)&P->idmax, 4 * N); da((void**)&P->quality, 4 * N); da((void**)&P->lastUV, 8 * N);
    da((void**)&P->interval, 4 * N); da((void**)&P->type, 4 * N); da((void**)&P->type2, 4 * N);
    da((void**)&P->dmap, (size_t)c->wG[1] * c->hG[1] + 64); da((void**)&P->action, N); da((void**)&P->sel_cnt, 16);
    da((void**)&P->host2, 4 * N); da((void**)&P->status2, 4 * N); da((void**)&P->u2, 4 * N); da((void**)&P->v2, 4 * N);
    da((void**)&P->color2, 32 * N); da((void**)&P->weights2, 32 * N); da((void**)&P->gradH2, 12 * N); da((void**)&P->energyTH2, 4 * N);
    da((void**)&P->idmin2, 4 * N); da((void**)&P->idmax2, 4 * N); da((void**)&P->quality2, 4 * N); da((void**)&P->lastUV2, 8 * N);
    da((void**)&P->interval2, 4 * N); da((void**)&P->keep, 4 * (N + 1)); da((void**)&P->pos, 4 * (N + 1));
    da((void**)&P->tables, 4 * 16 * LDSO_MAX_WINDOW); da((void**)&P->counts, 64);
    da((void**)&P->cand, 4 * N); da((void**)&P->act_idepth, 4 * N); da((void**)&P->act_ok, 4 * N);
    // act_state + 4096: the activation walk's segmented candidate arrays (n2_distmap.cu)
    da((void**)&P->act_state, 4 * (N * LDSO_MAX_WINDOW + 4096)); da((void**)&P->cr, 4 * 22 * N);
    P->src_cap = (int)(8 * N);
    da((void**)&P->src_cells, 4 * (size_t)P->src_cap); da((void**)&P->src_n, 32);   // [0] length, [1..4] corner values after a walk
    if (ok) { cudaMemset(P->src_n, 0, 32); cudaMemset(P->counts, 0, 64); }
    P->pinned_bytes = 4 * 32 * N + 4096;
    ok = ok && cudaMallocHost(&P->pinned, P->pinned_bytes) == cudaSuccess;
    ok = ok && cudaMallocHost((void**)&P->trace_cnt, 32) == cudaSuccess;
    ok = ok && cudaEventCreateWithFlags(&P->ev_ready, cudaEventDisableTiming) == cudaSuccess;
    ok = ok && cudaEventCreateWithFlags(&P->ev_done, cudaEventDisableTiming) == cudaSuccess;
    c->immature = P;
    if (!ok) {   // free what was allocated and leave no half-built pool behind
        immature_free(c);
        c->set_error("immature pool allocation failed", __FILE__, __LINE__);
        return nullptr;
    }
    return P;
}


void immature_free(ldso_gpu_ctx* c) {
    ImmatureDev* P = (ImmatureDev*)c->immature;
    if (!P) return;
    void* ptrs[] = {P->host, P->status, P->u, P->v, P->color, P->weights, P->gradH, P->energyTH, P->idmin, P->idmax, P->quality,
                    P->lastUV, P->interval, P->tables, P->counts, P->cand, P->act_idepth, P->act_ok, P->act_state, P->cr,
                    P->host2, P->status2, P->u2, P->v2, P->color2, P->weights2, P->gradH2, P->energyTH2, P->idmin2, P->idmax2,
                    P->quality2, P->lastUV2, P->interval2, P->keep, P->pos, P->type, P->type2, P->dmap, P->action, P->sel_cnt,
                    P->src_cells, P->src_n};
    for (void* p : ptrs) cudaFree(p);
    if (P->pinned) cudaFreeHost(P->pinned);
    if (P->trace_cnt) cudaFreeHost(P->trace_cnt);
    if (P->ev_ready) cudaEventDestroy(P->ev_ready);
    if (P->ev_done) cudaEventDestroy(P->ev_done);
    delete P;
    c->immature = nullptr;
}

extern "C" int ldso_gpu_immature_create(ldso_gpu_ctx* c, int slot, int n, const float* u, const float* v, float* color,
                                        float* weights, float* grad_h, float* energy_th) {
    LDSO_ENTER(c);
    ImmatureDev* P = immature_pool(c);
    if (!P) return -1;
    LDSO_REQUIRE(c, slot >= 0 && slot < c->cfg.n_frame_slots && n >= 0 && n <= P->cap);
    if (n == 0) return 0;
    LDSO_REQUIRE(c, u && v && color && weights && grad_h && energy_th);
    float* st = (float*)P->pinned;
    memcpy(st, u, 4 * (size_t)n); memcpy(st + n, v, 4 * (size_t)n);
    LDSO_CUDA_OK(c, ldso_copy(c, P->cr, st, 8 * (size_t)n, cudaMemcpyHostToDevice, c->stream));
    float eth = 8 * c->cfg.outlier_th;
    eth *= c->cfg.overall_energy_th_weight * c->cfg.overall_energy_th_weight;
    k6_create_kernel<<<(8 * n + IM_THREADS - 1) / IM_THREADS, IM_THREADS, 0, c->stream>>>(c->frame[slot], c->wG[0], n,
                                                                                         c->cfg.outlier_th_sum_component, eth, P->cr);
    c->launches++;
    LDSO_CUDA_OK(c, cudaGetLastError());
    LDSO_CUDA_OK(c, ldso_copy(c, st, P->cr + 2 * (size_t)n, 80 * (size_t)n, cudaMemcpyDeviceToHost, c->stream));
    LDSO_CUDA_OK(c, ldso_stream_sync(c->stream));
    for (int i = 0; i < n; i++) {
        const float* r = st + 20 * (size_t)i;
        memcpy(color + 8 * i, r, 32); memcpy(weights + 8 * i, r + 8, 32); memcpy(grad_h + 3 * i, r + 16, 12);
        energy_th[i] = r[19];
    }
    return 0;
}

extern "C" int ldso_gpu_immature_set(ldso_gpu_ctx* c, int n, const int* host, const float* u, const float* v, const float* color,
                                     const float* weights, const float* grad_h, const float* energy_th, const float* idepth_min,
                                     const float* idepth_max, const float* quality, const int* status) {
    LDSO_ENTER(c);
    ImmatureDev* P = immature_pool(c);
    if (!P) return -1;
    LDSO_REQUIRE(c, n >= 0 && n <= P->cap);
    P->n = n;
    if (n == 0) return 0;
    LDSO_REQUIRE(c, host && u && v && color && weights && grad_h && energy_th && idepth_min && idepth_max && quality && status);
    char* st = (char*)P->pinned;
    size_t off = 0;
    const size_t N = n;
    auto put = [&](void* dst, const void* src, size_t bytes) -> cudaError_t {
        memcpy(st + off, src, bytes);
        cudaError_t e = ldso_copy(c, dst, st + off, bytes, cudaMemcpyHostToDevice, c->stream);
        off += bytes;
        return e;
    };
    LDSO_CUDA_OK(c, put(P->host, host, 4 * N)); LDSO_CUDA_OK(c, put(P->status, status, 4 * N));
    LDSO_CUDA_OK(c, put(P->u, u, 4 * N)); LDSO_CUDA_OK(c, put(P->v, v, 4 * N));
    LDSO_CUDA_OK(c, put(P->color, color, 32 * N)); LDSO_CUDA_OK(c, put(P->weights, weights, 32 * N));
    LDSO_CUDA_OK(c, put(P->gradH, grad_h, 12 * N)); LDSO_CUDA_OK(c, put(P->energyTH, energy_th, 4 * N));
    LDSO_CUDA_OK(c, put(P->idmin, idepth_min, 4 * N)); LDSO_CUDA_OK(c, put(P->idmax, idepth_max, 4 * N));
    LDSO_CUDA_OK(c, put(P->quality, quality, 4 * N));
    LDSO_CUDA_OK(c, cudaMemsetAsync(P->lastUV, 0, 8 * N, c->stream));
    LDSO_CUDA_OK(c, cudaMemsetAsync(P->interval, 0, 4 * N, c->stream));
    im_fill_kernel<<<(n + 255) / 256, 256, 0, c->stream>>>(n, 1.f, P->type);   // my_type: see ldso_gpu_immature_set_types
    LDSO_CUDA_OK(c, ldso_stream_sync(c->stream));
    return 0;
}

static int trace_launch(ldso_gpu_ctx* c, int new_slot, int n_hosts, const float* KRKi, const float* Kt, const float* aff, bool async) {
    ImmatureDev* P = immature_pool(c);          // completes a pending trace first
    if (!P) return -1;
    LDSO_REQUIRE(c, new_slot >= 0 && new_slot < c->cfg.n_frame_slots && n_hosts >= 1 && n_hosts <= LDSO_MAX_WINDOW && KRKi && Kt && aff);
    for (int i = 0; i < 6; i++) P->trace_counts[i] = 0;
    if (P->n == 0) { P->trace_counts_fresh = true; return 0; }
    // async: the trace runs on the mapping stream, behind everything queued on the tracking stream so far (the new frame's
    // pyramid, earlier pool updates), and next to whatever the tracking stream gets afterwards (the next frame's tracker
    // occupies 32 of the SMs)
    static const bool one_stream = getenv("LDSO_TRACE_ON_TRACKING_STREAM") != nullptr;   // measurement knob: no second stream per sequence
    if (one_stream) async = false;
    cudaStream_t S = async ? c->stream_map : c->stream;
    if (async) {
        LDSO_CUDA_OK(c, cudaEventRecord(P->ev_ready, c->stream));
        LDSO_CUDA_OK(c, cudaStreamWaitEvent(S, P->ev_ready, 0));
    }
    K5Params p;
    for (int h = 0; h < n_hosts; h++) {
        float* T = p.tab + 16 * h;
        memcpy(T, KRKi + 9 * h, 36); memcpy(T + 9, Kt + 3 * h, 12); memcpy(T + 12, aff + 2 * h, 8);
        T[14] = T[15] = 0;
    }
    for (int i = 16 * n_hosts; i < 16 * LDSO_MAX_WINDOW; i++) p.tab[i] = 0;
    c->h2d_bytes += 64 * (long long)n_hosts;
    p.img = c->frame[new_slot]; p.n = P->n; p.w = c->wG[0]; p.h = c->hG[0];
    p.gn_its = c->cfg.trace_gn_iterations; p.min_radius = c->cfg.min_trace_test_radius;
    p.max_pix_search = c->cfg.max_pix_search; p.slack_interval = c->cfg.trace_slack_interval; p.stepsize = c->cfg.trace_stepsize;
    p.min_improvement = c->cfg.trace_min_improvement; p.huberTH = c->cfg.huber_th; p.gn_threshold = c->cfg.trace_gn_threshold;
    p.extra_slack = c->cfg.trace_extra_slack_on_th;
    p.host = P->host; p.status = P->status; p.u = P->u; p.v = P->v; p.color = P->color; p.weights = P->weights; p.gradH = P->gradH;
    p.energyTH = P->energyTH; p.idmin = P->idmin; p.idmax = P->idmax; p.quality = P->quality;
    p.lastUV = P->lastUV; p.interval = P->interval;
    P->trace_parity ^= 1;                      // two count blocks: each launch zeroes the one of the next launch
    p.counts = P->counts + 8 * P->trace_parity; p.counts_next = P->counts + 8 * (P->trace_parity ^ 1);
    prof_begin(c, S);
    k5_trace_kernel<<<(8 * P->n + IM_THREADS - 1) / IM_THREADS, IM_THREADS, 0, S>>>(p);
    P->trace_prof_idx = prof_end(c, S, LDSO_PROF_K5, P->n);
    c->launches++;
    LDSO_CUDA_OK(c, cudaGetLastError());
    // the status counts come back only when somebody asks (ldso_gpu_immature_trace_counts) or the profile needs its units
    P->trace_cnt_valid = false;
    if (c->profile_on) {
        LDSO_CUDA_OK(c, ldso_copy(c, P->trace_cnt, p.counts, 32, cudaMemcpyDeviceToHost, S));
        P->trace_cnt_valid = true;
    }
    LDSO_CUDA_OK(c, cudaEventRecord(P->ev_done, S));
    P->trace_stream = S; P->trace_slot = new_slot;
    P->trace_pending = true;
    P->trace_launched = true;
    return 0;
}

extern "C" int ldso_gpu_immature_trace(ldso_gpu_ctx* c, int new_slot, int n_hosts, const float* KRKi, const float* Kt,
                                       const float* aff, int counts_out[6]) {
    LDSO_ENTER(c);
    LDSO_REQUIRE(c, counts_out);
    if (trace_launch(c, new_slot, n_hosts, KRKi, Kt, aff, false)) return -1;
    return ldso_gpu_immature_trace_counts(c, counts_out);
}

extern "C" int ldso_gpu_immature_trace_async(ldso_gpu_ctx* c, int new_slot, int n_hosts, const float* KRKi, const float* Kt,
                                             const float* aff) {
    LDSO_ENTER(c);
    return trace_launch(c, new_slot, n_hosts, KRKi, Kt, aff, true);
}

extern "C" int ldso_gpu_immature_trace_counts(ldso_gpu_ctx* c, int counts_out[6]) {
    LDSO_ENTER(c);
    LDSO_REQUIRE(c, counts_out);
    if (immature_trace_finalize(c)) return -1;
    ImmatureDev* P = immature_pool(c);
    if (!P) return -1;
    if (!P->trace_counts_fresh && P->trace_launched) {   // the counts of the last trace are still on the device
        LDSO_CUDA_OK(c, ldso_copy(c, P->trace_cnt, P->counts + 8 * P->trace_parity, 32, cudaMemcpyDeviceToHost, c->stream));
        LDSO_CUDA_OK(c, ldso_stream_sync(c->stream));
        for (int i = 0; i < 6; i++) P->trace_counts[i] = P->trace_cnt[i];
        P->trace_counts_fresh = true;
    }
    for (int i = 0; i < 6; i++) counts_out[i] = P->trace_counts[i];
    return 0;
}

extern "C" int ldso_gpu_immature_get(ldso_gpu_ctx* c, float* idepth_min, float* idepth_max, float* quality, int* status,
                                     float* last_trace_uv, float* last_trace_pixel_interval) {
    LDSO_ENTER(c);
    ImmatureDev* P = immature_pool(c);
    if (!P) return -1;
    const size_t N = P->n;
    if (N == 0) return 0;
    char* st = (char*)P->pinned;
    size_t off = 0;
    struct Job { void* dst; size_t off, bytes; } jobs[6];
    int nj = 0;
    auto get = [&](void* dst, const void* src, size_t bytes) -> cudaError_t {
        if (!dst) return cudaSuccess;
        jobs[nj++] = {dst, off, bytes};
        cudaError_t e = ldso_copy(c, st + off, src, bytes, cudaMemcpyDeviceToHost, c->stream);
        off += bytes;
        return e;
    };
    LDSO_CUDA_OK(c, get(idepth_min, P->idmin, 4 * N)); LDSO_CUDA_OK(c, get(idepth_max, P->idmax, 4 * N));
    LDSO_CUDA_OK(c, get(quality, P->quality, 4 * N)); LDSO_CUDA_OK(c, get(status, P->status, 4 * N));
    LDSO_CUDA_OK(c, get(last_trace_uv, P->lastUV, 8 * N)); LDSO_CUDA_OK(c, get(last_trace_pixel_interval, P->interval, 4 * N));
    LDSO_CUDA_OK(c, ldso_stream_sync(c->stream));
    for (int i = 0; i < nj; i++) memcpy(jobs[i].dst, st + jobs[i].off, jobs[i].bytes);
    return 0;
}

extern "C" int ldso_gpu_immature_activate(ldso_gpu_ctx* c, int n_cand, const int* cand, int min_obs, float* idepth_out,
                                          int* ok_out, int* res_state_out) {
    LDSO_ENTER(c);
    ImmatureDev* P = immature_pool(c);
    if (!P) return -1;
    WindowDev& W = c->win;
    LDSO_REQUIRE(c, n_cand >= 0 && n_cand <= P->cap && W.n >= 1);
    if (n_cand == 0) return 0;
    LDSO_REQUIRE(c, cand && idepth_out && ok_out && res_state_out);
    for (int i = 0; i < n_cand; i++) LDSO_REQUIRE(c, cand[i] >= 0 && cand[i] < P->n);
    int* st = (int*)P->pinned;
    memcpy(st, cand, 4 * (size_t)n_cand);
    LDSO_CUDA_OK(c, ldso_copy(c, P->cand, st, 4 * (size_t)n_cand, cudaMemcpyHostToDevice, c->stream));
    K6Params p;
    for (int f = 0; f < LDSO_MAX_WINDOW; f++) p.frame[f] = f < W.n ? c->frame[W.slots[f]] : nullptr;
    p.n = W.n; p.n_cand = n_cand; p.w = c->wG[0]; p.h = c->hG[0]; p.gn_its = c->cfg.gn_its_on_point_activation; p.min_obs = min_obs;
    p.fxl = W.fxl; p.fyl = W.fyl; p.cxl = W.cxl; p.cyl = W.cyl; p.fxli = W.fxli; p.fyli = W.fyli; p.huberTH = c->cfg.huber_th;
    p.min_idepth_h = c->cfg.min_idepth_h_act;
    p.precalc = W.precalc; p.cand = P->cand; p.host = P->host; p.u = P->u; p.v = P->v; p.color = P->color; p.weights = P->weights;
    p.energyTH = P->energyTH; p.idmin = P->idmin; p.idmax = P->idmax;
    p.idepth_out = P->act_idepth; p.ok_out = P->act_ok; p.state_out = P->act_state;
    prof_begin(c, c->stream);
    k6_activate_kernel<<<(8 * n_cand + IM_THREADS - 1) / IM_THREADS, IM_THREADS, 0, c->stream>>>(p);
    prof_end(c, c->stream, LDSO_PROF_K6, n_cand);
    c->launches++;
    LDSO_CUDA_OK(c, cudaGetLastError());
    const size_t N = n_cand;
    LDSO_CUDA_OK(c, ldso_copy(c, st, P->act_idepth, 4 * N, cudaMemcpyDeviceToHost, c->stream));
    LDSO_CUDA_OK(c, ldso_copy(c, st + N, P->act_ok, 4 * N, cudaMemcpyDeviceToHost, c->stream));
    LDSO_CUDA_OK(c, ldso_copy(c, st + 2 * N, P->act_state, 4 * N * W.n, cudaMemcpyDeviceToHost, c->stream));
    LDSO_CUDA_OK(c, ldso_stream_sync(c->stream));
    memcpy(idepth_out, st, 4 * N); memcpy(ok_out, st + N, 4 * N); memcpy(res_state_out, st + 2 * N, 4 * N * W.n);
    return 0;
}

static PoolPtrs pool_ptrs(ImmatureDev* P, bool alt) {
    if (!alt) return {P->host, P->status, P->u, P->v, P->color, P->weights, P->gradH, P->energyTH, P->idmin, P->idmax, P->quality,
                      P->lastUV, P->interval, P->type};
    return {P->host2, P->status2, P->u2, P->v2, P->color2, P->weights2, P->gradH2, P->energyTH2, P->idmin2, P->idmax2, P->quality2,
            P->lastUV2, P->interval2, P->type2};
}

extern "C" int ldso_gpu_immature_compact(ldso_gpu_ctx* c, int drop_host, int drop_oob, int n_remove, const int* remove_idx,
                                         int* n_out) {
    LDSO_ENTER(c);
    ImmatureDev* P = immature_pool(c);
    if (!P) return -1;
    LDSO_REQUIRE(c, n_remove >= 0 && n_remove <= P->n && (n_remove == 0 || remove_idx) && n_out);
    const int n = P->n;
    if (n == 0) { *n_out = 0; return 0; }
    for (int i = 0; i < n_remove; i++) LDSO_REQUIRE(c, remove_idx[i] >= 0 && remove_idx[i] < n);
    cudaStream_t s = c->stream;
    LDSO_CUDA_OK(c, cudaMemsetAsync(P->keep, 0, 4 * (size_t)(n + 1), s));
    if (n_remove) {
        int* st = (int*)P->pinned;
        memcpy(st, remove_idx, 4 * (size_t)n_remove);
        LDSO_CUDA_OK(c, ldso_copy(c, P->cand, st, 4 * (size_t)n_remove, cudaMemcpyHostToDevice, s));
        im_mark_kernel<<<(n_remove + 255) / 256, 256, 0, s>>>(n_remove, P->cand, P->keep);
        c->launches++;
    }
    im_flag_kernel<<<(n + 255) / 256, 256, 0, s>>>(n, P->host, P->status, drop_host, drop_oob, P->keep);
    im_scan_kernel<<<1, 1024, 0, s>>>(n, P->keep, P->pos);
    im_scatter_kernel<<<(n + 255) / 256, 256, 0, s>>>(n, P->keep, P->pos, drop_host, pool_ptrs(P, false), pool_ptrs(P, true));
    c->launches += 3;
    LDSO_CUDA_OK(c, cudaGetLastError());
    int* total = (int*)P->pinned + P->cap;
    LDSO_CUDA_OK(c, ldso_copy(c, total, P->pos + n, 4, cudaMemcpyDeviceToHost, s));
    LDSO_CUDA_OK(c, ldso_stream_sync(s));
    std::swap(P->host, P->host2); std::swap(P->status, P->status2); std::swap(P->u, P->u2); std::swap(P->v, P->v2);
    std::swap(P->color, P->color2); std::swap(P->weights, P->weights2); std::swap(P->gradH, P->gradH2);
    std::swap(P->energyTH, P->energyTH2); std::swap(P->idmin, P->idmin2); std::swap(P->idmax, P->idmax2);
    std::swap(P->quality, P->quality2); std::swap(P->lastUV, P->lastUV2); std::swap(P->interval, P->interval2); std::swap(P->type, P->type2);
    P->n = *total;
    *n_out = P->n;
    return 0;
}

extern "C" int ldso_gpu_immature_append(ldso_gpu_ctx* c, int slot, int host_idx, int n, const float* u, const float* v,
                                        int* n_pool) {
    LDSO_ENTER(c);
    ImmatureDev* P = immature_pool(c);
    if (!P) return -1;
    LDSO_REQUIRE(c, slot >= 0 && slot < c->cfg.n_frame_slots && n >= 0 && P->n + n <= P->cap && n_pool);
    LDSO_REQUIRE(c, host_idx >= 0 && host_idx < LDSO_MAX_WINDOW && (n == 0 || (u && v)));
    if (n > 0) {
        float* st = (float*)ldso_staging(c, 8 * (size_t)n);   // queued: nothing waits here
        LDSO_REQUIRE(c, st != nullptr);
        memcpy(st, u, 4 * (size_t)n); memcpy(st + n, v, 4 * (size_t)n);
        LDSO_CUDA_OK(c, ldso_copy(c, P->cr, st, 8 * (size_t)n, cudaMemcpyHostToDevice, c->stream));
        float eth = 8 * c->cfg.outlier_th;
        eth *= c->cfg.overall_energy_th_weight * c->cfg.overall_energy_th_weight;
        k6_create_kernel<<<(8 * n + IM_THREADS - 1) / IM_THREADS, IM_THREADS, 0, c->stream>>>(c->frame[slot], c->wG[0], n,
                                                                                             c->cfg.outlier_th_sum_component, eth, P->cr);
        im_append_kernel<<<(n + 255) / 256, 256, 0, c->stream>>>(n, P->n, host_idx, P->cr, pool_ptrs(P, false));
        c->launches += 2;
        LDSO_CUDA_OK(c, cudaGetLastError());
        P->n += n;
    }
    *n_pool = P->n;
    return 0;
}
