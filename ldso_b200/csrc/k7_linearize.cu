// k7_linearize.cu — K7: PointFrameResidual::linearize (reference src/internal/Residuals.cc:15-217),
// applyRes / takeData (include/internal/Residuals.h:70-88,124-129), resetOOB (:63-67),
// fixLinearizationF (Residuals.cc:219-245) and the isNew bookkeeping of
// FullSystem::linearizeAll_Reductor (src/frontend/FullSystem.cc:1786-1820).
//
// Eight lanes cooperate on one residual, one per pixel of the 8-pixel pattern: each lane projects its
// pattern pixel, fetches 4 float4 texels of the target frame's level 0 and forms its row of the
// Jacobian.  The 2x2 inner-product blocks and the energy are summed in pattern order 0..7 (products transposed
// through shared memory, one lane per sum), so every stored value is bit-identical to a sequential strict-IEEE
// evaluation of the reference loop (the OUTLIER test `energy > frameEnergyTH` cannot flip on summation
// order).  Output is one 288-byte RawResidualJacobian record per residual, written as full 32-byte sectors.
#include "ctx.h"

namespace {

constexpr int K7_THREADS = 256;

struct K7Params {
    const float4* frame[LDSO_MAX_WINDOW];
    float eth[LDSO_MAX_WINDOW];
    int n, nr, w, h, fix, fixA, fixB;
    float fxl, fyl, cxl, cyl, fxli, fyli, huberTH, thsc;
    const float* precalc;
    const float *u, *v, *idepth, *idepth_zero, *color, *weights;
    const int *host, *r_point, *r_target;
    int *state, *new_state, *flags;
    float *energy, *new_energy, *new_energy_wo, *J, *JpJdF, *center, *max_rel_baseline, *num_good;
    double* stats;
};

constexpr int K7_GROUP_STRIDE = 8 * 12 + 8;   // floats per residual in the transpose buffer: 4 groups of a warp land on disjoint banks

__global__ void __launch_bounds__(K7_THREADS) k7_linearize_kernel(const K7Params p) {
    const int gid = blockIdx.x * K7_THREADS + threadIdx.x;
    const int r = gid >> 3, k = gid & 7;
    const int lane = threadIdx.x & 31, base = lane & ~7;
    bool live = false;
    int pt = 0, hst = 0, tgt = 0, fl = 0;
    if (r < p.nr) {
        fl = p.flags[r];
        live = !(fl & LDSO_RF_LINEARIZED);   // activeResiduals = every residual that is not linearised (FullSystem.cc:864)
        pt = p.r_point[r]; hst = p.host[pt]; tgt = p.r_target[r];
    }
    const int st_in = live ? p.state[r] : LDSO_RES_OOB;
    bool oob = !live || st_in == LDSO_RES_OOB;
    const float* pc = p.precalc + (size_t)(hst + p.n * tgt) * LDSO_PRECALC_STRIDE;
    const float wM3 = (float)(p.w - 3), hM3 = (float)(p.h - 3);
    float sh[32];   // the 32 floats shared by the record: Jpdxi[2][6], Jpdc[2][4], Jpdd[2], JIdx2, JabJIdx, Jab2
    float cKu = 0, cKv = 0, cId = 0;
    float pu = 0, pv = 0, pid = 0;
    if (!oob) {
        pu = p.u[pt]; pv = p.v[pt]; pid = p.idepth[pt];
        const float idz = p.idepth_zero[pt];
        // projectPoint at the FEJ point (ResidualProjections.h:57-84)
        const float K0 = (pu + 0 - p.cxl) * p.fxli, K1 = (pv + 0 - p.cyl) * p.fyli;
        const float x0 = pc[0] * K0 + pc[1] * K1 + pc[2] * 1 + pc[9] * idz;
        const float x1 = pc[3] * K0 + pc[4] * K1 + pc[5] * 1 + pc[10] * idz;
        const float x2 = pc[6] * K0 + pc[7] * K1 + pc[8] * 1 + pc[11] * idz;
        const float drescale = 1.0f / x2;
        const float new_idepth = idz * drescale;
        if (!(drescale > 0)) oob = true;
        const float u = x0 * drescale, v = x1 * drescale;
        cKu = u * p.fxl + p.cxl; cKv = v * p.fyl + p.cyl; cId = new_idepth;
        if (!(cKu > 1.1f && cKv > 1.1f && cKu < wM3 && cKv < hM3)) oob = true;
        // Residuals.cc:59-106
        sh[24] = drescale * (pc[9] - pc[11] * u) * 1.0f * p.fxl;
        sh[25] = drescale * (pc[10] - pc[11] * v) * 1.0f * p.fyl;
        float cx2 = drescale * (pc[6] * u - pc[0]);
        float cx3 = p.fxl * drescale * (pc[7] * u - pc[1]) * p.fyli;
        float cx0 = K0 * cx2, cx1 = K1 * cx3;
        float cy2 = p.fyl * drescale * (pc[6] * v - pc[3]) * p.fxli;
        float cy3 = drescale * (pc[7] * v - pc[4]);
        float cy0 = K0 * cy2, cy1 = K1 * cy3;
        sh[12] = (cx0 + u) * 50.0f; sh[13] = cx1 * 50.0f; sh[14] = (cx2 + 1) * 50.0f; sh[15] = cx3 * 50.0f;
        sh[16] = cy0 * 50.0f; sh[17] = (cy1 + v) * 50.0f; sh[18] = cy2 * 50.0f; sh[19] = (cy3 + 1) * 50.0f;
        sh[0] = new_idepth * p.fxl; sh[1] = 0; sh[2] = -new_idepth * u * p.fxl; sh[3] = -u * v * p.fxl;
        sh[4] = (1 + u * u) * p.fxl; sh[5] = -v * p.fxl;
        sh[6] = 0; sh[7] = new_idepth * p.fyl; sh[8] = -new_idepth * v * p.fyl; sh[9] = -(1 + v * v) * p.fyl;
        sh[10] = u * v * p.fyl; sh[11] = u * p.fyl;
    }
    // pattern pixel k at the current state (ResidualProjections.h:24-33, Residuals.cc:128-190)
    const int pdx = (k == 1 || k == 6) ? -1 : (k == 2) ? 1 : (k == 3) ? -2 : (k == 5) ? 2 : 0;
    const int pdy = (k == 0) ? -2 : (k == 1 || k == 2) ? -1 : (k == 6) ? 1 : (k == 7) ? 2 : 0;
    float resF = 0, h1 = 0, h2 = 0, ja0 = 0, ja1 = 0, e_k = 0;
    bool lane_oob = false;
    if (!oob) {
        const float up = pu + pdx, vp = pv + pdy;
        const float q0 = pc[12] * up + pc[13] * vp + pc[14] * 1 + pc[21] * pid;
        const float q1 = pc[15] * up + pc[16] * vp + pc[17] * 1 + pc[22] * pid;
        const float q2 = pc[18] * up + pc[19] * vp + pc[20] * 1 + pc[23] * pid;
        const float Ku = q0 / q2, Kv = q1 / q2;
        if (!(Ku > 1.1f && Kv > 1.1f && Ku < wM3 && Kv < hM3)) {
            lane_oob = true;
        } else {
            const int ix = (int)Ku, iy = (int)Kv;
            const float dx = Ku - ix, dy = Kv - iy, dxdy = dx * dy;
            const float4* bp = p.frame[tgt] + ix + iy * p.w;
            const float4 a = __ldg(bp + 1 + p.w), b = __ldg(bp + p.w), c4 = __ldg(bp + 1), d = __ldg(bp);
            const float w0 = dxdy, w1 = dy - dxdy, w2 = dx - dxdy, w3 = 1 - dx - dy + dxdy;
            const float hc0 = w0 * a.x + w1 * b.x + w2 * c4.x + w3 * d.x;
            float hc1 = w0 * a.y + w1 * b.y + w2 * c4.y + w3 * d.y;
            float hc2 = w0 * a.z + w1 * b.z + w2 * c4.z + w3 * d.z;
            const float col = p.color[8 * pt + k];
            const float residual = hc0 - (pc[24] * col + pc[25]);
            const float drdA = col - pc[26];
            if (!isfinite(hc0)) lane_oob = true;
            float wgt = sqrtf(p.thsc / (p.thsc + (hc1 * hc1 + hc2 * hc2)));
            wgt = 0.5f * (wgt + p.weights[8 * pt + k]);
            float hw = fabsf(residual) < p.huberTH ? 1 : p.huberTH / fabsf(residual);
            e_k = wgt * wgt * hw * residual * residual * (2 - hw);
            if (hw < 1) hw = sqrtf(hw);
            hw = hw * wgt;
            hc1 *= hw; hc2 *= hw;
            resF = residual * hw; h1 = hc1; h2 = hc2; ja0 = drdA * hw; ja1 = hw;
        }
    }
    const unsigned bal = __ballot_sync(0xffffffffu, lane_oob);
    if ((bal >> base) & 0xffu) oob = true;
    // ordered sums over the pattern: the 8 lanes of a residual hold 12 products each; they are transposed through shared
    // memory so that lane q sums product q (lanes 0..3 also product 8 + q) over the pattern in lane order — the
    // reference's `+=` over idx = 0..7 — and the 12 sums go back to every lane the same way.  ~40 instructions instead of
    // the 96 shuffles + 96 adds of an all-gather per sum.
    float drdA_l = 0.f;
    if (!oob && !lane_oob) drdA_l = p.color[8 * pt + k] - pc[26];   // drdA*drdA*hw*hw: the exact product, not recovered from ja0
    __shared__ __align__(16) float s_prod[K7_THREADS / 8][K7_GROUP_STRIDE];
    __shared__ __align__(16) float s_sum[K7_THREADS / 8][12];
    {
        float* mine = s_prod[threadIdx.x >> 3] + 12 * k;
        reinterpret_cast<float4*>(mine)[0] = make_float4(h1 * h1, h2 * h2, h1 * h2, ja0 * h1);
        reinterpret_cast<float4*>(mine)[1] = make_float4(ja0 * h2, ja1 * h1, ja1 * h2, drdA_l * drdA_l * ja1 * ja1);
        reinterpret_cast<float4*>(mine)[2] = make_float4(ja0 * ja1, ja1 * ja1, ja1 * ja1 * (h1 * h1 + h2 * h2), e_k);
        __syncwarp();
        const float* col = s_prod[threadIdx.x >> 3] + k;
        float sa = 0.f, sb = 0.f;
#pragma unroll
        for (int j = 0; j < 8; j++) sa += col[12 * j];
        if (k < 4) {
#pragma unroll
            for (int j = 0; j < 8; j++) sb += col[12 * j + 8];
        }
        s_sum[threadIdx.x >> 3][k] = sa;
        if (k < 4) s_sum[threadIdx.x >> 3][8 + k] = sb;
        __syncwarp();
    }
    const float4 q0 = reinterpret_cast<const float4*>(s_sum[threadIdx.x >> 3])[0];
    const float4 q1 = reinterpret_cast<const float4*>(s_sum[threadIdx.x >> 3])[1];
    const float4 q2 = reinterpret_cast<const float4*>(s_sum[threadIdx.x >> 3])[2];
    const float s00 = q0.x, s11 = q0.y, s10 = q0.z, a00 = q0.w, a01 = q1.x, a10 = q1.y, a11 = q1.z, b00 = q1.w;
    const float b01 = q2.x, b11 = q2.y, wJI2 = q2.z;
    float energyLeft = q2.w;

    double ret_energy = 0.0;
    int isIN = 0;
    if (live) {
        if (oob) {
            // state_NewState = OOB, return state_energy (Residuals.cc:19-23, 61-64, 134-137, 146-149)
            if (k == 0) {
                p.new_energy_wo[r] = -1.f;
                p.new_state[r] = LDSO_RES_OOB;
                ret_energy = (double)p.energy[r];
                if (p.fix && st_in != LDSO_RES_OOB) {   // applyRes(true) with NewState == OOB
                    p.flags[r] = fl & ~LDSO_RF_ACTIVE;
                    p.state[r] = LDSO_RES_OOB;
                    p.energy[r] = p.new_energy[r];
                }
            }
        } else {
            float* Jr = p.J + (size_t)r * LDSO_J_FLOATS;
            Jr[k] = resF; Jr[8 + k] = h1; Jr[16 + k] = h2;
            Jr[24 + k] = p.fixA ? 0.f : ja0;
            Jr[32 + k] = p.fixB ? 0.f : ja1;
            sh[20] = s00; sh[21] = s10; sh[22] = s11;
            sh[23] = a00; sh[26] = a01; sh[27] = a10; sh[28] = a11;
            sh[29] = b00; sh[30] = b01; sh[31] = b11;
            // record floats 40..71 = Jpdxi0[6] Jpdxi1[6] Jpdc0[4] Jpdc1[4] Jpdd[2] JIdx2[3] JabJIdx[4] Jab2[3]
            float4 o;
            switch (k) {
                case 0: o = make_float4(sh[0], sh[1], sh[2], sh[3]); break;
                case 1: o = make_float4(sh[4], sh[5], sh[6], sh[7]); break;
                case 2: o = make_float4(sh[8], sh[9], sh[10], sh[11]); break;
                case 3: o = make_float4(sh[12], sh[13], sh[14], sh[15]); break;
                case 4: o = make_float4(sh[16], sh[17], sh[18], sh[19]); break;
                case 5: o = make_float4(sh[24], sh[25], sh[20], sh[21]); break;
                case 6: o = make_float4(sh[22], sh[23], sh[26], sh[27]); break;
                default: o = make_float4(sh[28], sh[29], sh[30], sh[31]); break;
            }
            reinterpret_cast<float4*>(Jr + 40)[k] = o;
            const float th = fmaxf(p.eth[hst], p.eth[tgt]);
            const float ewo = energyLeft;
            int ns;
            if (energyLeft > th || wJI2 < 2) { energyLeft = th; ns = LDSO_RES_OUTLIER; }
            else ns = LDSO_RES_IN;
            if (k == 0) {
                p.new_energy_wo[r] = ewo;
                p.new_state[r] = ns;
                p.new_energy[r] = energyLeft;
                reinterpret_cast<float4*>(p.center)[r] = make_float4(cKu, cKv, cId, 0.f);
                ret_energy = (double)energyLeft;
                isIN = ns == LDSO_RES_IN;
            }
            if (p.fix) {
                // applyRes(true) + isNew bookkeeping (st_in != OOB here)
                if (ns == LDSO_RES_IN) {
                    // takeData: JpJdF = [Jpdxi^T (JIdx2 Jpdd); JabJIdx Jpdd]
                    const float t0 = s00 * sh[24] + s10 * sh[25], t1 = s10 * sh[24] + s11 * sh[25];
                    float val;
                    switch (k) {
                        case 0: val = sh[0] * t0 + sh[6] * t1; break;
                        case 1: val = sh[1] * t0 + sh[7] * t1; break;
                        case 2: val = sh[2] * t0 + sh[8] * t1; break;
                        case 3: val = sh[3] * t0 + sh[9] * t1; break;
                        case 4: val = sh[4] * t0 + sh[10] * t1; break;
                        case 5: val = sh[5] * t0 + sh[11] * t1; break;
                        case 6: val = a00 * sh[24] + a01 * sh[25]; break;
                        default: val = a10 * sh[24] + a11 * sh[25]; break;
                    }
                    p.JpJdF[8 * (size_t)r + k] = val;
                }
                if (k == 0) {
                    int nf = ns == LDSO_RES_IN ? (fl | LDSO_RF_ACTIVE) : (fl & ~LDSO_RF_ACTIVE);
                    p.flags[r] = nf;
                    p.state[r] = ns;
                    p.energy[r] = energyLeft;
                    if ((nf & LDSO_RF_ACTIVE) && (nf & LDSO_RF_NEW)) {   // FullSystem.cc:1799-1813
                        const float i0 = pc[12] * pu + pc[13] * pv + pc[14] * 1, i1 = pc[15] * pu + pc[16] * pv + pc[17] * 1,
                                    i2 = pc[18] * pu + pc[19] * pv + pc[20] * 1;
                        const float r0 = i0 + pc[21] * pid, r1 = i1 + pc[22] * pid, r2 = i2 + pc[23] * pid;
                        const float d0 = i0 / i2 - r0 / r2, d1 = i1 / i2 - r1 / r2;
                        const float relBS = (float)(0.01 * (double)sqrtf(d0 * d0 + d1 * d1));
                        atomicMax(reinterpret_cast<int*>(p.max_rel_baseline) + pt, __float_as_int(relBS));
                        atomicAdd(p.num_good + pt, 1.0f);
                    }
                }
            }
        }
    }
    // stats: [sum of returned energies, #IN]  (FullSystem.cc:1786-1790)
    double e = ret_energy;
    int cnt = isIN;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        e += __shfl_xor_sync(0xffffffffu, e, o);
        cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    }
    __shared__ double s_e[K7_THREADS / 32];
    __shared__ int s_c[K7_THREADS / 32];
    if (lane == 0) { s_e[threadIdx.x >> 5] = e; s_c[threadIdx.x >> 5] = cnt; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double te = 0; int tc = 0;
        for (int i = 0; i < K7_THREADS / 32; i++) { te += s_e[i]; tc += s_c[i]; }
        atomicAdd(p.stats, te);
        atomicAdd(p.stats + 1, (double)tc);
    }
}

// PointFrameResidual::applyRes(true) + takeData (include/internal/Residuals.h:70-88,124-129)
__global__ void k7_apply_res_kernel(int nr, int* __restrict__ state, const int* __restrict__ new_state, int* __restrict__ flags,
                                    float* __restrict__ energy, const float* __restrict__ new_energy, const float* __restrict__ J,
                                    float* __restrict__ JpJdF) {
    int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= nr) return;
    int fl = flags[r];
    if (fl & LDSO_RF_LINEARIZED) return;
    if (state[r] == LDSO_RES_OOB) return;
    int ns = new_state[r];
    if (ns == LDSO_RES_IN) {
        const float* j = J + (size_t)r * LDSO_J_FLOATS;
        float t0 = j[62] * j[60] + j[63] * j[61], t1 = j[63] * j[60] + j[64] * j[61];
        float* o = JpJdF + 8 * (size_t)r;
#pragma unroll
        for (int i = 0; i < 6; i++) o[i] = j[40 + i] * t0 + j[46 + i] * t1;
        o[6] = j[65] * j[60] + j[66] * j[61];
        o[7] = j[67] * j[60] + j[68] * j[61];
        flags[r] = fl | LDSO_RF_ACTIVE;
    } else {
        flags[r] = fl & ~LDSO_RF_ACTIVE;
    }
    state[r] = ns;
    energy[r] = new_energy[r];
}

// PointFrameResidual::resetOOB (include/internal/Residuals.h:63-67) for every non-linearised residual
__global__ void k7_reset_oob_kernel(int nr, int* __restrict__ state, int* __restrict__ new_state, const int* __restrict__ flags,
                                    float* __restrict__ energy, float* __restrict__ new_energy) {
    int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= nr || (flags[r] & LDSO_RF_LINEARIZED)) return;
    energy[r] = 0.f; new_energy[r] = 0.f;
    new_state[r] = LDSO_RES_OUTLIER;
    state[r] = LDSO_RES_IN;
}

// PointFrameResidual::fixLinearizationF (src/internal/Residuals.cc:219-245)
__global__ void k7_fix_lin_kernel(int nr, int n, const unsigned char* __restrict__ mask, const int* __restrict__ r_point,
                                  const int* __restrict__ r_target, const int* __restrict__ host, int* __restrict__ flags,
                                  const float* __restrict__ J, const float* __restrict__ adHTdelta, float4 cDelta,
                                  const float* __restrict__ idepth, const float* __restrict__ idepth_zero, float* __restrict__ rtz) {
    int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= nr || !mask[r]) return;
    const int pt = r_point[r];
    const float* j = J + (size_t)r * LDSO_J_FLOATS;
    const float* dp = adHTdelta + 8 * (host[pt] + n * r_target[r]);
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
    for (int i = 0; i < 8; i++) {
        float v = j[i];
        v = v - j[8 + i] * Jx;
        v = v - j[16 + i] * Jy;
        v = v - j[24 + i] * dp[6];
        v = v - j[32 + i] * dp[7];
        rtz[8 * (size_t)r + i] = v;
    }
    flags[r] |= LDSO_RF_LINEARIZED;
}

}  // namespace

int k7_linearize(ldso_gpu_ctx* c, int fix) {
    WindowDev& W = c->win;
    K7Params p;
    for (int i = 0; i < W.n; i++) { p.frame[i] = c->frame[W.slots[i]]; p.eth[i] = W.frameEnergyTH[i]; }
    p.n = W.n; p.nr = W.nr; p.w = c->wG[0]; p.h = c->hG[0]; p.fix = fix;
    p.fixA = c->cfg.affine_opt_mode_a < 0; p.fixB = c->cfg.affine_opt_mode_b < 0;
    p.fxl = W.fxl; p.fyl = W.fyl; p.cxl = W.cxl; p.cyl = W.cyl; p.fxli = W.fxli; p.fyli = W.fyli;
    p.huberTH = c->cfg.huber_th; p.thsc = c->cfg.outlier_th_sum_component;
    p.precalc = W.precalc; p.u = W.u; p.v = W.v; p.idepth = W.idepth; p.idepth_zero = W.idepth_zero; p.color = W.color;
    p.weights = W.weights; p.host = W.host; p.r_point = W.r_point; p.r_target = W.r_target; p.state = W.state;
    p.new_state = W.new_state; p.flags = W.flags; p.energy = W.energy; p.new_energy = W.new_energy;
    p.new_energy_wo = W.new_energy_wo; p.J = W.J; p.JpJdF = W.JpJdF; p.center = W.center;
    p.max_rel_baseline = W.max_rel_baseline; p.num_good = W.num_good; p.stats = W.stats;
    LDSO_CUDA_OK(c, cudaMemsetAsync(W.stats, 0, 4 * sizeof(double), c->stream));
    if (W.nr > 0) {
        int threads = W.nr * 8;
        prof_begin(c, c->stream);
        k7_linearize_kernel<<<(threads + K7_THREADS - 1) / K7_THREADS, K7_THREADS, 0, c->stream>>>(p);
        prof_end(c, c->stream, LDSO_PROF_K7, W.nr);
        c->launches++;
    }
    LDSO_CUDA_OK(c, cudaGetLastError());
    return 0;
}

int k7_apply_res(ldso_gpu_ctx* c) {
    WindowDev& W = c->win;
    if (W.nr == 0) return 0;
    k7_apply_res_kernel<<<(W.nr + 255) / 256, 256, 0, c->stream>>>(W.nr, W.state, W.new_state, W.flags, W.energy, W.new_energy,
                                                                   W.J, W.JpJdF);
    c->launches++;
    LDSO_CUDA_OK(c, cudaGetLastError());
    return 0;
}

int k7_reset_oob(ldso_gpu_ctx* c) {
    WindowDev& W = c->win;
    if (W.nr == 0) return 0;
    k7_reset_oob_kernel<<<(W.nr + 255) / 256, 256, 0, c->stream>>>(W.nr, W.state, W.new_state, W.flags, W.energy, W.new_energy);
    c->launches++;
    LDSO_CUDA_OK(c, cudaGetLastError());
    return 0;
}

int k7_fix_linearization(ldso_gpu_ctx* c, const unsigned char* mask_dev) {
    WindowDev& W = c->win;
    if (W.nr == 0) return 0;
    float4 cd = make_float4(W.cDeltaF[0], W.cDeltaF[1], W.cDeltaF[2], W.cDeltaF[3]);
    k7_fix_lin_kernel<<<(W.nr + 255) / 256, 256, 0, c->stream>>>(W.nr, W.n, mask_dev, W.r_point, W.r_target, W.host, W.flags, W.J,
                                                                 W.adHTdelta, cd, W.idepth, W.idepth_zero, W.rtz);
    c->launches++;
    LDSO_CUDA_OK(c, cudaGetLastError());
    return 0;
}
