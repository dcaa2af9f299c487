// k9_sc.cu — K9: AccumulatedSCHessianSSE::addPoint (reference
// src/internal/OptimizationBackend/AccumulatedSCHessian.cc:9-51) and K10:
// EnergyFunctional::resubstituteFPt (src/internal/OptimizationBackend/EnergyFunctional.cc:638-667).
//
// K9: one warp per point.  All residuals of a point share its host frame, and points are stored
// contiguously per host (allPoints order), so a CTA works on points of ONE host and keeps that host's
// slice of the Schur accumulators in shared memory: accD[h][t1][t2] (n^2 8x8 blocks), accE[h][t] (8x4),
// accEB[h][t] (8) plus accHcc/accbc.  The warp first folds the per-residual Hdd/bd/Hcd pieces written by
// K8 in residual order (bit-identical to the reference's sequential +=), derives HdiF / bdSumF, then for
// every active residual pair adds the 8x8 outer product (HdiF*JpJdF_1) JpJdF_2^T with the 64 entries
// spread over the lanes (conflict-free shared-memory atomics).  The CTA finally adds its non-zero
// entries into the fp64 global accumulators.
#include "ctx.h"

namespace {

constexpr int K9_WARPS = 8;

__global__ void __launch_bounds__(K9_WARPS * 32)
k9_sc_kernel(int n, const int* __restrict__ blocks, const int* __restrict__ res_begin, const int* __restrict__ flags,
             const int* __restrict__ r_target, const float* __restrict__ res_pt, const float* __restrict__ JpJdF,
             const float* __restrict__ prior, const float* __restrict__ idepth, const float* __restrict__ idepth_zero,
             float* __restrict__ pt_out, float* __restrict__ max_rel_baseline, int cap_p, double* __restrict__ acc_sc) {
    extern __shared__ float sm[];
    const int nn = n * n;
    float* sD = sm;                  // [n*n][64]  index (t1 + n*t2)
    float* sE = sD + nn * 64;        // [n][32]
    float* sEB = sE + n * 32;        // [n][8]
    float* sHcc = sEB + n * 8;       // [16]
    float* sbc = sHcc + 16;          // [4]
    const int total = nn * 64 + n * 40 + 20;
    for (int i = threadIdx.x; i < total; i += blockDim.x) sm[i] = 0.f;
    __syncthreads();
    const int h = blocks[3 * blockIdx.x], p0 = blocks[3 * blockIdx.x + 1], p1 = blocks[3 * blockIdx.x + 2];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    for (int p = p0 + wid; p < p1; p += K9_WARPS) {
        const int rb = res_begin[p], re = res_begin[p + 1];
        float HddA = 0, bdA = 0, HddL = 0, bdL = 0, HcdA[4] = {0, 0, 0, 0}, HcdL[4] = {0, 0, 0, 0};
        int ngood = 0;
        for (int r = rb; r < re; r++) {
            const int fl = flags[r];
            if (!(fl & LDSO_RF_ACTIVE)) continue;
            ngood++;
            const float4 a = __ldg(reinterpret_cast<const float4*>(res_pt + 8 * (size_t)r));
            const float4 b = __ldg(reinterpret_cast<const float4*>(res_pt + 8 * (size_t)r) + 1);
            if (b.z == 0.f) continue;   // not seen by K8 (cannot happen for an active residual)
            if (fl & LDSO_RF_LINEARIZED) {
                HddL += a.x; bdL += a.y; HcdL[0] += a.z; HcdL[1] += a.w; HcdL[2] += b.x; HcdL[3] += b.y;
            } else {
                HddA += a.x; bdA += a.y; HcdA[0] += a.z; HcdA[1] += a.w; HcdA[2] += b.x; HcdA[3] += b.y;
            }
        }
        float HdiF = 0.f, bdSum = 0.f, idh = 0.f;
        if (ngood > 0) {
            float H = HddA + HddL + prior[p];
            if (H < 1e-10) H = 1e-10;
            idh = H;
            HdiF = (float)(1.0 / (double)H);
            bdSum = bdA + bdL;
            bdSum += prior[p] * (idepth[p] - idepth_zero[p]);   // shiftPriorToZero
        }
        if (lane == 0) {
            pt_out[PT_HDD_A * cap_p + p] = HddA; pt_out[PT_BD_A * cap_p + p] = bdA;
            pt_out[PT_HDD_L * cap_p + p] = HddL; pt_out[PT_BD_L * cap_p + p] = bdL;
            for (int i = 0; i < 4; i++) {
                pt_out[(PT_HCD_A + i) * cap_p + p] = HcdA[i];
                pt_out[(PT_HCD_L + i) * cap_p + p] = HcdL[i];
            }
            pt_out[PT_HDI * cap_p + p] = HdiF; pt_out[PT_BDSUM * cap_p + p] = bdSum; pt_out[PT_IDH * cap_p + p] = idh;
            if (ngood == 0) max_rel_baseline[p] = 0.f;
        }
        if (ngood == 0) continue;
        const float Hcd[4] = {HcdA[0] + HcdL[0], HcdA[1] + HcdL[1], HcdA[2] + HcdL[2], HcdA[3] + HcdL[3]};
        if (lane < 16) atomicAdd(sHcc + lane, (HdiF * Hcd[lane >> 2]) * Hcd[lane & 3]);
        if (lane < 4) atomicAdd(sbc + lane, (bdSum * HdiF) * Hcd[lane]);
        for (int r1 = rb; r1 < re; r1++) {
            if (!(flags[r1] & LDSO_RF_ACTIVE)) continue;
            const int t1 = r_target[r1];
            const float* j1 = JpJdF + 8 * (size_t)r1;
            const float wa = HdiF * j1[lane >> 3];          // entries e = lane      -> row lane/8
            const float wb = HdiF * j1[4 + (lane >> 3)];    // entries e = lane + 32 -> row 4 + lane/8
            for (int r2 = rb; r2 < re; r2++) {
                if (!(flags[r2] & LDSO_RF_ACTIVE)) continue;
                const float c2 = JpJdF[8 * (size_t)r2 + (lane & 7)];
                float* d = sD + (t1 + n * r_target[r2]) * 64;
                atomicAdd(d + lane, wa * c2);
                atomicAdd(d + lane + 32, wb * c2);
            }
            atomicAdd(sE + t1 * 32 + lane, (HdiF * j1[lane >> 2]) * Hcd[lane & 3]);
            if (lane < 8) atomicAdd(sEB + t1 * 8 + lane, (HdiF * bdSum) * j1[lane]);
        }
    }
    __syncthreads();
    // flush to the fp64 accumulators: accD[(h + n t1 + n^2 t2)][64] | accE[(h + n t)][32] | accEB[(h + n t)][8] | Hcc | bc
    const size_t offE = (size_t)nn * n * 64, offEB = offE + (size_t)nn * 32, offH = offEB + (size_t)nn * 8;
    for (int i = threadIdx.x; i < total; i += blockDim.x) {
        const float v = sm[i];
        if (v == 0.f) continue;
        size_t g;
        if (i < nn * 64) {
            const int blk = i >> 6, e = i & 63, t1 = blk % n, t2 = blk / n;
            g = ((size_t)(h + n * t1 + nn * t2)) * 64 + e;
        } else if (i < nn * 64 + n * 32) {
            const int q = i - nn * 64, t = q >> 5;
            g = offE + (size_t)(h + n * t) * 32 + (q & 31);
        } else if (i < nn * 64 + n * 40) {
            const int q = i - nn * 64 - n * 32, t = q >> 3;
            g = offEB + (size_t)(h + n * t) * 8 + (q & 7);
        } else {
            g = offH + (i - nn * 64 - n * 40);
        }
        atomicAdd(acc_sc + g, (double)v);
    }
}

// K10: p->step = -(bdSumF - xc.Hcd - sum_r xAd[h,t].JpJdF_r) * HdiF / (1 + lambda)
__global__ void k10_resubstitute_kernel(int n, int np, const int* __restrict__ res_begin, const int* __restrict__ flags,
                                        const int* __restrict__ r_target, const int* __restrict__ host,
                                        const float* __restrict__ JpJdF, const float* __restrict__ xAd, float* __restrict__ pt_out,
                                        int cap_p, float lambda) {
    int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= np) return;
    const int rb = res_begin[p], re = res_begin[p + 1];
    int ngood = 0;
    for (int r = rb; r < re; r++) ngood += (flags[r] & LDSO_RF_ACTIVE) ? 1 : 0;
    if (ngood == 0) { pt_out[PT_STEP * cap_p + p] = 0.f; return; }
    float b = pt_out[PT_BDSUM * cap_p + p];
    const float* xc = xAd + (size_t)n * n * 8;
    float d = 0.f;
#pragma unroll
    for (int i = 0; i < 4; i++) d += xc[i] * (pt_out[(PT_HCD_A + i) * cap_p + p] + pt_out[(PT_HCD_L + i) * cap_p + p]);
    b -= d;
    const int h = host[p];
    for (int r = rb; r < re; r++) {
        if (!(flags[r] & LDSO_RF_ACTIVE)) continue;
        const float* xa = xAd + 8 * (h * n + r_target[r]);
        const float* jp = JpJdF + 8 * (size_t)r;
        float s = 0.f;
#pragma unroll
        for (int i = 0; i < 8; i++) s += xa[i] * jp[i];
        b -= s;
    }
    if (!isfinite(b)) return;   // the reference abandons the loop here; the point keeps its previous step
    pt_out[PT_STEP * cap_p + p] = (float)((double)(-b * pt_out[PT_HDI * cap_p + p]) / (1.0 + (double)lambda));
}

}  // namespace

int k9_accumulate_sc(ldso_gpu_ctx* c) {
    WindowDev& W = c->win;
    LDSO_CUDA_OK(c, cudaMemsetAsync(W.acc_sc, 0, sizeof(double) * W.acc_sc_size, c->stream));
    // CTA table: points of one host, 64 per CTA (built by window_set_points)
    const int nblk = W.n_sc_blocks;
    if (nblk > 0) {
        const int n = W.n;
        size_t smem = sizeof(float) * (size_t)(n * n * 64 + n * 40 + 20);
        static bool attr = false;
        if (!attr) { cudaFuncSetAttribute(k9_sc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024); attr = true; }
        k9_sc_kernel<<<nblk, K9_WARPS * 32, smem, c->stream>>>(n, W.sc_blocks, W.res_begin, W.flags, W.r_target,
                                                               W.res_pt, W.JpJdF, W.prior, W.idepth, W.idepth_zero, W.pt_out,
                                                               W.max_rel_baseline, W.cap_p, W.acc_sc);
        c->launches++;
    }
    LDSO_CUDA_OK(c, cudaGetLastError());
    return 0;
}

int k10_resubstitute(ldso_gpu_ctx* c, float lambda) {
    WindowDev& W = c->win;
    if (W.np == 0) return 0;
    k10_resubstitute_kernel<<<(W.np + 127) / 128, 128, 0, c->stream>>>(W.n, W.np, W.res_begin, W.flags, W.r_target, W.host, W.JpJdF,
                                                                       W.xAd, W.pt_out, W.cap_p, lambda);
    c->launches++;
    LDSO_CUDA_OK(c, cudaGetLastError());
    return 0;
}
