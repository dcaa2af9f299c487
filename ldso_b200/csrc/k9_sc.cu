// k9_sc.cu — K9: AccumulatedSCHessianSSE::addPoint (reference
// src/internal/OptimizationBackend/AccumulatedSCHessian.cc:9-51) and K10:
// EnergyFunctional::resubstituteFPt (src/internal/OptimizationBackend/EnergyFunctional.cc:638-667).
//
// K9 runs as two launches.
//  (a) k9_point_kernel, one thread per point: folds the per-residual Hdd/bd/Hcd pieces written by K8 in residual
//      order (bit-identical to the reference's sequential +=), derives HdiF / bdSumF / idepth_hessian and writes the
//      dense lookup ridx[point][target] = residual id (or -1) that (b) uses.
//  (b) k9_sc_kernel: all residuals of a point share its host frame and points are stored contiguously per host
//      (allPoints order), so a CTA takes up to 64 points of ONE host, stages them
//      {JpJdF per target, HdiF, bdSumF, Hcd} in shared memory and keeps its outputs in registers: a work item is
//      one row (t1, i) x one target t2 = 8 entries of accD[h][t1][t2] (n*n*8 items), plus n*8 items for
//      accE/accEB and 20 for accHcc/accbc.  No shared-memory atomics; one fp64 atomicAdd per output at the end.
#include "ctx.h"

namespace {

constexpr int K9_THREADS = 512;
constexpr int K9_TILE = 64;
constexpr int K9_MAXITEMS = 3;     // n = 12: 12*12*8 + 12*8 + 20 = 1268 items <= 3 * 512

__global__ void k9_point_kernel(int n, int np, const int* __restrict__ res_begin, const int* __restrict__ flags,
                                const int* __restrict__ r_target, const float* __restrict__ res_pt,
                                const float* __restrict__ prior, const float* __restrict__ idepth,
                                const float* __restrict__ idepth_zero, float* __restrict__ pt_out,
                                float* __restrict__ max_rel_baseline, int cap_p, int* __restrict__ ridx,
                                const unsigned char* __restrict__ marg, float prior_fac) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= np) return;
    if (marg && !marg[p]) return;   // marginalizePointsF touches the marginalised points only
    const int rb = res_begin[p], re = res_begin[p + 1];
    float HddA = 0, bdA = 0, HddL = 0, bdL = 0, HcdA[4] = {0, 0, 0, 0}, HcdL[4] = {0, 0, 0, 0};
    int ngood = 0;
    for (int t = 0; t < n; t++) ridx[(size_t)p * LDSO_MAX_WINDOW + t] = -1;
    for (int r = rb; r < re; r++) {
        const int fl = flags[r];
        if (!(fl & LDSO_RF_ACTIVE)) continue;
        ngood++;
        ridx[(size_t)p * LDSO_MAX_WINDOW + r_target[r]] = r;
        const float4 a = __ldg(reinterpret_cast<const float4*>(res_pt + 8 * (size_t)r));
        const float4 b = __ldg(reinterpret_cast<const float4*>(res_pt + 8 * (size_t)r) + 1);
        if (b.z == 0.f) continue;   // not seen by K8 (cannot happen for an active residual)
        if (marg || (fl & LDSO_RF_LINEARIZED)) {   // addPoint<2> books everything under the L fields and zeroes the A fields
            HddL += a.x; bdL += a.y; HcdL[0] += a.z; HcdL[1] += a.w; HcdL[2] += b.x; HcdL[3] += b.y;
        } else {
            HddA += a.x; bdA += a.y; HcdA[0] += a.z; HcdA[1] += a.w; HcdA[2] += b.x; HcdA[3] += b.y;
        }
    }
    float HdiF = 0.f, bdSum = 0.f, idh = 0.f;
    if (ngood > 0) {
        const float pr = marg ? prior[p] * prior_fac : prior[p];   // p->priorF *= setting_idepthFixPriorMargFac
        float H = HddA + HddL + pr;
        if (H < 1e-10) H = 1e-10;
        idh = H;
        HdiF = (float)(1.0 / (double)H);
        bdSum = bdA + bdL;
        if (!marg) bdSum += pr * (idepth[p] - idepth_zero[p]);   // shiftPriorToZero
    }
    pt_out[PT_HDD_A * cap_p + p] = HddA; pt_out[PT_BD_A * cap_p + p] = bdA;
    pt_out[PT_HDD_L * cap_p + p] = HddL; pt_out[PT_BD_L * cap_p + p] = bdL;
    for (int i = 0; i < 4; i++) {
        pt_out[(PT_HCD_A + i) * cap_p + p] = HcdA[i];
        pt_out[(PT_HCD_L + i) * cap_p + p] = HcdL[i];
    }
    pt_out[PT_HDI * cap_p + p] = HdiF; pt_out[PT_BDSUM * cap_p + p] = bdSum; pt_out[PT_IDH * cap_p + p] = idh;
    if (ngood == 0) max_rel_baseline[p] = 0.f;
}

__global__ void __launch_bounds__(K9_THREADS)
k9_sc_kernel(int n, const int* __restrict__ blocks, const int* __restrict__ ridx, const float* __restrict__ JpJdF,
             const float* __restrict__ pt_out, int cap_p, double* __restrict__ acc_sc,
             const unsigned char* __restrict__ marg) {
    extern __shared__ float sm[];
    float* sJ = sm;                               // [K9_TILE][n][8]
    float* sHdi = sJ + K9_TILE * n * 8;           // [K9_TILE]
    float* sBd = sHdi + K9_TILE;                  // [K9_TILE]
    float* sHcd = sBd + K9_TILE;                  // [K9_TILE][4]
    const int nn = n * n;
    const int h = blocks[3 * blockIdx.x], p0 = blocks[3 * blockIdx.x + 1], p1 = blocks[3 * blockIdx.x + 2];
    const int nD = nn * 8, nE = n * 8, nItems = nD + nE + 20;
    float acc[K9_MAXITEMS][8];
#pragma unroll
    for (int q = 0; q < K9_MAXITEMS; q++)
#pragma unroll
        for (int j = 0; j < 8; j++) acc[q][j] = 0.f;
    for (int base = p0; base < p1; base += K9_TILE) {
        const int cnt = min(K9_TILE, p1 - base);
        __syncthreads();
        for (int i = threadIdx.x; i < K9_TILE * n * 2; i += K9_THREADS) {   // one float4 (half a JpJdF row) per thread
            const int half = i & 1, t = (i >> 1) % n, q = (i >> 1) / n;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (q < cnt && (!marg || marg[base + q])) {
                const int r = ridx[(size_t)(base + q) * LDSO_MAX_WINDOW + t];
                if (r >= 0) v = __ldg(reinterpret_cast<const float4*>(JpJdF + 8 * (size_t)r) + half);
            }
            reinterpret_cast<float4*>(sJ)[i] = v;
        }
        for (int q = threadIdx.x; q < K9_TILE; q += K9_THREADS) {
            const int p = base + q;
            const bool in = q < cnt && (!marg || marg[p]);
            sHdi[q] = in ? pt_out[PT_HDI * cap_p + p] : 0.f;
            sBd[q] = in ? pt_out[PT_BDSUM * cap_p + p] : 0.f;
#pragma unroll
            for (int k = 0; k < 4; k++)
                sHcd[4 * q + k] = in ? pt_out[(PT_HCD_A + k) * cap_p + p] + pt_out[(PT_HCD_L + k) * cap_p + p] : 0.f;
        }
        __syncthreads();
#pragma unroll
        for (int q = 0; q < K9_MAXITEMS; q++) {
            const int it = threadIdx.x + q * K9_THREADS;
            if (it >= nItems) break;
            if (it < nD) {                        // accD[h][t1][t2] row i: (HdiF * J1_i) * J2_j
                const int t2 = it % n, ti = it / n, t1 = ti >> 3, i = ti & 7;
                for (int pp = 0; pp < cnt; pp++) {
                    const float a = sHdi[pp] * sJ[(pp * n + t1) * 8 + i];
                    const float4 b0 = reinterpret_cast<const float4*>(sJ + (pp * n + t2) * 8)[0];
                    const float4 b1 = reinterpret_cast<const float4*>(sJ + (pp * n + t2) * 8)[1];
                    acc[q][0] += a * b0.x; acc[q][1] += a * b0.y; acc[q][2] += a * b0.z; acc[q][3] += a * b0.w;
                    acc[q][4] += a * b1.x; acc[q][5] += a * b1.y; acc[q][6] += a * b1.z; acc[q][7] += a * b1.w;
                }
            } else if (it < nD + nE) {            // accE[h][t] row i (4 entries) and accEB[h][t][i]
                const int e = it - nD, t = e >> 3, i = e & 7;
                for (int pp = 0; pp < cnt; pp++) {
                    const float a = sHdi[pp] * sJ[(pp * n + t) * 8 + i];
                    acc[q][0] += a * sHcd[4 * pp]; acc[q][1] += a * sHcd[4 * pp + 1];
                    acc[q][2] += a * sHcd[4 * pp + 2]; acc[q][3] += a * sHcd[4 * pp + 3];
                    acc[q][4] += (sHdi[pp] * sBd[pp]) * sJ[(pp * n + t) * 8 + i];
                }
            } else {                              // accHcc[4][4], accbc[4]
                const int e = it - nD - nE;
                for (int pp = 0; pp < cnt; pp++) {
                    if (e < 16) acc[q][0] += (sHdi[pp] * sHcd[4 * pp + (e >> 2)]) * sHcd[4 * pp + (e & 3)];
                    else acc[q][0] += (sBd[pp] * sHdi[pp]) * sHcd[4 * pp + (e - 16)];
                }
            }
        }
    }
    // flush to the fp64 accumulators: accD[(h + n t1 + n^2 t2)][64] | accE[(h + n t)][32] | accEB[(h + n t)][8] | Hcc | bc
    const size_t offE = (size_t)nn * n * 64, offEB = offE + (size_t)nn * 32, offH = offEB + (size_t)nn * 8;
#pragma unroll
    for (int q = 0; q < K9_MAXITEMS; q++) {
        const int it = threadIdx.x + q * K9_THREADS;
        if (it >= nItems) break;
        if (it < nD) {
            const int t2 = it % n, ti = it / n, t1 = ti >> 3, i = ti & 7;
            double* g = acc_sc + ((size_t)(h + n * t1 + nn * t2)) * 64 + i * 8;
#pragma unroll
            for (int j = 0; j < 8; j++) if (acc[q][j] != 0.f) atomicAdd(g + j, (double)acc[q][j]);
        } else if (it < nD + nE) {
            const int e = it - nD, t = e >> 3, i = e & 7;
            double* g = acc_sc + offE + (size_t)(h + n * t) * 32 + i * 4;
#pragma unroll
            for (int j = 0; j < 4; j++) if (acc[q][j] != 0.f) atomicAdd(g + j, (double)acc[q][j]);
            if (acc[q][4] != 0.f) atomicAdd(acc_sc + offEB + (size_t)(h + n * t) * 8 + i, (double)acc[q][4]);
        } else {
            if (acc[q][0] != 0.f) atomicAdd(acc_sc + offH + (it - nD - nE), (double)acc[q][0]);
        }
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

int k9_accumulate_sc(ldso_gpu_ctx* c, const unsigned char* marg_mask, float prior_fac) {
    WindowDev& W = c->win;
    LDSO_CUDA_OK(c, cudaMemsetAsync(W.acc_sc, 0, sizeof(double) * W.acc_sc_size, c->stream));
    // CTA table: up to 64 consecutive points of one host per CTA (built by window_set_points)
    const int nblk = W.n_sc_blocks;
    if (nblk > 0) {
        const int n = W.n;
        LDSO_REQUIRE(c, n * n * 8 + n * 8 + 20 <= K9_MAXITEMS * K9_THREADS);
        const size_t smem = sizeof(float) * (size_t)(K9_TILE * n * 8 + K9_TILE * 6);
        prof_begin(c, c->stream);
        k9_point_kernel<<<(W.np + 127) / 128, 128, 0, c->stream>>>(n, W.np, W.res_begin, W.flags, W.r_target, W.res_pt, W.prior,
                                                                   W.idepth, W.idepth_zero, W.pt_out, W.max_rel_baseline, W.cap_p,
                                                                   W.ridx, marg_mask, prior_fac);
        k9_sc_kernel<<<nblk, K9_THREADS, smem, c->stream>>>(n, W.sc_blocks, W.ridx, W.JpJdF, W.pt_out, W.cap_p, W.acc_sc, marg_mask);
        prof_end(c, c->stream, LDSO_PROF_K9, W.nr);
        c->launches += 2;
    }
    LDSO_CUDA_OK(c, cudaGetLastError());
    return 0;
}

int k10_resubstitute(ldso_gpu_ctx* c, float lambda) {
    WindowDev& W = c->win;
    if (W.np == 0) return 0;
    prof_begin(c, c->stream);
    k10_resubstitute_kernel<<<(W.np + 127) / 128, 128, 0, c->stream>>>(W.n, W.np, W.res_begin, W.flags, W.r_target, W.host, W.JpJdF,
                                                                       W.xAd, W.pt_out, W.cap_p, lambda);
    prof_end(c, c->stream, LDSO_PROF_K10, W.np);
    c->launches++;
    LDSO_CUDA_OK(c, cudaGetLastError());
    return 0;
}
