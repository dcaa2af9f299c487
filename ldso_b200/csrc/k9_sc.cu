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
//      {JpJdF per target, HdiF, bdSumF, Hcd} in shared memory and keeps its outputs in registers: a thread owns a
//      4 x 8 register tile of one accD[h][t1][t2] block (2 n^2 tiles) over one of G point groups — 3 LDS.128 + 1 LDS
//      per 32 FMA; the first version's 1 x 8 rows were bound by their shared-memory reads — and the threads behind
//      them one accE/accEB row or accHcc/accbc entry each.  The G groups meet in shared memory at the end; no
//      shared-memory atomics, one fp64 atomicAdd per output per CTA.
#include "ctx.h"

namespace {

constexpr int K9_THREADS = 512;
constexpr int K9_TILE = 64;

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

// thread layout of k9_sc_kernel for a window of n frames: nD2 = 2 n^2 Schur items (one (t1, t2) pair x 4 of its 8 rows, all
// 8 columns: 32 sums in registers), replicated over G point groups; the threads behind them own one accE/accEB row or
// one accHcc/accbc entry each (n*8 + 20 <= 116 of them; G leaves at least 32 + ... threads, checked on the host)
__host__ __device__ inline int k9_groups(int n) { return max(1, min(8, (K9_THREADS - 32) / (2 * n * n))); }

__global__ void __launch_bounds__(K9_THREADS, 2)
k9_sc_kernel(int n, const int* __restrict__ blocks, const int* __restrict__ ridx, const float* __restrict__ JpJdF,
             const float* __restrict__ pt_out, int cap_p, double* __restrict__ acc_sc,
             const unsigned char* __restrict__ marg) {
    extern __shared__ __align__(16) float sm[];
    // two staging buffers {sJ [K9_TILE][n][8], sHdi [K9_TILE], sBd [K9_TILE], sHcd [K9_TILE][4]}: while a tile is being summed,
    // the JpJdF rows of the next one arrive by cp.async and its per-point scalars sit in registers (after the last tile the
    // shared memory is the cross-group fold scratch)
    const int bufF = K9_TILE * n * 8 + K9_TILE * 6;
    const int nn = n * n;
    const int h = blocks[3 * blockIdx.x], p0 = blocks[3 * blockIdx.x + 1], p1 = blocks[3 * blockIdx.x + 2];
    const int nD2 = nn * 2, G = k9_groups(n), nE = n * 8;
    const int tid = threadIdx.x;
    const bool isD = tid < nD2 * G;
    // Schur item of a D thread
    const int item = tid % nD2, g = tid / nD2;
    const int t2 = item % n, t1 = (item / n) >> 1, i0 = ((item / n) & 1) * 4;
    // accE / accEB / accHcc / accbc item of an E thread (e >= nE + 20: idle)
    const int e = tid - nD2 * G;
    float acc[4][8];
#pragma unroll
    for (int r = 0; r < 4; r++)
#pragma unroll
        for (int j = 0; j < 8; j++) acc[r][j] = 0.f;
    constexpr int PIECES = (K9_TILE * LDSO_MAX_WINDOW * 2 + K9_THREADS - 1) / K9_THREADS;   // float4 pieces of sJ per thread
    const int npieces = K9_TILE * n * 2;
    // residual id behind each of this thread's pieces of a tile (-1: nothing to stage, zero fill)
    auto load_rid = [&](int base, int (&rid)[PIECES]) {
        const int cnt = min(K9_TILE, p1 - base);
#pragma unroll
        for (int j = 0; j < PIECES; j++) {
            const int i = tid + j * K9_THREADS;
            rid[j] = -1;
            if (i < npieces) {
                const int t = (i >> 1) % n, q = (i >> 1) / n;
                if (q < cnt && (!marg || marg[base + q])) rid[j] = ridx[(size_t)(base + q) * LDSO_MAX_WINDOW + t];
            }
        }
    };
    auto issue = [&](int b, const int (&rid)[PIECES]) {
        float4* dst = reinterpret_cast<float4*>(sm + b * bufF);
#pragma unroll
        for (int j = 0; j < PIECES; j++) {
            const int i = tid + j * K9_THREADS;
            if (i >= npieces) continue;
            if (rid[j] >= 0) {
                const unsigned sa = (unsigned)__cvta_generic_to_shared(dst + i);
                const float4* src = reinterpret_cast<const float4*>(JpJdF + 8 * (size_t)rid[j]) + (i & 1);
                asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(src) : "memory");
            } else {
                dst[i] = make_float4(0.f, 0.f, 0.f, 0.f);
            }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    // the per-point scalars of a tile, one point per thread (tid < K9_TILE): HdiF, bdSumF, Hcd[4]
    auto load_scalars = [&](int base, float (&sc)[6]) {
        const int cnt = min(K9_TILE, p1 - base), p = base + tid;
        const bool in = tid < cnt && (!marg || marg[p]);
        sc[0] = in ? pt_out[PT_HDI * cap_p + p] : 0.f;
        sc[1] = in ? pt_out[PT_BDSUM * cap_p + p] : 0.f;
#pragma unroll
        for (int k = 0; k < 4; k++) sc[2 + k] = in ? pt_out[(PT_HCD_A + k) * cap_p + p] + pt_out[(PT_HCD_L + k) * cap_p + p] : 0.f;
    };
    auto store_scalars = [&](int b, const float (&sc)[6]) {
        float* s0 = sm + b * bufF + K9_TILE * n * 8;
        s0[tid] = sc[0]; s0[K9_TILE + tid] = sc[1];
#pragma unroll
        for (int k = 0; k < 4; k++) s0[2 * K9_TILE + 4 * tid + k] = sc[2 + k];
    };
    int rid[PIECES];
    float sc[6];
    load_rid(p0, rid);
    issue(0, rid);
    if (tid < K9_TILE) { load_scalars(p0, sc); store_scalars(0, sc); }
    if (p0 + K9_TILE < p1) load_rid(p0 + K9_TILE, rid);
    int b = 0;
    for (int base = p0; base < p1; base += K9_TILE, b ^= 1) {
        const int cnt = min(K9_TILE, p1 - base);
        const bool more = base + K9_TILE < p1;
        if (more) {
            issue(b ^ 1, rid);                                   // buffer b ^ 1 was last read two tiles ago (barrier below)
            if (tid < K9_TILE) load_scalars(base + K9_TILE, sc);
            if (base + 2 * K9_TILE < p1) load_rid(base + 2 * K9_TILE, rid);
            asm volatile("cp.async.wait_group 1;" ::: "memory");
        } else {
            asm volatile("cp.async.wait_group 0;" ::: "memory");
        }
        __syncthreads();
        const float* sJ = sm + b * bufF;
        const float* sHdi = sJ + K9_TILE * n * 8;
        const float* sBd = sHdi + K9_TILE;
        const float* sHcd = sBd + K9_TILE;
        // fp32 sums over the tile's points with fused multiply-adds (the oracle's mul/add sums differ in order anyway: 1e-4)
        if (isD) {                                // accD[h][t1][t2] rows i0..i0+3: (HdiF * J1_i) * J2_j; 3 LDS.128 + 1 LDS per 32 FMA
            const float* ja = sJ + t1 * 8 + i0;
            const float* jb = sJ + t2 * 8;
            for (int pp = g; pp < cnt; pp += G) {
                const float hdi = sHdi[pp];
                const float4 a4 = *reinterpret_cast<const float4*>(ja + pp * n * 8);
                const float4 b0 = reinterpret_cast<const float4*>(jb + pp * n * 8)[0];
                const float4 b1 = reinterpret_cast<const float4*>(jb + pp * n * 8)[1];
                const float a[4] = {hdi * a4.x, hdi * a4.y, hdi * a4.z, hdi * a4.w};
#pragma unroll
                for (int r = 0; r < 4; r++) {
                    acc[r][0] = __fmaf_rn(a[r], b0.x, acc[r][0]); acc[r][1] = __fmaf_rn(a[r], b0.y, acc[r][1]);
                    acc[r][2] = __fmaf_rn(a[r], b0.z, acc[r][2]); acc[r][3] = __fmaf_rn(a[r], b0.w, acc[r][3]);
                    acc[r][4] = __fmaf_rn(a[r], b1.x, acc[r][4]); acc[r][5] = __fmaf_rn(a[r], b1.y, acc[r][5]);
                    acc[r][6] = __fmaf_rn(a[r], b1.z, acc[r][6]); acc[r][7] = __fmaf_rn(a[r], b1.w, acc[r][7]);
                }
            }
        } else if (e < nE) {                      // accE[h][t] row i (4 entries) and accEB[h][t][i]
            const int t = e >> 3, i = e & 7;
            for (int pp = 0; pp < cnt; pp++) {
                const float jv = sJ[(pp * n + t) * 8 + i];
                const float av = sHdi[pp] * jv;
                const float4 hc = reinterpret_cast<const float4*>(sHcd)[pp];
                acc[0][0] = __fmaf_rn(av, hc.x, acc[0][0]); acc[0][1] = __fmaf_rn(av, hc.y, acc[0][1]);
                acc[0][2] = __fmaf_rn(av, hc.z, acc[0][2]); acc[0][3] = __fmaf_rn(av, hc.w, acc[0][3]);
                acc[0][4] = __fmaf_rn(sHdi[pp] * sBd[pp], jv, acc[0][4]);
            }
        } else if (e < nE + 20) {                 // accHcc[4][4], accbc[4]
            const int k = e - nE;
            for (int pp = 0; pp < cnt; pp++) {
                if (k < 16) acc[0][0] = __fmaf_rn(sHdi[pp] * sHcd[4 * pp + (k >> 2)], sHcd[4 * pp + (k & 3)], acc[0][0]);
                else acc[0][0] = __fmaf_rn(sBd[pp] * sHdi[pp], sHcd[4 * pp + (k - 16)], acc[0][0]);
            }
        }
        if (more && tid < K9_TILE) store_scalars(b ^ 1, sc);     // read after the next iteration's barrier
        __syncthreads();                                        // buffer b is free for the tile after the next
    }
    // flush to the fp64 accumulators: accD[(h + n t1 + n^2 t2)][64] | accE[(h + n t)][32] | accEB[(h + n t)][8] | Hcc | bc
    const size_t offE = (size_t)nn * n * 64, offEB = offE + (size_t)nn * 32, offH = offEB + (size_t)nn * 8;
    if (!isD) {
        if (e < nE) {
            const int t = e >> 3, i = e & 7;
            double* gE = acc_sc + offE + (size_t)(h + n * t) * 32 + i * 4;
#pragma unroll
            for (int j = 0; j < 4; j++) if (acc[0][j] != 0.f) atomicAdd(gE + j, (double)acc[0][j]);
            if (acc[0][4] != 0.f) atomicAdd(acc_sc + offEB + (size_t)(h + n * t) * 8 + i, (double)acc[0][4]);
        } else if (e < nE + 20) {
            if (acc[0][0] != 0.f) atomicAdd(acc_sc + offH + (e - nE), (double)acc[0][0]);
        }
    }
    // the G point groups of a Schur item meet in shared memory ([group][k = row*8 + col][item]: conflict-free both ways),
    // then every thread folds and adds a strided share of the nD2 x 32 sums: one fp64 atomic per output per CTA
    __syncthreads();
    float* fold = sm;
    if (isD) {
#pragma unroll
        for (int r = 0; r < 4; r++)
#pragma unroll
            for (int j = 0; j < 8; j++) fold[((g * 32) + r * 8 + j) * nD2 + item] = acc[r][j];
    }
    __syncthreads();
    for (int o = tid; o < nD2 * 32; o += K9_THREADS) {
        const int k = o / nD2, it = o - k * nD2;
        float v = 0.f;
        for (int gg = 0; gg < G; gg++) v += fold[(gg * 32 + k) * nD2 + it];
        if (v != 0.f) {
            const int c2 = it % n, c1 = (it / n) >> 1, row = ((it / n) & 1) * 4 + (k >> 3), col = k & 7;
            atomicAdd(acc_sc + ((size_t)(h + n * c1 + nn * c2)) * 64 + row * 8 + col, (double)v);
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
        const int G = k9_groups(n);
        LDSO_REQUIRE(c, 2 * n * n * G + n * 8 + 20 <= K9_THREADS);
        const size_t smem = sizeof(float) * std::max((size_t)2 * (K9_TILE * n * 8 + K9_TILE * 6), (size_t)G * 32 * 2 * n * n);
        {
            const int rc = ldso_raise_smem(c, (const void*)k9_sc_kernel, smem, 0);
            if (rc) return rc;
        }
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
