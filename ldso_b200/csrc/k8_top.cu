// k8_top.cu — K8: AccumulatedTopHessianSSE::addPoint<0>, <1> and <2>
// (reference src/internal/OptimizationBackend/AccumulatedTopHessian.cc:9-118) with AccumulatorApprox
// (MatrixAccumulators.h:749-1102).
//
// The reference walks points -> residuals and scatters each residual's 91-float Gauss-Newton block into
// acc[host + n*target].  Here the host pre-sorts residual ids by accumulator key (mode, host, target)
// and cuts the sorted list into chunks of <= 128 residuals of one key; one CTA reduces one chunk:
// a thread forms its residual's 91 products in registers (same expression order as the reference),
// warp shuffles + shared memory reduce them in fp32, and 91 threads add the CTA sum into the fp64
// accumulator of the key.  Active (mode 0) and linearised (mode 1) residuals are disjoint sets, so a
// single launch serves accumulateAF_MT and accumulateLF_MT.  Per-residual Hdd/bd/Hcd contributions are
// written out for the per-point sums of K9.  Reads 288 B (+32 B res_toZeroF in mode 1) per residual.
#include "ctx.h"

namespace {

constexpr int K8_THREADS = 128;
constexpr int NV = 91;

__global__ void __launch_bounds__(K8_THREADS)
k8_top_kernel(int n, const int* __restrict__ chunk, const int* __restrict__ order, const int* __restrict__ flags,
              const int* __restrict__ r_point, const int* __restrict__ r_target, const int* __restrict__ host,
              const float* __restrict__ J, const float* __restrict__ rtz, const float* __restrict__ adHTdelta, float4 cDelta,
              const float* __restrict__ idepth, const float* __restrict__ idepth_zero, float* __restrict__ res_pt,
              double* __restrict__ acc_top, int* __restrict__ acc_cnt, int marg) {
    const int key = chunk[3 * blockIdx.x], start = chunk[3 * blockIdx.x + 1], count = chunk[3 * blockIdx.x + 2];
    const int nn = n * n;
    const int mode = marg ? 2 : key >= nn ? 1 : 0;   // marg: addPoint<2>, the list holds only the marginalised points' residuals
    float acc[NV];
#pragma unroll
    for (int i = 0; i < NV; i++) acc[i] = 0.f;
    int used = 0;
    if (threadIdx.x < count) {
        const int r = order[start + threadIdx.x];
        const int fl = flags[r];
        float* rp = res_pt + 8 * (size_t)r;
        if (fl & LDSO_RF_ACTIVE) {
            used = 1;
            const float4* j4 = reinterpret_cast<const float4*>(J + (size_t)r * LDSO_J_FLOATS);
            float j[LDSO_J_FLOATS];
#pragma unroll
            for (int i = 0; i < LDSO_J_FLOATS / 4; i++) {
                float4 t = __ldg(j4 + i);
                j[4 * i] = t.x; j[4 * i + 1] = t.y; j[4 * i + 2] = t.z; j[4 * i + 3] = t.w;
            }
            float res[8];
            if (mode == 0) {
#pragma unroll
                for (int i = 0; i < 8; i++) res[i] = j[i];
            } else if (mode == 2) {
#pragma unroll
                for (int i = 0; i < 8; i++) res[i] = rtz[8 * (size_t)r + i];
            } else {
                // resApprox = res_toZeroF + J * delta (AccumulatedTopHessian.cc:47-64)
                const int pt = r_point[r];
                const float* dp = adHTdelta + 8 * (host[pt] + n * r_target[r]);
                const float dd = idepth[pt] - idepth_zero[pt];
                float dx = 0, dy = 0, cx = 0, cy = 0;
#pragma unroll
                for (int i = 0; i < 6; i++) { dx += j[40 + i] * dp[i]; dy += j[46 + i] * dp[i]; }
                const float cd[4] = {cDelta.x, cDelta.y, cDelta.z, cDelta.w};
#pragma unroll
                for (int i = 0; i < 4; i++) { cx += j[52 + i] * cd[i]; cy += j[56 + i] * cd[i]; }
                const float Jx = dx + cx + j[60] * dd, Jy = dy + cy + j[61] * dd;
                const float da = dp[6], db = dp[7];
#pragma unroll
                for (int i = 0; i < 8; i++) {
                    float v = rtz[8 * (size_t)r + i];
                    v = v + j[8 + i] * Jx;
                    v = v + j[16 + i] * Jy;
                    v = v + j[24 + i] * da;
                    v = v + j[32 + i] * db;
                    res[i] = v;
                }
            }
            float JI_r0 = 0, JI_r1 = 0, Jab_r0 = 0, Jab_r1 = 0, rr = 0;
#pragma unroll
            for (int i = 0; i < 8; i++) {
                JI_r0 += res[i] * j[8 + i];
                JI_r1 += res[i] * j[16 + i];
                Jab_r0 += res[i] * j[24 + i];
                Jab_r1 += res[i] * j[32 + i];
                rr += res[i] * res[i];
            }
            // x = [Jpdc[0] | Jpdxi[0]], y = [Jpdc[1] | Jpdxi[1]]
            float x[10], y[10];
#pragma unroll
            for (int i = 0; i < 4; i++) { x[i] = j[52 + i]; y[i] = j[56 + i]; }
#pragma unroll
            for (int i = 0; i < 6; i++) { x[4 + i] = j[40 + i]; y[4 + i] = j[46 + i]; }
            const float a = j[62], b = j[63], cc = j[64];
            int idx = 0;
#pragma unroll
            for (int rr_ = 0; rr_ < 10; rr_++)
#pragma unroll
                for (int c_ = rr_; c_ < 10; c_++)
                    acc[idx++] = a * x[c_] * x[rr_] + cc * y[c_] * y[rr_] + b * (x[c_] * y[rr_] + y[c_] * x[rr_]);
#pragma unroll
            for (int rr_ = 0; rr_ < 10; rr_++) {
                acc[55 + 3 * rr_ + 0] = x[rr_] * j[65] + y[rr_] * j[66];
                acc[55 + 3 * rr_ + 1] = x[rr_] * j[67] + y[rr_] * j[68];
                acc[55 + 3 * rr_ + 2] = x[rr_] * JI_r0 + y[rr_] * JI_r1;
            }
            acc[85] = j[69]; acc[86] = j[70]; acc[87] = Jab_r0; acc[88] = j[71]; acc[89] = Jab_r1; acc[90] = rr;
            // per-point pieces (AccumulatedTopHessian.cc:92-96)
            const float t0 = j[62] * j[60] + j[63] * j[61], t1 = j[63] * j[60] + j[64] * j[61];
            rp[0] = t0 * j[60] + t1 * j[61];               // Hdd
            rp[1] = JI_r0 * j[60] + JI_r1 * j[61];         // bd
#pragma unroll
            for (int i = 0; i < 4; i++) rp[2 + i] = j[52 + i] * t0 + j[56 + i] * t1;   // Hcd
            rp[6] = 1.f;
        } else {
            rp[6] = 0.f;
        }
    }
    // CTA reduction
    __shared__ float s[K8_THREADS / 32][WIN_TOP_STRIDE];
    __shared__ int s_used[K8_THREADS / 32];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
    for (int i = 0; i < NV; i++) {
        float v = acc[i];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0) s[wid][i] = v;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) used += __shfl_xor_sync(0xffffffffu, used, o);
    if (lane == 0) s_used[wid] = used;
    __syncthreads();
    if (threadIdx.x < NV) {
        float v = 0.f;
#pragma unroll
        for (int wq = 0; wq < K8_THREADS / 32; wq++) v += s[wq][threadIdx.x];
        if (v != 0.f) atomicAdd(acc_top + (size_t)key * WIN_TOP_STRIDE + threadIdx.x, (double)v);
    } else if (threadIdx.x == NV) {
        int u = 0;
        for (int wq = 0; wq < K8_THREADS / 32; wq++) u += s_used[wq];
        if (u) atomicAdd(acc_cnt + key, u);
    }
}

// EnergyFunctional::calcLEnergyPt (EnergyFunctional.cc:751-806): one thread per point, per-point fp32 sum in the
// reference's order, warp fold, one fp64 atomic per warp.
__global__ void __launch_bounds__(128)
k8_lenergy_kernel(int n, int np, const int* __restrict__ res_begin, const int* __restrict__ flags, const int* __restrict__ r_target,
                  const int* __restrict__ host, const float* __restrict__ J, const float* __restrict__ rtz,
                  const float* __restrict__ adHTdelta, float4 cDelta, const float* __restrict__ idepth,
                  const float* __restrict__ idepth_zero, const float* __restrict__ prior, double* __restrict__ out) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    float E = 0.f;
    if (p < np) {
        const float dd = idepth[p] - idepth_zero[p];
        const float cd[4] = {cDelta.x, cDelta.y, cDelta.z, cDelta.w};
        const int h = host[p];
        for (int r = res_begin[p]; r < res_begin[p + 1]; r++) {
            const int fl = flags[r];
            if (!(fl & LDSO_RF_LINEARIZED) || !(fl & LDSO_RF_ACTIVE)) continue;
            const float* j = J + (size_t)r * LDSO_J_FLOATS;
            const float* dp = adHTdelta + 8 * (h + n * r_target[r]);
            float dx = 0, dy = 0, cx = 0, cy = 0;
#pragma unroll
            for (int i = 0; i < 6; i++) { dx += j[40 + i] * dp[i]; dy += j[46 + i] * dp[i]; }
#pragma unroll
            for (int i = 0; i < 4; i++) { cx += j[52 + i] * cd[i]; cy += j[56 + i] * cd[i]; }
            const float Jx = dx + cx + j[60] * dd, Jy = dy + cy + j[61] * dd;
            const float da = dp[6], db = dp[7];
#pragma unroll
            for (int i = 0; i < 8; i++) {
                float Jd = j[8 + i] * Jx;
                Jd = Jd + j[16 + i] * Jy;
                Jd = Jd + j[24 + i] * da;
                Jd = Jd + j[32 + i] * db;
                float r0 = rtz[8 * (size_t)r + i];
                r0 = r0 + r0;
                r0 = r0 + Jd;
                E += Jd * r0;
            }
        }
        E += dd * dd * prior[p];
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) E += __shfl_xor_sync(0xffffffffu, E, o);
    if ((threadIdx.x & 31) == 0 && E != 0.f) atomicAdd(out, (double)E);
}

}  // namespace

int k8_accumulate_top(ldso_gpu_ctx* c, int marg) {
    WindowDev& W = c->win;
    const int nn = W.n * W.n;
    LDSO_CUDA_OK(c, cudaMemsetAsync(W.acc_top, 0, sizeof(double) * 2 * nn * WIN_TOP_STRIDE, c->stream));
    LDSO_CUDA_OK(c, cudaMemsetAsync(W.acc_top_cnt, 0, sizeof(int) * 2 * nn, c->stream));
    if (W.n_chunks > 0) {
        float4 cd = make_float4(W.cDeltaF[0], W.cDeltaF[1], W.cDeltaF[2], W.cDeltaF[3]);
        prof_begin(c, c->stream);
        k8_top_kernel<<<W.n_chunks, K8_THREADS, 0, c->stream>>>(W.n, W.chunk, W.order, W.flags, W.r_point, W.r_target, W.host, W.J,
                                                                W.rtz, W.adHTdelta, cd, W.idepth, W.idepth_zero, W.res_pt,
                                                                W.acc_top, W.acc_top_cnt, marg);
        prof_end(c, c->stream, LDSO_PROF_K8, W.nr);
        c->launches++;
    }
    LDSO_CUDA_OK(c, cudaGetLastError());
    return 0;
}

int k8_linearized_energy(ldso_gpu_ctx* c, double* e_dev) {
    WindowDev& W = c->win;
    LDSO_CUDA_OK(c, cudaMemsetAsync(e_dev, 0, sizeof(double), c->stream));
    if (W.np > 0) {
        float4 cd = make_float4(W.cDeltaF[0], W.cDeltaF[1], W.cDeltaF[2], W.cDeltaF[3]);
        k8_lenergy_kernel<<<(W.np + 127) / 128, 128, 0, c->stream>>>(W.n, W.np, W.res_begin, W.flags, W.r_target, W.host, W.J, W.rtz,
                                                                     W.adHTdelta, cd, W.idepth, W.idepth_zero, W.prior, e_dev);
        c->launches++;
    }
    LDSO_CUDA_OK(c, cudaGetLastError());
    return 0;
}
