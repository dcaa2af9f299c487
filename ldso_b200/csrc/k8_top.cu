// k8_top.cu — K8: AccumulatedTopHessianSSE::addPoint<0>, <1> and <2>
// (reference src/internal/OptimizationBackend/AccumulatedTopHessian.cc:9-118) with AccumulatorApprox
// (MatrixAccumulators.h:749-1102).
//
// The reference walks points -> residuals and scatters each residual's 91-float Gauss-Newton block into
// acc[host + n*target].  Here the host pre-sorts residual ids by accumulator key (mode, host, target) and cuts the
// sorted list into chunks of one key; one CTA reduces one chunk: a thread forms its residuals' 91 products in
// registers, warp shuffles + shared memory reduce them in fp32, and 91 threads add the CTA sum into the fp64
// accumulator of the key.  Active (mode 0) and linearised (mode 1) residuals are disjoint sets, so a single launch
// serves accumulateAF_MT and accumulateLF_MT.  Per-residual Hdd/bd/Hcd contributions are written out for the
// per-point sums of K9.  Reads 288 B (+32 B res_toZeroF in mode 1) and writes 32 B per residual.
//
// Two launch shapes, chosen by build_order (window_api.cu) from the size of the work list:
//   * chunks of <= 128 residuals, thread per residual, records loaded straight to registers (one window: ~100 CTAs,
//     latency matters, nothing to pipeline);
//   * chunks of up to 1024 residuals (batched windows): each warp streams 32-record tiles of its chunk through two
//     private shared-memory buffers with cp.async (18 x 16 B per record, contiguous -> few L1 wavefronts, the next
//     tile in flight while the current one is multiplied), lane l then reads record l with conflict-free LDS.128
//     (row stride 84 floats).  profiles/k8_lab holds the design study and its timings.
//
// The 55 Hessian entries use p = a x + b y, q = b x + c y, entry = x_c p_r + y_c q_r with fused multiply-adds (2 FMA
// per entry; the reference spells a x_c x_r + c y_c y_r + b (x_c y_r + y_c x_r) in SSE mul/add).  Per term this
// differs by an ulp or so; the sums are compared against the oracle at 1e-4 either way (summation order).
#include "ctx.h"

#include <cuda_pipeline.h>

namespace {

constexpr int K8_THREADS = 128;
constexpr int NV = 91;
constexpr int K8_ROW = 84;                                              // tile row stride (floats): 72 J + 8 res_toZeroF + 4 pad, conflict-free LDS.128
constexpr int K8_TILE_FLOATS = 32 * K8_ROW;
static_assert(NV * (K8_THREADS + 1) <= (K8_THREADS / 32) * 2 * K8_TILE_FLOATS, "fold scratch fits in the tile buffers");
constexpr size_t K8_STREAM_SMEM = (size_t)(K8_THREADS / 32) * 2 * K8_TILE_FLOATS * 4;

struct K8Args {
    int n, marg;
    const int *flags, *r_point, *r_target, *host;
    const float *J, *rtz, *adHTdelta, *idepth, *idepth_zero;
    float4 cDelta;
    float* res_pt;
    double* acc_top;
    int* acc_cnt;
};

// one active residual: its record j (72 floats, in registers) and, for modes 1/2, its res_toZeroF z -> 91 sums into acc,
// per-point pieces to res_pt.  dd = idepth - idepth_zero of the residual's point (mode 1 only).
__device__ __forceinline__ void k8_residual(const K8Args& A, const float* j, const float* z, float dd, const float* dp, int r, int mode, float* acc) {
    float res[8];
    if (mode == 0) {
#pragma unroll
        for (int i = 0; i < 8; i++) res[i] = j[i];
    } else if (mode == 2) {
#pragma unroll
        for (int i = 0; i < 8; i++) res[i] = z[i];
    } else {
        // resApprox = res_toZeroF + J * delta (AccumulatedTopHessian.cc:47-64)
        float dx = 0, dy = 0, cx = 0, cy = 0;
#pragma unroll
        for (int i = 0; i < 6; i++) { dx += j[40 + i] * dp[i]; dy += j[46 + i] * dp[i]; }
        const float cd[4] = {A.cDelta.x, A.cDelta.y, A.cDelta.z, A.cDelta.w};
#pragma unroll
        for (int i = 0; i < 4; i++) { cx += j[52 + i] * cd[i]; cy += j[56 + i] * cd[i]; }
        const float Jx = dx + cx + j[60] * dd, Jy = dy + cy + j[61] * dd;
        const float da = dp[6], db = dp[7];
#pragma unroll
        for (int i = 0; i < 8; i++) {
            float v = z[i];
            v = v + j[8 + i] * Jx;
            v = v + j[16 + i] * Jy;
            v = v + j[24 + i] * da;
            v = v + j[32 + i] * db;
            res[i] = v;
        }
    }
    // these five feed the per-point pieces below, which are bit-exact against the oracle: reference mul/add order
    float JI_r0 = 0, JI_r1 = 0, Jab_r0 = 0, Jab_r1 = 0, rr = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) {
        JI_r0 += res[i] * j[8 + i];
        JI_r1 += res[i] * j[16 + i];
        Jab_r0 += res[i] * j[24 + i];
        Jab_r1 += res[i] * j[32 + i];
        rr += res[i] * res[i];
    }
    // x = [Jpdc[0] | Jpdxi[0]], y = [Jpdc[1] | Jpdxi[1]]; p, q = the 2x2 weight [a b; b c] applied to (x, y)
    float x[10], y[10], p[10], q[10];
#pragma unroll
    for (int i = 0; i < 4; i++) { x[i] = j[52 + i]; y[i] = j[56 + i]; }
#pragma unroll
    for (int i = 0; i < 6; i++) { x[4 + i] = j[40 + i]; y[4 + i] = j[46 + i]; }
    const float a = j[62], b = j[63], cc = j[64];
#pragma unroll
    for (int i = 0; i < 10; i++) { p[i] = __fmaf_rn(b, y[i], a * x[i]); q[i] = __fmaf_rn(cc, y[i], b * x[i]); }
    int idx = 0;
#pragma unroll
    for (int r_ = 0; r_ < 10; r_++)
#pragma unroll
        for (int c_ = r_; c_ < 10; c_++) { acc[idx] = __fmaf_rn(x[c_], p[r_], __fmaf_rn(y[c_], q[r_], acc[idx])); idx++; }
#pragma unroll
    for (int r_ = 0; r_ < 10; r_++) {
        acc[55 + 3 * r_ + 0] = __fmaf_rn(x[r_], j[65], __fmaf_rn(y[r_], j[66], acc[55 + 3 * r_ + 0]));
        acc[55 + 3 * r_ + 1] = __fmaf_rn(x[r_], j[67], __fmaf_rn(y[r_], j[68], acc[55 + 3 * r_ + 1]));
        acc[55 + 3 * r_ + 2] = __fmaf_rn(x[r_], JI_r0, __fmaf_rn(y[r_], JI_r1, acc[55 + 3 * r_ + 2]));
    }
    acc[85] += j[69]; acc[86] += j[70]; acc[87] += Jab_r0; acc[88] += j[71]; acc[89] += Jab_r1; acc[90] += rr;
    // per-point pieces (AccumulatedTopHessian.cc:92-96): kept in the reference's mul/add order, K9 sums them per point
    const float t0 = j[62] * j[60] + j[63] * j[61], t1 = j[63] * j[60] + j[64] * j[61];
    float4 o0, o1;
    o0.x = t0 * j[60] + t1 * j[61];                 // Hdd
    o0.y = JI_r0 * j[60] + JI_r1 * j[61];           // bd
    o0.z = j[52] * t0 + j[56] * t1;                 // Hcd
    o0.w = j[53] * t0 + j[57] * t1;
    o1.x = j[54] * t0 + j[58] * t1;
    o1.y = j[55] * t0 + j[59] * t1;
    o1.z = 1.f;
    o1.w = 0.f;
    float4* rp = reinterpret_cast<float4*>(A.res_pt + 8 * (size_t)r);
    rp[0] = o0; rp[1] = o1;
}

// fold the CTA's per-thread sums and add them to the key's fp64 accumulator.  The 128 x 91 partial sums go through
// shared memory transposed (row stride 129: writes and reads both conflict-free): 91 stores per thread, then thread i
// sums column i.  5x fewer instructions than 91 five-step shuffle trees per warp, which at 128 residuals per CTA cost
// more than forming the products.
constexpr int K8_FOLD_STRIDE = K8_THREADS + 1;
constexpr int K8_FOLD_FLOATS = NV * K8_FOLD_STRIDE;
__device__ __forceinline__ void k8_cta_fold(const K8Args& A, float* acc, int used, int key, float* scratch) {
    __shared__ int s_used[K8_THREADS / 32];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
    for (int i = 0; i < NV; i++) scratch[i * K8_FOLD_STRIDE + threadIdx.x] = acc[i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) used += __shfl_xor_sync(0xffffffffu, used, o);
    if (lane == 0) s_used[wid] = used;
    __syncthreads();
    if (threadIdx.x < NV) {
        const float* col = scratch + threadIdx.x * K8_FOLD_STRIDE;
        float v0 = 0.f, v1 = 0.f, v2 = 0.f, v3 = 0.f;
#pragma unroll 8
        for (int k = 0; k < K8_THREADS; k += 4) { v0 += col[k]; v1 += col[k + 1]; v2 += col[k + 2]; v3 += col[k + 3]; }
        const float v = (v0 + v1) + (v2 + v3);
        if (v != 0.f) atomicAdd(A.acc_top + (size_t)key * WIN_TOP_STRIDE + threadIdx.x, (double)v);
    } else if (threadIdx.x == NV) {
        int u = 0;
        for (int wq = 0; wq < K8_THREADS / 32; wq++) u += s_used[wq];
        if (u) atomicAdd(A.acc_cnt + key, u);
    }
}

// chunks of <= 128 residuals: thread per residual
__global__ void __launch_bounds__(K8_THREADS)
k8_top_kernel(K8Args A, const int* __restrict__ chunk, const int* __restrict__ order) {
    const int key = chunk[3 * blockIdx.x], start = chunk[3 * blockIdx.x + 1], count = chunk[3 * blockIdx.x + 2];
    const int nn = A.n * A.n;
    const int mode = A.marg ? 2 : key >= nn ? 1 : 0;   // marg: addPoint<2>, the list holds only the marginalised points' residuals
    const float* dp = A.adHTdelta + 8 * (key >= nn ? key - nn : key);       // key = mode n^2 + host + n target
    float acc[NV];
#pragma unroll
    for (int i = 0; i < NV; i++) acc[i] = 0.f;
    int used = 0;
    if (threadIdx.x < count) {
        const int r = order[start + threadIdx.x];
        if (A.flags[r] & LDSO_RF_ACTIVE) {
            used = 1;
            const float4* j4 = reinterpret_cast<const float4*>(A.J + (size_t)r * LDSO_J_FLOATS);
            float j[LDSO_J_FLOATS], z[8] = {0, 0, 0, 0, 0, 0, 0, 0};
            float dd = 0.f;
#pragma unroll
            for (int i = 0; i < LDSO_J_FLOATS / 4; i++) {
                float4 t = __ldg(j4 + i);
                j[4 * i] = t.x; j[4 * i + 1] = t.y; j[4 * i + 2] = t.z; j[4 * i + 3] = t.w;
            }
            if (mode) {
                const float4* z4 = reinterpret_cast<const float4*>(A.rtz + 8 * (size_t)r);
                const float4 z0 = __ldg(z4), z1 = __ldg(z4 + 1);
                z[0] = z0.x; z[1] = z0.y; z[2] = z0.z; z[3] = z0.w; z[4] = z1.x; z[5] = z1.y; z[6] = z1.z; z[7] = z1.w;
                if (mode == 1) { const int pt = A.r_point[r]; dd = A.idepth[pt] - A.idepth_zero[pt]; }
            }
            k8_residual(A, j, z, dd, dp, r, mode, acc);
        } else {
            A.res_pt[8 * (size_t)r + 6] = 0.f;
        }
    }
    __shared__ float scratch[K8_FOLD_FLOATS];
    k8_cta_fold(A, acc, used, key, scratch);
}

// chunks of up to 1024 residuals: warp w takes tiles w, w+4, ... of 32 records.  Three tiles deep per warp: the record
// ids of tile i+2 are being loaded, the records of tile i+1 (288 B of J, + 32 B res_toZeroF in modes 1/2) are in
// flight to the other shared-memory buffer (cp.async) together with their flags / point ids (registers), tile i is
// multiplied.  No global load sits on the critical path of the multiply.
__global__ void __launch_bounds__(K8_THREADS)
k8_top_stream_kernel(K8Args A, const int* __restrict__ chunk, const int* __restrict__ order) {
    extern __shared__ __align__(16) float dyn[];           // [4 warps][2][32 x K8_ROW] floats
    const int key = chunk[3 * blockIdx.x], start = chunk[3 * blockIdx.x + 1], count = chunk[3 * blockIdx.x + 2];
    const int nn = A.n * A.n;
    const int mode = A.marg ? 2 : key >= nn ? 1 : 0;
    const float* dp = A.adHTdelta + 8 * (key >= nn ? key - nn : key);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    float* tiles = dyn + (size_t)wid * 2 * K8_TILE_FLOATS;
    float acc[NV];
#pragma unroll
    for (int i = 0; i < NV; i++) acc[i] = 0.f;
    int used = 0;
    auto load_id = [&](int base) -> int { return base + lane < count ? __ldg(order + start + base + lane) : -1; };
    // start the copy of a tile whose record ids (one per lane, -1 past the end) are r.  Half a warp copies the first
    // 256 B of a record per instruction (16 lanes x 16 B, two records per instruction), two lanes per record copy the
    // last 32 B (and the 32 B of res_toZeroF): every address is a lane constant plus an unrolled immediate, the record
    // id comes by shuffle — three instructions per 16-byte piece.
    const int hi = lane >> 4, lo = lane & 15, h2 = lane >> 1, l2 = lane & 1;
    auto issue = [&](int r, int buf) {
        const int rc = max(r, 0);                          // past-the-end lanes re-copy record 0 (never read)
        float* my = tiles + buf * K8_TILE_FLOATS;
        float* dst = my + hi * K8_ROW + 4 * lo;
        const float* src = A.J + 4 * lo;
#pragma unroll
        for (int i = 0; i < 16; i++) {
            const int rr = __shfl_sync(0xffffffffu, rc, 2 * i + hi);
            __pipeline_memcpy_async(dst + 2 * i * K8_ROW, src + (size_t)rr * LDSO_J_FLOATS, 16);
        }
#pragma unroll
        for (int i = 0; i < 2; i++) {
            const int rr = __shfl_sync(0xffffffffu, rc, 16 * i + h2);
            __pipeline_memcpy_async(my + (16 * i + h2) * K8_ROW + 64 + 4 * l2, A.J + (size_t)rr * LDSO_J_FLOATS + 64 + 4 * l2, 16);
            if (mode) __pipeline_memcpy_async(my + (16 * i + h2) * K8_ROW + 72 + 4 * l2, A.rtz + 8 * (size_t)rr + 4 * l2, 16);
        }
        __pipeline_commit();
    };
    int buf = 0, base = wid * 32;
    int r0 = load_id(base), r1 = load_id(base + K8_THREADS), r2;
    int fl0 = 0, fl1 = 0, pt0 = 0, pt1 = 0;
    if (base < count) {
        issue(r0, 0);
        if (r0 >= 0) { fl0 = __ldg(A.flags + r0); if (mode == 1) pt0 = __ldg(A.r_point + r0); }
    }
    for (; base < count; base += K8_THREADS) {
        r2 = load_id(base + 2 * K8_THREADS);
        if (base + K8_THREADS < count) {
            issue(r1, buf ^ 1);
            if (r1 >= 0) { fl1 = __ldg(A.flags + r1); if (mode == 1) pt1 = __ldg(A.r_point + r1); }
        }
        float dd = 0.f;
        if (mode == 1 && r0 >= 0) dd = __ldg(A.idepth + pt0) - __ldg(A.idepth_zero + pt0);
        if (base + K8_THREADS < count) __pipeline_wait_prior(1); else __pipeline_wait_prior(0);
        __syncwarp();
        if (r0 >= 0) {
            if (fl0 & LDSO_RF_ACTIVE) {
                used++;
                const float* my = tiles + buf * K8_TILE_FLOATS + lane * K8_ROW;
                float j[LDSO_J_FLOATS], z[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
                for (int i = 0; i < LDSO_J_FLOATS / 4; i++) {
                    const float4 t = *reinterpret_cast<const float4*>(my + 4 * i);
                    j[4 * i] = t.x; j[4 * i + 1] = t.y; j[4 * i + 2] = t.z; j[4 * i + 3] = t.w;
                }
                if (mode) {
                    const float4 z0 = *reinterpret_cast<const float4*>(my + 72), z1 = *reinterpret_cast<const float4*>(my + 76);
                    z[0] = z0.x; z[1] = z0.y; z[2] = z0.z; z[3] = z0.w; z[4] = z1.x; z[5] = z1.y; z[6] = z1.z; z[7] = z1.w;
                }
                k8_residual(A, j, z, dd, dp, r0, mode, acc);
            } else {
                A.res_pt[8 * (size_t)r0 + 6] = 0.f;
            }
        }
        __syncwarp();
        r0 = r1; r1 = r2; fl0 = fl1; pt0 = pt1;
        buf ^= 1;
    }
    __syncthreads();                                        // every warp is done with its tiles: the buffers become the fold scratch
    k8_cta_fold(A, acc, used, key, dyn);
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
        K8Args A;
        A.n = W.n; A.marg = marg;
        A.flags = W.flags; A.r_point = W.r_point; A.r_target = W.r_target; A.host = W.host;
        A.J = W.J; A.rtz = W.rtz; A.adHTdelta = W.adHTdelta; A.idepth = W.idepth; A.idepth_zero = W.idepth_zero;
        A.cDelta = make_float4(W.cDeltaF[0], W.cDeltaF[1], W.cDeltaF[2], W.cDeltaF[3]);
        A.res_pt = W.res_pt; A.acc_top = W.acc_top; A.acc_cnt = W.acc_top_cnt;
        prof_begin(c, c->stream);
        if (W.chunk_len > K8_THREADS) {
            if (!c->k8_attr_set) {   // per ctx = per device
                LDSO_CUDA_OK(c, cudaFuncSetAttribute(k8_top_stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)K8_STREAM_SMEM));
                c->k8_attr_set = true;
            }
            k8_top_stream_kernel<<<W.n_chunks, K8_THREADS, K8_STREAM_SMEM, c->stream>>>(A, W.chunk, W.order);
        } else {
            k8_top_kernel<<<W.n_chunks, K8_THREADS, 0, c->stream>>>(A, W.chunk, W.order);
        }
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
