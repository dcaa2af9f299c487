// k1_pyramid.cu — K1: FrameHessian::makeImages (reference src/internal/FrameHessian.cc:59-115).
//
// One launch builds every pyramid level of a frame.  Each CTA owns a 64x32 level-0 tile, stages it in
// shared memory with a halo of 2^(L-1) pixels, box-filters it hierarchically (the reference's
// 0.25f*(a+b+c+d) association order, so the result is bit-identical), takes central differences per
// level from shared memory and writes ONE float4 {I, dx, dy, absSquaredGrad} per pixel per level —
// the reference's dIp[] (12 B AoS) and absSquaredGrad[] (4 B) fused into a single 16-B texel that the
// bilinear gathers of K3/K5/K7 fetch with one LDG.128.  Input is read once from HBM (halo re-reads
// hit L2), output is written once: 4 + 16*sum_l(w_l h_l)/(w h) ~ 25.3 B per level-0 pixel.
//
// The reference's gradient loop runs over the flat index range [w, w(h-1)) (FrameHessian.cc:95), so
// column 0 reads the previous row's last pixel and column w-1 the next row's first one.  Those two
// columns are patched by k1_wrap_columns_kernel, which reads the I channel written by the tile pass.
#include "ctx.h"

namespace {

constexpr int K1_TW = 64, K1_TH = 32, K1_THREADS = 256;

struct K1Params {
    int levels;
    int w[LDSO_MAX_LEVELS], h[LDSO_MAX_LEVELS];
    unsigned int off[LDSO_MAX_LEVELS];   // level offset inside the slot, in texels
    int gammaWeights;
};

__device__ __forceinline__ float clamp_grad(float d) {
    // FrameHessian.cc:100-101
    return (isnan(d) || fabsf(d) > 255.0f) ? 0.f : d;
}

__device__ __forceinline__ float abs_sq_grad(float I, float dx, float dy, const float* __restrict__ B, int gammaWeights) {
    // FrameHessian.cc:106-111, include/internal/CalibHessian.h:102-111.  No FMA contraction: the
    // selection maps derived from this value must be bit-exact.
    float a = __fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy));
    if (gammaWeights == 1) {
        int c = (int)__fadd_rn(I, 0.5f);
        c = c < 5 ? 5 : (c > 250 ? 250 : c);
        float gw = __fsub_rn(__ldg(B + c + 1), __ldg(B + c));
        a = __fmul_rn(a, __fmul_rn(gw, gw));
    }
    return a;
}

// L = number of pyramid levels, a template parameter so that the shared-memory tile geometry is a compile-time constant:
// the fill / box / store loops unroll completely and all global loads of a thread are in flight at once
template <int L>
__global__ void __launch_bounds__(K1_THREADS) k1_pyramid_kernel(const float* __restrict__ img, float4* __restrict__ out,
                                                                 const float* __restrict__ B, const K1Params p) {
    extern __shared__ float sm[];
    const int tid = threadIdx.x;
    const int X0 = blockIdx.x * K1_TW, Y0 = blockIdx.y * K1_TH;
    int soff[LDSO_MAX_LEVELS], sw[LDSO_MAX_LEVELS], sh[LDSO_MAX_LEVELS];
    {
        int acc = 0;
#pragma unroll
        for (int l = 0; l < LDSO_MAX_LEVELS; l++) {
            if (l < L) {
                int hl = 1 << (L - 1 - l);
                sw[l] = (K1_TW >> l) + 2 * hl;
                sh[l] = (K1_TH >> l) + 2 * hl;
                soff[l] = acc;
                acc += sw[l] * sh[l];
            }
        }
    }
    // level 0: tile + halo, zero outside the image
    {
        const int h0 = 1 << (L - 1);
        const int w = p.w[0], h = p.h[0];
        const int n = sw[0] * sh[0];
#pragma unroll
        for (int k_ = 0; k_ < (n + K1_THREADS - 1) / K1_THREADS; k_++) {
            const int i = tid + k_ * K1_THREADS;
            if (i >= n) continue;
            int sy = i / sw[0], sx = i - sy * sw[0];
            int gx = X0 - h0 + sx, gy = Y0 - h0 + sy;
            float v = 0.f;
            if (gx >= 0 && gx < w && gy >= 0 && gy < h) v = __ldg(img + (size_t)gy * w + gx);
            sm[i] = v;
        }
    }
    __syncthreads();
    // levels 1..L-1: hierarchical 2x2 box, FrameHessian.cc:86-92
#pragma unroll
    for (int l = 1; l < L; l++) {
        const float* src = sm + soff[l - 1];
        float* dst = sm + soff[l];
        const int swp = sw[l - 1];
        const int n = sw[l] * sh[l];
#pragma unroll
        for (int k_ = 0; k_ < (n + K1_THREADS - 1) / K1_THREADS; k_++) {
            const int i = tid + k_ * K1_THREADS;
            if (i >= n) continue;
            int sy = i / sw[l], sx = i - sy * sw[l];
            const float* q = src + (2 * sy) * swp + 2 * sx;
            float s = __fadd_rn(__fadd_rn(__fadd_rn(q[0], q[1]), q[swp]), q[swp + 1]);
            dst[i] = __fmul_rn(0.25f, s);
        }
        __syncthreads();
    }
    // gradients + fused texel store, FrameHessian.cc:95-114
#pragma unroll
    for (int l = 0; l < L; l++) {
        const int hl = 1 << (L - 1 - l);
        const int tw = K1_TW >> l, th = K1_TH >> l;
        const int wl = p.w[l], hgt = p.h[l];
        const float* s = sm + soff[l];
        const int swl = sw[l];
        float4* o = out + p.off[l];
        const int n = tw * th;
#pragma unroll
        for (int k_ = 0; k_ < (n + K1_THREADS - 1) / K1_THREADS; k_++) {
            const int i = tid + k_ * K1_THREADS;
            if (i >= n) continue;
            int ty = i / tw, tx = i - ty * tw;
            int gx = (X0 >> l) + tx, gy = (Y0 >> l) + ty;
            if (gx >= wl || gy >= hgt) continue;
            const float* q = s + (ty + hl) * swl + (tx + hl);
            float I = q[0];
            float dx = 0.f, dy = 0.f, ag = 0.f;
            if (gy > 0 && gy < hgt - 1) {
                dx = clamp_grad(__fmul_rn(0.5f, __fsub_rn(q[1], q[-1])));
                dy = clamp_grad(__fmul_rn(0.5f, __fsub_rn(q[swl], q[-swl])));
                ag = abs_sq_grad(I, dx, dy, B, p.gammaWeights);
            }
            o[(size_t)gy * wl + gx] = make_float4(I, dx, dy, ag);
        }
    }
}

// columns 0 and w-1 of rows 1..h-2: neighbours wrap to the adjacent row (flat-index loop of the reference)
__global__ void k1_wrap_columns_kernel(float4* __restrict__ out, const float* __restrict__ B, const K1Params p) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    for (int l = 0; l < p.levels; l++) {
        int rows = p.h[l] - 2;
        if (i < 2 * rows) {
            int y = 1 + (i >> 1);
            int wl = p.w[l];
            int x = (i & 1) ? wl - 1 : 0;
            float4* o = out + p.off[l];
            size_t idx = (size_t)y * wl + x;
            float4 t = o[idx];
            float left = o[idx - 1].x, right = o[idx + 1].x;
            float dx = clamp_grad(__fmul_rn(0.5f, __fsub_rn(right, left)));
            t.y = dx;
            t.w = abs_sq_grad(t.x, dx, t.z, B, p.gammaWeights);
            o[idx] = t;
            return;
        }
        i -= 2 * rows;
    }
}

}  // namespace

int k1_make_images(ldso_gpu_ctx* c, int slot, const float* img_dev) {
    if (immature_trace_fence(c, slot)) return -1;
    K1Params p;
    p.levels = c->levels;
    p.gammaWeights = c->cfg.gamma_weights_pixel_select;
    int smem_floats = 0, border = 0;
    for (int l = 0; l < c->levels; l++) {
        p.w[l] = c->wG[l];
        p.h[l] = c->hG[l];
        p.off[l] = (unsigned int)c->lvl_off[l];
        int hl = 1 << (c->levels - 1 - l);
        smem_floats += ((K1_TW >> l) + 2 * hl) * ((K1_TH >> l) + 2 * hl);
        border += 2 * (c->hG[l] - 2);
    }
    dim3 grid((c->wG[0] + K1_TW - 1) / K1_TW, (c->hG[0] + K1_TH - 1) / K1_TH);
    const size_t smem = smem_floats * sizeof(float);
    prof_begin(c, c->stream);
#define K1_LAUNCH(LV)                                                                                                          \
    case LV: {                                                                                                                 \
        if (!c->k1_attr_set) { /* per ctx = per device: the attribute belongs to the device the ctx lives on */             \
            LDSO_CUDA_OK(c, cudaFuncSetAttribute(k1_pyramid_kernel<LV>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));  \
            c->k1_attr_set = true;                                                                                             \
        }                                                                                                                      \
        k1_pyramid_kernel<LV><<<grid, K1_THREADS, smem, c->stream>>>(img_dev, c->frame[slot], c->gammaB_dev, p);             \
    } break;
    switch (c->levels) {
        K1_LAUNCH(1) K1_LAUNCH(2) K1_LAUNCH(3) K1_LAUNCH(4) K1_LAUNCH(5) K1_LAUNCH(6)
        default: LDSO_REQUIRE(c, c->levels >= 1 && c->levels <= 6);
    }
#undef K1_LAUNCH
    k1_wrap_columns_kernel<<<(border + 127) / 128, 128, 0, c->stream>>>(c->frame[slot], c->gammaB_dev, p);
    prof_end(c, c->stream, LDSO_PROF_K1, 1);
    c->launches += 2;
    LDSO_CUDA_OK(c, cudaGetLastError());
    return 0;
}
