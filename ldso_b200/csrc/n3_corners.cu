// n3_corners.cu — N3: FeatureDetector::DetectCorners (reference src/frontend/FeatureDetector.cc:34-124) up to the corner
// flags, ShiTomasiScore (include/frontend/FeatureDetector.h:50-81).  The descriptors (IC_Angle, rotated BRIEF) stay on the
// host.
//
//   n3_cell_kernel     one warp per grid cell: cell maximum of absSquaredGrad, Shi-Tomasi score of every pixel above the
//                      gradient threshold (a lane sums the 8x8 box of its candidate in the reference's order, so the
//                      score is bit-identical), the global score maximum through an integer atomicMax on the float's
//                      bits, and the K best candidates of the cell by repeated warp arg-max (ties to the earlier
//                      candidate, i.e. a stable sort);
//   n3_compact_kernel  exclusive scan of the per-cell counts, features written in the reference's order (gx outer, gy
//                      inner, score descending inside a cell);
//   n3_corner_kernel   thread per feature: score > 0.01 * max score, then the pairwise 5-px suppression, evaluated per
//                      feature from the same pair rule the reference's double loop applies (the loop never skips a
//                      feature that was already suppressed, so the outcome does not depend on the visiting order).
#include "ctx.h"

namespace {

struct N3Geom { int w, h, g, nx, ny, skip, K; };

__device__ __forceinline__ float shi_tomasi(const float4* __restrict__ img, int w, int h, int u, int v) {
    const int x_min = u - 4, x_max = u + 4, y_min = v - 4, y_max = v + 4;
    if (x_min < 1 || x_max >= w - 1 || y_min < 1 || y_max >= h - 1) return 0.f;
    float dXX = 0.f, dYY = 0.f, dXY = 0.f;
    for (int y = y_min; y < y_max; ++y)
        for (int x = x_min; x < x_max; ++x) {
            const float4 t = __ldg(img + y * w + x);
            dXX += t.y * t.y; dYY += t.z * t.z; dXY += t.y * t.z;
        }
    // the reference divides in double and rounds to float; its sqrt is std::sqrt(float) (using namespace std, NumTypes.h:12)
    dXX = (float)((double)dXX / (2.0 * 64)); dYY = (float)((double)dYY / (2.0 * 64)); dXY = (float)((double)dXY / (2.0 * 64));
    const float s = dXX + dYY, q = dXX * dYY - dXY * dXY;
    return 0.5f * (s - sqrtf(s * s - 4 * q));
}

__global__ void __launch_bounds__(32)
n3_cell_kernel(const N3Geom G, const float4* __restrict__ img, float* __restrict__ slot_xys, int* __restrict__ cnt,
               unsigned int* __restrict__ max_score_bits) {
    extern __shared__ float s_score[];               // [g*g] in candidate order q = x * g + y; -inf = not a candidate
    const int cell = blockIdx.x, lane = threadIdx.x;
    const int gx = G.skip + cell / G.ny, gy = G.skip + cell % G.ny;
    const int x0 = gx * G.g, y0 = gy * G.g, gg = G.g * G.g;
    float mx = 0.f;
    for (int q = lane; q < gg; q += 32) {
        const int x = q / G.g, y = q - x * G.g;
        mx = fmaxf(mx, __ldg(&img[(y0 + y) * G.w + x0 + x].w));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    const float gradTH = (0.5f * mx) > 5.f ? 0.5f * mx : 5.f;
    float smax = 0.f;
    for (int q = lane; q < gg; q += 32) {
        const int x = q / G.g, y = q - x * G.g;
        float s = -INFINITY;
        if (__ldg(&img[(y0 + y) * G.w + x0 + x].w) > gradTH) {
            s = shi_tomasi(img, G.w, G.h, x0 + x, y0 + y);
            smax = fmaxf(smax, s);
            if (!(s == s)) s = -3.0e38f;             // NaN (negative discriminant by rounding) sorts last
        }
        s_score[q] = s;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) smax = fmaxf(smax, __shfl_xor_sync(0xffffffffu, smax, o));
    if (lane == 0 && smax > 0.f) atomicMax(max_score_bits, __float_as_uint(smax));
    __syncwarp();
    int picked = 0;
    for (int k = 0; k < G.K; k++) {
        float best = -INFINITY;
        int bq = 0x7fffffff;
        for (int q = lane; q < gg; q += 32) {
            const float s = s_score[q];
            if (s > best) { best = s; bq = q; }      // strict: the earliest q of a lane wins its ties
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ob = __shfl_xor_sync(0xffffffffu, best, o);
            const int oq = __shfl_xor_sync(0xffffffffu, bq, o);
            if (ob > best || (ob == best && oq < bq)) { best = ob; bq = oq; }
        }
        if (bq == 0x7fffffff) break;
        if (lane == 0) {
            const int x = bq / G.g, y = bq - x * G.g;
            float* o = slot_xys + 3 * ((size_t)cell * G.K + k);
            o[0] = (float)(x0 + x); o[1] = (float)(y0 + y); o[2] = best;
            s_score[bq] = -INFINITY;
        }
        picked++;
        __syncwarp();
    }
    if (lane == 0) cnt[cell] = picked;
}

__global__ void __launch_bounds__(1024)
n3_compact_kernel(int ncells, int K, const float* __restrict__ slot_xys, const int* __restrict__ cnt, float* __restrict__ out_xys,
                  int* __restrict__ n_out) {
    __shared__ int s_warp[32];
    __shared__ int s_base;
    if (threadIdx.x == 0) s_base = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    for (int base = 0; base < ncells; base += 1024) {
        const int c = base + threadIdx.x;
        const int n = c < ncells ? cnt[c] : 0;
        int incl = n;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
        if (lane == 31) s_warp[wid] = incl;
        __syncthreads();
        int woff = 0, total = 0;
        for (int q = 0; q < 32; q++) { if (q < wid) woff += s_warp[q]; total += s_warp[q]; }
        const int off = s_base + woff + incl - n;
        for (int k = 0; k < n; k++)
            for (int j = 0; j < 3; j++) out_xys[3 * (size_t)(off + k) + j] = slot_xys[3 * ((size_t)c * K + k) + j];
        __syncthreads();
        if (threadIdx.x == 0) s_base += total;
        __syncthreads();
    }
    if (threadIdx.x == 0) *n_out = s_base;
}

__global__ void n3_corner_kernel(const int* __restrict__ n_ptr, const float* __restrict__ xys, const unsigned int* __restrict__ max_score_bits,
                                 unsigned char* __restrict__ is_corner, int* __restrict__ n_corners) {
    const int n = *n_ptr;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float scoreTH = (float)(0.01 * (double)__uint_as_float(*max_score_bits));
    const float xi = xys[3 * i], yi = xys[3 * i + 1], si = xys[3 * i + 2];
    bool corner = si > scoreTH;
    if (corner)
        for (int j = 0; j < n; j++) {
            if (j == i) continue;
            const float sj = xys[3 * j + 2];
            if (!(sj > scoreTH)) continue;
            const float dx = xi - xys[3 * j], dy = yi - xys[3 * j + 1];
            if (sqrtf(dx * dx + dy * dy) < 5.f) {
                // pair (a, b) with a < b: a survives only if s_a > s_b, otherwise b survives
                if (i < j ? !(si > sj) : (sj > si)) { corner = false; break; }
            }
        }
    is_corner[i] = corner ? 1 : 0;
    if (corner) atomicAdd(n_corners, 1);
}

struct N3Dev { float *slots = nullptr, *out = nullptr; int* cnt = nullptr; unsigned int* misc = nullptr; unsigned char* flag = nullptr;
               size_t cap_slots = 0; int cap_cells = 0; };

}  // namespace

void corners_free(ldso_gpu_ctx* c) {
    N3Dev* D = (N3Dev*)c->corners;
    if (!D) return;
    cudaFree(D->slots); cudaFree(D->out); cudaFree(D->cnt); cudaFree(D->misc); cudaFree(D->flag);
    delete D;
    c->corners = nullptr;
}

extern "C" int ldso_gpu_detect_corners(ldso_gpu_ctx* c, int slot, int nFeatures, int max_out, float* ox, float* oy, float* oscore,
                                       unsigned char* is_corner, int* n_out, int* n_corners) {
    LDSO_ENTER(c);
    LDSO_REQUIRE(c, slot >= 0 && slot < c->cfg.n_frame_slots && nFeatures > 0 && max_out > 0 && ox && oy && oscore && is_corner && n_out && n_corners);
    const int w = c->wG[0], h = c->hG[0];
    N3Geom G;
    G.w = w; G.h = h;
    G.g = int(sqrtf(w * h / nFeatures) + 0.5);
    LDSO_REQUIRE(c, G.g >= 1 && G.g * G.g * sizeof(float) <= 48 * 1024);
    const int gridX = w / G.g + 1, gridY = h / G.g + 1;
    const float nfeatInGrid = float(nFeatures) / (w * h) * (G.g * G.g);
    G.skip = (15 * 2 / G.g) + 1;                      // HALF_PATCH_SIZE = 15 (FeatureDetector.h:20)
    G.nx = gridX - 2 * G.skip; G.ny = gridY - 2 * G.skip;
    G.K = (int)floorf(nfeatInGrid) + 1;
    *n_out = 0; *n_corners = 0;
    if (G.nx <= 0 || G.ny <= 0) return 0;
    const int ncells = G.nx * G.ny;
    if (!c->corners) c->corners = new N3Dev();
    N3Dev* D = (N3Dev*)c->corners;
    const size_t need = (size_t)ncells * G.K;
    if (need > D->cap_slots || ncells > D->cap_cells) {
        cudaFree(D->slots); cudaFree(D->out); cudaFree(D->cnt); cudaFree(D->flag); cudaFree(D->misc);
        LDSO_CUDA_OK(c, cudaMalloc((void**)&D->slots, sizeof(float) * 3 * need));
        LDSO_CUDA_OK(c, cudaMalloc((void**)&D->out, sizeof(float) * 3 * need));
        LDSO_CUDA_OK(c, cudaMalloc((void**)&D->cnt, sizeof(int) * ncells));
        LDSO_CUDA_OK(c, cudaMalloc((void**)&D->flag, need));
        LDSO_CUDA_OK(c, cudaMalloc((void**)&D->misc, 16));
        D->cap_slots = need; D->cap_cells = ncells;
    }
    cudaStream_t s = c->stream;
    LDSO_CUDA_OK(c, cudaMemsetAsync(D->misc, 0, 16, s));   // [0] max score bits, [1] n, [2] corners
    const float4* img = c->frame[slot] + c->lvl_off[0];
    n3_cell_kernel<<<ncells, 32, sizeof(float) * G.g * G.g, s>>>(G, img, D->slots, D->cnt, D->misc);
    n3_compact_kernel<<<1, 1024, 0, s>>>(ncells, G.K, D->slots, D->cnt, D->out, (int*)D->misc + 1);
    n3_corner_kernel<<<(int)((need + 127) / 128), 128, 0, s>>>((int*)D->misc + 1, D->out, D->misc, D->flag, (int*)D->misc + 2);
    c->launches += 3;
    LDSO_CUDA_OK(c, cudaGetLastError());
    // read-back: 13 bytes per feature
    char* st = (char*)c->scratch_host;
    LDSO_CUDA_OK(c, ldso_copy(c, st, D->misc, 16, cudaMemcpyDeviceToHost, s));
    LDSO_CUDA_OK(c, ldso_copy(c, st + 16, D->out, sizeof(float) * 3 * need, cudaMemcpyDeviceToHost, s));
    LDSO_CUDA_OK(c, ldso_copy(c, st + 16 + sizeof(float) * 3 * need, D->flag, need, cudaMemcpyDeviceToHost, s));
    LDSO_CUDA_OK(c, ldso_stream_sync(s));
    const int n = ((int*)st)[1];
    const float* xys = (const float*)(st + 16);
    const unsigned char* fl = (const unsigned char*)(st + 16 + sizeof(float) * 3 * need);
    int m = 0, nc = 0;
    for (int i = 0; i < n && m < max_out; i++, m++) {
        ox[m] = xys[3 * i]; oy[m] = xys[3 * i + 1]; oscore[m] = xys[3 * i + 2]; is_corner[m] = fl[i];
        nc += fl[i];
    }
    *n_out = m; *n_corners = nc;
    return 0;
}
