// a2_selector.cu — A2: PixelSelector::makeHists / select / makeMaps (reference src/frontend/PixelSelector2.cc).
//
// The reference's select() is one sequential scan whose only cross-block dependency is n2, the running
// count of level-0 selections, through dir = directions[randomPattern[n2] & 0xF] (SURVEY.md trap T8).
// Decomposition used here (bit-exact, verified against the oracle):
//   a2_hist / a2_smooth   per 32x32 cell gradient histogram -> 50 % quantile + 7, 3x3 mean, squared  (:36-109)
//   a2_cand0              one thread per pot-block: the level-0 winner for EACH of the 16 directions
//                         (first-occurrence arg-max in scan order, strict >) + a 16-bit "has winner" mask
//   host scan             n2(b) = sum_{b'<b} mask[b'] bit dir(n2(b')) over the pot-blocks in the reference's
//                         nested 4pot/2pot/pot order — 2 bytes per block down, 1 byte per block up
//   a2_final              one thread per 4pot-block: writes level-0 picks, and where a 2pot (4pot) block had
//                         no finer winner runs the level-1 (level-2) arg-max exactly as the reference does
//   a2_thin_*             the random thinning of makeMaps (:151-164) needs each selected pixel's row-major
//                         rank: per-row counts, scan, apply
// All float tests are evaluated with the reference's expressions (library built with -fmad=false).
#include "ctx.h"

#include <algorithm>
#include <cmath>
#include <vector>

namespace {

__constant__ float c_dirs[16][2] = {{0, 1.0000f},       {0.3827f, 0.9239f}, {0.1951f, 0.9808f},  {0.9239f, 0.3827f},
                                    {0.7071f, 0.7071f},  {0.3827f, -0.9239f}, {0.8315f, 0.5556f},  {0.8315f, -0.5556f},
                                    {0.5556f, -0.8315f}, {0.9808f, 0.1951f},  {0.9239f, -0.3827f}, {0.7071f, -0.7071f},
                                    {0.5556f, 0.8315f},  {0.9808f, -0.1951f}, {1.0000f, 0.0000f},  {0.1951f, -0.9808f}};

__global__ void a2_hist_kernel(const float4* __restrict__ l0, int w, int h, int w32, float cut, float add, float* __restrict__ ths) {
    __shared__ int hist[100];
    const int cx = blockIdx.x % w32, cy = blockIdx.x / w32;
    if (threadIdx.x < 100) hist[threadIdx.x] = 0;
    __syncthreads();
    for (int k = threadIdx.x; k < 1024; k += blockDim.x) {
        int i = k & 31, j = k >> 5;
        int it = i + 32 * cx, jt = j + 32 * cy;
        if (it > w - 2 || jt > h - 2 || it < 1 || jt < 1) continue;
        int g = (int)sqrtf(l0[it + jt * w].w);
        if (g > 48) g = 48;
        atomicAdd(&hist[g + 1], 1);
        atomicAdd(&hist[0], 1);
    }
    __syncthreads();
    if (threadIdx.x == 0) {   // computeHistQuantil (:28-35)
        int th = (int)(hist[0] * cut + 0.5f);
        int q = 90;
        for (int i = 0; i < 90; i++) {
            th -= hist[i + 1];
            if (th < 0) { q = i; break; }
        }
        ths[cx + cy * w32] = q + add;
    }
}

__global__ void a2_smooth_kernel(const float* __restrict__ ths, float* __restrict__ sm, int w32, int h32) {
    int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= w32 * h32) return;
    int x = k % w32, y = k / w32;
    float sum = 0, num = 0;
    if (x > 0) {
        if (y > 0) { num++; sum += ths[x - 1 + (y - 1) * w32]; }
        if (y < h32 - 1) { num++; sum += ths[x - 1 + (y + 1) * w32]; }
        num++; sum += ths[x - 1 + y * w32];
    }
    if (x < w32 - 1) {
        if (y > 0) { num++; sum += ths[x + 1 + (y - 1) * w32]; }
        if (y < h32 - 1) { num++; sum += ths[x + 1 + (y + 1) * w32]; }
        num++; sum += ths[x + 1 + y * w32];
    }
    if (y > 0) { num++; sum += ths[x + (y - 1) * w32]; }
    if (y < h32 - 1) { num++; sum += ths[x + (y + 1) * w32]; }
    num++; sum += ths[x + y * w32];
    sm[k] = (sum / num) * (sum / num);
}

struct SelGeom {
    int w, h, w1, w2, pot, n4x, n4y, thsStep, dirDist;
    float thFactor, dw1, dw2;
};

// origin of pot-block slot b = i4*16 + s3*4 + s2; false when the slot lies outside the image
__device__ __forceinline__ bool slot_origin(const SelGeom& g, int b, int& x, int& y) {
    int i4 = b >> 4, s3 = (b >> 2) & 3, s2 = b & 3;
    x = (i4 % g.n4x) * 4 * g.pot + (s3 & 1) * 2 * g.pot + (s2 & 1) * g.pot;
    y = (i4 / g.n4x) * 4 * g.pot + (s3 >> 1) * 2 * g.pot + (s2 >> 1) * g.pot;
    return x < g.w && y < g.h;
}

// lpb (1, 8 or 32) consecutive lanes share one pot-block: each lane keeps the first strict maximum of the pixels it
// scans (q = lane, lane + lpb, ... in the reference's y1-major order), then a (value, scan position) arg-max per
// direction across the group restores the reference's first-occurrence winner.
__global__ void a2_cand0_kernel(const SelGeom g, const float4* __restrict__ l0, const float* __restrict__ thsS,
                                unsigned short* __restrict__ mask, int* __restrict__ cand, int nslots, int lpb) {
    const int gid = blockIdx.x * blockDim.x + threadIdx.x;
    const int b = gid / lpb, sub = gid % lpb;
    int x0 = 0, y0 = 0;
    float best[16];
    int bidx[16], bq[16];
#pragma unroll
    for (int d = 0; d < 16; d++) { best[d] = 0.f; bidx[d] = -1; bq[d] = 0; }
    const bool live = b < nslots && slot_origin(g, b, x0, y0);
    if (live) {
        const int pp = g.pot * g.pot;
        for (int q = sub; q < pp; q += lpb) {
            const int y1 = q / g.pot, x1 = q - y1 * g.pot;
            const int xf = x0 + x1, yf = y0 + y1;
            if (xf >= g.w || yf >= g.h) continue;
            if (xf < 4 || xf >= g.w - 5 || yf < 4 || yf > g.h - 4) continue;
            const int idx = xf + g.w * yf;
            const float4 t = __ldg(l0 + idx);
            const float th0 = thsS[(xf >> 5) + (yf >> 5) * g.thsStep];
            if (t.w > th0 * g.thFactor) {
#pragma unroll
                for (int d = 0; d < 16; d++) {
                    const float dn = g.dirDist ? fabsf(t.y * c_dirs[d][0] + t.z * c_dirs[d][1]) : t.w;
                    if (dn > best[d]) { best[d] = dn; bidx[d] = idx; bq[d] = q; }
                }
            }
        }
    }
    for (int o = lpb >> 1; o > 0; o >>= 1) {
#pragma unroll
        for (int d = 0; d < 16; d++) {
            const float ov = __shfl_xor_sync(0xffffffffu, best[d], o);
            const int oq = __shfl_xor_sync(0xffffffffu, bq[d], o);
            const int oi = __shfl_xor_sync(0xffffffffu, bidx[d], o);
            if (oi >= 0 && (bidx[d] < 0 || ov > best[d] || (ov == best[d] && oq < bq[d]))) { best[d] = ov; bq[d] = oq; bidx[d] = oi; }
        }
    }
    if (b >= nslots) return;
    if (sub == 0) {
        unsigned m = 0;
#pragma unroll
        for (int d = 0; d < 16; d++) m |= (bidx[d] > 0 ? 1u : 0u) << d;
        mask[b] = (unsigned short)m;
    }
    // the 16 winners of the block, spread over the group's lanes
    for (int d = sub; d < 16; d += lpb) {
        int v = -1;
#pragma unroll
        for (int e = 0; e < 16; e++) if (e == d) v = bidx[e];
        cand[(size_t)b * 16 + d] = v;
    }
}

// first-occurrence arg-max across the warp: larger value wins, ties go to the earlier scan position
__device__ __forceinline__ void warp_argmax(float& v, int& q, int& idx) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, v, o);
        const int oq = __shfl_xor_sync(0xffffffffu, q, o);
        const int oi = __shfl_xor_sync(0xffffffffu, idx, o);
        if (oi >= 0 && (idx < 0 || ov > v || (ov == v && oq < q))) { v = ov; q = oq; idx = oi; }
    }
}

// One CTA per 4pot-block, one warp per 2pot-block.  The reference scans the pixels of a 2pot-block in order and keeps
// the first strict maximum; its level-2 candidates only matter when the whole 4pot-block produced neither a level-0
// nor a level-1 pick (dead4), so the scan parallelises as per-lane strict maxima + a (value, scan position) arg-max.
__global__ void __launch_bounds__(128)
a2_final_kernel(const SelGeom g, const float4* __restrict__ l0, const float4* __restrict__ l1, const float4* __restrict__ l2,
                const float* __restrict__ thsS, const unsigned short* __restrict__ mask, const int* __restrict__ cand,
                const unsigned char* __restrict__ dsel, float* __restrict__ map, int* __restrict__ counts) {
    __shared__ int s_dead[4], s_i4[4];
    __shared__ float s_v4[4];
    const int i4 = blockIdx.x, b4 = i4 << 4, s3 = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int b3 = b4 + (s3 << 2);
    int xa, ya;
    const bool valid3 = slot_origin(g, b3, xa, ya);
    int i3 = -1, q3 = 0, i4b = -1, q4 = 0;
    float v3 = 0.f, v4 = 0.f;
    bool hit0 = false;
    if (valid3) {
        bool hme = false;
        if (lane < 4) {
            const int b = b3 + lane, d = dsel[b];
            if ((mask[b] >> d) & 1) { map[cand[(size_t)b * 16 + d]] = 1.f; hme = true; }
        }
        hit0 = __any_sync(0xffffffffu, hme);
        if (!hit0) {
            const float d3x = c_dirs[dsel[b3]][0], d3y = c_dirs[dsel[b3]][1];
            const float d4x = c_dirs[dsel[b4]][0], d4y = c_dirs[dsel[b4]][1];
            const int pp = g.pot * g.pot;
            for (int q = lane; q < 4 * pp; q += 32) {
                const int s2 = q / pp, r = q - s2 * pp, y1 = r / g.pot, x1 = r - y1 * g.pot;
                int x0, y0;
                if (!slot_origin(g, b3 + s2, x0, y0)) continue;
                const int xf = x0 + x1, yf = y0 + y1;
                if (xf >= g.w || yf >= g.h) continue;
                if (xf < 4 || xf >= g.w - 5 || yf < 4 || yf > g.h - 4) continue;
                const int idx = xf + g.w * yf;
                const float th0 = thsS[(xf >> 5) + (yf >> 5) * g.thsStep];
                const float th1 = th0 * g.dw1;
                const float th2 = th1 * g.dw2;
                const float4 t = __ldg(l0 + idx);
                const float ag1 = l1[(int)(xf * 0.5f + 0.25f) + (int)(yf * 0.5f + 0.25f) * g.w1].w;
                if (ag1 > th1 * g.thFactor) {
                    const float dn = g.dirDist ? fabsf(t.y * d3x + t.z * d3y) : ag1;
                    if (dn > v3) { v3 = dn; i3 = idx; q3 = q; }
                }
                const float ag2 = l2[(int)((double)(xf * 0.25f) + 0.125) + (int)((double)(yf * 0.25f) + 0.125) * g.w2].w;
                if (ag2 > th2 * g.thFactor) {
                    const float dn = g.dirDist ? fabsf(t.y * d4x + t.z * d4y) : ag2;
                    if (dn > v4) { v4 = dn; i4b = idx; q4 = q; }
                }
            }
            warp_argmax(v3, q3, i3);
            warp_argmax(v4, q4, i4b);
        }
    }
    if (lane == 0) {
        const bool has3 = i3 > 0;
        if (has3) { map[i3] = 2.f; atomicAdd(counts + 1, 1); }
        s_dead[s3] = (hit0 || has3) ? 1 : 0;
        s_v4[s3] = v4; s_i4[s3] = i4b;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        if (!(s_dead[0] | s_dead[1] | s_dead[2] | s_dead[3])) {
            float bv = 0.f;
            int bi = -1;
            for (int k = 0; k < 4; k++)
                if (s_i4[k] >= 0 && s_v4[k] > bv) { bv = s_v4[k]; bi = s_i4[k]; }
            if (bi > 0) { map[bi] = 4.f; atomicAdd(counts + 2, 1); }
        }
    }
}

// ---- thinning: row-major rank of every selected pixel ---------------------------------------------
__global__ void a2_thin_count_kernel(const float* __restrict__ map, int w, int h, int* __restrict__ row_cnt) {
    int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (row >= h) return;
    int cnt = 0;
    for (int x = lane; x < w; x += 32) cnt += map[(size_t)row * w + x] != 0.f;
    for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    if (lane == 0) row_cnt[row] = cnt;
}
__global__ void a2_thin_scan_kernel(int* __restrict__ row_cnt, int h) {   // single CTA, exclusive scan in place
    __shared__ int s[1024];
    __shared__ int carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int st = 0; st < h; st += blockDim.x) {
        int i = st + threadIdx.x;
        int v = i < h ? row_cnt[i] : 0;
        s[threadIdx.x] = v;
        __syncthreads();
        for (int o = 1; o < blockDim.x; o <<= 1) {
            int t = threadIdx.x >= o ? s[threadIdx.x - o] : 0;
            __syncthreads();
            s[threadIdx.x] += t;
            __syncthreads();
        }
        if (i < h) row_cnt[i] = carry + s[threadIdx.x] - v;
        __syncthreads();
        if (threadIdx.x == blockDim.x - 1) carry += s[threadIdx.x];
        __syncthreads();
    }
}
__global__ void a2_thin_apply_kernel(float* __restrict__ map, int w, int h, const int* __restrict__ row_off,
                                     const unsigned char* __restrict__ rp, unsigned char charTH, int* __restrict__ removed) {
    int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (row >= h) return;
    int off = row_off[row], rem = 0;
    for (int x0 = 0; x0 < w; x0 += 32) {
        int x = x0 + lane;
        bool s = x < w && map[(size_t)row * w + x] != 0.f;
        unsigned m = __ballot_sync(0xffffffffu, s);
        if (s) {
            int rn = off + __popc(m & ((1u << lane) - 1u));
            if (rp[rn] > charTH) { map[(size_t)row * w + x] = 0.f; rem++; }
        }
        off += __popc(m);
    }
    for (int o = 16; o > 0; o >>= 1) rem += __shfl_xor_sync(0xffffffffu, rem, o);
    if (lane == 0 && rem) atomicAdd(removed, rem);
}

// ---- selected pixels as a row-major list (inside the border) ----------------------------------------------
__global__ void a2_list_count_kernel(const float* __restrict__ map, int w, int h, int border, int* __restrict__ row_cnt) {
    int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (row > h) return;
    int cnt = 0;
    if (row >= border && row < h - border)
        for (int x = border + lane; x < w - border; x += 32) cnt += map[(size_t)row * w + x] != 0.f;
    for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    if (lane == 0) row_cnt[row] = cnt;      // row h: sentinel, becomes the total after the exclusive scan
}
__global__ void a2_list_emit_kernel(const float* __restrict__ map, int w, int h, int border, const int* __restrict__ row_off,
                                    float* __restrict__ list, int max_list) {
    int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (row >= h || row < border || row >= h - border) return;
    int off = row_off[row];
    for (int x0 = border; x0 < w - border; x0 += 32) {
        int x = x0 + lane;
        float m = x < w - border ? map[(size_t)row * w + x] : 0.f;
        unsigned bal = __ballot_sync(0xffffffffu, m != 0.f);
        if (m != 0.f) {
            int k = off + __popc(bal & ((1u << lane) - 1u));
            if (k < max_list) { list[3 * k] = (float)x; list[3 * k + 1] = (float)row; list[3 * k + 2] = m; }
        }
        off += __popc(bal);
    }
}

}  // namespace

// ---- host side ---------------------------------------------------------------------------------------
struct SelectorDev {
    std::vector<unsigned char> rp_host;
    unsigned char* rp = nullptr;
    float *ths = nullptr, *thsS = nullptr, *map = nullptr;
    unsigned short* mask = nullptr;
    int* cand = nullptr;
    unsigned char* dsel = nullptr;
    int *row_cnt = nullptr, *counts = nullptr;
    float* list = nullptr;   // [max_list][3] selected (u, v, type)
    int max_list = 0;
    unsigned short* mask_pinned = nullptr;
    unsigned char* dsel_pinned = nullptr;
    int* counts_pinned = nullptr;
    int max_slots = 0, ths_n = 0;
};

static int selector_init(ldso_gpu_ctx* c) {
    if (c->selector) return 0;
    SelectorDev* S = new SelectorDev();
    c->selector = S;
    const int w = c->wG[0], h = c->hG[0];
    const size_t wh = (size_t)w * h;
    // PixelSelector::PixelSelector (:9-19): srand(3141592); rand() & 0xFF.  glibc rand() is random() on the TYPE_3
    // state; random_r on a private state reproduces the sequence without touching the process-global generator.
    S->rp_host.resize(wh);
    {
        struct random_data rd;
        char state[128];
        memset(&rd, 0, sizeof(rd));
        initstate_r(3141592u, state, sizeof(state), &rd);
        for (size_t i = 0; i < wh; i++) {
            int32_t r;
            random_r(&rd, &r);
            S->rp_host[i] = (unsigned char)(r & 0xFF);
        }
    }
    S->max_slots = ((w + 3) / 4) * ((h + 3) / 4) * 16;
    S->ths_n = (w / 32 + 1) * (h / 32 + 1) + 100;
    LDSO_CUDA_OK(c, cudaMalloc(&S->rp, wh));
    LDSO_CUDA_OK(c, cudaMemcpy(S->rp, S->rp_host.data(), wh, cudaMemcpyHostToDevice));
    LDSO_CUDA_OK(c, cudaMalloc(&S->ths, S->ths_n * sizeof(float)));
    LDSO_CUDA_OK(c, cudaMalloc(&S->thsS, S->ths_n * sizeof(float)));
    LDSO_CUDA_OK(c, cudaMemset(S->ths, 0, S->ths_n * sizeof(float)));      // trap T3: defined (zero) off the 32-grid
    LDSO_CUDA_OK(c, cudaMemset(S->thsS, 0, S->ths_n * sizeof(float)));
    LDSO_CUDA_OK(c, cudaMalloc(&S->map, wh * sizeof(float)));
    LDSO_CUDA_OK(c, cudaMalloc(&S->mask, (size_t)S->max_slots * sizeof(unsigned short)));
    LDSO_CUDA_OK(c, cudaMalloc(&S->cand, (size_t)S->max_slots * 16 * sizeof(int)));
    LDSO_CUDA_OK(c, cudaMalloc(&S->dsel, (size_t)S->max_slots));
    LDSO_CUDA_OK(c, cudaMalloc(&S->row_cnt, (h + 2) * sizeof(int)));
    S->max_list = (int)std::min(wh, c->scratch_host ? wh / 2 : wh);
    LDSO_CUDA_OK(c, cudaMalloc(&S->list, (size_t)S->max_list * 3 * sizeof(float)));
    LDSO_CUDA_OK(c, cudaMalloc(&S->counts, 8 * sizeof(int)));
    LDSO_CUDA_OK(c, cudaMallocHost(&S->mask_pinned, (size_t)S->max_slots * sizeof(unsigned short)));
    LDSO_CUDA_OK(c, cudaMallocHost(&S->dsel_pinned, (size_t)S->max_slots));
    LDSO_CUDA_OK(c, cudaMallocHost(&S->counts_pinned, 8 * sizeof(int)));
    return 0;
}

void selector_free(ldso_gpu_ctx* c) {
    SelectorDev* S = (SelectorDev*)c->selector;
    if (!S) return;
    cudaFree(S->rp); cudaFree(S->ths); cudaFree(S->thsS); cudaFree(S->map); cudaFree(S->mask); cudaFree(S->cand);
    cudaFree(S->dsel); cudaFree(S->row_cnt); cudaFree(S->counts); cudaFree(S->list);
    cudaFreeHost(S->mask_pinned); cudaFreeHost(S->dsel_pinned); cudaFreeHost(S->counts_pinned);
    delete S;
    c->selector = nullptr;
}

// PixelSelector::select (:171-316): n[0..2] = n2, n3, n4; the map stays on the device
static int select_pass(ldso_gpu_ctx* c, SelectorDev* S, int slot, int pot, float thFactor, int n[3]) {
    cudaStream_t s = c->stream;
    SelGeom g;
    g.w = c->wG[0]; g.h = c->hG[0]; g.w1 = c->wG[1]; g.w2 = c->wG[2]; g.pot = pot;
    g.n4x = (g.w + 4 * pot - 1) / (4 * pot); g.n4y = (g.h + 4 * pot - 1) / (4 * pot);
    g.thsStep = g.w / 32; g.dirDist = c->cfg.select_direction_distribution != 0;
    g.thFactor = thFactor; g.dw1 = c->cfg.grad_downweight_per_level; g.dw2 = g.dw1 * g.dw1;
    const int nslots = g.n4x * g.n4y * 16;
    LDSO_REQUIRE(c, nslots <= S->max_slots);
    const float4* l0 = c->frame[slot] + c->lvl_off[0];
    const float4* l1 = c->frame[slot] + c->lvl_off[1];
    const float4* l2 = c->frame[slot] + c->lvl_off[2];
    LDSO_CUDA_OK(c, cudaMemsetAsync(S->map, 0, sizeof(float) * (size_t)g.w * g.h, s));
    LDSO_CUDA_OK(c, cudaMemsetAsync(S->counts, 0, 8 * sizeof(int), s));
    prof_begin(c, s);
    const int lpb = pot * pot >= 64 ? 32 : pot * pot >= 16 ? 8 : 1;
    a2_cand0_kernel<<<(int)(((size_t)nslots * lpb + 127) / 128), 128, 0, s>>>(g, l0, S->thsS, S->mask, S->cand, nslots, lpb);
    prof_end(c, s, LDSO_PROF_A2, (double)g.w * g.h);
    LDSO_CUDA_OK(c, ldso_copy(c, S->mask_pinned, S->mask, (size_t)nslots * sizeof(unsigned short), cudaMemcpyDeviceToHost, s));
    LDSO_CUDA_OK(c, ldso_stream_sync(s));
    // the ordered scan: which direction each pot-block sees, and the running level-0 count
    int n2 = 0;
    const unsigned char* rp = S->rp_host.data();
    for (int b = 0; b < nslots; b++) {
        const int d = rp[n2] & 0xF;
        S->dsel_pinned[b] = (unsigned char)d;
        n2 += (S->mask_pinned[b] >> d) & 1;
    }
    LDSO_CUDA_OK(c, ldso_copy(c, S->dsel, S->dsel_pinned, nslots, cudaMemcpyHostToDevice, s));
    const int n4blocks = g.n4x * g.n4y;
    prof_begin(c, s);
    a2_final_kernel<<<n4blocks, 128, 0, s>>>(g, l0, l1, l2, S->thsS, S->mask, S->cand, S->dsel, S->map, S->counts);
    prof_end(c, s, LDSO_PROF_A2, 0);
    c->launches += 2;
    LDSO_CUDA_OK(c, ldso_copy(c, S->counts_pinned, S->counts, 8 * sizeof(int), cudaMemcpyDeviceToHost, s));
    LDSO_CUDA_OK(c, ldso_stream_sync(s));
    n[0] = n2; n[1] = S->counts_pinned[1]; n[2] = S->counts_pinned[2];
    return 0;
}

// makeMaps with the map left on the device (S->map)
// exact_count = false: *n_selected is an upper bound (the pixels the random thinning removes are not counted: no read-back)
static int make_maps_device(ldso_gpu_ctx* c, int slot, float density, int recursions_left, float th_factor, int* potential_io,
                            int* n_selected, bool exact_count = true) {
    LDSO_REQUIRE(c, slot >= 0 && slot < c->cfg.n_frame_slots && potential_io && n_selected);
    LDSO_REQUIRE(c, c->levels >= 3 && *potential_io >= 1);
    int rc = selector_init(c);
    if (rc) return rc;
    SelectorDev* S = (SelectorDev*)c->selector;
    cudaStream_t s = c->stream;
    const int w = c->wG[0], h = c->hG[0], w32 = w / 32, h32 = h / 32;
    // makeHists (:36-109)
    if (w32 * h32 > 0) {
        a2_hist_kernel<<<w32 * h32, 256, 0, s>>>(c->frame[slot], w, h, w32, c->cfg.min_grad_hist_cut, c->cfg.min_grad_hist_add, S->ths);
        a2_smooth_kernel<<<(w32 * h32 + 127) / 128, 128, 0, s>>>(S->ths, S->thsS, w32, h32);
        c->launches += 2;
    }
    // makeMaps (:111-169)
    int pot = *potential_io, n[3];
    float numHave, quotia;
    int ideal;
    for (;;) {
        if ((rc = select_pass(c, S, slot, pot, th_factor, n))) return rc;
        numHave = (float)(n[0] + n[1] + n[2]);
        quotia = density / numHave;
        float K = numHave * (pot + 1) * (pot + 1);
        ideal = (int)(sqrtf(K / density) - 1);
        if (ideal < 1) ideal = 1;
        if (recursions_left > 0 && quotia > 1.25 && pot > 1) {
            if (ideal >= pot) ideal = pot - 1;
            pot = ideal; recursions_left--;
            continue;
        }
        if (recursions_left > 0 && quotia < 0.25) {
            if (ideal <= pot) ideal = pot + 1;
            pot = ideal; recursions_left--;
            continue;
        }
        break;
    }
    int numHaveSub = (int)numHave;
    if (quotia < 0.95) {
        unsigned char charTH = (unsigned char)(255 * quotia);
        a2_thin_count_kernel<<<(h + 7) / 8, 256, 0, s>>>(S->map, w, h, S->row_cnt);
        a2_thin_scan_kernel<<<1, 1024, 0, s>>>(S->row_cnt, h);
        a2_thin_apply_kernel<<<(h + 7) / 8, 256, 0, s>>>(S->map, w, h, S->row_cnt, S->rp, charTH, S->counts + 4);
        c->launches += 3;
        if (exact_count) {
            LDSO_CUDA_OK(c, ldso_copy(c, S->counts_pinned, S->counts, 8 * sizeof(int), cudaMemcpyDeviceToHost, s));
            LDSO_CUDA_OK(c, ldso_stream_sync(s));
            numHaveSub -= S->counts_pinned[4];
        }
    }
    *potential_io = ideal;
    *n_selected = numHaveSub;
    LDSO_CUDA_OK(c, cudaGetLastError());
    return 0;
}

extern "C" int ldso_gpu_select_make_maps(ldso_gpu_ctx* c, int slot, float density, int recursions_left, float th_factor,
                                         int* potential_io, float* map_out, int* n_selected) {
    LDSO_ENTER(c);
    LDSO_REQUIRE(c, map_out != nullptr);
    int rc = make_maps_device(c, slot, density, recursions_left, th_factor, potential_io, n_selected);
    if (rc) return rc;
    SelectorDev* S = (SelectorDev*)c->selector;
    const int w = c->wG[0], h = c->hG[0];
    float* hmap = (float*)c->scratch_host;
    LDSO_CUDA_OK(c, ldso_copy(c, hmap, S->map, sizeof(float) * (size_t)w * h, cudaMemcpyDeviceToHost, c->stream));
    LDSO_CUDA_OK(c, ldso_stream_sync(c->stream));
    memcpy(map_out, hmap, sizeof(float) * (size_t)w * h);
    return 0;
}

// makeMaps + the pixel loop of FullSystem::makeNewTraces (src/frontend/FullSystem.cc:1440-1460): the selected pixels
// inside the border, in row-major order, as (u, v, type) lists — 12 bytes per selected point cross the bus instead of
// the whole map
extern "C" int ldso_gpu_select_points(ldso_gpu_ctx* c, int slot, float density, int recursions_left, float th_factor,
                                      int* potential_io, int border, int max_points, float* u_out, float* v_out,
                                      float* type_out, int* n_points) {
    LDSO_ENTER(c);
    LDSO_REQUIRE(c, u_out && v_out && type_out && n_points && border >= 0 && max_points >= 0);
    int nsel = 0;
    int rc = make_maps_device(c, slot, density, recursions_left, th_factor, potential_io, &nsel, false);   // the list below carries its own length
    if (rc) return rc;
    SelectorDev* S = (SelectorDev*)c->selector;
    cudaStream_t s = c->stream;
    const int w = c->wG[0], h = c->hG[0];
    a2_list_count_kernel<<<(h + 8) / 8, 256, 0, s>>>(S->map, w, h, border, S->row_cnt);
    a2_thin_scan_kernel<<<1, 1024, 0, s>>>(S->row_cnt, h + 1);
    a2_list_emit_kernel<<<(h + 7) / 8, 256, 0, s>>>(S->map, w, h, border, S->row_cnt, S->list, S->max_list);
    c->launches += 3;
    LDSO_CUDA_OK(c, cudaGetLastError());
    // the list holds at most the nsel selected pixels: its length and that many entries come back in ONE round trip
    int* total = (int*)S->counts_pinned;
    const int bound = std::min(std::min(nsel, max_points), S->max_list);
    float* hl = (float*)c->scratch_host;
    LDSO_CUDA_OK(c, ldso_copy(c, total, S->row_cnt + h, sizeof(int), cudaMemcpyDeviceToHost, s));
    if (bound > 0) LDSO_CUDA_OK(c, ldso_copy(c, hl, S->list, sizeof(float) * 3 * (size_t)bound, cudaMemcpyDeviceToHost, s));
    LDSO_CUDA_OK(c, ldso_stream_sync(s));
    int n = std::min(*total, bound);
    *n_points = n;
    if (n == 0) return 0;
    for (int i = 0; i < n; i++) { u_out[i] = hl[3 * i]; v_out[i] = hl[3 * i + 1]; type_out[i] = hl[3 * i + 2]; }
    return 0;
}
