// k2_coarse_depth.cu — K2: CoarseTracker::makeCoarseDepthL0 (reference src/frontend/CoarseTracker.cc:357-538).
//
// Stages (all on the ctx's tracking stream, no host round trip until the final pc_n read-back):
//   scatter   idepth*w and w of every qualifying point into the level-0 maps            (:363-382)
//   pool      2x2 SUM up the pyramid, (a+b)+c+d association as in the reference          (:384-409)
//   dilate    fill empty cells from 4 diagonal (levels 0,1) or axis (levels >=2) nbrs,
//             reading the pre-dilation weight copy so the result is order-independent    (:412-495)
//   compact   normalise and emit interior cells with weight>0 in ROW-MAJOR order into a
//             float4 {u,v,idepth,color} point cloud per level                             (:498-537)
// pc_n[] and the (u,v) sequence are bit-exact against the reference; idepth sums can differ in the last
// bit only where >=3 points scatter into one pixel (atomicAdd order).
#include "ctx.h"

namespace {

__global__ void k2_scatter_kernel(int n, const float* __restrict__ u, const float* __restrict__ v, const float* __restrict__ id,
                                  const float* __restrict__ hdi, float* __restrict__ idepth0, float* __restrict__ wsum0, int w,
                                  int h) {
    int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    int x = (int)(u[k] + 0.5f);
    int y = (int)(v[k] + 0.5f);
    if (x < 0 || y < 0 || x >= w || y >= h) return;   // the reference would write out of bounds here
    float weight = sqrtf((float)(1e-3 / ((double)hdi[k] + 1e-12)));
    atomicAdd(idepth0 + x + w * y, id[k] * weight);
    atomicAdd(wsum0 + x + w * y, weight);
}

__global__ void k2_pool_kernel(const float* __restrict__ id_lm, const float* __restrict__ ws_lm, float* __restrict__ id_l,
                               float* __restrict__ ws_l, int wl, int hl, int wlm1) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= wl * hl) return;
    int y = i / wl, x = i - y * wl;
    int bidx = 2 * x + 2 * y * wlm1;
    id_l[i] = __fadd_rn(__fadd_rn(__fadd_rn(id_lm[bidx], id_lm[bidx + 1]), id_lm[bidx + wlm1]), id_lm[bidx + wlm1 + 1]);
    ws_l[i] = __fadd_rn(__fadd_rn(__fadd_rn(ws_lm[bidx], ws_lm[bidx + 1]), ws_lm[bidx + wlm1]), ws_lm[bidx + wlm1 + 1]);
}

// all levels in one launch (blockIdx.y = level): the planes of different levels do not depend on each other here
struct K2Planes {
    int levels;
    int w[LDSO_MAX_LEVELS], h[LDSO_MAX_LEVELS];
    float* idepth[LDSO_MAX_LEVELS];
    float* ws[LDSO_MAX_LEVELS];
    float* ws_bak[LDSO_MAX_LEVELS];
};
__global__ void k2_backup_all_kernel(const K2Planes p) {
    const int l = blockIdx.y, i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < p.w[l] * p.h[l]) p.ws_bak[l][i] = p.ws[l][i];
}
// reads ws_bak (pre-dilation copy), writes idepth/ws only where ws_bak<=0, reads idepth only where ws_bak>0
__global__ void k2_dilate_all_kernel(const K2Planes p) {
    const int l = blockIdx.y;
    float* __restrict__ idepth = p.idepth[l];
    float* __restrict__ ws = p.ws[l];
    const float* __restrict__ ws_bak = p.ws_bak[l];
    const int wl = p.w[l], hl = p.h[l], diagonal = l < 2 ? 1 : 0;
    int i = blockIdx.x * blockDim.x + threadIdx.x + wl;
    if (i >= wl * hl - wl) return;
    if (ws_bak[i] > 0) return;
    int o0, o1, o2, o3;
    if (diagonal) { o0 = 1 + wl; o1 = -1 - wl; o2 = wl - 1; o3 = -wl + 1; }
    else { o0 = 1; o1 = -1; o2 = wl; o3 = -wl; }
    float sum = 0, num = 0, numn = 0;
    float b;
    const int n = wl * hl;
    // the diagonal neighbours of the first cell of row 1 and of the last cell of row h-2 fall one element outside the
    // plane (the reference reads past its arrays there, CoarseTracker.cc:437-466; those two cells never reach the point
    // cloud): a neighbour outside [0, w h) counts as weight 0
    auto wgt = [&](int j) { return (j >= 0 && j < n) ? ws_bak[j] : 0.f; };
    b = wgt(i + o0); if (b > 0) { sum = __fadd_rn(sum, idepth[i + o0]); num = __fadd_rn(num, b); numn++; }
    b = wgt(i + o1); if (b > 0) { sum = __fadd_rn(sum, idepth[i + o1]); num = __fadd_rn(num, b); numn++; }
    b = wgt(i + o2); if (b > 0) { sum = __fadd_rn(sum, idepth[i + o2]); num = __fadd_rn(num, b); numn++; }
    b = wgt(i + o3); if (b > 0) { sum = __fadd_rn(sum, idepth[i + o3]); num = __fadd_rn(num, b); numn++; }
    if (numn > 0) {
        idepth[i] = __fdiv_rn(sum, numn);
        ws[i] = __fdiv_rn(num, numn);
    }
}

struct K2Levels {
    int levels;
    int w[LDSO_MAX_LEVELS], h[LDSO_MAX_LEVELS];
    int row_base[LDSO_MAX_LEVELS + 1];   // first compaction row of each level in the flat row list
    float* idepth[LDSO_MAX_LEVELS];
    float* ws[LDSO_MAX_LEVELS];
    float4* pc[LDSO_MAX_LEVELS];
    const float4* ref[LDSO_MAX_LEVELS];
};

__device__ __forceinline__ bool k2_valid_values(float wsum, float idepth, float color, float& nid) {
    if (!(wsum > 0)) return false;
    nid = __fdiv_rn(idepth, wsum);
    return isfinite(color) && (nid > 0);
}
// one warp per interior row (y in [2,h-2)) of every level: count
__global__ void k2_count_rows_kernel(const K2Levels p, int* __restrict__ row_cnt) {
    int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    int lane = threadIdx.x & 31;
    if (row >= p.row_base[p.levels]) return;
    int l = 0;
    while (row >= p.row_base[l + 1]) l++;
    int y = row - p.row_base[l] + 2;
    int wl = p.w[l];
    int cnt = 0;
    const float* __restrict__ ws = p.ws[l];
    const float* __restrict__ idp = p.idepth[l];
    const float4* __restrict__ ref = p.ref[l];
    for (int x0 = 2; x0 < wl - 2; x0 += 128) {          // four loads of every plane in flight per lane
        float w4[4], d4[4], c4[4];
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const int x = x0 + 32 * k + lane;
            w4[k] = 0.f; d4[k] = 0.f; c4[k] = 0.f;
            if (x < wl - 2) { const int i = x + y * wl; w4[k] = ws[i]; d4[k] = idp[i]; c4[k] = ref[i].x; }
        }
#pragma unroll
        for (int k = 0; k < 4; k++) {
            float nid;
            cnt += k2_valid_values(w4[k], d4[k], c4[k], nid) ? 1 : 0;
        }
    }
    for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    if (lane == 0) row_cnt[row] = cnt;
}

// one CTA per level: exclusive scan of that level's row counts (in place), total -> pc_n[l]
__global__ void k2_scan_rows_kernel(const K2Levels p, int* __restrict__ row_cnt, int* __restrict__ pc_n) {
    __shared__ int warp_tot[32];
    __shared__ int carry;
    int l = blockIdx.x;
    int base = p.row_base[l], n = p.row_base[l + 1] - base;
    int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int start = 0; start < n; start += blockDim.x) {
        int i = start + threadIdx.x;
        int v = i < n ? row_cnt[base + i] : 0;
        int incl = v;
        for (int o = 1; o < 32; o <<= 1) {
            int t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) warp_tot[wid] = incl;
        __syncthreads();
        if (wid == 0) {
            int t = lane < (blockDim.x >> 5) ? warp_tot[lane] : 0;
            int ti = t;
            for (int o = 1; o < 32; o <<= 1) {
                int s = __shfl_up_sync(0xffffffffu, ti, o);
                if (lane >= o) ti += s;
            }
            warp_tot[lane] = ti - t;   // exclusive
        }
        __syncthreads();
        int excl = carry + warp_tot[wid] + incl - v;
        if (i < n) row_cnt[base + i] = excl;
        __syncthreads();
        if (threadIdx.x == blockDim.x - 1) carry = excl + v;
        __syncthreads();
    }
    if (threadIdx.x == 0) pc_n[l] = carry;
}

// one warp per row: ordered emit (CoarseTracker.cc:513-533), also normalises the idepth map
__global__ void k2_emit_rows_kernel(const K2Levels p, const int* __restrict__ row_off) {
    int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    int lane = threadIdx.x & 31;
    if (row >= p.row_base[p.levels]) return;
    int l = 0;
    while (row >= p.row_base[l + 1]) l++;
    int y = row - p.row_base[l] + 2;
    int wl = p.w[l];
    int off = row_off[row];
    float4* pc = p.pc[l];
    const float* __restrict__ ws = p.ws[l];
    float* idp = p.idepth[l];
    const float4* __restrict__ ref = p.ref[l];
    for (int x0 = 2; x0 < wl - 2; x0 += 128) {          // four loads of every plane in flight per lane; a cell is read before it is rewritten
        float w4[4], d4[4], c4[4];
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const int x = x0 + 32 * k + lane;
            w4[k] = 0.f; d4[k] = 0.f; c4[k] = 0.f;
            if (x < wl - 2) { const int i = x + y * wl; w4[k] = ws[i]; d4[k] = idp[i]; c4[k] = ref[i].x; }
        }
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const int x = x0 + 32 * k + lane;
            bool ok = false;
            float nid = 0;
            if (x < wl - 2) {
                ok = k2_valid_values(w4[k], d4[k], c4[k], nid);
                idp[x + y * wl] = ok ? nid : -1.f;
            }
            const unsigned m = __ballot_sync(0xffffffffu, ok);
            if (ok) pc[off + __popc(m & ((1u << lane) - 1u))] = make_float4((float)x, (float)y, nid, c4[k]);
            off += __popc(m);
        }
    }
}

}  // namespace

int k2_set_ref(ldso_gpu_ctx* c, int tracker, int ref_slot, int n, const float* u, const float* v, const float* id,
               const float* hdi) {
    TrackerDev& T = c->trk[tracker];
    cudaStream_t s = c->stream;
    const int L = c->levels;
    // upload the point list (4 arrays) through the ctx's staging ring: queued, nothing waits
    float* dp = (float*)c->img_dev;   // reuse of the staged-image buffer: n <= w*h/4 guaranteed by max_points check
    if (n > 0) {
        float* hp = (float*)ldso_staging(c, 4 * (size_t)n * sizeof(float));
        LDSO_REQUIRE(c, hp != nullptr);
        memcpy(hp, u, n * sizeof(float));
        memcpy(hp + n, v, n * sizeof(float));
        memcpy(hp + 2 * n, id, n * sizeof(float));
        memcpy(hp + 3 * n, hdi, n * sizeof(float));
        LDSO_CUDA_OK(c, ldso_copy(c, dp, hp, 4 * (size_t)n * sizeof(float), cudaMemcpyHostToDevice, s));
    }
    prof_begin(c, s);
    size_t n0 = (size_t)T.w[0] * T.h[0];
    LDSO_CUDA_OK(c, cudaMemsetAsync(T.idepth[0], 0, n0 * sizeof(float), s));
    LDSO_CUDA_OK(c, cudaMemsetAsync(T.wsum[0], 0, n0 * sizeof(float), s));
    if (n > 0) {
        k2_scatter_kernel<<<(n + 255) / 256, 256, 0, s>>>(n, dp, dp + n, dp + 2 * n, dp + 3 * n, T.idepth[0], T.wsum[0], T.w[0],
                                                          T.h[0]);
        c->launches++;
    }
    for (int l = 1; l < L; l++) {
        int nl = T.w[l] * T.h[l];
        k2_pool_kernel<<<(nl + 255) / 256, 256, 0, s>>>(T.idepth[l - 1], T.wsum[l - 1], T.idepth[l], T.wsum[l], T.w[l], T.h[l],
                                                        T.w[l - 1]);
        c->launches++;
    }
    {
        // weightSums_bak = weightSums, then the dilation (CoarseTracker.cc:425-470): two launches over all levels
        K2Planes q;
        q.levels = L;
        for (int l = 0; l < L; l++) { q.w[l] = T.w[l]; q.h[l] = T.h[l]; q.idepth[l] = T.idepth[l]; q.ws[l] = T.wsum[l]; q.ws_bak[l] = T.wsum_bak[l]; }
        const dim3 grid((unsigned)((n0 + 255) / 256), (unsigned)L);
        k2_backup_all_kernel<<<grid, 256, 0, s>>>(q);
        k2_dilate_all_kernel<<<grid, 256, 0, s>>>(q);
        c->launches += 2;
    }
    K2Levels p;
    p.levels = L;
    for (int l = 0; l < L; l++) {
        p.w[l] = T.w[l]; p.h[l] = T.h[l];
        p.row_base[l] = T.row_base[l];
        p.idepth[l] = T.idepth[l]; p.ws[l] = T.wsum[l]; p.pc[l] = T.pc[l];
        p.ref[l] = c->frame[ref_slot] + c->lvl_off[l];
    }
    p.row_base[L] = T.row_base[L];
    int rows = T.row_base[L];
    k2_count_rows_kernel<<<(rows + 7) / 8, 256, 0, s>>>(p, T.row_cnt);
    k2_scan_rows_kernel<<<L, 256, 0, s>>>(p, T.row_cnt, T.pc_n_dev);
    k2_emit_rows_kernel<<<(rows + 7) / 8, 256, 0, s>>>(p, T.row_cnt);
    c->launches += 3;
    prof_end(c, s, LDSO_PROF_K2, n);
    LDSO_CUDA_OK(c, cudaGetLastError());
    // the point counts come back without a wait: the cluster tracker reads them on the device (LmLevel::n_dev), a host reader
    // (calc_res, get_pc, the other LM modes) waits for the event first (trk_pc_n_sync)
    LDSO_CUDA_OK(c, ldso_copy(c, T.pc_n_pinned, T.pc_n_dev, LDSO_MAX_LEVELS * sizeof(int), cudaMemcpyDeviceToHost, s));
    LDSO_CUDA_OK(c, cudaEventRecord(T.ev_pc_n, s));
    T.pc_n_valid = false;
    T.ref_slot = ref_slot;
    T.last_lvl = -1;
    return 0;
}

int trk_pc_n_sync(ldso_gpu_ctx* c, TrackerDev& T) {
    if (T.pc_n_valid) return 0;
    LDSO_CUDA_OK(c, ldso_event_sync(T.ev_pc_n));
    for (int l = 0; l < LDSO_MAX_LEVELS; l++) T.pc_n[l] = T.pc_n_pinned[l];
    T.pc_n_valid = true;
    return 0;
}
