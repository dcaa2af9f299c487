// n2_distmap.cu — N2: CoarseDistanceMap::makeDistanceMap / growDistBFS / addIntoDistFinal (reference
// src/frontend/CoarseTracker.cc:795-930) and the candidate walk of FullSystem::activatePointsMT
// (src/frontend/FullSystem.cc:1226-1298).
//
// The level-1 grid (w/2 x h/2: 76.8 k cells at 640x480) is held as ONE BYTE per cell in the shared memory of a single
// CTA: ring number 0..39, 255 standing for the reference's 1000.f.
//   * makeDistanceMap: the reference keeps explicit BFS lists; from a map that holds only 0 and 1000 the frontier of
//     ring k is exactly the set of cells with value k - 1, so ring k is a stencil pass "value > k and a non-border
//     neighbour (8-connected for odd k, 4-connected for even k) holds k - 1 -> k".  Writes only turn values > k into k
//     and reads only test for k - 1, so the pass runs in place.  39 passes, 1024 threads, no global traffic.
//   * the candidate walk is order dependent (every accepted point lowers the map for the ones after it) and stays
//     sequential in pool order, but only over ACCEPTED points: all threads classify the pool in parallel first, then
//     one warp tests 32 points at a time against the current map, accepts the first that passes, grows its rings with
//     an explicit frontier (byte claims through atomicCAS on the containing word) and re-tests the lanes behind it.
//     A rejected point can never become acceptable later (the map only decreases), so skipping them is exact.
#include "immature.h"

namespace {

constexpr int N2_THREADS = 1024;
constexpr int N2_FCAP = 512;        // ring k of one source has at most 8k <= 312 cells

__device__ __forceinline__ void project(const float* __restrict__ tab, int h, float u, float v, float idepth, float& px, float& py, float& p0) {
    const float* K = tab + 12 * h;
    p0 = (K[0] * u + K[1] * v + K[2] * 1.f) + K[9] * idepth;
    const float p1 = (K[3] * u + K[4] * v + K[5] * 1.f) + K[10] * idepth;
    const float p2 = (K[6] * u + K[7] * v + K[8] * 1.f) + K[11] * idepth;
    px = p0 / p2; py = p1 / p2;
}

__device__ __forceinline__ void fill_map(unsigned char* D, int cells16) {
    uint4* d4 = reinterpret_cast<uint4*>(D);
    for (int i = threadIdx.x; i < cells16; i += blockDim.x) d4[i] = make_uint4(~0u, ~0u, ~0u, ~0u);
}

__global__ void __launch_bounds__(N2_THREADS)
n2_distmap_kernel(int w1, int h1, int n_points, const float* __restrict__ tab, const int* __restrict__ host,
                  const float* __restrict__ u, const float* __restrict__ v, const float* __restrict__ idepth,
                  unsigned char* __restrict__ dmap) {
    extern __shared__ __align__(16) unsigned char D[];
    const int cells = w1 * h1, cells16 = (cells + 15) / 16;
    fill_map(D, cells16);
    __syncthreads();
    for (int i = threadIdx.x; i < n_points; i += N2_THREADS) {
        const int h = host[i];
        if (h < 0) continue;
        float px, py, p0;
        project(tab, h, u[i], v[i], idepth[i], px, py, p0);
        const int x = (int)(px + 0.5f), y = (int)(py + 0.5f);
        if (x > 0 && y > 0 && x < w1 && y < h1) D[x + w1 * y] = 0;
    }
    __syncthreads();
    for (int k = 1; k < 40; k++) {
        const int want = k - 1;
        const bool diag = (k & 1) != 0;
        int changed = 0;
        for (int i = threadIdx.x; i < cells; i += N2_THREADS) {
            if (D[i] <= k) continue;
            const int y = i / w1, x = i - y * w1;
            bool hit = false;
            // neighbour q propagates when it is not a border cell and holds k - 1
#pragma unroll
            for (int dy = -1; dy <= 1; dy++)
#pragma unroll
                for (int dx = -1; dx <= 1; dx++) {
                    if ((dx == 0 && dy == 0) || (!diag && dx != 0 && dy != 0)) continue;
                    const int qx = x + dx, qy = y + dy;
                    if (qx <= 0 || qy <= 0 || qx >= w1 - 1 || qy >= h1 - 1) continue;
                    hit = hit || D[qx + w1 * qy] == want;
                }
            if (hit) { D[i] = (unsigned char)k; changed = 1; }
        }
        if (!__syncthreads_or(changed)) break;
    }
    uint4* o4 = reinterpret_cast<uint4*>(dmap);
    const uint4* d4 = reinterpret_cast<const uint4*>(D);
    for (int i = threadIdx.x; i < cells16; i += N2_THREADS) o4[i] = d4[i];
}

__device__ __forceinline__ bool claim(unsigned* Dw, int idx, unsigned k) {
    unsigned* w = Dw + (idx >> 2);
    const int sh = (idx & 3) * 8;
    unsigned old = *reinterpret_cast<volatile unsigned*>(w);
    while (true) {
        if (((old >> sh) & 255u) <= k) return false;
        const unsigned nw = (old & ~(255u << sh)) | (k << sh);
        const unsigned prev = atomicCAS(w, old, nw);
        if (prev == old) return true;
        old = prev;
    }
}

struct N2Sel {
    int w1, h1, n, n_hosts;
    const float* tab;                 // [n_hosts][12] KRKi | Kt, then n_hosts skip bytes and n_hosts flag bytes (as floats)
    const int *host, *status;
    const float *u, *v, *idmin, *idmax, *quality, *interval, *type;
    float minActDist, minQuality;
    int* cell; float* frac; float* thr;   // scratch [n]
    unsigned char* action;            // [n]
    int* cnt;                         // [2]
    unsigned char* dmap;
};

__global__ void __launch_bounds__(N2_THREADS) n2_select_kernel(const N2Sel p) {
    extern __shared__ __align__(16) unsigned char D[];
    __shared__ short2 fl[2][N2_FCAP];
    __shared__ int fcnt, s_acc, s_del;
    const int cells = p.w1 * p.h1, cells16 = (cells + 15) / 16;
    {
        uint4* d4 = reinterpret_cast<uint4*>(D);
        const uint4* g4 = reinterpret_cast<const uint4*>(p.dmap);
        for (int i = threadIdx.x; i < cells16; i += N2_THREADS) d4[i] = g4[i];
    }
    if (threadIdx.x == 0) { s_acc = 0; s_del = 0; }
    __syncthreads();
    // ---- classification, all threads (FullSystem.cc:1243-1296 without the map test)
    const float* skip = p.tab + 12 * p.n_hosts;
    const float* flagged = skip + p.n_hosts;
    int ndel = 0;
    for (int i = threadIdx.x; i < p.n; i += N2_THREADS) {
        const int h = p.host[i];
        int act = 0, cell = -1;
        float fr = 0.f, th = 0.f;
        if (skip[h] == 0.f) {
            const int st = p.status[i];
            const float imax = p.idmax[i], imin = p.idmin[i];
            if (!isfinite(imax) || st == 2) {
                act = 1;
            } else {
                const bool can = (st == 0 || st == 3 || st == 4 || st == 1) && p.interval[i] < 8 && p.quality[i] > p.minQuality &&
                                 (imax + imin) > 0;
                if (!can) {
                    if (flagged[h] != 0.f || st == 1) act = 1;
                } else {
                    float px, py, p0;
                    project(p.tab, h, p.u[i], p.v[i], 0.5f * (imax + imin), px, py, p0);
                    const int x = (int)(px + 0.5f), y = (int)(py + 0.5f);
                    if (x > 0 && y > 0 && x < p.w1 && y < p.h1) {
                        cell = x + p.w1 * y;
                        fr = p0 - floorf(p0);   // the reference takes the fraction of the homogeneous ptp[0] (FullSystem.cc:1279-1280)
                        th = p.minActDist * p.type[i];
                    } else {
                        act = 1;
                    }
                }
            }
        }
        p.action[i] = (unsigned char)act;
        p.cell[i] = cell; p.frac[i] = fr; p.thr[i] = th;
        ndel += act;
    }
    if (ndel) atomicAdd(&s_del, ndel);
    __syncthreads();
    // ---- the ordered walk, warp 0
    if (threadIdx.x < 32) {
        const int lane = threadIdx.x;
        unsigned* Dw = reinterpret_cast<unsigned*>(D);
        int nacc = 0;
        for (int base = 0; base < p.n; base += 32) {
            const int i = base + lane;
            const int cell = i < p.n ? p.cell[i] : -1;
            const float fr = i < p.n ? p.frac[i] : 0.f, th = i < p.n ? p.thr[i] : 0.f;
            int from = 0;
            while (true) {
                bool pass = false;
                if (cell >= 0 && lane >= from) {
                    const unsigned char d = D[cell];
                    const float dist = (d == 255 ? 1000.f : (float)d) + fr;
                    pass = dist >= th;
                }
                const unsigned m = __ballot_sync(0xffffffffu, pass);
                if (!m) break;
                const int j = __ffs(m) - 1;
                const int c0 = __shfl_sync(0xffffffffu, cell, j);
                if (lane == j) p.action[i] = 2;
                nacc++;
                from = j + 1;
                // addIntoDistFinal + growDistBFS(1)
                const int sy = c0 / p.w1, sx = c0 - sy * p.w1;
                if (lane == 0) { D[c0] = 0; fl[0][0] = make_short2((short)sx, (short)sy); }
                int cur = 0, nf = 1;
                __syncwarp();
                for (int k = 1; k < 40 && nf > 0; k++) {
                    if (lane == 0) fcnt = 0;
                    __syncwarp();
                    const bool diag = (k & 1) != 0;
                    for (int it = lane; it < nf; it += 32) {
                        const short2 q = fl[cur][it];
                        const int x = q.x, y = q.y;
                        if (x == 0 || y == 0 || x == p.w1 - 1 || y == p.h1 - 1) continue;
#pragma unroll
                        for (int dy = -1; dy <= 1; dy++)
#pragma unroll
                            for (int dx = -1; dx <= 1; dx++) {
                                if ((dx == 0 && dy == 0) || (!diag && dx != 0 && dy != 0)) continue;
                                const int nx = x + dx, ny = y + dy;
                                if (claim(Dw, nx + p.w1 * ny, (unsigned)k)) {
                                    const int pos = atomicAdd(&fcnt, 1);
                                    if (pos < N2_FCAP) fl[cur ^ 1][pos] = make_short2((short)nx, (short)ny);
                                }
                            }
                    }
                    __syncwarp();
                    nf = min(*reinterpret_cast<volatile int*>(&fcnt), N2_FCAP);
                    cur ^= 1;
                    __syncwarp();
                }
                if (from >= 32) break;
            }
        }
        if (lane == 0) s_acc = nacc;
    }
    __syncthreads();
    if (threadIdx.x == 0) { p.cnt[0] = s_acc; p.cnt[1] = s_del; }
    uint4* o4 = reinterpret_cast<uint4*>(p.dmap);
    const uint4* d4 = reinterpret_cast<const uint4*>(D);
    for (int i = threadIdx.x; i < cells16; i += N2_THREADS) o4[i] = d4[i];
}

__global__ void n2_types_kernel(int n, const float* __restrict__ src, float* __restrict__ dst) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) dst[i] = src[i];
}

size_t map_bytes(const ldso_gpu_ctx* c) { return (((size_t)c->wG[1] * c->hG[1] + 15) / 16) * 16; }

int set_smem(ldso_gpu_ctx* c, const void* fn, size_t bytes) {
    LDSO_REQUIRE(c, bytes <= 200 * 1024);   // level-1 grids up to ~200 k cells (1232x368 -> 113 k)
    LDSO_CUDA_OK(c, cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    return 0;
}

// per-host tables -> device (P->tables is sized for K5's 16 floats per host: 12 + skip + flag fit)
int put_tables(ldso_gpu_ctx* c, ImmatureDev* P, int n_hosts, const float* KRKi, const float* Kt, const unsigned char* skip,
               const unsigned char* flagged, float* st) {
    for (int h = 0; h < n_hosts; h++) {
        for (int k = 0; k < 9; k++) st[12 * h + k] = KRKi[9 * h + k];
        for (int k = 0; k < 3; k++) st[12 * h + 9 + k] = Kt[3 * h + k];
        st[12 * n_hosts + h] = skip && skip[h] ? 1.f : 0.f;
        st[13 * n_hosts + h] = flagged && flagged[h] ? 1.f : 0.f;
    }
    LDSO_CUDA_OK(c, ldso_copy(c, P->tables, st, sizeof(float) * 14 * n_hosts, cudaMemcpyHostToDevice, c->stream));
    return 0;
}

}  // namespace

extern "C" int ldso_gpu_immature_set_types(ldso_gpu_ctx* c, int first, int n, const float* type) {
    ImmatureDev* P = immature_pool(c);
    if (!P) return -1;
    LDSO_REQUIRE(c, first >= 0 && n >= 0 && first + n <= P->n && (n == 0 || type));
    if (n == 0) return 0;
    float* st = (float*)P->pinned;
    memcpy(st, type, sizeof(float) * n);
    LDSO_CUDA_OK(c, ldso_copy(c, P->cr, st, sizeof(float) * n, cudaMemcpyHostToDevice, c->stream));
    n2_types_kernel<<<(n + 255) / 256, 256, 0, c->stream>>>(n, P->cr, P->type + first);
    c->launches++;
    LDSO_CUDA_OK(c, cudaGetLastError());
    LDSO_CUDA_OK(c, cudaStreamSynchronize(c->stream));   // the pinned staging is reused by the next entry point
    return 0;
}

extern "C" int ldso_gpu_distmap_make(ldso_gpu_ctx* c, int n_hosts, const float* KRKi, const float* Kt, int n_points,
                                     const int* host, const float* u, const float* v, const float* idepth) {
    ImmatureDev* P = immature_pool(c);
    if (!P) return -1;
    LDSO_REQUIRE(c, n_hosts >= 1 && n_hosts <= LDSO_MAX_WINDOW && KRKi && Kt && n_points >= 0 && c->levels >= 2);
    LDSO_REQUIRE(c, n_points == 0 || (host && u && v && idepth));
    LDSO_REQUIRE(c, (size_t)n_points * 4 <= (size_t)22 * P->cap && sizeof(float) * (4 * (size_t)n_points + 16 * LDSO_MAX_WINDOW) <= P->pinned_bytes);
    for (int i = 0; i < n_points; i++) LDSO_REQUIRE(c, host[i] < n_hosts);
    float* st = (float*)P->pinned;
    int rc = put_tables(c, P, n_hosts, KRKi, Kt, nullptr, nullptr, st);
    if (rc) return rc;
    float* sp = st + 16 * LDSO_MAX_WINDOW;
    const size_t N = n_points;
    if (N) {
        memcpy(sp, host, 4 * N); memcpy(sp + N, u, 4 * N); memcpy(sp + 2 * N, v, 4 * N); memcpy(sp + 3 * N, idepth, 4 * N);
        LDSO_CUDA_OK(c, ldso_copy(c, P->cr, sp, 16 * N, cudaMemcpyHostToDevice, c->stream));
    }
    const size_t smem = map_bytes(c);
    rc = set_smem(c, (const void*)n2_distmap_kernel, smem);
    if (rc) return rc;
    prof_begin(c, c->stream);
    n2_distmap_kernel<<<1, N2_THREADS, smem, c->stream>>>(c->wG[1], c->hG[1], n_points, P->tables, (const int*)P->cr, P->cr + N,
                                                          P->cr + 2 * N, P->cr + 3 * N, P->dmap);
    prof_end(c, c->stream, LDSO_PROF_N2, c->wG[1] * c->hG[1]);
    c->launches++;
    LDSO_CUDA_OK(c, cudaGetLastError());
    LDSO_CUDA_OK(c, cudaStreamSynchronize(c->stream));   // the pinned staging is reused by the next entry point
    return 0;
}

extern "C" int ldso_gpu_distmap_get(ldso_gpu_ctx* c, float* dist) {
    ImmatureDev* P = immature_pool(c);
    if (!P) return -1;
    LDSO_REQUIRE(c, dist && c->levels >= 2);
    const size_t cells = (size_t)c->wG[1] * c->hG[1];
    LDSO_REQUIRE(c, cells <= P->pinned_bytes);
    unsigned char* st = (unsigned char*)P->pinned;
    LDSO_CUDA_OK(c, ldso_copy(c, st, P->dmap, cells, cudaMemcpyDeviceToHost, c->stream));
    LDSO_CUDA_OK(c, cudaStreamSynchronize(c->stream));
    for (size_t i = 0; i < cells; i++) dist[i] = st[i] == 255 ? 1000.f : (float)st[i];
    return 0;
}

extern "C" int ldso_gpu_immature_select_activation(ldso_gpu_ctx* c, int n_hosts, const float* KRKi, const float* Kt,
                                                   const unsigned char* host_skip, const unsigned char* host_flagged,
                                                   float current_min_act_dist, float min_trace_quality,
                                                   unsigned char* action_out, int counts_out[2]) {
    ImmatureDev* P = immature_pool(c);
    if (!P) return -1;
    LDSO_REQUIRE(c, n_hosts >= 1 && n_hosts <= LDSO_MAX_WINDOW && KRKi && Kt && host_skip && host_flagged && counts_out);
    LDSO_REQUIRE(c, (action_out || P->n == 0) && c->levels >= 2);
    counts_out[0] = counts_out[1] = 0;
    if (P->n == 0) return 0;
    float* st = (float*)P->pinned;
    int rc = put_tables(c, P, n_hosts, KRKi, Kt, host_skip, host_flagged, st);
    if (rc) return rc;
    N2Sel p;
    p.w1 = c->wG[1]; p.h1 = c->hG[1]; p.n = P->n; p.n_hosts = n_hosts; p.tab = P->tables;
    p.host = P->host; p.status = P->status; p.u = P->u; p.v = P->v; p.idmin = P->idmin; p.idmax = P->idmax;
    p.quality = P->quality; p.interval = P->interval; p.type = P->type;
    p.minActDist = current_min_act_dist; p.minQuality = min_trace_quality;
    p.cell = P->act_state; p.frac = (float*)(P->act_state + P->cap); p.thr = (float*)(P->act_state + 2 * (size_t)P->cap);
    p.action = P->action; p.cnt = P->sel_cnt; p.dmap = P->dmap;
    const size_t smem = map_bytes(c);
    rc = set_smem(c, (const void*)n2_select_kernel, smem);
    if (rc) return rc;
    prof_begin(c, c->stream);
    n2_select_kernel<<<1, N2_THREADS, smem, c->stream>>>(p);
    prof_end(c, c->stream, LDSO_PROF_N2SEL, P->n);
    c->launches++;
    LDSO_CUDA_OK(c, cudaGetLastError());
    unsigned char* sa = (unsigned char*)P->pinned + 4096;
    LDSO_REQUIRE(c, (size_t)P->n + 4096 <= P->pinned_bytes);
    LDSO_CUDA_OK(c, ldso_copy(c, sa, P->action, P->n, cudaMemcpyDeviceToHost, c->stream));
    LDSO_CUDA_OK(c, ldso_copy(c, st, P->sel_cnt, 8, cudaMemcpyDeviceToHost, c->stream));
    LDSO_CUDA_OK(c, cudaStreamSynchronize(c->stream));
    memcpy(action_out, sa, P->n);
    counts_out[0] = ((int*)st)[0]; counts_out[1] = ((int*)st)[1];
    return 0;
}
