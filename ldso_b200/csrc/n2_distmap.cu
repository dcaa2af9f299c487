// n2_distmap.cu — N2: CoarseDistanceMap::makeDistanceMap / growDistBFS / addIntoDistFinal (reference
// src/frontend/CoarseTracker.cc:795-930) and the candidate walk of FullSystem::activatePointsMT
// (src/frontend/FullSystem.cc:1226-1298).
//
// The level-1 grid (w/2 x h/2: 76.8 k cells at 640x480) is held as ONE BYTE per cell: ring number 0..39, 255 standing
// for the reference's 1000.f.
//   * The reference grows explicit BFS lists ring by ring; ring k spreads over 8 neighbours when k is odd and over 4
//     when k is even, never out of a border cell, and claims a cell only if it still holds more than k.  Because the
//     neighbourhood of ring k depends on k alone, the grown map is the pointwise minimum over the sources of the
//     single-source ring distance, which for interior cells has the closed form
//         d(dx, dy) = max(a, k*),  a = max(|dx|,|dy|),  k* = least k with k + ceil(k/2) >= |dx| + |dy|
//     (monotone shortest paths stay inside the bounding box of two interior cells, so the border rule never bites).
//     A border cell c is claimed by the first interior neighbour q that can reach it: min over q of D[q] + 1, where a
//     diagonal q only counts if D[q] + 1 is odd.  Sources on the border hold 0 and spread nothing.
//     (tests/test_distmap.py pins the oracle's literal BFS against this form; the GPU is compared with the oracle.)
//   * makeDistanceMap: n2_project_kernel turns the active points into a source-cell list, n2_map_kernel (one CTA per
//     16x16 tile, sources filtered to those within 39 cells of the tile, in chunks through shared memory) takes the
//     minimum for interior cells, n2_border_kernel finishes the border.  No ring iteration, no single-CTA pass.
//   * the candidate walk is order dependent (every accepted point lowers the map for the ones after it) and stays
//     sequential in pool order, but only over ACCEPTED points: all threads classify the pool in parallel first, then
//     one warp tests 32 points at a time against a shared-memory copy of the map, accepts the first that passes, stamps
//     its ring distances over the radius that can still change a decision (thresholds are <= 4 * minActDist) and
//     re-tests the lanes behind it.  A rejected point can never become acceptable later (the map only decreases), so
//     skipping them is exact.  Accepted cells are appended to the source list and the two map kernels rebuild the
//     full-radius fwdWarpedIDDistFinal afterwards.
#include "immature.h"

namespace {

constexpr int N2_THREADS = 1024;
constexpr int N2_TILE = 16;         // n2_map_kernel: cells per tile edge
constexpr int N2_RMAX = 39;         // rings grown by growDistBFS

__device__ __forceinline__ void project(const float* __restrict__ tab, int h, float u, float v, float idepth, float& px, float& py, float& p0) {
    const float* K = tab + 12 * h;
    p0 = (K[0] * u + K[1] * v + K[2] * 1.f) + K[9] * idepth;
    const float p1 = (K[3] * u + K[4] * v + K[5] * 1.f) + K[10] * idepth;
    const float p2 = (K[6] * u + K[7] * v + K[8] * 1.f) + K[11] * idepth;
    px = p0 / p2; py = p1 / p2;
}

// single-source ring distance between interior cells (see the header comment); >= 40 means "not reached"
__device__ __forceinline__ int ring_dist(int dx, int dy) {
    dx = abs(dx); dy = abs(dy);
    const int a = max(dx, dy), s = dx + dy;
    int k = (2 * s) / 3;
    if (k + ((k + 1) >> 1) < s) k++;
    return max(k, a);
}

// value a border cell takes from its interior neighbours: min D[q] + 1, diagonal q only when D[q] + 1 is odd
__device__ __forceinline__ int border_value(const unsigned char* D, int x, int y, int w1, int h1) {
    int best = 255;
#pragma unroll
    for (int dy = -1; dy <= 1; dy++)
#pragma unroll
        for (int dx = -1; dx <= 1; dx++) {
            if (dx == 0 && dy == 0) continue;
            const int qx = x + dx, qy = y + dy;
            if (qx <= 0 || qy <= 0 || qx >= w1 - 1 || qy >= h1 - 1) continue;
            const int j = D[qx + w1 * qy];
            if (j >= N2_RMAX) continue;
            const int cand = j + 1;
            if (dx != 0 && dy != 0 && !(cand & 1)) continue;
            best = min(best, cand);
        }
    return best;
}

// CoarseDistanceMap::makeDistanceMap, first half (CoarseTracker.cc:795-835): active points -> level-1 source cells
__global__ void n2_project_kernel(int w1, int h1, int n_points, const float* __restrict__ tab, const int* __restrict__ host,
                                  const float* __restrict__ u, const float* __restrict__ v, const float* __restrict__ idepth,
                                  int* __restrict__ cells, int* __restrict__ src_n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) *src_n = n_points;
    if (i >= n_points) return;
    int cell = -1;
    const int h = host[i];
    if (h >= 0) {
        float px, py, p0;
        project(tab, h, u[i], v[i], idepth[i], px, py, p0);
        const int x = (int)(px + 0.5f), y = (int)(py + 0.5f);
        if (x > 0 && y > 0 && x < w1 && y < h1) cell = x + w1 * y;
    }
    cells[i] = cell;
}

// growDistBFS over all sources (CoarseTracker.cc:837-923) for the interior cells of one 16x16 tile
__global__ void __launch_bounds__(N2_TILE * N2_TILE)
n2_map_kernel(int w1, int h1, const int* __restrict__ cells, const int* __restrict__ src_n, unsigned char* __restrict__ dmap) {
    __shared__ short2 s_src[N2_TILE * N2_TILE];
    __shared__ int s_warp[N2_TILE * N2_TILE / 32];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int X0 = blockIdx.x * N2_TILE, Y0 = blockIdx.y * N2_TILE;
    const int x = X0 + (tid & (N2_TILE - 1)), y = Y0 + tid / N2_TILE;
    const int n = *src_n;
    int best = 255;
    for (int base = 0; base < n; base += N2_TILE * N2_TILE) {
        const int i = base + tid;
        const int c = i < n ? __ldg(cells + i) : -1;
        int sx = 0, sy = 0;
        bool near = false;
        if (c >= 0) {
            sy = c / w1; sx = c - sy * w1;
            near = sx > 0 && sy > 0 && sx < w1 - 1 && sy < h1 - 1 &&            // border sources spread nothing
                   sx >= X0 - N2_RMAX && sx < X0 + N2_TILE + N2_RMAX && sy >= Y0 - N2_RMAX && sy < Y0 + N2_TILE + N2_RMAX;
        }
        const unsigned m = __ballot_sync(0xffffffffu, near);
        if (lane == 0) s_warp[wid] = __popc(m);
        __syncthreads();
        int off = 0, total = 0;
#pragma unroll
        for (int q = 0; q < N2_TILE * N2_TILE / 32; q++) { if (q < wid) off += s_warp[q]; total += s_warp[q]; }
        if (near) s_src[off + __popc(m & ((1u << lane) - 1))] = make_short2((short)sx, (short)sy);
        __syncthreads();
        for (int k = 0; k < total; k++) {
            const short2 q = s_src[k];
            best = min(best, ring_dist(x - q.x, y - q.y));
        }
        __syncthreads();
    }
    if (x > 0 && y > 0 && x < w1 - 1 && y < h1 - 1) dmap[x + w1 * y] = (unsigned char)(best <= N2_RMAX ? best : 255);
}

// the border ring of the map, then the border sources themselves
__global__ void __launch_bounds__(N2_THREADS)
n2_border_kernel(int w1, int h1, const int* __restrict__ cells, const int* __restrict__ src_n, unsigned char* __restrict__ dmap,
                 const int* __restrict__ corner) {
    const int nb = 2 * w1 + 2 * (h1 - 2);
    for (int b = threadIdx.x; b < nb; b += N2_THREADS) {
        int x, y;
        if (b < w1) { x = b; y = 0; }
        else if (b < 2 * w1) { x = b - w1; y = h1 - 1; }
        else { const int r = b - 2 * w1; y = 1 + (r >> 1); x = (r & 1) ? w1 - 1 : 0; }
        int val = border_value(dmap, x, y, w1, h1);
        if (corner && (x == 0 || x == w1 - 1) && (y == 0 || y == h1 - 1)) val = min(val, corner[(x ? 1 : 0) + (y ? 2 : 0)]);
        dmap[x + w1 * y] = (unsigned char)val;
    }
    __syncthreads();
    const int n = *src_n;
    for (int i = threadIdx.x; i < n; i += N2_THREADS) {
        const int c = cells[i];
        if (c < 0) continue;
        const int y = c / w1, x = c - y * w1;
        if (x == 0 || y == 0 || x == w1 - 1 || y == h1 - 1) dmap[c] = 0;
    }
}

struct N2Sel {
    int w1, h1, n, n_hosts;
    const float* tab;                 // [n_hosts][12] KRKi | Kt, then n_hosts skip bytes and n_hosts flag bytes (as floats)
    const int *host, *status;
    const float *u, *v, *idmin, *idmax, *quality, *interval, *type;
    float minActDist, minQuality;
    int* cell; float* frac; float* thr; int* idx; int* blk_cnt;   // candidates of classification CTA b at [256 b ...), blk_cnt[b] of them
    int* ccell; float* cfrac; float* cthr; int* cidx;   // the same, gaps closed: the candidates in pool order
    unsigned char* action;            // [n]
    int* cnt;                         // [4]: accepted, deleted, bits of the largest candidate threshold, spare (zeroed before the launch)
    const unsigned char* dmap;
    int* src_cells; int* src_n;       // source list of the map kernels: accepted cells are appended
    int* corner;                      // [4] values of the four corner cells after the walk (see n2_walk_kernel)
};

// ---- the candidate walk ------------------------------------------------------------------------------------------------
// n2_classify_kernel (grid over the pool): FullSystem.cc:1243-1296 without the ordered part — delete / keep / candidate, the
// candidate's level-1 cell, the fraction the reference adds to the map value and its threshold; a candidate that already
// fails against the map as it stands can never be accepted (the walk only lowers the map) and is dropped here.
__global__ void __launch_bounds__(256) n2_classify_kernel(const N2Sel p) {
    __shared__ int s_del, s_wc[8];
    __shared__ unsigned s_thmax;
    if (threadIdx.x == 0) { s_del = 0; s_thmax = 0u; }
    __syncthreads();
    const float* skip = p.tab + 12 * p.n_hosts;
    const float* flagged = skip + p.n_hosts;
    const int i = blockIdx.x * blockDim.x + threadIdx.x, lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    int cell = -1;
    float fr = 0.f, th = 0.f;
    if (i < p.n) {
        const int h = p.host[i];
        int act = 0;
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
                        const unsigned char d = p.dmap[cell];
                        if ((d == 255 ? 1000.f : (float)d) + fr >= th) atomicMax(&s_thmax, __float_as_uint(th));
                        else cell = -1;
                    } else {
                        act = 1;
                    }
                }
            }
        }
        p.action[i] = (unsigned char)act;
        if (act) atomicAdd(&s_del, 1);
    }
    // the CTA's candidates, in pool order, into its segment [256 * blockIdx.x ...) of the candidate arrays
    const unsigned m = __ballot_sync(0xffffffffu, cell >= 0);
    if (lane == 0) s_wc[wid] = __popc(m);
    __syncthreads();
    int rank = __popc(m & ((1u << lane) - 1)), total = 0;
#pragma unroll
    for (int q = 0; q < 8; q++) { if (q < wid) rank += s_wc[q]; total += s_wc[q]; }
    if (cell >= 0) {
        const int o = blockIdx.x * 256 + rank;
        p.cell[o] = cell; p.frac[o] = fr; p.thr[o] = th; p.idx[o] = i;
    }
    if (threadIdx.x == 0) {
        p.blk_cnt[blockIdx.x] = total;
        if (s_del) atomicAdd(p.cnt + 1, s_del);
        if (s_thmax) atomicMax(reinterpret_cast<unsigned*>(p.cnt + 2), s_thmax);
    }
}

constexpr int N2_SEL_THREADS = 256;
constexpr int N2_TPL_ROWS = 2 * N2_RMAX + 1;                       // 79
constexpr int N2_TPL_STRIDE = ((4 + N2_TPL_ROWS + 4 + 3) / 4) * 4 + 4;   // bytes per template row: 4 of padding, the row, >= 4 of padding
size_t tpl_bytes() { return (size_t)N2_TPL_ROWS * N2_TPL_STRIDE; }

// value a border cell (not a corner) takes after a stamp
__device__ __forceinline__ void border_fix(unsigned char* D, int x, int y, int w1, int h1) {
    if ((x == 0 || x == w1 - 1) && (y == 0 || y == h1 - 1)) return;   // corners: followed by the walker's lanes 0..3
    const int v = border_value(D, x, y, w1, h1);
    if (v < D[x + w1 * y]) D[x + w1 * y] = (unsigned char)v;
}

// n2_walk_kernel (one CTA; after the set-up four warps): the ordered walk over the surviving candidates.
//   * the candidates are compacted in pool order (CTA-wide scan) so the warp tests 32 real candidates per step;
//   * CoarseDistanceMap::addIntoDistFinal + growDistBFS(1) (CoarseTracker.cc:925-930) on the shared-memory map is a
//     pointwise minimum with the ring-distance TEMPLATE of the source, out to ring R (rings beyond the largest activation
//     threshold cannot change a decision of the walk): the template is built once per launch in shared memory and the
//     stamp is taken four cells at a time (__vminu4 of an aligned map word and the template bytes under it, fetched from
//     two template words by a funnel shift); border cells follow from their interior neighbours where the box reaches them;
//   * warp 0 decides which candidate is accepted next (32 candidates per test); four warps (one per scheduler: a lone warp
//     issues one dependent instruction every ~5 cycles, and the stamp is ~300 map words) stamp it; they meet at a named
//     barrier of 128 threads before and after each stamp.
__global__ void __launch_bounds__(N2_SEL_THREADS) n2_walk_kernel(const N2Sel p) {
    extern __shared__ __align__(16) unsigned char smem[];
    __shared__ int s_scan[N2_SEL_THREADS / 32], s_off[N2_SEL_THREADS], s_cnt[N2_SEL_THREADS], s_c0[2], s_list[2][32];
    __shared__ float s_inv[32];
    if (threadIdx.x < 32) s_inv[threadIdx.x] = 1.0f / (float)max((int)threadIdx.x, 1);
    const int cells = p.w1 * p.h1, cells16 = (cells + 15) / 16;
    unsigned char* D = smem;
    unsigned char* T = smem + (size_t)cells16 * 16;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    {
        uint4* d4 = reinterpret_cast<uint4*>(D);
        const uint4* g4 = reinterpret_cast<const uint4*>(p.dmap);
        for (int i = tid; i < cells16; i += N2_SEL_THREADS) d4[i] = g4[i];
    }
    const float tm = __uint_as_float((unsigned)p.cnt[2]);
    const int R = tm >= (float)N2_RMAX ? N2_RMAX : min(N2_RMAX, (int)ceilf(tm) + 1);
    // template rows -R..R; byte 4 + (dx + R) of a row holds ring_dist(dx, dy) (255 beyond ring R and in the padding)
    for (int i = tid; i < (2 * R + 1) * N2_TPL_STRIDE; i += N2_SEL_THREADS) {
        const int row = i / N2_TPL_STRIDE, b = i - row * N2_TPL_STRIDE;
        const int dx = b - 4 - R, dy = row - R;
        int d = 255;
        if (dx >= -R && dx <= R) { d = ring_dist(dx, dy); if (d > R) d = 255; }
        T[i] = (unsigned char)d;
    }
    // the CTAs of the classification left their candidates in segments of 256: scan the segment lengths and close the gaps
    const int nblk = (p.n + 255) / 256;                  // <= N2_SEL_THREADS (checked by the host)
    const int mine = tid < nblk ? p.blk_cnt[tid] : 0;
    int incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
    if (lane == 31) s_scan[wid] = incl;
    __syncthreads();
    int off = incl - mine, K = 0;
#pragma unroll
    for (int q = 0; q < N2_SEL_THREADS / 32; q++) { if (q < wid) off += s_scan[q]; K += s_scan[q]; }
    s_off[tid] = off; s_cnt[tid] = mine;
    __syncthreads();
    for (int b = wid; b < nblk; b += N2_SEL_THREADS / 32) {
        const int cb = s_cnt[b], ob = s_off[b];
        for (int e = lane; e < cb; e += 32) {
            const int src = b * 256 + e;
            p.ccell[ob + e] = p.cell[src]; p.cfrac[ob + e] = p.frac[src]; p.cthr[ob + e] = p.thr[src]; p.cidx[ob + e] = p.idx[src];
        }
    }
    __syncthreads();
    constexpr int WT = N2_SEL_THREADS;                   // warp 0 decides, the whole CTA stamps
    // ---- the walk
    const int w1 = p.w1, h1 = p.h1;
    const unsigned* T32 = reinterpret_cast<const unsigned*>(T);
    unsigned* D32 = reinterpret_cast<unsigned*>(D);
    const bool words = (w1 & 3) == 0;
    const int n0 = *p.src_n;
    int nacc = 0;
    // A corner cell has ONE interior neighbour, a diagonal one, which can claim it only on an odd ring.  The
    // reference's incremental BFS therefore leaves a corner at the best value that neighbour ever offered while the
    // walk lowered it — not a function of the final map.  Lanes 0..3 of warp 0 follow one corner each.
    const int kx = (lane & 1) ? w1 - 1 : 0, ky = (lane & 2) ? h1 - 1 : 0;
    const int qx = (lane & 1) ? w1 - 2 : 1, qy = (lane & 2) ? h1 - 2 : 1;
    int cv = 255, qv = 255;
    if (tid < 4) { cv = D[kx + w1 * ky]; qv = D[qx + w1 * qy]; }
    // walker state (warp 0): the chunk of 32 candidates under test, the next chunk in flight
    int base = 0, from = 0;
    int cell = -1, idx = 0, cell_n = -1, idx_n = 0;
    float fr = 0.f, th = 0.f, fr_n = 0.f, th_n = 0.f;
    if (wid == 0) {
        if (lane < K) { cell = p.ccell[lane]; fr = p.cfrac[lane]; th = p.cthr[lane]; idx = p.cidx[lane]; }
        if (32 + lane < K) { cell_n = p.ccell[32 + lane]; fr_n = p.cfrac[32 + lane]; th_n = p.cthr[32 + lane]; idx_n = p.cidx[32 + lane]; }
    }
    for (int par = 0;; par ^= 1) {
        if (wid == 0) {
            // The candidates accepted next, in pool order (a rejected one can never pass later: the map only decreases).
            // All lanes that pass against the map as it stands are resolved at once, in registers: lane j is accepted unless
            // an earlier lane accepted in this round would have stamped its cell below its threshold — the very test the
            // reference makes after each addIntoDistFinal, with the closed-form ring distance between two interior cells.
            // A passing candidate on a border cell takes its value from its neighbours instead: then only the first
            // passing lane is accepted and the rest of the chunk is tested again after the stamp.
            int count = -1;
            while (base < K) {
                bool pass = false;
                int cx_ = 0, cy_ = 0;
                if (cell >= 0 && lane >= from) {
                    const unsigned char d = D[cell];
                    pass = (d == 255 ? 1000.f : (float)d) + fr >= th;
                    cy_ = cell / w1; cx_ = cell - cy_ * w1;
                }
                const unsigned m = __ballot_sync(0xffffffffu, pass);
                if (m) {
                    const bool onborder = pass && (cx_ == 0 || cy_ == 0 || cx_ == w1 - 1 || cy_ == h1 - 1);
                    unsigned acc = 0;
                    if (__ballot_sync(0xffffffffu, onborder)) {
                        acc = 1u << (__ffs(m) - 1);
                        from = __ffs(m);
                    } else {
                        bool blocked = false;
                        for (unsigned rem = m; rem; rem &= rem - 1) {
                            const int j = __ffs(rem) - 1;
                            if (__shfl_sync(0xffffffffu, (int)blocked, j)) continue;
                            acc |= 1u << j;
                            const int sjx = __shfl_sync(0xffffffffu, cx_, j), sjy = __shfl_sync(0xffffffffu, cy_, j);
                            if (pass && lane > j) {
                                const int d = ring_dist(cx_ - sjx, cy_ - sjy);
                                if (d <= R && !((float)d + fr >= th)) blocked = true;
                            }
                        }
                        from = 32;                      // every lane of the chunk is decided
                    }
                    count = __popc(acc);
                    const bool mine = (acc >> lane) & 1u;
                    if (mine) {
                        const int r = __popc(acc & ((1u << lane) - 1));
                        p.action[idx] = 2; p.src_cells[n0 + nacc + r] = cell;
                        s_list[par][r] = cx_ | (cy_ << 16);
                    }
                    // a source whose box reaches the border ring changes border cells, whose values depend on the order of the
                    // stamps (a diagonal neighbour counts on odd rings only): such a round is stamped one source at a time
                    const bool edge = mine && (cx_ - R <= 1 || cy_ - R <= 1 || cx_ + R >= w1 - 2 || cy_ + R >= h1 - 2);
                    if (__ballot_sync(0xffffffffu, edge)) count |= 0x100;
                    nacc += count & 0xff;
                    break;
                }
                from = 32;
                if (from >= 32) {
                    base += 32; from = 0;
                    cell = cell_n; fr = fr_n; th = th_n; idx = idx_n;
                    const int inext = base + 32 + lane;
                    cell_n = -1;
                    if (inext < K) { cell_n = p.ccell[inext]; fr_n = p.cfrac[inext]; th_n = p.cthr[inext]; idx_n = p.cidx[inext]; }
                }
            }
            if (count > 0 && from >= 32) {              // the chunk is finished: the next one is tested after the stamps
                base += 32; from = 0;
                cell = cell_n; fr = fr_n; th = th_n; idx = idx_n;
                const int inext = base + 32 + lane;
                cell_n = -1;
                if (inext < K) { cell_n = p.ccell[inext]; fr_n = p.cfrac[inext]; th_n = p.cthr[inext]; idx_n = p.cidx[inext]; }
            }
            if (lane == 0) s_c0[par] = count;
        }
        __syncthreads();
        const int round = s_c0[par];
        if (round < 0) break;
        const int count = round & 0xff;
        if (!(round & 0x100) && words) {
            // ---- no source of the round reaches the border ring: the sources are stamped side by side, one warp each (warp 0
            // follows the corners); where two boxes overlap the word-wise minimum is taken with a compare-and-swap
            if (wid == 0) {
                if (lane < 4)
                    for (int k = 0; k < count; k++) {
                        const int c = s_list[par][k], sx = c & 0xffff, sy = c >> 16;
                        qv = min(qv, ring_dist(qx - sx, qy - sy));
                        const int cand = qv + 1;
                        if ((cand & 1) && cand <= N2_RMAX) cv = min(cv, cand);
                    }
                if (lane < 4 && cv < D[kx + w1 * ky]) D[kx + w1 * ky] = (unsigned char)cv;
            } else {
                for (int k = wid - 1; k < count; k += N2_SEL_THREADS / 32 - 1) {
                    const int c = s_list[par][k], sx = c & 0xffff, sy = c >> 16;
                    const int x0 = sx - R, ylo = sy - R;                    // the box is interior: no clamping
                    const int wlo = x0 >> 2, nw = ((sx + R) >> 2) - wlo + 1, slots = (2 * R + 1) * nw;
                    const float inv_nw = s_inv[nw];
#pragma unroll 2
                    for (int t = lane; t < slots; t += 32) {
                        const int ry = (int)(((float)t + 0.5f) * inv_nw), wq = t - ry * nw;
                        const int y = ylo + ry, wx = wlo + wq;
                        const int b = 4 + 4 * wx - x0;
                        const unsigned* trow = T32 + (ry * N2_TPL_STRIDE >> 2) + (b >> 2);
                        const unsigned tv = __funnelshift_r(trow[0], trow[1], 8 * (b & 3));
                        unsigned* dw = D32 + (y * w1 >> 2) + wx;
                        unsigned dv = *dw;
                        for (;;) {
                            const unsigned mv = __vminu4(dv, tv);
                            if (mv == dv) break;
                            const unsigned seen = atomicCAS(dw, dv, mv);
                            if (seen == dv) break;
                            dv = seen;
                        }
                    }
                }
            }
            __syncthreads();
            continue;
        }
        for (int k = 0; k < count; k++) {
            // ---- stamp source k of the round (all warps), one source at a time
            const int cpk = s_list[par][k];
            const int sx = cpk & 0xffff, sy = cpk >> 16, c0 = sx + w1 * sy;
            const bool interior = sx > 0 && sy > 0 && sx < w1 - 1 && sy < h1 - 1;
            if (tid == 0) D[c0] = 0;
            if (tid < 4) {
                if (sx == kx && sy == ky) cv = 0;
                if (interior) {
                    qv = min(qv, ring_dist(qx - sx, qy - sy));
                    const int cand = qv + 1;
                    if ((cand & 1) && cand <= N2_RMAX) cv = min(cv, cand);
                }
                if (cv < D[kx + w1 * ky]) D[kx + w1 * ky] = (unsigned char)cv;
            }
            if (interior && R > 0) {
                const int ylo = max(1, sy - R), yhi = min(h1 - 2, sy + R);
                if (words) {
                    const int x0 = sx - R;
                    const int wlo = x0 <= 0 ? 0 : (x0 >> 2), whi = min(sx + R, w1 - 1) >> 2, wlast = (w1 >> 2) - 1;
                    const int nw = whi - wlo + 1, slots = (yhi - ylo + 1) * nw;
                    const float inv_nw = 1.0f / (float)nw;                  // (t + 0.5) / nw stays 0.5 / nw away from an integer: exact floor
#pragma unroll 2
                    for (int t = tid; t < slots; t += WT) {
                        const int ry = (int)(((float)t + 0.5f) * inv_nw), wq = t - ry * nw;
                        const int y = ylo + ry, wx = wlo + wq;
                        const int b = 4 + 4 * wx - x0;                       // template byte under the word's first cell (>= 1)
                        const unsigned* trow = T32 + ((y - sy + R) * N2_TPL_STRIDE >> 2) + (b >> 2);
                        unsigned tv = __funnelshift_r(trow[0], trow[1], 8 * (b & 3));
                        if (wx == 0) tv |= 0xffu;                            // column 0 and column w1 - 1 are border cells
                        if (wx == wlast) tv |= 0xff000000u;
                        const int wi = (y * w1 >> 2) + wx;
                        const unsigned dv = D32[wi], mv = __vminu4(dv, tv);
                        if (mv != dv) D32[wi] = mv;
                    }
                } else {
                    const int xlo = max(1, sx - R), xhi = min(w1 - 2, sx + R);
                    const int bw = xhi - xlo + 1, nc = bw * (yhi - ylo + 1);
                    for (int t = tid; t < nc; t += WT) {
                        const int ty = t / bw, x = xlo + (t - ty * bw), y = ylo + ty;
                        const int d = ring_dist(x - sx, y - sy);
                        const int di = x + w1 * y;
                        if (d <= R && D[di] > d) D[di] = (unsigned char)d;
                    }
                }
                if (sx - R <= 1 || sy - R <= 1 || sx + R >= w1 - 2 || sy + R >= h1 - 2) {   // uniform over the CTA
                    // the box reaches the border ring: those cells follow from their interior neighbours
                    __syncthreads();
                    const int bx0 = max(0, sx - R - 1), bx1 = min(w1 - 1, sx + R + 1), by0 = max(0, sy - R - 1), by1 = min(h1 - 1, sy + R + 1);
                    if (by0 == 0) for (int x = bx0 + tid; x <= bx1; x += WT) border_fix(D, x, 0, w1, h1);
                    if (by1 == h1 - 1) for (int x = bx0 + tid; x <= bx1; x += WT) border_fix(D, x, h1 - 1, w1, h1);
                    if (bx0 == 0) for (int y = max(1, by0) + tid; y <= min(h1 - 2, by1); y += WT) border_fix(D, 0, y, w1, h1);
                    if (bx1 == w1 - 1) for (int y = max(1, by0) + tid; y <= min(h1 - 2, by1); y += WT) border_fix(D, w1 - 1, y, w1, h1);
                }
            }
            __syncthreads();
        }
    }
    if (tid == 0) { *p.src_n = n0 + nacc; p.cnt[0] = nacc; }
    if (tid < 4) p.corner[lane] = cv;
}

__global__ void n2_types_kernel(int n, const float* __restrict__ src, float* __restrict__ dst) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) dst[i] = src[i];
}

size_t map_bytes(const ldso_gpu_ctx* c) { return (((size_t)c->wG[1] * c->hG[1] + 15) / 16) * 16; }

int set_smem(ldso_gpu_ctx* c, const void* fn, size_t bytes) {
    LDSO_REQUIRE(c, bytes <= 220 * 1024);   // level-1 grids up to ~215 k cells (1232x368 -> 113 k) + the 7 KB stamp template
    return ldso_raise_smem(c, fn, bytes, 1);
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

// interior cells by tiles, then the border ring: the full fwdWarpedIDDistFinal of the current source list
int launch_map(ldso_gpu_ctx* c, ImmatureDev* P, bool after_walk) {
    const int w1 = c->wG[1], h1 = c->hG[1];
    dim3 grid((w1 + N2_TILE - 1) / N2_TILE, (h1 + N2_TILE - 1) / N2_TILE);
    n2_map_kernel<<<grid, N2_TILE * N2_TILE, 0, c->stream>>>(w1, h1, P->src_cells, P->src_n, P->dmap);
    n2_border_kernel<<<1, N2_THREADS, 0, c->stream>>>(w1, h1, P->src_cells, P->src_n, P->dmap, after_walk ? P->src_n + 1 : nullptr);
    c->launches += 2;
    LDSO_CUDA_OK(c, cudaGetLastError());
    return 0;
}

}  // namespace

extern "C" int ldso_gpu_immature_set_types(ldso_gpu_ctx* c, int first, int n, const float* type) {
    LDSO_ENTER(c);
    ImmatureDev* P = immature_pool(c);
    if (!P) return -1;
    LDSO_REQUIRE(c, first >= 0 && n >= 0 && first + n <= P->n && (n == 0 || type));
    if (n == 0) return 0;
    float* st = (float*)ldso_staging(c, sizeof(float) * n);   // queued: nothing waits here; the copy goes straight into place
    LDSO_REQUIRE(c, st != nullptr);
    memcpy(st, type, sizeof(float) * n);
    LDSO_CUDA_OK(c, ldso_copy(c, P->type + first, st, sizeof(float) * n, cudaMemcpyHostToDevice, c->stream));
    return 0;
}

extern "C" int ldso_gpu_distmap_make(ldso_gpu_ctx* c, int n_hosts, const float* KRKi, const float* Kt, int n_points,
                                     const int* host, const float* u, const float* v, const float* idepth) {
    LDSO_ENTER(c);
    ImmatureDev* P = immature_pool(c);
    if (!P) return -1;
    LDSO_REQUIRE(c, n_hosts >= 1 && n_hosts <= LDSO_MAX_WINDOW && KRKi && Kt && n_points >= 0 && c->levels >= 2);
    LDSO_REQUIRE(c, n_points == 0 || (host && u && v && idepth));
    LDSO_REQUIRE(c, (size_t)n_points * 4 <= (size_t)22 * P->cap && sizeof(float) * (4 * (size_t)n_points + 16 * LDSO_MAX_WINDOW) <= P->pinned_bytes);
    for (int i = 0; i < n_points; i++) LDSO_REQUIRE(c, host[i] < n_hosts);
    float* st = (float*)ldso_staging(c, sizeof(float) * (4 * (size_t)n_points + 16 * LDSO_MAX_WINDOW));   // queued: nothing waits here
    LDSO_REQUIRE(c, st != nullptr);
    int rc = put_tables(c, P, n_hosts, KRKi, Kt, nullptr, nullptr, st);
    if (rc) return rc;
    float* sp = st + 16 * LDSO_MAX_WINDOW;
    const size_t N = n_points;
    if (N) {
        memcpy(sp, host, 4 * N); memcpy(sp + N, u, 4 * N); memcpy(sp + 2 * N, v, 4 * N); memcpy(sp + 3 * N, idepth, 4 * N);
        LDSO_CUDA_OK(c, ldso_copy(c, P->cr, sp, 16 * N, cudaMemcpyHostToDevice, c->stream));
    }
    LDSO_REQUIRE(c, n_points <= P->src_cap);
    P->src_n_host = n_points;
    P->map_stale = false;
    prof_begin(c, c->stream);
    n2_project_kernel<<<(n_points + 255) / 256 + 1, 256, 0, c->stream>>>(c->wG[1], c->hG[1], n_points, P->tables, (const int*)P->cr,
                                                                        P->cr + N, P->cr + 2 * N, P->cr + 3 * N, P->src_cells, P->src_n);
    rc = launch_map(c, P, false);
    prof_end(c, c->stream, LDSO_PROF_N2, c->wG[1] * c->hG[1]);
    c->launches++;
    return rc;
}

extern "C" int ldso_gpu_distmap_get(ldso_gpu_ctx* c, float* dist) {
    LDSO_ENTER(c);
    ImmatureDev* P = immature_pool(c);
    if (!P) return -1;
    LDSO_REQUIRE(c, dist && c->levels >= 2);
    const size_t cells = (size_t)c->wG[1] * c->hG[1];
    LDSO_REQUIRE(c, cells <= P->pinned_bytes);
    if (P->map_stale) {    // fwdWarpedIDDistFinal after the walk of activatePointsMT: nothing on the path reads it, only debug code
        const int rc = launch_map(c, P, true);
        if (rc) return rc;
        P->map_stale = false;
    }
    unsigned char* st = (unsigned char*)P->pinned;
    LDSO_CUDA_OK(c, ldso_copy(c, st, P->dmap, cells, cudaMemcpyDeviceToHost, c->stream));
    LDSO_CUDA_OK(c, ldso_stream_sync(c->stream));
    for (size_t i = 0; i < cells; i++) dist[i] = st[i] == 255 ? 1000.f : (float)st[i];
    return 0;
}

extern "C" int ldso_gpu_immature_select_activation(ldso_gpu_ctx* c, int n_hosts, const float* KRKi, const float* Kt,
                                                   const unsigned char* host_skip, const unsigned char* host_flagged,
                                                   float current_min_act_dist, float min_trace_quality,
                                                   unsigned char* action_out, int counts_out[2]) {
    LDSO_ENTER(c);
    ImmatureDev* P = immature_pool(c);
    if (!P) return -1;
    LDSO_REQUIRE(c, n_hosts >= 1 && n_hosts <= LDSO_MAX_WINDOW && KRKi && Kt && host_skip && host_flagged && counts_out);
    LDSO_REQUIRE(c, (action_out || P->n == 0) && c->levels >= 2);
    counts_out[0] = counts_out[1] = 0;
    if (P->n == 0) return 0;
    float* st = (float*)ldso_staging(c, sizeof(float) * 16 * LDSO_MAX_WINDOW);
    LDSO_REQUIRE(c, st != nullptr);
    if (P->map_stale) {    // a second walk on the same map: it starts from the map the first walk left
        const int rcm = launch_map(c, P, true);
        if (rcm) return rcm;
        P->map_stale = false;
    }
    int rc = put_tables(c, P, n_hosts, KRKi, Kt, host_skip, host_flagged, st);
    if (rc) return rc;
    N2Sel p;
    p.w1 = c->wG[1]; p.h1 = c->hG[1]; p.n = P->n; p.n_hosts = n_hosts; p.tab = P->tables;
    p.host = P->host; p.status = P->status; p.u = P->u; p.v = P->v; p.idmin = P->idmin; p.idmax = P->idmax;
    p.quality = P->quality; p.interval = P->interval; p.type = P->type;
    p.minActDist = current_min_act_dist; p.minQuality = min_trace_quality;
    LDSO_REQUIRE(c, P->n <= 256 * N2_SEL_THREADS && LDSO_MAX_WINDOW >= 9);
    {
        int* a = P->act_state;                         // [cap][LDSO_MAX_WINDOW] ints of scratch; segments need n rounded up to 256 <= cap + 255
        const size_t S = ((size_t)P->cap + 255) / 256 * 256;
        LDSO_REQUIRE(c, 8 * S + 256 <= (size_t)P->cap * LDSO_MAX_WINDOW + 4096);
        p.cell = a; p.frac = (float*)(a + S); p.thr = (float*)(a + 2 * S); p.idx = a + 3 * S;
        p.ccell = a + 4 * S; p.cfrac = (float*)(a + 5 * S); p.cthr = (float*)(a + 6 * S); p.cidx = a + 7 * S; p.blk_cnt = a + 8 * S;
    }
    p.action = P->action; p.cnt = P->sel_cnt; p.dmap = P->dmap; p.src_cells = P->src_cells; p.src_n = P->src_n; p.corner = P->src_n + 1;
    LDSO_REQUIRE(c, P->src_n_host + P->n <= P->src_cap);
    P->src_n_host += P->n;   // upper bound of the list length after this walk
    const size_t smem = map_bytes(c) + tpl_bytes();
    rc = set_smem(c, (const void*)n2_walk_kernel, smem);
    if (rc) return rc;
    LDSO_CUDA_OK(c, cudaMemsetAsync(P->sel_cnt, 0, 16, c->stream));
    prof_begin(c, c->stream);
    n2_classify_kernel<<<(P->n + 255) / 256, 256, 0, c->stream>>>(p);
    n2_walk_kernel<<<1, N2_SEL_THREADS, smem, c->stream>>>(p);
    prof_end(c, c->stream, LDSO_PROF_N2SEL, P->n);
    c->launches += 2;
    LDSO_CUDA_OK(c, cudaGetLastError());
    P->map_stale = true;   // the accepted cells joined the source list: the full-radius map is rebuilt when somebody reads it (ldso_gpu_distmap_get)
    unsigned char* sa = (unsigned char*)P->pinned + 4096;
    LDSO_REQUIRE(c, (size_t)P->n + 4096 <= P->pinned_bytes);
    LDSO_CUDA_OK(c, ldso_copy(c, sa, P->action, P->n, cudaMemcpyDeviceToHost, c->stream));
    int* cnt_h = (int*)P->pinned;
    LDSO_CUDA_OK(c, ldso_copy(c, cnt_h, P->sel_cnt, 8, cudaMemcpyDeviceToHost, c->stream));
    LDSO_CUDA_OK(c, ldso_stream_sync(c->stream));
    memcpy(action_out, sa, P->n);
    counts_out[0] = cnt_h[0]; counts_out[1] = cnt_h[1];
    return 0;
}
