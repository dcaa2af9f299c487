// n1_initializer.cu — N1: CoarseInitializer (reference src/frontend/CoarseInitializer.cc, include/frontend/CoarseInitializer.h)
// and makePixelStatus / gridMaxSelection (include/frontend/PixelSelector2.h:60-278).
//
// Point state (struct Pnt) is SoA per pyramid level on the device.  One calcResAndGS (:186-342) is two launches and one
// read-back:
//   n1_residual_kernel  calculateAllResiduals (:391-518) + calculateAlphaEnergy (:376-389): thread per point, the 8 pattern
//                       pixels in a register loop (4-texel gathers on the float4 pyramid), JbBuffer_new written as SoA,
//                       the 45 entries of Accumulator9 + E + alpha energy reduced warp -> CTA -> per-CTA partial; the CTA
//                       that arrives last folds all partials in fp64 in CTA order (deterministic) and decides alphaOpt
//                       (:273-281) on the device, so the second launch needs no host round trip;
//   n1_sc_kernel        calculateSC (:349-374) + the two sums of calcEC (:520-536): same reduction, the last CTA publishes
//                       both result blocks.
// doStep / applyStep / propagateDown / propagateUp (children of a parent summed in index order: same float sums as the
// reference's scatter loop) are one thread per point.  Two steps are inherently sequential in the reference and stay on the
// host, on a small read-back of {iR, idepth, isGood}: optReg (:538-567: each point takes the median over neighbours that may
// already have been regularised in the same sweep) and the neighbour fill of resetPoints on the coarsest level (:747-759).
// The 10-NN / parent graph (makeNN :839-906) is built once per setFirst on the host (include/ldso_knn.h), as SURVEY.md 8f
// keeps it.  The LM loop of trackFrame (:45-183) runs on the host over these launches.
#include "ctx.h"
#include "../../include/ldso_knn.h"

#include <algorithm>

using ldso_sm::SE3;

namespace {

constexpr int N1_THREADS = 128;
constexpr int N1_NRED = 48;          // 45 (9x9 upper) + 3 scalars
enum { F_U, F_V, F_ID, F_IDN, F_IR, F_E0, F_E1, F_EN0, F_EN1, F_LH, F_LHN, F_MAXSTEP, F_TYPE, F_NUM };

struct InitLevel {
    int n = 0, cap = 0, blocks = 0;
    float* f = nullptr;                  // [F_NUM][cap]
    unsigned char *good = nullptr, *good_new = nullptr;
    float *Jb = nullptr, *Jb_new = nullptr;   // [10][cap]
    int* nb = nullptr;                   // [10][cap]
    int* parent = nullptr;               // [cap]
    int *child_start = nullptr, *child_idx = nullptr;   // CSR over the points of the level below that name this one as parent
    std::vector<int> nb_host, parent_host;               // [n][10], [n]
    float* F(int k) const { return f + (size_t)k * cap; }
};

struct InitDev {
    int levels = 0;
    int w[LDSO_MAX_LEVELS], h[LDSO_MAX_LEVELS];
    double fx[LDSO_MAX_LEVELS], fy[LDSO_MAX_LEVELS], cx[LDSO_MAX_LEVELS], cy[LDSO_MAX_LEVELS];
    double Ki[LDSO_MAX_LEVELS][9];
    InitLevel lv[LDSO_MAX_LEVELS];
    int first_slot = -1, new_slot = -1;
    SE3 thisToNext;
    double aff_a = 0, aff_b = 0;
    bool snapped = false, fixAffine = true;
    int frameID = -1, snappedAt = 0;
    float alphaK = 2.5f * 2.5f, alphaW = 150 * 150, regWeight = 0.8f, couplingWeight = 1;
    int sparsityFactor = 5;
    // reduction workspace
    float* partials = nullptr;           // [max_blocks][2][N1_NRED]
    unsigned int* counter = nullptr;     // [2]
    double* result = nullptr;            // [2][64]: block 0 = residual pass, block 1 = SC pass
    double* result_pinned = nullptr;
    unsigned char* map_dev = nullptr;    // makePixelStatus map (level >= 1) + count
    int* count_dev = nullptr;
    void* pinned = nullptr;              // host staging for point arrays
    size_t pinned_bytes = 0;
    int max_blocks = 0;
};

struct ResParams {
    int n, cap, w, h;
    float RKi[9], t[3], aff0, aff1, fx, fy, cx, cy, huberTH;
    double tsq;                          // refToNew.translation().squaredNorm()
    float alphaW, alphaK, couplingWeight, outlierTH20;   // outlierTH20 = patternNum * setting_outlierTH * 20 (:485, :715)
    int npts;
};

__device__ __forceinline__ void pattern(int k, int& pdx, int& pdy) {   // src/Setting.cc:275
    pdx = (k == 1 || k == 6) ? -1 : (k == 2) ? 1 : (k == 3) ? -2 : (k == 5) ? 2 : 0;
    pdy = (k == 0) ? -2 : (k == 1 || k == 2) ? -1 : (k == 6) ? 1 : (k == 7) ? 2 : 0;
}
// getInterpolatedElement33 / 31 (include/internal/GlobalFuncs.h:89-103, :145-159) on float4 texels
__device__ __forceinline__ float3 interp33(const float4* img, float x, float y, int w) {
    const int ix = (int)x, iy = (int)y;
    const float dx = x - ix, dy = y - iy, dxdy = dx * dy;
    const float4* bp = img + ix + iy * w;
    const float4 a = __ldg(bp + 1 + w), b = __ldg(bp + w), c = __ldg(bp + 1), d = __ldg(bp);
    const float w0 = dxdy, w1 = dy - dxdy, w2 = dx - dxdy, w3 = 1 - dx - dy + dxdy;
    return make_float3(w0 * a.x + w1 * b.x + w2 * c.x + w3 * d.x, w0 * a.y + w1 * b.y + w2 * c.y + w3 * d.y,
                       w0 * a.z + w1 * b.z + w2 * c.z + w3 * d.z);
}
__device__ __forceinline__ float interp31(const float4* img, float x, float y, int w) {
    const int ix = (int)x, iy = (int)y;
    const float dx = x - ix, dy = y - iy, dxdy = dx * dy;
    const float4* bp = img + ix + iy * w;
    return dxdy * __ldg(&bp[1 + w].x) + (dy - dxdy) * __ldg(&bp[w].x) + (dx - dxdy) * __ldg(&bp[1].x) +
           (1 - dx - dy + dxdy) * __ldg(&bp[0].x);
}

// CTA reduction of acc[N1_NRED] into partials[blockIdx][pass]; returns true in the CTA that arrived last
__device__ __forceinline__ bool reduce_to_partials(float* acc, float* partials, unsigned int* counter, int pass) {
    __shared__ float s[N1_THREADS / 32][N1_NRED];
    __shared__ bool s_last;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < N1_NRED; k++) {
        float v = acc[k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0) s[wid][k] = v;
    }
    __syncthreads();
    if (threadIdx.x < N1_NRED) {
        float v = 0.f;
#pragma unroll
        for (int q = 0; q < N1_THREADS / 32; q++) v += s[q][threadIdx.x];
        partials[((size_t)blockIdx.x * 2 + pass) * N1_NRED + threadIdx.x] = v;
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) s_last = atomicAdd(counter + pass, 1u) == gridDim.x - 1;
    __syncthreads();
    return s_last;
}
// the last CTA: fp64 sum over CTAs in CTA order
__device__ __forceinline__ void fold_partials(const float* partials, int pass, double* out) {
    __threadfence();
    if (threadIdx.x < N1_NRED) {
        double v = 0;
        const volatile float* pv = partials;
        for (unsigned b = 0; b < gridDim.x; b++) v += (double)pv[((size_t)b * 2 + pass) * N1_NRED + threadIdx.x];
        out[threadIdx.x] = v;
    }
}

// ---- calculateAllResiduals (:391-518) + calculateAlphaEnergy (:376-389) ------------------------------------------
// acc layout: [0..44] Accumulator9 upper triangle (dp0..dp7, r), [45] accE, [46] accEAlpha
__global__ void __launch_bounds__(N1_THREADS)
n1_residual_kernel(const ResParams p, const float4* __restrict__ colorRef, const float4* __restrict__ colorNew, float* __restrict__ f,
                   const unsigned char* __restrict__ good, unsigned char* __restrict__ good_new, float* __restrict__ Jb_new,
                   float* partials, unsigned int* counter, double* result) {
    float acc[N1_NRED];
#pragma unroll
    for (int k = 0; k < N1_NRED; k++) acc[k] = 0.f;
    const int i = blockIdx.x * N1_THREADS + threadIdx.x;
    const int cap = p.cap;
    if (i < p.n) {
        float maxstepP = 1e10f;
        const float e0 = f[F_E0 * cap + i], e1 = f[F_E1 * cap + i];
        const float idn = f[F_IDN * cap + i];
        bool ok = good[i] != 0;
        bool gnew = false;
        if (ok) {
            const float pu = f[F_U * cap + i], pv = f[F_V * cap + i];
            float dp[9][8];
            float jb[10];
            float energy = 0;
#pragma unroll
            for (int k = 0; k < 10; k++) jb[k] = 0.f;
#pragma unroll
            for (int idx = 0; idx < 8; idx++) {
#pragma unroll
                for (int k = 0; k < 9; k++) dp[k][idx] = 0.f;
                if (!ok) continue;                      // the reference breaks out of the pattern loop (:431-445)
                int dx, dy;
                pattern(idx, dx, dy);
                const float px = pu + dx, py = pv + dy;
                const float pt0 = (p.RKi[0] * px + p.RKi[1] * py + p.RKi[2] * 1.f) + p.t[0] * idn;
                const float pt1 = (p.RKi[3] * px + p.RKi[4] * py + p.RKi[5] * 1.f) + p.t[1] * idn;
                const float pt2 = (p.RKi[6] * px + p.RKi[7] * py + p.RKi[8] * 1.f) + p.t[2] * idn;
                const float u = pt0 / pt2, v = pt1 / pt2;
                const float Ku = p.fx * u + p.cx, Kv = p.fy * v + p.cy;
                const float new_idepth = idn / pt2;
                if (!(Ku > 1 && Kv > 1 && Ku < p.w - 2 && Kv < p.h - 2 && new_idepth > 0)) { ok = false; continue; }
                const float3 hit = interp33(colorNew, Ku, Kv, p.w);
                const float rlR = interp31(colorRef, px, py, p.w);
                if (!isfinite(rlR) || !isfinite(hit.x)) { ok = false; continue; }
                const float residual = hit.x - p.aff0 * rlR - p.aff1;
                float hw = fabsf(residual) < p.huberTH ? 1.f : p.huberTH / fabsf(residual);
                energy += hw * residual * residual * (2 - hw);
                const float dxdd = (p.t[0] - p.t[2] * u) / pt2;
                const float dydd = (p.t[1] - p.t[2] * v) / pt2;
                if (hw < 1) hw = sqrtf(hw);
                const float dxInterp = hw * hit.y * p.fx;
                const float dyInterp = hw * hit.z * p.fy;
                dp[0][idx] = new_idepth * dxInterp;
                dp[1][idx] = new_idepth * dyInterp;
                dp[2][idx] = -new_idepth * (u * dxInterp + v * dyInterp);
                dp[3][idx] = -u * v * dxInterp - (1 + v * v) * dyInterp;
                dp[4][idx] = (1 + u * u) * dxInterp + u * v * dyInterp;
                dp[5][idx] = -v * dxInterp + u * dyInterp;
                dp[6][idx] = -hw * p.aff0 * rlR;
                dp[7][idx] = -hw * 1;
                const float dd = dxInterp * dxdd + dyInterp * dydd;
                dp[8][idx] = hw * residual;
                const float a = dxdd * p.fx, b = dydd * p.fy;
                const float maxstep = 1.0f / sqrtf(a * a + b * b);
                if (maxstep < maxstepP) maxstepP = maxstep;
#pragma unroll
                for (int k = 0; k < 9; k++) jb[k] += dp[k][idx] * dd;
                jb[9] += dd * dd;
            }
            // JbBuffer_new holds whatever the pattern loop reached, as in the reference
#pragma unroll
            for (int k = 0; k < 10; k++) Jb_new[(size_t)k * cap + i] = jb[k];
            if (ok && !(energy > p.outlierTH20)) {
                gnew = true;
                acc[45] = energy;
                f[F_EN0 * cap + i] = energy;
                // acc9.updateSSE over the two groups of four pattern pixels (:499-509)
                int q = 0;
#pragma unroll
                for (int r = 0; r < 9; r++)
#pragma unroll
                    for (int c2 = r; c2 < 9; c2++) {
                        float sacc = 0.f;
#pragma unroll
                        for (int idx = 0; idx < 8; idx++) sacc += dp[r][idx] * dp[c2][idx];
                        acc[q++] = sacc;
                    }
            }
        }
        if (!gnew) {
            acc[45] = e0;                                // accE.updateSingle(point->energy[0]) (:398, :486)
            f[F_EN0 * cap + i] = e0; f[F_EN1 * cap + i] = e1;
            acc[46] = e1;                                // calculateAlphaEnergy (:381-383)
        } else {
            const float ea = (idn - 1) * (idn - 1);      // (:384-387)
            f[F_EN1 * cap + i] = ea;
            acc[46] = ea;
        }
        f[F_MAXSTEP * cap + i] = maxstepP;
        good_new[i] = gnew ? 1 : 0;
    }
    if (reduce_to_partials(acc, partials, counter, 0)) {
        fold_partials(partials, 0, result);
        __syncthreads();
        if (threadIdx.x == 0) {
            // calcResAndGS :260-281
            float alphaEnergy = (float)(p.tsq * p.npts);
            alphaEnergy += (float)result[46];
            alphaEnergy *= p.alphaW;
            float alphaOpt;
            if (alphaEnergy > p.alphaK * p.npts) { alphaOpt = 0; alphaEnergy = p.alphaK * p.npts; }
            else alphaOpt = p.alphaW;
            result[48] = alphaOpt; result[49] = alphaEnergy;
            counter[0] = 0;
        }
    }
}

// ---- calculateSC (:349-374) + calcEC sums (:520-536) ---------------------------------------------------------------
// acc layout: [0..44] Accumulator9 upper triangle of the weighted Schur terms, [45] sum rOld^2, [46] sum rNew^2
__global__ void __launch_bounds__(N1_THREADS)
n1_sc_kernel(int n, int cap, float couplingWeight, float* __restrict__ f, const unsigned char* __restrict__ good_new,
             float* __restrict__ Jb_new, float* partials, unsigned int* counter, double* result, double* result_host) {
    float acc[N1_NRED];
#pragma unroll
    for (int k = 0; k < N1_NRED; k++) acc[k] = 0.f;
    const int i = blockIdx.x * N1_THREADS + threadIdx.x;
    const float alphaOpt = (float)((volatile double*)result)[48];
    if (i < n && good_new[i]) {
        float J[9];
#pragma unroll
        for (int k = 0; k < 8; k++) J[k] = Jb_new[(size_t)k * cap + i];
        float j8 = Jb_new[(size_t)8 * cap + i], j9 = Jb_new[(size_t)9 * cap + i];
        const float idn = f[F_IDN * cap + i], iR = f[F_IR * cap + i], id = f[F_ID * cap + i];
        f[F_LHN * cap + i] = j9;
        j8 += alphaOpt * (idn - 1);
        j9 += alphaOpt;
        if (alphaOpt == 0) {
            j8 += couplingWeight * (idn - iR);
            j9 += couplingWeight;
        }
        j9 = 1 / (1 + j9);
        Jb_new[(size_t)8 * cap + i] = j8;
        Jb_new[(size_t)9 * cap + i] = j9;
        J[8] = j8;
        int q = 0;
#pragma unroll
        for (int r = 0; r < 9; r++) {
            const float Jw = J[r] * j9;                  // updateSingleWeighted (MatrixAccumulators.h:1493-1615)
#pragma unroll
            for (int c2 = r; c2 < 9; c2++) acc[q++] = Jw * J[c2];
        }
        const float rOld = id - iR, rNew = idn - iR;
        acc[45] = rOld * rOld; acc[46] = rNew * rNew;
    }
    if (reduce_to_partials(acc, partials, counter, 1)) {
        fold_partials(partials, 1, result + 64);
        __syncthreads();
        if (threadIdx.x == 0) counter[1] = 0;
        // publish both result blocks through mapped pinned memory
        for (int k = threadIdx.x; k < 128; k += N1_THREADS) result_host[k] = result[k];
    }
}

// ---- per-point passes ---------------------------------------------------------------------------------------------
// doStep (:763-789)
__global__ void n1_do_step_kernel(int n, int cap, float lambda, const float* __restrict__ inc8, float* __restrict__ f,
                                  const unsigned char* __restrict__ good, const float* __restrict__ Jb) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n || !good[i]) return;
    float dot = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) dot += Jb[(size_t)k * cap + i] * inc8[k];
    const float b = Jb[(size_t)8 * cap + i] + dot;
    float step = -b * Jb[(size_t)9 * cap + i] / (1 + lambda);
    float maxstep = 0.25f * f[F_MAXSTEP * cap + i];
    if (maxstep > 1e10f) maxstep = 1e10f;
    if (step > maxstep) step = maxstep;
    if (step < -maxstep) step = -maxstep;
    float newIdepth = f[F_ID * cap + i] + step;
    if (newIdepth < 1e-3f) newIdepth = 1e-3f;
    if (newIdepth > 50) newIdepth = 50;
    f[F_IDN * cap + i] = newIdepth;
}
// applyStep (:791-809); the JbBuffer swap is a pointer swap on the host
__global__ void n1_apply_step_kernel(int n, int cap, float* __restrict__ f, unsigned char* __restrict__ good,
                                     const unsigned char* __restrict__ good_new) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (!good[i]) {
        const float iR = f[F_IR * cap + i];
        f[F_ID * cap + i] = iR; f[F_IDN * cap + i] = iR;
        return;
    }
    f[F_E0 * cap + i] = f[F_EN0 * cap + i]; f[F_E1 * cap + i] = f[F_EN1 * cap + i];
    good[i] = good_new[i];
    f[F_ID * cap + i] = f[F_IDN * cap + i];
    f[F_LH * cap + i] = f[F_LHN * cap + i];
}
// resetPoints without the neighbour fill (:739-745); mode 1: the !snapped reset of trackFrame (:55-66); mode 2: optReg
// while not snapped (:541-545)
__global__ void n1_reset_kernel(int n, int cap, int mode, float* __restrict__ f) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (mode == 0) { f[F_E0 * cap + i] = 0; f[F_E1 * cap + i] = 0; f[F_IDN * cap + i] = f[F_ID * cap + i]; }
    else if (mode == 1) { f[F_IR * cap + i] = 1; f[F_IDN * cap + i] = 1; f[F_LH * cap + i] = 0; }
    else f[F_IR * cap + i] = 1;
}
// propagateDown (:606-629): fine level `f` reads its parents in the coarse level `fc`
__global__ void n1_prop_down_kernel(int n, int cap, float* __restrict__ f, unsigned char* __restrict__ good,
                                    const int* __restrict__ parent, int capc, const float* __restrict__ fc,
                                    const unsigned char* __restrict__ goodc) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int pa = parent[i];
    if (pa < 0) return;
    const float plh = fc[F_LH * capc + pa], piR = fc[F_IR * capc + pa];
    if (!goodc[pa] || plh < 0.1f) return;
    if (!good[i]) {
        f[F_IR * cap + i] = piR; f[F_ID * cap + i] = piR; f[F_IDN * cap + i] = piR;
        good[i] = 1;
        f[F_LH * cap + i] = 0;
    } else {
        const float lh = f[F_LH * cap + i];
        const float newiR = (f[F_IR * cap + i] * lh * 2 + piR * plh) / (lh * 2 + plh);
        f[F_IR * cap + i] = newiR; f[F_ID * cap + i] = newiR; f[F_IDN * cap + i] = newiR;
    }
}
// propagateUp (:570-601): one thread per parent walks its children in index order, which is the order the reference's
// scatter loop adds them in
__global__ void n1_prop_up_kernel(int nparent, int capp, float* __restrict__ fp, unsigned char* __restrict__ goodp,
                                  const int* __restrict__ child_start, const int* __restrict__ child_idx, int capc,
                                  const float* __restrict__ fc, const unsigned char* __restrict__ goodc) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nparent) return;
    float iR = 0, sum = 0;
    for (int k = child_start[i]; k < child_start[i + 1]; k++) {
        const int ch = child_idx[k];
        if (!goodc[ch]) continue;
        const float lh = fc[F_LH * capc + ch];
        iR += fc[F_IR * capc + ch] * lh;
        sum += lh;
    }
    if (sum > 0) {
        const float v = iR / sum;
        fp[F_ID * capp + i] = v; fp[F_IR * capp + i] = v;
        goodp[i] = 1;
    } else {
        fp[F_IR * capp + i] = 0;                         // parent->iR = 0 (:582) stays when nothing was added
    }
}

// gridMaxSelection (include/frontend/PixelSelector2.h:63-236): one thread per pot x pot cell; the four winners of a cell
// lie inside the cell, so cells never collide
__global__ void n1_grid_max_kernel(const float4* __restrict__ grads, unsigned char* __restrict__ map, int w, int h, int pot,
                                   float THFac, int* __restrict__ count) {
    const int ncx = (w - pot - 1 + pot - 1) / pot, ncy = (h - pot - 1 + pot - 1) / pot;   // x = 1, 1 + pot, ... < w - pot
    const int cidx = blockIdx.x * blockDim.x + threadIdx.x;
    if (cidx >= ncx * ncy) return;
    const int x = 1 + (cidx % ncx) * pot, y = 1 + (cidx / ncx) * pot;
    int bestXXID = -1, bestYYID = -1, bestXYID = -1, bestYXID = -1;
    float bestXX = 0, bestYY = 0, bestXY = 0, bestYX = 0;
    const float TH = THFac * 10.0f * 0.75f;
    for (int dx = 0; dx < pot; dx++)
        for (int dy = 0; dy < pot; dy++) {
            const int idx = dx + dy * w;
            const float4 g = __ldg(grads + x + y * w + idx);
            const float sqgd = g.y * g.y + g.z * g.z;
            if (sqgd > TH * TH) {
                const float agx = fabsf(g.y);
                if (agx > bestXX) { bestXX = agx; bestXXID = idx; }
                const float agy = fabsf(g.z);
                if (agy > bestYY) { bestYY = agy; bestYYID = idx; }
                const float gxpy = fabsf(g.y - g.z);
                if (gxpy > bestXY) { bestXY = gxpy; bestXYID = idx; }
                const float gxmy = fabsf(g.y + g.z);
                if (gxmy > bestYX) { bestYX = gxmy; bestYXID = idx; }
            }
        }
    unsigned char* map0 = map + x + y * w;
    const int ids[4] = {bestXXID, bestYYID, bestXYID, bestYXID};
    int good = 0;
#pragma unroll
    for (int k = 0; k < 4; k++)
        if (ids[k] >= 0) {
            if (!map0[ids[k]]) good++;
            map0[ids[k]] = 1;
        }
    if (good) atomicAdd(count, good);
}

}  // namespace

// ---- host side ------------------------------------------------------------------------------------------------------
namespace {

void level_free(InitLevel& L) {
    void* ptrs[] = {L.f, L.good, L.good_new, L.Jb, L.Jb_new, L.nb, L.parent, L.child_start, L.child_idx};
    for (void* q : ptrs) cudaFree(q);
    L = InitLevel();
}

InitDev* init_state(ldso_gpu_ctx* c) {
    if (c->initializer) return (InitDev*)c->initializer;
    InitDev* I = new InitDev();
    c->initializer = I;
    I->levels = c->levels;
    const size_t px = (size_t)c->wG[0] * c->hG[0];
    I->max_blocks = (int)((px + N1_THREADS - 1) / N1_THREADS);
    bool ok = true;
    ok = ok && cudaMalloc((void**)&I->partials, sizeof(float) * 2 * N1_NRED * (size_t)I->max_blocks) == cudaSuccess;
    ok = ok && cudaMalloc((void**)&I->counter, 16) == cudaSuccess;
    ok = ok && cudaMalloc((void**)&I->result, sizeof(double) * 128) == cudaSuccess;
    ok = ok && cudaMalloc((void**)&I->map_dev, px) == cudaSuccess;
    ok = ok && cudaMalloc((void**)&I->count_dev, 16) == cudaSuccess;
    ok = ok && cudaHostAlloc((void**)&I->result_pinned, sizeof(double) * 128, cudaHostAllocMapped) == cudaSuccess;
    I->pinned_bytes = 16 * px + 4096;
    ok = ok && cudaMallocHost(&I->pinned, I->pinned_bytes) == cudaSuccess;
    if (ok) { cudaMemset(I->counter, 0, 16); cudaMemset(I->result, 0, sizeof(double) * 128); }
    if (!ok) { c->set_error("initializer allocation failed", __FILE__, __LINE__); return nullptr; }
    return I;
}

// CoarseInitializer::makeK (:811-837)
void make_k(ldso_gpu_ctx* c, InitDev* I) {
    I->w[0] = c->wG[0]; I->h[0] = c->hG[0];
    I->fx[0] = c->cfg.fx; I->fy[0] = c->cfg.fy; I->cx[0] = c->cfg.cx; I->cy[0] = c->cfg.cy;
    for (int l = 1; l < I->levels; l++) {
        I->w[l] = I->w[0] >> l; I->h[l] = I->h[0] >> l;
        I->fx[l] = I->fx[l - 1] * 0.5; I->fy[l] = I->fy[l - 1] * 0.5;
        I->cx[l] = (I->cx[0] + 0.5) / ((int)1 << l) - 0.5;
        I->cy[l] = (I->cy[0] + 0.5) / ((int)1 << l) - 0.5;
    }
    for (int l = 0; l < I->levels; l++) {
        const double K[9] = {I->fx[l], 0, I->cx[l], 0, I->fy[l], I->cy[l], 0, 0, 1};
        ldso_sm::mat3_inv(K, I->Ki[l]);
    }
}

// makePixelStatus (PixelSelector2.h:239-277) around the grid-max kernel; the map comes back to the host (setFirst builds
// the point list and the neighbour graph there)
int make_pixel_status(ldso_gpu_ctx* c, InitDev* I, int slot, int lvl, float desiredDensity, unsigned char* map_host, int recsLeft = 5,
                      float THFac = 1) {
    const int w = I->w[lvl], h = I->h[lvl];
    if (I->sparsityFactor < 1) I->sparsityFactor = 1;
    const int pot = I->sparsityFactor;
    int numGoodPoints = 0;
    LDSO_CUDA_OK(c, cudaMemsetAsync(I->map_dev, 0, (size_t)w * h, c->stream));
    LDSO_CUDA_OK(c, cudaMemsetAsync(I->count_dev, 0, 4, c->stream));
    const int ncx = (w - pot - 1 + pot - 1) / pot, ncy = (h - pot - 1 + pot - 1) / pot;
    if (ncx > 0 && ncy > 0) {
        n1_grid_max_kernel<<<(ncx * ncy + 127) / 128, 128, 0, c->stream>>>(c->frame[slot] + c->lvl_off[lvl], I->map_dev, w, h, pot, THFac,
                                                                           I->count_dev);
        c->launches++;
        LDSO_CUDA_OK(c, cudaGetLastError());
    }
    int* cnt = (int*)I->pinned;
    LDSO_CUDA_OK(c, ldso_copy(c, cnt, I->count_dev, 4, cudaMemcpyDeviceToHost, c->stream));
    LDSO_CUDA_OK(c, ldso_stream_sync(c->stream));
    numGoodPoints = *cnt;
    const float quotia = numGoodPoints / (float)(desiredDensity);
    int newSparsity = (I->sparsityFactor * sqrtf(quotia)) + 0.7f;
    if (newSparsity < 1) newSparsity = 1;
    const float oldTHFac = THFac;
    if (newSparsity == 1 && I->sparsityFactor == 1) THFac = 0.5;
    if ((abs(newSparsity - I->sparsityFactor) < 1 && THFac == oldTHFac) || (quotia > 0.8 && 1.0f / quotia > 0.8) || recsLeft == 0) {
        I->sparsityFactor = newSparsity;
        unsigned char* st = (unsigned char*)I->pinned + 64;
        LDSO_CUDA_OK(c, ldso_copy(c, st, I->map_dev, (size_t)w * h, cudaMemcpyDeviceToHost, c->stream));
        LDSO_CUDA_OK(c, ldso_stream_sync(c->stream));
        memcpy(map_host, st, (size_t)w * h);
        return 0;
    }
    I->sparsityFactor = newSparsity;
    return make_pixel_status(c, I, slot, lvl, desiredDensity, map_host, recsLeft - 1, THFac);
}

inline int nblk(int n) { return (n + N1_THREADS - 1) / N1_THREADS; }

// host <-> device movement of single point fields through the pinned staging
int pull(ldso_gpu_ctx* c, InitDev* I, InitLevel& L, std::initializer_list<int> fields, bool with_good, float** out, unsigned char** good_out) {
    float* st = (float*)I->pinned;
    int k = 0;
    for (int fld : fields) {
        LDSO_CUDA_OK(c, ldso_copy(c, st + (size_t)k * L.n, L.F(fld), sizeof(float) * L.n, cudaMemcpyDeviceToHost, c->stream));
        out[k] = st + (size_t)k * L.n;
        k++;
    }
    if (with_good) {
        unsigned char* g = (unsigned char*)(st + (size_t)k * L.n);
        LDSO_CUDA_OK(c, ldso_copy(c, g, L.good, L.n, cudaMemcpyDeviceToHost, c->stream));
        *good_out = g;
    }
    LDSO_CUDA_OK(c, ldso_stream_sync(c->stream));
    return 0;
}

// optReg (:538-567)
int opt_reg(ldso_gpu_ctx* c, InitDev* I, int lvl) {
    InitLevel& L = I->lv[lvl];
    if (L.n == 0) return 0;
    if (!I->snapped) {
        n1_reset_kernel<<<nblk(L.n), N1_THREADS, 0, c->stream>>>(L.n, L.cap, 2, L.f);
        c->launches++;
        return 0;
    }
    float* a[2];
    unsigned char* good;
    int rc = pull(c, I, L, {F_IR, F_ID}, true, a, &good);
    if (rc) return rc;
    float *iR = a[0], *idepth = a[1];
    const int* nb = L.nb_host.data();
    for (int i = 0; i < L.n; i++) {
        if (!good[i]) continue;
        float idnn[10];
        int nnn = 0;
        for (int j = 0; j < 10; j++) {
            const int o = nb[10 * i + j];
            if (o == -1 || !good[o]) continue;
            idnn[nnn++] = iR[o];
        }
        if (nnn > 2) {
            std::nth_element(idnn, idnn + nnn / 2, idnn + nnn);
            iR[i] = (1 - I->regWeight) * idepth[i] + I->regWeight * idnn[nnn / 2];
        }
    }
    LDSO_CUDA_OK(c, ldso_copy(c, L.F(F_IR), iR, sizeof(float) * L.n, cudaMemcpyHostToDevice, c->stream));
    LDSO_CUDA_OK(c, ldso_stream_sync(c->stream));   // the staging is reused
    return 0;
}

// resetPoints (:739-761)
int reset_points(ldso_gpu_ctx* c, InitDev* I, int lvl) {
    InitLevel& L = I->lv[lvl];
    if (L.n == 0) return 0;
    n1_reset_kernel<<<nblk(L.n), N1_THREADS, 0, c->stream>>>(L.n, L.cap, 0, L.f);
    c->launches++;
    if (lvl != I->levels - 1) return 0;
    float* a[1];
    unsigned char* good;
    int rc = pull(c, I, L, {F_IR}, true, a, &good);
    if (rc) return rc;
    float* iR = a[0];
    const int* nb = L.nb_host.data();
    std::vector<int> changed;
    for (int i = 0; i < L.n; i++) {
        if (good[i]) continue;
        float snd = 0, sn = 0;
        for (int n = 0; n < 10; n++) {
            const int o = nb[10 * i + n];
            if (o == -1 || !good[o]) continue;
            snd += iR[o];
            sn += 1;
        }
        if (sn > 0) { good[i] = 1; iR[i] = snd / sn; changed.push_back(i); }
    }
    if (!changed.empty()) {
        // iR = idepth = idepth_new = snd / sn for the revived points: the three rows are rewritten from host copies
        float* a3[2];
        unsigned char* g2;
        std::vector<float> iRv(iR, iR + L.n);
        std::vector<unsigned char> gv(good, good + L.n);
        rc = pull(c, I, L, {F_ID, F_IDN}, false, a3, &g2);
        if (rc) return rc;
        for (int i : changed) { a3[0][i] = iRv[i]; a3[1][i] = iRv[i]; }
        float* st = (float*)I->pinned;
        memcpy(st + 2 * (size_t)L.n, iRv.data(), sizeof(float) * L.n);
        unsigned char* gs = (unsigned char*)(st + 3 * (size_t)L.n);
        memcpy(gs, gv.data(), L.n);
        LDSO_CUDA_OK(c, ldso_copy(c, L.F(F_ID), a3[0], sizeof(float) * L.n, cudaMemcpyHostToDevice, c->stream));
        LDSO_CUDA_OK(c, ldso_copy(c, L.F(F_IDN), a3[1], sizeof(float) * L.n, cudaMemcpyHostToDevice, c->stream));
        LDSO_CUDA_OK(c, ldso_copy(c, L.F(F_IR), st + 2 * (size_t)L.n, sizeof(float) * L.n, cudaMemcpyHostToDevice, c->stream));
        LDSO_CUDA_OK(c, ldso_copy(c, L.good, gs, L.n, cudaMemcpyHostToDevice, c->stream));
        LDSO_CUDA_OK(c, ldso_stream_sync(c->stream));
    }
    return 0;
}

struct GS { float H[64], b[8], Hsc[64], bsc[8], res[3], ec[2]; };

// calcResAndGS (:186-342) + the sums of calcEC
int calc_res_and_gs(ldso_gpu_ctx* c, InitDev* I, int lvl, const SE3& refToNew, double aff_a, double aff_b, GS* out) {
    InitLevel& L = I->lv[lvl];
    ResParams p;
    p.n = L.n; p.cap = L.cap; p.w = I->w[lvl]; p.h = I->h[lvl]; p.npts = L.n;
    double R[9], RKid[9];
    refToNew.rotationMatrix(R);
    ldso_sm::mat3_mul(R, I->Ki[lvl], RKid);
    for (int k = 0; k < 9; k++) p.RKi[k] = (float)RKid[k];
    for (int k = 0; k < 3; k++) p.t[k] = (float)refToNew.t[k];
    p.aff0 = (float)exp(aff_a); p.aff1 = (float)aff_b;
    p.fx = (float)I->fx[lvl]; p.fy = (float)I->fy[lvl]; p.cx = (float)I->cx[lvl]; p.cy = (float)I->cy[lvl];
    p.huberTH = c->cfg.huber_th;
    p.tsq = refToNew.t[0] * refToNew.t[0] + refToNew.t[1] * refToNew.t[1] + refToNew.t[2] * refToNew.t[2];
    p.alphaW = I->alphaW; p.alphaK = I->alphaK; p.couplingWeight = I->couplingWeight;
    p.outlierTH20 = 8 * c->cfg.outlier_th * 20;
    const int blocks = std::max(1, nblk(L.n));
    LDSO_REQUIRE(c, blocks <= I->max_blocks);
    double* result_host_dev = nullptr;
    LDSO_CUDA_OK(c, cudaHostGetDevicePointer((void**)&result_host_dev, I->result_pinned, 0));
    const float4* ref = c->frame[I->first_slot] + c->lvl_off[lvl];
    const float4* nw = c->frame[I->new_slot] + c->lvl_off[lvl];
    n1_residual_kernel<<<blocks, N1_THREADS, 0, c->stream>>>(p, ref, nw, L.f, L.good, L.good_new, L.Jb_new, I->partials, I->counter,
                                                              I->result);
    n1_sc_kernel<<<blocks, N1_THREADS, 0, c->stream>>>(L.n, L.cap, I->couplingWeight, L.f, L.good_new, L.Jb_new, I->partials, I->counter,
                                                        I->result, result_host_dev);
    c->launches += 2;
    LDSO_CUDA_OK(c, cudaGetLastError());
    LDSO_CUDA_OK(c, ldso_stream_sync(c->stream));
    c->d2h_bytes += 128 * sizeof(double);
    const double* r0 = I->result_pinned;
    const double* r1 = I->result_pinned + 64;
    const float alphaOpt = (float)r0[48];
    int q = 0;
    for (int r = 0; r < 9; r++)
        for (int cc = r; cc < 9; cc++, q++) {
            const float v = (float)r0[q], vs = (float)r1[q];
            if (r < 8 && cc < 8) { out->H[8 * r + cc] = out->H[8 * cc + r] = v; out->Hsc[8 * r + cc] = out->Hsc[8 * cc + r] = vs; }
            else if (r < 8) { out->b[r] = v; out->bsc[r] = vs; }
        }
    const int npts = L.n;
    out->H[0] += alphaOpt * npts; out->H[9] += alphaOpt * npts; out->H[18] += alphaOpt * npts;
    double lg[6];
    refToNew.log(lg);
    for (int k = 0; k < 3; k++) out->b[k] += (float)lg[k] * alphaOpt * npts;
    out->res[0] = (float)r0[45]; out->res[1] = (float)r0[49]; out->res[2] = (float)npts;
    out->ec[0] = (float)r1[45]; out->ec[1] = (float)r1[46];
    return 0;
}

}  // namespace

void initializer_free(ldso_gpu_ctx* c) {
    InitDev* I = (InitDev*)c->initializer;
    if (!I) return;
    for (int l = 0; l < LDSO_MAX_LEVELS; l++) level_free(I->lv[l]);
    cudaFree(I->partials); cudaFree(I->counter); cudaFree(I->result); cudaFree(I->map_dev); cudaFree(I->count_dev);
    if (I->result_pinned) cudaFreeHost(I->result_pinned);
    if (I->pinned) cudaFreeHost(I->pinned);
    delete I;
    c->initializer = nullptr;
}

// ---- setFirst (:656-737) --------------------------------------------------------------------------------------------
extern "C" int ldso_gpu_init_set_first(ldso_gpu_ctx* c, int slot, int* num_points_out) {
    LDSO_ENTER(c);
    LDSO_REQUIRE(c, slot >= 0 && slot < c->cfg.n_frame_slots && num_points_out);
    InitDev* I = init_state(c);
    if (!I) return -1;
    make_k(c, I);
    I->first_slot = slot;
    const int w0 = I->w[0], h0 = I->h[0];
    std::vector<float> statusMap((size_t)w0 * h0, 0.f);
    std::vector<unsigned char> statusMapB((size_t)w0 * h0, 0);
    const float densities[] = {0.03f, 0.05f, 0.15f, 0.5f, 1.f};
    std::vector<float> us[LDSO_MAX_LEVELS], vs[LDSO_MAX_LEVELS], types[LDSO_MAX_LEVELS];
    for (int lvl = 0; lvl < I->levels; lvl++) {
        const int wl = I->w[lvl], hl = I->h[lvl];
        if (lvl == 0) {
            int potential = 3, nsel = 0;
            int rc = ldso_gpu_select_make_maps(c, slot, densities[0] * w0 * h0, 1, 2.f, &potential, statusMap.data(), &nsel);
            if (rc) return rc;
        } else {
            int rc = make_pixel_status(c, I, slot, lvl, densities[lvl] * w0 * h0, statusMapB.data());
            if (rc) return rc;
        }
        for (int y = 3; y < hl - 4; y++)                 // patternPadding + 1 .. h - patternPadding - 2 (:690-691)
            for (int x = 3; x < wl - 4; x++)
                if ((lvl != 0 && statusMapB[x + y * wl]) || (lvl == 0 && statusMap[x + y * wl] != 0)) {
                    us[lvl].push_back(x + 0.1); vs[lvl].push_back(y + 0.1);
                    types[lvl].push_back(lvl != 0 ? 1.f : statusMap[x + y * wl]);
                }
        num_points_out[lvl] = (int)us[lvl].size();
    }
    // makeNN (:839-906)
    std::vector<ldso_knn::Grid> grids(I->levels);
    for (int l = 0; l < I->levels; l++) grids[l].build((int)us[l].size(), us[l].data(), vs[l].data(), I->w[l], I->h[l], 8.f);
    for (int l = 0; l < I->levels; l++) {
        InitLevel& L = I->lv[l];
        level_free(L);
        const int n = (int)us[l].size();
        L.n = n; L.cap = std::max(n, 1);
        L.nb_host.assign((size_t)10 * n, -1);
        L.parent_host.assign(n, -1);
        for (int i = 0; i < n; i++) {
            int idx[10]; float dist[10];
            const int found = grids[l].query(us[l][i], vs[l][i], 10, idx, dist);
            for (int k = 0; k < found; k++) L.nb_host[10 * (size_t)i + k] = idx[k];
            if (l < I->levels - 1 && !us[l + 1].empty()) {
                int pi; float pd;
                grids[l + 1].query(us[l][i] * 0.5f - 0.25f, vs[l][i] * 0.5f - 0.25f, 1, &pi, &pd);
                L.parent_host[i] = pi;
            }
        }
        bool ok = true;
        auto da = [&](void** q, size_t bytes) { ok = ok && cudaMalloc(q, bytes) == cudaSuccess; };
        const size_t cap = L.cap;
        da((void**)&L.f, sizeof(float) * F_NUM * cap); da((void**)&L.good, cap); da((void**)&L.good_new, cap);
        da((void**)&L.Jb, sizeof(float) * 10 * cap); da((void**)&L.Jb_new, sizeof(float) * 10 * cap);
        da((void**)&L.nb, sizeof(int) * 10 * cap); da((void**)&L.parent, sizeof(int) * cap);
        if (!ok) { c->set_error("initializer level allocation failed", __FILE__, __LINE__); return -1; }
        // initial state (:695-715): idepth = iR = 1, isGood, energies and Hessians zero
        std::vector<float> f((size_t)F_NUM * cap, 0.f);
        for (int i = 0; i < n; i++) {
            f[F_U * cap + i] = us[l][i]; f[F_V * cap + i] = vs[l][i]; f[F_ID * cap + i] = 1; f[F_IR * cap + i] = 1;
            f[F_TYPE * cap + i] = types[l][i];
        }
        std::vector<int> nbT((size_t)10 * cap, -1);
        for (int i = 0; i < n; i++)
            for (int k = 0; k < 10; k++) nbT[(size_t)k * cap + i] = L.nb_host[10 * (size_t)i + k];
        std::vector<unsigned char> ones(cap, 1);
        LDSO_CUDA_OK(c, cudaMemcpyAsync(L.f, f.data(), sizeof(float) * F_NUM * cap, cudaMemcpyHostToDevice, c->stream));
        LDSO_CUDA_OK(c, cudaMemcpyAsync(L.good, ones.data(), cap, cudaMemcpyHostToDevice, c->stream));
        LDSO_CUDA_OK(c, cudaMemsetAsync(L.good_new, 0, cap, c->stream));
        LDSO_CUDA_OK(c, cudaMemsetAsync(L.Jb, 0, sizeof(float) * 10 * cap, c->stream));
        LDSO_CUDA_OK(c, cudaMemsetAsync(L.Jb_new, 0, sizeof(float) * 10 * cap, c->stream));
        LDSO_CUDA_OK(c, cudaMemcpyAsync(L.nb, nbT.data(), sizeof(int) * 10 * cap, cudaMemcpyHostToDevice, c->stream));
        if (n) LDSO_CUDA_OK(c, cudaMemcpyAsync(L.parent, L.parent_host.data(), sizeof(int) * n, cudaMemcpyHostToDevice, c->stream));
        LDSO_CUDA_OK(c, ldso_stream_sync(c->stream));
        c->h2d_bytes += (long long)(sizeof(float) * F_NUM * cap + cap + sizeof(int) * 11 * cap);
    }
    // children lists for propagateUp: level l + 1 gets the CSR of level l's parent links (children in index order)
    for (int l = 0; l + 1 < I->levels; l++) {
        InitLevel& C = I->lv[l];
        InitLevel& P = I->lv[l + 1];
        std::vector<int> start(P.n + 1, 0), idx(std::max(C.n, 1), 0);
        for (int i = 0; i < C.n; i++) if (C.parent_host[i] >= 0) start[C.parent_host[i] + 1]++;
        for (int i = 0; i < P.n; i++) start[i + 1] += start[i];
        std::vector<int> fill(start.begin(), start.end() - 1);
        for (int i = 0; i < C.n; i++) if (C.parent_host[i] >= 0) idx[fill[C.parent_host[i]]++] = i;
        bool ok = cudaMalloc((void**)&P.child_start, sizeof(int) * (P.n + 1)) == cudaSuccess &&
                  cudaMalloc((void**)&P.child_idx, sizeof(int) * idx.size()) == cudaSuccess;
        if (!ok) { c->set_error("initializer child list allocation failed", __FILE__, __LINE__); return -1; }
        LDSO_CUDA_OK(c, cudaMemcpy(P.child_start, start.data(), sizeof(int) * (P.n + 1), cudaMemcpyHostToDevice));
        LDSO_CUDA_OK(c, cudaMemcpy(P.child_idx, idx.data(), sizeof(int) * idx.size(), cudaMemcpyHostToDevice));
    }
    I->thisToNext = SE3();
    I->aff_a = I->aff_b = 0;
    I->snapped = false;
    I->frameID = I->snappedAt = 0;
    return 0;
}

// ---- trackFrame (:45-183) -------------------------------------------------------------------------------------------
extern "C" int ldso_gpu_init_track_frame(ldso_gpu_ctx* c, int slot, float exposure_first, float exposure_new, double pose_out[12],
                                         float aff_out[2], int* snapped_out, int* ready, float latest_res[3],
                                         int iters_out[LDSO_MAX_LEVELS]) {
    LDSO_ENTER(c);
    InitDev* I = (InitDev*)c->initializer;
    LDSO_REQUIRE(c, I && I->first_slot >= 0 && slot >= 0 && slot < c->cfg.n_frame_slots && pose_out && aff_out && snapped_out && ready);
    I->new_slot = slot;
    const int maxIterations[] = {5, 5, 10, 30, 50};
    I->alphaK = 2.5f * 2.5f; I->alphaW = 150 * 150; I->regWeight = 0.8f; I->couplingWeight = 1;
    const int Lv = I->levels;
    cudaStream_t s = c->stream;
    if (!I->snapped) {
        I->thisToNext.t[0] = I->thisToNext.t[1] = I->thisToNext.t[2] = 0;
        for (int lvl = 0; lvl < Lv; lvl++) {
            InitLevel& L = I->lv[lvl];
            if (L.n) { n1_reset_kernel<<<nblk(L.n), N1_THREADS, 0, s>>>(L.n, L.cap, 1, L.f); c->launches++; }
        }
    }
    SE3 cur = I->thisToNext;
    double aff_a = I->aff_a, aff_b = I->aff_b;
    if (exposure_first > 0 && exposure_new > 0) { aff_a = logf(exposure_new / exposure_first); aff_b = 0; }
    const float wM[8] = {LDSO_SCALE_XI_ROT, LDSO_SCALE_XI_ROT, LDSO_SCALE_XI_ROT, LDSO_SCALE_XI_TRANS, LDSO_SCALE_XI_TRANS,
                         LDSO_SCALE_XI_TRANS, LDSO_SCALE_A, LDSO_SCALE_B};   // (:29-32)
    float* inc_dev = (float*)(I->result + 120);          // 8 floats at the tail of the result block
    float latestRes[3] = {0, 0, 0};
    int rc;
    for (int lvl = Lv - 1; lvl >= 0; lvl--) {
        InitLevel& L = I->lv[lvl];
        if (lvl < Lv - 1) {
            // propagateDown(lvl + 1) (:606-631)
            InitLevel& Pn = I->lv[lvl + 1];
            if (L.n && Pn.n) {
                n1_prop_down_kernel<<<nblk(L.n), N1_THREADS, 0, s>>>(L.n, L.cap, L.f, L.good, L.parent, Pn.cap, Pn.f, Pn.good);
                c->launches++;
            }
            if ((rc = opt_reg(c, I, lvl))) return rc;
        }
        if ((rc = reset_points(c, I, lvl))) return rc;
        GS g, gn;
        if ((rc = calc_res_and_gs(c, I, lvl, cur, aff_a, aff_b, &g))) return rc;
        auto apply_step = [&]() {
            if (L.n) { n1_apply_step_kernel<<<nblk(L.n), N1_THREADS, 0, s>>>(L.n, L.cap, L.f, L.good, L.good_new); c->launches++; }
            std::swap(L.Jb, L.Jb_new);
        };
        apply_step();
        float lambda = 0.1f;
        const float eps = 1e-4f;
        int fails = 0, iteration = 0;
        while (true) {
            float Hl[64], bl[8];
            for (int k = 0; k < 64; k++) Hl[k] = g.H[k];
            for (int k = 0; k < 8; k++) Hl[9 * k] *= (1 + lambda);
            for (int k = 0; k < 64; k++) Hl[k] -= g.Hsc[k] * (1 / (1 + lambda));
            for (int k = 0; k < 8; k++) bl[k] = g.b[k] - g.bsc[k] * (1 / (1 + lambda));
            const float sc = 0.01f / (I->w[lvl] * I->h[lvl]);
            for (int r = 0; r < 8; r++) {
                for (int q = 0; q < 8; q++) Hl[8 * r + q] = wM[r] * Hl[8 * r + q] * wM[q] * sc;
                bl[r] = wM[r] * bl[r] * sc;
            }
            float inc[8];
            const int nd = I->fixAffine ? 6 : 8;
            {
                double Hd[64], bd[8], x[8];
                for (int r = 0; r < nd; r++) { for (int q = 0; q < nd; q++) Hd[nd * r + q] = Hl[8 * r + q]; bd[r] = bl[r]; }
                ldso_sm::ldlt_solve(nd, Hd, bd, x);
                for (int k = 0; k < 8; k++) inc[k] = k < nd ? -(wM[k] * (float)x[k]) : 0.f;
            }
            double incd[6];
            for (int k = 0; k < 6; k++) incd[k] = inc[k];
            const SE3 nw = SE3::exp(incd) * cur;
            const double aff_na = aff_a + inc[6], aff_nb = aff_b + inc[7];
            // doStep (:763-789)
            LDSO_CUDA_OK(c, ldso_copy(c, inc_dev, inc, sizeof(inc), cudaMemcpyHostToDevice, s));
            if (L.n) { n1_do_step_kernel<<<nblk(L.n), N1_THREADS, 0, s>>>(L.n, L.cap, lambda, inc_dev, L.f, L.good, L.Jb); c->launches++; }
            if ((rc = calc_res_and_gs(c, I, lvl, nw, aff_na, aff_nb, &gn))) return rc;
            // calcEC (:520-536)
            const float reg0 = I->snapped ? I->couplingWeight * gn.ec[0] : 0.f, reg1 = I->snapped ? I->couplingWeight * gn.ec[1] : 0.f;
            const float eTotalNew = (gn.res[0] + gn.res[1] + reg1);
            const float eTotalOld = (g.res[0] + g.res[1] + reg0);
            const bool accept = eTotalOld > eTotalNew;
            if (accept) {
                if (gn.res[1] == I->alphaK * (float)L.n) I->snapped = true;
                g = gn;
                aff_a = aff_na; aff_b = aff_nb;
                cur = nw;
                apply_step();
                if ((rc = opt_reg(c, I, lvl))) return rc;
                lambda *= 0.5f;
                fails = 0;
                if (lambda < 0.0001f) lambda = 0.0001f;
            } else {
                fails++;
                lambda *= 4;
                if (lambda > 10000) lambda = 10000;
            }
            float n2 = 0;
            for (int k = 0; k < 8; k++) n2 += inc[k] * inc[k];
            if (!(sqrtf(n2) > eps) || iteration >= maxIterations[lvl] || fails >= 2) break;
            iteration++;
        }
        if (iters_out) iters_out[lvl] = iteration;
        memcpy(latestRes, g.res, sizeof(latestRes));
    }
    I->thisToNext = cur;
    I->aff_a = aff_a; I->aff_b = aff_b;
    for (int i = 0; i < Lv - 1; i++) {
        // propagateUp(i) (:570-604)
        InitLevel& C = I->lv[i];
        InitLevel& P = I->lv[i + 1];
        if (P.n) {
            n1_prop_up_kernel<<<nblk(P.n), N1_THREADS, 0, s>>>(P.n, P.cap, P.f, P.good, P.child_start, P.child_idx, C.cap, C.f, C.good);
            c->launches++;
        }
        if ((rc = opt_reg(c, I, i + 1))) return rc;
    }
    LDSO_CUDA_OK(c, cudaGetLastError());
    LDSO_CUDA_OK(c, ldso_stream_sync(s));
    I->frameID++;
    if (!I->snapped) I->snappedAt = 0;
    if (I->snapped && I->snappedAt == 0) I->snappedAt = I->frameID;
    I->thisToNext.toRt(pose_out);
    aff_out[0] = (float)I->aff_a; aff_out[1] = (float)I->aff_b;
    *snapped_out = I->snapped ? 1 : 0;
    *ready = (I->snapped && I->frameID > I->snappedAt + 5) ? 1 : 0;
    if (latest_res) memcpy(latest_res, latestRes, sizeof(latestRes));
    return 0;
}

extern "C" int ldso_gpu_init_get_points(ldso_gpu_ctx* c, int lvl, int* n, float* u, float* v, float* idepth, float* iR,
                                        unsigned char* is_good, float* my_type, float* last_hessian, float* energy) {
    LDSO_ENTER(c);
    InitDev* I = (InitDev*)c->initializer;
    LDSO_REQUIRE(c, I && lvl >= 0 && lvl < I->levels && n);
    InitLevel& L = I->lv[lvl];
    *n = L.n;
    if (L.n == 0) return 0;
    const int fields[] = {F_U, F_V, F_ID, F_IR, F_TYPE, F_LH, F_E0};
    float* outs[] = {u, v, idepth, iR, my_type, last_hessian, energy};
    float* st = (float*)I->pinned;
    for (int k = 0; k < 7; k++) {
        if (!outs[k]) continue;
        LDSO_CUDA_OK(c, ldso_copy(c, st, L.F(fields[k]), sizeof(float) * L.n, cudaMemcpyDeviceToHost, c->stream));
        LDSO_CUDA_OK(c, ldso_stream_sync(c->stream));
        memcpy(outs[k], st, sizeof(float) * L.n);
    }
    if (is_good) {
        LDSO_CUDA_OK(c, ldso_copy(c, st, L.good, L.n, cudaMemcpyDeviceToHost, c->stream));
        LDSO_CUDA_OK(c, ldso_stream_sync(c->stream));
        memcpy(is_good, st, L.n);
    }
    return 0;
}

extern "C" int ldso_gpu_init_get_graph(ldso_gpu_ctx* c, int lvl, int* neighbours, int* parent) {
    LDSO_ENTER(c);
    InitDev* I = (InitDev*)c->initializer;
    LDSO_REQUIRE(c, I && lvl >= 0 && lvl < I->levels && neighbours && parent);
    InitLevel& L = I->lv[lvl];
    if (L.n) { memcpy(neighbours, L.nb_host.data(), sizeof(int) * 10 * L.n); memcpy(parent, L.parent_host.data(), sizeof(int) * L.n); }
    return 0;
}

extern "C" int ldso_gpu_init_calc_res_and_gs(ldso_gpu_ctx* c, int slot, int lvl, const double ref_to_new[12], float aff_a, float aff_b,
                                             float H[64], float b[8], float Hsc[64], float bsc[8], float res[3]) {
    LDSO_ENTER(c);
    InitDev* I = (InitDev*)c->initializer;
    LDSO_REQUIRE(c, I && I->first_slot >= 0 && lvl >= 0 && lvl < I->levels && slot >= 0 && slot < c->cfg.n_frame_slots);
    I->new_slot = slot;
    I->alphaK = 2.5f * 2.5f; I->alphaW = 150 * 150; I->regWeight = 0.8f; I->couplingWeight = 1;
    GS g;
    int rc = reset_points(c, I, lvl);        // trackFrame's resetPoints(lvl) (:84): idepth_new = idepth (setFirst leaves it unset)
    if (rc) return rc;
    rc = calc_res_and_gs(c, I, lvl, SE3::fromRt(ref_to_new), aff_a, aff_b, &g);
    if (rc) return rc;
    memcpy(H, g.H, sizeof(g.H)); memcpy(b, g.b, sizeof(g.b)); memcpy(Hsc, g.Hsc, sizeof(g.Hsc)); memcpy(bsc, g.bsc, sizeof(g.bsc));
    memcpy(res, g.res, sizeof(g.res));
    return 0;
}
