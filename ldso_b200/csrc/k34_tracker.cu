// k34_tracker.cu — K3 + K4 fused: CoarseTracker::calcRes (reference src/frontend/CoarseTracker.cc:540-673)
// and CoarseTracker::calcGSSSE (:675-741) + Accumulator9 (MatrixAccumulators.h:1104-1643) in ONE launch.
//
// The reference warps every point of the level's cloud into the new frame, appends the survivors to
// eight SoA "buf_warped_*" streams, and then re-reads those streams to accumulate the 9x9 system.
// Here one thread warps one point, fetches its four float4 texels, and accumulates the 45 weighted
// products, the huber energy and the flow statistics in registers; warp shuffles and a shared-memory
// pass reduce them per CTA in fp32; the last CTA to finish folds the per-CTA partials in fp64 in a
// fixed order (deterministic) and publishes [45 H | E | flowT | flowRT] + 4 integer counts.
// The warped buffers never exist.  Algorithmic traffic: 16 B point + 4x16 B texels per point.
#include "ctx.h"

#include <cstdlib>

#include <algorithm>
#include <atomic>
#include <cooperative_groups.h>
#include <immintrin.h>

namespace {

constexpr int TRK_THREADS = 128;
constexpr int NF = 48;   // float values reduced: 45 H + E + flowT + flowRT
constexpr int NI = 4;    // int values reduced: numTermsInE, numSaturated, numWarped, flowNum

struct TrkParams {
    float RKi[9], t[3], Ki[9];
    float fx, fy, cx, cy;
    int w, h, n, lvl;
    float aff0, aff1;   // affLL (CoarseTracker.cc:556-557)
    float aJ, b0;       // a and lastRef_aff_g2l.b of calcGSSSE (:681-685)
    float huberTH, cutoffTH, maxEnergy;
};

__device__ __forceinline__ float4 lerp_texel(const float4* __restrict__ img, float x, float y, int width) {
    // getInterpolatedElement33 (include/internal/GlobalFuncs.h:89-103), same weight expressions
    int ix = (int)x, iy = (int)y;
    float dx = x - ix, dy = y - iy, dxdy = dx * dy;
    const float4* bp = img + ix + iy * width;
    float4 a = __ldg(bp + 1 + width), b = __ldg(bp + width), c = __ldg(bp + 1), d = __ldg(bp);
    float w0 = dxdy, w1 = dy - dxdy, w2 = dx - dxdy, w3 = 1 - dx - dy + dxdy;
    float4 r;
    r.x = w0 * a.x + w1 * b.x + w2 * c.x + w3 * d.x;
    r.y = w0 * a.y + w1 * b.y + w2 * c.y + w3 * d.y;
    r.z = w0 * a.z + w1 * b.z + w2 * c.z + w3 * d.z;
    r.w = 0.f;
    return r;
}

// one tracker point: calcRes (CoarseTracker.cc:566-661) + the calcGSSSE products (:692-723), added into acc / counters
__device__ __forceinline__ void eval_point(const TrkParams& p, const float4 q, const int i, const float4* __restrict__ newImg,
                                           float (&acc)[NF], int& cE, int& cSat, int& cWarp, int& cFlow) {
    const float x = q.x, y = q.y, id = q.z, refColor = q.w;
    // r = RKi*(x,y,1); pt = r + t*id — same association as the reference's Eigen expression, and the
    // library is compiled with -fmad=false, so every per-point value below is bit-identical to a
    // strict-IEEE evaluation of CoarseTracker.cc:575-585 (threshold tests cannot flip on FMA rounding).
    const float r0 = p.RKi[0] * x + p.RKi[1] * y + p.RKi[2];
    const float r1 = p.RKi[3] * x + p.RKi[4] * y + p.RKi[5];
    const float r2 = p.RKi[6] * x + p.RKi[7] * y + p.RKi[8];
    float pt0 = r0 + p.t[0] * id, pt1 = r1 + p.t[1] * id, pt2 = r2 + p.t[2] * id;
    float u = pt0 / pt2, v = pt1 / pt2;
    float Ku = p.fx * u + p.cx, Kv = p.fy * v + p.cy;
    float new_idepth = id / pt2;

    if (p.lvl == 0 && (i & 31) == 0) {   // CoarseTracker.cc:587-617
        float k0 = p.Ki[0] * x + p.Ki[1] * y + p.Ki[2], k1 = p.Ki[3] * x + p.Ki[4] * y + p.Ki[5],
              k2 = p.Ki[6] * x + p.Ki[7] * y + p.Ki[8];
        float zT = k2 + p.t[2] * id, zT2 = k2 - p.t[2] * id;
        float KuT = p.fx * ((k0 + p.t[0] * id) / zT) + p.cx, KvT = p.fy * ((k1 + p.t[1] * id) / zT) + p.cy;
        float KuT2 = p.fx * ((k0 - p.t[0] * id) / zT2) + p.cx, KvT2 = p.fy * ((k1 - p.t[1] * id) / zT2) + p.cy;
        float z3 = r2 - p.t[2] * id;
        float Ku3 = p.fx * ((r0 - p.t[0] * id) / z3) + p.cx, Kv3 = p.fy * ((r1 - p.t[1] * id) / z3) + p.cy;
        acc[46] += ((KuT - x) * (KuT - x) + (KvT - y) * (KvT - y)) + ((KuT2 - x) * (KuT2 - x) + (KvT2 - y) * (KvT2 - y));
        acc[47] += ((Ku - x) * (Ku - x) + (Kv - y) * (Kv - y)) + ((Ku3 - x) * (Ku3 - x) + (Kv3 - y) * (Kv3 - y));
        cFlow += 2;
    }

    if (Ku > 2 && Kv > 2 && Ku < p.w - 3 && Kv < p.h - 3 && new_idepth > 0) {
        float4 hit = lerp_texel(newImg, Ku, Kv, p.w);
        if (isfinite(hit.x)) {
            float residual = hit.x - (p.aff0 * refColor + p.aff1);
            float ar = fabsf(residual);
            float hw = ar < p.huberTH ? 1.f : p.huberTH / ar;
            cE += 1;
            if (ar > p.cutoffTH) {
                acc[45] += p.maxEnergy;
                cSat += 1;
            } else {
                acc[45] += hw * residual * residual * (2 - hw);
                cWarp += 1;
                // calcGSSSE Jacobian (CoarseTracker.cc:692-723)
                float dx = hit.y * p.fx, dy = hit.z * p.fy;
                float J[9];
                J[0] = new_idepth * dx;
                J[1] = new_idepth * dy;
                J[2] = -(new_idepth * (u * dx + v * dy));
                J[3] = -(u * v * dx + dy * (1 + v * v));
                J[4] = u * v * dy + dx * (1 + u * u);
                J[5] = u * dy - v * dx;
                J[6] = p.aJ * (p.b0 - refColor);
                J[7] = -1.f;
                J[8] = residual;
                int k = 0;
#pragma unroll
                for (int r = 0; r < 9; r++) {
                    float Jw = J[r] * hw;
#pragma unroll
                    for (int cidx = r; cidx < 9; cidx++) acc[k++] += Jw * J[cidx];
                }
            }
        }
    }
}

__global__ void __launch_bounds__(TRK_THREADS)
k34_track_eval_kernel(const float4* __restrict__ pc, const float4* __restrict__ newImg, const TrkParams p,
                      float* __restrict__ partials, int* __restrict__ partials_i, unsigned int* __restrict__ counter,
                      double* __restrict__ result, double seq) {
    float acc[NF];
#pragma unroll
    for (int k = 0; k < NF; k++) acc[k] = 0.f;
    int cE = 0, cSat = 0, cWarp = 0, cFlow = 0;

    const int i = blockIdx.x * TRK_THREADS + threadIdx.x;
    if (i < p.n) eval_point(p, __ldg(pc + i), i, newImg, acc, cE, cSat, cWarp, cFlow);

    // warp reduction (fp32)
#pragma unroll
    for (int k = 0; k < NF; k++) {
        float v = acc[k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        acc[k] = v;
    }
    unsigned packed = (unsigned)cE | ((unsigned)cSat << 8) | ((unsigned)cWarp << 16) | ((unsigned)cFlow << 24);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) packed += __shfl_xor_sync(0xffffffffu, packed, o);   // each field <= 64 < 256

    __shared__ float s_f[TRK_THREADS / 32][NF];
    __shared__ unsigned s_i[TRK_THREADS / 32];
    __shared__ bool s_last;
    __shared__ double s_res[NF + NI];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (lane == 0) {
#pragma unroll
        for (int k = 0; k < NF; k++) s_f[wid][k] = acc[k];
        s_i[wid] = packed;
    }
    __syncthreads();
    if (threadIdx.x < NF) {
        float v = 0.f;
#pragma unroll
        for (int wq = 0; wq < TRK_THREADS / 32; wq++) v += s_f[wq][threadIdx.x];
        partials[blockIdx.x * TRK_NRED + threadIdx.x] = v;
    } else if (threadIdx.x < NF + NI) {
        int f = threadIdx.x - NF;
        int v = 0;
#pragma unroll
        for (int wq = 0; wq < TRK_THREADS / 32; wq++) v += (s_i[wq] >> (8 * f)) & 0xff;
        partials_i[blockIdx.x * NI + f] = v;
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned prev = atomicAdd(counter, 1u);
        s_last = (prev == gridDim.x - 1);
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    // last CTA: fp64 fold over CTAs.  Warp w owns values w, w+4, ...; lanes stride over CTAs, then a fixed
    // shuffle tree — the summation order depends only on gridDim, so results are run-to-run deterministic.
    for (int k = wid; k < NF + NI; k += TRK_THREADS / 32) {
        double s = 0.0;
        if (k < NF) {
            for (unsigned b = lane; b < gridDim.x; b += 32) s += (double)__ldcg(partials + b * TRK_NRED + k);
        } else {
            for (unsigned b = lane; b < gridDim.x; b += 32) s += (double)__ldcg(partials_i + b * NI + (k - NF));
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (lane == 0) s_res[k] = s;
    }
    // `result` is mapped host memory: one warp publishes the sums (a few combined PCIe writes, one system fence),
    // then the sequence number the host is polling for
    __syncthreads();
    if (wid == 0) {
        for (int k = lane; k < NF + NI; k += 32) result[k] = s_res[k];
        __threadfence_system();
        __syncwarp();
        if (lane == 0) {
            *counter = 0u;
            *reinterpret_cast<volatile double*>(result + 63) = seq;
        }
    }
}

}  // namespace

// =====================================================================================================
// Device-resident LM loop of one pyramid level (CoarseTracker::trackNewestCoarse, CoarseTracker.cc:86-249).
// One cooperative launch runs every evaluation of the level: all CTAs evaluate the points, publish per-CTA partial
// sums, meet at ONE grid barrier per evaluation, and then every CTA folds the partials (fp64, fixed order) and runs
// the same scalar LM transition on its thread 0 (accept/reject, 8x8 LDLT, SE3 exponential) — identical inputs give
// identical next parameters in every CTA, so no second barrier or broadcast is needed.  Partials are double-buffered
// by evaluation parity.  The host sees one mailbox message per level instead of one per evaluation.
namespace {

namespace cg = cooperative_groups;

// CTAs of the cooperative launch: the grid barrier and the all-CTA fold both grow with the grid, the evaluation
// itself is a few hundred points per CTA either way
constexpr int LM_MAX_CTAS = 64;   // measured on B200 (configs[1]): 32 -> 198.5 us per frame, 48/64/96 -> 192.5 us
constexpr int LM_THREADS = 128;   // threads per CTA of the device-resident LM kernel (the fold uses its first four warps)

struct LmConst {                // the level being tracked; lives in shared memory, rewritten by thread 0 at each level start
    float Ki[9], fx, fy, cx, cy;
    int w, h, n, lvl, maxIt, solveMode, dbg;   // dbg (cfg.reserved[0], measurement only): 1 = fixed 12 evaluations without the LM transition, 2 = also without the fold
    float huberTH, coarseCutoff, optA, optB;
    float refExposure, newExposure, refAffA, refAffB;
    double q[4], t[3];          // initial refToNew
    float affA, affB;           // initial aff_g2l
    const float4 *pc, *img;     // reference point cloud and new-frame image of this level
};

struct LmLevel {
    float Ki[9], fx, fy, cx, cy;
    int w, h, n, maxIt;
    const int* n_dev;           // if set: the level's point count is read on the device (the host has not waited for it)
    float minResForAbort;
    const float4 *pc, *img;
};
struct LmAll {                  // kernel parameter of the all-levels tracker
    LmConst base;               // everything that does not depend on the level + the initial pose
    LmLevel lv[5];
    int coarsest;
};

struct LmState {                // plain data: lives in shared memory, only thread 0 touches it
    double q[4], t[3];          // current accepted pose
    double cq[4], ct[3];        // candidate pose
    float a, b, ca, cb;
    double H[2][64], bv[2][8], resOld[6], inc[8];   // [hsel] = system at the accepted pose, [hsel ^ 1] = latest evaluation
    float lambda, cutoffRepeat;
    int it, phase, nEval, nAccept, hsel;
    // across levels (CoarseTracker.cc:71-260)
    int lvl, repeated, status;
    float lastRes[5], flow[3];
    double work;                // point evaluations so far (profile units)
    double meanOld, accFlow[3]; // cluster kernel: E / n and the raw flow sums / count of the evaluation the pose stands on
};

__device__ inline ldso_sm::SE3 se3_of(const double* q, const double* t) {
    ldso_sm::SE3 s;
    s.qw = q[0]; s.qx = q[1]; s.qy = q[2]; s.qz = q[3];
    s.t[0] = t[0]; s.t[1] = t[1]; s.t[2] = t[2];
    return s;
}

// the per-evaluation parameters k34_eval builds on the host (CoarseTracker.cc:546-564, 681-685)
__device__ void lm_make_params(const LmConst& k, const double* q, const double* t, float a, float b, float cutoffTH,
                               TrkParams& p) {
    double Rd[9];
    se3_of(q, t).rotationMatrix(Rd);
    float Rf[9];
    for (int i = 0; i < 9; i++) Rf[i] = (float)Rd[i];
    ldso_sm::mat3_mul<float>(Rf, k.Ki, p.RKi);
    for (int i = 0; i < 3; i++) p.t[i] = (float)t[i];
    for (int i = 0; i < 9; i++) p.Ki[i] = k.Ki[i];
    p.fx = k.fx; p.fy = k.fy; p.cx = k.cx; p.cy = k.cy; p.w = k.w; p.h = k.h; p.n = k.n; p.lvl = k.lvl;
    float eF = k.refExposure, eT = k.newExposure;
    if (eF == 0 || eT == 0) eT = eF = 1;
    const float aa = expf(a - k.refAffA) * eT / eF;   // AffLight::fromToVecExposure (include/AffLight.h:27-35)
    p.aff0 = aa; p.aff1 = b - aa * k.refAffB;
    p.aJ = p.aff0; p.b0 = k.refAffB;
    p.huberTH = k.huberTH; p.cutoffTH = cutoffTH;
    p.maxEnergy = 2 * k.huberTH * cutoffTH - k.huberTH * k.huberTH;
}

// LDLT with diagonal pivoting on the leading n x n block, same operation order as ldso_sm::ldlt_solve
__device__ __noinline__ void lm_ldlt(int n, const double* M8, const double* rhs, double* x) {
    double A[64];
    int perm[8];
    for (int i = 0; i < n; i++)
        for (int j = 0; j <= i; j++) A[i * n + j] = A[j * n + i] = M8[i * 8 + j];
    for (int i = 0; i < n; i++) perm[i] = i;
    for (int k = 0; k < n; k++) {
        int p = k;
        double best = fabs(A[k * n + k]);
        for (int i = k + 1; i < n; i++)
            if (fabs(A[i * n + i]) > best) { best = fabs(A[i * n + i]); p = i; }
        if (p != k) {
            for (int j = 0; j < n; j++) { double s = A[k * n + j]; A[k * n + j] = A[p * n + j]; A[p * n + j] = s; }
            for (int j = 0; j < n; j++) { double s = A[j * n + k]; A[j * n + k] = A[j * n + p]; A[j * n + p] = s; }
            int s = perm[k]; perm[k] = perm[p]; perm[p] = s;
        }
        const double d = A[k * n + k];
        if (d == 0) continue;
        for (int i = k + 1; i < n; i++) {
            const double l = A[i * n + k] / d;
            for (int j = k + 1; j <= i; j++) A[i * n + j] -= l * A[k * n + j];
            A[i * n + k] = l;
        }
        for (int i = k + 1; i < n; i++)
            for (int j = k + 1; j < i; j++) A[j * n + i] = A[i * n + j];
    }
    double y[8];
    for (int i = 0; i < n; i++) y[i] = rhs[perm[i]];
    for (int i = 0; i < n; i++)
        for (int j = 0; j < i; j++) y[i] -= A[i * n + j] * y[j];
    for (int i = 0; i < n; i++) {
        const double d = A[i * n + i];
        y[i] = (d == 0) ? 0 : y[i] / d;
    }
    for (int i = n - 1; i >= 0; i--)
        for (int j = i + 1; j < n; j++) y[i] -= A[j * n + i] * y[j];
    for (int i = 0; i < n; i++) x[perm[i]] = y[i];
}

// fixed-affine variants of solve_step (setting_affineOptModeA/B < 0): rare, kept out of line the common case is solved by lm_ldlt8_warp
__device__ __noinline__ void lm_solve_step_fixed(const double* H, const double* b, float lambda, int mode, bool fixA, bool fixB,
                                                 double* inc) {
    double Hl[64], rhs[8];
    for (int i = 0; i < 8; i++) {
        for (int j = 0; j < 8; j++) Hl[i * 8 + j] = (mode == 1 && j < i) ? 0.0 : H[i * 8 + j];
        Hl[i * 8 + i] *= (1 + lambda);
        rhs[i] = -b[i];
    }
    if (fixA && fixB) { lm_ldlt(6, Hl, rhs, inc); inc[6] = inc[7] = 0; }
    else if (!fixA && fixB) { lm_ldlt(7, Hl, rhs, inc); inc[7] = 0; }
    else if (fixA && !fixB) {
        double Hs[64], rs[8], xs[8];
        for (int i = 0; i < 64; i++) Hs[i] = Hl[i];
        for (int i = 0; i < 8; i++) Hs[i * 8 + 6] = Hs[i * 8 + 7];
        for (int j = 0; j < 8; j++) Hs[6 * 8 + j] = Hs[7 * 8 + j];
        for (int i = 0; i < 8; i++) rs[i] = -b[i];
        rs[6] = -b[7];
        lm_ldlt(7, Hs, rs, xs);
        for (int i = 0; i < 6; i++) inc[i] = xs[i];
        inc[6] = 0; inc[7] = xs[6];
    }
}

// fill_vec6 / fill_Hb of api.cu (CoarseTracker.cc:664-672, 725-740) on the folded sums
__device__ void lm_vec6(const double* r, double rs[6]) {
    rs[0] = (float)r[45];
    rs[1] = (int)r[NF + 0];
    const float num = (float)(int)r[NF + 3];
    rs[2] = (float)r[46] / (num + 0.1);
    rs[3] = 0;
    rs[4] = (float)r[47] / (num + 0.1);
    rs[5] = (int)r[NF + 1] / (float)(int)r[NF + 0];
}
// entry t (0..44) of the 9x9 upper triangle, one thread each
__device__ void lm_Hb_entry(const double* r, double* H, double* b, int t) {
    const int n = ((int)r[NF + 2] + 3) & ~3;
    const double invn = (double)(1.0f / n);
    int rr = 0, rem = t;
    while (rem >= 9 - rr) { rem -= 9 - rr; rr++; }
    const int cc = rr + rem;
    auto sc = [](int i) { return i < 3 ? (double)LDSO_SCALE_XI_TRANS : i < 6 ? (double)LDSO_SCALE_XI_ROT : i == 6 ? (double)LDSO_SCALE_A : (double)LDSO_SCALE_B; };
    const double v = r[t] * invn;
    if (rr < 8 && cc < 8) H[rr * 8 + cc] = H[cc * 8 + rr] = v * sc(rr) * sc(cc);
    else if (rr < 8) b[rr] = v * sc(rr);
}

enum { LM_FIRST = 0, LM_ITER = 1, LM_CONT = 0, LM_DONE = 1, LM_SOLVE = 2 };

// ---- the LM transition between two evaluations, split so that the heavy pieces run on many threads ------------
// (1) lm_decide, thread 0: accept / reject, lambda, termination (CoarseTracker.cc:97-104, 213-247).  Hn/bn = the
//     system of the evaluation that just finished, already formed by 45 threads.
// (2) lm_ldlt8_warp, warp 0: the damped 8x8 solve.
// (3) lm_candidate, thread 0: extrapolation, SE3 exponential, the next evaluation's parameters (:118-211).
__device__ int lm_decide(const LmConst& k, LmState& s, const double* r, TrkParams& p) {
    s.nEval++;
    if (s.phase == LM_FIRST) {
        lm_vec6(r, s.resOld);
        if (s.resOld[5] > 0.6 && s.cutoffRepeat < 50) {
            s.cutoffRepeat *= 2;
            lm_make_params(k, s.q, s.t, s.a, s.b, k.coarseCutoff * s.cutoffRepeat, p);
            return LM_CONT;
        }
        s.hsel ^= 1;            // take the new system
        s.lambda = 0.01f;
        s.it = 0;
        s.phase = LM_ITER;
        return s.it >= k.maxIt ? LM_DONE : LM_SOLVE;
    }
    double resNew[6];
    lm_vec6(r, resNew);
    if ((resNew[0] / resNew[1]) < (s.resOld[0] / s.resOld[1])) {
        s.hsel ^= 1;
        for (int i = 0; i < 6; i++) s.resOld[i] = resNew[i];
        for (int i = 0; i < 4; i++) s.q[i] = s.cq[i];
        for (int i = 0; i < 3; i++) s.t[i] = s.ct[i];
        s.a = s.ca; s.b = s.cb;
        s.lambda *= 0.5;
        s.nAccept++;
    } else {
        s.lambda *= 4.0;
        if (s.lambda < 0.001f) s.lambda = 0.001f;
    }
    double nrm = 0;
    for (int i = 0; i < 8; i++) nrm += s.inc[i] * s.inc[i];
    if (!(sqrt(nrm) > 1e-3)) return LM_DONE;
    s.it++;
    return s.it >= k.maxIt ? LM_DONE : LM_SOLVE;
}

// Register-resident 8x8 LDL^T without pivoting, one thread, fully unrolled (all indices static).  The damped tracker
// system is symmetric positive definite, so the factorisation exists without pivoting and agrees with Eigen's pivoted
// LDLT (CoarseTracker.cc:159-179) to fp64 rounding (~1e-13 relative, eight orders below the fp32 noise of H).  A pivot
// that is not safely positive sends the caller to the pivoted warp version below, which mirrors the oracle's
// operation order and handles semi-definite systems.  Dependent chain: 8 x (div, mul, fma) + 16 substitution steps,
// about 1 us against 6 us for the shared-memory warp version.
__device__ __forceinline__ bool lm_ldlt8_fast(const double* H, const double* bv, float lambda, int mode, double* inc) {
    double a[8][8];   // lower triangle used
    double dmax = 0;
#pragma unroll
    for (int i = 0; i < 8; i++)
#pragma unroll
        for (int j = 0; j <= i; j++) {
            double v = (mode == 1 && j < i) ? 0.0 : H[i * 8 + j];
            if (i == j) { v *= (1 + lambda); dmax = fmax(dmax, fabs(v)); }
            a[i][j] = v;
        }
    double y[8], inv[8];
#pragma unroll
    for (int i = 0; i < 8; i++) y[i] = -bv[i];
    bool ok = true;
#pragma unroll
    for (int k = 0; k < 8; k++) {
        const double d = a[k][k];
        ok = ok && (d > 1e-12 * dmax);
        inv[k] = 1.0 / d;
        double l[8];
#pragma unroll
        for (int i = k + 1; i < 8; i++) l[i] = a[i][k] * inv[k];
#pragma unroll
        for (int i = k + 1; i < 8; i++)
#pragma unroll
            for (int j = k + 1; j <= i; j++) a[i][j] -= l[i] * a[j][k];
#pragma unroll
        for (int i = k + 1; i < 8; i++) a[i][k] = l[i];
    }
    if (!ok) return false;
#pragma unroll
    for (int i = 0; i < 8; i++)
#pragma unroll
        for (int j = 0; j < i; j++) y[i] -= a[i][j] * y[j];
#pragma unroll
    for (int i = 0; i < 8; i++) y[i] *= inv[i];
#pragma unroll
    for (int i = 7; i >= 0; i--)
#pragma unroll
        for (int j = i + 1; j < 8; j++) y[i] -= a[j][i] * y[j];
#pragma unroll
    for (int i = 0; i < 8; i++) inc[i] = y[i];
    return true;
}

// LDLT with diagonal pivoting of the damped system on one warp: lane l owns matrix elements l and l + 32 of the 8x8
// array in shared memory.  Every element sees the same operations in the same order as in ldso_sm::ldlt_solve
// (the lower triangle is eliminated, the upper one mirrors it), so the factor is bit-identical to the serial code.
__device__ void lm_ldlt8_warp(const double* H, const double* bv, float lambda, int mode, double* A, double* inc, int lane) {
    const int e0 = lane, e1 = lane + 32;
    const int i0 = e0 >> 3, j0 = e0 & 7, i1 = e1 >> 3, j1 = e1 & 7;
    {   // Hl = H with the diagonal times (1 + lambda) (trap T4: mode 1 drops the strict lower triangle), symmetrised from
        // its lower triangle as Eigen's LDLT reads it
        auto hl = [&](int i, int j) {
            const int r = j <= i ? i : j, c = j <= i ? j : i;   // lower-triangle source (r >= c)
            double v = (mode == 1 && c < r) ? 0.0 : H[r * 8 + c];
            if (r == c) v *= (1 + lambda);
            return v;
        };
        A[e0] = hl(i0, j0); A[e1] = hl(i1, j1);
    }
    int perm = lane < 8 ? lane : 0;   // lane i < 8 holds perm[i]
    __syncwarp();
    for (int kk = 0; kk < 8; kk++) {
        // pivot: first index of the largest |diagonal| among rows kk..7 (lane i looks at row i)
        double best = (lane < 8 && lane >= kk) ? fabs(A[lane * 9]) : -1.0;
        int p = lane;
#pragma unroll
        for (int o = 4; o > 0; o >>= 1) {
            const double ov = __shfl_xor_sync(0xffffffffu, best, o);
            const int op = __shfl_xor_sync(0xffffffffu, p, o);
            if (ov > best || (ov == best && op < p)) { best = ov; p = op; }
        }
        p = __shfl_sync(0xffffffffu, p, 0);
        if (p != kk) {   // uniform
            auto pi = [&](int i) { return i == kk ? p : i == p ? kk : i; };
            const double a0 = A[pi(i0) * 8 + pi(j0)], a1 = A[pi(i1) * 8 + pi(j1)];
            const int pk = __shfl_sync(0xffffffffu, perm, kk), pp = __shfl_sync(0xffffffffu, perm, p);
            if (lane == kk) perm = pp;
            else if (lane == p) perm = pk;
            __syncwarp();
            A[e0] = a0; A[e1] = a1;
            __syncwarp();
        }
        const double d = A[kk * 9];
        if (d == 0) continue;
        // one division per row, on the lane that owns the row index
        const double lrow = (lane < 8 && lane > kk) ? A[lane * 8 + kk] / d : 0.0;
        const double l0 = __shfl_sync(0xffffffffu, lrow, j0 <= i0 ? i0 : j0);
        const double l1 = __shfl_sync(0xffffffffu, lrow, j1 <= i1 ? i1 : j1);
        auto upd = [&](int i, int j, double cur, double l) {
            if (i > kk && j == kk) return l;
            if (i > kk && j > kk && j <= i) return cur - l * A[kk * 8 + j];
            if (i > kk && j > i) return A[j * 8 + i] - l * A[kk * 8 + i];
            return cur;
        };
        const double n0 = upd(i0, j0, A[e0], l0), n1 = upd(i1, j1, A[e1], l1);
        __syncwarp();
        A[e0] = n0; A[e1] = n1;
        __syncwarp();
    }
    // substitutions: lane i < 8 holds y[i]
    double y = 0;
    {
        const int pm = perm;                       // rhs[perm[i]] with rhs = -b
        if (lane < 8) y = -bv[pm];
    }
    for (int j = 0; j < 8; j++) {                  // forward: y[i] -= L[i][j] * y[j], j ascending
        const double yj = __shfl_sync(0xffffffffu, y, j);
        if (lane < 8 && lane > j) y -= A[lane * 8 + j] * yj;
    }
    if (lane < 8) { const double d = A[lane * 9]; y = (d == 0) ? 0 : y / d; }
    for (int i = 7; i >= 0; i--) {                 // backward: y[i] -= sum_{j > i} L[j][i] * y[j], j ascending, i descending
        double acc = __shfl_sync(0xffffffffu, y, i);
        for (int j = i + 1; j < 8; j++) acc -= A[j * 8 + i] * __shfl_sync(0xffffffffu, y, j);
        if (lane == i) y = acc;
    }
    if (lane < 8) inc[perm] = y;
    __syncwarp();
}

__device__ void lm_candidate(const LmConst& k, LmState& s, TrkParams& p) {
    double incS[8];
    const float extrapLimit = 0.001f;
    float extrap = 1;
    if (s.lambda < extrapLimit) extrap = sqrtf(sqrt(extrapLimit / s.lambda));
    double sq = 0;
    for (int i = 0; i < 8; i++) {
        s.inc[i] *= extrap;
        incS[i] = s.inc[i] * (i < 3 ? LDSO_SCALE_XI_TRANS : i < 6 ? LDSO_SCALE_XI_ROT : i == 6 ? LDSO_SCALE_A : LDSO_SCALE_B);
        sq += incS[i] * incS[i];
    }
    if (!isfinite(sq))
        for (int i = 0; i < 8; i++) s.inc[i] = incS[i] = 0;
    const ldso_sm::SE3 cand = ldso_sm::SE3::exp(incS) * se3_of(s.q, s.t);
    s.cq[0] = cand.qw; s.cq[1] = cand.qx; s.cq[2] = cand.qy; s.cq[3] = cand.qz;
    s.ct[0] = cand.t[0]; s.ct[1] = cand.t[1]; s.ct[2] = cand.t[2];
    s.ca = s.a; s.cb = s.b;
    s.ca += incS[6];
    s.cb += incS[7];
    lm_make_params(k, s.cq, s.ct, s.ca, s.cb, k.coarseCutoff * s.cutoffRepeat, p);
}

__global__ void __launch_bounds__(LM_THREADS)
k34_track_level_kernel(const LmAll A, float* __restrict__ partials, int* __restrict__ partials_i, double* __restrict__ mailbox,
                       double seq) {
    cg::grid_group grid = cg::this_grid();
    __shared__ LmConst k;
    __shared__ TrkParams sp;
    __shared__ LmState st;
    __shared__ double s_res[NF + NI];
    __shared__ float s_f[LM_THREADS / 32][NF];
    __shared__ int s_i[LM_THREADS / 32][NI];
    __shared__ double s_A[64];
    __shared__ int s_cmd;
    __shared__ long long s_t[8];   // dbg == 3: SM clocks spent per phase (CTA 0, thread 0)
    long long tk = 0;
    if (threadIdx.x < 8) s_t[threadIdx.x] = 0;
#define LM_TICK(slot) do { if (k.dbg == 3 && threadIdx.x == 0) { const long long now = clock64(); s_t[slot] += now - tk; tk = now; } } while (0)
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int G = gridDim.x;
    if (threadIdx.x == 0) {
        k = A.base;
        for (int i = 0; i < 4; i++) st.q[i] = k.q[i];
        for (int i = 0; i < 3; i++) st.t[i] = k.t[i];
        st.a = k.affA; st.b = k.affB;
        st.nEval = 0; st.nAccept = 0; st.hsel = 0; st.lvl = A.coarsest; st.repeated = 0; st.status = 0; st.work = 0;
        for (int i = 0; i < 5; i++) st.lastRes[i] = __int_as_float(0x7fc00000);
        for (int i = 0; i < 3; i++) st.flow[i] = 1000.f;
    }
    __syncthreads();
    tk = clock64();
    int gev = 0;                 // evaluations since launch: selects the partials buffer across level boundaries too
  while (true) {                 // one pass per pyramid level, coarsest first (CoarseTracker.cc:71)
    if (threadIdx.x == 0) {
        const LmLevel& L = A.lv[st.lvl];
        for (int i = 0; i < 9; i++) k.Ki[i] = L.Ki[i];
        k.fx = L.fx; k.fy = L.fy; k.cx = L.cx; k.cy = L.cy; k.w = L.w; k.h = L.h; k.n = L.n_dev ? *L.n_dev : L.n; k.lvl = st.lvl; k.maxIt = L.maxIt;
        k.pc = L.pc; k.img = L.img;
        st.cutoffRepeat = 1; st.phase = LM_FIRST; st.it = 0; st.lambda = 0.01f;
        for (int i = 0; i < 8; i++) st.inc[i] = 0;
        lm_make_params(k, st.q, st.t, st.a, st.b, k.coarseCutoff, sp);
    }
    __syncthreads();
    const float4* __restrict__ pc = k.pc;
    const float4* __restrict__ newImg = k.img;
    const int npts = k.n;
    for (int ev = 0;; ev++, gev++) {
        float acc[NF];
#pragma unroll
        for (int q = 0; q < NF; q++) acc[q] = 0.f;
        int cnt[NI] = {0, 0, 0, 0};
        for (int i = blockIdx.x * LM_THREADS + threadIdx.x; i < npts; i += G * LM_THREADS)
            eval_point(sp, __ldg(pc + i), i, newImg, acc, cnt[0], cnt[1], cnt[2], cnt[3]);
#pragma unroll
        for (int q = 0; q < NF; q++) {
            float v = acc[q];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            acc[q] = v;
        }
#pragma unroll
        for (int q = 0; q < NI; q++)
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) cnt[q] += __shfl_xor_sync(0xffffffffu, cnt[q], o);
        if (lane == 0) {
#pragma unroll
            for (int q = 0; q < NF; q++) s_f[wid][q] = acc[q];
#pragma unroll
            for (int q = 0; q < NI; q++) s_i[wid][q] = cnt[q];
        }
        __syncthreads();
        float* pf = partials + (size_t)(gev & 1) * G * TRK_NRED;
        int* pi = partials_i + (size_t)(gev & 1) * G * NI;
        if (threadIdx.x < NF) {
            float v = 0.f;
#pragma unroll
            for (int wq = 0; wq < LM_THREADS / 32; wq++) v += s_f[wq][threadIdx.x];
            pf[threadIdx.x * G + blockIdx.x] = v;            // [value][CTA]: the fold reads whole rows
        } else if (threadIdx.x < NF + NI) {
            const int f = threadIdx.x - NF;
            int v = 0;
#pragma unroll
            for (int wq = 0; wq < LM_THREADS / 32; wq++) v += s_i[wq][f];
            pi[f * G + blockIdx.x] = v;
        }
        LM_TICK(0);
        grid.sync();
        LM_TICK(1);
        if (k.dbg == 2) { if (ev == 11) { gev++; break; } continue; }
        // every CTA folds all partials in the same fixed order (lanes stride over CTAs, then a shuffle tree); warp w owns
        // values w, w + 4, ...: 13 independent load streams per lane, so the fold costs about one L2 round trip
        if (wid < 4) {
            double acc13[13];
#pragma unroll
            for (int qq = 0; qq < 13; qq++) acc13[qq] = 0.0;
            for (int b = lane; b < G; b += 32) {
#pragma unroll
                for (int qq = 0; qq < 13; qq++) {
                    const int q = wid + 4 * qq;
                    if (qq < 12) acc13[qq] += (double)__ldcg(pf + q * G + b);
                    else acc13[qq] += (double)__ldcg(pi + (q - NF) * G + b);
                }
            }
#pragma unroll
            for (int qq = 0; qq < 13; qq++) {
                double v = acc13[qq];
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                if (lane == 0) s_res[wid + 4 * qq] = v;
            }
        }
        __syncthreads();
        LM_TICK(2);
        if (k.dbg == 1) { if (ev == 11) { gev++; break; } continue; }
        if (threadIdx.x < 45) lm_Hb_entry(s_res, st.H[st.hsel ^ 1], st.bv[st.hsel ^ 1], threadIdx.x);
        __syncthreads();
        const bool fixed = k.optA < 0 || k.optB < 0;
        if (threadIdx.x == 0) {
            // decide -> solve -> candidate on one thread without a barrier in between (the common case)
            st.work += npts;
            int cmd = lm_decide(k, st, s_res, sp);
            LM_TICK(3);
            if (cmd == LM_SOLVE && !fixed && k.dbg != 4 &&
                lm_ldlt8_fast(st.H[st.hsel], st.bv[st.hsel], st.lambda, k.solveMode, st.inc)) {
                LM_TICK(4);
                lm_candidate(k, st, sp);
                LM_TICK(5);
                cmd = LM_CONT;
            }
            s_cmd = cmd;
        }
        __syncthreads();
        if (s_cmd == LM_DONE) { gev++; break; }
        if (s_cmd == LM_SOLVE) {   // fixed affine parameters, or a pivot the fast factorisation refused
            if (!fixed && wid == 0) lm_ldlt8_warp(st.H[st.hsel], st.bv[st.hsel], st.lambda, k.solveMode, s_A, st.inc, lane);
            if (fixed && threadIdx.x == 0)
                lm_solve_step_fixed(st.H[st.hsel], st.bv[st.hsel], st.lambda, k.solveMode, k.optA < 0, k.optB < 0, st.inc);
            __syncthreads();
            LM_TICK(4);
            if (threadIdx.x == 0) lm_candidate(k, st, sp);
            __syncthreads();
            LM_TICK(5);
        }
    }
    // ---- between levels (CoarseTracker.cc:249-259): residual of the level, abort test, one repeat of a level whose
    // cutoff had to be relaxed; every CTA takes the same decision from the same folded sums
    __syncthreads();
    if (threadIdx.x == 0) {
        const int lvl = st.lvl;
        if (k.dbg != 1 && k.dbg != 2) {
            st.lastRes[lvl] = sqrtf((float)(st.resOld[0] / st.resOld[1]));
            st.flow[0] = (float)st.resOld[2]; st.flow[1] = (float)st.resOld[3]; st.flow[2] = (float)st.resOld[4];
        }
        int next = lvl;
        if ((double)st.lastRes[lvl] > 1.5 * (double)A.lv[lvl].minResForAbort) { st.status = 1; next = -1; }
        else {
            if (st.cutoffRepeat > 1 && !st.repeated) { next++; st.repeated = 1; }
            next--;
        }
        st.lvl = next;
    }
    __syncthreads();
    if (st.lvl < 0) break;
  }
    if (blockIdx.x != 0 || wid != 0) return;
    // message to the host: pose (q, t), affine, lastResiduals[5], flow[3], nEval, nAccept, status, work — then the sequence number
    double v = 0.0;
    if (lane < 4) v = st.q[lane];
    else if (lane < 7) v = st.t[lane - 4];
    else if (lane == 7) v = st.a;
    else if (lane == 8) v = st.b;
    else if (lane < 14) v = st.lastRes[lane - 9];
    else if (lane < 17) v = st.flow[lane - 14];
    else if (lane == 17) v = st.nEval;
    else if (lane == 18) v = st.nAccept;
    else if (lane == 19) v = st.status;
    else if (lane == 20) v = st.work;
    if (lane < 21) mailbox[lane] = v;
    if (k.dbg == 3 && lane < 8) mailbox[24 + lane] = (double)s_t[lane];
    __threadfence_system();
    __syncwarp();
    if (lane == 0) *reinterpret_cast<volatile double*>(mailbox + 63) = seq;
}


// =====================================================================================================
// The same pyramid walk on ONE thread-block cluster (sm_90+: distributed shared memory, hardware cluster barrier).
// CLUSTER CTAs of CL_THREADS threads cover the level's points; per evaluation every CTA reduces its points to 52 values,
// PUSHES them into slot [rank] of every CTA's shared-memory table (st.shared::cluster), and the cluster meets at one
// barrier.cluster; then every CTA folds the table in fp64 in rank order (identical inputs, identical order: every CTA
// holds the same sums, no broadcast) and runs the scalar LM transition.  Against k34_track_level_kernel this removes the
// grid barrier and the round trip of the partials through L2, it is an ordinary (non-cooperative) launch, so the trackers
// of several sequences run side by side on the 148 SMs, and the warp reduction is a transpose-reduce (57 shuffles per
// warp instead of 260).  The table is double-buffered by evaluation parity: a CTA can be at most one evaluation ahead.
constexpr int CL_THREADS = 256;
constexpr int CL_MAX = 16;       // non-portable cluster size (needs cudaFuncAttributeNonPortableClusterSizeAllowed)
constexpr int NV = NF + NI;      // 52

// sum of acc[q] over the warp lands on lane q (q < 32) / lanes q - 32 and q - 16 (32 <= q < 48): each step halves the
// values a lane holds and doubles the lanes summed into them
__device__ __forceinline__ void warp_transpose_reduce48(float (&v)[NF], const int lane, float& out_lo, float& out_hi) {
#pragma unroll
    for (int w = 16; w >= 1; w >>= 1) {
        const bool up = (lane & w) != 0;
#pragma unroll
        for (int i = 0; i < w; i++) {
            const float send = up ? v[i] : v[i + w];
            const float keep = up ? v[i + w] : v[i];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, w);
        }
    }
    out_lo = v[0];
#pragma unroll
    for (int w = 8; w >= 1; w >>= 1) {
        const bool up = (lane & w) != 0;
#pragma unroll
        for (int i = 0; i < w; i++) {
            const float send = up ? v[32 + i] : v[32 + i + w];
            const float keep = up ? v[32 + i + w] : v[32 + i];
            v[32 + i] = keep + __shfl_xor_sync(0xffffffffu, send, w);
        }
    }
    out_hi = v[32] + __shfl_xor_sync(0xffffffffu, v[32], 16);   // lanes l and l ^ 16 both hold value 32 + (l & 15)
}

// ---- branch-free evaluation of one tracker point (same expressions as eval_point; an invalid point adds exact zeros) ----
struct PtProj {
    float u, v, Ku, Kv, new_idepth, refColor;
    bool inb;
};
__device__ __forceinline__ PtProj cl_project(const TrkParams& p, const float4 q, const bool on) {
    PtProj o;
    const float x = q.x, y = q.y, id = q.z;
    const float r0 = p.RKi[0] * x + p.RKi[1] * y + p.RKi[2];
    const float r1 = p.RKi[3] * x + p.RKi[4] * y + p.RKi[5];
    const float r2 = p.RKi[6] * x + p.RKi[7] * y + p.RKi[8];
    const float pt0 = r0 + p.t[0] * id, pt1 = r1 + p.t[1] * id, pt2 = r2 + p.t[2] * id;
    o.u = pt0 / pt2; o.v = pt1 / pt2;
    o.Ku = p.fx * o.u + p.cx; o.Kv = p.fy * o.v + p.cy;
    o.new_idepth = id / pt2;
    o.refColor = q.w;
    o.inb = on && o.Ku > 2 && o.Kv > 2 && o.Ku < p.w - 3 && o.Kv < p.h - 3 && o.new_idepth > 0;
    return o;
}
__device__ __forceinline__ void cl_accumulate(const TrkParams& p, const PtProj& o, const float4 hit, float (&acc)[NF], int& cE, int& cSat,
                                              int& cWarp) {
    const bool fin = o.inb && isfinite(hit.x);
    const float residual = hit.x - (p.aff0 * o.refColor + p.aff1);
    const float ar = fabsf(residual);
    const float hw = ar < p.huberTH ? 1.f : p.huberTH / ar;
    const bool sat = fin && ar > p.cutoffTH, good = fin && !sat;
    cE += fin; cSat += sat; cWarp += good;
    acc[45] += sat ? p.maxEnergy : (good ? hw * residual * residual * (2 - hw) : 0.f);
    const float dx = hit.y * p.fx, dy = hit.z * p.fy;
    float J[9];
    J[0] = o.new_idepth * dx;
    J[1] = o.new_idepth * dy;
    J[2] = -(o.new_idepth * (o.u * dx + o.v * dy));
    J[3] = -(o.u * o.v * dx + dy * (1 + o.v * o.v));
    J[4] = o.u * o.v * dy + dx * (1 + o.u * o.u);
    J[5] = o.u * dy - o.v * dx;
    J[6] = p.aJ * (p.b0 - o.refColor);
    J[7] = -1.f;
    J[8] = residual;
#pragma unroll
    for (int r = 0; r < 9; r++) J[r] = good ? J[r] : 0.f;    // a dropped point may carry NaN / inf: select, do not multiply by 0
    int kk = 0;
#pragma unroll
    for (int r = 0; r < 9; r++) {
        const float Jw = J[r] * hw;
#pragma unroll
        for (int cidx = r; cidx < 9; cidx++) { acc[kk] = __fmaf_rn(Jw, J[cidx], acc[kk]); kk++; }   // a sum across points: fused is fine (and closer)
    }
}
// the flow indicators of CoarseTracker.cc:587-617 (level 0, every 32nd point)
__device__ __forceinline__ void cl_flow(const TrkParams& p, const float4 q, float (&acc)[NF], int& cFlow) {
    const float x = q.x, y = q.y, id = q.z;
    const float r0 = p.RKi[0] * x + p.RKi[1] * y + p.RKi[2];
    const float r1 = p.RKi[3] * x + p.RKi[4] * y + p.RKi[5];
    const float r2 = p.RKi[6] * x + p.RKi[7] * y + p.RKi[8];
    const float pt0 = r0 + p.t[0] * id, pt1 = r1 + p.t[1] * id, pt2 = r2 + p.t[2] * id;
    const float Ku = p.fx * (pt0 / pt2) + p.cx, Kv = p.fy * (pt1 / pt2) + p.cy;
    const float k0 = p.Ki[0] * x + p.Ki[1] * y + p.Ki[2], k1 = p.Ki[3] * x + p.Ki[4] * y + p.Ki[5], k2 = p.Ki[6] * x + p.Ki[7] * y + p.Ki[8];
    const float zT = k2 + p.t[2] * id, zT2 = k2 - p.t[2] * id;
    const float KuT = p.fx * ((k0 + p.t[0] * id) / zT) + p.cx, KvT = p.fy * ((k1 + p.t[1] * id) / zT) + p.cy;
    const float KuT2 = p.fx * ((k0 - p.t[0] * id) / zT2) + p.cx, KvT2 = p.fy * ((k1 - p.t[1] * id) / zT2) + p.cy;
    const float z3 = r2 - p.t[2] * id;
    const float Ku3 = p.fx * ((r0 - p.t[0] * id) / z3) + p.cx, Kv3 = p.fy * ((r1 - p.t[1] * id) / z3) + p.cy;
    acc[46] += ((KuT - x) * (KuT - x) + (KvT - y) * (KvT - y)) + ((KuT2 - x) * (KuT2 - x) + (KvT2 - y) * (KvT2 - y));
    acc[47] += ((Ku - x) * (Ku - x) + (Kv - y) * (Kv - y)) + ((Ku3 - x) * (Ku3 - x) + (Kv3 - y) * (Kv3 - y));
    cFlow += 2;
}

// ---- the scalar LM transition, trimmed for latency (one thread, registers; same decisions as lm_decide / lm_candidate) ----
// 1 / d to fp64 rounding from the hardware's 20-bit estimate and two Newton steps (the division routine costs ~3x as much
// and sits on the dependent chain of the factorisation eight times)
__device__ __forceinline__ double cl_rcp(double d) {
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(d));
    r = __fma_rn(__fma_rn(-d, r, 1.0), r, r);
    r = __fma_rn(__fma_rn(-d, r, 1.0), r, r);
    return r;
}
// damped 8x8 solve, register resident, no pivoting (see lm_ldlt8_fast), reciprocals by cl_rcp
__device__ __forceinline__ bool cl_ldlt8(const double* H, const double* bv, float lambda, int mode, double* inc) {
    double a[8][8];
    double dmax = 0;
#pragma unroll
    for (int i = 0; i < 8; i++)
#pragma unroll
        for (int j = 0; j <= i; j++) {
            double v = (mode == 1 && j < i) ? 0.0 : H[i * 8 + j];
            if (i == j) { v *= (1 + lambda); dmax = fmax(dmax, fabs(v)); }
            a[i][j] = v;
        }
    double y[8], inv[8];
#pragma unroll
    for (int i = 0; i < 8; i++) y[i] = -bv[i];
    bool ok = true;
#pragma unroll
    for (int k = 0; k < 8; k++) {
        const double d = a[k][k];
        ok = ok && (d > 1e-12 * dmax);
        inv[k] = cl_rcp(d);
        double l[8];
#pragma unroll
        for (int i = k + 1; i < 8; i++) l[i] = a[i][k] * inv[k];
#pragma unroll
        for (int i = k + 1; i < 8; i++)
#pragma unroll
            for (int j = k + 1; j <= i; j++) a[i][j] = __fma_rn(-l[i], a[j][k], a[i][j]);
#pragma unroll
        for (int i = k + 1; i < 8; i++) a[i][k] = l[i];
    }
    if (!ok) return false;
#pragma unroll
    for (int i = 0; i < 8; i++)
#pragma unroll
        for (int j = 0; j < i; j++) y[i] = __fma_rn(-a[i][j], y[j], y[i]);
#pragma unroll
    for (int i = 0; i < 8; i++) y[i] *= inv[i];
#pragma unroll
    for (int i = 7; i >= 0; i--)
#pragma unroll
        for (int j = i + 1; j < 8; j++) y[i] = __fma_rn(-a[j][i], y[j], y[i]);
#pragma unroll
    for (int i = 0; i < 8; i++) inc[i] = y[i];
    return true;
}
// SE3::exp for the small rotations of an LM increment: the four coefficient functions as even power series in theta
// (|theta| < 0.25: the 9th term is below 1e-19), exact functions beyond
__device__ __forceinline__ ldso_sm::SE3 cl_se3_exp(const double* a) {
    const double* om = a + 3;
    const double th2 = om[0] * om[0] + om[1] * om[1] + om[2] * om[2];
    if (!(th2 < 0.0625)) return ldso_sm::SE3::exp(a);
    // sin(th/2)/th, cos(th/2), (1 - cos th)/th^2, (th - sin th)/th^3
    double im = 0, re = 0, A = 0, B = 0;
    {
        // coefficients (-1)^k / (2^(2k+1) (2k+1)!), (-1)^k / (4^k (2k)!), (-1)^k / (2k+2)!, (-1)^k / (2k+3)!
        const double ci[9] = {0.5, -1.0 / 48, 1.0 / 3840, -1.0 / 645120, 1.0 / 185794560, -1.0 / 81749606400.0, 1.0 / 51011754393600.0,
                              -1.0 / 42849873690624000.0, 1.0 / 46620662575398912000.0};
        const double cr[9] = {1.0, -1.0 / 8, 1.0 / 384, -1.0 / 46080, 1.0 / 10321920, -1.0 / 3715891200.0, 1.0 / 1961990553600.0,
                              -1.0 / 1428329123020800.0, 1.0 / 1371195958099968000.0};
        const double cA[9] = {1.0 / 2, -1.0 / 24, 1.0 / 720, -1.0 / 40320, 1.0 / 3628800, -1.0 / 479001600, 1.0 / 87178291200.0,
                              -1.0 / 20922789888000.0, 1.0 / 6402373705728000.0};
        const double cB[9] = {1.0 / 6, -1.0 / 120, 1.0 / 5040, -1.0 / 362880, 1.0 / 39916800, -1.0 / 6227020800.0, 1.0 / 1307674368000.0,
                              -1.0 / 355687428096000.0, 1.0 / 121645100408832000.0};
#pragma unroll
        for (int k = 8; k >= 0; k--) {
            im = __fma_rn(im, th2, ci[k]); re = __fma_rn(re, th2, cr[k]);
            A = __fma_rn(A, th2, cA[k]); B = __fma_rn(B, th2, cB[k]);
        }
    }
    ldso_sm::SE3 r;
    r.qw = re; r.qx = im * om[0]; r.qy = im * om[1]; r.qz = im * om[2];
    double O[9], O2[9], V[9];
    ldso_sm::hat(om, O);
    ldso_sm::mat3_mul(O, O, O2);
    for (int i = 0; i < 9; i++) V[i] = A * O[i] + B * O2[i];
    V[0] += 1; V[4] += 1; V[8] += 1;
    ldso_sm::mat3_vec(V, a, r.t);
    return r;
}

// upper-triangle index t (0..44) of the 9x9 system -> (row, column)
__constant__ unsigned char c_tri_r[45] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 1, 1, 1, 1, 2, 2, 2, 2, 2, 2, 2, 3, 3, 3, 3, 3, 3, 4, 4, 4, 4, 4, 5, 5, 5, 5, 6, 6, 6, 7, 7, 8};
__constant__ unsigned char c_tri_c[45] = {0, 1, 2, 3, 4, 5, 6, 7, 8, 1, 2, 3, 4, 5, 6, 7, 8, 2, 3, 4, 5, 6, 7, 8, 3, 4, 5, 6, 7, 8, 4, 5, 6, 7, 8, 5, 6, 7, 8, 6, 7, 8, 7, 8, 8};
__device__ __forceinline__ double cl_scale(int i) {
    return i < 3 ? (double)LDSO_SCALE_XI_TRANS : i < 6 ? (double)LDSO_SCALE_XI_ROT : i == 6 ? (double)LDSO_SCALE_A : (double)LDSO_SCALE_B;
}
// lm_decide with one division per evaluation: the accepted mean energy is kept, the flow indicators and the saturation
// ratio are formed only where they are read (CoarseTracker.cc:94-99, 213-247, 664-672)
__device__ __forceinline__ int cl_decide(const LmConst& k, LmState& s, const double* r, TrkParams& p) {
    s.nEval++;
    const double E = (float)r[45], nE = (int)r[NF + 0];
    auto take = [&]() { s.meanOld = E / nE; s.accFlow[0] = r[46]; s.accFlow[1] = r[47]; s.accFlow[2] = r[NF + 3]; };
    if (s.phase == LM_FIRST) {
        const double satRatio = (int)r[NF + 1] / (float)(int)r[NF + 0];
        if (satRatio > 0.6 && s.cutoffRepeat < 50) {
            s.cutoffRepeat *= 2;
            lm_make_params(k, s.q, s.t, s.a, s.b, k.coarseCutoff * s.cutoffRepeat, p);
            return LM_CONT;
        }
        take();
        s.hsel ^= 1;
        s.lambda = 0.01f;
        s.it = 0;
        s.phase = LM_ITER;
        return s.it >= k.maxIt ? LM_DONE : LM_SOLVE;
    }
    if (E / nE < s.meanOld) {
        take();
        s.hsel ^= 1;
        for (int i = 0; i < 4; i++) s.q[i] = s.cq[i];
        for (int i = 0; i < 3; i++) s.t[i] = s.ct[i];
        s.a = s.ca; s.b = s.cb;
        s.lambda *= 0.5;
        s.nAccept++;
    } else {
        s.lambda *= 4.0;
        if (s.lambda < 0.001f) s.lambda = 0.001f;
    }
    double nrm = 0;
    for (int i = 0; i < 8; i++) nrm += s.inc[i] * s.inc[i];
    if (!(nrm > 1e-6)) return LM_DONE;     // !(inc.norm() > 1e-3)
    s.it++;
    return s.it >= k.maxIt ? LM_DONE : LM_SOLVE;
}
__device__ __forceinline__ void cl_candidate(const LmConst& k, LmState& s, TrkParams& p) {   // lm_candidate over cl_se3_exp
    double incS[8], inc[8];
    const float extrapLimit = 0.001f;
    float extrap = 1;
    if (s.lambda < extrapLimit) extrap = sqrtf(sqrt(extrapLimit / s.lambda));
    double sq = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) {
        inc[i] = s.inc[i] * extrap;
        incS[i] = inc[i] * (i < 3 ? LDSO_SCALE_XI_TRANS : i < 6 ? LDSO_SCALE_XI_ROT : i == 6 ? LDSO_SCALE_A : LDSO_SCALE_B);
        sq += incS[i] * incS[i];
    }
    if (!isfinite(sq))
        for (int i = 0; i < 8; i++) inc[i] = incS[i] = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) s.inc[i] = inc[i];
    const ldso_sm::SE3 cand = cl_se3_exp(incS) * se3_of(s.q, s.t);
    s.cq[0] = cand.qw; s.cq[1] = cand.qx; s.cq[2] = cand.qy; s.cq[3] = cand.qz;
    s.ct[0] = cand.t[0]; s.ct[1] = cand.t[1]; s.ct[2] = cand.t[2];
    s.ca = s.a; s.cb = s.b;
    s.ca += incS[6];
    s.cb += incS[7];
    lm_make_params(k, s.cq, s.ct, s.ca, s.cb, k.coarseCutoff * s.cutoffRepeat, p);
}

constexpr int CL_SLOTS = 4;      // points per thread staged in shared memory per level (beyond: streamed from global memory)

__global__ void __launch_bounds__(CL_THREADS)
k34_track_cluster_kernel(const LmAll A, double* __restrict__ mailbox, double seq) {
    cg::cluster_group cluster = cg::this_cluster();
    const int C = (int)cluster.num_blocks(), rank = (int)cluster.block_rank();
    __shared__ LmConst k;
    __shared__ TrkParams sp;
    __shared__ LmState st;
    __shared__ double s_res[NV];
    __shared__ float s_f[CL_THREADS / 32][NV];
    __shared__ float s_own[NV];
    __shared__ float s_part[2][CL_MAX][NV];      // [evaluation parity][source CTA][value]; counts travel as exact floats (< 2^24)
    __shared__ float4 s_pc[CL_SLOTS][CL_THREADS];
    __shared__ double s_A[64];
    __shared__ int s_cmd;
    __shared__ long long s_t[8];
    __shared__ unsigned char s_tri[48];          // (row << 4) | column of the 45 upper-triangle entries (lane-varying index: not constant memory)
    long long tk = 0;
    if (threadIdx.x < 8) s_t[threadIdx.x] = 0;
    if (threadIdx.x < 45) s_tri[threadIdx.x] = (unsigned char)((c_tri_r[threadIdx.x] << 4) | c_tri_c[threadIdx.x]);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (threadIdx.x == 0) {
        k = A.base;
        for (int i = 0; i < 4; i++) st.q[i] = k.q[i];
        for (int i = 0; i < 3; i++) st.t[i] = k.t[i];
        st.a = k.affA; st.b = k.affB;
        st.nEval = 0; st.nAccept = 0; st.hsel = 0; st.lvl = A.coarsest; st.repeated = 0; st.status = 0; st.work = 0;
        for (int i = 0; i < 5; i++) st.lastRes[i] = __int_as_float(0x7fc00000);
        for (int i = 0; i < 3; i++) st.flow[i] = 1000.f;
    }
    cluster.sync();              // every CTA of the cluster is resident before anyone writes into a neighbour's shared memory
    tk = clock64();
    int gev = 0;
    const int stride = C * CL_THREADS, first = rank * CL_THREADS + threadIdx.x;
  while (true) {                 // one pass per pyramid level, coarsest first (CoarseTracker.cc:71)
    if (threadIdx.x == 0) {
        const LmLevel& L = A.lv[st.lvl];
        for (int i = 0; i < 9; i++) k.Ki[i] = L.Ki[i];
        k.fx = L.fx; k.fy = L.fy; k.cx = L.cx; k.cy = L.cy; k.w = L.w; k.h = L.h; k.n = L.n_dev ? *L.n_dev : L.n; k.lvl = st.lvl; k.maxIt = L.maxIt;
        k.pc = L.pc; k.img = L.img;
        st.cutoffRepeat = 1; st.phase = LM_FIRST; st.it = 0; st.lambda = 0.01f;
        for (int i = 0; i < 8; i++) st.inc[i] = 0;
        lm_make_params(k, st.q, st.t, st.a, st.b, k.coarseCutoff, sp);
    }
    __syncthreads();
    const float4* __restrict__ pc = k.pc;
    const float4* __restrict__ newImg = k.img;
    const int npts = k.n, lvl0 = k.lvl == 0;
    // this thread's points of the level stay in shared memory for all its evaluations
#pragma unroll
    for (int j = 0; j < CL_SLOTS; j++) {
        const int i = first + j * stride;
        s_pc[j][threadIdx.x] = i < npts ? __ldg(pc + i) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    for (int ev = 0;; ev++, gev++) {
        float acc[NF];
#pragma unroll
        for (int q = 0; q < NF; q++) acc[q] = 0.f;
        int cnt[NI] = {0, 0, 0, 0};
        const TrkParams P = sp;      // registers / uniform
#pragma unroll
        for (int j = 0; j < CL_SLOTS; j += 2) {
            // two points in flight: projections, then eight texel loads, then the products
            const int i0 = first + j * stride, i1 = i0 + stride;
            const int w0 = i0 - lane;                        // first point of this warp in slot j (warps own 32 consecutive points)
            if (w0 >= npts) break;                           // warp-uniform: this and every later slot is empty for the warp
            const float4 q0 = s_pc[j][threadIdx.x];
            const PtProj o0 = cl_project(P, q0, i0 < npts);
            if (w0 + stride < npts) {                        // warp-uniform: the second slot of the pair has points too
                const float4 q1 = s_pc[j + 1][threadIdx.x];
                const PtProj o1 = cl_project(P, q1, i1 < npts);
                const float4 h0 = lerp_texel(newImg, o0.inb ? o0.Ku : 3.f, o0.inb ? o0.Kv : 3.f, P.w);
                const float4 h1 = lerp_texel(newImg, o1.inb ? o1.Ku : 3.f, o1.inb ? o1.Kv : 3.f, P.w);
                cl_accumulate(P, o0, h0, acc, cnt[0], cnt[1], cnt[2]);
                cl_accumulate(P, o1, h1, acc, cnt[0], cnt[1], cnt[2]);
                if (lvl0 && (i1 & 31) == 0 && i1 < npts) cl_flow(P, q1, acc, cnt[3]);
            } else {
                const float4 h0 = lerp_texel(newImg, o0.inb ? o0.Ku : 3.f, o0.inb ? o0.Kv : 3.f, P.w);
                cl_accumulate(P, o0, h0, acc, cnt[0], cnt[1], cnt[2]);
            }
            if (lvl0 && (i0 & 31) == 0 && i0 < npts) cl_flow(P, q0, acc, cnt[3]);
        }
        for (int i = first + CL_SLOTS * stride; i < npts; i += stride)    // clouds beyond CL_SLOTS points per thread
            eval_point(P, __ldg(pc + i), i, newImg, acc, cnt[0], cnt[1], cnt[2], cnt[3]);
        float lo, hi;
        warp_transpose_reduce48(acc, lane, lo, hi);
        // counts: two 16-bit fields per word (a warp sees at most 32 x points-per-thread of them)
        unsigned c01 = (unsigned)cnt[0] | ((unsigned)cnt[1] << 16), c23 = (unsigned)cnt[2] | ((unsigned)cnt[3] << 16);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { c01 += __shfl_xor_sync(0xffffffffu, c01, o); c23 += __shfl_xor_sync(0xffffffffu, c23, o); }
        s_f[wid][lane] = lo;
        if (lane < 16) s_f[wid][32 + lane] = hi;
        if (lane == 0) {
            s_f[wid][NF + 0] = (float)(c01 & 0xffffu); s_f[wid][NF + 1] = (float)(c01 >> 16);
            s_f[wid][NF + 2] = (float)(c23 & 0xffffu); s_f[wid][NF + 3] = (float)(c23 >> 16);
        }
        __syncthreads();
        if (threadIdx.x < NV) {
            float v = 0.f;
#pragma unroll
            for (int wq = 0; wq < CL_THREADS / 32; wq++) v += s_f[wq][threadIdx.x];
            s_own[threadIdx.x] = v;
        }
        __syncthreads();
        float (*tab)[NV] = s_part[gev & 1];
        for (int idx = threadIdx.x; idx < NV * C; idx += CL_THREADS) {
            const int dst = idx / NV, q = idx - dst * NV;
            float* remote = cluster.map_shared_rank(&tab[rank][q], dst);
            *remote = s_own[q];
        }
        if (k.dbg == 3 && threadIdx.x == 0) { const long long now = clock64(); s_t[0] += now - tk; tk = now; }
        cluster.sync();          // barrier.cluster arrive.release / wait.acquire: the pushed partials are visible
        if (k.dbg == 3 && threadIdx.x == 0) { const long long now = clock64(); s_t[1] += now - tk; tk = now; }
        if (threadIdx.x < NV) {
            // fold in rank order; the threads of the 45 H / b entries also fold the point count their entry is divided by
            // (CoarseTracker.cc:724-726), so the scaled system is written here, without a second pass over s_res
            double v = 0.0, cn = 0.0;
            for (int r = 0; r < C; r++) { v += (double)tab[r][threadIdx.x]; cn += (double)tab[r][NF + 2]; }
            s_res[threadIdx.x] = v;
            if (threadIdx.x < 45) {
                const int n = ((int)cn + 3) & ~3;
                const double w = v * (double)(1.0f / n);
                const int rr = s_tri[threadIdx.x] >> 4, cc = s_tri[threadIdx.x] & 15;
                double* H = st.H[st.hsel ^ 1];
                if (rr < 8 && cc < 8) H[rr * 8 + cc] = H[cc * 8 + rr] = w * cl_scale(rr) * cl_scale(cc);
                else if (rr < 8) st.bv[st.hsel ^ 1][rr] = w * cl_scale(rr);
            }
        }
        __syncthreads();
        if (k.dbg == 3 && threadIdx.x == 0) { const long long now = clock64(); s_t[2] += now - tk; tk = now; }
        const bool fixed = k.optA < 0 || k.optB < 0;
        if (threadIdx.x == 0) {
            st.work += npts;
            int cmd = cl_decide(k, st, s_res, sp);
            if (k.dbg == 3) { const long long now = clock64(); s_t[3] += now - tk; tk = now; }
            if (cmd == LM_SOLVE && !fixed && cl_ldlt8(st.H[st.hsel], st.bv[st.hsel], st.lambda, k.solveMode, st.inc)) {
                if (k.dbg == 3) { const long long now = clock64(); s_t[4] += now - tk; tk = now; }
                cl_candidate(k, st, sp);
                if (k.dbg == 3) { const long long now = clock64(); s_t[5] += now - tk; tk = now; }
                cmd = LM_CONT;
            }
            s_cmd = cmd;
        }
        __syncthreads();
        if (s_cmd == LM_DONE) { gev++; break; }
        if (s_cmd == LM_SOLVE) {   // fixed affine parameters, or a pivot the fast factorisation refused
            if (!fixed && wid == 0) lm_ldlt8_warp(st.H[st.hsel], st.bv[st.hsel], st.lambda, k.solveMode, s_A, st.inc, lane);
            if (fixed && threadIdx.x == 0)
                lm_solve_step_fixed(st.H[st.hsel], st.bv[st.hsel], st.lambda, k.solveMode, k.optA < 0, k.optB < 0, st.inc);
            __syncthreads();
            if (threadIdx.x == 0) cl_candidate(k, st, sp);
            __syncthreads();
        }
    }
    // between levels (CoarseTracker.cc:249-259), as in k34_track_level_kernel: every CTA takes the same decision
    __syncthreads();
    if (threadIdx.x == 0) {
        const int lvl = st.lvl;
        st.lastRes[lvl] = sqrtf((float)st.meanOld);
        const float num = (float)(int)st.accFlow[2];           // lm_vec6: rs[2], rs[3], rs[4]
        st.flow[0] = (float)((float)st.accFlow[0] / (num + 0.1)); st.flow[1] = 0.f; st.flow[2] = (float)((float)st.accFlow[1] / (num + 0.1));
        int next = lvl;
        if ((double)st.lastRes[lvl] > 1.5 * (double)A.lv[lvl].minResForAbort) { st.status = 1; next = -1; }
        else {
            if (st.cutoffRepeat > 1 && !st.repeated) { next++; st.repeated = 1; }
            next--;
        }
        st.lvl = next;
    }
    __syncthreads();
    if (st.lvl < 0) break;
  }
    if (rank != 0 || wid != 0) return;
    double v = 0.0;
    if (lane < 4) v = st.q[lane];
    else if (lane < 7) v = st.t[lane - 4];
    else if (lane == 7) v = st.a;
    else if (lane == 8) v = st.b;
    else if (lane < 14) v = st.lastRes[lane - 9];
    else if (lane < 17) v = st.flow[lane - 14];
    else if (lane == 17) v = st.nEval;
    else if (lane == 18) v = st.nAccept;
    else if (lane == 19) v = st.status;
    else if (lane == 20) v = st.work;
    if (lane < 21) mailbox[lane] = v;
    if (k.dbg == 3 && lane < 8) mailbox[24 + lane] = (double)s_t[lane];
    __threadfence_system();
    __syncwarp();
    if (lane == 0) *reinterpret_cast<volatile double*>(mailbox + 63) = seq;
}

}  // namespace

// include/AffLight.h:27-35 (float arithmetic, returned as double in the reference)
static inline void from_to_vec_exposure(float exposureF, float exposureT, float aF, float bF, float aT, float bT, float* a_out,
                                        float* b_out) {
    if (exposureF == 0 || exposureT == 0) exposureT = exposureF = 1;
    float a = exp(aT - aF) * exposureT / exposureF;
    float b = bT - a * bF;
    *a_out = a;
    *b_out = b;
}

int k34_eval(ldso_gpu_ctx* c, int tracker, int lvl, const ldso_sm::SE3& refToNew, float aff_a, float aff_b, float cutoffTH,
             TrackEvalOut* out) {
    TrackerDev& T = c->trk[tracker];
    if (trk_pc_n_sync(c, T)) return -1;
    TrkParams p;
    double Rd[9];
    refToNew.rotationMatrix(Rd);
    float Rf[9];
    for (int i = 0; i < 9; i++) Rf[i] = (float)Rd[i];
    ldso_sm::mat3_mul<float>(Rf, T.Ki[lvl], p.RKi);   // CoarseTracker.cc:553
    for (int i = 0; i < 3; i++) p.t[i] = (float)refToNew.t[i];
    for (int i = 0; i < 9; i++) p.Ki[i] = T.Ki[lvl][i];
    p.fx = T.fx[lvl]; p.fy = T.fy[lvl]; p.cx = T.cx[lvl]; p.cy = T.cy[lvl];
    p.w = T.w[lvl]; p.h = T.h[lvl]; p.n = T.pc_n[lvl]; p.lvl = lvl;
    from_to_vec_exposure(T.ref_exposure, T.new_exposure, T.ref_aff_a, T.ref_aff_b, aff_a, aff_b, &p.aff0, &p.aff1);
    p.aJ = p.aff0;
    p.b0 = T.ref_aff_b;
    p.huberTH = c->cfg.huber_th;
    p.cutoffTH = cutoffTH;
    p.maxEnergy = 2 * p.huberTH * cutoffTH - p.huberTH * p.huberTH;   // :563-564
    volatile double* res = c->trk_mailbox;
    if (p.n <= 0) {
        memset(out, 0, sizeof(*out));
        return 0;
    }
    int blocks = (p.n + TRK_THREADS - 1) / TRK_THREADS;
    if (blocks > c->trk_max_blocks) { c->set_error("tracker point cloud too large", __FILE__, __LINE__); return -2; }
    prof_begin(c, c->stream);
    k34_track_eval_kernel<<<blocks, TRK_THREADS, 0, c->stream>>>(T.pc[lvl], c->frame[T.new_slot] + c->lvl_off[lvl], p,
                                                                 c->trk_partials, c->trk_partials_i, c->trk_counter,
                                                                 c->trk_mailbox_dev, (double)(++c->trk_seq));
    c->launches++;
    prof_end(c, c->stream, LDSO_PROF_K34, p.n);
    LDSO_CUDA_OK(c, cudaGetLastError());
    // the kernel's last CTA writes the 52 sums and then the sequence number straight into mapped pinned memory:
    // poll for it instead of paying a D2H copy + stream synchronisation per LM evaluation
    const double want = (double)c->trk_seq;
    for (unsigned spin = 1; res[63] != want; spin++) {
        ldso_cpu_relax();
        if ((spin & 0xfffff) == 0) {   // every ~1M polls make sure the stream is still alive
            cudaError_t q = cudaStreamQuery(c->stream);
            if (q != cudaSuccess && q != cudaErrorNotReady) LDSO_CUDA_OK(c, q);
            if (q == cudaSuccess && res[63] != want) { c->set_error("tracker kernel finished without publishing", __FILE__, __LINE__); return -1; }
        }
    }
    std::atomic_thread_fence(std::memory_order_acquire);
    c->d2h_bytes += (NF + NI + 1) * sizeof(double);
    for (int k = 0; k < NF; k++) out->sums[k] = res[k];
    for (int k = 0; k < NI; k++) out->counts[k] = (int)res[NF + k];
    return 0;
}

// the device-resident tracker publishes one message per frame into the mapped pinned mailbox: poll it (no D2H copy, no sync)
static int track_all_wait(ldso_gpu_ctx* c, double seq, int prof_idx, int dbg, ldso_sm::SE3& cur, float& aff_a, float& aff_b,
                          float last_residuals[5], float flow[3], int* status, int* nEval, int* nAccept) {
    volatile double* res = c->trk_mailbox;
    if (ldso_wait_block()) LDSO_CUDA_OK(c, ldso_event_sync(c->ev_trk));   // sleep until the kernel is done instead of polling
    for (unsigned spin = 1; res[63] != seq; spin++) {
        ldso_cpu_relax();
        if ((spin & 0xfffff) == 0) {
            cudaError_t q = cudaStreamQuery(c->stream);
            if (q != cudaSuccess && q != cudaErrorNotReady) LDSO_CUDA_OK(c, q);
            if (q == cudaSuccess && res[63] != seq) { c->set_error("tracker kernel finished without publishing", __FILE__, __LINE__); return -1; }
        }
    }
    std::atomic_thread_fence(std::memory_order_acquire);
    c->d2h_bytes += 22 * sizeof(double);
    cur.qw = res[0]; cur.qx = res[1]; cur.qy = res[2]; cur.qz = res[3];
    cur.t[0] = res[4]; cur.t[1] = res[5]; cur.t[2] = res[6];
    aff_a = (float)res[7]; aff_b = (float)res[8];
    for (int i = 0; i < 5; i++) last_residuals[i] = (float)res[9 + i];
    for (int i = 0; i < 3; i++) flow[i] = (float)res[14 + i];
    *nEval = (int)res[17]; *nAccept = (int)res[18]; *status = (int)res[19];
    if (prof_idx >= 0) c->prof_pending[prof_idx].units = res[20];   // point evaluations of this launch
    if (dbg == 3) for (int i = 0; i < 8; i++) c->lm_dbg_clocks[i] += res[24 + i];
    return 0;
}

// trackNewestCoarse's level loop on the device (see k34_track_level_kernel): ONE cooperative launch walks the pyramid
// from coarsest_lvl down, the host polls the mapped mailbox once.  In/out: pose, affine.
int k34_track_all_wait(ldso_gpu_ctx* c, ldso_sm::SE3& cur, float& aff_a, float& aff_b, float last_residuals[5], float flow[3], int* status,
                       int* nEval, int* nAccept) {
    const auto& P = c->trk_pending;
    return track_all_wait(c, P.seq, P.prof_idx, P.dbg, cur, aff_a, aff_b, last_residuals, flow, status, nEval, nAccept);
}

int k34_track_all_launch(ldso_gpu_ctx* c, int tracker, int coarsest_lvl, const int* maxIt, const float* min_res_for_abort,
                         const ldso_sm::SE3& cur, float aff_a, float aff_b) {
    TrackerDev& T = c->trk[tracker];
    LmAll A;
    memset(&A, 0, sizeof(A));
    LmConst& k = A.base;
    k.solveMode = c->cfg.tracker_solve_mode;
    k.huberTH = c->cfg.huber_th; k.coarseCutoff = c->cfg.coarse_cutoff_th;
    k.optA = c->cfg.affine_opt_mode_a; k.optB = c->cfg.affine_opt_mode_b;
    k.refExposure = T.ref_exposure; k.newExposure = T.new_exposure; k.refAffA = T.ref_aff_a; k.refAffB = T.ref_aff_b;
    k.q[0] = cur.qw; k.q[1] = cur.qx; k.q[2] = cur.qy; k.q[3] = cur.qz;
    k.t[0] = cur.t[0]; k.t[1] = cur.t[1]; k.t[2] = cur.t[2];
    k.affA = aff_a; k.affB = aff_b;
    k.dbg = c->cfg.reserved[0];
    A.coarsest = coarsest_lvl;
    int nmax = 1;
    const bool n_on_device = c->cfg.tracker_lm_mode != 2;   // the cluster kernel's grid does not depend on the point counts
    if (!n_on_device && trk_pc_n_sync(c, T)) return -1;
    for (int lvl = 0; lvl <= coarsest_lvl; lvl++) {
        LmLevel& L = A.lv[lvl];
        for (int i = 0; i < 9; i++) L.Ki[i] = T.Ki[lvl][i];
        L.fx = T.fx[lvl]; L.fy = T.fy[lvl]; L.cx = T.cx[lvl]; L.cy = T.cy[lvl];
        L.w = T.w[lvl]; L.h = T.h[lvl]; L.maxIt = maxIt[lvl];
        L.n = n_on_device ? 0 : T.pc_n[lvl]; L.n_dev = n_on_device ? T.pc_n_dev + lvl : nullptr;
        L.minResForAbort = min_res_for_abort[lvl];
        L.pc = T.pc[lvl];
        L.img = c->frame[T.new_slot] + c->lvl_off[lvl];
        nmax = std::max(nmax, L.n);
    }
    if (c->cfg.tracker_lm_mode != 2) {
        // one thread-block cluster per tracker (ordinary launch: trackers of other sequences run beside it)
        if (c->trk_cluster == 0) {
            // 8 CTAs (the portable cluster size): measured on B200, one sequence tracks as fast on 8 CTAs as on 16 (the
            // evaluation is latency-, not throughput-bound), and 8-CTA clusters of several sequences pack better into the
            // GPCs (8 sequences: 15.2 k frames/s against 12.6 k with 16-CTA clusters)
            int want = 8;
            if (const char* e = getenv("LDSO_LM_CLUSTER")) want = atoi(e);   // tuning knob: 1, 2, 4, 8 or 16 CTAs
            LDSO_CUDA_OK(c, cudaFuncSetAttribute(k34_track_cluster_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
            for (; want >= 1; want >>= 1) {
                cudaLaunchConfig_t q = {};
                cudaLaunchAttribute at[1];
                at[0].id = cudaLaunchAttributeClusterDimension;
                at[0].val.clusterDim.x = want; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
                q.gridDim = dim3(want); q.blockDim = dim3(CL_THREADS); q.attrs = at; q.numAttrs = 1;
                int nclusters = 0;
                if (cudaOccupancyMaxActiveClusters(&nclusters, k34_track_cluster_kernel, &q) == cudaSuccess && nclusters > 0) break;
                cudaGetLastError();
            }
            if (want < 1) { c->set_error("no cluster size of the tracker kernel fits this device", __FILE__, __LINE__); return -1; }
            c->trk_cluster = want;
        }
        cudaLaunchConfig_t q = {};
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = c->trk_cluster; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        q.gridDim = dim3(c->trk_cluster); q.blockDim = dim3(CL_THREADS); q.stream = c->stream; q.attrs = at; q.numAttrs = 1;
        double* mbx = c->trk_mailbox_dev;
        const double sq = (double)(++c->trk_seq);
        prof_begin(c, c->stream);
        LDSO_CUDA_OK(c, cudaLaunchKernelEx(&q, k34_track_cluster_kernel, A, mbx, sq));
        c->launches++;
        c->trk_pending.seq = sq; c->trk_pending.prof_idx = prof_end(c, c->stream, LDSO_PROF_K34, 0); c->trk_pending.dbg = k.dbg;
        if (ldso_wait_block()) LDSO_CUDA_OK(c, cudaEventRecord(c->ev_trk, c->stream));
        return 0;
    }
    if (c->trk_coop_max == 0) {
        int per_sm = 0, sms = 0, dev = 0, coop = 0;
        LDSO_CUDA_OK(c, cudaGetDevice(&dev));
        LDSO_CUDA_OK(c, cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev));
        if (!coop) { c->set_error("device lacks cooperative launch", __FILE__, __LINE__); return -1; }
        LDSO_CUDA_OK(c, cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        LDSO_CUDA_OK(c, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k34_track_level_kernel, LM_THREADS, 0));
        int cap = LM_MAX_CTAS;
        if (const char* e = getenv("LDSO_LM_CTAS")) cap = atoi(e);   // tuning knob
        c->trk_coop_max = std::max(1, std::min(per_sm * sms, cap));
    }
    int blocks = std::max(1, (nmax + LM_THREADS - 1) / LM_THREADS);   // sized for the finest level; coarser levels leave CTAs idle
    blocks = std::min(blocks, std::min(c->trk_coop_max, c->trk_max_blocks / 2));
    float* pf = c->trk_partials;
    int* pi = c->trk_partials_i;
    double* mb = c->trk_mailbox_dev;
    double seq = (double)(++c->trk_seq);
    void* args[] = {(void*)&A, (void*)&pf, (void*)&pi, (void*)&mb, (void*)&seq};
    prof_begin(c, c->stream);
    LDSO_CUDA_OK(c, cudaLaunchCooperativeKernel((void*)k34_track_level_kernel, dim3(blocks), dim3(LM_THREADS), args, 0, c->stream));
    c->launches++;
    c->trk_pending.seq = seq; c->trk_pending.prof_idx = prof_end(c, c->stream, LDSO_PROF_K34, 0); c->trk_pending.dbg = k.dbg;
    if (ldso_wait_block()) LDSO_CUDA_OK(c, cudaEventRecord(c->ev_trk, c->stream));
    return 0;
}

// =====================================================================================================
// Allocation-free variant of the fused K3+K4 evaluation (ldso_gpu_tracker_calc_res_noalloc, include/ldso_gpu.h).
// No counterpart in the reference: added for callers that must not allocate
// (CUDA-graph capture, tight serving loops).  The kernel is k34_track_eval_kernel unchanged; its reduction scratch
// (per-CTA fp32 partials, integer partials, the arrival counter and the 64-double result block) lives in ONE file-scope
// device array instead of the ctx's cudaMalloc'ed workspace.  Every GPU that loads the library holds the array until
// the process exits.  All calls on a device share it: calls that overlap on the GPU race.
#define NOALLOC_MAX_CTAS (LDSO_NOALLOC_MAX_POINTS / TRK_THREADS)
namespace {
constexpr size_t NA_OFF_PART = 0;                                                     // float [NOALLOC_MAX_CTAS][TRK_NRED]
constexpr size_t NA_OFF_PART_I = NA_OFF_PART + sizeof(float) * NOALLOC_MAX_CTAS * TRK_NRED;   // int [NOALLOC_MAX_CTAS][NI]
constexpr size_t NA_OFF_RESULT = NA_OFF_PART_I + sizeof(int) * NOALLOC_MAX_CTAS * NI;         // double [64]
constexpr size_t NA_OFF_COUNTER = NA_OFF_RESULT + sizeof(double) * 64;                        // unsigned
constexpr size_t NA_BYTES = ((NA_OFF_COUNTER + sizeof(unsigned) + 255) / 256) * 256;
static_assert(NA_OFF_PART_I % 16 == 0 && NA_OFF_RESULT % 16 == 0 && NA_OFF_COUNTER % 4 == 0, "scratch carve-up alignment");
}  // namespace
static __device__ __align__(256) unsigned char g_noalloc_scratch[NA_BYTES];

// In host code `g_noalloc_scratch` is only a stand-in for the symbol, not a device address: the address is resolved once
// per device, with that device current, and cached.
static int noalloc_scratch(ldso_gpu_ctx* c, unsigned char** out) {
    constexpr int MAXDEV = 64;
    static std::atomic<unsigned char*> cache[MAXDEV];
    int dev = -1;
    LDSO_CUDA_OK(c, cudaGetDevice(&dev));
    if (dev != c->cfg.device) { c->set_error("the ctx's device is not the calling thread's current device", __FILE__, __LINE__); return -2; }
    if (dev < 0 || dev >= MAXDEV) { c->set_error("device ordinal out of range", __FILE__, __LINE__); return -2; }
    unsigned char* p = cache[dev].load(std::memory_order_acquire);
    if (!p) {
        void* sym = nullptr;
        LDSO_CUDA_OK(c, cudaGetSymbolAddress(&sym, g_noalloc_scratch));
        if (!sym) { c->set_error("cudaGetSymbolAddress returned a null address", __FILE__, __LINE__); return -1; }
        p = static_cast<unsigned char*>(sym);
        cache[dev].store(p, std::memory_order_release);   // two racing first calls resolve the same address
    }
    *out = p;
    return 0;
}

extern "C" int ldso_gpu_tracker_calc_res_noalloc(ldso_gpu_ctx* c, int tracker, int new_slot, float new_exposure, int lvl,
                                                 const double ref_to_new[12], float aff_a, float aff_b, float cutoff_th,
                                                 float out6[6], double sums52[52]) {
    LDSO_ENTER(c);
    LDSO_REQUIRE(c, tracker >= 0 && tracker < LDSO_NUM_TRACKERS && lvl >= 0 && lvl < c->levels);
    LDSO_REQUIRE(c, new_slot >= 0 && new_slot < c->cfg.n_frame_slots && ref_to_new && out6);
    if (trk_pc_n_sync(c, c->trk[tracker])) return -1;   // (a wait, not an allocation)
    const TrackerDev& T = c->trk[tracker];
    LDSO_REQUIRE(c, T.ref_slot >= 0);
    const int n = T.pc_n[lvl];
    if (n > LDSO_NOALLOC_MAX_POINTS) {
        char msg[160];
        snprintf(msg, sizeof(msg), "ldso_gpu_tracker_calc_res_noalloc: %d tracker points exceed LDSO_NOALLOC_MAX_POINTS = %d", n,
                 LDSO_NOALLOC_MAX_POINTS);
        c->set_error(msg, __FILE__, __LINE__);
        return -2;
    }
    double res[NF + NI];
    for (int k = 0; k < NF + NI; k++) res[k] = 0.0;
    if (n > 0) {
        unsigned char* base = nullptr;
        int rc = noalloc_scratch(c, &base);
        if (rc) return rc;
        float* partials = reinterpret_cast<float*>(base + NA_OFF_PART);
        int* partials_i = reinterpret_cast<int*>(base + NA_OFF_PART_I);
        double* result = reinterpret_cast<double*>(base + NA_OFF_RESULT);
        unsigned* counter = reinterpret_cast<unsigned*>(base + NA_OFF_COUNTER);
        // same parameter block as k34_eval (CoarseTracker.cc:546-564, 681-685), without touching the tracker's state
        TrkParams p;
        double Rd[9];
        float Rf[9];
        ldso_sm::SE3::fromRt(ref_to_new).rotationMatrix(Rd);
        for (int i = 0; i < 9; i++) Rf[i] = (float)Rd[i];
        ldso_sm::mat3_mul<float>(Rf, T.Ki[lvl], p.RKi);
        for (int i = 0; i < 3; i++) p.t[i] = (float)ref_to_new[9 + i];
        for (int i = 0; i < 9; i++) p.Ki[i] = T.Ki[lvl][i];
        p.fx = T.fx[lvl]; p.fy = T.fy[lvl]; p.cx = T.cx[lvl]; p.cy = T.cy[lvl];
        p.w = T.w[lvl]; p.h = T.h[lvl]; p.n = n; p.lvl = lvl;
        from_to_vec_exposure(T.ref_exposure, new_exposure, T.ref_aff_a, T.ref_aff_b, aff_a, aff_b, &p.aff0, &p.aff1);
        p.aJ = p.aff0;
        p.b0 = T.ref_aff_b;
        p.huberTH = c->cfg.huber_th;
        p.cutoffTH = cutoff_th;
        p.maxEnergy = 2 * p.huberTH * cutoff_th - p.huberTH * p.huberTH;
        const int blocks = (n + TRK_THREADS - 1) / TRK_THREADS;
        // the array keeps the previous call's contents: the arrival counter and the result block must start at zero
        LDSO_CUDA_OK(c, cudaMemsetAsync(result, 0, sizeof(double) * 64 + sizeof(unsigned), c->stream));
        k34_track_eval_kernel<<<blocks, TRK_THREADS, 0, c->stream>>>(T.pc[lvl], c->frame[new_slot] + c->lvl_off[lvl], p, partials,
                                                                     partials_i, counter, result, 1.0);
        c->launches++;
        LDSO_CUDA_OK(c, cudaGetLastError());
        // read-back through the ctx's pinned 64-double block (allocated in ldso_gpu_create): no allocation here either
        LDSO_CUDA_OK(c, ldso_copy(c, c->trk_result_pinned, result, sizeof(double) * 64, cudaMemcpyDeviceToHost, c->stream));
        LDSO_CUDA_OK(c, ldso_stream_sync(c->stream));
        if (c->trk_result_pinned[63] != 1.0) { c->set_error("tracker kernel finished without publishing", __FILE__, __LINE__); return -1; }
        for (int k = 0; k < NF + NI; k++) res[k] = c->trk_result_pinned[k];
    }
    // CoarseTracker.cc:664-672, as fill_vec6 in api.cu
    const float num = (float)(int)res[NF + 3];
    out6[0] = (float)res[45];
    out6[1] = (float)(int)res[NF + 0];
    out6[2] = (float)((float)res[46] / (num + 0.1));
    out6[3] = 0;
    out6[4] = (float)((float)res[47] / (num + 0.1));
    out6[5] = (float)(double)((int)res[NF + 1] / (float)(int)res[NF + 0]);
    if (sums52) for (int k = 0; k < NF + NI; k++) sums52[k] = res[k];
    return 0;
}
