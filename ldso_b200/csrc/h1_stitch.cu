// h1_stitch.cu — device version of the accumulator stitching the reference does on the host:
// AccumulatedTopHessianSSE::stitchDoubleInternal + the symmetrisation of stitchDoubleMT
// (src/internal/OptimizationBackend/AccumulatedTopHessian.cc:129-255) and
// AccumulatedSCHessianSSE::stitchDoubleInternal + stitchDoubleMT (AccumulatedSCHessian.cc:53-150).
//
// The accumulators already live on the device in fp64, so multiplying them with the adjoints there and reading
// back three (8n+4)^2 systems is cheaper than shipping n^3 8x8 blocks to the host and looping over them in
// scalar code.  One 64-thread CTA per accumulator block; contributions go into the dense systems with fp64
// atomics, a last launch applies priors and the reference's mirror/symmetrise pass.
#include "ctx.h"

#define CPARS 4

namespace {

struct StitchParams {
    int n;
    const double *adH, *adT;      // [n*n][64]
    const double* acc_top;        // [2][n*n][WIN_TOP_STRIDE]
    const int* cnt;               // [2][n*n]
    const double* acc_sc;
    const double* prior;          // cPrior[4] cDelta[4] fprior[8n] fdelta_prior[8n]
    double* out;                  // [3] x (D*D + D): A, L, SC
};

__device__ __forceinline__ double m13(const double* a, int r, int c) {
    // AccumulatorApprox::finish layout (MatrixAccumulators.h:771-800): symmetric 13x13 from 91 packed entries
    if (r > c) { const int t = r; r = c; c = t; }
    if (c < 10) return a[r * 10 - r * (r - 1) / 2 + (c - r)];
    if (r < 10) return a[55 + 3 * r + (c - 10)];
    const int i = r - 10, j = c - 10;   // 85 86 87 / 88 89 / 90
    return a[85 + (i == 0 ? j : i == 1 ? 2 + j : 5)];
}

__global__ void __launch_bounds__(64) stitch_top_kernel(const StitchParams p) {
    const int n = p.n, nn = n * n, D = 8 * n + CPARS;
    const int mode = blockIdx.x / nn, k = blockIdx.x % nn;
    if (p.cnt[mode * nn + k] == 0) return;
    const int h = k % n, t = k / n, hI = CPARS + 8 * h, tI = CPARS + 8 * t;
    const double* a = p.acc_top + (size_t)(mode * nn + k) * WIN_TOP_STRIDE;
    const double* AH = p.adH + 64 * k;
    const double* AT = p.adT + 64 * k;
    double* H = p.out + (size_t)mode * (D * D + D);
    double* b = H + D * D;
    __shared__ double sAH[64], sAT[64], sAHP[64], sATP[64], sM[13][13];
    const int tid = threadIdx.x, i = tid >> 3, j = tid & 7;
    sAH[tid] = AH[tid]; sAT[tid] = AT[tid];
    for (int e = tid; e < 169; e += 64) sM[e / 13][e % 13] = m13(a, e / 13, e % 13);
    __syncthreads();
    double s1 = 0, s2 = 0;
    for (int l = 0; l < 8; l++) { s1 += sAH[i * 8 + l] * sM[CPARS + l][CPARS + j]; s2 += sAT[i * 8 + l] * sM[CPARS + l][CPARS + j]; }
    sAHP[tid] = s1; sATP[tid] = s2;
    __syncthreads();
    double hh = 0, tt = 0, ht = 0;
    for (int l = 0; l < 8; l++) {
        hh += sAHP[i * 8 + l] * sAH[j * 8 + l];
        tt += sATP[i * 8 + l] * sAT[j * 8 + l];
        ht += sAHP[i * 8 + l] * sAT[j * 8 + l];
    }
    atomicAdd(&H[(hI + i) * D + hI + j], hh);
    atomicAdd(&H[(tI + i) * D + tI + j], tt);
    atomicAdd(&H[(hI + i) * D + tI + j], ht);
    if (j < CPARS) {
        double sh = 0, st = 0;
        for (int l = 0; l < 8; l++) { sh += sAH[i * 8 + l] * sM[CPARS + l][j]; st += sAT[i * 8 + l] * sM[CPARS + l][j]; }
        atomicAdd(&H[(hI + i) * D + j], sh);
        atomicAdd(&H[(tI + i) * D + j], st);
    } else if (j == CPARS) {
        double bh = 0, bt = 0;
        for (int l = 0; l < 8; l++) { bh += sAH[i * 8 + l] * sM[CPARS + l][8 + CPARS]; bt += sAT[i * 8 + l] * sM[CPARS + l][8 + CPARS]; }
        atomicAdd(&b[hI + i], bh);
        atomicAdd(&b[tI + i], bt);
    }
    if (tid < 16) atomicAdd(&H[(tid >> 2) * D + (tid & 3)], sM[tid >> 2][tid & 3]);
    else if (tid < 20) atomicAdd(&b[tid - 16], sM[tid - 16][8 + CPARS]);
}

__global__ void __launch_bounds__(64) stitch_sc_kernel(const StitchParams p) {
    const int n = p.n, nn = n * n, D = 8 * n + CPARS;
    const int ij = blockIdx.x % nn, k = blockIdx.x / nn;
    const int i = ij % n, j = ij / n, iI = CPARS + 8 * i, jI = CPARS + 8 * j, kI = CPARS + 8 * k, ik = i + n * k;
    double* H = p.out + (size_t)2 * (D * D + D);
    double* b = H + D * D;
    const double* accD = p.acc_sc;
    const double* accE = accD + (size_t)nn * n * 64;
    const double* accEB = accE + (size_t)nn * 32;
    __shared__ double sAHij[64], sATij[64], sAHik[64], sATik[64], sD[64], sAHD[64], sATD[64], sE[32], sEB[8];
    const int tid = threadIdx.x, r = tid >> 3, cidx = tid & 7;
    const double d = accD[(size_t)(ij + k * nn) * 64 + tid];
    const int any = __syncthreads_or(d != 0.0);
    if (!any && k != 0) return;
    sAHij[tid] = p.adH[64 * ij + tid]; sATij[tid] = p.adT[64 * ij + tid];
    sAHik[tid] = p.adH[64 * ik + tid]; sATik[tid] = p.adT[64 * ik + tid];
    sD[tid] = d;
    if (tid < 32) sE[tid] = accE[(size_t)ij * 32 + tid];
    if (tid < 8) sEB[tid] = accEB[(size_t)ij * 8 + tid];
    __syncthreads();
    if (k == 0) {   // the (i,j) part that does not depend on k: E and EB
        if (cidx < CPARS) {
            double sh = 0, st = 0;
            for (int l = 0; l < 8; l++) { sh += sAHij[r * 8 + l] * sE[l * 4 + cidx]; st += sATij[r * 8 + l] * sE[l * 4 + cidx]; }
            atomicAdd(&H[(iI + r) * D + cidx], sh);
            atomicAdd(&H[(jI + r) * D + cidx], st);
        } else if (cidx == CPARS) {
            double bh = 0, bt = 0;
            for (int l = 0; l < 8; l++) { bh += sAHij[r * 8 + l] * sEB[l]; bt += sATij[r * 8 + l] * sEB[l]; }
            atomicAdd(&b[iI + r], bh);
            atomicAdd(&b[jI + r], bt);
        }
        if (ij == 0) {
            const double* accHcc = accEB + (size_t)nn * 8;
            if (tid >= 40 && tid < 56) atomicAdd(&H[((tid - 40) >> 2) * D + ((tid - 40) & 3)], accHcc[tid - 40]);
            else if (tid >= 56 && tid < 60) atomicAdd(&b[tid - 56], accHcc[16 + tid - 56]);
        }
    }
    if (!any) return;
    double s1 = 0, s2 = 0;
    for (int l = 0; l < 8; l++) { s1 += sAHij[r * 8 + l] * sD[l * 8 + cidx]; s2 += sATij[r * 8 + l] * sD[l * 8 + cidx]; }
    sAHD[tid] = s1; sATD[tid] = s2;
    __syncthreads();
    double ii = 0, jk = 0, ji = 0, ikv = 0;
    for (int l = 0; l < 8; l++) {
        ii += sAHD[r * 8 + l] * sAHik[cidx * 8 + l];
        jk += sATD[r * 8 + l] * sATik[cidx * 8 + l];
        ji += sATD[r * 8 + l] * sAHik[cidx * 8 + l];
        ikv += sAHD[r * 8 + l] * sATik[cidx * 8 + l];
    }
    atomicAdd(&H[(iI + r) * D + iI + cidx], ii);
    atomicAdd(&H[(jI + r) * D + kI + cidx], jk);
    atomicAdd(&H[(jI + r) * D + iI + cidx], ji);
    atomicAdd(&H[(iI + r) * D + kI + cidx], ikv);
}

// priors (mode L only) and the mirror passes; one CTA per system
__global__ void __launch_bounds__(256) stitch_final_kernel(const StitchParams p) {
    const int n = p.n, D = 8 * n + CPARS, sys = blockIdx.x;
    double* H = p.out + (size_t)sys * (D * D + D);
    double* b = H + D * D;
    const int tid = threadIdx.x;
    if (sys == 1) {
        const double *cPrior = p.prior, *cDelta = p.prior + 4, *fprior = p.prior + 8, *fdp = p.prior + 8 + 8 * n;
        if (tid < CPARS) { H[tid * D + tid] += cPrior[tid]; b[tid] += cPrior[tid] * cDelta[tid]; }
        for (int e = tid; e < 8 * n; e += 256) { H[(CPARS + e) * D + CPARS + e] += fprior[e]; b[CPARS + e] += fprior[e] * fdp[e]; }
        __syncthreads();
    }
    // S(a, hI + q) = S(hI + q, a)
    for (int e = tid; e < n * CPARS * 8; e += 256) {
        const int h = e / 32, a = (e >> 3) & 3, q = e & 7;
        H[a * D + CPARS + 8 * h + q] = H[(CPARS + 8 * h + q) * D + a];
    }
    if (sys == 2) return;
    // for t > h: S(hI+i, tI+j) += S(tI+j, hI+i); S(tI+i, hI+j) = S(hI+j, tI+i)
    for (int e = tid; e < n * n * 64; e += 256) {
        const int pair = e >> 6, h = pair % n, t = pair / n, i = (e >> 3) & 7, j = e & 7;
        if (t <= h) continue;
        const int hI = CPARS + 8 * h, tI = CPARS + 8 * t;
        const double v = H[(hI + i) * D + tI + j] + H[(tI + j) * D + hI + i];
        H[(hI + i) * D + tI + j] = v;
        H[(tI + j) * D + hI + i] = v;
    }
}

}  // namespace

int h1_stitch(ldso_gpu_ctx* c) {
    WindowDev& W = c->win;
    const int n = W.n, nn = n * n, D = 8 * n + CPARS;
    StitchParams p;
    p.n = n; p.adH = W.adH_dev; p.adT = W.adT_dev; p.acc_top = W.acc_top; p.cnt = W.acc_top_cnt; p.acc_sc = W.acc_sc;
    p.prior = W.prior_dev; p.out = W.stitched;
    LDSO_CUDA_OK(c, cudaMemsetAsync(W.stitched, 0, sizeof(double) * 3 * (D * D + D), c->stream));
    stitch_top_kernel<<<2 * nn, 64, 0, c->stream>>>(p);
    stitch_sc_kernel<<<nn * n, 64, 0, c->stream>>>(p);
    stitch_final_kernel<<<3, 256, 0, c->stream>>>(p);
    c->launches += 3;
    LDSO_CUDA_OK(c, cudaGetLastError());
    return 0;
}
