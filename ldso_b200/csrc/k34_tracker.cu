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

#include <atomic>
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

__global__ void __launch_bounds__(TRK_THREADS)
k34_track_eval_kernel(const float4* __restrict__ pc, const float4* __restrict__ newImg, const TrkParams p,
                      float* __restrict__ partials, int* __restrict__ partials_i, unsigned int* __restrict__ counter,
                      double* __restrict__ result, double seq) {
    float acc[NF];
#pragma unroll
    for (int k = 0; k < NF; k++) acc[k] = 0.f;
    int cE = 0, cSat = 0, cWarp = 0, cFlow = 0;

    const int i = blockIdx.x * TRK_THREADS + threadIdx.x;
    if (i < p.n) {
        const float4 q = __ldg(pc + i);
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
            acc[46] = ((KuT - x) * (KuT - x) + (KvT - y) * (KvT - y)) + ((KuT2 - x) * (KuT2 - x) + (KvT2 - y) * (KvT2 - y));
            acc[47] = ((Ku - x) * (Ku - x) + (Kv - y) * (Kv - y)) + ((Ku3 - x) * (Ku3 - x) + (Kv3 - y) * (Kv3 - y));
            cFlow = 2;
        }

        if (Ku > 2 && Kv > 2 && Ku < p.w - 3 && Kv < p.h - 3 && new_idepth > 0) {
            float4 hit = lerp_texel(newImg, Ku, Kv, p.w);
            if (isfinite(hit.x)) {
                float residual = hit.x - (p.aff0 * refColor + p.aff1);
                float ar = fabsf(residual);
                float hw = ar < p.huberTH ? 1.f : p.huberTH / ar;
                cE = 1;
                if (ar > p.cutoffTH) {
                    acc[45] = p.maxEnergy;
                    cSat = 1;
                } else {
                    acc[45] = hw * residual * residual * (2 - hw);
                    cWarp = 1;
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
                        for (int cidx = r; cidx < 9; cidx++) acc[k++] = Jw * J[cidx];
                    }
                }
            }
        }
    }

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
        _mm_pause();
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
