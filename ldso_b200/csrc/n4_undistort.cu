// n4_undistort.cu — N4: Undistort::undistort (reference src/frontend/Undistort.cc:366-458) with
// PhotometricUndistorter::processFrame (:190-230) folded in.
//
// One thread per output pixel: look up the source position in remapX / remapY, fetch the four raw 8- or 16-bit neighbours,
// push each through the inverse response G and the inverse vignette (the reference does this for the whole raw image
// first; doing it per fetched neighbour gives the same floats), interpolate with the reference's expression.  The float
// image lands in the staging buffer K1 reads (it stays in L2), so the host sends 1 byte per pixel instead of 4.
#include "ctx.h"

namespace {

struct UndistortDev {
    int w_org = 0, h_org = 0, g_depth = 0, mode = 0;
    bool has_G = false, has_vig = false, has_remap = false;
    float *G = nullptr, *vig = nullptr, *remap_x = nullptr, *remap_y = nullptr;
    unsigned char* raw = nullptr;        // device copy of the raw frame
    unsigned char* raw_pinned = nullptr;
};

struct N4Params {
    int w, h, w_org, h_org, bpp, photometric;   // photometric: 0 factor * raw, 1 G[raw], 2 G[raw] * vignetteMapInv
    float factor;
};

template <int BPP>
__device__ __forceinline__ float source_value(const unsigned char* __restrict__ raw, const float* __restrict__ G,
                                              const float* __restrict__ vig, int idx, const N4Params& p) {
    const unsigned v = BPP == 1 ? (unsigned)raw[idx] : (unsigned)reinterpret_cast<const unsigned short*>(raw)[idx];
    if (p.photometric == 0) return p.factor * (float)v;
    float d = __ldg(G + v);
    if (p.photometric == 2) d *= __ldg(vig + idx);
    return d;
}

template <int BPP>
__global__ void __launch_bounds__(256)
n4_undistort_kernel(const N4Params p, const unsigned char* __restrict__ raw, const float* __restrict__ G, const float* __restrict__ vig,
                    const float* __restrict__ remap_x, const float* __restrict__ remap_y, float* __restrict__ out) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= p.w * p.h) return;
    if (!remap_x) {                                   // passthrough (:452-454)
        out[idx] = source_value<BPP>(raw, G, vig, idx, p);
        return;
    }
    float xx = __ldg(remap_x + idx), yy = __ldg(remap_y + idx);
    if (xx < 0) { out[idx] = 0; return; }
    const int xxi = (int)xx, yyi = (int)yy;
    xx -= xxi; yy -= yyi;
    const float xxyy = xx * yy;
    const int s = xxi + yyi * p.w_org;
    const float s00 = source_value<BPP>(raw, G, vig, s, p), s01 = source_value<BPP>(raw, G, vig, s + 1, p);
    const float s10 = source_value<BPP>(raw, G, vig, s + p.w_org, p), s11 = source_value<BPP>(raw, G, vig, s + 1 + p.w_org, p);
    out[idx] = xxyy * s11 + (yy - xxyy) * s10 + (xx - xxyy) * s01 + (1 - xx - yy + xxyy) * s00;
}

UndistortDev* state(ldso_gpu_ctx* c) {
    if (!c->undistort) c->undistort = new UndistortDev();
    return (UndistortDev*)c->undistort;
}

void release(UndistortDev* U) {
    cudaFree(U->G); cudaFree(U->vig); cudaFree(U->remap_x); cudaFree(U->remap_y); cudaFree(U->raw);
    if (U->raw_pinned) cudaFreeHost(U->raw_pinned);
    *U = UndistortDev();
}

}  // namespace

void undistort_free(ldso_gpu_ctx* c) {
    if (!c->undistort) return;
    release((UndistortDev*)c->undistort);
    delete (UndistortDev*)c->undistort;
    c->undistort = nullptr;
}

extern "C" int ldso_gpu_undistort_set(ldso_gpu_ctx* c, int w_org, int h_org, int g_depth, const float* G, const float* vignette_inv,
                                      const float* remap_x, const float* remap_y, int photometric_mode) {
    LDSO_ENTER(c);
    LDSO_REQUIRE(c, w_org > 0 && h_org > 0 && (G == nullptr || g_depth == 256 || g_depth == 65536));
    LDSO_REQUIRE(c, (remap_x == nullptr) == (remap_y == nullptr));
    LDSO_REQUIRE(c, remap_x != nullptr || (w_org == c->cfg.w && h_org == c->cfg.h));
    UndistortDev* U = state(c);
    release(U);
    U->w_org = w_org; U->h_org = h_org; U->g_depth = g_depth; U->mode = photometric_mode;
    const size_t no = (size_t)w_org * h_org, n = (size_t)c->cfg.w * c->cfg.h;
    auto up = [&](float** dst, const float* src, size_t count) -> int {
        LDSO_CUDA_OK(c, cudaMalloc((void**)dst, sizeof(float) * count));
        LDSO_CUDA_OK(c, cudaMemcpy(*dst, src, sizeof(float) * count, cudaMemcpyHostToDevice));
        c->h2d_bytes += (long long)(sizeof(float) * count);
        return 0;
    };
    int rc = 0;
    if (G) { rc |= up(&U->G, G, g_depth); U->has_G = true; }
    if (vignette_inv) { rc |= up(&U->vig, vignette_inv, no); U->has_vig = true; }
    if (remap_x) { rc |= up(&U->remap_x, remap_x, n); rc |= up(&U->remap_y, remap_y, n); U->has_remap = true; }
    if (rc) return -1;
    LDSO_CUDA_OK(c, cudaMalloc((void**)&U->raw, 2 * no + 16));
    LDSO_CUDA_OK(c, cudaMallocHost((void**)&U->raw_pinned, 2 * no + 16));
    return 0;
}

// wait = false: the (page-locked) raw image stays valid until the tracking stream has been waited for; nothing waits here
static int make_images_raw_impl(ldso_gpu_ctx* c, int slot, const void* raw, int bytes_per_pixel, float exposure_time, float factor, bool wait) {
    UndistortDev* U = (UndistortDev*)c->undistort;
    LDSO_REQUIRE(c, U && U->raw && raw && slot >= 0 && slot < c->cfg.n_frame_slots && (bytes_per_pixel == 1 || bytes_per_pixel == 2));
    LDSO_REQUIRE(c, !(U->has_G && bytes_per_pixel == 2 && U->g_depth != 65536));
    const size_t bytes = (size_t)U->w_org * U->h_org * bytes_per_pixel;
    cudaPointerAttributes attr;
    const bool pinned = cudaPointerGetAttributes(&attr, raw) == cudaSuccess && attr.type == cudaMemoryTypeHost;
    if (!pinned) {
        (void)cudaGetLastError();
        LDSO_CUDA_OK(c, ldso_stream_sync(c->stream));
        memcpy(U->raw_pinned, raw, bytes);
    }
    LDSO_CUDA_OK(c, ldso_copy(c, U->raw, pinned ? raw : (const void*)U->raw_pinned, bytes, cudaMemcpyHostToDevice, c->stream));
    if (pinned && wait) LDSO_CUDA_OK(c, cudaEventRecord(c->ev_h2d, c->stream));
    N4Params p;
    p.w = c->cfg.w; p.h = c->cfg.h; p.w_org = U->w_org; p.h_org = U->h_org; p.bpp = bytes_per_pixel; p.factor = factor;
    // PhotometricUndistorter::processFrame (:197-199): no calibration, no exposure or calibration switched off -> factor * raw
    p.photometric = (!U->has_G || exposure_time <= 0 || U->mode == 0) ? 0 : (U->mode == 2 && U->has_vig ? 2 : 1);
    const int n = p.w * p.h;
    if (bytes_per_pixel == 1)
        n4_undistort_kernel<1><<<(n + 255) / 256, 256, 0, c->stream>>>(p, U->raw, U->G, U->vig, U->remap_x, U->remap_y, c->img_dev);
    else
        n4_undistort_kernel<2><<<(n + 255) / 256, 256, 0, c->stream>>>(p, U->raw, U->G, U->vig, U->remap_x, U->remap_y, c->img_dev);
    c->launches++;
    LDSO_CUDA_OK(c, cudaGetLastError());
    const int rc = k1_make_images(c, slot, c->img_dev);
    if (pinned && wait) LDSO_CUDA_OK(c, ldso_event_sync(c->ev_h2d));
    return rc;
}

extern "C" int ldso_gpu_frame_make_images_raw(ldso_gpu_ctx* c, int slot, const void* raw, int bytes_per_pixel, float exposure_time,
                                              float factor) {
    LDSO_ENTER(c);
    return make_images_raw_impl(c, slot, raw, bytes_per_pixel, exposure_time, factor, true);
}
int n4_make_images_raw_async(ldso_gpu_ctx* c, int slot, const void* raw, int bytes_per_pixel, float exposure_time, float factor) {
    return make_images_raw_impl(c, slot, raw, bytes_per_pixel, exposure_time, factor, false);
}

extern "C" int ldso_gpu_undistort_get_image(ldso_gpu_ctx* c, float* image_out) {
    LDSO_ENTER(c);
    LDSO_REQUIRE(c, image_out);
    const size_t bytes = sizeof(float) * (size_t)c->cfg.w * c->cfg.h;
    LDSO_CUDA_OK(c, ldso_copy(c, c->scratch_host, c->img_dev, bytes, cudaMemcpyDeviceToHost, c->stream));
    LDSO_CUDA_OK(c, ldso_stream_sync(c->stream));
    memcpy(image_out, c->scratch_host, bytes);
    return 0;
}
