// window.h — device-resident sliding window (points, residuals, accumulators) of one sequence.
// Internal to libldso_gpu.so; the public face is the ldso_gpu_window_* group of include/ldso_gpu.h.
#pragma once
#include <cuda_runtime.h>
#include <vector>

#include "../../include/ldso_gpu.h"

#define WIN_TOP_STRIDE 96    // 91 unique AccumulatorApprox entries (55 + 30 + 6), padded
#define WIN_PT_ROWS 16       // per-point output rows, see PT_* below

// rows of WindowDev::pt_out ([row][cap_p])
enum { PT_HDD_A = 0, PT_BD_A = 1, PT_HCD_A = 2 /*..5*/, PT_HDD_L = 6, PT_BD_L = 7, PT_HCD_L = 8 /*..11*/, PT_HDI = 12,
       PT_BDSUM = 13, PT_IDH = 14, PT_STEP = 15 };

struct WindowDev {
    int n = 0, np = 0, nr = 0;
    int cap_p = 0, cap_r = 0;
    int slots[LDSO_MAX_WINDOW];
    float frameEnergyTH[LDSO_MAX_WINDOW];
    float fxl = 0, fyl = 0, cxl = 0, cyl = 0, fxli = 0, fyli = 0;
    // device: frames
    float* precalc = nullptr;      // [n*n][32]
    // pools the per-upload arrays below are carved from (one H2D copy / one memset per upload): see carve_points / carve_residuals
    float* pool_p = nullptr;       // points: u v idepth idepth_zero prior host res_begin color[8] weights[8], stride q = np rounded up
    int* pool_ri = nullptr;        // residual ints: r_target r_point state flags new_state, stride r = nr rounded up
    float* pool_rz = nullptr;      // residual floats that start at zero: energy new_energy new_energy_wo rtz[8] JpJdF[8] center[4] res_pt[8] J[72]
    // device: points (SoA)
    float *u = nullptr, *v = nullptr, *idepth = nullptr, *idepth_zero = nullptr, *prior = nullptr;
    float *color = nullptr, *weights = nullptr;   // [cap_p][8]
    int *host = nullptr, *res_begin = nullptr;
    float* pt_out = nullptr;       // [WIN_PT_ROWS][cap_p]
    float* max_rel_baseline = nullptr;   // [cap_p] (atomicMax on the bit pattern: values are >= 0)
    float* num_good = nullptr;     // [cap_p]
    // device: residuals
    int *r_point = nullptr, *r_target = nullptr, *state = nullptr, *new_state = nullptr, *flags = nullptr;
    float *energy = nullptr, *new_energy = nullptr, *new_energy_wo = nullptr;
    float* J = nullptr;            // [cap_r][72]  AoS record = RawResidualJacobian
    float* rtz = nullptr;          // [cap_r][8]   res_toZeroF
    float* JpJdF = nullptr;        // [cap_r][8]
    float* center = nullptr;       // [cap_r][4]   centerProjectedTo (+pad)
    float* res_pt = nullptr;       // [cap_r][8]   per-residual Hdd, bd, Hcd[4] contribution + mode tag
    int* order = nullptr;          // residual ids sorted by accumulator key (mode * n^2 + host + n * target)
    int* chunk = nullptr;          // [n_chunks][3] = key, start, count
    int n_chunks = 0;
    int chunk_len = 128;           // longest chunk in the list: 128 (thread per residual) or up to 1024 (k8_top_stream_kernel)
    int* sc_blocks = nullptr;      // [n_sc_blocks][3] = host, first point, end point (K9 CTA table)
    int n_sc_blocks = 0;
    unsigned char* pmask = nullptr;   // [cap_p] point mask of marginalizePointsF
    int* ridx = nullptr;           // [cap_p][LDSO_MAX_WINDOW] residual id of (point, target) or -1 (written by K9a)
    // device: small tables
    float* adHTdelta = nullptr;    // [n*n][8]
    float* xAd = nullptr;          // [n*n][8] + xc[4]
    double* stats = nullptr;       // [4]
    // device: accumulators (double, filled with atomicAdd of per-CTA fp32 partial sums)
    double* acc_top = nullptr;     // [2][n*n][WIN_TOP_STRIDE]
    int* acc_top_cnt = nullptr;    // [2][n*n]
    double* acc_sc = nullptr;      // accD [n^3][64] | accE [n^2][32] | accEB [n^2][8] | accHcc [16] | accbc [4]
    size_t acc_sc_size = 0;
    double *adH_dev = nullptr, *adT_dev = nullptr;   // [n*n][64] adjoints (setAdjointsF)
    double* prior_dev = nullptr;   // cPrior[4] cDelta[4] fprior[8n] fdelta_prior[8n]
    double* stitched = nullptr;    // [3] x ((8n+4)^2 + (8n+4)): H,b of the active, linearised and Schur systems
    // host mirrors
    std::vector<int> h_host, h_res_begin, h_target, h_point, h_host_begin;
    std::vector<unsigned char> h_lin;
    std::vector<double> adHost, adTarget, fprior, fdelta_prior;
    std::vector<float> adHostF, adTargetF;
    bool ad_dev_valid = false;     // adH_dev / adT_dev hold adHost / adTarget
    float cDeltaF[4] = {0, 0, 0, 0};
    double cPrior[4] = {0, 0, 0, 0};
    // pinned staging
    void* pinned = nullptr;
    size_t pinned_bytes = 0, pinned_off = 0;
};

struct ldso_gpu_ctx;
int win_alloc(ldso_gpu_ctx* c);
void win_free(ldso_gpu_ctx* c);
int k7_linearize(ldso_gpu_ctx* c, int fix);
int k7_apply_res(ldso_gpu_ctx* c);
int k7_reset_oob(ldso_gpu_ctx* c);
int k7_fix_linearization(ldso_gpu_ctx* c, const unsigned char* mask_dev);
int k8_accumulate_top(ldso_gpu_ctx* c, int marg = 0);
int k9_accumulate_sc(ldso_gpu_ctx* c, const unsigned char* marg_mask = nullptr, float prior_fac = 1.f);
int k8_linearized_energy(ldso_gpu_ctx* c, double* e_out);
int k10_resubstitute(ldso_gpu_ctx* c, float lambda);
int h1_stitch(ldso_gpu_ctx* c);
