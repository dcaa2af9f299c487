// immature.h — device-resident pool of immature points (K5 / K6 / N2), internal to libldso_gpu.so
#pragma once
#include "ctx.h"

struct ImmatureDev {
    int cap = 0, n = 0;
    int *host = nullptr, *status = nullptr;
    float *u = nullptr, *v = nullptr, *color = nullptr, *weights = nullptr, *gradH = nullptr, *energyTH = nullptr;
    float *idmin = nullptr, *idmax = nullptr, *quality = nullptr, *lastUV = nullptr, *interval = nullptr;
    float* type = nullptr;                   // ImmaturePoint::my_type (1 / 2 / 4: the selector level the pixel came from)
    // second set of pool arrays: compaction writes here, then the two sets swap
    int *host2 = nullptr, *status2 = nullptr;
    float *u2 = nullptr, *v2 = nullptr, *color2 = nullptr, *weights2 = nullptr, *gradH2 = nullptr, *energyTH2 = nullptr;
    float *idmin2 = nullptr, *idmax2 = nullptr, *quality2 = nullptr, *lastUV2 = nullptr, *interval2 = nullptr;
    float* type2 = nullptr;
    int *keep = nullptr, *pos = nullptr;     // [cap + 1] compaction flags / exclusive scan
    float* tables = nullptr;                 // per-host KRKi[9] Kt[3] aff[2] (stride 16)
    int* counts = nullptr;                   // [8]
    // activation scratch
    int* cand = nullptr;                     // [cap]
    float* act_idepth = nullptr;             // [cap]
    int* act_ok = nullptr;                   // [cap]
    int* act_state = nullptr;                // [cap][LDSO_MAX_WINDOW]
    // create scratch: u, v in; color[8], weights[8], gradH[3], energyTH out
    float* cr = nullptr;                     // [cap][22]
    // N2: CoarseDistanceMap (level-1 grid, one byte per cell: BFS ring 0..39, 255 = the reference's 1000) and the
    // activation walk's outputs
    unsigned char* dmap = nullptr;           // [w1 * h1] (+ padding to a multiple of 16)
    unsigned char* action = nullptr;         // [cap] 0 keep, 1 delete, 2 candidate
    int* sel_cnt = nullptr;                  // [4] candidates, deleted
    int* src_cells = nullptr;                // [src_cap] level-1 cell of every source of the map (active points, then accepted ones), -1 = none
    int* src_n = nullptr;                    // device: length of the list
    bool map_stale = false;                  // accepted cells were appended to src_cells after dmap was built
    int src_cap = 0, src_n_host = 0;         // host: capacity and an upper bound of the current length
    void* pinned = nullptr;
    int* trace_cnt = nullptr;                // pinned [8]: status counts + step counts of the last trace
    cudaEvent_t ev_ready = nullptr, ev_done = nullptr;   // tracking stream -> trace stream, trace done
    cudaStream_t trace_stream = nullptr;     // stream of the pending trace
    int trace_slot = -1;                     // frame slot the pending trace reads
    bool trace_pending = false;              // a trace was launched without waiting (ldso_gpu_immature_trace_async)
    int trace_parity = 0;                    // which of the two device count blocks the last trace wrote
    bool trace_cnt_valid = false, trace_counts_fresh = false, trace_launched = false;
    int trace_prof_idx = -1;
    int trace_counts[6] = {0, 0, 0, 0, 0, 0};
    size_t pinned_bytes = 0;
};

struct PoolPtrs {
    int *host, *status;
    float *u, *v, *color, *weights, *gradH, *energyTH, *idmin, *idmax, *quality, *lastUV, *interval, *type;
};

ImmatureDev* immature_pool(ldso_gpu_ctx* c);   // allocates on first use; nullptr (and the ctx error set) on failure
