// oracle/orc_selector.cpp — TEST INFRASTRUCTURE (see orc_common.h).  PARITY UNPINNED.
// CPU restatement of PixelSelector (src/frontend/PixelSelector2.cc): constructor (:9-19), computeHistQuantil
// (:28-35), makeHists (:36-109), makeMaps (:111-169), select (:171-316).
// Decision recorded in SURVEY.md trap T3: ths / thsSmoothed are zero-filled at allocation.
#include <mutex>
#include "orc_common.h"

#include <algorithm>

using namespace orc;

namespace {

struct Selector {
    std::vector<unsigned char> randomPattern;
    std::vector<float> ths, thsSmoothed;
    int thsStep = 0;
    int histFrame = -1;
    long long histStamp = -1;
};

Selector& sel_of(orc_ctx* c) {
    if (c->selector) return *(Selector*)c->selector.get();
    std::shared_ptr<Selector> s = std::make_shared<Selector>();
    int w = c->wG[0], h = c->hG[0];
    s->randomPattern.resize((size_t)w * h);
    {   // the reference draws from the process-global libc generator (PixelSelector2.cc:11-13); contexts created from
        // several threads must not interleave their draws
        static std::mutex rand_mu;
        std::lock_guard<std::mutex> lk(rand_mu);
        std::srand(3141592);
        for (int i = 0; i < w * h; i++) s->randomPattern[i] = rand() & 0xFF;
    }
    s->ths.assign((size_t)(w / 32 + 1) * (h / 32 + 1) + 100, 0.f);
    s->thsSmoothed.assign((size_t)(w / 32 + 1) * (h / 32 + 1) + 100, 0.f);
    c->selector = s;
    return *s;
}

int computeHistQuantil(int* hist, float below) {
    int th = hist[0] * below + 0.5f;
    for (int i = 0; i < 90; i++) {
        th -= hist[i + 1];
        if (th < 0) return i;
    }
    return 90;
}

void makeHists(orc_ctx* c, Selector& S, int slot) {
    const float* mapmax0 = c->frames[slot].absSquaredGrad[0].data();
    int w = c->wG[0], h = c->hG[0];
    int w32 = w / 32, h32 = h / 32;
    S.thsStep = w32;
    int hist0[100];
    for (int y = 0; y < h32; y++)
        for (int x = 0; x < w32; x++) {
            const float* map0 = mapmax0 + 32 * x + 32 * y * w;
            memset(hist0, 0, sizeof(int) * 50);
            for (int j = 0; j < 32; j++)
                for (int i = 0; i < 32; i++) {
                    int it = i + 32 * x, jt = j + 32 * y;
                    if (it > w - 2 || jt > h - 2 || it < 1 || jt < 1) continue;
                    int g = sqrtf(map0[i + j * w]);
                    if (g > 48) g = 48;
                    hist0[g + 1]++;
                    hist0[0]++;
                }
            // NB: the reference memsets 50 ints but computeHistQuantil reads hist[1..90]; entries 50..90 of its
            // heap buffer are never written after `new int[]` — zero here (the quantile is reached before bin 50)
            for (int i = 50; i < 100; i++) hist0[i] = 0;
            S.ths[x + y * w32] = computeHistQuantil(hist0, c->cfg.min_grad_hist_cut) + c->cfg.min_grad_hist_add;
        }
    for (int y = 0; y < h32; y++)
        for (int x = 0; x < w32; x++) {
            float sum = 0, num = 0;
            if (x > 0) {
                if (y > 0) { num++; sum += S.ths[x - 1 + (y - 1) * w32]; }
                if (y < h32 - 1) { num++; sum += S.ths[x - 1 + (y + 1) * w32]; }
                num++; sum += S.ths[x - 1 + (y)*w32];
            }
            if (x < w32 - 1) {
                if (y > 0) { num++; sum += S.ths[x + 1 + (y - 1) * w32]; }
                if (y < h32 - 1) { num++; sum += S.ths[x + 1 + (y + 1) * w32]; }
                num++; sum += S.ths[x + 1 + (y)*w32];
            }
            if (y > 0) { num++; sum += S.ths[x + (y - 1) * w32]; }
            if (y < h32 - 1) { num++; sum += S.ths[x + (y + 1) * w32]; }
            num++; sum += S.ths[x + y * w32];
            S.thsSmoothed[x + y * w32] = (sum / num) * (sum / num);
        }
}

void select(orc_ctx* c, Selector& S, int slot, float* map_out, int pot, float thFactor, int n[3]) {
    const Frame& f = c->frames[slot];
    const float* map0 = f.dIp[0].data();
    const float* mapmax0 = f.absSquaredGrad[0].data();
    const float* mapmax1 = f.absSquaredGrad[1].data();
    const float* mapmax2 = f.absSquaredGrad[2].data();
    int w = c->wG[0], w1 = c->wG[1], w2 = c->wG[2], h = c->hG[0];
    static const float directions[16][2] = {{0, 1.0000},       {0.3827, 0.9239}, {0.1951, 0.9808},  {0.9239, 0.3827},
                                            {0.7071, 0.7071},  {0.3827, -0.9239}, {0.8315, 0.5556},  {0.8315, -0.5556},
                                            {0.5556, -0.8315}, {0.9808, 0.1951},  {0.9239, -0.3827}, {0.7071, -0.7071},
                                            {0.5556, 0.8315},  {0.9808, -0.1951}, {1.0000, 0.0000},  {0.1951, -0.9808}};
    memset(map_out, 0, (size_t)w * h * sizeof(float));
    float dw1 = c->cfg.grad_downweight_per_level;
    float dw2 = dw1 * dw1;
    const bool dirDist = c->cfg.select_direction_distribution != 0;
    int n3 = 0, n2 = 0, n4 = 0;
    for (int y4 = 0; y4 < h; y4 += (4 * pot))
        for (int x4 = 0; x4 < w; x4 += (4 * pot)) {
            int my3 = std::min((4 * pot), h - y4), mx3 = std::min((4 * pot), w - x4);
            int bestIdx4 = -1;
            float bestVal4 = 0;
            const float* dir4 = directions[S.randomPattern[n2] & 0xF];
            for (int y3 = 0; y3 < my3; y3 += (2 * pot))
                for (int x3 = 0; x3 < mx3; x3 += (2 * pot)) {
                    int x34 = x3 + x4, y34 = y3 + y4;
                    int my2 = std::min((2 * pot), h - y34), mx2 = std::min((2 * pot), w - x34);
                    int bestIdx3 = -1;
                    float bestVal3 = 0;
                    const float* dir3 = directions[S.randomPattern[n2] & 0xF];
                    for (int y2 = 0; y2 < my2; y2 += pot)
                        for (int x2 = 0; x2 < mx2; x2 += pot) {
                            int x234 = x2 + x34, y234 = y2 + y34;
                            int my1 = std::min(pot, h - y234), mx1 = std::min(pot, w - x234);
                            int bestIdx2 = -1;
                            float bestVal2 = 0;
                            const float* dir2 = directions[S.randomPattern[n2] & 0xF];
                            for (int y1 = 0; y1 < my1; y1 += 1)
                                for (int x1 = 0; x1 < mx1; x1 += 1) {
                                    int idx = x1 + x234 + w * (y1 + y234);
                                    int xf = x1 + x234, yf = y1 + y234;
                                    if (xf < 4 || xf >= w - 5 || yf < 4 || yf > h - 4) continue;
                                    float pixelTH0 = S.thsSmoothed[(xf >> 5) + (yf >> 5) * S.thsStep];
                                    float pixelTH1 = pixelTH0 * dw1;
                                    float pixelTH2 = pixelTH1 * dw2;
                                    float ag0 = mapmax0[idx];
                                    if (ag0 > pixelTH0 * thFactor) {
                                        float dirNorm = fabsf((float)(map0[3 * idx + 1] * dir2[0] + map0[3 * idx + 2] * dir2[1]));
                                        if (!dirDist) dirNorm = ag0;
                                        if (dirNorm > bestVal2) { bestVal2 = dirNorm; bestIdx2 = idx; bestIdx3 = -2; bestIdx4 = -2; }
                                    }
                                    if (bestIdx3 == -2) continue;
                                    float ag1 = mapmax1[(int)(xf * 0.5f + 0.25f) + (int)(yf * 0.5f + 0.25f) * w1];
                                    if (ag1 > pixelTH1 * thFactor) {
                                        float dirNorm = fabsf((float)(map0[3 * idx + 1] * dir3[0] + map0[3 * idx + 2] * dir3[1]));
                                        if (!dirDist) dirNorm = ag1;
                                        if (dirNorm > bestVal3) { bestVal3 = dirNorm; bestIdx3 = idx; bestIdx4 = -2; }
                                    }
                                    if (bestIdx4 == -2) continue;
                                    float ag2 = mapmax2[(int)(xf * 0.25f + 0.125) + (int)(yf * 0.25f + 0.125) * w2];
                                    if (ag2 > pixelTH2 * thFactor) {
                                        float dirNorm = fabsf((float)(map0[3 * idx + 1] * dir4[0] + map0[3 * idx + 2] * dir4[1]));
                                        if (!dirDist) dirNorm = ag2;
                                        if (dirNorm > bestVal4) { bestVal4 = dirNorm; bestIdx4 = idx; }
                                    }
                                }
                            if (bestIdx2 > 0) { map_out[bestIdx2] = 1; bestVal3 = 1e10; n2++; }
                        }
                    if (bestIdx3 > 0) { map_out[bestIdx3] = 2; bestVal4 = 1e10; n3++; }
                }
            if (bestIdx4 > 0) { map_out[bestIdx4] = 4; n4++; }
        }
    n[0] = n2; n[1] = n3; n[2] = n4;
}

int makeMaps(orc_ctx* c, Selector& S, int slot, float* map_out, float density, int recursionsLeft, float thFactor, int& currentPotential) {
    float numHave = 0, numWant = density, quotia;
    int idealPotential = currentPotential;
    int n[3];
    select(c, S, slot, map_out, currentPotential, thFactor, n);
    numHave = n[0] + n[1] + n[2];
    quotia = numWant / numHave;
    float K = numHave * (currentPotential + 1) * (currentPotential + 1);
    idealPotential = sqrtf(K / numWant) - 1;
    if (idealPotential < 1) idealPotential = 1;
    if (recursionsLeft > 0 && quotia > 1.25 && currentPotential > 1) {
        if (idealPotential >= currentPotential) idealPotential = currentPotential - 1;
        currentPotential = idealPotential;
        return makeMaps(c, S, slot, map_out, density, recursionsLeft - 1, thFactor, currentPotential);
    } else if (recursionsLeft > 0 && quotia < 0.25) {
        if (idealPotential <= currentPotential) idealPotential = currentPotential + 1;
        currentPotential = idealPotential;
        return makeMaps(c, S, slot, map_out, density, recursionsLeft - 1, thFactor, currentPotential);
    }
    int numHaveSub = numHave;
    if (quotia < 0.95) {
        int wh = c->wG[0] * c->hG[0];
        int rn = 0;
        unsigned char charTH = 255 * quotia;
        for (int i = 0; i < wh; i++) {
            if (map_out[i] != 0) {
                if (S.randomPattern[rn] > charTH) { map_out[i] = 0; numHaveSub--; }
                rn++;
            }
        }
    }
    currentPotential = idealPotential;
    return numHaveSub;
}

}  // namespace

extern "C" int orc_select_make_maps(orc_ctx* c, int slot, float density, int recursions_left, float th_factor, int* potential_io,
                                    float* map_out, int* n_selected) {
    if (c->levels < 3) return -2;
    Selector& S = sel_of(c);
    makeHists(c, S, slot);   // the reference caches per frame (gradHistFrame); recomputing is idempotent
    int pot = *potential_io;
    *n_selected = makeMaps(c, S, slot, map_out, density, recursions_left, th_factor, pot);
    *potential_io = pot;
    return 0;
}

// makeMaps + the pixel loop of FullSystem::makeNewTraces (src/frontend/FullSystem.cc:1440-1460)
extern "C" int orc_select_points(orc_ctx* c, int slot, float density, int recursions_left, float th_factor, int* potential_io,
                                 int border, int max_points, float* u_out, float* v_out, float* type_out, int* n_points) {
    const int w = c->wG[0], h = c->hG[0];
    std::vector<float> map((size_t)w * h);
    int nsel = 0;
    int rc = orc_select_make_maps(c, slot, density, recursions_left, th_factor, potential_io, map.data(), &nsel);
    if (rc) return rc;
    int n = 0;
    for (int y = border; y < h - border; y++)
        for (int x = border; x < w - border; x++) {
            const float m = map[(size_t)y * w + x];
            if (m == 0) continue;
            if (n < max_points) { u_out[n] = (float)x; v_out[n] = (float)y; type_out[n] = m; n++; }
        }
    *n_points = n;
    return 0;
}

// =================================================================================================
// N3 — FeatureDetector::DetectCorners (src/frontend/FeatureDetector.cc:34-124) up to the corner flags;
// ShiTomasiScore: include/frontend/FeatureDetector.h:50-81
namespace {
float shiTomasiScore(const float* dIp, int wl, int hl, float u, float v) {
    const int halfbox = 4;
    float dXX = 0.0, dYY = 0.0, dXY = 0.0;
    const int box_size = 2 * halfbox, box_area = box_size * box_size;
    const int x_min = u - halfbox, x_max = u + halfbox, y_min = v - halfbox, y_max = v + halfbox;
    if (x_min < 1 || x_max >= wl - 1 || y_min < 1 || y_max >= hl - 1) return 0.0;
    for (int y = y_min; y < y_max; ++y)
        for (int x = x_min; x < x_max; ++x) {
            const float dx = dIp[3 * (y * wl + x) + 1], dy = dIp[3 * (y * wl + x) + 2];
            dXX += dx * dx; dYY += dy * dy; dXY += dx * dy;
        }
    dXX = dXX / (2.0 * box_area); dYY = dYY / (2.0 * box_area); dXY = dXY / (2.0 * box_area);
    // `using namespace std` is in effect in the reference (include/NumTypes.h:12), so its unqualified sqrt(float) is std::sqrt(float)
    return 0.5 * (dXX + dYY - sqrtf((dXX + dYY) * (dXX + dYY) - 4 * (dXX * dYY - dXY * dXY)));
}
}  // namespace

extern "C" int orc_detect_corners(orc_ctx* c, int slot, int nFeatures, int max_out, float* ox, float* oy, float* oscore,
                                  unsigned char* is_corner, int* n_out, int* n_corners) {
    const int w = c->wG[0], h = c->hG[0];
    const float* grad = c->frames[slot].absSquaredGrad[0].data();
    const float* dIp = c->frames[slot].dIp[0].data();
    const int HALF_PATCH_SIZE = 15;
    int gridsize = int(sqrtf(w * h / nFeatures) + 0.5);
    int gridX = w / gridsize + 1, gridY = h / gridsize + 1;
    float nfeatInGrid = float(nFeatures) / (w * h) * (gridsize * gridsize);
    float maxScore = 0, scoreTH = 0;
    int skip = (HALF_PATCH_SIZE * 2 / gridsize) + 1;
    struct F { float x, y, score; bool corner; };
    std::vector<F> feats;
    for (int gx = skip; gx < gridX - skip; gx++)
        for (int gy = skip; gy < gridY - skip; gy++) {
            float maxGrad = 0, gradTH = 0;
            const float* gradData = &grad[gy * gridsize * w + gx * gridsize];
            for (int x = 0; x < gridsize; x++)
                for (int y = 0; y < gridsize; y++)
                    if (gradData[y * w + x] > maxGrad) maxGrad = gradData[y * w + x];
            std::vector<std::pair<int, float>> candidate;
            gradTH = (0.5 * maxGrad) > 5 ? 0.5 * maxGrad : 5;
            int picked = 0;
            for (int x = 0; x < gridsize; x++)
                for (int y = 0; y < gridsize; y++) {
                    int idx = y * gridsize + x;
                    if (gradData[y * w + x] > gradTH) {
                        int realX = gx * gridsize + x, realY = gy * gridsize + y;
                        float s = shiTomasiScore(dIp, w, h, realX, realY);
                        candidate.push_back(std::pair<int, float>(idx, s));
                        if (s > maxScore) maxScore = s;
                    }
                }
            std::stable_sort(candidate.begin(), candidate.end(),
                             [](const std::pair<int, float>& p1, const std::pair<int, float>& p2) { return p1.second > p2.second; });
            for (auto& p : candidate) {
                int x = p.first % gridsize, y = p.first / gridsize;
                feats.push_back(F{(float)(gx * gridsize + x), (float)(gy * gridsize + y), p.second, false});
                picked++;
                if (picked > (nfeatInGrid)) break;
            }
        }
    scoreTH = 0.01 * maxScore;
    std::vector<int> corners;
    for (size_t i = 0; i < feats.size(); i++)
        if (feats[i].score > scoreTH) { feats[i].corner = true; corners.push_back((int)i); }
    for (size_t i = 0; i < corners.size(); i++)
        for (size_t j = i + 1; j < corners.size(); j++) {
            F &f1 = feats[corners[i]], &f2 = feats[corners[j]];
            const float dx = f1.x - f2.x, dy = f1.y - f2.y;
            if (sqrtf(dx * dx + dy * dy) < 5) {
                if (f1.score > f2.score) f2.corner = false;
                else f1.corner = false;
            }
        }
    int n = 0, nc = 0;
    for (const F& f : feats) {
        if (n >= max_out) break;
        ox[n] = f.x; oy[n] = f.y; oscore[n] = f.score; is_corner[n] = f.corner ? 1 : 0;
        nc += f.corner ? 1 : 0;
        n++;
    }
    *n_out = n; *n_corners = nc;
    return 0;
}
