/* ldso_knn.h — exact k-nearest-neighbour search on a planar point set (host code).
 *
 * Stands in for the nanoflann kd-tree of CoarseInitializer::makeNN (src/frontend/CoarseInitializer.cc:839-906,
 * src/frontend/nanoflann.h): squared Euclidean distance in float (L2_Simple_Adaptor, d0*d0 + d1*d1), results sorted
 * by ascending distance.  nanoflann does not specify which of several equidistant points it reports first; here
 * ties resolve to the lower point index, so the result is a pure function of the point set.
 * A uniform bucket grid is searched ring by ring until the k-th best distance is closer than the next ring. */
#ifndef LDSO_KNN_H_
#define LDSO_KNN_H_
#include <algorithm>
#include <cmath>
#include <utility>
#include <vector>

namespace ldso_knn {

struct Grid {
    float cell = 8.f;
    int gw = 0, gh = 0, n = 0;
    const float *u = nullptr, *v = nullptr;
    std::vector<int> start, items;   // CSR: points of every bucket in ascending index order

    void build(int n_, const float* u_, const float* v_, int w, int h, float cell_) {
        n = n_; u = u_; v = v_; cell = cell_;
        gw = (int)(w / cell) + 2; gh = (int)(h / cell) + 2;
        start.assign((size_t)gw * gh + 1, 0);
        for (int i = 0; i < n; i++) start[bucket(u[i], v[i]) + 1]++;
        for (size_t b = 0; b < (size_t)gw * gh; b++) start[b + 1] += start[b];
        items.resize(n);
        std::vector<int> fill(start.begin(), start.end() - 1);
        for (int i = 0; i < n; i++) items[fill[bucket(u[i], v[i])]++] = i;
    }
    int bx(float x) const { int b = (int)std::floor(x / cell); return b < 0 ? 0 : (b >= gw ? gw - 1 : b); }
    int by(float y) const { int b = (int)std::floor(y / cell); return b < 0 ? 0 : (b >= gh ? gh - 1 : b); }
    int bucket(float x, float y) const { return bx(x) + gw * by(y); }

    /* the k nearest points of (qx, qy): indices and squared distances, ascending (distance, index).
     * Returns how many were found (min(k, n)). */
    int query(float qx, float qy, int k, int* idx_out, float* dist_out) const {
        std::vector<std::pair<float, int>> best;   // kept sorted, at most k entries
        best.reserve(k + 1);
        const int cx = bx(qx), cy = by(qy);
        const int rmax = std::max(gw, gh);
        for (int r = 0; r <= rmax; r++) {
            if ((int)best.size() == k && r >= 1) {
                // every point in ring r is at least (r - 1) * cell + margin away, where margin is the distance of the
                // query to the border of its own bucket; use the conservative (r - 1) * cell
                const float lo = (r - 1) * cell;
                if (lo > 0 && lo * lo > best.back().first) break;
            }
            for (int yy = cy - r; yy <= cy + r; yy++) {
                if (yy < 0 || yy >= gh) continue;
                const bool edge_row = (yy == cy - r || yy == cy + r);
                for (int xx = cx - r; xx <= cx + r; xx += (edge_row ? 1 : 2 * r > 0 ? 2 * r : 1)) {
                    if (xx < 0 || xx >= gw) continue;
                    const int b = xx + gw * yy;
                    for (int s = start[b]; s < start[b + 1]; s++) {
                        const int i = items[s];
                        const float d0 = qx - u[i], d1 = qy - v[i];
                        const float d = d0 * d0 + d1 * d1;
                        if ((int)best.size() == k && !(d < best.back().first || (d == best.back().first && i < best.back().second)))
                            continue;
                        std::pair<float, int> e(d, i);
                        best.insert(std::upper_bound(best.begin(), best.end(), e), e);
                        if ((int)best.size() > k) best.pop_back();
                    }
                }
            }
        }
        for (size_t j = 0; j < best.size(); j++) { idx_out[j] = best[j].second; dist_out[j] = best[j].first; }
        return (int)best.size();
    }
};

}  // namespace ldso_knn
#endif /* LDSO_KNN_H_ */
