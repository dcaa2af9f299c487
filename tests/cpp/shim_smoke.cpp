// shim_smoke.cpp — drives include/ldso_shim.hpp the way the reference's FullSystem drives its own classes:
// makeImages on two frames, setCoarseTrackingRef + trackNewestCoarse, PixelSelector::selectPoints, ImmaturePool add +
// traceOn.  Input: a blob written by tests/test_cpp_shim.py; output: one line per result, compared there with the
// ctypes path through the same library.
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "ldso_shim.hpp"

using namespace ldso_b200;

template <class T> static std::vector<T> rd(FILE* f, size_t n) {
    std::vector<T> v(n);
    if (n && fread(v.data(), sizeof(T), n, f) != n) { fprintf(stderr, "short read\n"); exit(2); }
    return v;
}

int main(int argc, char** argv) {
    if (argc < 2) return 2;
    FILE* f = fopen(argv[1], "rb");
    if (!f) return 2;
    auto hdr = rd<int>(f, 3);                       // w, h, n_points
    auto cal = rd<float>(f, 4);                     // fx fy cx cy
    const int w = hdr[0], h = hdr[1], n = hdr[2];
    auto img0 = rd<float>(f, (size_t)w * h), img1 = rd<float>(f, (size_t)w * h);
    auto u = rd<float>(f, n), v = rd<float>(f, n), idp = rd<float>(f, n), hdi = rd<float>(f, n);
    auto tab = rd<float>(f, 14);                    // host->new KRKi[9], Kt[3], aff[2]
    fclose(f);
    try {
        Context c(w, h, cal[0], cal[1], cal[2], cal[3]);
        c.cfg.affine_opt_mode_a = 0; c.cfg.affine_opt_mode_b = 0;   // mode = 1 (run_dso_tum_mono.cc:283-288)
        c.create();
        FrameHessian ref, cur;
        ref.makeImages(c, 0, img0.data());
        cur.makeImages(c, 1, img1.data());
        CoarseTracker tracker(0);
        tracker.makeK(c, cal[0], cal[1], cal[2], cal[3]);
        tracker.setCoarseTrackingRef(c, ref, 0, u, v, idp, hdi);
        SE3 lastToNew;
        AffLight aff;
        const double minRes[5] = {NAN, NAN, NAN, NAN, NAN};
        const bool ok = tracker.trackNewestCoarse(c, cur, lastToNew, aff, c.pyrLevelsUsed() - 1, minRes);
        double Rt[12];
        lastToNew.toRt(Rt);
        printf("track %d", ok ? 1 : 0);
        for (int i = 0; i < 12; i++) printf(" %.17g", Rt[i]);
        printf(" %.9g %.9g", aff.a, aff.b);
        for (int i = 0; i < 5; i++) printf(" %.9g", tracker.lastResiduals[i]);
        printf("\n");
        PixelSelector sel;
        std::vector<float> su, sv, st;
        const int nsel = sel.selectPoints(c, ref, 300.0f, 5, su, sv, st);
        printf("select %d %d\n", nsel, sel.currentPotential);
        ImmaturePool pool;
        pool.clear(c);
        pool.add(c, ref, 0, su, sv);
        int counts[6];
        pool.traceOn(c, cur, 1, tab.data(), tab.data() + 9, tab.data() + 12, counts);
        printf("trace %d %d %d %d %d %d\n", counts[0], counts[1], counts[2], counts[3], counts[4], counts[5]);
        std::vector<float> idmin, idmax, q;
        std::vector<int> status;
        pool.sync(c, idmin, idmax, q, status);
        double sum = 0;
        int good = 0;
        for (int i = 0; i < pool.size; i++)
            if (status[i] == (int)ImmaturePointStatus::IPS_GOOD) { sum += idmin[i] + idmax[i]; good++; }
        printf("depth %d %.9g\n", good, sum);
        // CoarseInitializer::setFirst + one trackFrame, CoarseDistanceMap::makeDistanceMap (identity warp of the tracker points)
        CoarseInitializer init;
        init.setFirst(c, ref);
        const bool ready = init.trackFrame(c, cur);
        double iRt[12];
        init.thisToNext.toRt(iRt);
        auto pts = init.points(c, 0);
        double irs = 0;
        for (auto& p : pts) irs += p.iR;
        printf("init %d %d %d %d %.9g %.9g %.9g %.9g\n", init.numPoints[0], init.numPoints[1], ready ? 1 : 0, init.snapped ? 1 : 0, iRt[9],
               iRt[10], iRt[11], irs);
        CoarseDistanceMap dm;
        const float K1[9] = {0.5f, 0, -0.25f, 0, 0.5f, -0.25f, 0, 0, 1}, Kt0[3] = {0, 0, 0};
        std::vector<int> hostIdx(n, 0);
        dm.makeDistanceMap(c, 1, K1, Kt0, hostIdx, u, v, idp);
        std::vector<float> D;
        dm.fwdWarpedIDDistFinal(c, D, w / 2, h / 2);
        double dsum = 0;
        for (float d : D) dsum += d;
        printf("distmap %.9g\n", dsum);
    } catch (const std::exception& e) {
        fprintf(stderr, "error: %s\n", e.what());
        return 1;
    }
    return 0;
}
