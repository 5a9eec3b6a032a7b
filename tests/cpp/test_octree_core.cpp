// Host-side check of mc-slam_b200/csrc/octree_core.h (the block-parallel octree formulation used by the CUDA kernel)
// against the oracle's std::list restatement of DistributeOctTree (oracle/mcs_oracle.cpp, ORBextractor.cpp:554-778).
// The kernel code is executed with a sequential "block" context: every par() phase loops over the virtual threads.
// Build: g++ -O2 -std=c++17 tests/cpp/test_octree_core.cpp -Loracle -loracle -Wl,-rpath,$PWD/oracle
#include <algorithm>
#include <array>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <set>
#include <vector>

#include "../../mc-slam_b200/csrc/octree_core.h"

extern "C" int orc_distribute_octtree(const float* xyr, int n, int minX, int maxX, int minY, int maxY, int N, int32_t* keep);

struct HostCtx {
    int nt, cur = 0;
    explicit HostCtx(int n) : nt(n) {}
    int nthreads() const { return nt; }
    int tid() const { return cur; }
    template <typename F> void par(F f) { for (cur = 0; cur < nt; cur++) f(); cur = 0; }
    template <typename F> void serial(F f) { cur = 0; f(); }
    int exclusive_scan(int* a, int n) { int s = 0; for (int i = 0; i < n; i++) { int v = a[i]; a[i] = s; s += v; } return s; }
    void atomic_add(int* p, int v) { *p += v; }
    void atomic_max(unsigned* p, unsigned v) { if (v > *p) *p = v; }
    void atomic_min(unsigned* p, unsigned v) { if (v < *p) *p = v; }
    void atomic_max64(unsigned long long* p, unsigned long long v) { if (v > *p) *p = v; }
};

int main(int argc, char** argv) {
    int trials = argc > 1 ? atoi(argv[1]) : 300;
    std::mt19937 rng(777);
    int fails = 0;
    for (int t = 0; t < trials; t++) {
        // level geometry like ComputeKeyPointsOctTree (:780-803)
        int cols = 80 + rng() % (t % 7 == 0 ? 1900 : 700), rows = 80 + rng() % (t % 7 == 0 ? 1200 : 500);
        if (t % 11 == 0) { cols = 400 + rng() % 900; rows = cols / 2 - rng() % 60; }  // wide: n_ini = 2
        const int minB = 16, maxBX = cols - 16, maxBY = rows - 16;
        const int roi_w = maxBX - minB, roi_h = maxBY - minB;
        const int nCols = (int)((float)roi_w / 35.f), nRows = (int)((float)roi_h / 35.f);
        if (nCols < 1 || nRows < 1) continue;
        const int wCell = (int)std::ceil((float)roi_w / nCols), hCell = (int)std::ceil((float)roi_h / nRows);
        int N = 1 + rng() % (t % 5 == 0 ? 900 : 250);
        // candidates: distinct pixels in [3, roi-3), emitted in the reference's order (cell row-major, raster in cell)
        double density = (t % 3 == 0) ? 0.002 : (t % 3 == 1 ? 0.02 : 0.1);
        if (t % 13 == 0) density = 0.0002;
        int score_range = 1 + rng() % (t % 2 ? 6 : 200);  // few distinct scores => many response ties
        std::vector<std::array<int, 3>> pts;
        for (int y = 3; y < roi_h - 3; y++)
            for (int x = 3; x < roi_w - 3; x++)
                if ((rng() % 100000) < density * 100000) pts.push_back({x, y, 7 + (int)(rng() % score_range)});
        if (t % 17 == 0 && pts.size() > 40) {  // clustered: many keys in one corner
            for (auto& p : pts) { p[0] = 3 + p[0] % 40; p[1] = 3 + p[1] % 40; }
            std::set<std::pair<int,int>> seen; std::vector<std::array<int,3>> u;
            for (auto& p : pts) if (seen.insert({p[0], p[1]}).second) u.push_back(p);
            pts = u;
        }
        auto cellkey = [&](const std::array<int, 3>& p) {
            int j = (p[0] - 3) / wCell, i = (p[1] - 3) / hCell;
            return std::array<int, 4>{i, j, p[1], p[0]};
        };
        std::sort(pts.begin(), pts.end(), [&](auto& a, auto& b) { return cellkey(a) < cellkey(b); });
        const int n = (int)pts.size();
        std::vector<float> xyr(n * 3 + 3);
        std::vector<uint32_t> cand(n + 1);
        // device candidate order is arbitrary: shuffle to prove order-independence
        std::vector<int> perm(n);
        for (int i = 0; i < n; i++) perm[i] = i;
        std::shuffle(perm.begin(), perm.end(), rng);
        for (int i = 0; i < n; i++) {
            xyr[i * 3] = (float)pts[i][0]; xyr[i * 3 + 1] = (float)pts[i][1]; xyr[i * 3 + 2] = (float)pts[i][2];
            cand[perm[i]] = mcs_oct::pack_cand(pts[i][0], pts[i][1], pts[i][2]);
        }
        std::vector<int32_t> keep(n + 8);
        int nk = orc_distribute_octtree(xyr.data(), n, minB, maxBX, minB, maxBY, N, keep.data());

        mcs_oct::Params P;
        P.roi_w = roi_w; P.roi_h = roi_h; P.N = N;
        P.n_ini = std::max(1, (int)std::round((float)roi_w / roi_h));
        P.hX = (float)roi_w / P.n_ini;
        P.w_cell = wCell; P.h_cell = hCell; P.n_cols = nCols;
        P.cap = std::max(N, 4 * P.n_ini) + 4;
        std::vector<char> smem(mcs_oct::Shared::bytes(P.cap) + 64);
        mcs_oct::Shared S;
        S.carve(smem.data(), P.cap);
        std::vector<uint16_t> knode(n + 1);
        std::vector<uint32_t> out(P.cap, 0xFFFFFFFFu);
        HostCtx ctx(1 + rng() % 96);
        int no = mcs_oct::distribute(ctx, P, S, cand.data(), n, knode.data(), out.data());
        bool ok = (no == nk);
        for (int i = 0; ok && i < nk; i++) {
            uint32_t c = out[i];
            const auto& p = pts[keep[i]];
            ok = mcs_oct::cand_x(c) == p[0] && mcs_oct::cand_y(c) == p[1] && mcs_oct::cand_score(c) == p[2];
        }
        if (!ok) {
            fails++;
            printf("FAIL trial %d: cols %d rows %d N %d keys %d oracle %d core %d\n", t, cols, rows, N, n, nk, no);
            if (fails > 5) break;
        }
    }
    if (fails) { printf("%d FAILURES\n", fails); return 1; }
    printf("OK %d trials\n", trials);
    return 0;
}
