// Checks mc-slam_b200/csrc/introsort_emul.h against the real libstdc++ std::sort on tie-heavy inputs, using the
// reference's comparator shape (size, then UL.x; ORBextractor.cpp:537-552).  Exit code 0 = all permutations identical.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <random>
#include <vector>

#include "../../mc-slam_b200/csrc/introsort_emul.h"

struct Nd { int ulx; int id; };
static bool cmp(std::pair<int, Nd*>& a, std::pair<int, Nd*>& b) {
    if (a.first < b.first) return true;
    if (a.first > b.first) return false;
    return a.second->ulx < b.second->ulx;
}

int main(int argc, char** argv) {
    int trials = argc > 1 ? atoi(argv[1]) : 20000;
    std::mt19937 rng(12345);
    long long checked = 0;
    for (int t = 0; t < trials; t++) {
        int n = 1 + rng() % (t % 50 == 0 ? 3000 : 400);
        int size_range = 1 + rng() % (t % 3 == 0 ? 3 : 40);
        int x_range = 1 + rng() % (t % 5 == 0 ? 2 : 30);
        std::vector<Nd> nodes(n);
        std::vector<std::pair<int, Nd*>> v(n);
        std::vector<uint32_t> k(n);
        std::vector<uint16_t> p(n);
        int mode = rng() % 6;
        for (int i = 0; i < n; i++) {
            int size = 2 + rng() % size_range;
            int ulx = (rng() % x_range) * 7;
            if (mode == 1) size = 2 + (i * size_range) / n;             // ascending
            if (mode == 2) size = 2 + ((n - 1 - i) * size_range) / n;   // descending
            if (mode == 3) { size = 5; ulx = 0; }                       // all equal
            if (mode == 4) size = 2 + ((i % 2) ? i : n - i) % size_range;  // organ-pipe-ish (provokes heap fallback)
            nodes[i] = {ulx, i};
            v[i] = {size, &nodes[i]};
            k[i] = ((uint32_t)size << 12) | (uint32_t)ulx;
            p[i] = (uint16_t)i;
        }
        std::sort(v.begin(), v.end(), cmp);
        mcs_sort::sort(k.data(), p.data(), n);
        for (int i = 0; i < n; i++) {
            if (v[i].second->id != p[i]) {
                printf("MISMATCH trial %d n %d at %d: std %d emul %d\n", t, n, i, v[i].second->id, p[i]);
                return 1;
            }
        }
        checked += n;
    }
    // adversarial: McIlroy's "antiqsort" run against std::sort itself builds an input that drives introsort to its
    // depth limit, so the heap-sort fallback (and its tie order) is exercised too.
    for (int n : {64, 257, 1000, 4096}) {
        std::vector<int> val(n, n - 1), ptr(n);
        int nsolid = 0, candidate = 0;
        const int gas = n - 1;
        for (int i = 0; i < n; i++) ptr[i] = i;
        std::sort(ptr.begin(), ptr.end(), [&](int x, int y) {
            if (val[x] == gas && val[y] == gas) { if (x == candidate) val[x] = nsolid++; else val[y] = nsolid++; }
            if (val[x] == gas) candidate = x; else if (val[y] == gas) candidate = y;
            return val[x] < val[y];
        });
        std::vector<int> kk = val;
        std::vector<Nd> nodes(n);
        std::vector<std::pair<int, Nd*>> v(n);
        std::vector<uint32_t> k(n);
        std::vector<uint16_t> p(n);
        for (int i = 0; i < n; i++) {
            nodes[i] = {0, i};
            v[i] = {kk[i] / 2 + 2, &nodes[i]};
            k[i] = ((uint32_t)(kk[i] / 2 + 2) << 12);
            p[i] = (uint16_t)i;
        }
        std::sort(v.begin(), v.end(), cmp);
        mcs_sort::sort(k.data(), p.data(), n);
        for (int i = 0; i < n; i++)
            if (v[i].second->id != p[i]) { printf("MISMATCH adversarial n %d at %d\n", n, i); return 1; }
    }
    printf("OK %d trials, %lld elements\n", trials, checked);
    return 0;
}
