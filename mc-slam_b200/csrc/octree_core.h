// octree_core.h — order-preserving octree keypoint distribution, formulated for one CUDA thread block.
//
// Reproduces ORBextractor::DistributeOctTree (/root/reference/MCSlam/src/ORBextractor.cpp:554-778, DivideNode :479-535,
// compareNodes :537-552) including the ORDER of the returned keypoints, which fixes descriptor row order downstream.
//
// The reference walks a std::list, push_front-ing children and erasing parents.  That is re-expressed here without a
// linked list: the node table is kept IN LIST ORDER and rebuilt every round —
//     new list = reverse(children in creation order) ++ (untouched nodes in their old order)
// so child positions come from exclusive scans over the nodes split this round (in processing order) and survivor
// positions from a scan over the old list.  Breadth-first rounds (:596-684) process splittable nodes in list order; the
// size-sorted rounds (:688-753) process them from the back of libstdc++'s std::sort order (introsort_emul.h) and stop at
// the first prefix whose list size reaches N (:745).  Keys never move: each key carries the id (= list position) of its
// node and is relabelled once per round; per-node key counts are smem atomics.  The per-node winner ("max response, first
// wins", :757-775) is an atomicMax on score followed by an atomicMin on the key's rank in the reference's candidate order
// (cell row-major, raster inside a cell, :804-871).
//
// The algorithm is written against a small execution context so the SAME code runs (a) as a CUDA block (octree kernel in
// extract_kernels.cuh) and (b) single-threaded on the host for tests/cpp/test_octree_core.cpp, which checks it against a
// std::list restatement of the reference on random inputs.  Ctx provides:
//     int  nthreads();  int tid();          — GPU: blockDim.x / threadIdx.x; host: set by the loop driver
//     template<F> void par(F f);            — run f() for every thread, then barrier
//     template<F> void serial(F f);         — run f() on thread 0, then barrier
//     int  exclusive_scan(int* a, int n);   — in-place block-wide exclusive scan, returns the total (collective)
//     void atomic_add(int*, int), atomic_max64(unsigned long long*, unsigned long long)
#ifndef MCS_OCTREE_CORE_H
#define MCS_OCTREE_CORE_H

#include <stdint.h>

#include "introsort_emul.h"

namespace mcs_oct {

// candidate record: x (12 bits) | y (12 bits) << 12 | score (8 bits) << 24, ROI coordinates (before +minBorder)
MCS_HD uint32_t pack_cand(int x, int y, int score) { return (uint32_t)x | ((uint32_t)y << 12) | ((uint32_t)score << 24); }
MCS_HD int cand_x(uint32_t c) { return (int)(c & 0xFFFu); }
MCS_HD int cand_y(uint32_t c) { return (int)((c >> 12) & 0xFFFu); }
MCS_HD int cand_score(uint32_t c) { return (int)(c >> 24); }

struct Geo {
    short ulx, uly, brx, bry;
};

struct Params {
    int roi_w, roi_h;  // maxBorderX - minBorderX, maxBorderY - minBorderY
    int N;             // mnFeaturesPerLevel[level]
    int n_ini;         // round(roi_w / roi_h), >= 1
    float hX;          // (float)roi_w / n_ini
    int w_cell, h_cell, n_cols;  // FAST cell grid, for the candidate-order rank
    int cap;           // node-table capacity (>= max(N, 4*n_ini) + 4)
};

// Shared-memory carve-up.  All arrays have `cap` entries unless noted.
struct Shared {
    Geo* geo[2];
    int* cnt[2];
    int* cc;          // [cap*4] child key counts of the nodes that may be split this round (cc2[0] or cc2[1])
    int* cc2[2];      // two buffers: the relabelling pass of a round already counts the NEXT round's children into the other one
    uint16_t* map;    // [cap*4] (old node, quadrant) -> new node id; quadrant slot 0 for unsplit nodes
    uint8_t* flag;    // 1 = node is split this round
    uint16_t* proc;   // nodes in processing order
    uint16_t* exp[2]; // expandable children (count > 1) in creation order: vSizeAndPointerToNode
    int* s0;          // scan scratch
    int* s1;
    int* s2;
    uint32_t* sortk;  // phase-2 sort keys
    unsigned long long* best64;  // final selection per node: max of (score << 32 | ~rank) = highest score, then first in the reference's order
    int* vars;        // [16] block-uniform scalars

    static MCS_HD size_t bytes(int cap) {
        size_t b = 0;
        b += 2 * sizeof(Geo) * cap + 2 * sizeof(int) * cap;  // geo, cnt
        b += 2 * sizeof(int) * cap * 4;                      // cc2
        b += sizeof(uint16_t) * cap * 4;                     // map
        b += sizeof(uint16_t) * cap * 3;                     // proc, exp x2
        b += sizeof(int) * cap * 3;                          // s0..s2
        b += sizeof(uint32_t) * cap * 3;                     // sortk, best, bestrank
        b += ((cap + 3) & ~3);                               // flag
        b += 16 * sizeof(int);
        return (b + 15) & ~(size_t)15;
    }
    MCS_HD void carve(void* base, int cap) {
        char* p = (char*)base;
        geo[0] = (Geo*)p; p += sizeof(Geo) * cap;
        geo[1] = (Geo*)p; p += sizeof(Geo) * cap;
        cnt[0] = (int*)p; p += sizeof(int) * cap;
        cnt[1] = (int*)p; p += sizeof(int) * cap;
        cc2[0] = (int*)p; p += sizeof(int) * cap * 4;
        cc2[1] = (int*)p; p += sizeof(int) * cap * 4;
        cc = cc2[0];
        s0 = (int*)p; p += sizeof(int) * cap;
        s1 = (int*)p; p += sizeof(int) * cap;
        s2 = (int*)p; p += sizeof(int) * cap;
        sortk = (uint32_t*)p; p += sizeof(uint32_t) * cap;
        best64 = (unsigned long long*)p; p += sizeof(unsigned long long) * cap;  // offset 72 * cap: 8-byte aligned
        vars = (int*)p; p += sizeof(int) * 16;
        map = (uint16_t*)p; p += sizeof(uint16_t) * cap * 4;
        proc = (uint16_t*)p; p += sizeof(uint16_t) * cap;
        exp[0] = (uint16_t*)p; p += sizeof(uint16_t) * cap;
        exp[1] = (uint16_t*)p; p += sizeof(uint16_t) * cap;
        flag = (uint8_t*)p;
    }
};

enum { V_NLIST = 0, V_NEXP, V_PHASE, V_FINISH, V_NPROC, V_CUR, V_PREVSIZE, V_T, V_S, V_E, V_OVERFLOW };

MCS_HD int quadrant(const Geo& g, int x, int y) {
    // DivideNode: halfX = ceil((UR.x-UL.x)/2); n1 = left/top, n2 = right/top, n3 = left/bottom, n4 = right/bottom
    const int midx = g.ulx + ((g.brx - g.ulx + 1) >> 1);
    const int midy = g.uly + ((g.bry - g.uly + 1) >> 1);
    return (x < midx ? 0 : 1) + (y < midy ? 0 : 2);
}
MCS_HD Geo child_geo(const Geo& g, int q) {
    const short midx = (short)(g.ulx + ((g.brx - g.ulx + 1) >> 1));
    const short midy = (short)(g.uly + ((g.bry - g.uly + 1) >> 1));
    Geo c;
    c.ulx = (q & 1) ? midx : g.ulx;
    c.brx = (q & 1) ? g.brx : midx;
    c.uly = (q & 2) ? midy : g.uly;
    c.bry = (q & 2) ? g.bry : midy;
    return c;
}

// rank of a candidate in the reference's vToDistributeKeys order
// The two cell divisions as multiplications: t / n == (t * ceil(2^20 / n)) >> 20 for t * n < 2^20 (coordinates < 4096, cells <= 69 px).
struct RankDiv {
    unsigned mw, mh;
    MCS_HD explicit RankDiv(const Params& P) : mw(((1u << 20) + P.w_cell - 1) / P.w_cell), mh(((1u << 20) + P.h_cell - 1) / P.h_cell) {}
};
MCS_HD unsigned cand_rank(const Params& P, const RankDiv& D, int x, int y) {
    const int j = (int)(((unsigned)(x - 3) * D.mw) >> 20), i = (int)(((unsigned)(y - 3) * D.mh) >> 20);
    return ((unsigned)(i * P.n_cols + j) << 14) | ((unsigned)(y - 3 - i * P.h_cell) << 7) | (unsigned)(x - 3 - j * P.w_cell);
}

// Runs the whole distribution.  cand[n_keys] (global), knode[n_keys] (global scratch), out[<= cap] = winning candidate
// records in final list order.  Returns the number of output keypoints (block-uniform); -1 on table overflow.
template <typename Ctx>
MCS_HD int distribute(Ctx& ctx, const Params& P, Shared& S, const uint32_t* cand, int n_keys, uint16_t* knode, uint32_t* out) {
    const int NT = ctx.nthreads();
    int* V = S.vars;
    if (n_keys <= 0) return 0;
    const RankDiv rdiv(P);

    // ---- roots (:558-582) --------------------------------------------------------------------------------------
    ctx.par([&] {
        for (int i = ctx.tid(); i < P.n_ini; i += NT) S.cnt[1][i] = 0;
        if (ctx.tid() == 0) {
            for (int i = 0; i < 16; i++) V[i] = 0;
        }
    });
    ctx.par([&] {
        for (int k0 = ctx.tid(); k0 < n_keys; k0 += 4 * NT) {
            uint32_t c[4];
            for (int u = 0; u < 4; u++) c[u] = k0 + u * NT < n_keys ? cand[k0 + u * NT] : 0u;
            for (int u = 0; u < 4; u++) {
                if (k0 + u * NT >= n_keys) break;
                int r = (int)((float)cand_x(c[u]) / P.hX);
                r = r < 0 ? 0 : (r >= P.n_ini ? P.n_ini - 1 : r);
                knode[k0 + u * NT] = (uint16_t)r;
                ctx.atomic_add(&S.cnt[1][r], 1);
            }
        }
    });
    ctx.serial([&] {
        // erase empty roots (:586-597); remaining roots keep their order.  map root -> list position.
        int n = 0;
        for (int i = 0; i < P.n_ini; i++) {
            if (S.cnt[1][i] > 0) {
                Geo g;
                g.ulx = (short)(int)(P.hX * (float)i);
                g.brx = (short)(int)(P.hX * (float)(i + 1));
                g.uly = 0;
                g.bry = (short)P.roi_h;
                S.geo[0][n] = g;
                S.cnt[0][n] = S.cnt[1][i];
                S.map[i * 4] = (uint16_t)n;
                n++;
            }
        }
        V[V_NLIST] = n;
        V[V_PHASE] = 1;
        V[V_CUR] = 0;
    });
    if (P.n_ini > 1) {
        ctx.par([&] {
            for (int k = ctx.tid(); k < n_keys; k += NT) knode[k] = S.map[knode[k] * 4];
        });
    }

    // ---- rounds ------------------------------------------------------------------------------------------------
    for (int round = 0; round < 64; round++) {
        const int cur = V[V_CUR], nxt = cur ^ 1;
        const int n_list = V[V_NLIST], phase = V[V_PHASE], n_prev = V[V_NEXP];
        Geo* geo = S.geo[cur];
        int* cnt = S.cnt[cur];
        int n_cand;  // nodes that may be split this round, in processing order in S.proc
        if (phase == 1) {
            // splittable nodes in list order (:605-612: bNoMore nodes are skipped)
            ctx.par([&] {
                for (int i = ctx.tid(); i < n_list; i += NT) S.s0[i] = cnt[i] > 1 ? 1 : 0;
            });
            n_cand = ctx.exclusive_scan(S.s0, n_list);
            ctx.par([&] {
                for (int i = ctx.tid(); i < n_list; i += NT) {
                    const bool split = cnt[i] > 1;
                    S.flag[i] = split ? 1 : 0;
                    if (split) S.proc[S.s0[i]] = (uint16_t)i;
                }
            });
        } else {
            // size-sorted expansion (:688-753): sort vPrevSizeAndPointerToNode, walk it from the back
            n_cand = n_prev;
            ctx.par([&] {
                for (int i = ctx.tid(); i < n_list; i += NT) S.flag[i] = 0;
            });
            ctx.par([&] {
                for (int i = ctx.tid(); i < n_prev; i += NT) {
                    const int nd = S.exp[cur][i];
                    S.sortk[i] = ((uint32_t)cnt[nd] << 12) | (uint32_t)geo[nd].ulx;
                    S.proc[i] = (uint16_t)nd;
                    S.flag[nd] = 1;
                }
            });
            ctx.serial([&] { mcs_sort::sort(S.sortk, S.proc, n_prev); });
            ctx.par([&] {  // reverse: processing order is from the back of the sorted vector
                for (int i = ctx.tid(); i < n_prev / 2; i += NT) {
                    uint16_t a = S.proc[i], b = S.proc[n_prev - 1 - i];
                    S.proc[i] = b;
                    S.proc[n_prev - 1 - i] = a;
                }
            });
        }
        // Child key counts of every node that may be split.  Round 0 counts them with a pass over the keys; every later round finds
        // them ready: the nodes that may be split in a round are children created in the round before with more than one key (phase 1
        // splits every such node, phase 2 takes its candidates from exactly that list, :688-753), so the relabelling pass of the
        // round before, which has each key's new node in hand, counted the key's quadrant inside that node as well (cc2[round & 1]).
        int* const ccur = S.cc2[round & 1];
        int* const cnext = S.cc2[(round & 1) ^ 1];
        if (round == 0) {
            ctx.par([&] {
                for (int i = ctx.tid(); i < n_cand; i += NT) {
                    int* c = &ccur[(int)S.proc[i] * 4];
                    c[0] = c[1] = c[2] = c[3] = 0;
                }
            });
            // four keys per step with all their loads issued first: the passes over the keys are bound by the latency of the
            // knode[] / cand[] reads (global memory), not by arithmetic
            ctx.par([&] {
                for (int k0 = ctx.tid(); k0 < n_keys; k0 += 4 * NT) {
                    int nd[4];
                    uint32_t c[4];
                    for (int u = 0; u < 4; u++) {
                        const int k = k0 + u * NT;
                        nd[u] = k < n_keys ? (int)knode[k] : -1;
                        c[u] = k < n_keys ? cand[k] : 0u;
                    }
                    for (int u = 0; u < 4; u++)
                        if (nd[u] >= 0 && S.flag[nd[u]]) ctx.atomic_add(&ccur[nd[u] * 4 + quadrant(geo[nd[u]], cand_x(c[u]), cand_y(c[u]))], 1);
                }
            });
        }
        // per processed node: number of non-empty children m, expandable children e
        ctx.par([&] {
            for (int p = ctx.tid(); p < n_cand; p += NT) {
                const int* c = &ccur[(int)S.proc[p] * 4];
                S.s0[p] = (c[0] > 0) + (c[1] > 0) + (c[2] > 0) + (c[3] > 0);
                S.s1[p] = (c[0] > 1) + (c[1] > 1) + (c[2] > 1) + (c[3] > 1);
            }
        });
        int n_proc = n_cand;
        if (phase == 2) {
            // stop after the first node whose split brings the list to >= N (:745); size after p = n_list + sum(m - 1)
            ctx.par([&] {
                for (int p = ctx.tid(); p < n_cand; p += NT) S.s2[p] = S.s0[p] - 1;
                if (ctx.tid() == 0) V[V_NPROC] = n_cand;
            });
            ctx.exclusive_scan(S.s2, n_cand);
            ctx.par([&] {
                for (int p = ctx.tid(); p < n_cand; p += NT) {
                    const int after = n_list + S.s2[p] + S.s0[p] - 1;
                    const int before = n_list + S.s2[p];
                    if (after >= P.N && before < P.N) V[V_NPROC] = p + 1;  // sizes are non-decreasing: unique crossing
                }
            });
            n_proc = V[V_NPROC];
            if (n_proc < n_cand) {
                ctx.par([&] {
                    for (int p = n_proc + ctx.tid(); p < n_cand; p += NT) S.flag[S.proc[p]] = 0;
                });
            }
        }
        const int T = ctx.exclusive_scan(S.s0, n_proc);  // creation index base per processed node
        const int E = ctx.exclusive_scan(S.s1, n_proc);  // base into the new expand list
        // survivors keep their relative order behind the new children
        ctx.par([&] {
            for (int i = ctx.tid(); i < n_list; i += NT) S.s2[i] = S.flag[i] ? 0 : 1;
        });
        const int Sv = ctx.exclusive_scan(S.s2, n_list);
        const int n_new = T + Sv;
        if (n_new > P.cap) return -1;
        Geo* ngeo = S.geo[nxt];
        int* ncnt = S.cnt[nxt];
        ctx.par([&] {
            for (int p = ctx.tid(); p < n_proc; p += NT) {
                const int nd = S.proc[p];
                const int* c = &ccur[nd * 4];
                int e = S.s0[p], x = S.s1[p];
                for (int q = 0; q < 4; q++) {
                    if (c[q] > 0) {
                        const int id = T - 1 - e;  // push_front: later children sit nearer the front
                        ngeo[id] = child_geo(geo[nd], q);
                        ncnt[id] = c[q];
                        S.map[nd * 4 + q] = (uint16_t)id;
                        if (c[q] > 1) {
                            S.exp[nxt][x++] = (uint16_t)id;
                            int* cn = &cnext[id * 4];  // next round's counts of this child, filled by the relabelling pass below
                            cn[0] = cn[1] = cn[2] = cn[3] = 0;
                        }
                        e++;
                    }
                }
            }
            for (int i = ctx.tid(); i < n_list; i += NT) {
                if (!S.flag[i]) {
                    const int id = T + S.s2[i];
                    ngeo[id] = geo[i];
                    ncnt[id] = cnt[i];
                    S.map[i * 4] = (uint16_t)id;
                }
            }
        });
        // Relabel the keys (node ids are list positions and the list was just rebuilt).  The same pass prepares what comes next with
        // the key's new node in hand: in the last round the per-node winner (:757-775: highest response, the first in the reference's
        // candidate order among equals) as one 64-bit atomic max of (score << 32 | ~rank); in every other round the child counts of the
        // next round (see above).
        const bool finish = (n_new >= P.N) || (n_new == n_list);  // :684 / :748
        if (finish) {
            ctx.par([&] {
                for (int i = ctx.tid(); i < n_new; i += NT) S.best64[i] = 0ull;
            });
        }
        ctx.par([&] {
            for (int k0 = ctx.tid(); k0 < n_keys; k0 += 4 * NT) {
                int nd[4];
                uint32_t c[4];
                for (int u = 0; u < 4; u++) {
                    const int k = k0 + u * NT;
                    nd[u] = k < n_keys ? (int)knode[k] : -1;
                    c[u] = k < n_keys ? cand[k] : 0u;
                }
                for (int u = 0; u < 4; u++) {
                    if (nd[u] < 0) continue;
                    const bool split = S.flag[nd[u]] != 0;
                    const int x = cand_x(c[u]), y = cand_y(c[u]);
                    const int q = split ? quadrant(geo[nd[u]], x, y) : 0;
                    const int id = S.map[nd[u] * 4 + q];
                    knode[k0 + u * NT] = (uint16_t)id;
                    if (finish)
                        ctx.atomic_max64(&S.best64[id], ((unsigned long long)cand_score(c[u]) << 32) | (0xFFFFFFFFu - cand_rank(P, rdiv, x, y)));
                    else if (split && ncnt[id] > 1)
                        ctx.atomic_add(&cnext[id * 4 + quadrant(ngeo[id], x, y)], 1);
                }
            }
            if (ctx.tid() == 0) {
                V[V_CUR] = nxt;
                V[V_NLIST] = n_new;
                V[V_NEXP] = E;
                if (!finish && phase == 1 && n_new + E * 3 > P.N) V[V_PHASE] = 2;  // :688
                V[V_FINISH] = finish ? 1 : 0;
            }
        });
        if (finish) break;
    }

    // ---- the best key of every node, in list order (:757-775): the winner's record is rebuilt from its key ----------------------
    const int n_out = V[V_NLIST];
    if (!V[V_FINISH]) {
        // 64 rounds without reaching a stop condition cannot happen (every round splits at least one node or stops); keep the
        // selection defined anyway
        ctx.par([&] {
            for (int i = ctx.tid(); i < n_out; i += NT) S.best64[i] = 0ull;
        });
        ctx.par([&] {
            for (int k = ctx.tid(); k < n_keys; k += NT) {
                const uint32_t c = cand[k];
                ctx.atomic_max64(&S.best64[knode[k]], ((unsigned long long)cand_score(c) << 32) | (0xFFFFFFFFu - cand_rank(P, rdiv, cand_x(c), cand_y(c))));
            }
        });
    }
    ctx.par([&] {
        for (int i = ctx.tid(); i < n_out; i += NT) {
            const unsigned long long b = S.best64[i];
            const unsigned rank = 0xFFFFFFFFu - (unsigned)(b & 0xFFFFFFFFull);
            const int cell = (int)(rank >> 14), ci = cell / P.n_cols, cj = cell - ci * P.n_cols;
            const int y = (int)((rank >> 7) & 0x7Fu) + 3 + ci * P.h_cell, x = (int)(rank & 0x7Fu) + 3 + cj * P.w_cell;
            out[i] = (uint32_t)x | ((uint32_t)y << 12) | ((uint32_t)(b >> 32) << 24);
        }
    });
    return n_out;
}

}  // namespace mcs_oct
#endif
