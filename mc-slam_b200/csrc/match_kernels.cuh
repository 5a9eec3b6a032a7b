// match_kernels.cuh — sm_100a kernels for the intra-rig stages of MultiCameraFrame / FrontEnd:
//   K6 hamming_match_kernel : BruteForceMatch (/root/reference/MCSlam/src/MultiCameraFrame.cpp:1024-1086) for every camera
//                             pair i<j of every frame: 256-bit XOR + popc, running top-2, ratio + distance gate fused;
//                             no distance matrix ever reaches HBM.
//   K7 merge_kernel         : computeIntraMatches(matches, old) merge step (:1100-1288) incl. the optional epipolar gate.
//   K8 triangulate_kernel   : FrontEnd::obtainLfFeatures triangulation block (MCSlam/src/FrontEnd.cpp:249-370): mask
//                             filter, normalised coordinates, N-view DLT (cv::sfm::triangulatePoints), depth gate, uv_ref,
//                             representative descriptor (MultiCameraFrame.cpp:530-567).
//   + standalone N-view (MultiCameraFrame.cpp:411-441) and 2-view f32 (:481-507) DLT kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "extract_kernels.cuh"

namespace mcs {

#ifndef MCS_MATCH_NT
#define MCS_MATCH_NT 128
#endif
#ifndef MCS_MATCH_CHUNK
#define MCS_MATCH_CHUNK 1024
#endif
#ifndef MCS_MATCH_UNROLL
#define MCS_MATCH_UNROLL 4
#endif
#define MCS_PRAGMA_(x) _Pragma(#x)
#define MCS_UNROLL(n) MCS_PRAGMA_(unroll n)
// MCS_MATCH_CHUNK train descriptors are staged in shared memory per pass (1024 = 32 KB)

struct PairList {
    int n_pairs;
    unsigned char a[28], b[28];  // camera pairs i<j in the reference's loop order (:1110-1111)
};

// 256-bit Hamming distance in bits 16.. of a sort key.  POPC issues at a quarter of the LOP3 rate, so the 8 XOR words go through a
// carry-save adder tree first (ones / twos / fours bit-planes): 4 POPC instead of 8.
__device__ __forceinline__ uint32_t hamming_key(const uint4 qa, const uint4 qb, const uint4 ta, const uint4 tb) {
    const uint32_t x0 = qa.x ^ ta.x, x1 = qa.y ^ ta.y, x2 = qa.z ^ ta.z, x3 = qa.w ^ ta.w;
    const uint32_t x4 = qb.x ^ tb.x, x5 = qb.y ^ tb.y, x6 = qb.z ^ tb.z, x7 = qb.w ^ tb.w;
    const uint32_t s1 = x0 ^ x1 ^ x2, c1 = (x0 & x1) | (x2 & (x0 ^ x1));
    const uint32_t s2 = x3 ^ x4 ^ x5, c2 = (x3 & x4) | (x5 & (x3 ^ x4));
    const uint32_t s3 = s1 ^ s2 ^ x6, c3 = (s1 & s2) | (x6 & (s1 ^ s2));
    const uint32_t t1 = c1 ^ c2 ^ c3, f1 = (c1 & c2) | (c3 & (c1 ^ c2));
    return (uint32_t)(__popc(s3) + __popc(x7)) * 65536u + (uint32_t)__popc(t1) * 131072u + (uint32_t)__popc(f1) * 262144u;
}

// match_train[(frame * n_pairs + pair) * kp_cap + q] = accepted train index of query q, or -1.
// match_d0 (optional) gets the best distance for diagnostics/tests.
__global__ void __launch_bounds__(MCS_MATCH_NT) hamming_match_kernel(int num_cams, int kp_cap, PairList pairs, int frame0,
                                                                     const uint8_t* __restrict__ desc, const int* __restrict__ kp_count,
                                                                     float dist_thresh, float ratio, int* __restrict__ match_train,
                                                                     int* __restrict__ match_d0, int* __restrict__ match_d1) {
    extern __shared__ __align__(16) uint4 s_train[];  // [MCS_MATCH_CHUNK][2]
    const int frame = frame0 + blockIdx.z, pair = blockIdx.y;  // outputs are indexed by blockIdx.z (frame0 = scratch calls)
    const int ca = frame * num_cams + pairs.a[pair], cb = frame * num_cams + pairs.b[pair];
    const int nq = kp_count[ca], nt = kp_count[cb];
    const int q0i = blockIdx.x * MCS_MATCH_NT;
    if (q0i >= nq) return;
    const int q = q0i + threadIdx.x;
    const bool active = q < nq;
    uint4 qa = make_uint4(0, 0, 0, 0), qb = qa;
    if (active) {
        const uint4* qp = reinterpret_cast<const uint4*>(desc + ((size_t)ca * kp_cap + q) * 32);
        qa = __ldg(qp);
        qb = __ldg(qp + 1);
    }
    // running top-2 as packed keys (distance << 16 | train index): the order of keys is knnMatch's order (distance, then lower
    // train index first), and best / second-best update with three min/max instead of compare-and-select chains
    uint32_t k0 = 0xFFFFFFFFu, k1 = 0xFFFFFFFFu;
    const uint4* tbase = reinterpret_cast<const uint4*>(desc + (size_t)cb * kp_cap * 32);
    for (int c0 = 0; c0 < nt; c0 += MCS_MATCH_CHUNK) {
        const int m = min(MCS_MATCH_CHUNK, nt - c0);
        __syncthreads();
        for (int t = threadIdx.x; t < m * 2; t += MCS_MATCH_NT) s_train[t] = __ldg(tbase + (size_t)c0 * 2 + t);
        __syncthreads();
        if (active) {
            MCS_UNROLL(MCS_MATCH_UNROLL)
            for (int j = 0; j < m; j++) {
                const uint4 ta = s_train[2 * j], tb = s_train[2 * j + 1];
                const uint32_t key = hamming_key(qa, qb, ta, tb) + (uint32_t)(c0 + j);
                const uint32_t hi = max(key, k0);
                k0 = min(key, k0);
                k1 = min(k1, hi);
            }
        }
    }
    const int d0 = nt >= 1 ? (int)(k0 >> 16) : (1 << 20), d1 = nt >= 2 ? (int)(k1 >> 16) : (1 << 20);
    const int i0 = nt >= 1 ? (int)(k0 & 0xFFFFu) : -1;
    if (!active) return;
    // m[0].distance < neigh_ratio * m[1].distance  and not  m[0].distance > dist_thresh   (:1061-1063), f32 arithmetic.
    // With fewer than 2 train descriptors the reference reads m[1] out of range; here: no match.
    bool ok = nt >= 2 && ((float)d0 < __fmul_rn(ratio, (float)d1)) && !((float)d0 > dist_thresh);
    const size_t o = ((size_t)blockIdx.z * pairs.n_pairs + pair) * kp_cap + q;
    match_train[o] = ok ? i0 : -1;
    if (match_d0) match_d0[o] = d0;
    if (match_d1) match_d1[o] = d1;
}

// ------------------------------------------------------------------------------------------------------------
// K6x  EXPERIMENT, not the default path (MCS_MATCH_IMMA=1): the same BruteForceMatch with the distances from the integer tensor path.
// north_star asks for XOR + __popc and no tensor cores in the matcher, so hamming_match_kernel above stays the product path; this kernel
// exists to MEASURE what an exact integer formulation would buy (profiles/README.md, round 2).
//   popc(q ^ t) = popc(q) + popc(t) - 2 * <q, t>   with the bits of q and t as 0 / 1 bytes, and <q, t> from
//   mma.sync.m16n8k32.s32.u8.u8.s32 (IMMA.16832.U8.U8 on sm_100a; the b1 xor.popc form is emulated by ptxas with 8 of these plus masking).
// One warp = 16 queries; their 256 bits live expanded in 32 A-fragment registers for the whole kernel.  The block expands MCS_IMMA_TC
// train descriptors at a time into shared memory in B-fragment order (256 B per descriptor, conflict-free 64-bit loads) together with
// the word K_t = popc(t) << 16 | train index, so that  key = K_t + (popc(q) << 16) - (dot << 17)  is the packed (distance, index) key
// of hamming_match_kernel: same top-2 by min / max, same tie order, same accept rule => identical outputs.  Any fixed permutation of
// the 256 bit positions cancels in the dot product; the one used here: k-step s = descriptor word s, thread-in-group t = byte t of
// that word, low nibble in the k = 0..15 half, high nibble in the k = 16..31 half.
// ------------------------------------------------------------------------------------------------------------
#define MCS_IMMA_TC 128   // train descriptors per shared-memory pass
#define MCS_IMMA_NT 256   // 8 warps x 16 queries

__device__ __forceinline__ uint32_t bits4_to_bytes(uint32_t n) { return ((n & 0xFu) * 0x00204081u) & 0x01010101u; }

__global__ void __launch_bounds__(MCS_IMMA_NT) hamming_match_imma_kernel(int num_cams, int kp_cap, PairList pairs, int frame0,
                                                                         const uint8_t* __restrict__ desc, const int* __restrict__ kp_count,
                                                                         float dist_thresh, float ratio, int* __restrict__ match_train,
                                                                         int* __restrict__ match_d0, int* __restrict__ match_d1) {
    __shared__ __align__(16) uint32_t s_b[MCS_IMMA_TC / 8][8][8][4][2];  // [tile][k-step][column][thread-in-group][low, high nibble]
    __shared__ uint32_t s_k[MCS_IMMA_TC];                                // popc(t) << 16 | train index
    const int frame = frame0 + blockIdx.z, pair = blockIdx.y;
    const int ca = frame * num_cams + pairs.a[pair], cb = frame * num_cams + pairs.b[pair];
    const int nq = kp_count[ca], nt = kp_count[cb];
    const int q0i = blockIdx.x * (MCS_IMMA_NT / 2);
    if (q0i >= nq) return;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, t = lane & 3;
    const int qa = q0i + warp * 16 + g, qb = qa + 8;  // the two query rows of this thread's fragments
    uint32_t a[8][4];
    uint32_t rq0 = 0, rq1 = 0;  // popc(q) << 16
    {
        uint32_t w0[8], w1[8];
#pragma unroll
        for (int s4 = 0; s4 < 2; s4++) {
            uint4 v0 = make_uint4(0, 0, 0, 0), v1 = v0;
            if (qa < nq) v0 = __ldg(reinterpret_cast<const uint4*>(desc + ((size_t)ca * kp_cap + qa) * 32) + s4);
            if (qb < nq) v1 = __ldg(reinterpret_cast<const uint4*>(desc + ((size_t)ca * kp_cap + qb) * 32) + s4);
            w0[4 * s4] = v0.x; w0[4 * s4 + 1] = v0.y; w0[4 * s4 + 2] = v0.z; w0[4 * s4 + 3] = v0.w;
            w1[4 * s4] = v1.x; w1[4 * s4 + 1] = v1.y; w1[4 * s4 + 2] = v1.z; w1[4 * s4 + 3] = v1.w;
        }
#pragma unroll
        for (int s = 0; s < 8; s++) {
            rq0 += (uint32_t)__popc(w0[s]) << 16;
            rq1 += (uint32_t)__popc(w1[s]) << 16;
            const uint32_t b0 = w0[s] >> (8 * t), b1 = w1[s] >> (8 * t);
            a[s][0] = bits4_to_bytes(b0);
            a[s][1] = bits4_to_bytes(b1);
            a[s][2] = bits4_to_bytes(b0 >> 4);
            a[s][3] = bits4_to_bytes(b1 >> 4);
        }
    }
    uint32_t k00 = 0xFFFFFFFFu, k01 = 0xFFFFFFFFu, k10 = 0xFFFFFFFFu, k11 = 0xFFFFFFFFu;  // top-2 keys of row qa / row qb (this thread's columns)
    const uint32_t* tbase = reinterpret_cast<const uint32_t*>(desc + (size_t)cb * kp_cap * 32);
    for (int c0 = 0; c0 < nt; c0 += MCS_IMMA_TC) {
        __syncthreads();
        for (int j = threadIdx.x; j < MCS_IMMA_TC; j += MCS_IMMA_NT) s_k[j] = 0x3FFF0000u;  // past the end: never among the best two
        __syncthreads();
        // expand: item = (train j, word s); 8 expanded words = [t][low, high]
        for (int it = threadIdx.x; it < MCS_IMMA_TC * 8; it += MCS_IMMA_NT) {
            const int s = it / MCS_IMMA_TC, j = it - s * MCS_IMMA_TC;  // train fastest: the 128-bit stores of a quarter-warp conflict 2-way at most
            uint32_t w = 0;
            const bool live = c0 + j < nt;
            if (live) w = __ldg(tbase + (size_t)(c0 + j) * 8 + s);
            uint4 lo, hi;  // t = 0, 1 | t = 2, 3
            lo.x = bits4_to_bytes(w); lo.y = bits4_to_bytes(w >> 4); lo.z = bits4_to_bytes(w >> 8); lo.w = bits4_to_bytes(w >> 12);
            hi.x = bits4_to_bytes(w >> 16); hi.y = bits4_to_bytes(w >> 20); hi.z = bits4_to_bytes(w >> 24); hi.w = bits4_to_bytes(w >> 28);
            uint4* dst = reinterpret_cast<uint4*>(&s_b[j >> 3][s][j & 7][0][0]);
            dst[0] = lo;
            dst[1] = hi;
            if (live) atomicAdd(&s_k[j], ((uint32_t)__popc(w) << 16) + (s == 0 ? (uint32_t)(c0 + j) - 0x3FFF0000u : 0u));
        }
        __syncthreads();
        const int ntile = (min(MCS_IMMA_TC, nt - c0) + 7) >> 3;
        for (int tile = 0; tile < ntile; tile++) {
            int d0 = 0, d1 = 0, d2 = 0, d3 = 0;
#pragma unroll
            for (int s = 0; s < 8; s++) {
                const uint2 b = *reinterpret_cast<const uint2*>(&s_b[tile][s][g][t][0]);
                asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.u8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                             : "+r"(d0), "+r"(d1), "+r"(d2), "+r"(d3)
                             : "r"(a[s][0]), "r"(a[s][1]), "r"(a[s][2]), "r"(a[s][3]), "r"(b.x), "r"(b.y));
            }
            // d0: (row qa, column 2t), d1: (qa, 2t + 1), d2: (qb, 2t), d3: (qb, 2t + 1)
            const uint2 kt = *reinterpret_cast<const uint2*>(&s_k[tile * 8 + 2 * t]);
            uint32_t key, hi;
            key = kt.x + rq0 - ((uint32_t)d0 << 17); hi = max(key, k00); k00 = min(key, k00); k01 = min(k01, hi);
            key = kt.y + rq0 - ((uint32_t)d1 << 17); hi = max(key, k00); k00 = min(key, k00); k01 = min(k01, hi);
            key = kt.x + rq1 - ((uint32_t)d2 << 17); hi = max(key, k10); k10 = min(key, k10); k11 = min(k11, hi);
            key = kt.y + rq1 - ((uint32_t)d3 << 17); hi = max(key, k10); k10 = min(key, k10); k11 = min(k11, hi);
        }
    }
    // merge the four column groups of a row: top-2 of the union of two top-2 lists
#pragma unroll
    for (int o = 1; o <= 2; o <<= 1) {
        uint32_t b0 = __shfl_xor_sync(0xFFFFFFFFu, k00, o), b1 = __shfl_xor_sync(0xFFFFFFFFu, k01, o);
        k01 = min(max(k00, b0), min(k01, b1));
        k00 = min(k00, b0);
        b0 = __shfl_xor_sync(0xFFFFFFFFu, k10, o);
        b1 = __shfl_xor_sync(0xFFFFFFFFu, k11, o);
        k11 = min(max(k10, b0), min(k11, b1));
        k10 = min(k10, b0);
    }
    if (t != 0) return;
#pragma unroll
    for (int r = 0; r < 2; r++) {
        const int q = r ? qb : qa;
        if (q >= nq) continue;
        const uint32_t k0 = r ? k10 : k00, k1 = r ? k11 : k01;
        const int e0 = nt >= 1 ? (int)(k0 >> 16) : (1 << 20), e1 = nt >= 2 ? (int)(k1 >> 16) : (1 << 20);
        const int i0 = nt >= 1 ? (int)(k0 & 0xFFFFu) : -1;
        const bool ok = nt >= 2 && ((float)e0 < __fmul_rn(ratio, (float)e1)) && !((float)e0 > dist_thresh);
        const size_t o = ((size_t)blockIdx.z * pairs.n_pairs + pair) * kp_cap + q;
        match_train[o] = ok ? i0 : -1;
        if (match_d0) match_d0[o] = e0;
        if (match_d1) match_d1[o] = e1;
    }
}

// ------------------------------------------------------------------------------------------------------------
// K6b  knn2: DescriptorMatcher("BruteForce-Hamming")->knnMatch(query, train, k = 2) as the inter-frame matchers call it
// (/root/reference/MCSlam/src/FrontEnd.cpp:3134-3135 findMatchesMono, :3199-3201 findInterMatches), for many independent
// (query, train) sets per launch.  Output per query row: (trainIdx0, dist0, trainIdx1, dist1), -1 where the train set is
// shorter than k; ties keep the lower train index first like cv::batchDistance's sort.  The ratio / distance / 3-D gates
// stay with the caller because the two reference call sites apply different ones.
// ------------------------------------------------------------------------------------------------------------
struct Knn2Set {
    int q_ofs, nq, t_ofs, nt;  // row offsets into the query / train descriptor arrays; with `counts`: nq, nt index it
    int out_ofs, pad0, pad1, pad2;
};

__global__ void __launch_bounds__(MCS_MATCH_NT) knn2_kernel(const uint8_t* __restrict__ qdesc, const uint8_t* __restrict__ tdesc,
                                                            const Knn2Set* __restrict__ sets, const int* __restrict__ counts,
                                                            int4* __restrict__ out) {
    extern __shared__ __align__(16) uint4 s_train[];  // [MCS_MATCH_CHUNK][2]
    const Knn2Set S = sets[blockIdx.y];
    const int nq = counts ? counts[S.nq] : S.nq, nt = counts ? counts[S.nt] : S.nt;
    const int q0i = blockIdx.x * MCS_MATCH_NT;
    if (q0i >= nq) return;
    const int q = q0i + threadIdx.x;
    const bool active = q < nq;
    uint4 qa = make_uint4(0, 0, 0, 0), qb = qa;
    if (active) {
        const uint4* qp = reinterpret_cast<const uint4*>(qdesc + ((size_t)S.q_ofs + q) * 32);
        qa = __ldg(qp);
        qb = __ldg(qp + 1);
    }
    uint32_t k0 = 0xFFFFFFFFu, k1 = 0xFFFFFFFFu;
    const uint4* tbase = reinterpret_cast<const uint4*>(tdesc + (size_t)S.t_ofs * 32);
    for (int c0 = 0; c0 < nt; c0 += MCS_MATCH_CHUNK) {
        const int m = min(MCS_MATCH_CHUNK, nt - c0);
        __syncthreads();
        for (int t = threadIdx.x; t < m * 2; t += MCS_MATCH_NT) s_train[t] = __ldg(tbase + (size_t)c0 * 2 + t);
        __syncthreads();
        if (active) {
#pragma unroll 4
            for (int j = 0; j < m; j++) {
                const uint32_t key = hamming_key(qa, qb, s_train[2 * j], s_train[2 * j + 1]) + (uint32_t)(c0 + j);
                const uint32_t hi = max(key, k0);
                k0 = min(key, k0);
                k1 = min(k1, hi);
            }
        }
    }
    if (!active) return;
    int4 r;
    r.x = nt >= 1 ? (int)(k0 & 0xFFFFu) : -1;
    r.y = nt >= 1 ? (int)(k0 >> 16) : -1;
    r.z = nt >= 2 ? (int)(k1 & 0xFFFFu) : -1;
    r.w = nt >= 2 ? (int)(k1 >> 16) : -1;
    out[(size_t)S.out_ofs + q] = r;
}

// ------------------------------------------------------------------------------------------------------------
// K6c  ORBextractor::getMatches_distRatio (/root/reference/MCSlam/src/ORBextractor.cpp:1228-1290) for many candidate-list pairs
// ("groups": the features of one vocabulary node in image A and in image B) per launch; one block per group.
// Phase 1, thread per query of i_A: best / second-best candidate of i_B with strict '<' (first candidate wins ties, :1253-1262);
// accepted when best <= th_low and (double)best / (double)second <= max_ratio (:1264-1265; second = 1e9 with one candidate).
// Phase 2, warp 0, queries in list order: a B feature claimed twice keeps the closer A feature at the position of the first claim
// (:1268-1284); the linear search of the reference runs 32 entries per step.
// Outputs of group g start at a_ofs[g] (a group makes at most |i_A| matches); group_count / group_dups give the match count and
// the number of duplicate-claim distance re-evaluations (BookK bookkeeping).
// ------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ int hamming256(const uint4 a0, const uint4 a1, const uint4 b0, const uint4 b1) {
    return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) + __popc(a1.x ^ b1.x) +
           __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

__global__ void __launch_bounds__(128) subset_match_kernel(const uint8_t* __restrict__ A, const uint8_t* __restrict__ B,
                                                           const int* __restrict__ a_ofs, const uint32_t* __restrict__ a_idx,
                                                           const int* __restrict__ b_ofs, const uint32_t* __restrict__ b_idx, int th_low,
                                                           double max_ratio, uint32_t* __restrict__ tmp_b, int* __restrict__ tmp_d,
                                                           uint32_t* __restrict__ match_a, uint32_t* __restrict__ match_b,
                                                           int* __restrict__ match_d, int* __restrict__ group_count,
                                                           int* __restrict__ group_dups) {
    const int g = blockIdx.x;
    const int a0 = a_ofs[g], na = a_ofs[g + 1] - a0, b0 = b_ofs[g], nb = b_ofs[g + 1] - b0;
    for (int i = threadIdx.x; i < na; i += blockDim.x) {
        const uint4* qa = reinterpret_cast<const uint4*>(A + (size_t)a_idx[a0 + i] * 32);
        const uint4 q0 = __ldg(qa), q1 = __ldg(qa + 1);
        int best_j = -1, best1 = 1 << 30, best2 = 1 << 30;
        for (int j = 0; j < nb; j++) {
            const uint4* tb = reinterpret_cast<const uint4*>(B + (size_t)__ldg(b_idx + b0 + j) * 32);
            const int d = hamming256(q0, q1, __ldg(tb), __ldg(tb + 1));
            if (d < best1) { best_j = j; best2 = best1; best1 = d; }
            else if (d < best2) best2 = d;
        }
        const double d2 = best2 == (1 << 30) ? 1e9 : (double)best2;
        const bool ok = best_j >= 0 && best1 <= th_low && __ddiv_rn((double)best1, d2) <= max_ratio;
        tmp_b[a0 + i] = ok ? b_idx[b0 + best_j] : 0xFFFFFFFFu;
        tmp_d[a0 + i] = best1;
    }
    __syncthreads();
    if (threadIdx.x >= 32) return;
    const int lane = threadIdx.x;
    int n = 0, dups = 0;
    for (int i = 0; i < na; i++) {
        const uint32_t idxB = tmp_b[a0 + i];
        if (idxB == 0xFFFFFFFFu) continue;  // warp-uniform
        int pos = -1;
        for (int k0 = 0; k0 < n && pos < 0; k0 += 32) {
            const int k = k0 + lane;
            const unsigned hit = __ballot_sync(0xFFFFFFFFu, k < n && match_b[a0 + k] == idxB);
            if (hit) pos = k0 + __ffs(hit) - 1;
        }
        if (lane == 0) {
            const int d1 = tmp_d[a0 + i];
            if (pos < 0) {
                match_b[a0 + n] = idxB;
                match_a[a0 + n] = a_idx[a0 + i];
                match_d[a0 + n] = d1;
            } else if (d1 < match_d[a0 + pos]) {  // distance of the current holder to idxB == its own best distance
                match_a[a0 + pos] = a_idx[a0 + i];
                match_d[a0 + pos] = d1;
            }
        }
        if (pos < 0) n++; else dups++;
        __syncwarp();
    }
    if (lane == 0) {
        group_count[g] = n;
        group_dups[g] = dups;
    }
}

// ------------------------------------------------------------------------------------------------------------
// MultiCameraFrame::UndistortKeyPoints (:300-347, the RECTIFY = false branch of extractFeatureSingle): image_kps_undist =
// cv::undistortPoints(image_kps, K32, dist32, noArray(), K32) with K and the coefficients narrowed to f32 first (:318-322).  OpenCV's
// undistortPoints: f64, 5 fixed-point iterations of the inverse radial-tangential model, re-projection through K, narrowed to f32.
// A camera is copied unchanged when `dist_coeffs_[cam].at<float>(0) == 0.0` (:302) — the reference reads the first f64 coefficient
// through a float pointer, i.e. it tests the LOW 32 bits of the double; `copy_cam` carries that decision from the host.
// ------------------------------------------------------------------------------------------------------------
struct UndistPtsCam {
    double fx, fy, cx, cy;  // (double)(float)K
    double k[5];            // (double)(float)dist: k1 k2 p1 p2 k3
    int copy_cam, pad;
};

__global__ void __launch_bounds__(128) undistort_points_kernel(const KeypointPod* __restrict__ in, KeypointPod* __restrict__ out,
                                                               const int* __restrict__ kp_count, int kp_cap,
                                                               const UndistPtsCam* __restrict__ cams, int num_cams, int img0) {
    const int img = blockIdx.y, q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= kp_count[img]) return;
    const UndistPtsCam c = cams[(img0 + img) % num_cams];
    KeypointPod p = in[(size_t)img * kp_cap + q];
    if (!c.copy_cam) {
        const double ifx = 1. / c.fx, ify = 1. / c.fy;
        double x = ((double)p.x - c.cx) * ifx, y = ((double)p.y - c.cy) * ify;
        const double x0 = x, y0 = y;
        for (int it = 0; it < 5; it++) {
            const double r2 = x * x + y * y;
            const double icdist = (1 + ((0 * r2 + 0) * r2 + 0) * r2) / (1 + ((c.k[4] * r2 + c.k[1]) * r2 + c.k[0]) * r2);
            if (icdist < 0) { x = x0; y = y0; break; }
            const double dX = 2 * c.k[2] * x * y + c.k[3] * (r2 + 2 * x * x) + 0 * r2 + 0 * r2 * r2;
            const double dY = c.k[2] * (r2 + 2 * y * y) + 2 * c.k[3] * x * y + 0 * r2 + 0 * r2 * r2;
            x = (x0 - dX) * icdist;
            y = (y0 - dY) * icdist;
        }
        const double xx = c.fx * x + 0 * y + c.cx, yy = 0 * x + c.fy * y + c.cy, ww = 1. / (0 * x + 0 * y + 1.0);
        p.x = (float)(xx * ww);
        p.y = (float)(yy * ww);
    }
    out[(size_t)img * kp_cap + q] = p;
}

// ------------------------------------------------------------------------------------------------------------
// DBoW2 vocabulary transform (external dependency of the reference, not vendored: DBoW2 as forked for ORB-SLAM2 with
// loadFromTextFile; called as orb_vocabulary->transform(vec_desc, bowVec, featVec, 4) at MultiCameraFrame.cpp:257 and
// FrontEnd.cpp:533).  TemplatedVocabulary::transform(feature, word_id, weight, nid, levelsup): from the root, at every level move
// to the child whose descriptor is nearest in Hamming distance (first child wins ties, strict '<'), remember the node reached at
// level L - levelsup, stop at a leaf; word id / weight are the leaf's.  One thread per feature; children are stored contiguously.
// ------------------------------------------------------------------------------------------------------------
struct VocabNode {
    int child_begin, child_count;  // children are nodes [child_begin, child_begin + child_count) of the reordered table
    int word_id;                   // >= 0 for leaves
    int orig_id;                   // node id in the file's numbering (what FeatureVector keys use)
};

__global__ void __launch_bounds__(128) vocab_transform_kernel(const VocabNode* __restrict__ nodes, const uint4* __restrict__ node_desc,
                                                              const double* __restrict__ node_weight, int L, int levelsup,
                                                              const uint8_t* __restrict__ desc, const int* __restrict__ counts, int stride,
                                                              int* __restrict__ word_id, int* __restrict__ node_id,
                                                              double* __restrict__ weight) {
    const int set = blockIdx.y, q = blockIdx.x * blockDim.x + threadIdx.x;
    const int n = counts ? counts[set] : stride;
    if (q >= n) return;
    const size_t o = (size_t)set * stride + q;
    const uint4* f = reinterpret_cast<const uint4*>(desc + o * 32);
    const uint4 f0 = __ldg(f), f1 = __ldg(f + 1);
    const int nid_level = L - levelsup;
    int cur = 0, level = 0, nid = 0;
    do {
        level++;
        const VocabNode nd = nodes[cur];
        int best = nd.child_begin, best_d = 1 << 30;
        for (int c = 0; c < nd.child_count; c++) {
            const int id = nd.child_begin + c;
            const int d = hamming256(f0, f1, __ldg(node_desc + 2 * id), __ldg(node_desc + 2 * id + 1));
            if (d < best_d) { best_d = d; best = id; }
        }
        cur = best;
        if (level == nid_level) nid = nodes[cur].orig_id;
    } while (nodes[cur].child_count > 0);
    word_id[o] = nodes[cur].word_id;
    node_id[o] = nid_level <= 0 ? 0 : nid;
    weight[o] = node_weight[cur];
}

// ------------------------------------------------------------------------------------------------------------
// K6d  inner loop of the BoW-guided computeIntraMatches(matches, words_) (/root/reference/MCSlam/src/MultiCameraFrame.cpp:716-751):
// for a vocabulary node seen by cameras cam1 < cam2, per cam1 feature of the node the best / second-best cam2 feature of the node
// whose undistorted keypoint lies within y_gate rows (|dy| >= 50 skipped, :731), strict '<' so the first candidate wins.  The
// order-dependent row bookkeeping (:752-816) stays on the host; one block per (node, camera pair) group, raw results out.
// best_j: position inside the group's cam2 list or -1; d1 / d2: distances, 1000000000 standing for the reference's 1e9.
// ------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(64) bow_best2_kernel(const uint8_t* __restrict__ desc, const KeypointPod* __restrict__ kps_undist, int kp_cap,
                                                       const int* __restrict__ img_a, const int* __restrict__ img_b,
                                                       const int* __restrict__ a_ofs, const uint32_t* __restrict__ a_idx,
                                                       const int* __restrict__ b_ofs, const uint32_t* __restrict__ b_idx, float y_gate,
                                                       int* __restrict__ best_j, int* __restrict__ d1, int* __restrict__ d2) {
    const int g = blockIdx.x;
    const int a0 = a_ofs[g], na = a_ofs[g + 1] - a0, b0 = b_ofs[g], nb = b_ofs[g + 1] - b0;
    const size_t ia = (size_t)img_a[g] * kp_cap, ib = (size_t)img_b[g] * kp_cap;
    for (int i = threadIdx.x; i < na; i += blockDim.x) {
        const uint32_t fa = a_idx[a0 + i];
        const uint4* qa = reinterpret_cast<const uint4*>(desc + (ia + fa) * 32);
        const uint4 q0 = __ldg(qa), q1 = __ldg(qa + 1);
        const float ya = kps_undist[ia + fa].y;
        int bj = -1, best1 = 1000000000, best2 = 1000000000;
        for (int j = 0; j < nb; j++) {
            const uint32_t fb = __ldg(b_idx + b0 + j);
            if (fabsf(__fsub_rn(ya, kps_undist[ib + fb].y)) >= y_gate) continue;
            const uint4* tb = reinterpret_cast<const uint4*>(desc + (ib + fb) * 32);
            const int d = hamming256(q0, q1, __ldg(tb), __ldg(tb + 1));
            if (d < best1) { bj = j; best2 = best1; best1 = d; }
            else if (d < best2) best2 = d;
        }
        best_j[a0 + i] = bj;
        d1[a0 + i] = best1;
        d2[a0 + i] = best2;
    }
}

// ------------------------------------------------------------------------------------------------------------
// K7  merge: computeIntraMatches(matches, old) after the pair matcher (:1222-1266).  The reference walks the matches of the pairs
// in order and mutates match_inv_idx / the row table as it goes, so later matches and later pairs see earlier state.  Pairs stay
// sequential (one block per frame walks them), but inside one pair (i, j) the walk decomposes:
//   * a match (fi, fj) reads a = inv_i[fi] and b = inv_j[fj].  inv_i[fi] is only ever written by fi's own match, so a is the
//     pre-pair value.  inv_j[fj] is shared by the matches with the same fj ("group").
//   * b == -1: the group's FIRST match in query order acts — a == -1: new row;  a != -1: R[a][j] = fj, inv_j[fj] = a (no check,
//     :1260-1264; a row shared by several fi through stale inverse entries keeps the LAST writer in query order).  Every later
//     match of the group then sees b != -1 with R[b][i] occupied and does nothing.
//   * b != -1 (pre-pair): only matches with a == -1 can act, and only while R[b][i] == -1: the first in query order over all
//     groups that share the row b takes the slot (:1238-1247).
//   * new row ids are handed out in query order: block-wide exclusive scan of the "creates a row" flags.
// Three block-wide passes per pair instead of a one-thread walk; bit-identical rows (same order, same ids).
// Shared memory: match_inv_idx (u16, 0xFFFF = -1), per-train-feature group minimum, per-row first / last claimant, per-query
// decision, and (when it fits) the row table itself.
// Fmat: per pair the f64 fundamental matrix built on the host (:1126-1141); sigma2 = mvLevelSigma2.
// ------------------------------------------------------------------------------------------------------------
#define MCS_MERGE_NT 256
#define MCS_MERGE_NT_WIDE 1024

struct MergeSmem {
    size_t inv, grp, rfirst, rlast, val, kind, rows, total;
};
__host__ __device__ inline MergeSmem merge_smem_layout(int num_cams, int kp_cap, int row_cap, bool rows_in_smem) {
    MergeSmem L;
    size_t o = 0;
    auto take = [&o](size_t bytes) { const size_t at = o; o += (bytes + 15) & ~(size_t)15; return at; };
    L.inv = take((size_t)num_cams * kp_cap * 2);
    L.grp = take((size_t)kp_cap * 2);
    L.rfirst = take((size_t)row_cap * 2);
    L.rlast = take((size_t)row_cap * 2);
    L.val = take((size_t)kp_cap * 2);
    L.kind = take((size_t)kp_cap);
    L.rows = take(rows_in_smem ? (size_t)row_cap * num_cams * 4 : 0);
    L.total = o;
    return L;
}

// 16-bit atomic min / max in shared memory (CAS on the containing word; the claim tables see little contention)
__device__ __forceinline__ void atomic_min_u16(uint16_t* p, uint32_t v) {
    uint32_t* w = reinterpret_cast<uint32_t*>(reinterpret_cast<uintptr_t>(p) & ~(uintptr_t)3);
    const int sh = (int)(reinterpret_cast<uintptr_t>(p) & 2) * 8;
    uint32_t old = *w;
    while (((old >> sh) & 0xFFFFu) > v) {
        const uint32_t prev = atomicCAS(w, old, (old & ~(0xFFFFu << sh)) | (v << sh));
        if (prev == old) break;
        old = prev;
    }
}
__device__ __forceinline__ void atomic_max_u16(uint16_t* p, uint32_t v) {
    uint32_t* w = reinterpret_cast<uint32_t*>(reinterpret_cast<uintptr_t>(p) & ~(uintptr_t)3);
    const int sh = (int)(reinterpret_cast<uintptr_t>(p) & 2) * 8;
    uint32_t old = *w;
    while (((old >> sh) & 0xFFFFu) < v) {
        const uint32_t prev = atomicCAS(w, old, (old & ~(0xFFFFu << sh)) | (v << sh));
        if (prev == old) break;
        old = prev;
    }
}

__device__ __forceinline__ bool epipolar_ok(const KeypointPod& p1, const KeypointPod& p2, const double* __restrict__ F,
                                            const float* __restrict__ sigma2) {
    // :1177-1205 — products in f64 (no FMA contraction), narrowed to f32 exactly where the reference does
    float a = (float)__dadd_rn(__dadd_rn(__dmul_rn((double)p2.x, F[0]), __dmul_rn((double)p2.y, F[3])), F[6]);
    float b = (float)__dadd_rn(__dadd_rn(__dmul_rn((double)p2.x, F[1]), __dmul_rn((double)p2.y, F[4])), F[7]);
    float c = (float)__dadd_rn(__dadd_rn(__dmul_rn((double)p2.x, F[2]), __dmul_rn((double)p2.y, F[5])), F[8]);
    float den = __fadd_rn(__fmul_rn(a, a), __fmul_rn(b, b));
    den = den != 0.f ? (float)__ddiv_rn(1.0, __dsqrt_rn((double)den)) : 1.f;
    a = __fmul_rn(a, den);
    b = __fmul_rn(b, den);
    c = __fmul_rn(c, den);
    den = __fadd_rn(__fmul_rn(a, a), __fmul_rn(b, b));
    const float num = __fadd_rn(__fadd_rn(__fmul_rn(a, p1.x), __fmul_rn(b, p1.y)), c);
    if (den == 0.f) return false;
    const float dsqr = __fdiv_rn(__fmul_rn(num, num), den);
    const float thr = (float)__dmul_rn(3.84, (double)sigma2[p1.octave]);
    return dsqr < thr;
}

// NT = MCS_MERGE_NT inside the chunked pipeline (one block per frame, many frames per launch); MCS_MERGE_NT_WIDE for launches of a
// few frames, where the block's latency is the stage's: one query per thread instead of four shortens every pass of every pair.
template <int NT>
__global__ void __launch_bounds__(NT) merge_kernel(int num_cams, int kp_cap, int row_cap, PairList pairs,
                                                   const int* __restrict__ kp_count, const int* __restrict__ match_train,
                                                   const KeypointPod* __restrict__ kps, int use_epipolar,
                                                   const double* __restrict__ Fmat, const float* __restrict__ sigma2,
                                                   int* __restrict__ rows, int* __restrict__ row_count, int rows_in_smem) {
    enum { K_NONE = 0, K_VALID = 1, K_NEW = 2, K_ATT_A = 3, K_ATT_B = 4 };
    constexpr int BIG = 0xFFFF;
    extern __shared__ __align__(16) unsigned char s_raw[];
    __shared__ int s_wsum[NT / 32];
    const MergeSmem L = merge_smem_layout(num_cams, kp_cap, row_cap, rows_in_smem != 0);
    uint16_t* s_inv = reinterpret_cast<uint16_t*>(s_raw + L.inv);
    uint16_t* s_grp = reinterpret_cast<uint16_t*>(s_raw + L.grp);        // [fj] smallest query index of the valid matches onto fj
    uint16_t* s_rfirst = reinterpret_cast<uint16_t*>(s_raw + L.rfirst);  // [row] smallest query index claiming the row's column i
    uint16_t* s_rlast = reinterpret_cast<uint16_t*>(s_raw + L.rlast);    // [row] largest (query index + 1) writing the row's column j
    uint16_t* s_val = reinterpret_cast<uint16_t*>(s_raw + L.val);  // [q] row the decision refers to
    uint8_t* s_kind = s_raw + L.kind;                              // [q] decision of the current pair
    const int frame = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    int* G = rows + (size_t)frame * row_cap * num_cams;
    int* R = rows_in_smem ? reinterpret_cast<int*>(s_raw + L.rows) : G;
    for (int t = tid; t < num_cams * kp_cap; t += NT) s_inv[t] = 0xFFFFu;
    for (int t = tid; t < kp_cap; t += NT) { s_grp[t] = BIG; s_kind[t] = K_NONE; }
    for (int t = tid; t < row_cap; t += NT) { s_rfirst[t] = BIG; s_rlast[t] = 0; }
    // every thread owns the same contiguous range of query indices in all pairs (needed for the ordered scan)
    const int IT = (kp_cap + NT - 1) / NT;
    const int q_lo = tid * IT;
    int M = 0;
    __syncthreads();
    for (int p = 0; p < pairs.n_pairs; p++) {
        const int i = pairs.a[p], j = pairs.b[p];
        const int nq = kp_count[frame * num_cams + i];
        const int q_hi = min(q_lo + IT, nq);
        const int* mt = match_train + ((size_t)frame * pairs.n_pairs + p) * kp_cap;
        uint16_t* invi = s_inv + i * kp_cap;
        uint16_t* invj = s_inv + j * kp_cap;
        const KeypointPod* ki = kps + (size_t)(frame * num_cams + i) * kp_cap;
        const KeypointPod* kj = kps + (size_t)(frame * num_cams + j) * kp_cap;
        // pass A: valid matches (epipolar gate when `old`), group minimum per train feature
        for (int q = q_lo; q < q_hi; q++) {
            const int fj = __ldg(&mt[q]);
            bool ok = fj >= 0;
            if (ok && use_epipolar) ok = epipolar_ok(ki[q], kj[fj], Fmat + p * 9, sigma2);
            s_kind[q] = ok ? K_VALID : K_NONE;
            if (ok) atomic_min_u16(&s_grp[fj], (uint32_t)q);
        }
        __syncthreads();
        // pass B: decisions against the pre-pair state
        for (int q = q_lo; q < q_hi; q++) {
            if (s_kind[q] == K_NONE) continue;
            const int fj = __ldg(&mt[q]);
            const int a = invi[q], b = invj[fj];
            int kind = K_VALID, val = 0;  // valid but without effect unless one of the cases below applies
            if (b == 0xFFFF) {
                if (s_grp[fj] == q) {
                    if (a == 0xFFFF) kind = K_NEW;
                    else { kind = K_ATT_A; val = a; atomic_max_u16(&s_rlast[a], (uint32_t)(q + 1)); }
                }
            } else if (a == 0xFFFF && R[(size_t)b * num_cams + i] == -1) {
                kind = K_ATT_B; val = b; atomic_min_u16(&s_rfirst[b], (uint32_t)q);
            }
            s_kind[q] = (uint8_t)kind;
            s_val[q] = (uint16_t)val;
        }
        // ordered ids of the new rows
        int my_new = 0;
        __syncthreads();
        for (int q = q_lo; q < q_hi; q++) my_new += s_kind[q] == K_NEW;
        int incl = my_new;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xFFFFFFFFu, incl, o);
            if (lane >= o) incl += v;
        }
        if (lane == 31) s_wsum[warp] = incl;
        __syncthreads();
        // offsets of the warps: one shuffle scan over the (at most 32) warp sums instead of a loop over them in every thread
        int wincl = lane < NT / 32 ? s_wsum[lane] : 0;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xFFFFFFFFu, wincl, o);
            if (lane >= o) wincl += v;
        }
        const int total_new = __shfl_sync(0xFFFFFFFFu, wincl, 31);
        const int wexcl = __shfl_sync(0xFFFFFFFFu, wincl, warp) - s_wsum[warp];
        int base = M + wexcl + incl - my_new;
        // pass C: apply the decisions (they touch disjoint state)
        for (int q = q_lo; q < q_hi; q++) {
            const int kind = s_kind[q];
            if (kind < K_NEW) continue;
            const int fj = __ldg(&mt[q]);
            if (kind == K_NEW) {
                const int id = base++;
                if (id < row_cap) {
                    int* r = R + (size_t)id * num_cams;
                    for (int c = 0; c < num_cams; c++) r[c] = -1;
                    r[i] = q;
                    r[j] = fj;
                    invi[q] = (uint16_t)id;
                    invj[fj] = (uint16_t)id;
                }
            } else if (kind == K_ATT_A) {
                const int a = s_val[q];
                invj[fj] = (uint16_t)a;
                if (s_rlast[a] == q + 1) R[(size_t)a * num_cams + j] = fj;
            } else {
                const int b = s_val[q];
                if (s_rfirst[b] == q) {
                    R[(size_t)b * num_cams + i] = q;
                    invi[q] = (uint16_t)b;
                }
            }
        }
        M += total_new;
        __syncthreads();
        // reset the claim tables for the next pair
        for (int q = q_lo; q < q_hi; q++) {
            const int kind = s_kind[q];
            if (kind == K_NONE) continue;
            s_grp[__ldg(&mt[q])] = BIG;
            if (kind == K_ATT_A) s_rlast[s_val[q]] = 0;
            else if (kind == K_ATT_B) s_rfirst[s_val[q]] = BIG;
            s_kind[q] = K_NONE;
        }
        __syncthreads();
    }
    if (tid == 0) row_count[frame] = M;
    if (rows_in_smem) {
        const int n = min(M, row_cap) * num_cams;
        for (int t = tid; t < n; t += NT) G[t] = R[t];
    }
}

// ------------------------------------------------------------------------------------------------------------
// One-sided (Hestenes) Jacobi SVD: right singular vector of the smallest singular value of A (m x n) = cv::SVD::solveZ.
// Per-thread, arrays in local memory.  T = double for the sfm path, float for MultiCameraFrame::triangulate.
// ------------------------------------------------------------------------------------------------------------
template <typename T, int MAXM, int MAXN>
__device__ void null_vector(T (&A)[MAXM * MAXN], T (&V)[MAXN * MAXN], int m, int n, T* z, T eps) {
    for (int i = 0; i < n * n; i++) V[i] = 0;
    for (int i = 0; i < n; i++) V[i * n + i] = 1;
    for (int sweep = 0; sweep < 60; sweep++) {
        bool changed = false;
        for (int p = 0; p < n - 1; p++)
            for (int q = p + 1; q < n; q++) {
                T a = 0, b = 0, g = 0;
                for (int k = 0; k < m; k++) {
                    const T x = A[k * n + p], y = A[k * n + q];
                    a += x * x;
                    b += y * y;
                    g += x * y;
                }
                if (fabs((double)g) <= (double)eps * sqrt((double)a * (double)b) || g == 0) continue;
                changed = true;
                const T zeta = (b - a) / (2 * g);
                const T t = (zeta >= 0 ? (T)1 : (T)-1) / ((T)fabs((double)zeta) + (T)sqrt((double)(1 + zeta * zeta)));
                const T c = (T)1 / (T)sqrt((double)(1 + t * t)), s = c * t;
                for (int k = 0; k < m; k++) {
                    const T x = A[k * n + p], y = A[k * n + q];
                    A[k * n + p] = c * x - s * y;
                    A[k * n + q] = s * x + c * y;
                }
                for (int k = 0; k < n; k++) {
                    const T x = V[k * n + p], y = V[k * n + q];
                    V[k * n + p] = c * x - s * y;
                    V[k * n + q] = s * x + c * y;
                }
            }
        if (!changed) break;
    }
    int best = 0;
    T bestn = -1;
    for (int j = 0; j < n; j++) {
        T s = 0;
        for (int k = 0; k < m; k++) s += A[k * n + j] * A[k * n + j];
        if (bestn < 0 || s < bestn) {
            bestn = s;
            best = j;
        }
    }
    for (int k = 0; k < n; k++) z[k] = V[k * n + best];
}

struct CalibDev {  // per camera: K (9), [R|t] (12), K[R|t] (12), all f64 row-major
    double K[9];
    double Rt[12];
    double P[12];
};

__device__ __forceinline__ int hamming256(const uint8_t* a, const uint8_t* b) {
    const uint4* pa = reinterpret_cast<const uint4*>(a);
    const uint4* pb = reinterpret_cast<const uint4*>(b);
    const uint4 a0 = pa[0], a1 = pa[1], b0 = pb[0], b1 = pb[1];
    return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) + __popc(a1.x ^ b1.x) +
           __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

// Same algorithm and operation order as null_vector (so results stay bit-identical to the oracle's f64 Jacobi), with compile-time
// sizes: exact-size arrays keep the per-thread footprint small (L1-resident), and for the small systems (UNROLL) the pair loops are
// unrolled so A and V live in registers.
template <typename T, int M, int N, bool UNROLL>
__device__ __forceinline__ void null_vector_fixed(T (&A)[M * N], T* z, T eps) {
    T V[N * N];
#pragma unroll
    for (int i = 0; i < N * N; i++) V[i] = 0;
#pragma unroll
    for (int i = 0; i < N; i++) V[i * N + i] = 1;
#pragma unroll 1
    for (int sweep = 0; sweep < 60; sweep++) {
        bool changed = false;
#pragma unroll(UNROLL ? N : 1)
        for (int p = 0; p < N - 1; p++) {
#pragma unroll(UNROLL ? N : 1)
            for (int q = p + 1; q < N; q++) {
                T a = 0, b = 0, g = 0;
#pragma unroll
                for (int k = 0; k < M; k++) {
                    const T x = A[k * N + p], y = A[k * N + q];
                    a += x * x;
                    b += y * y;
                    g += x * y;
                }
                if (fabs((double)g) <= (double)eps * sqrt((double)a * (double)b) || g == 0) continue;
                changed = true;
                const T zeta = (b - a) / (2 * g);
                const T t = (zeta >= 0 ? (T)1 : (T)-1) / ((T)fabs((double)zeta) + (T)sqrt((double)(1 + zeta * zeta)));
                const T c = (T)1 / (T)sqrt((double)(1 + t * t)), sn = c * t;
#pragma unroll
                for (int k = 0; k < M; k++) {
                    const T x = A[k * N + p], y = A[k * N + q];
                    A[k * N + p] = c * x - sn * y;
                    A[k * N + q] = sn * x + c * y;
                }
#pragma unroll
                for (int k = 0; k < N; k++) {
                    const T x = V[k * N + p], y = V[k * N + q];
                    V[k * N + p] = c * x - sn * y;
                    V[k * N + q] = sn * x + c * y;
                }
            }
        }
        if (!changed) break;
    }
    int best = 0;
    T bestn = -1;
#pragma unroll
    for (int j = 0; j < N; j++) {
        T sq = 0;
#pragma unroll
        for (int k = 0; k < M; k++) sq += A[k * N + j] * A[k * N + j];
        if (bestn < 0 || sq < bestn) {
            bestn = sq;
            best = j;
        }
    }
#pragma unroll
    for (int k = 0; k < 4; k++) {
        T v = 0;
#pragma unroll
        for (int j = 0; j < N; j++) v = (j == best) ? V[k * N + j] : v;
        z[k] = v;
    }
}

// 2-view DLT (cv::sfm::triangulatePoints with two views): 4x4 system rows x*P[2]-P[0], y*P[2]-P[1].
// Kept out of line with scalar arguments: inlining it next to the design-matrix path made cicc -O3 (12.9) read the caller's
// xs/ys/Pm arrays before they were written (observed on sm_100a: NaN for every 2-view row), see tools/dbg.
__device__ __noinline__ void dlt_2views(const double* P0, const double* P1, double x0, double y0, double x1, double y1, double* X3) {
    double A[16], z[4];
#pragma unroll
    for (int i = 0; i < 4; i++) {
        A[0 * 4 + i] = x0 * P0[8 + i] - P0[i];
        A[1 * 4 + i] = y0 * P0[8 + i] - P0[4 + i];
        A[2 * 4 + i] = x1 * P1[8 + i] - P1[i];
        A[3 * 4 + i] = y1 * P1[8 + i] - P1[4 + i];
    }
    null_vector_fixed<double, 4, 4, true>(A, z, 2.220446049250313e-16);
    X3[0] = z[0] / z[3];
    X3[1] = z[1] / z[3];
    X3[2] = z[2] / z[3];
}

// N-view DLT, (3n) x (4+n) design [-P_i | x_i in column 4+i] (MultiCameraFrame.cpp:419-433), n = NV >= 2.
template <int NV>
__device__ __noinline__ void dlt_design_fixed(const double* const* Pm, const double* xs, const double* ys, double* X3) {
    constexpr int M = 3 * NV, N = 4 + NV;
    double A[M * N], z[4];
#pragma unroll
    for (int i = 0; i < M * N; i++) A[i] = 0;
#pragma unroll
    for (int i = 0; i < NV; i++) {
#pragma unroll
        for (int jj = 0; jj < 3; jj++)
#pragma unroll
            for (int ii = 0; ii < 4; ii++) A[(3 * i + jj) * N + ii] = -Pm[i][jj * 4 + ii];
        A[(3 * i + 0) * N + 4 + i] = xs[i];
        A[(3 * i + 1) * N + 4 + i] = ys[i];
        A[(3 * i + 2) * N + 4 + i] = 1.0;
    }
    null_vector_fixed<double, M, N, (NV <= 3)>(A, z, 2.220446049250313e-16);
    X3[0] = z[0] / z[3];
    X3[1] = z[1] / z[3];
    X3[2] = z[2] / z[3];
}

template <int MAXC>
__device__ __forceinline__ void dlt_nviews(const double* const* Pm, const double* xs, const double* ys, int nv, bool force_design, double* X3) {
    if (nv == 2 && !force_design) {
        dlt_2views(Pm[0], Pm[1], xs[0], ys[0], xs[1], ys[1], X3);
        return;
    }
    switch (nv) {
        case 2: dlt_design_fixed<2>(Pm, xs, ys, X3); break;
        case 3: dlt_design_fixed<3>(Pm, xs, ys, X3); break;
        case 4: dlt_design_fixed<4>(Pm, xs, ys, X3); break;
        case 5: dlt_design_fixed<5>(Pm, xs, ys, X3); break;
        case 6: dlt_design_fixed<6>(Pm, xs, ys, X3); break;
        case 7: dlt_design_fixed<7>(Pm, xs, ys, X3); break;
        default: dlt_design_fixed<(MAXC < 8 ? MAXC : 8)>(Pm, xs, ys, X3); break;
    }
}

struct LandmarkOut {
    int* rows;          // [frames][row_cap][num_cams]
    uint8_t* accepted;  // [frames][row_cap]
    int* n_views;
    double* xyz;        // [frames][row_cap][3]
    float* uv_ref;      // [frames][row_cap][2]
    int* rep_cam;
    int* n_accepted;    // [frames]
};

// ------------------------------------------------------------------------------------------------------------
// K8  triangulation block of FrontEnd::obtainLfFeatures (FrontEnd.cpp:249-370), two kernels:
//   tri_classify_kernel : one thread per IntraMatch row — mask filter (:265-273), n_rays, defaults; rows with >= 2 views are
//                         appended to the work list of their view count, so that
//   tri_solve_kernel<NV>: runs warps whose rows all have NV views (no divergence between system sizes, registers and shared
//                         memory sized for exactly NV).  NV = 2 keeps the 4x4 system in registers; NV >= 3 keeps the
//                         (3NV) x (4+NV) design matrix and V in shared memory as [element][thread] (conflict-free), columns p, q
//                         of the current rotation in registers.  Same operation order as the oracle's one-sided Jacobi.
// ------------------------------------------------------------------------------------------------------------
#define MCS_TRI_NT 32

template <int MAXC>
__global__ void __launch_bounds__(128) tri_classify_kernel(int num_cams, int kp_cap, int row_cap, int width, int height,
                                                           const int* __restrict__ rows_in, const int* __restrict__ row_count,
                                                           const KeypointPod* __restrict__ kps, const float* __restrict__ masks, int mask_u8,
                                                           LandmarkOut out, int* __restrict__ lists, int* __restrict__ list_count,
                                                           int list_stride) {
    const int frame = blockIdx.y;
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    const int M = min(row_count[frame], row_cap);
    if (r >= M) return;
    const size_t ro = (size_t)frame * row_cap + r;
    const int* row = rows_in + ro * num_cams;
    int nv = 0, first = -1;
    for (int c = 0; c < num_cams; c++) {
        int f = row[c];
        if (f != -1) {
            bool keep = true;
            if (masks) {
                const KeypointPod p = kps[(size_t)(frame * num_cams + c) * kp_cap + f];
                int my = (int)p.y, mx = (int)p.x;  // Mat::at<float>(float, float): truncation
                my = min(max(my, 0), height - 1);
                mx = min(max(mx, 0), width - 1);
                const size_t at = ((size_t)(frame * num_cams + c) * height + my) * width + mx;
                // :266 on the f32 mask; the u8 form carries the caller's decision: 0 = static pixel (mask < 0.7), else dynamic
                keep = mask_u8 ? reinterpret_cast<const uint8_t*>(masks)[at] == 0 : (double)masks[at] < 0.7;
            }
            if (keep) {
                if (nv == 0) first = c;
                nv++;
            } else {
                f = -1;
            }
        }
        out.rows[ro * num_cams + c] = f;
    }
    out.n_views[ro] = nv;
    out.accepted[ro] = 0;
    out.rep_cam[ro] = nv == 1 ? first : -1;
    out.xyz[ro * 3] = 0;
    out.xyz[ro * 3 + 1] = 0;
    out.xyz[ro * 3 + 2] = 0;
    out.uv_ref[ro * 2] = 0;
    out.uv_ref[ro * 2 + 1] = 0;
    if (nv >= 2) lists[(size_t)nv * list_stride + atomicAdd(&list_count[nv], 1)] = (int)ro;
}

// null vector of the (3NV) x (4+NV) design in shared memory, element e of thread t at sm[e * MCS_TRI_NT + t]
template <int NV>
__device__ __forceinline__ void null_vector_smem(double* sA, double* sV, double* z) {
    constexpr int M = 3 * NV, N = 4 + NV, T = MCS_TRI_NT;
#ifdef MCS_TRI_ROT_REF
    const double eps = 2.220446049250313e-16;
#endif
    // only rows 0..3 of V are ever read (z = V[0..3][best]) and a rotation updates every row independently, so V is kept as 4 x N
#pragma unroll 1
    for (int i = 0; i < 4 * N; i++) sV[i * T] = 0;
#pragma unroll
    for (int i = 0; i < 4; i++) sV[(i * N + i) * T] = 1;
#pragma unroll 1
    for (int sweep = 0; sweep < 60; sweep++) {
        bool changed = false;
#pragma unroll 1
        for (int p = 0; p < N - 1; p++) {
#pragma unroll 1
            for (int q = p + 1; q < N; q++) {
                double xk[M], yk[M];
                double a = 0, b = 0, g = 0;
#pragma unroll
                for (int k = 0; k < M; k++) {
                    xk[k] = sA[(k * N + p) * T];
                    yk[k] = sA[(k * N + q) * T];
                }
                // summation order of tri_solve4_kernel (four lanes per row: matrix rows k = j, j + 4, ... summed per lane j, then
                // (s0 + s1) + (s2 + s3)), so that the two kernels give bit-identical points and a call's size does not show in them.
                // Both use explicit fma() in the dot products and the rotation update (the build has no contraction, --fmad=false,
                // for the sake of the reference's f32 expressions elsewhere): a third fewer f64 instructions in the chain.
                {
                    double pa[4] = {0, 0, 0, 0}, pb[4] = {0, 0, 0, 0}, pg[4] = {0, 0, 0, 0};
#pragma unroll
                    for (int k = 0; k < M; k++) {
                        pa[k & 3] = fma(xk[k], xk[k], pa[k & 3]);
                        pb[k & 3] = fma(yk[k], yk[k], pb[k & 3]);
                        pg[k & 3] = fma(xk[k], yk[k], pg[k & 3]);
                    }
                    a = (pa[0] + pa[1]) + (pa[2] + pa[3]);
                    b = (pb[0] + pb[1]) + (pb[2] + pb[3]);
                    g = (pg[0] + pg[1]) + (pg[2] + pg[3]);
                }
                // The rotation of null_vector() — skip when |g| <= eps sqrt(a b); zeta = (b - a) / 2g, t = sign(zeta) / (|zeta| +
                // sqrt(1 + zeta^2)), c = 1 / sqrt(1 + t^2) — in a form with a shorter dependent chain (this kernel is latency-bound and
                // three square roots and three divisions in a row were most of it): the fraction for t extended by |2g|, the test
                // squared, c by rsqrt — one square root, one division, one reciprocal square root.  The same rotation to a few ulp.
                // Stopping rule 8 eps instead of eps: at eps the test sits inside the rounding noise of g itself (15 products), so
                // now and then a row keeps "rotating" by less than an ulp until the 60-sweep cap — which row is a matter of
                // rounding luck (the literal form has such rows too), and one such row holds its warp, and with it a small launch,
                // for 10x the normal time (measured: one row of the bench's frame 6, triangulation stage of an 8-frame call
                // 0.18 -> 0.35 ms).  At 8 eps every row of the workload stops after 5-7 sweeps.  Points move by <= 4e-13 relative
                // against the oracle, which keeps the literal rule (tests: 1e-6; MCS_TRI_ROT_REF builds the literal form).
#ifdef MCS_TRI_ROT_REF
                if (fabs(g) <= eps * sqrt(a * b) || g == 0) continue;
                changed = true;
                const double zeta = (b - a) / (2 * g);
                const double t = (zeta >= 0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1 + zeta * zeta));
                const double c = 1.0 / sqrt(1 + t * t), sn = c * t;
#else
                constexpr double tol2 = (8 * 2.220446049250313e-16) * (8 * 2.220446049250313e-16);
                if (g * g <= tol2 * (a * b) || g == 0) continue;
                changed = true;
                const double d = b - a, h = 2 * g;
                const double tt = fabs(h) / (fabs(d) + sqrt(fma(d, d, h * h)));
                const double t = (d != 0 && (d < 0) != (h < 0)) ? -tt : tt;
                const double c = rsqrt(fma(t, t, 1.0)), sn = c * t;
#endif
#pragma unroll
                for (int k = 0; k < M; k++) {
                    sA[(k * N + p) * T] = fma(c, xk[k], -(sn * yk[k]));
                    sA[(k * N + q) * T] = fma(sn, xk[k], c * yk[k]);
                }
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    const double x = sV[(k * N + p) * T], y = sV[(k * N + q) * T];
                    sV[(k * N + p) * T] = fma(c, x, -(sn * y));
                    sV[(k * N + q) * T] = fma(sn, x, c * y);
                }
            }
        }
        if (!changed) break;
    }
    int best = 0;
    double bestn = -1;
#pragma unroll 1
    for (int j = 0; j < N; j++) {
        double ps[4] = {0, 0, 0, 0};
#pragma unroll
        for (int k = 0; k < M; k++) {
            const double v = sA[(k * N + j) * T];
            ps[k & 3] = fma(v, v, ps[k & 3]);
        }
        const double sq = (ps[0] + ps[1]) + (ps[2] + ps[3]);
        if (bestn < 0 || sq < bestn) {
            bestn = sq;
            best = j;
        }
    }
#pragma unroll
    for (int k = 0; k < 4; k++) z[k] = sV[(k * N + best) * T];
}

template <int NV, int NT = MCS_TRI_NT>
__device__ __forceinline__ void tri_solve_class(double* s_tri, int block, int nblocks, int num_cams, int kp_cap, int row_cap,
                                                const CalibDev* __restrict__ calib, const KeypointPod* __restrict__ kps,
                                                const uint8_t* __restrict__ desc, const LandmarkOut& out, const int* __restrict__ lists,
                                                const int* __restrict__ list_count, int list_stride) {
    const int count = list_count[NV];
    const int* list = lists + (size_t)NV * list_stride;
    for (int e = block * NT + threadIdx.x; e < count; e += nblocks * NT) {
        const size_t ro = (size_t)list[e];
        const int frame = (int)(ro / row_cap);
        int views[NV], feats[NV];
        {
            int n = 0;
            for (int c = 0; c < num_cams; c++) {
                const int f = out.rows[ro * num_cams + c];
                if (f != -1 && n < NV) {
                    views[n] = c;
                    feats[n] = f;
                    n++;
                }
            }
        }
        double xs[NV], ys[NV];
#pragma unroll
        for (int k = 0; k < NV; k++) {
            const int c = views[k];
            const KeypointPod p = kps[(size_t)(frame * num_cams + c) * kp_cap + feats[k]];
            xs[k] = ((double)p.x - calib[c].K[2]) / calib[c].K[0];  // :292-298
            ys[k] = ((double)p.y - calib[c].K[5]) / calib[c].K[4];
        }
        double X[3];
        if (NV == 2) {
            dlt_2views(calib[views[0]].Rt, calib[views[1]].Rt, xs[0], ys[0], xs[1], ys[1], X);
        } else {
            constexpr int M = 3 * NV, N = 4 + NV, T = MCS_TRI_NT;
            double* sA = s_tri + threadIdx.x;
            double* sV = sA + M * N * T;
#pragma unroll 1
            for (int i = 0; i < M * N; i++) sA[i * T] = 0;
#pragma unroll
            for (int i = 0; i < NV; i++) {
                const double* Pm = calib[views[i]].Rt;
#pragma unroll
                for (int jj = 0; jj < 3; jj++)
#pragma unroll
                    for (int ii = 0; ii < 4; ii++) sA[((3 * i + jj) * N + ii) * T] = -Pm[jj * 4 + ii];
                sA[((3 * i + 0) * N + 4 + i) * T] = xs[i];
                sA[((3 * i + 1) * N + 4 + i) * T] = ys[i];
                sA[((3 * i + 2) * N + 4 + i) * T] = 1.0;
            }
            double z[4];
            null_vector_smem<NV>(sA, sV, z);
            X[0] = z[0] / z[3];
            X[1] = z[1] / z[3];
            X[2] = z[2] / z[3];
        }
        // representative descriptor: least median Hamming distance, first wins (MultiCameraFrame.cpp:530-567)
        int bestMedian = 0x7FFFFFFF, bestIdx = 0;
#pragma unroll
        for (int a = 0; a < NV; a++) {
            int d[NV];
#pragma unroll
            for (int b = 0; b < NV; b++)
                d[b] = a == b ? 0
                              : hamming256(desc + ((size_t)(frame * num_cams + views[a]) * kp_cap + feats[a]) * 32,
                                           desc + ((size_t)(frame * num_cams + views[b]) * kp_cap + feats[b]) * 32);
#pragma unroll
            for (int i = 1; i < NV; i++) {  // insertion sort (values only: ties are indistinguishable)
                const int key = d[i];
                int j = i - 1;
                while (j >= 0 && d[j] > key) {
                    d[j + 1] = d[j];
                    j--;
                }
                d[j + 1] = key;
            }
            const int median = d[(int)(0.5 * (NV - 1))];
            if (median < bestMedian) {
                bestMedian = median;
                bestIdx = a;
            }
        }
        out.rep_cam[ro] = views[bestIdx];
        float u = 0, v = 0;
        if (X[2] < 40 && X[2] > 0.5) {  // :308
            out.accepted[ro] = 1;
            atomicAdd(&out.n_accepted[frame], 1);
            const double* K0 = calib[0].K;
            const double px = K0[0] * X[0] + K0[1] * X[1] + K0[2] * X[2];
            const double py = K0[3] * X[0] + K0[4] * X[1] + K0[5] * X[2];
            const double pz = K0[6] * X[0] + K0[7] * X[1] + K0[8] * X[2];
            u = (float)(px / pz);
            v = (float)(py / pz);
        }
        out.xyz[ro * 3] = X[0];
        out.xyz[ro * 3 + 1] = X[1];
        out.xyz[ro * 3 + 2] = X[2];
        out.uv_ref[ro * 2] = u;
        out.uv_ref[ro * 2 + 1] = v;
    }
}

// 2-view rows (the majority): 4x4 systems, no shared memory, wide blocks.  Launched on a second stream next to tri_solve_kernel.
__global__ void __launch_bounds__(128) tri_solve2_kernel(int num_cams, int kp_cap, int row_cap, const CalibDev* __restrict__ calib,
                                                         const KeypointPod* __restrict__ kps, const uint8_t* __restrict__ desc,
                                                         LandmarkOut out, const int* __restrict__ lists,
                                                         const int* __restrict__ list_count, int list_stride) {
    tri_solve_class<2, 128>(nullptr, blockIdx.x, gridDim.x, num_cams, kp_cap, row_cap, calib, kps, desc, out, lists, list_count, list_stride);
}

// One launch for every view count >= 3.  The solves are long, latency-bound and need a whole shared-memory slice each, so the grid is
// as many one-warp blocks as fit the SMs, and every block pulls tiles of 32 rows from one queue that runs over the view-count classes
// from the largest systems down (they run longest): no block idles while another class still has work.
__global__ void __launch_bounds__(MCS_TRI_NT) tri_solve_kernel(int num_cams, int kp_cap, int row_cap, const CalibDev* __restrict__ calib,
                                                               const KeypointPod* __restrict__ kps, const uint8_t* __restrict__ desc,
                                                               LandmarkOut out, const int* __restrict__ lists,
                                                               const int* __restrict__ list_count, int list_stride,
                                                               int* __restrict__ tile_counter) {
    extern __shared__ __align__(16) double s_tri[];
    __shared__ int s_tile;
    constexpr int ONE = 1 << 20;  // "grid size" that makes tri_solve_class handle exactly one tile
    for (;;) {
        __syncthreads();
        if (threadIdx.x == 0) s_tile = atomicAdd(tile_counter, 1);
        __syncthreads();
        int t = s_tile, nv = num_cams;
        for (; nv >= 3; nv--) {
            const int tiles = (list_count[nv] + MCS_TRI_NT - 1) / MCS_TRI_NT;
            if (t < tiles) break;
            t -= tiles;
        }
        if (nv < 3) return;
        switch (nv) {
            case 3: tri_solve_class<3>(s_tri, t, ONE, num_cams, kp_cap, row_cap, calib, kps, desc, out, lists, list_count, list_stride); break;
            case 4: tri_solve_class<4>(s_tri, t, ONE, num_cams, kp_cap, row_cap, calib, kps, desc, out, lists, list_count, list_stride); break;
            case 5: tri_solve_class<5>(s_tri, t, ONE, num_cams, kp_cap, row_cap, calib, kps, desc, out, lists, list_count, list_stride); break;
            case 6: tri_solve_class<6>(s_tri, t, ONE, num_cams, kp_cap, row_cap, calib, kps, desc, out, lists, list_count, list_stride); break;
            case 7: tri_solve_class<7>(s_tri, t, ONE, num_cams, kp_cap, row_cap, calib, kps, desc, out, lists, list_count, list_stride); break;
            default: tri_solve_class<8>(s_tri, t, ONE, num_cams, kp_cap, row_cap, calib, kps, desc, out, lists, list_count, list_stride); break;
        }
    }
}

// ---- four lanes per row ---------------------------------------------------------------------------------------------------
// The solve of one row is a chain of 6-7 sweeps x 21-36 dependent rotations, and with one row per thread every rotation walks the
// 3NV matrix rows in sequence (dot products, then the update).  Here four adjacent lanes share a row: lane j of the group owns matrix
// rows k = j, j + 4, ... (and row j of the 4 x N block of V), the three dot products are partial sums closed by two xor-shuffles (fp
// addition commutes, so the four lanes hold bit-identical sums and take identical decisions), the scalar part is computed
// redundantly.  A tile is 8 rows per warp; layout [column][k][row of the tile]: the 32 lanes of a load touch 32 consecutive doubles.
// Control flow is warp-uniform (a converged or empty row only re-tests its pairs: |g| stays under the threshold, nothing changes),
// so every shuffle runs with the full mask.  Pair order and stopping rule as in null_vector_smem; the summation order inside a dot
// product differs (four partial sums), i.e. points move in the last bits only.
#define MCS_TRI4_ROWS 8
template <int NV>
__device__ __forceinline__ void null_vector_smem4(double* sA, double* sV, int r, int j, double* z) {
    constexpr int M = 3 * NV, N = 4 + NV, MI = (M + 3) / 4, MP = MI * 4, R = MCS_TRI4_ROWS;
    constexpr double tol2 = (8 * 2.220446049250313e-16) * (8 * 2.220446049250313e-16);
    double* a0 = sA + j * R + r;  // element (col, 4 i + j) at a0[(col * MP + 4 i) * R]
    double* v0 = sV + j * R + r;  // V[j][col] at v0[col * 4 * R]
#pragma unroll 1
    for (int sweep = 0; sweep < 60; sweep++) {
        bool changed = false;
#pragma unroll 1
        for (int p = 0; p < N - 1; p++) {
#pragma unroll 1
            for (int q = p + 1; q < N; q++) {
                double xk[MI], yk[MI];
                double a = 0, b = 0, g = 0;
#pragma unroll
                for (int i = 0; i < MI; i++) {
                    xk[i] = a0[(p * MP + 4 * i) * R];
                    yk[i] = a0[(q * MP + 4 * i) * R];
                }
#pragma unroll
                for (int i = 0; i < MI; i++) {
                    a = fma(xk[i], xk[i], a);
                    b = fma(yk[i], yk[i], b);
                    g = fma(xk[i], yk[i], g);
                }
#pragma unroll
                for (int o = 1; o < 4; o <<= 1) {
                    a += __shfl_xor_sync(0xFFFFFFFFu, a, o);
                    b += __shfl_xor_sync(0xFFFFFFFFu, b, o);
                    g += __shfl_xor_sync(0xFFFFFFFFu, g, o);
                }
                if (g * g <= tol2 * (a * b) || g == 0) continue;
                changed = true;
                const double d = b - a, h = 2 * g;
                const double tt = fabs(h) / (fabs(d) + sqrt(fma(d, d, h * h)));
                const double t = (d != 0 && (d < 0) != (h < 0)) ? -tt : tt;
                const double c = rsqrt(fma(t, t, 1.0)), sn = c * t;
#pragma unroll
                for (int i = 0; i < MI; i++) {
                    a0[(p * MP + 4 * i) * R] = fma(c, xk[i], -(sn * yk[i]));
                    a0[(q * MP + 4 * i) * R] = fma(sn, xk[i], c * yk[i]);
                }
                const double x = v0[p * 4 * R], y = v0[q * 4 * R];
                v0[p * 4 * R] = fma(c, x, -(sn * y));
                v0[q * 4 * R] = fma(sn, x, c * y);
            }
        }
        if (!__any_sync(0xFFFFFFFFu, changed)) break;
    }
    int best = 0;
    double bestn = -1;
#pragma unroll 1
    for (int col = 0; col < N; col++) {
        double sq = 0;
#pragma unroll
        for (int i = 0; i < MI; i++) {
            const double v = a0[(col * MP + 4 * i) * R];
            sq = fma(v, v, sq);
        }
        sq += __shfl_xor_sync(0xFFFFFFFFu, sq, 1);
        sq += __shfl_xor_sync(0xFFFFFFFFu, sq, 2);
        if (bestn < 0 || sq < bestn) {
            bestn = sq;
            best = col;
        }
    }
    const double mine = v0[best * 4 * R];  // V[j][best]
    const int lane0 = (threadIdx.x & 31) & ~3;
#pragma unroll
    for (int k = 0; k < 4; k++) z[k] = __shfl_sync(0xFFFFFFFFu, mine, lane0 + k);
}

// one tile (8 rows) of view-count class NV; the whole warp calls it
template <int NV>
__device__ __forceinline__ void tri_solve_tile4(double* s_tri, int tile, int num_cams, int kp_cap, int row_cap, const CalibDev* __restrict__ calib,
                                                const KeypointPod* __restrict__ kps, const uint8_t* __restrict__ desc, const LandmarkOut& out,
                                                const int* __restrict__ lists, const int* __restrict__ list_count, int list_stride) {
    constexpr int M = 3 * NV, N = 4 + NV, MI = (M + 3) / 4, MP = MI * 4, R = MCS_TRI4_ROWS;
    const int lane = threadIdx.x & 31, r = lane >> 2, j = lane & 3;
    const int count = list_count[NV];
    const int e = tile * R + r;
    const bool valid = e < count;
    const size_t ro = valid ? (size_t)lists[(size_t)NV * list_stride + e] : 0;
    const int frame = (int)(ro / row_cap);
    int views[NV], feats[NV];
    double xs[NV], ys[NV];
#pragma unroll
    for (int k = 0; k < NV; k++) { views[k] = 0; feats[k] = 0; xs[k] = 0; ys[k] = 0; }
    if (valid) {
        int n = 0;
        for (int c = 0; c < num_cams; c++) {
            const int f = out.rows[ro * num_cams + c];
            if (f != -1 && n < NV) {
                views[n] = c;
                feats[n] = f;
                n++;
            }
        }
#pragma unroll
        for (int k = 0; k < NV; k++) {
            const int c = views[k];
            const KeypointPod p = kps[(size_t)(frame * num_cams + c) * kp_cap + feats[k]];
            xs[k] = ((double)p.x - calib[c].K[2]) / calib[c].K[0];  // :292-298
            ys[k] = ((double)p.y - calib[c].K[5]) / calib[c].K[4];
        }
    }
    double* sA = s_tri;
    double* sV = s_tri + N * MP * R;
    double* a0 = sA + j * R + r;
    double* v0 = sV + j * R + r;
#pragma unroll 1
    for (int i = 0; i < N * MI; i++) a0[i * 4 * R] = 0;  // own elements: (col, 4 i' + j) for all col, i'
#pragma unroll 1
    for (int col = 0; col < N; col++) v0[col * 4 * R] = col == j ? 1.0 : 0.0;
    __syncwarp();
    if (valid && j == 0) {
        double* sr = sA + r;  // element (col, k) of this row at sr[(col * MP + k) * R]
#pragma unroll
        for (int i = 0; i < NV; i++) {
            const double* Pm = calib[views[i]].Rt;
#pragma unroll
            for (int jj = 0; jj < 3; jj++)
#pragma unroll
                for (int ii = 0; ii < 4; ii++) sr[(ii * MP + 3 * i + jj) * R] = -Pm[jj * 4 + ii];
            sr[((4 + i) * MP + 3 * i + 0) * R] = xs[i];
            sr[((4 + i) * MP + 3 * i + 1) * R] = ys[i];
            sr[((4 + i) * MP + 3 * i + 2) * R] = 1.0;
        }
    }
    __syncwarp();
    double z[4];
    null_vector_smem4<NV>(sA, sV, r, j, z);
    __syncwarp();
    if (!valid || j != 0) return;
    double X[3];
    X[0] = z[0] / z[3];
    X[1] = z[1] / z[3];
    X[2] = z[2] / z[3];
    // representative descriptor: least median Hamming distance, first wins (MultiCameraFrame.cpp:530-567)
    int bestMedian = 0x7FFFFFFF, bestIdx = 0;
#pragma unroll
    for (int a = 0; a < NV; a++) {
        int dd[NV];
#pragma unroll
        for (int b = 0; b < NV; b++)
            dd[b] = a == b ? 0
                           : hamming256(desc + ((size_t)(frame * num_cams + views[a]) * kp_cap + feats[a]) * 32,
                                        desc + ((size_t)(frame * num_cams + views[b]) * kp_cap + feats[b]) * 32);
#pragma unroll
        for (int i = 1; i < NV; i++) {  // insertion sort (values only: ties are indistinguishable)
            const int key = dd[i];
            int jx = i - 1;
            while (jx >= 0 && dd[jx] > key) {
                dd[jx + 1] = dd[jx];
                jx--;
            }
            dd[jx + 1] = key;
        }
        const int median = dd[(int)(0.5 * (NV - 1))];
        if (median < bestMedian) {
            bestMedian = median;
            bestIdx = a;
        }
    }
    out.rep_cam[ro] = views[bestIdx];
    float u = 0, v = 0;
    if (X[2] < 40 && X[2] > 0.5) {  // :308
        out.accepted[ro] = 1;
        atomicAdd(&out.n_accepted[frame], 1);
        const double* K0 = calib[0].K;
        const double px = K0[0] * X[0] + K0[1] * X[1] + K0[2] * X[2];
        const double py = K0[3] * X[0] + K0[4] * X[1] + K0[5] * X[2];
        const double pz = K0[6] * X[0] + K0[7] * X[1] + K0[8] * X[2];
        u = (float)(px / pz);
        v = (float)(py / pz);
    }
    out.xyz[ro * 3] = X[0];
    out.xyz[ro * 3 + 1] = X[1];
    out.xyz[ro * 3 + 2] = X[2];
    out.uv_ref[ro * 2] = u;
    out.uv_ref[ro * 2 + 1] = v;
}

// bytes of shared memory of one tri_solve4_kernel block for rigs of up to `nv` cameras
__host__ __device__ inline size_t tri4_smem_bytes(int nv) {
    const int M = 3 * nv, N = 4 + nv, MP = ((M + 3) / 4) * 4;
    return nv <= 2 ? 0 : (size_t)(N * MP + 4 * N) * MCS_TRI4_ROWS * sizeof(double);
}

__global__ void __launch_bounds__(32) tri_solve4_kernel(int num_cams, int kp_cap, int row_cap, const CalibDev* __restrict__ calib,
                                                        const KeypointPod* __restrict__ kps, const uint8_t* __restrict__ desc,
                                                        LandmarkOut out, const int* __restrict__ lists,
                                                        const int* __restrict__ list_count, int list_stride,
                                                        int* __restrict__ tile_counter) {
    extern __shared__ __align__(16) double s_tri[];
    for (;;) {
        int t = 0;
        if ((threadIdx.x & 31) == 0) t = atomicAdd(tile_counter, 1);
        t = __shfl_sync(0xFFFFFFFFu, t, 0);
        int nv = num_cams;
        for (; nv >= 3; nv--) {
            const int tiles = (list_count[nv] + MCS_TRI4_ROWS - 1) / MCS_TRI4_ROWS;
            if (t < tiles) break;
            t -= tiles;
        }
        if (nv < 3) return;
        switch (nv) {
            case 3: tri_solve_tile4<3>(s_tri, t, num_cams, kp_cap, row_cap, calib, kps, desc, out, lists, list_count, list_stride); break;
            case 4: tri_solve_tile4<4>(s_tri, t, num_cams, kp_cap, row_cap, calib, kps, desc, out, lists, list_count, list_stride); break;
            case 5: tri_solve_tile4<5>(s_tri, t, num_cams, kp_cap, row_cap, calib, kps, desc, out, lists, list_count, list_stride); break;
            case 6: tri_solve_tile4<6>(s_tri, t, num_cams, kp_cap, row_cap, calib, kps, desc, out, lists, list_count, list_stride); break;
            case 7: tri_solve_tile4<7>(s_tri, t, num_cams, kp_cap, row_cap, calib, kps, desc, out, lists, list_count, list_stride); break;
            default: tri_solve_tile4<8>(s_tri, t, num_cams, kp_cap, row_cap, calib, kps, desc, out, lists, list_count, list_stride); break;
        }
        __syncwarp();
    }
}

// MultiCameraFrame::triangulateIntraMatches(x, view_inds, pt) (:411-441): always the (3n) x (4+n) design, K[R|t], pixels.
template <int MAXC>
__global__ void triangulate_nview_kernel(const CalibDev* __restrict__ calib, const int* __restrict__ view_inds, int nv,
                                         const double* __restrict__ xs, const double* __restrict__ ys, int npts, double* __restrict__ xyz) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= npts) return;
    const double* Pm[MAXC];
    double x[MAXC], y[MAXC];
    for (int k = 0; k < nv; k++) {
        Pm[k] = calib[view_inds[k]].P;
        x[k] = xs[(size_t)i * nv + k];
        y[k] = ys[(size_t)i * nv + k];
    }
    double X[3];
    dlt_nviews<MAXC>(Pm, x, y, nv, true, X);
    xyz[i * 3] = X[0];
    xyz[i * 3 + 1] = X[1];
    xyz[i * 3 + 2] = X[2];
}

// MultiCameraFrame::triangulate(kps_1, kps_2, pts3D, cam1, cam2) (:481-507): 4x4 f32 DLT.
__global__ void triangulate_pair_f32_kernel(const CalibDev* __restrict__ calib, int cam1, int cam2, const float* __restrict__ p1,
                                            const float* __restrict__ p2, int npts, float* __restrict__ xyz) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= npts) return;
    const double* P1 = calib[cam1].P;
    const double* P2 = calib[cam2].P;
    float A[16], V[16], z[4];
    for (int c = 0; c < 4; c++) {
        A[0 * 4 + c] = (float)((double)p1[i * 2] * P1[8 + c] - P1[c]);
        A[1 * 4 + c] = (float)((double)p1[i * 2 + 1] * P1[8 + c] - P1[4 + c]);
        A[2 * 4 + c] = (float)((double)p2[i * 2] * P2[8 + c] - P2[c]);
        A[3 * 4 + c] = (float)((double)p2[i * 2 + 1] * P2[8 + c] - P2[4 + c]);
    }
    null_vector<float, 4, 4>(A, V, 4, 4, z, 1.1920929e-07f);
    xyz[i * 3] = z[0] / z[3];
    xyz[i * 3 + 1] = z[1] / z[3];
    xyz[i * 3 + 2] = z[2] / z[3];
}

}  // namespace mcs
