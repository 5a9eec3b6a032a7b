// extract_kernels.cuh — sm_100a kernels for ORBextractor::operator() (/root/reference/MCSlam/src/ORBextractor.cpp:1085-1171).
//   K1 resize_level_kernel   : one pyramid level from the previous one (ComputePyramid :1173-1198, cv::resize INTER_LINEAR)
//   K2 blur_kernel           : 7x7 sigma-2 fixed-point Gaussian of every level (:1132-1133)
//   K3 fast_cells_kernel     : per-cell FAST-9/16 + score + NMS + iniTh/minTh fallback (:804-871)
//   K4 octree_kernel         : DistributeOctTree (:554-778), one block per (image, level)  -> octree_core.h
//   K5 describe_kernel       : post-octree fix-up (:879-889), orientation (:470-477 / IC_Angle :75-102 behind a flag),
//                              BRIEF-256 (:106-145), pt *= scale (:1149-1151), final level-major order (:1122-1163)
// All integer arithmetic is bit-exact with the OpenCV primitives the reference calls (checked by tests/ against the CPU restatement and cv2 golden vectors).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "octree_core.h"
#include "plan.h"

namespace mcs {

struct ImageSet {           // where the pixels of image `img`, level `l` live
    const uint8_t* l0;      // level 0 (may alias the caller's buffer)
    int l0_pitch;
    size_t l0_stride;       // bytes between images
    uint8_t* pyr;           // levels >= 1 (and level 0 when staged): pyr + img * pyr_stride + plane_ofs
};

__device__ __forceinline__ const uint8_t* level_ptr(const McsPlan& P, const ImageSet& S, int img, int l, int& pitch) {
    if (l == 0) {
        pitch = S.l0_pitch;
        return S.l0 + (size_t)img * S.l0_stride;
    }
    pitch = P.L[l].pitch;
    return S.pyr + (size_t)img * P.pyr_stride + P.L[l].plane_ofs;
}

__device__ __forceinline__ int reflect101(int p, int len) {
    if (p < 0) p = -p;
    if (p >= len) p = 2 * len - 2 - p;
    return min(max(p, 0), len - 1);  // exact for |overshoot| < len; far-out tile padding (never used) is clamped
}

// ------------------------------------------------------------------------------------------------------------
// K1  cv::resize(prev, cur, INTER_LINEAR) on 8U: 11-bit fixed-point coefficients, two-stage rounding.
// xtab[dx] = {sx, a0 | a1 << 16}, ytab[dy] = {sy, b0 | b1 << 16} are built on the host exactly as OpenCV does.
// One thread produces 4 horizontally adjacent pixels (one 32-bit store).
// ------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) resize_level_kernel(const __grid_constant__ McsPlan P, ImageSet S, int level,
                                                           const int2* __restrict__ xtab, const int2* __restrict__ ytab) {
    const McsLevel& D = P.L[level];
    const McsLevel& Sv = P.L[level - 1];
    const int img = blockIdx.z;
    const int dy = blockIdx.y * blockDim.y + threadIdx.y;
    const int dx0 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (dy >= D.h || dx0 >= D.w) return;
    int spitch;
    const uint8_t* src = level_ptr(P, S, img, level - 1, spitch);
    uint8_t* dst = S.pyr + (size_t)img * P.pyr_stride + D.plane_ofs;
    const int2 yt = __ldg(&ytab[D.ytab_ofs + dy]);
    const int sh = Sv.h, sw = Sv.w;
    int sy0 = yt.x, sy1 = yt.x + 1;
    sy0 = sy0 < 0 ? 0 : (sy0 < sh ? sy0 : sh - 1);
    sy1 = sy1 < 0 ? 0 : (sy1 < sh ? sy1 : sh - 1);
    const int b0 = (short)(yt.y & 0xFFFF), b1 = yt.y >> 16;
    const uint8_t* r0 = src + (size_t)sy0 * spitch;
    const uint8_t* r1 = src + (size_t)sy1 * spitch;
    uint32_t outw = 0;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int dx = dx0 + k;
        if (dx < D.w) {
            const int2 xt = __ldg(&xtab[D.xtab_ofs + dx]);
            const int sx = xt.x, sx1 = min(sx + 1, sw - 1);
            const int a0 = (short)(xt.y & 0xFFFF), a1 = xt.y >> 16;
            const int h0 = r0[sx] * a0 + r0[sx1] * a1;
            const int h1 = r1[sx] * a0 + r1[sx1] * a1;
            int v = (((b0 * (h0 >> 4)) >> 16) + ((b1 * (h1 >> 4)) >> 16) + 2) >> 2;
            v = min(max(v, 0), 255);
            outw |= (uint32_t)v << (8 * k);
        }
    }
    *reinterpret_cast<uint32_t*>(dst + (size_t)dy * D.pitch + dx0) = outw;  // pitch % 64 == 0: row padding absorbs the tail
}

// ------------------------------------------------------------------------------------------------------------
// K2  cv::GaussianBlur(7x7, sigma 2, BORDER_REFLECT_101) on 8U: 8.8 fixed-point kernel {18,34,48,56,48,34,18}/256,
// exact integer accumulation, one rounding (v + 32768) >> 16.  Tile 64x16 + 3 px halo staged in shared memory,
// horizontal pass into a u16 shared tile, vertical pass from it.  All levels of all images in one launch.
// ------------------------------------------------------------------------------------------------------------
__constant__ int c_gauss7[7];  // set by the host from the bit-exact kernel generator

__global__ void __launch_bounds__(256) blur_kernel(const __grid_constant__ McsPlan P, ImageSet S, uint8_t* __restrict__ blurred) {
    constexpr int TW = MCS_BLUR_TW, TH = MCS_BLUR_TH, SW = TW + 6, SH = TH + 6;
    __shared__ uint8_t s_in[SH][SW + 2];
    __shared__ uint16_t s_h[SH][TW];
    const int img = blockIdx.y;
    int level = 0;
#pragma unroll 1
    for (int l = 1; l < P.nlevels; l++)
        if ((int)blockIdx.x >= P.L[l].tile_base) level = l;
    const McsLevel& Lv = P.L[level];
    const int t = blockIdx.x - Lv.tile_base;
    const int tx = (t % Lv.tiles_x) * TW, ty = (t / Lv.tiles_x) * TH;
    int pitch;
    const uint8_t* src = level_ptr(P, S, img, level, pitch);
    const int w = Lv.w, h = Lv.h;
    for (int i = threadIdx.x; i < SW * SH; i += blockDim.x) {
        const int r = i / SW, c = i - r * SW;
        const int gy = reflect101(ty + r - 3, h), gx = reflect101(tx + c - 3, w);
        s_in[r][c] = src[(size_t)gy * pitch + gx];
    }
    __syncthreads();
    const int k0 = c_gauss7[0], k1 = c_gauss7[1], k2 = c_gauss7[2], k3 = c_gauss7[3];
    for (int i = threadIdx.x; i < TW * SH; i += blockDim.x) {
        const int r = i / TW, c = i - r * TW;
        const uint8_t* p = &s_in[r][c];
        s_h[r][c] = (uint16_t)(k0 * (p[0] + p[6]) + k1 * (p[1] + p[5]) + k2 * (p[2] + p[4]) + k3 * p[3]);
    }
    __syncthreads();
    uint8_t* dst = blurred + (size_t)img * P.pyr_stride + Lv.plane_ofs;
    // 4 pixels per thread -> one 32-bit store; TW*TH/4 = 256 threads
    for (int i = threadIdx.x; i < (TW / 4) * TH; i += blockDim.x) {
        const int r = i / (TW / 4), c4 = (i - r * (TW / 4)) * 4;
        const int gy = ty + r, gx = tx + c4;
        if (gy >= h || gx >= w) continue;
        uint32_t outw = 0;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const int c = c4 + k;
            const uint32_t a = (uint32_t)k0 * (s_h[r][c] + s_h[r + 6][c]) + (uint32_t)k1 * (s_h[r + 1][c] + s_h[r + 5][c]) +
                               (uint32_t)k2 * (s_h[r + 2][c] + s_h[r + 4][c]) + (uint32_t)k3 * s_h[r + 3][c];
            outw |= min((a + 32768u) >> 16, 255u) << (8 * k);
        }
        *reinterpret_cast<uint32_t*>(dst + (size_t)gy * Lv.pitch + gx) = outw;  // pitch % 64 == 0 and gx % 4 == 0
    }
}

// ------------------------------------------------------------------------------------------------------------
// K3  per-cell FAST.  One block = one 35-ish px cell of ComputeKeyPointsOctTree's grid = exactly the sub-image the
// reference hands to cv::FAST (rowRange(iniY,maxY).colRange(iniX,maxX), :825), so the 3 px non-computed margin, the
// "scores outside the interior are 0" NMS rule and the per-cell threshold fallback (:842-845) hold by construction.
//
// Threshold-free formulation: strength s(p) = max over the 16 nine-pixel arcs of min(|v - ring|) (darker or brighter
// arc) - 1.  p is a FAST corner at threshold t  <=>  s(p) >= t, and cv::FAST's score is s(p) for every t.  With
// strict-'>' NMS a pixel survives at t  <=>  s(p) >= t  and  s(p) > s(q) for its 8 in-interior neighbours q (neighbours
// below t score 0 < s(p)).  So one pass computes s, one pass the local maxima; the cell keeps maxima >= iniTh if it has
// any, else maxima >= minTh.
//
// The arc minima use packed 16-bit lanes (d + 256 keeps both lanes positive) and 3-input min/max (VIMNMX3).
// ------------------------------------------------------------------------------------------------------------
#define MCS_FAST_NT 128
#define MCS_FAST_TP 84   // shared tile pitch (bytes): <= 75 px + 3 alignment + slack, multiple of 4
#define MCS_FAST_MAXC 76 // max sub-image side
#define MCS_FAST_SP 72   // score tile pitch
#define MCS_FAST_MAXSURV 1296

__device__ __forceinline__ int fast_strength(const uint8_t* c, int tp) {
    // ring offsets of FAST 9/16 in OpenCV order: (0,3)(1,3)(2,2)(3,1)(3,0)(3,-1)(2,-2)(1,-3)(0,-3)(-1,-3)(-2,-2)(-3,-1)(-3,0)(-3,1)(-2,2)(-1,3)
    const int v = c[0];
    const uint32_t vb = (uint32_t)(v + 256) * 0x00010001u;
    uint32_t W0[8], W1[8];
    W0[0] = vb - ((uint32_t)c[3 * tp] | ((uint32_t)c[3 * tp + 1] << 16));
    W0[1] = vb - ((uint32_t)c[2 * tp + 2] | ((uint32_t)c[tp + 3] << 16));
    W0[2] = vb - ((uint32_t)c[3] | ((uint32_t)c[-tp + 3] << 16));
    W0[3] = vb - ((uint32_t)c[-2 * tp + 2] | ((uint32_t)c[-3 * tp + 1] << 16));
    W0[4] = vb - ((uint32_t)c[-3 * tp] | ((uint32_t)c[-3 * tp - 1] << 16));
    W0[5] = vb - ((uint32_t)c[-2 * tp - 2] | ((uint32_t)c[-tp - 3] << 16));
    W0[6] = vb - ((uint32_t)c[-3] | ((uint32_t)c[tp - 3] << 16));
    W0[7] = vb - ((uint32_t)c[2 * tp - 2] | ((uint32_t)c[3 * tp - 1] << 16));
#pragma unroll
    for (int j = 0; j < 8; j++) W1[j] = __byte_perm(W0[j], W0[(j + 1) & 7], 0x5432);
    uint32_t M0[8], X0[8];
#pragma unroll
    for (int j = 0; j < 8; j++) {
        M0[j] = __vimin3_u16x2(W0[j], W1[j], W0[(j + 1) & 7]);
        X0[j] = __vimax3_u16x2(W0[j], W1[j], W0[(j + 1) & 7]);
    }
    uint32_t amin = 0, amax = 0xFFFFFFFFu;  // max over arcs of arc-min ; min over arcs of arc-max (per 16-bit lane)
#pragma unroll
    for (int j = 0; j < 8; j++) {
        const uint32_t M1 = __byte_perm(M0[(j + 1) & 7], M0[(j + 2) & 7], 0x5432);
        const uint32_t X1 = __byte_perm(X0[(j + 1) & 7], X0[(j + 2) & 7], 0x5432);
        amin = __vmaxu2(amin, __vimin3_u16x2(M0[j], M1, M0[(j + 3) & 7]));
        amax = __vminu2(amax, __vimax3_u16x2(X0[j], X1, X0[(j + 3) & 7]));
    }
    const int dark = (int)max(amin & 0xFFFFu, amin >> 16) - 256;     // max arc-min of  (v - x)
    const int bright = 256 - (int)min(amax & 0xFFFFu, amax >> 16);   // max arc-min of -(v - x)
    return max(dark, bright) - 1;
}

__global__ void __launch_bounds__(MCS_FAST_NT) fast_cells_kernel(const __grid_constant__ McsPlan P, ImageSet S,
                                                                 uint32_t* __restrict__ cand, int* __restrict__ cand_count) {
    __shared__ __align__(16) uint8_t s_tile[MCS_FAST_MAXC * MCS_FAST_TP];
    __shared__ uint8_t s_score[(MCS_FAST_MAXC - 4) * MCS_FAST_SP];
    __shared__ uint32_t s_surv[MCS_FAST_MAXSURV];
    __shared__ int s_nsurv, s_cellmax;
    const int img = blockIdx.y;
    int level = 0;
#pragma unroll 1
    for (int l = 1; l < P.nlevels; l++)
        if ((int)blockIdx.x >= P.L[l].cell_base) level = l;
    const McsLevel& Lv = P.L[level];
    const int ci = blockIdx.x - Lv.cell_base;
    const int i = ci / Lv.n_cols, j = ci - i * Lv.n_cols;
    const int maxBX = Lv.w - MCS_MIN_BORDER, maxBY = Lv.h - MCS_MIN_BORDER;
    const int iniX = MCS_MIN_BORDER + j * Lv.w_cell, iniY = MCS_MIN_BORDER + i * Lv.h_cell;
    if (iniY >= maxBY - 3 || iniX >= maxBX - 6) return;  // :809, :818
    const int maxX = min(iniX + Lv.w_cell + 6, maxBX), maxY = min(iniY + Lv.h_cell + 6, maxBY);
    const int cw = maxX - iniX, ch = maxY - iniY;  // sub-image handed to cv::FAST
    const int iw = cw - 6, ih = ch - 6;            // computed interior
    if (threadIdx.x == 0) { s_nsurv = 0; s_cellmax = 0; }
    if (iw <= 0 || ih <= 0) return;
    int pitch;
    const uint8_t* src = level_ptr(P, S, img, level, pitch);
    // stage the sub-image with aligned 32-bit loads
    const int a0 = iniX & ~3, off = iniX - a0, nw = (off + cw + 3) >> 2;
    for (int t = threadIdx.x; t < nw * ch; t += MCS_FAST_NT) {
        const int r = t / nw, c = t - r * nw;
        const uint32_t wv = __ldg(reinterpret_cast<const uint32_t*>(src + (size_t)(iniY + r) * pitch + a0) + c);
        *reinterpret_cast<uint32_t*>(&s_tile[r * MCS_FAST_TP + c * 4]) = wv;
    }
    // zero score tile incl. 1 px frame: (iw + 2) x (ih + 2)
    for (int t = threadIdx.x; t < (ih + 2) * MCS_FAST_SP; t += MCS_FAST_NT) s_score[t] = 0;
    __syncthreads();
    const int th = min(P.min_th, P.ini_th);  // weakest threshold any of the two FAST calls can use
    for (int t = threadIdx.x; t < iw * ih; t += MCS_FAST_NT) {
        const int ly = t / iw, lx = t - ly * iw;
        const uint8_t* c = &s_tile[(ly + 3) * MCS_FAST_TP + lx + 3 + off];
        const int v = c[0];
        // necessary condition for a 9-arc: of each opposite ring pair at least one pixel is on the arc
        const int p0 = c[3 * MCS_FAST_TP], p8 = c[-3 * MCS_FAST_TP], p4 = c[3], p12 = c[-3];
        const int hi = v + th, lo = v - th;
        const bool brightOK = (p0 > hi || p8 > hi) && (p4 > hi || p12 > hi);
        const bool darkOK = (p0 < lo || p8 < lo) && (p4 < lo || p12 < lo);
        if (brightOK || darkOK) {
            const int s = fast_strength(c, MCS_FAST_TP);
            if (s >= th && s > 0) s_score[(ly + 1) * MCS_FAST_SP + lx + 1] = (uint8_t)min(s, 255);
        }
    }
    __syncthreads();
    for (int t = threadIdx.x; t < iw * ih; t += MCS_FAST_NT) {
        const int ly = t / iw, lx = t - ly * iw;
        const uint8_t* q = &s_score[(ly + 1) * MCS_FAST_SP + lx + 1];
        const int s = q[0];
        if (s == 0) continue;
        if (s > q[-1] && s > q[1] && s > q[-MCS_FAST_SP - 1] && s > q[-MCS_FAST_SP] && s > q[-MCS_FAST_SP + 1] &&
            s > q[MCS_FAST_SP - 1] && s > q[MCS_FAST_SP] && s > q[MCS_FAST_SP + 1]) {
            const int slot = atomicAdd(&s_nsurv, 1);
            // keypoint in ROI coordinates: FAST x in the sub-image (lx + 3) shifted by j*wCell (:864-865)
            if (slot < MCS_FAST_MAXSURV) s_surv[slot] = mcs_oct::pack_cand(lx + 3 + j * Lv.w_cell, ly + 3 + i * Lv.h_cell, s);
            atomicMax(&s_cellmax, s);
        }
    }
    __syncthreads();
    const int ns = min(s_nsurv, MCS_FAST_MAXSURV);
    const int keep_th = s_cellmax >= P.ini_th ? P.ini_th : P.min_th;  // FAST(iniTh) non-empty ? keep it : FAST(minTh)
    uint32_t* out = cand + (size_t)img * P.cand_stride + Lv.cand_ofs;
    int* cnt = cand_count + img * P.nlevels + level;
    for (int base = 0; base < ns; base += MCS_FAST_NT) {
        const int t = base + threadIdx.x;
        const uint32_t rec = t < ns ? s_surv[t] : 0;
        const bool keep = t < ns && mcs_oct::cand_score(rec) >= keep_th;
        const unsigned m = __ballot_sync(0xFFFFFFFFu, keep);
        if (m) {
            const int lane = threadIdx.x & 31;
            int pos = 0;
            if (lane == 0) pos = atomicAdd(cnt, __popc(m));
            pos = __shfl_sync(0xFFFFFFFFu, pos, 0) + __popc(m & ((1u << lane) - 1));
            if (keep && pos < Lv.cand_cap) out[pos] = rec;
        }
    }
}

// ------------------------------------------------------------------------------------------------------------
// K4  DistributeOctTree.  One block per (level, image); algorithm in octree_core.h, executed with a CUDA block context.
// ------------------------------------------------------------------------------------------------------------
#define MCS_OCT_NT 256

struct CudaBlockCtx {
    __device__ __forceinline__ int nthreads() const { return blockDim.x; }
    __device__ __forceinline__ int tid() const { return threadIdx.x; }
    template <typename F>
    __device__ __forceinline__ void par(F f) {
        f();
        __syncthreads();
    }
    template <typename F>
    __device__ __forceinline__ void serial(F f) {
        if (threadIdx.x == 0) f();
        __syncthreads();
    }
    __device__ __forceinline__ void atomic_add(int* p, int v) { atomicAdd(p, v); }
    __device__ __forceinline__ void atomic_max(unsigned* p, unsigned v) { atomicMax(p, v); }
    __device__ __forceinline__ void atomic_min(unsigned* p, unsigned v) { atomicMin(p, v); }
    // in-place exclusive scan of a[0..n) in shared memory; collective; returns the total
    __device__ int exclusive_scan(int* a, int n) {
        __shared__ int s_ws[34];
        const int NT = blockDim.x, t = threadIdx.x, lane = t & 31, w = t >> 5, nwarps = NT >> 5;
        const int per = (n + NT - 1) / NT;
        const int beg = min(t * per, n), end = min(beg + per, n);
        int sum = 0;
        for (int i = beg; i < end; i++) sum += a[i];
        int incl = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xFFFFFFFFu, incl, o);
            if (lane >= o) incl += v;
        }
        if (lane == 31) s_ws[w] = incl;
        __syncthreads();
        if (w == 0) {
            const int ws = lane < nwarps ? s_ws[lane] : 0;
            int wi = ws;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int v = __shfl_up_sync(0xFFFFFFFFu, wi, o);
                if (lane >= o) wi += v;
            }
            s_ws[lane] = wi - ws;
            if (lane == 31) s_ws[32] = wi;
        }
        __syncthreads();
        int base = s_ws[w] + incl - sum;
        for (int i = beg; i < end; i++) {
            const int v = a[i];
            a[i] = base;
            base += v;
        }
        const int total = s_ws[32];
        __syncthreads();
        return total;
    }
};

// sel[img][sel_ofs(level) + k] = k-th retained candidate of the level in final list order; sel_count[img][level] = how many.
__global__ void __launch_bounds__(MCS_OCT_NT) octree_kernel(const __grid_constant__ McsPlan P, const uint32_t* __restrict__ cand,
                                                            const int* __restrict__ cand_count, uint16_t* __restrict__ knode,
                                                            uint32_t* __restrict__ sel, int* __restrict__ sel_count) {
    extern __shared__ __align__(16) unsigned char s_oct[];
    const int level = blockIdx.x, img = blockIdx.y;
    const McsLevel& Lv = P.L[level];
    int n_keys = cand_count[img * P.nlevels + level];
    n_keys = min(n_keys, Lv.cand_cap);
    int n_out = 0;
    if (n_keys > 0 && Lv.n_cols > 0) {
        mcs_oct::Params op;
        op.roi_w = Lv.roi_w;
        op.roi_h = Lv.roi_h;
        op.N = Lv.quota;
        op.n_ini = Lv.n_ini;
        op.hX = Lv.hX;
        op.w_cell = Lv.w_cell;
        op.h_cell = Lv.h_cell;
        op.n_cols = Lv.n_cols;
        op.cap = Lv.oct_cap;
        mcs_oct::Shared sh;
        sh.carve(s_oct, op.cap);
        CudaBlockCtx ctx;
        const size_t kofs = (size_t)img * P.cand_stride + Lv.cand_ofs;
        n_out = mcs_oct::distribute(ctx, op, sh, cand + kofs, n_keys, knode + kofs, sel + (size_t)img * P.sel_stride + Lv.sel_ofs);
    }
    if (threadIdx.x == 0) sel_count[img * P.nlevels + level] = n_out;  // -1 = node table overflow (reported by the host)
}

// ------------------------------------------------------------------------------------------------------------
// K5  keypoint fix-up + orientation + BRIEF-256 + final ordering.  One warp per keypoint.
// Output slot = (sum of earlier levels' counts) + position in the level's octree order (operator() :1122-1163 with
// vLappingArea = {0,0}: every keypoint is "mono", written front to back).  A non-trivial lapping area is applied by
// lapping_permute_kernel afterwards.
// ------------------------------------------------------------------------------------------------------------
__constant__ signed char c_pattern[1024];  // 512 (x, y) pairs, see include/mcs_orb_pattern.inc

struct KeypointPod {
    float x, y, size, angle, response;
    int octave, class_id;
};

// cv::fastAtan2 (OpenCV core mathfuncs_core: atan_f32), degrees
__device__ __forceinline__ float fast_atan2_deg(float y, float x) {
    const float sc = (float)(180.0 / 3.14159265358979323846);
    const float p1 = 0.9997878412794807f * sc, p3 = -0.3258083974640975f * sc, p5 = 0.1555786518463281f * sc,
                p7 = -0.04432655554792128f * sc;
    const float ax = fabsf(x), ay = fabsf(y);
    float a, c, c2;
    if (ax >= ay) {
        c = __fdiv_rn(ay, __fadd_rn(ax, 2.2204460492503131e-16f));
        c2 = __fmul_rn(c, c);
        a = __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c);
    } else {
        c = __fdiv_rn(ax, __fadd_rn(ay, 2.2204460492503131e-16f));
        c2 = __fmul_rn(c, c);
        a = __fsub_rn(90.f, __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c));
    }
    if (x < 0) a = __fsub_rn(180.f, a);
    if (y < 0) a = __fsub_rn(360.f, a);
    return a;
}

#define MCS_DESC_WARPS 8
__global__ void __launch_bounds__(MCS_DESC_WARPS * 32) describe_kernel(const __grid_constant__ McsPlan P, ImageSet S,
                                                                       const uint8_t* __restrict__ blurred,
                                                                       const uint32_t* __restrict__ sel,
                                                                       const int* __restrict__ sel_count, KeypointPod* __restrict__ kps,
                                                                       uint8_t* __restrict__ desc, int* __restrict__ kp_count) {
    const int img = blockIdx.y;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int k = blockIdx.x * MCS_DESC_WARPS + warp;  // output slot
    // locate the level: prefix over the level counts (negative = overflow flag, treated as 0 here)
    int level = -1, before = 0, total = 0;
#pragma unroll 1
    for (int l = 0; l < P.nlevels; l++) {
        const int c = max(sel_count[img * P.nlevels + l], 0);
        if (level < 0 && k < total + c) {
            level = l;
            before = total;
        }
        total += c;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) kp_count[img] = min(total, P.kp_cap);
    if (level < 0 || k >= P.kp_cap) return;
    const McsLevel& Lv = P.L[level];
    const uint32_t rec = sel[(size_t)img * P.sel_stride + Lv.sel_ofs + (k - before)];
    const int px = mcs_oct::cand_x(rec) + MCS_MIN_BORDER, py = mcs_oct::cand_y(rec) + MCS_MIN_BORDER;  // :883-884
    float angle = 0.0f;
    int rx0, ry0, rx1, ry1;  // sampling offsets of this lane's 16 pattern points are computed below
    if (P.use_angle) {
        // IC_Angle (:75-102) on the un-blurred level; lanes split the 31 rows of the circular patch
        int pitch;
        const uint8_t* im = level_ptr(P, S, img, level, pitch);
        int m10 = 0, m01 = 0;
        if (lane < 31) {
            const int v = lane - 15;
            const int d = P.umax[v < 0 ? -v : v];
            const uint8_t* row = im + (size_t)(py + v) * pitch + px;
            int rs = 0;
            for (int u = -d; u <= d; u++) {
                const int val = row[u];
                m10 += u * val;
                rs += val;
            }
            m01 = v * rs;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            m10 += __shfl_xor_sync(0xFFFFFFFFu, m10, o);
            m01 += __shfl_xor_sync(0xFFFFFFFFu, m01, o);
        }
        angle = fast_atan2_deg((float)m01, (float)m10);
    }
    const uint8_t* bl = blurred + (size_t)img * P.pyr_stride + Lv.plane_ofs + (size_t)py * Lv.pitch + px;
    const int bp = Lv.pitch;
    int val = 0;
    const signed char* pat = c_pattern + lane * 32;  // 16 points = 8 tests = 1 descriptor byte per lane
    if (!P.use_angle) {
#pragma unroll
        for (int t = 0; t < 8; t++) {
            const int t0 = bl[pat[4 * t + 1] * bp + pat[4 * t]];
            const int t1 = bl[pat[4 * t + 3] * bp + pat[4 * t + 2]];
            val |= (t0 < t1) << t;
        }
    } else {
        const float ang = __fmul_rn(angle, (float)(3.14159265358979323846 / 180.f));  // factorPI (:105)
        const float a = (float)cos((double)ang), b = (float)sin((double)ang);
#pragma unroll
        for (int t = 0; t < 8; t++) {
            const float x0 = pat[4 * t], y0 = pat[4 * t + 1], x1 = pat[4 * t + 2], y1 = pat[4 * t + 3];
            ry0 = __float2int_rn(__fadd_rn(__fmul_rn(x0, b), __fmul_rn(y0, a)));
            rx0 = __float2int_rn(__fsub_rn(__fmul_rn(x0, a), __fmul_rn(y0, b)));
            ry1 = __float2int_rn(__fadd_rn(__fmul_rn(x1, b), __fmul_rn(y1, a)));
            rx1 = __float2int_rn(__fsub_rn(__fmul_rn(x1, a), __fmul_rn(y1, b)));
            val |= ((int)bl[ry0 * bp + rx0] < (int)bl[ry1 * bp + rx1]) << t;
        }
    }
    // 32 descriptor bytes -> 8 lanes x 4 bytes
    int w4 = val & 0xFF;
    w4 |= (__shfl_down_sync(0xFFFFFFFFu, val, 1) & 0xFF) << 8;
    w4 |= (__shfl_down_sync(0xFFFFFFFFu, val, 2) & 0xFF) << 16;
    w4 |= (__shfl_down_sync(0xFFFFFFFFu, val, 3) & 0xFF) << 24;
    const size_t slot = (size_t)img * P.kp_cap + k;
    if ((lane & 3) == 0) reinterpret_cast<uint32_t*>(desc + slot * 32)[lane >> 2] = (uint32_t)w4;
    if (lane == 0) {
        KeypointPod kp;
        kp.x = (float)px;
        kp.y = (float)py;
        if (level != 0) {  // :1149-1151  pt *= scale (f32)
            kp.x = __fmul_rn(kp.x, Lv.scale);
            kp.y = __fmul_rn(kp.y, Lv.scale);
        }
        kp.size = (float)Lv.scaled_patch;
        kp.angle = angle;
        kp.response = (float)mcs_oct::cand_score(rec);
        kp.octave = level;
        kp.class_id = -1;
        kps[slot] = kp;
    }
}

// operator() :1153-1162 with a non-trivial vLappingArea: keypoints with lap0 <= pt.x <= lap1 go to the BACK (filled
// downwards), the others to the front, both in encounter order.  One block per image; in-place via a staging copy.
__global__ void __launch_bounds__(256) lapping_permute_kernel(const __grid_constant__ McsPlan P, int lap0, int lap1,
                                                              KeypointPod* __restrict__ kps, uint8_t* __restrict__ desc,
                                                              KeypointPod* __restrict__ kps_tmp, uint8_t* __restrict__ desc_tmp,
                                                              const int* __restrict__ kp_count, int* __restrict__ mono_count) {
    __shared__ int s_flags[1024];
    __shared__ int s_carry;
    const int img = blockIdx.x;
    const int n = kp_count[img];
    KeypointPod* K = kps + (size_t)img * P.kp_cap;
    KeypointPod* KT = kps_tmp + (size_t)img * P.kp_cap;
    uint4* D = reinterpret_cast<uint4*>(desc + (size_t)img * P.kp_cap * 32);
    uint4* DT = reinterpret_cast<uint4*>(desc_tmp + (size_t)img * P.kp_cap * 32);
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        KT[i] = K[i];
        DT[2 * i] = D[2 * i];
        DT[2 * i + 1] = D[2 * i + 1];
    }
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    CudaBlockCtx ctx;
    for (int base = 0; base < n; base += 1024) {
        const int m = min(1024, n - base);
        for (int i = threadIdx.x; i < m; i += blockDim.x) {
            const float x = KT[base + i].x;
            s_flags[i] = (x >= (float)lap0 && x <= (float)lap1) ? 1 : 0;
        }
        __syncthreads();
        const int carry = s_carry;
        int stereo_here = ctx.exclusive_scan(s_flags, m);
        for (int i = threadIdx.x; i < m; i += blockDim.x) {
            const float x = KT[base + i].x;
            const bool st = (x >= (float)lap0 && x <= (float)lap1);
            const int stereo_before = carry + s_flags[i];
            const int g = base + i;
            const int dst = st ? (n - 1 - stereo_before) : (g - stereo_before);
            K[dst] = KT[g];
            D[2 * dst] = DT[2 * g];
            D[2 * dst + 1] = DT[2 * g + 1];
        }
        __syncthreads();
        if (threadIdx.x == 0) s_carry = carry + stereo_here;
        __syncthreads();
    }
    if (threadIdx.x == 0) mono_count[img] = n - s_carry;
}

}  // namespace mcs
