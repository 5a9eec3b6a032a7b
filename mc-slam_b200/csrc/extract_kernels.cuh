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
// K0  MultiCameraFrame::setData input conversion (/root/reference/MCSlam/src/MultiCameraFrame.cpp:109-116): the app holds CV_32F
// images in [0, 1] (imread / 255, MCDataUtils/src/DatasetReader.cpp:699-711).  multiply(img, 255) in f32, convertTo(CV_8U) =
// saturate_cast<uchar>(cvRound(v)) (round half to even), and for 3 interleaved channels cvtColor(BGR2GRAY) on the 8-bit values:
// (B * 3735 + G * 19235 + R * 9798 + 2^14) >> 15.  4 pixels per thread, 128-bit loads, one 32-bit store: HBM-bound.
// ------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t f32_to_u8(float v) {
    const int r = __float2int_rn(__fmul_rn(v, 255.0f));  // NaN -> 0 like saturate_cast of cvRound's INT_MIN
    return (uint32_t)min(max(r, 0), 255);
}

template <int CH>
__global__ void __launch_bounds__(256) stage_f32_kernel(const float* __restrict__ src, uint8_t* __restrict__ dst, size_t n_quads) {
    for (size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x; q < n_quads; q += (size_t)gridDim.x * blockDim.x) {
        const float4* p = reinterpret_cast<const float4*>(src) + q * CH;
        uint32_t o;
        if (CH == 1) {
            const float4 a = __ldcs(p);
            o = f32_to_u8(a.x) | (f32_to_u8(a.y) << 8) | (f32_to_u8(a.z) << 16) | (f32_to_u8(a.w) << 24);
        } else {
            const float4 a = __ldcs(p), b = __ldcs(p + 1), c = __ldcs(p + 2);
            const float v[12] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w, c.x, c.y, c.z, c.w};
            o = 0;
#pragma unroll
            for (int k = 0; k < 4; k++) {
                const uint32_t B = f32_to_u8(v[3 * k]), G = f32_to_u8(v[3 * k + 1]), R = f32_to_u8(v[3 * k + 2]);
                o |= ((B * 3735u + G * 19235u + R * 9798u + (1u << 14)) >> 15) << (8 * k);
            }
        }
        reinterpret_cast<uint32_t*>(dst)[q] = o;
    }
}

// ------------------------------------------------------------------------------------------------------------
// K0b  cv::undistort of setData's RECTIFY branch (/root/reference/MCSlam/src/MultiCameraFrame.cpp:123-136) as a gather: the host
// builds cv::initUndistortRectifyMap's CV_16SC2 fixed-point maps once per camera (map.x = sx | sy << 16 as int16, map.y = fy * 32 + fx,
// 5-bit fractions); remap(INTER_LINEAR, BORDER_CONSTANT 0) is 4 taps with the integer weights (32 - fy)(32 - fx) * 32 ... (the
// entries of OpenCV's 15-bit table are exactly these products) and (v + 2^14) >> 15.  4 output pixels per thread; the result goes
// straight into the pyramid's level-0 plane.  The f32 variant (segmentation masks, :131-134) uses the float weights in OpenCV's order.
// ------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) undistort_u8_kernel(const uint8_t* __restrict__ src, size_t src_stride, uint8_t* __restrict__ dst,
                                                           size_t dst_stride, int dpitch, const uint2* __restrict__ maps, int num_cams,
                                                           int img0, int w, int h) {
    const int g = blockIdx.x * 32 + threadIdx.x, y = blockIdx.y * 8 + threadIdx.y;  // g = group of 4 pixels
    if (4 * g >= w || y >= h) return;
    const int img = blockIdx.z;
    const uint8_t* S = src + (size_t)img * src_stride;
    const uint2* M = maps + (size_t)((img0 + img) % num_cams) * w * h + (size_t)y * w + 4 * g;
    uint32_t out = 0;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        if (4 * g + k >= w) break;
        const uint2 m = __ldg(M + k);
        const int sx = (short)(m.x & 0xFFFFu), sy = (short)(m.x >> 16), fx = (int)(m.y & 31u), fy = (int)(m.y >> 5);
        const bool x0 = sx >= 0 && sx < w, x1 = sx + 1 >= 0 && sx + 1 < w, y0 = sy >= 0 && sy < h, y1 = sy + 1 >= 0 && sy + 1 < h;
        const uint8_t* r0 = S + (size_t)sy * w + sx;
        const int p00 = (x0 && y0) ? r0[0] : 0, p01 = (x1 && y0) ? r0[1] : 0, p10 = (x0 && y1) ? r0[w] : 0, p11 = (x1 && y1) ? r0[w + 1] : 0;
        const int v = p00 * ((32 - fy) * (32 - fx) * 32) + p01 * ((32 - fy) * fx * 32) + p10 * (fy * (32 - fx) * 32) + p11 * (fy * fx * 32);
        out |= (uint32_t)((v + (1 << 14)) >> 15) << (8 * k);
    }
    *reinterpret_cast<uint32_t*>(dst + (size_t)img * dst_stride + (size_t)y * dpitch + 4 * g) = out;  // dpitch % 64 == 0
}

__global__ void __launch_bounds__(256) undistort_f32_kernel(const float* __restrict__ src, float* __restrict__ dst,
                                                            const uint2* __restrict__ maps, int num_cams, int w, int h) {
    const int x = blockIdx.x * 32 + threadIdx.x, y = blockIdx.y * 8 + threadIdx.y;
    if (x >= w || y >= h) return;
    const int img = blockIdx.z;
    const float* S = src + (size_t)img * w * h;
    const uint2 m = __ldg(maps + (size_t)(img % num_cams) * w * h + (size_t)y * w + x);
    const int sx = (short)(m.x & 0xFFFFu), sy = (short)(m.x >> 16);
    const float fx = __fmul_rn((float)(m.y & 31u), 1.f / 32), fy = __fmul_rn((float)(m.y >> 5), 1.f / 32);
    const bool x0 = sx >= 0 && sx < w, x1 = sx + 1 >= 0 && sx + 1 < w, y0 = sy >= 0 && sy < h, y1 = sy + 1 >= 0 && sy + 1 < h;
    const float* r0 = S + (size_t)sy * w + sx;
    const float p00 = (x0 && y0) ? r0[0] : 0.f, p01 = (x1 && y0) ? r0[1] : 0.f, p10 = (x0 && y1) ? r0[w] : 0.f, p11 = (x1 && y1) ? r0[w + 1] : 0.f;
    const float w00 = __fmul_rn(1.f - fy, 1.f - fx), w01 = __fmul_rn(1.f - fy, fx), w10 = __fmul_rn(fy, 1.f - fx), w11 = __fmul_rn(fy, fx);
    dst[(size_t)img * w * h + (size_t)y * w + x] =
        __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(p00, w00), __fmul_rn(p01, w01)), __fmul_rn(p10, w10)), __fmul_rn(p11, w11));
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

// Word-load variant used when the level-to-level scale is <= 2 (the 4 output pixels of a thread then read at most 12 consecutive
// source bytes per row): 3 aligned 32-bit loads per source row instead of 8 byte loads, the 2-tap horizontal filter as one DP2A
// (coefficients a0 | a1 << 16 from the table, pixels b0 | b1 << 8 cut out with a funnel shift).  Same arithmetic as above.
__global__ void __launch_bounds__(256) resize_level_words_kernel(const __grid_constant__ McsPlan P, ImageSet S, int level,
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
    const int sy0 = min(max(yt.x, 0), sh - 1), sy1 = min(max(yt.x + 1, 0), sh - 1);
    const int b0 = (short)(yt.y & 0xFFFF), b1 = yt.y >> 16;
    const uint32_t* r0 = reinterpret_cast<const uint32_t*>(src + (uint32_t)(sy0 * spitch));
    const uint32_t* r1 = reinterpret_cast<const uint32_t*>(src + (uint32_t)(sy1 * spitch));
    // 4 table entries {sx, a0 | a1 << 16} as two 128-bit loads (xtab_ofs is even and the table is padded past the row end)
    const int4* xp = reinterpret_cast<const int4*>(xtab + D.xtab_ofs + dx0);
    const int4 xa = __ldg(xp), xb = __ldg(xp + 1);
    const int sx[4] = {xa.x, xa.z, xb.x, xb.z};
    const uint32_t cf[4] = {(uint32_t)xa.y, (uint32_t)xa.w, (uint32_t)xb.y, (uint32_t)xb.w};
    const int lastw = ((sw + 3) >> 2) - 1;
    const int wi = sx[0] >> 2, wi1 = min(wi + 1, lastw), wi2 = min(wi + 2, lastw);
    const uint32_t u0 = __ldg(r0 + wi), u1 = __ldg(r0 + wi1), u2 = __ldg(r0 + wi2);
    const uint32_t v0 = __ldg(r1 + wi), v1 = __ldg(r1 + wi1), v2 = __ldg(r1 + wi2);
    uint32_t outw = 0;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        // bytes o, o+1 of the 12 loaded ones (o <= 10 for scale <= 2; beyond the row end o+1 only ever meets coefficient 0):
        // one PRMT on the word pair that holds them.  Pixels past the row end (table padding) give garbage that is never stored.
        const int o = min(max(sx[k] - 4 * wi, 0), 10);
        const bool hi = o >= 4;
        const uint32_t selv = (uint32_t)((o & 3) + (o >= 8 ? 4 : 0)) * 0x11u + 0x10u;
        const uint32_t pu = __byte_perm(hi ? u1 : u0, hi ? u2 : u1, selv);
        const uint32_t pv = __byte_perm(hi ? v1 : v0, hi ? v2 : v1, selv);
        const int h0 = (int)__dp2a_lo(cf[k], pu, 0u);
        const int h1 = (int)__dp2a_lo(cf[k], pv, 0u);
        int v = (((b0 * (h0 >> 4)) >> 16) + ((b1 * (h1 >> 4)) >> 16) + 2) >> 2;
        v = min(max(v, 0), 255);
        outw |= (uint32_t)v << (8 * k);
    }
    *reinterpret_cast<uint32_t*>(dst + (uint32_t)(dy * D.pitch + dx0)) = outw;  // pitch % 64 == 0: row padding absorbs the tail
}

// Third form of K1, used when the host could build the selector tables (build_resize_word_tables: scale <= 2 and non-negative
// coefficients that sum to <= 2049, which also makes the 0..255 saturation a no-op).  Per output word the table holds the first
// source word, the byte shift that aligns source pixel sx[0] to byte 0 of an 8-byte window, four PRMT selectors relative to that
// window and the four coefficient pairs; per output row the two clamped source rows and the vertical coefficients.  No per-pixel
// index arithmetic is left in the kernel and all launch constants arrive as scalar arguments.
struct ResizeSelArgs {
    const uint8_t* src;   // level l-1 of image 0
    uint8_t* dst;         // level l of image 0
    size_t src_stride, dst_stride;  // bytes between images
    int spitch, dpitch;
    int lastw;            // last valid 32-bit word of a source row
    int sh;               // source rows
    int dwords, dh;       // output words per row, output rows
    const int4* xw;       // [dwords][2]: {wi, selectors, c0, c1}, {c2, c3, shift, 0}
    const int4* yw;       // [dh]: {sy0, sy1, b0, b1}
};

// Each thread produces MCS_RS_ROWS consecutive output rows of its 4-pixel word.  Those rows read a contiguous band of at most NR
// source rows (<= 6 for the 1.2 pyramid, host-checked per level): the thread loads the band once (all loads in flight together, one
// 64-bit row pointer each, words at +0 / +4 / +8), runs the horizontal pass once per source row and parks the four sums as one
// 16-byte shared-memory slot per (row, thread), because the output rows pick their two source rows by a run-time index
// (ytab) that registers cannot be indexed with.  The vertical pass is two LDS.128, eight IMAD.HI and a pack per output row.
// First version (one output row per thread): fmaheavy pipe 71 % busy (DP2A + IMAD.HI + IMAD.WIDE address math + IMAD.MOV copies),
// 28 instructions per pixel; this one runs the horizontal pass on ~1.4 instead of 2 source rows per output row and sets the column
// tables up once per 16 pixels.  CLAMP: level 0 may alias a caller's tight buffer, where the two words after the last word of a row
// must not be touched for the last row of the allocation; the library's own planes are padded.
#define MCS_RS_ROWS 4
template <int NR, bool CLAMP>
__global__ void __launch_bounds__(256) resize_level_sel_kernel(const ResizeSelArgs A) {
    __shared__ uint4 s_h[NR][256];
    const int g = blockIdx.x * 32 + threadIdx.x, dy0 = (blockIdx.y * 8 + threadIdx.y) * MCS_RS_ROWS;
    if (g >= A.dwords || dy0 >= A.dh) return;
    const int tid = threadIdx.y * 32 + threadIdx.x;
    const int4 xa = __ldg(A.xw + 2 * g), xb = __ldg(A.xw + 2 * g + 1);
    int4 yt[MCS_RS_ROWS];
#pragma unroll
    for (int r = 0; r < MCS_RS_ROWS; r++) yt[r] = __ldg(A.yw + min(dy0 + r, A.dh - 1));  // {sy0, sy1, b0, b1}, rows clamped on the host
    const int wi = xa.x;
    const uint8_t* col = A.src + blockIdx.z * A.src_stride + 4 * wi;
    const int o1 = CLAMP ? 4 * (min(wi + 1, A.lastw) - wi) : 4, o2 = CLAMP ? 4 * (min(wi + 2, A.lastw) - wi) : 8;
    const int smin = yt[0].x;
    uint32_t u[NR][3];
#pragma unroll
    for (int i = 0; i < NR; i++) {
        const uint8_t* rp = col + (uint32_t)(min(smin + i, A.sh - 1) * A.spitch);
        u[i][0] = __ldg(reinterpret_cast<const uint32_t*>(rp));
        u[i][1] = __ldg(reinterpret_cast<const uint32_t*>(rp + o1));
        u[i][2] = __ldg(reinterpret_cast<const uint32_t*>(rp + o2));
    }
    const uint32_t sh = (uint32_t)xb.z;
    const uint32_t cf[4] = {(uint32_t)xa.z, (uint32_t)xa.w, (uint32_t)xb.x, (uint32_t)xb.y};
    const uint32_t sl[4] = {(uint32_t)xa.y, (uint32_t)xa.y >> 8, (uint32_t)xa.y >> 16, (uint32_t)xa.y >> 24};
    // horizontal pass: (sum of two taps) >> 4, as cv::resize stores it between its passes
#pragma unroll
    for (int i = 0; i < NR; i++) {
        const uint32_t ua = __funnelshift_r(u[i][0], u[i][1], sh), ub = __funnelshift_r(u[i][1], u[i][2], sh);
        uint4 hv;
        hv.x = __dp2a_lo(cf[0], __byte_perm(ua, ub, sl[0]), 0u) >> 4;
        hv.y = __dp2a_lo(cf[1], __byte_perm(ua, ub, sl[1]), 0u) >> 4;
        hv.z = __dp2a_lo(cf[2], __byte_perm(ua, ub, sl[2]), 0u) >> 4;
        hv.w = __dp2a_lo(cf[3], __byte_perm(ua, ub, sl[3]), 0u) >> 4;
        s_h[i][tid] = hv;
    }
    uint8_t* drow = A.dst + blockIdx.z * A.dst_stride + (uint32_t)(dy0 * A.dpitch + 4 * g);
#pragma unroll
    for (int r = 0; r < MCS_RS_ROWS; r++) {
        if (dy0 + r >= A.dh) break;
        const uint4 h0 = s_h[yt[r].x - smin][tid], h1 = s_h[yt[r].y - smin][tid];  // own slots only: no barrier
        // (b * h) >> 16 as the high half of (b << 16) * h: coefficients and sums are non-negative (build_resize_word_tables)
        const uint32_t b0 = (uint32_t)yt[r].z << 16, b1 = (uint32_t)yt[r].w << 16;
        const uint32_t v0 = (__umulhi(b0, h0.x) + __umulhi(b1, h1.x) + 2u) >> 2;  // in 0..255 by construction of the tables
        const uint32_t v1 = (__umulhi(b0, h0.y) + __umulhi(b1, h1.y) + 2u) >> 2;
        const uint32_t v2 = (__umulhi(b0, h0.z) + __umulhi(b1, h1.z) + 2u) >> 2;
        const uint32_t v3 = (__umulhi(b0, h0.w) + __umulhi(b1, h1.w) + 2u) >> 2;
        *reinterpret_cast<uint32_t*>(drow + (uint32_t)(r * A.dpitch)) = __byte_perm(__byte_perm(v0, v1, 0x0040), __byte_perm(v2, v3, 0x0040), 0x5410);
    }
}

// ------------------------------------------------------------------------------------------------------------
// K2  cv::GaussianBlur(7x7, sigma 2, BORDER_REFLECT_101) on 8U: 8.8 fixed-point kernel {18,34,48,56,48,34,18}/256,
// exact integer accumulation, one rounding (v + 32768) >> 16.  Tile 64x16 + 3 px halo staged in shared memory,
// horizontal pass into a u16 shared tile, vertical pass from it.  All levels of all images in one launch.
// ------------------------------------------------------------------------------------------------------------
__constant__ int c_gauss7[7];  // set by the host from the bit-exact kernel generator

// Tile 128 x 32 outputs (+3 px halo) per block.  Pass H: 7-tap row filter as two DP4A per pixel on byte windows cut out of the
// staged 32-bit words with funnel shifts; results (<= 65280, u16) are stored as vertical pairs (row 2p | row 2p+1 << 16).  Pass V:
// 7-tap column filter as four DP2A per pixel on those pairs; one thread makes 4 x 2 pixels and stores two 32-bit words.
__global__ void __launch_bounds__(256) blur_kernel(const __grid_constant__ McsPlan P, ImageSet S, uint8_t* __restrict__ blurred) {
    constexpr int TW = MCS_BLUR_TW, TH = MCS_BLUR_TH, NWI = TW / 4 + 2, RI = TH + 6, NP = RI / 2;
    static_assert(TW == 128 && TH == 32, "tile geometry is baked into the index math");
    __shared__ uint32_t s_in[RI][NWI];          // input words: column 0 = pixels x0-4 .. x0-1
    __shared__ __align__(16) uint32_t s_pair[NP][TW];  // H-filtered rows as vertical pairs
    const int img = blockIdx.y;
    int level = 0;
#pragma unroll 1
    for (int l = 1; l < P.nlevels; l++)
        if ((int)blockIdx.x >= P.L[l].tile_base) level = l;
    const McsLevel& Lv = P.L[level];
    const int t = blockIdx.x - Lv.tile_base;
    const int x0 = (t % Lv.tiles_x) * TW, y0 = (t / Lv.tiles_x) * TH;
    int pitch;
    const uint8_t* src = level_ptr(P, S, img, level, pitch);
    const int w = Lv.w, h = Lv.h;
    for (int i = threadIdx.x; i < RI * NWI; i += 256) {
        const int r = i / NWI, c = i - r * NWI;
        const int gy = reflect101(y0 + r - 3, h), gx = x0 - 4 + 4 * c;
        const uint8_t* row = src + (size_t)gy * pitch;
        uint32_t v;
        if (gx >= 0 && gx + 3 < w) {
            v = __ldg(reinterpret_cast<const uint32_t*>(row + gx));
        } else {  // words touching the left/right border: BORDER_REFLECT_101 per byte
            v = (uint32_t)row[reflect101(gx, w)] | ((uint32_t)row[reflect101(gx + 1, w)] << 8) |
                ((uint32_t)row[reflect101(gx + 2, w)] << 16) | ((uint32_t)row[reflect101(gx + 3, w)] << 24);
        }
        s_in[r][c] = v;
    }
    __syncthreads();
    const uint32_t wA = (uint32_t)c_gauss7[0] | ((uint32_t)c_gauss7[1] << 8) | ((uint32_t)c_gauss7[2] << 16) | ((uint32_t)c_gauss7[3] << 24);
    const uint32_t wB = (uint32_t)c_gauss7[4] | ((uint32_t)c_gauss7[5] << 8) | ((uint32_t)c_gauss7[6] << 16);
    // pass H: item = (row pair p, word g) -> 4 pixels x 2 rows
    for (int i = threadIdx.x; i < NP * (TW / 4); i += 256) {
        const int p = i >> 5, g = i & 31;
        uint32_t hv[2][4];
#pragma unroll
        for (int rr = 0; rr < 2; rr++) {
            const uint32_t W0 = s_in[2 * p + rr][g], W1 = s_in[2 * p + rr][g + 1], W2 = s_in[2 * p + rr][g + 2];
            // pixel e of the word needs bytes (4g + e + 1) .. (4g + e + 7) of the staged row
            hv[rr][0] = __dp4a(__funnelshift_r(W1, W2, 8), wB, __dp4a(__funnelshift_r(W0, W1, 8), wA, 0u));
            hv[rr][1] = __dp4a(__funnelshift_r(W1, W2, 16), wB, __dp4a(__funnelshift_r(W0, W1, 16), wA, 0u));
            hv[rr][2] = __dp4a(__funnelshift_r(W1, W2, 24), wB, __dp4a(__funnelshift_r(W0, W1, 24), wA, 0u));
            hv[rr][3] = __dp4a(W2, wB, __dp4a(W1, wA, 0u));
        }
        uint4 o;
        o.x = hv[0][0] | (hv[1][0] << 16);
        o.y = hv[0][1] | (hv[1][1] << 16);
        o.z = hv[0][2] | (hv[1][2] << 16);
        o.w = hv[0][3] | (hv[1][3] << 16);
        *reinterpret_cast<uint4*>(&s_pair[p][4 * g]) = o;
    }
    __syncthreads();
    const uint32_t k0 = c_gauss7[0], k1 = c_gauss7[1], k2 = c_gauss7[2], k3 = c_gauss7[3], k4 = c_gauss7[4], k5 = c_gauss7[5], k6 = c_gauss7[6];
    const uint32_t e0 = k0 | (k1 << 8), e1 = k2 | (k3 << 8), e2 = k4 | (k5 << 8), e3 = k6;               // even output row 2q
    const uint32_t o0 = k0 << 8, o1 = k1 | (k2 << 8), o2 = k3 | (k4 << 8), o3 = k5 | (k6 << 8);          // odd output row 2q+1
    uint8_t* dst = blurred + (size_t)img * P.pyr_stride + Lv.plane_ofs;
    // pass V: item = (output row pair q, word g)
    for (int i = threadIdx.x; i < (TH / 2) * (TW / 4); i += 256) {
        const int q = i >> 5, g = i & 31;
        const int gx = x0 + 4 * g, gy = y0 + 2 * q;
        if (gx >= w || gy >= h) continue;
        const uint4 a = *reinterpret_cast<const uint4*>(&s_pair[q][4 * g]);
        const uint4 b = *reinterpret_cast<const uint4*>(&s_pair[q + 1][4 * g]);
        const uint4 c = *reinterpret_cast<const uint4*>(&s_pair[q + 2][4 * g]);
        const uint4 d = *reinterpret_cast<const uint4*>(&s_pair[q + 3][4 * g]);
        const uint32_t pa[4] = {a.x, a.y, a.z, a.w}, pb[4] = {b.x, b.y, b.z, b.w}, pc[4] = {c.x, c.y, c.z, c.w}, pd[4] = {d.x, d.y, d.z, d.w};
        uint32_t oe = 0, oo = 0;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const uint32_t ve = __dp2a_lo(pd[k], e3, __dp2a_lo(pc[k], e2, __dp2a_lo(pb[k], e1, __dp2a_lo(pa[k], e0, 32768u))));
            const uint32_t vo = __dp2a_lo(pd[k], o3, __dp2a_lo(pc[k], o2, __dp2a_lo(pb[k], o1, __dp2a_lo(pa[k], o0, 32768u))));
            oe |= (ve >> 16) << (8 * k);
            oo |= (vo >> 16) << (8 * k);
        }
        *reinterpret_cast<uint32_t*>(dst + (size_t)gy * Lv.pitch + gx) = oe;  // pitch % 64 == 0 and gx % 4 == 0
        if (gy + 1 < h) *reinterpret_cast<uint32_t*>(dst + (size_t)(gy + 1) * Lv.pitch + gx) = oo;
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
// ring offsets of FAST 9/16 in OpenCV order: (0,3)(1,3)(2,2)(3,1)(3,0)(3,-1)(2,-2)(1,-3)(0,-3)(-1,-3)(-2,-2)(-3,-1)(-3,0)(-3,1)(-2,2)(-1,3)
// ------------------------------------------------------------------------------------------------------------
#ifndef MCS_FAST_NT
#define MCS_FAST_NT 128
#endif
#define MCS_FAST_TPW 21  // tile pitch in 32-bit words: sub-image <= 75 px + 1 + staging slack (compile-time so ring offsets are immediates)
#define MCS_FAST_SP 72   // score tile pitch (bytes)

// One polarity only (sgn = +1: ring brighter than the centre, -1: darker): max over the 16 arcs of min over the arc of
// sgn * (ring - v), minus 1.  Ring pixels are packed as ANTIPODAL pairs P_k = (e_k, e_{k+8}), e = 256 + sgn * (ring - v), so that
// "next ring position" is the next word and wrapping past position 7 is a half-word swap: 8 swaps instead of 24 byte permutes.
__device__ __forceinline__ int fast_strength_pol(const uint8_t* c, int tp, int sgn) {
    const int v = c[0];
    const uint32_t base = (uint32_t)(256 - sgn * v) * 0x00010001u;
    uint32_t Pk[10];
    Pk[0] = base + (uint32_t)sgn * ((uint32_t)c[3 * tp] | ((uint32_t)c[-3 * tp] << 16));           // ( 0, 3) | ( 0,-3)
    Pk[1] = base + (uint32_t)sgn * ((uint32_t)c[3 * tp + 1] | ((uint32_t)c[-3 * tp - 1] << 16));   // ( 1, 3) | (-1,-3)
    Pk[2] = base + (uint32_t)sgn * ((uint32_t)c[2 * tp + 2] | ((uint32_t)c[-2 * tp - 2] << 16));   // ( 2, 2) | (-2,-2)
    Pk[3] = base + (uint32_t)sgn * ((uint32_t)c[tp + 3] | ((uint32_t)c[-tp - 3] << 16));           // ( 3, 1) | (-3,-1)
    Pk[4] = base + (uint32_t)sgn * ((uint32_t)c[3] | ((uint32_t)c[-3] << 16));                     // ( 3, 0) | (-3, 0)
    Pk[5] = base + (uint32_t)sgn * ((uint32_t)c[-tp + 3] | ((uint32_t)c[tp - 3] << 16));           // ( 3,-1) | (-3, 1)
    Pk[6] = base + (uint32_t)sgn * ((uint32_t)c[-2 * tp + 2] | ((uint32_t)c[2 * tp - 2] << 16));   // ( 2,-2) | (-2, 2)
    Pk[7] = base + (uint32_t)sgn * ((uint32_t)c[-3 * tp + 1] | ((uint32_t)c[3 * tp - 1] << 16));   // ( 1,-3) | (-1, 3)
    Pk[8] = __byte_perm(Pk[0], 0, 0x1032);
    Pk[9] = __byte_perm(Pk[1], 0, 0x1032);
    uint32_t Q[14];
#pragma unroll
    for (int k = 0; k < 8; k++) Q[k] = __vimin3_u16x2(Pk[k], Pk[k + 1], Pk[k + 2]);  // min of 3 consecutive ring positions
#pragma unroll
    for (int k = 0; k < 6; k++) Q[8 + k] = __byte_perm(Q[k], 0, 0x1032);
    uint32_t R[8];
#pragma unroll
    for (int k = 0; k < 8; k++) R[k] = __vimin3_u16x2(Q[k], Q[k + 3], Q[k + 6]);     // min of 9 consecutive ring positions
    const uint32_t m = __vimax3_u16x2(__vimax3_u16x2(R[0], R[1], R[2]), __vimax3_u16x2(R[3], R[4], R[5]), __vmaxu2(R[6], R[7]));
    return (int)max(m & 0xFFFFu, m >> 16) - 257;
}

// Shared-memory carve-up (sizes from the plan's largest cell, so small-cell configurations keep high occupancy):
//   tile  [ch][fast_tpw] u32   sub-image, staged so that sub-image column c sits at byte c + 1 (interior column 0 = byte 4: aligned)
//   score [ih + 2][fast_sp] u8 corner scores with a 1 px zero frame
//   flags [ih][G] u8           pass-1 result per group of 4 pixels: bit 2e = bright pre-test of pixel e, bit 2e+1 = dark
//   queue u16                  one entry per (pixel, polarity) that passed: corner << 15 | dark << 14 | ly << 7 | lx
//   surv  u32                  NMS survivors (packed candidates)
// c_magic20[n] = ceil(2^20 / n): t / n == (t * c_magic20[n]) >> 20 for the small t, n of the cell loops (checked on the host for
// every plan, build_cell_table)
#define MCS_M20(n) (((1u << 20) + (n) - 1u) / (n))
__constant__ uint32_t c_magic20[32] = {0,           MCS_M20(1),  MCS_M20(2),  MCS_M20(3),  MCS_M20(4),  MCS_M20(5),  MCS_M20(6),  MCS_M20(7),
                                       MCS_M20(8),  MCS_M20(9),  MCS_M20(10), MCS_M20(11), MCS_M20(12), MCS_M20(13), MCS_M20(14), MCS_M20(15),
                                       MCS_M20(16), MCS_M20(17), MCS_M20(18), MCS_M20(19), MCS_M20(20), MCS_M20(21), MCS_M20(22), MCS_M20(23),
                                       MCS_M20(24), MCS_M20(25), MCS_M20(26), MCS_M20(27), MCS_M20(28), MCS_M20(29), MCS_M20(30), MCS_M20(31)};

__global__ void __launch_bounds__(MCS_FAST_NT) fast_cells_kernel(const __grid_constant__ McsPlan P, ImageSet S,
                                                                 const uint4* __restrict__ cell_tab, uint32_t* __restrict__ cand,
                                                                 int* __restrict__ cand_count) {
    extern __shared__ __align__(16) unsigned char s_fast[];
    __shared__ int s_wsum[MCS_FAST_NT / 32], s_nsurv;
    uint32_t* s_tile = reinterpret_cast<uint32_t*>(s_fast);
    uint8_t* s_score = s_fast + P.fast_ofs_score;
    uint8_t* s_flags = s_fast + P.fast_ofs_flags;
    uint16_t* s_queue = reinterpret_cast<uint16_t*>(s_fast + P.fast_ofs_queue);
    uint32_t* s_surv = reinterpret_cast<uint32_t*>(s_fast + P.fast_ofs_surv);
    constexpr int TPW = MCS_FAST_TPW, TP = TPW * 4, SP = MCS_FAST_SP;
    const int img = blockIdx.y;
    // host-built cell record (build_cell_table): level, sub-image origin and size, candidate coordinate offset; cells the
    // reference skips (:809, :818) are not in the table
    const uint4 ct = __ldg(cell_tab + blockIdx.x);
    const int level = (int)ct.x;
    const McsLevel& Lv = P.L[level];
    const int iniX = (int)(ct.y & 0xFFFFu), iniY = (int)(ct.y >> 16);
    const int cw = (int)(ct.z & 0xFFFFu), ch = (int)(ct.z >> 16);  // sub-image handed to cv::FAST
    const int iw = cw - 6, ih = ch - 6;                            // computed interior
    const int xoff = (int)(ct.w & 0xFFFFu), yoff = (int)(ct.w >> 16);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int pitch;
    const uint8_t* src = level_ptr(P, S, img, level, pitch);
    const int G = (iw + 3) >> 2;  // groups of 4 interior pixels per row
    // stage the sub-image: shared word c of row r = image bytes (iniX - 1 + 4c) .. +3, cut out of two aligned 32-bit loads
    {
        const int nws = G + 2, gb = iniX - 1, aw = gb >> 2, sh = (gb & 3) * 8;
        const uint32_t mg = c_magic20[nws];  // t / nws == (t * mg) >> 20 (t < 76 * 21)
        for (int t = threadIdx.x; t < nws * ch; t += MCS_FAST_NT) {
            const int r = (int)(((uint32_t)t * mg) >> 20), c = t - r * nws;
            const uint32_t* g = reinterpret_cast<const uint32_t*>(src + (size_t)(iniY + r) * pitch) + aw + c;
            s_tile[r * TPW + c] = __funnelshift_r(__ldg(g), __ldg(g + 1), sh);
        }
    }
    const int ngroups = G * ih;
    const uint32_t mgG = c_magic20[G];
    uint32_t* out = cand + (size_t)img * P.cand_stride + Lv.cand_ofs;
    int* cnt = cand_count + img * P.nlevels + level;
    // FAST(cell, iniThFAST); only if that finds nothing, FAST(cell, minThFAST)  (:825, :842-845)
#pragma unroll 1
    for (int phase = 0; phase < 2; phase++) {
        const int th = phase == 0 ? P.ini_th : P.min_th;
        if (threadIdx.x == 0) s_nsurv = 0;
        for (int t = threadIdx.x; t < (ih + 2) * (SP >> 2); t += MCS_FAST_NT) reinterpret_cast<uint32_t*>(s_score)[t] = 0;
        __syncthreads();
        // pass 1, every interior pixel, 4 per thread on packed 16-bit lanes: necessary condition for a 9-arc — of each opposite
        // ring pair at least one pixel is on the arc, tested on the axis and the diagonal pairs (4 of the 8):
        //   bright: min over pairs of max(pair) > v + th        dark: max over pairs of min(pair) < v - th
        const uint32_t K1 = (uint32_t)(0x8000 - th - 1) * 0x00010001u;  // bit 15 of (a + K1 - b) per lane  <=>  a - b > th
        int my_count = 0;
        for (int gi = threadIdx.x; gi < ngroups; gi += MCS_FAST_NT) {
            const int ly = (int)(((uint32_t)gi * mgG) >> 20), g = gi - ly * G;
            const uint32_t* row = s_tile + (ly + 3) * TPW + g;
            const uint32_t W0 = row[0], W1 = row[1], W2 = row[2];
            const uint32_t Up = row[1 - 3 * TPW], Dn = row[1 + 3 * TPW];
            const uint32_t Lf = __funnelshift_r(W0, W1, 8), Rt = __funnelshift_r(W1, W2, 24);
            // diagonal ring pixels (+-2, +-2): rows y-2 and y+2, columns x-2 and x+2
            const uint32_t A0 = row[-2 * TPW], A1 = row[1 - 2 * TPW], A2 = row[2 - 2 * TPW];
            const uint32_t B0 = row[2 * TPW], B1 = row[1 + 2 * TPW], B2 = row[2 + 2 * TPW];
            const uint32_t UL = __funnelshift_r(A0, A1, 16), UR = __funnelshift_r(A1, A2, 16);
            const uint32_t DL = __funnelshift_r(B0, B1, 16), DR = __funnelshift_r(B1, B2, 16);
            uint32_t f = 0;
#pragma unroll
            for (int hpair = 0; hpair < 2; hpair++) {
                const uint32_t sel = hpair ? 0x4342u : 0x4140u;
                const uint32_t c2 = __byte_perm(W1, 0, sel), u2 = __byte_perm(Up, 0, sel), d2 = __byte_perm(Dn, 0, sel);
                const uint32_t l2 = __byte_perm(Lf, 0, sel), r2 = __byte_perm(Rt, 0, sel);
                const uint32_t ul = __byte_perm(UL, 0, sel), dr = __byte_perm(DR, 0, sel);
                const uint32_t ur = __byte_perm(UR, 0, sel), dl = __byte_perm(DL, 0, sel);
                // the four opposite pairs (0,8) (4,12) (2,10) (6,14)
                const uint32_t bmin = __vimin3_u16x2(__vminu2(__vmaxu2(u2, d2), __vmaxu2(l2, r2)), __vmaxu2(ul, dr), __vmaxu2(ur, dl));
                const uint32_t dmax = __vimax3_u16x2(__vmaxu2(__vminu2(u2, d2), __vminu2(l2, r2)), __vminu2(ul, dr), __vminu2(ur, dl));
                const uint32_t Bf = bmin + K1 - c2, Df = c2 + K1 - dmax;
                const uint32_t t2 = ((Bf >> 15) & 0x00010001u) | ((Df >> 14) & 0x00020002u);
                f |= ((t2 | (t2 >> 14)) & 0xFu) << (4 * hpair);
            }
            const int nvalid = min(4, iw - 4 * g);
            f &= (1u << (2 * nvalid)) - 1u;
            s_flags[gi] = (uint8_t)f;
            my_count += __popc(f);
        }
        // block-wide exclusive scan of the per-thread entry counts -> queue offsets (deterministic order, no atomics)
        int incl = my_count;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xFFFFFFFFu, incl, o);
            if (lane >= o) incl += v;
        }
        if (lane == 31) s_wsum[warp] = incl;
        __syncthreads();
        int pos = incl - my_count, nq = 0;
#pragma unroll
        for (int w = 0; w < MCS_FAST_NT / 32; w++) {
            const int ws = s_wsum[w];
            if (w < warp) pos += ws;
            nq += ws;
        }
        for (int gi = threadIdx.x; gi < ngroups; gi += MCS_FAST_NT) {
            uint32_t f = s_flags[gi];
            if (!f) continue;
            const int ly = (int)(((uint32_t)gi * mgG) >> 20), g = gi - ly * G;
            const uint32_t base = (uint32_t)((ly << 7) | (4 * g));
#pragma unroll
            for (int b = 0; b < 8; b++) {  // predicated stores: no divergent loop
                if (f & (1u << b)) {
                    s_queue[pos] = (uint16_t)(((b & 1) << 14) | (base + (b >> 1)));
                    pos++;
                }
            }
        }
        __syncthreads();
        // pass 2, queued (pixel, polarity) entries on dense warps: arc strength = cv::FAST's score; corner at th <=> strength >= th.
        // A pixel cannot be a bright and a dark corner at once (two 9-arcs do not fit in 16), so at most one entry per pixel scores.
        // Corners are re-queued densely at the front of s_queue (their slot is never ahead of an entry that is still unread: the
        // barrier inside the round orders "everyone has read its entry" before "corner slots are written").
        int ncorner = 0;
        for (int q0 = 0; q0 < nq; q0 += MCS_FAST_NT) {
            const int q = q0 + threadIdx.x;
            int e = 0, sc = 0;
            bool corner = false;
            if (q < nq) {
                e = s_queue[q];
                const int ly = (e >> 7) & 0x7F, lx = e & 0x7F;
                const uint8_t* c = reinterpret_cast<const uint8_t*>(s_tile) + (ly + 3) * TP + lx + 4;
                sc = fast_strength_pol(c, TP, (e & 0x4000) ? -1 : 1);
                corner = sc >= th && sc > 0;
                if (corner) s_score[(ly + 1) * SP + lx + 1] = (uint8_t)min(sc, 255);
            }
            const unsigned m = __ballot_sync(0xFFFFFFFFu, corner);
            if (lane == 0) s_wsum[warp] = __popc(m);
            __syncthreads();
            int ofs = ncorner + __popc(m & ((1u << lane) - 1));
#pragma unroll
            for (int w = 0; w < MCS_FAST_NT / 32; w++) {
                const int ws = s_wsum[w];
                if (w < warp) ofs += ws;
                ncorner += ws;
            }
            if (corner) s_queue[ofs] = (uint16_t)(e & 0x3FFF);
            __syncthreads();
        }
        // pass 3, corners only: strict-'>' 8-neighbour NMS inside the computed interior (scores outside count as 0)
        for (int q = threadIdx.x; q < ncorner; q += MCS_FAST_NT) {
            const int e = s_queue[q];
            const int ly = (e >> 7) & 0x7F, lx = e & 0x7F;
            const uint8_t* qp = &s_score[(ly + 1) * SP + lx + 1];
            const int sc = qp[0];
            const int n0 = max(max((int)qp[-SP - 1], (int)qp[-SP]), (int)qp[-SP + 1]);
            const int n1 = max(max((int)qp[SP - 1], (int)qp[SP]), (int)qp[SP + 1]);
            const int n2 = max((int)qp[-1], (int)qp[1]);
            if (sc > max(max(n0, n1), n2)) {
                const int slot = atomicAdd(&s_nsurv, 1);
                // keypoint in ROI coordinates: FAST x in the sub-image (lx + 3) shifted by j*wCell (:864-865)
                if (slot < P.fast_maxsurv) s_surv[slot] = mcs_oct::pack_cand(lx + xoff, ly + yoff, sc);
            }
        }
        __syncthreads();
        const int ns = min(s_nsurv, P.fast_maxsurv);
        if (ns > 0) {
            for (int base = 0; base < ns; base += MCS_FAST_NT) {
                const int t = base + threadIdx.x;
                const bool keep = t < ns;
                const unsigned m = __ballot_sync(0xFFFFFFFFu, keep);
                if (m) {
                    int p0 = 0;
                    if (lane == 0) p0 = atomicAdd(cnt, __popc(m));
                    p0 = __shfl_sync(0xFFFFFFFFu, p0, 0) + __popc(m & ((1u << lane) - 1));
                    if (keep && p0 < Lv.cand_cap) out[p0] = s_surv[t];
                }
            }
            return;  // block-uniform
        }
        if (P.min_th >= P.ini_th) return;  // the retry could only find a subset of nothing
        __syncthreads();
    }
}

// ------------------------------------------------------------------------------------------------------------
// K3s  per-cell FAST on STRIPS of cells.  One block = up to 5 horizontally adjacent cells of one cell row.  The computed
// interiors of neighbouring cells tile the level without overlap (cell j computes columns [iniX_j + 3, iniX_j + wCell + 3)), so the
// corner score of a pixel is the same whichever cell it is computed for; what is per cell is (a) the NMS rule "neighbours outside
// the cell's interior score 0" and (b) the iniTh -> minTh retry of a cell that returned nothing (:842-845).  The strip kernel
// therefore stages one tile for the whole strip (16-byte copies at the image's own alignment, no shifting), runs the pre-test and
// the arc strength once per pixel, and applies (a) with a cell-boundary mask in the NMS and (b) by re-scanning only the cells that
// stayed empty.  Against one block per cell this removes the 3 px halo columns between cells from the staging, most of the
// per-block prologue and every block-wide scan:
//   * each warp compacts the (pixel, polarity) entries of its own 32 pixel groups into a private stack in shared memory and
//     evaluates them 32 at a time as soon as 32 are pending (dense warps, no block barrier, no flag array);
//   * corners go to a block list through one shared atomic per warp batch; the NMS walks that list and appends survivors to the
//     level's candidate list with one global atomic per warp batch (the list is unordered; the octree kernel orders it).
// Tile x coordinates are image x - (iniX & ~15); the score tile uses the same coordinates with one zero row above and below.
// ------------------------------------------------------------------------------------------------------------
#define MCS_FS_NT 256
#define MCS_FS_TP 208  // tile and score pitch in bytes: 15 (alignment) + 187 (interior) + 6 (halo)
#define MCS_FS_WQ 288  // entries per warp stack: < 32 pending + 32 groups x 8
#define MCS_FS_MAXW 187

// Plain atomics as PTX: the callers are already warp-aggregated (one lane per batch), so nvcc's automatic same-address aggregation
// (VOTE / FLO / POPC / SHFL around every atomicAdd, 6.7 % of the kernel's instructions in the first ncu capture) is pure overhead.
__device__ __forceinline__ int fs_atom_add_shared(int* p, int v) {
    int old;
    asm volatile("atom.shared.add.u32 %0, [%1], %2;" : "=r"(old) : "r"((uint32_t)__cvta_generic_to_shared(p)), "r"(v) : "memory");
    return old;
}
__device__ __forceinline__ void fs_red_or_shared(unsigned* p, unsigned v) {
    asm volatile("red.shared.or.b32 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(p)), "r"(v) : "memory");
}
__device__ __forceinline__ int fs_atom_add_global(int* p, int v) {
    int old;
    asm volatile("atom.global.add.u32 %0, [%1], %2;" : "=r"(old) : "l"(p), "r"(v) : "memory");
    return old;
}

struct FastStripCtx {
    const uint8_t* tile;
    uint8_t* score;
    uint16_t* wq;
    uint16_t* corner;
    int* ncorner;
    int ih, lane;
    int cap;  // entries the corner list holds; the count keeps running past it
};

// one batch of the warp's stack: arc strength (= cv::FAST's score), corner <=> strength >= th
__device__ __forceinline__ void fs_eval(const FastStripCtx& X, bool active, int e, int th) {
    constexpr int TP = MCS_FS_TP;
    bool corner = false;
    if (active) {
        const int ly = (e >> 8) & 0x7F, tx = e & 0xFF;
        const int sc = fast_strength_pol(X.tile + (ly + 3) * TP + tx, TP, (e & 0x8000) ? -1 : 1);
        corner = sc >= th && sc > 0;
        if (corner) X.score[(ly + 1) * TP + tx] = (uint8_t)min(sc, 255);
    }
    const unsigned m = __ballot_sync(0xFFFFFFFFu, corner);
    if (m) {
        int base = 0;
        if (X.lane == 0) base = fs_atom_add_shared(X.ncorner, __popc(m));
        base = __shfl_sync(0xFFFFFFFFu, base, 0);
        const int at = base + __popc(m & ((1u << X.lane) - 1u));
        if (corner && at < X.cap) X.corner[at] = (uint16_t)(e & 0x7FFF);
    }
}

// pre-test + strength of the interior columns [xa, xb) (tile coordinates) of all ih rows at threshold th
__device__ __forceinline__ void fs_scan(const FastStripCtx& X, int th, int xa, int xb, uint32_t mg, int warp) {
    constexpr int TPW = MCS_FS_TP / 4;
    const int ga = xa >> 2, nG = ((xb - 1) >> 2) - ga + 1;
    const int ngroups = nG * X.ih;
    const uint32_t K1 = (uint32_t)(0x8000 - th - 1) * 0x00010001u;  // bit 15 of (a + K1 - b) per lane  <=>  a - b > th
    int pending = 0;
    for (int gb = warp * 32; gb < ngroups; gb += MCS_FS_NT) {
        const int gi = gb + X.lane;
        uint32_t f = 0, ebase = 0;
        if (gi < ngroups) {
            const int ly = (int)(((uint32_t)gi * mg) >> 20), g = ga + gi - ly * nG;
            const uint32_t* row = reinterpret_cast<const uint32_t*>(X.tile) + (ly + 3) * TPW + g;
            const uint32_t W0 = row[-1], W1 = row[0], W2 = row[1];
            const uint32_t Up = row[-3 * TPW], Dn = row[3 * TPW];
            const uint32_t Lf = __funnelshift_r(W0, W1, 8), Rt = __funnelshift_r(W1, W2, 24);
            const uint32_t A0 = row[-1 - 2 * TPW], A1 = row[-2 * TPW], A2 = row[1 - 2 * TPW];
            const uint32_t B0 = row[-1 + 2 * TPW], B1 = row[2 * TPW], B2 = row[1 + 2 * TPW];
            const uint32_t UL = __funnelshift_r(A0, A1, 16), UR = __funnelshift_r(A1, A2, 16);
            const uint32_t DL = __funnelshift_r(B0, B1, 16), DR = __funnelshift_r(B1, B2, 16);
#pragma unroll
            for (int hpair = 0; hpair < 2; hpair++) {
                const uint32_t sel = hpair ? 0x4342u : 0x4140u;
                const uint32_t c2 = __byte_perm(W1, 0, sel), u2 = __byte_perm(Up, 0, sel), d2 = __byte_perm(Dn, 0, sel);
                const uint32_t l2 = __byte_perm(Lf, 0, sel), r2 = __byte_perm(Rt, 0, sel);
                const uint32_t ul = __byte_perm(UL, 0, sel), dr = __byte_perm(DR, 0, sel);
                const uint32_t ur = __byte_perm(UR, 0, sel), dl = __byte_perm(DL, 0, sel);
                // the four opposite pairs (0,8) (4,12) (2,10) (6,14): a 9-arc holds at least one pixel of each
                const uint32_t bmin = __vimin3_u16x2(__vminu2(__vmaxu2(u2, d2), __vmaxu2(l2, r2)), __vmaxu2(ul, dr), __vmaxu2(ur, dl));
                const uint32_t dmax = __vimax3_u16x2(__vmaxu2(__vminu2(u2, d2), __vminu2(l2, r2)), __vminu2(ul, dr), __vminu2(ur, dl));
                const uint32_t Bf = bmin + K1 - c2, Df = c2 + K1 - dmax;
                const uint32_t t2 = ((Bf >> 15) & 0x00010001u) | ((Df >> 14) & 0x00020002u);
                f |= ((t2 | (t2 >> 14)) & 0xFu) << (4 * hpair);
            }
            // pixel e of the group is inside [xa, xb)  <=>  lo <= e < hi
            const int lo = max(xa - 4 * g, 0), hi = min(xb - 4 * g, 4);
            f &= ((1u << (2 * hi)) - 1u) & ~((1u << (2 * lo)) - 1u);
            ebase = (uint32_t)((ly << 8) | (4 * g));
        }
        const int cnt = __popc(f);
        int incl = cnt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xFFFFFFFFu, incl, o);
            if (X.lane >= o) incl += v;
        }
        const int total = __shfl_sync(0xFFFFFFFFu, incl, 31);
        if (total == 0) continue;  // warp-uniform
        int pos = pending + incl - cnt;
        if (f) {
#pragma unroll
            for (int e = 0; e < 4; e++) {  // predicated stores; one entry per pixel, bit 15 = dark
                const uint32_t fb = (f >> (2 * e)) & 3u;
                if (fb) {
                    X.wq[pos] = (uint16_t)(((fb & 2u) << 14) | (ebase + e));
                    pos++;
                }
            }
            const uint32_t both = f & (f >> 1) & 0x55u;  // rare: both pre-tests passed, the dark entry went in; add the bright one
            if (both) {
#pragma unroll
                for (int e = 0; e < 4; e++)
                    if (both & (1u << (2 * e))) {
                        X.wq[pos] = (uint16_t)(ebase + e);
                        pos++;
                    }
            }
        }
        pending += total;
        __syncwarp();
        while (pending >= 32) {
            pending -= 32;
            fs_eval(X, true, X.wq[pending + X.lane], th);
            __syncwarp();
        }
    }
    if (pending > 0) {
        const bool act = X.lane < pending;
        fs_eval(X, act, act ? X.wq[X.lane] : 0, th);
        __syncwarp();
    }
}

// The 4-pair necessary test on the four pixels of one tile word (bits 2e / 2e + 1 of the result: bright / dark pre-test of pixel e).
// c, up, dn: the words at rows 0, -3, +3; lf, rt: the pixels at x - 3, x + 3; ul .. dr: the diagonal ring pixels (+-2, +-2).
__device__ __forceinline__ uint32_t fs_pretest_word(uint32_t c, uint32_t up, uint32_t dn, uint32_t lf, uint32_t rt, uint32_t ul, uint32_t ur, uint32_t dl,
                                                    uint32_t dr, uint32_t K1) {
    uint32_t f = 0;
#pragma unroll
    for (int hpair = 0; hpair < 2; hpair++) {
        const uint32_t sel = hpair ? 0x4342u : 0x4140u;
        const uint32_t c2 = __byte_perm(c, 0, sel), u2 = __byte_perm(up, 0, sel), d2 = __byte_perm(dn, 0, sel);
        const uint32_t l2 = __byte_perm(lf, 0, sel), r2 = __byte_perm(rt, 0, sel);
        const uint32_t a2 = __byte_perm(ul, 0, sel), b2 = __byte_perm(dr, 0, sel);
        const uint32_t e2 = __byte_perm(ur, 0, sel), g2 = __byte_perm(dl, 0, sel);
        // the four opposite pairs (0,8) (4,12) (2,10) (6,14): a 9-arc holds at least one pixel of each
        const uint32_t bmin = __vimin3_u16x2(__vminu2(__vmaxu2(u2, d2), __vmaxu2(l2, r2)), __vmaxu2(a2, b2), __vmaxu2(e2, g2));
        const uint32_t dmax = __vimax3_u16x2(__vmaxu2(__vminu2(u2, d2), __vminu2(l2, r2)), __vminu2(a2, b2), __vminu2(e2, g2));
        const uint32_t Bf = bmin + K1 - c2, Df = c2 + K1 - dmax;
        const uint32_t t2 = ((Bf >> 15) & 0x00010001u) | ((Df >> 14) & 0x00020002u);
        f |= ((t2 | (t2 >> 14)) & 0xFu) << (4 * hpair);
    }
    return f;
}

#define MCS_FS_WQ8 544  // entries per warp stack of fs_scan8: < 32 pending + 32 units x 16

// fs_scan with EIGHT pixels (two adjacent tile words) per thread and step: the two words share 6 of their 22 loads and 2 of their 8
// funnel shifts, and the index arithmetic, the [xa, xb) mask and the warp scan of the compaction are paid once per 8 pixels.
// Units start at an even word, so the five word pairs of a unit are 64-bit loads (a warp's 32 units read 64 consecutive words in two
// wavefronts, as 32-bit loads of stride 2 they would conflict two-way).  mu = ceil(2^20 / units per row); units per row =
// ceil(words from the even word at or before xa to the word of xb - 1, / 2) — fs_units_per_row, also used by the host for the strip record.
__host__ __device__ __forceinline__ int fs_units_per_row(int xa, int xb) { return ((((xb - 1) >> 2) - ((xa >> 2) & ~1) + 1) + 1) >> 1; }

__device__ __forceinline__ void fs_scan8(const FastStripCtx& X, int th, int xa, int xb, uint32_t mu, int warp) {
    constexpr int TPW = MCS_FS_TP / 4;
    const int ga = (xa >> 2) & ~1, nU = fs_units_per_row(xa, xb);
    const int nunits = nU * X.ih;
    const uint32_t K1 = (uint32_t)(0x8000 - th - 1) * 0x00010001u;  // bit 15 of (a + K1 - b) per lane  <=>  a - b > th
    int pending = 0;
    for (int ub = warp * 32; ub < nunits; ub += MCS_FS_NT) {
        const int ui = ub + X.lane;
        uint32_t f = 0, ebase = 0;
        if (ui < nunits) {
            const int ly = (int)(((uint32_t)ui * mu) >> 20), g = ga + 2 * (ui - ly * nU);
            const uint32_t* row = reinterpret_cast<const uint32_t*>(X.tile) + (ly + 3) * TPW + g;  // 8-byte aligned: TPW and g are even
            const uint2 W = *reinterpret_cast<const uint2*>(row), A = *reinterpret_cast<const uint2*>(row - 2 * TPW),
                        B = *reinterpret_cast<const uint2*>(row + 2 * TPW), U = *reinterpret_cast<const uint2*>(row - 3 * TPW),
                        D = *reinterpret_cast<const uint2*>(row + 3 * TPW);
            const uint32_t UR1 = __funnelshift_r(A.x, A.y, 16), DR1 = __funnelshift_r(B.x, B.y, 16);  // = UL / DL of the second word
            {
                const uint32_t W0 = row[-1], A0 = row[-1 - 2 * TPW], B0 = row[-1 + 2 * TPW];
                f = fs_pretest_word(W.x, U.x, D.x, __funnelshift_r(W0, W.x, 8), __funnelshift_r(W.x, W.y, 24), __funnelshift_r(A0, A.x, 16), UR1,
                                    __funnelshift_r(B0, B.x, 16), DR1, K1);
            }
            {
                const uint32_t W3 = row[2], A3 = row[2 - 2 * TPW], B3 = row[2 + 2 * TPW];
                f |= fs_pretest_word(W.y, U.y, D.y, __funnelshift_r(W.x, W.y, 8), __funnelshift_r(W.y, W3, 24), UR1, __funnelshift_r(A.y, A3, 16), DR1,
                                     __funnelshift_r(B.y, B3, 16), K1)
                     << 8;
            }
            // pixel e of the unit is inside [xa, xb)  <=>  lo <= e < hi
            const int lo = max(xa - 4 * g, 0), hi = min(xb - 4 * g, 8);
            f &= ((1u << (2 * hi)) - 1u) & ~((1u << (2 * lo)) - 1u);
            ebase = (uint32_t)((ly << 8) | (4 * g));
        }
        const int cnt = __popc(f);
        int incl = cnt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xFFFFFFFFu, incl, o);
            if (X.lane >= o) incl += v;
        }
        const int total = __shfl_sync(0xFFFFFFFFu, incl, 31);
        if (total == 0) continue;  // warp-uniform
        int pos = pending + incl - cnt;
        if (f) {
#pragma unroll
            for (int e = 0; e < 8; e++) {  // predicated stores; one entry per pixel, bit 15 = dark
                const uint32_t fb = (f >> (2 * e)) & 3u;
                if (fb) {
                    X.wq[pos] = (uint16_t)(((fb & 2u) << 14) | (ebase + e));
                    pos++;
                }
            }
            const uint32_t both = f & (f >> 1) & 0x5555u;  // rare: both pre-tests passed, the dark entry went in; add the bright one
            if (both) {
#pragma unroll
                for (int e = 0; e < 8; e++)
                    if (both & (1u << (2 * e))) {
                        X.wq[pos] = (uint16_t)(ebase + e);
                        pos++;
                    }
            }
        }
        pending += total;
        __syncwarp();
        while (pending >= 32) {
            pending -= 32;
            fs_eval(X, true, X.wq[pending + X.lane], th);
            __syncwarp();
        }
    }
    if (pending > 0) {
        const bool act = X.lane < pending;
        fs_eval(X, act, act ? X.wq[X.lane] : 0, th);
        __syncwarp();
    }
}

#define MCS_FT_CORNERS 2048
struct FastTmaMaps {
    CUtensorMap map[MCS_PLAN_MAX_LEVELS];
    int box_rows[MCS_PLAN_MAX_LEVELS];
};

__device__ __forceinline__ uint32_t ft_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
// Bounded: a tile that never arrives (a bad tensor map) traps the kernel instead of hanging the device.
__device__ __forceinline__ void ft_mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t done = 0;
#pragma unroll 1
    for (int spin = 0; spin < (1 << 22); spin++) {
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}"
            : "=r"(done)
            : "r"(bar), "r"(parity)
            : "memory");
        if (done) return;
    }
    __trap();
}

#ifndef MCS_FS_MINB
#define MCS_FS_MINB 6
#endif
// TMA = true (the default): the tile comes as ONE cp.async.bulk.tensor.3d issued by thread 0 onto an mbarrier (box 208 x box_rows[level]
// of the level's (x, y, image) tensor map; rows of the box below the image arrive as zeros) instead of the per-thread LDG.128 + STS.128
// loop with its index division; the score tile is cleared while the tile is in flight.  img0 = image index of blockIdx.y = 0 in the
// tensor maps (which are encoded on image 0 of the workspace).
// PX8: the pre-test handles eight pixels per thread and step (fs_scan8) instead of four.
template <bool TMA, bool PX8>
__global__ void __launch_bounds__(MCS_FS_NT, MCS_FS_MINB) fast_strips_kernel(const __grid_constant__ McsPlan P, ImageSet S,
                                                                    const __grid_constant__ FastTmaMaps M, int img0,
                                                                    const uint4* __restrict__ strip_tab, uint32_t* __restrict__ cand,
                                                                    int* __restrict__ cand_count) {
    // no static shared memory: the dynamic block then starts at the (1 KB aligned) base of the CTA's window, which makes the tile a valid
    // TMA destination (128-byte aligned) without a run-time fix-up
    extern __shared__ __align__(128) unsigned char s_fs[];
    uint64_t& s_bar = *reinterpret_cast<uint64_t*>(s_fs + P.fs_ofs_ctl);
    int& s_ncorner = *reinterpret_cast<int*>(s_fs + P.fs_ofs_ctl + 8);
    unsigned& s_cellmask = *reinterpret_cast<unsigned*>(s_fs + P.fs_ofs_ctl + 12);  // bit c: cell c of the strip has a keypoint
    constexpr int TP = MCS_FS_TP;
    const int img = blockIdx.y;
    // host-built strip record (build_strip_table)
    const uint4 ra = __ldg(strip_tab + 2 * blockIdx.x), rb = __ldg(strip_tab + 2 * blockIdx.x + 1);
    const int level = (int)(ra.x & 0xFFu), ncells = (int)((ra.x >> 8) & 0xFFu), wcell = (int)(ra.x >> 16);
    const int iniX = (int)(ra.y & 0xFFFFu), iniY = (int)(ra.y >> 16);
    const int wint = (int)(ra.z & 0xFFFFu), ih = (int)(ra.z >> 16);  // computed interior of the strip
    const int xoff = (int)(ra.w & 0xFFFFu), yoff = (int)(ra.w >> 16);
    const uint32_t mg_groups = PX8 ? rb.w : rb.x, mg_quads = rb.y, mg_cell = rb.z;
    const McsLevel& Lv = P.L[level];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int ox = iniX & ~15;         // image x of tile byte 0
    const int x0 = iniX + 3 - ox;      // tile x of interior column 0 (3..18)
    const int ch = ih + 6;
    uint8_t* s_tile = s_fs;
    uint8_t* s_score = s_fs + P.fs_ofs_score;
    FastStripCtx X;
    X.tile = s_tile;
    X.score = s_score;
    X.wq = reinterpret_cast<uint16_t*>(s_fs + P.fs_ofs_wq) + warp * (PX8 ? MCS_FS_WQ8 : MCS_FS_WQ);
    X.corner = reinterpret_cast<uint16_t*>(s_fs + P.fs_ofs_corner);
    X.ncorner = &s_ncorner;
    X.ih = ih;
    X.lane = lane;
    X.cap = 0x7FFFFFFF;  // the list is sized for every pixel of the largest strip
    if constexpr (TMA) {
        if (threadIdx.x == 0) {
            const uint32_t bar = ft_smem_u32(&s_bar);
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(1) : "memory");
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(TP * M.box_rows[level]) : "memory");
            asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(
                             ft_smem_u32(s_tile)),
                         "l"(reinterpret_cast<uint64_t>(&M.map[level])), "r"(bar), "r"(ox), "r"(iniY), "r"(img0 + img)
                         : "memory");
            s_ncorner = 0;
            s_cellmask = 0;
        }
        for (int t = threadIdx.x; t < (ih + 2) * (TP / 16); t += MCS_FS_NT) reinterpret_cast<uint4*>(s_score)[t] = make_uint4(0, 0, 0, 0);
        __syncthreads();  // the barrier is initialised, the score tile is zero
        ft_mbar_wait(ft_smem_u32(&s_bar), 0);
    } else {   // stage the sub-image rows [iniY, iniY + ch) as 16-byte copies (rows are 16-byte aligned: make_set), zero the score tile
        int pitch;
        const uint8_t* src = level_ptr(P, S, img, level, pitch);
        const int nq = (x0 + wint + 3 + 15) >> 4;
        const uint8_t* g0 = src + (size_t)iniY * pitch + ox;
        for (int t = threadIdx.x; t < nq * ch; t += MCS_FS_NT) {
            const int r = (int)(((uint32_t)t * mg_quads) >> 20), c = t - r * nq;
            *reinterpret_cast<uint4*>(s_tile + r * TP + c * 16) = __ldg(reinterpret_cast<const uint4*>(g0 + (size_t)r * pitch) + c);
        }
        for (int t = threadIdx.x; t < (ih + 2) * (TP / 16); t += MCS_FS_NT) reinterpret_cast<uint4*>(s_score)[t] = make_uint4(0, 0, 0, 0);
        if (threadIdx.x == 0) {
            s_ncorner = 0;
            s_cellmask = 0;
        }
        __syncthreads();
    }
    uint32_t* out = cand + (size_t)img * P.cand_stride + Lv.cand_ofs;
    int* cnt = cand_count + img * P.nlevels + level;
    // strict-'>' 8-neighbour NMS over the corner list; a neighbour in another cell of the strip scores 0 for this pixel
    auto nms = [&]() {
        const int nc = s_ncorner;
        for (int q0 = 0; q0 < nc; q0 += MCS_FS_NT) {
            const int q = q0 + threadIdx.x;
            bool surv = false;
            uint32_t rec = 0;
            if (q < nc) {
                const int e = X.corner[q];
                const int ly = (e >> 8) & 0x7F, tx = e & 0xFF;
                const uint8_t* qp = &s_score[(ly + 1) * TP + tx];
                const int sc = qp[0];
                const int xi = tx - x0;
                const int ci = (int)(((uint32_t)xi * mg_cell) >> 20), xc = xi - ci * wcell;
                int n = max((int)qp[-TP], (int)qp[TP]);
                if (xc != 0) n = max(n, max(max((int)qp[-TP - 1], (int)qp[-1]), (int)qp[TP - 1]));
                if (xc != wcell - 1) n = max(n, max(max((int)qp[-TP + 1], (int)qp[1]), (int)qp[TP + 1]));
                surv = sc > n;
                if (surv) {
                    fs_red_or_shared(&s_cellmask, 1u << ci);
                    // keypoint in ROI coordinates: FAST x in the cell's sub-image shifted by j * wCell (:864-865)
                    rec = mcs_oct::pack_cand(xi + xoff, ly + yoff, sc);
                }
            }
            const unsigned m = __ballot_sync(0xFFFFFFFFu, surv);
            if (m) {
                int p0 = 0;
                if (lane == 0) p0 = fs_atom_add_global(cnt, __popc(m));
                p0 = __shfl_sync(0xFFFFFFFFu, p0, 0) + __popc(m & ((1u << lane) - 1u));
                if (surv && p0 < Lv.cand_cap) out[p0] = rec;
            }
        }
    };
    // FAST(cell, iniThFAST) for every cell of the strip
    if constexpr (PX8) fs_scan8(X, P.ini_th, x0, x0 + wint, mg_groups, warp);
    else fs_scan(X, P.ini_th, x0, x0 + wint, mg_groups, warp);
    __syncthreads();
    nms();
    if (P.min_th >= P.ini_th) return;  // the retry could only find a subset of nothing
    __syncthreads();
    // only for the cells that returned nothing: FAST(cell, minThFAST)  (:842-845)
    const unsigned failed = ~s_cellmask & ((1u << ncells) - 1u);
    if (!failed) return;  // block-uniform
    if (threadIdx.x == 0) s_ncorner = 0;
    __syncthreads();
    for (int c = 0; c < ncells; c++) {
        if (!(failed & (1u << c))) continue;
        const int xa = x0 + c * wcell, xb = min(xa + wcell, x0 + wint);
        const int nG = ((xb - 1) >> 2) - (xa >> 2) + 1;  // <= 19
        if constexpr (PX8) fs_scan8(X, P.min_th, xa, xb, c_magic20[fs_units_per_row(xa, xb)], warp);
        else fs_scan(X, P.min_th, xa, xb, c_magic20[nG], warp);
    }
    __syncthreads();
    nms();
}

// ------------------------------------------------------------------------------------------------------------
// K3c  FAST strips, persistent blocks with two TMA tile buffers.  NOT the default (MCS_FAST_TMA=2): measured 1.33-1.36 ms per 64 frames
// against 1.16 ms for fast_strips_kernel<true, true> — the loop-carried state at the 40-register cap makes ptxas re-materialise values
// inside the scan loops, and prefetching the next tile buys nothing because five other resident blocks already hide the load
// (profiles/README.md, round 2).  Kept for A/B runs.  Same arithmetic (4-pixel pre-test) and the same strip records as
// fast_strips_kernel; what changes is how a block gets its pixels and how long it lives:
//   * every (level) plane is described by a 3-D tensor map (x, y, image) with a 208 x box_rows[level] box; ONE elected thread
//     issues cp.async.bulk.tensor.3d for a whole strip tile onto an mbarrier — no per-thread LDG.128 + STS.128 staging loop, no index
//     division, out-of-image rows of the box arrive as zeros;
//   * a block takes work items (strip, image) — blockIdx.x first, then from a global counter (a fixed stride gives a block only
//     the strips of one residue class and measured 30 % slower than the plain kernel) — with two tile buffers: the tile of item
//     k + 1 is in flight while item k is scanned, so the load latency every short-lived block of fast_strips_kernel paid is hidden;
//   * the score tile is zeroed once per block; after each NMS only the bytes written for corners are cleared again.  The corner
//     list has a fixed capacity (MCS_FT_CORNERS); a strip with more corners than that (dense noise) runs its NMS as a scan of the
//     score tile instead — same survivors, the candidate list is unordered anyway.
// ------------------------------------------------------------------------------------------------------------
#ifndef MCS_FT_MINB
#define MCS_FT_MINB 6
#endif
__global__ void __launch_bounds__(MCS_FS_NT, MCS_FT_MINB) fast_strips_tma_kernel(const __grid_constant__ McsPlan P, const __grid_constant__ FastTmaMaps M,
                                                                        const uint4* __restrict__ strip_tab, int n_strips, int n_items, int img0,
                                                                        uint32_t* __restrict__ cand, int* __restrict__ cand_count,
                                                                        int* __restrict__ work_ctr) {
    // no static shared memory in this kernel: the dynamic block then starts at the (1 KB aligned) base of the CTA's window, so the TMA
    // destinations are 128-byte aligned without a run-time fix-up and every address below is base + a constant-bank offset
    extern __shared__ __align__(128) unsigned char s_ft[];
    constexpr int TP = MCS_FS_TP;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint8_t* s_score = s_ft + P.ft_ofs_score;
    const int score_quads = P.ft_score_bytes >> 4;
    uint64_t* s_bar = reinterpret_cast<uint64_t*>(s_ft + P.ft_ofs_ctl);         // [2] tile-arrived barriers
    int* s_ctl = reinterpret_cast<int*>(s_ft + P.ft_ofs_ctl + 16);
    int& s_ncorner = s_ctl[0];
    int* s_item = s_ctl + 2;              // [2][2] (work item, image) whose tile is in (or on its way to) buffer 0 / 1; item >= n_items: none left
    unsigned* s_cellmask3 = reinterpret_cast<unsigned*>(s_ctl + 6);  // [3] bit c: cell c of the strip has a keypoint; item k uses word k % 3 (loop head)
    FastStripCtx X;
    X.score = s_score;
    X.wq = reinterpret_cast<uint16_t*>(s_ft + P.ft_ofs_wq) + warp * MCS_FS_WQ;
    X.corner = reinterpret_cast<uint16_t*>(s_ft + P.ft_ofs_corner);
    X.ncorner = &s_ncorner;
    X.lane = lane;
    X.cap = MCS_FT_CORNERS;
    // one elected thread: tile of work item `w` -> buffer `b`
    auto issue = [&](int w, int b) {
        const int img = w / n_strips, s = w - img * n_strips;
        s_item[2 * b] = w;
        s_item[2 * b + 1] = img;
        const uint4 ra = __ldg(strip_tab + 2 * s);
        const int level = (int)(ra.x & 0xFFu), iniX = (int)(ra.y & 0xFFFFu), iniY = (int)(ra.y >> 16);
        const uint32_t bar = ft_smem_u32(&s_bar[b]), dst = ft_smem_u32(s_ft + b * P.ft_tile_bytes);
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(TP * M.box_rows[level]) : "memory");
        asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(dst),
                     "l"(reinterpret_cast<uint64_t>(&M.map[level])), "r"(bar), "r"(iniX & ~15), "r"(iniY), "r"(img0 + img)
                     : "memory");
    };
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(ft_smem_u32(&s_bar[0])), "r"(1) : "memory");
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(ft_smem_u32(&s_bar[1])), "r"(1) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        s_ncorner = 0;
        s_cellmask3[0] = 0;
        if ((int)blockIdx.x < n_items) issue(blockIdx.x, 0);
        else s_item[0] = n_items;
    }
    for (int t = threadIdx.x; t < score_quads; t += MCS_FS_NT) reinterpret_cast<uint4*>(s_score)[t] = make_uint4(0, 0, 0, 0);
    __syncthreads();
    int k3 = 0;
    for (int k = 0;; k++, k3 = k3 == 2 ? 0 : k3 + 1) {
        const int b = k & 1;
        const int2 wi = *reinterpret_cast<const int2*>(s_item + 2 * b);  // written by thread 0 at the head of item k - 1 (or before the first barrier); barriers since
        const int w = wi.x;
        if (w >= n_items) break;
        if (threadIdx.x == 0) {
            // The mask word of item k + 1 is cleared here, not at the end of item k: thread 0 may reach this point while other
            // warps still read the word of item k - 1 (the `failed` test below has no barrier after it), hence three words.
            s_cellmask3[k3 == 2 ? 0 : k3 + 1] = 0;
            const int wn = (int)gridDim.x + fs_atom_add_global(work_ctr, 1);
            if (wn < n_items) issue(wn, b ^ 1);  // tile buffer b ^ 1 was last read before the barrier that ended item k - 1
            else s_item[2 * (b ^ 1)] = n_items;
        }
        unsigned* s_cellmask = &s_cellmask3[k3];
        const int img = wi.y, s = w - img * n_strips;
        const uint4 ra = __ldg(strip_tab + 2 * s), rb = __ldg(strip_tab + 2 * s + 1);
        const int level = (int)(ra.x & 0xFFu), ncells = (int)((ra.x >> 8) & 0xFFu), wcell = (int)(ra.x >> 16);
        const int iniX = (int)(ra.y & 0xFFFFu);
        const int wint = (int)(ra.z & 0xFFFFu), ih = (int)(ra.z >> 16);
        const int xoff = (int)(ra.w & 0xFFFFu), yoff = (int)(ra.w >> 16);
        const uint32_t mg_groups = rb.x, mg_cell = rb.z;
        const McsLevel& Lv = P.L[level];
        const int x0 = (iniX & 15) + 3;
        X.tile = s_ft + b * P.ft_tile_bytes;
        X.ih = ih;
        uint32_t* out = cand + (size_t)img * P.cand_stride + Lv.cand_ofs;
        int* cnt = cand_count + img * P.nlevels + level;
        // NMS of one pixel (tile coordinates) whose score is sc > 0; a neighbour in another cell of the strip scores 0 for it
        auto nms_pixel = [&](int ly, int tx, int sc, uint32_t& rec) -> bool {
            const uint8_t* qp = &s_score[(ly + 1) * TP + tx];
            const int xi = tx - x0;
            const int ci = (int)(((uint32_t)xi * mg_cell) >> 20), xc = xi - ci * wcell;
            int n = max((int)qp[-TP], (int)qp[TP]);
            if (xc != 0) n = max(n, max(max((int)qp[-TP - 1], (int)qp[-1]), (int)qp[TP - 1]));
            if (xc != wcell - 1) n = max(n, max(max((int)qp[-TP + 1], (int)qp[1]), (int)qp[TP + 1]));
            if (sc <= n) return false;
            fs_red_or_shared(s_cellmask, 1u << ci);
            rec = mcs_oct::pack_cand(xi + xoff, ly + yoff, sc);  // ROI coordinates: FAST x in the cell's sub-image shifted by j * wCell (:864-865)
            return true;
        };
        auto append = [&](bool surv, uint32_t rec) {
            const unsigned m = __ballot_sync(0xFFFFFFFFu, surv);
            if (m) {
                int p0 = 0;
                if (lane == 0) p0 = fs_atom_add_global(cnt, __popc(m));
                p0 = __shfl_sync(0xFFFFFFFFu, p0, 0) + __popc(m & ((1u << lane) - 1u));
                if (surv && p0 < Lv.cand_cap) out[p0] = rec;
            }
        };
        // strict-'>' 8-neighbour NMS, then the score bytes and the corner count go back to zero for the next scan
        auto nms_and_clear = [&]() {
            const int nc = s_ncorner;
            if (nc <= MCS_FT_CORNERS) {
                for (int q0 = 0; q0 < nc; q0 += MCS_FS_NT) {
                    const int q = q0 + threadIdx.x;
                    bool surv = false;
                    uint32_t rec = 0;
                    if (q < nc) {
                        const int e = X.corner[q];
                        const int ly = (e >> 8) & 0x7F, tx = e & 0xFF;
                        surv = nms_pixel(ly, tx, s_score[(ly + 1) * TP + tx], rec);
                    }
                    append(surv, rec);
                }
                __syncthreads();
                if (threadIdx.x == 0) s_ncorner = 0;
                for (int q = threadIdx.x; q < nc; q += MCS_FS_NT) {
                    const int e = X.corner[q];
                    s_score[(((e >> 8) & 0x7F) + 1) * TP + (e & 0xFF)] = 0;
                }
            } else {  // more corners than the list holds: walk the score tile itself
                const int npix = ih * wint;
                for (int p0 = 0; p0 < npix; p0 += MCS_FS_NT) {
                    const int p = p0 + threadIdx.x;
                    bool surv = false;
                    uint32_t rec = 0;
                    if (p < npix) {
                        const int ly = p / wint, tx = x0 + p - ly * wint;
                        const int sc = s_score[(ly + 1) * TP + tx];
                        if (sc > 0) surv = nms_pixel(ly, tx, sc, rec);
                    }
                    append(surv, rec);
                }
                __syncthreads();
                if (threadIdx.x == 0) s_ncorner = 0;
                for (int t = threadIdx.x; t < score_quads; t += MCS_FS_NT) reinterpret_cast<uint4*>(s_score)[t] = make_uint4(0, 0, 0, 0);
            }
        };
        ft_mbar_wait(ft_smem_u32(&s_bar[b]), (uint32_t)(k >> 1) & 1u);
        // FAST(cell, iniThFAST) for every cell of the strip
        fs_scan(X, P.ini_th, x0, x0 + wint, mg_groups, warp);
        __syncthreads();
        nms_and_clear();
        __syncthreads();  // tile buffer b is free, the score tile is zero again
        const unsigned failed = ~*s_cellmask & ((1u << ncells) - 1u);
        if (P.min_th < P.ini_th && failed) {  // block-uniform.  Only for the cells that returned nothing: FAST(cell, minThFAST)  (:842-845)
            for (int c = 0; c < ncells; c++) {
                if (!(failed & (1u << c))) continue;
                const int xa = x0 + c * wcell, xb = min(xa + wcell, x0 + wint);
                const int nG = ((xb - 1) >> 2) - (xa >> 2) + 1;  // <= 19
                fs_scan(X, P.min_th, xa, xb, c_magic20[nG], warp);
            }
            __syncthreads();
            nms_and_clear();
            __syncthreads();
        }
    }
}

// ------------------------------------------------------------------------------------------------------------
// K4  DistributeOctTree.  One block per (level, image); algorithm in octree_core.h, executed with a CUDA block context.
// ------------------------------------------------------------------------------------------------------------
#ifndef MCS_OCT_NT
#define MCS_OCT_NT 128  // 128 beats 256 inside the chunked pipeline: more blocks per SM hide the barrier-heavy serial phases
#endif
#define MCS_OCT_NT_MAX 512  // the block size is a launch parameter (Engine::oct_nt): large quotas (config 4) use wider blocks

// GTAB = false: the node tables live in the block's shared memory (every configuration the reference ships).  GTAB = true: a level
// quota too large for that (one pyramid level holding ~2,150+ features) puts them in a per-block slice of global memory instead;
// the same algorithm, generic atomics, __syncthreads() orders the block's global accesses.
template <bool GTAB = false>
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
    // shared-memory tables (mcs_oct::Shared): plain PTX reductions skip nvcc's same-address aggregation wrapper
    __device__ __forceinline__ void atomic_add(int* p, int v) {
        if (GTAB) atomicAdd(p, v);
        else asm volatile("red.shared.add.u32 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(p)), "r"(v) : "memory");
    }
    __device__ __forceinline__ void atomic_max(unsigned* p, unsigned v) {
        if (GTAB) atomicMax(p, v);
        else asm volatile("red.shared.max.u32 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(p)), "r"(v) : "memory");
    }
    __device__ __forceinline__ void atomic_min(unsigned* p, unsigned v) {
        if (GTAB) atomicMin(p, v);
        else asm volatile("red.shared.min.u32 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(p)), "r"(v) : "memory");
    }
    __device__ __forceinline__ void atomic_max64(unsigned long long* p, unsigned long long v) {
        if (GTAB) atomicMax(p, v);
        else asm volatile("red.shared.max.u64 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(p)), "l"(v) : "memory");
    }
    // in-place exclusive scan of a[0..n) (in the tables' memory); collective; returns the total
    __device__ int exclusive_scan(int* a, int n) {
        __shared__ int s_ws[34];
        const int NT = blockDim.x, t = threadIdx.x, lane = t & 31, w = t >> 5, nwarps = NT >> 5;
        const int per = (n + NT - 1) >> (31 - __clz(NT));  // block sizes are powers of two
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
template <bool GTAB>
__global__ void __maxnreg__(48) octree_kernel(const __grid_constant__ McsPlan P, const uint32_t* __restrict__ cand,
                                                            const int* __restrict__ cand_count, uint16_t* __restrict__ knode,
                                                            uint32_t* __restrict__ sel, int* __restrict__ sel_count,
                                                            unsigned char* __restrict__ g_tab, size_t g_tab_stride) {
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
        sh.carve(GTAB ? g_tab + ((size_t)img * P.nlevels + level) * g_tab_stride : s_oct, op.cap);
        CudaBlockCtx<GTAB> ctx;
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
// 512 (x, y) pairs, see include/mcs_orb_pattern.inc.  Global (L1-cached), not __constant__: every lane reads a different
// 32-byte slice, which the constant cache would serialise 32 ways.
__device__ __align__(16) signed char g_pattern[1024];
// un-steered pattern for describe_fused_kernel: word 8 * lane + slot = ofs0 | ofs1 << 10 | bit << 20 with ofs = y * 28 + x + 377, the
// byte offset into a 28-byte-pitch blurred patch whose centre is (13, 13); bit = which of the lane's 8 tests the slot holds
// (filled by the host next to g_pattern, slot order chosen against shared-memory bank conflicts)
__device__ __align__(16) uint32_t g_pattern_ofs28[256];

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
    // 16 points = 8 tests = 1 descriptor byte per lane; test t = bytes (x0, y0, x1, y1) of word t
    const uint4 pw0 = __ldg(reinterpret_cast<const uint4*>(g_pattern) + lane * 2);
    const uint4 pw1 = __ldg(reinterpret_cast<const uint4*>(g_pattern) + lane * 2 + 1);
    const uint32_t pw[8] = {pw0.x, pw0.y, pw0.z, pw0.w, pw1.x, pw1.y, pw1.z, pw1.w};
    if (!P.use_angle) {
#pragma unroll
        for (int t = 0; t < 8; t++) {
            const int x0 = (int)(signed char)(pw[t]), y0 = (int)(signed char)(pw[t] >> 8);
            const int x1 = (int)(signed char)(pw[t] >> 16), y1 = (int)(signed char)(pw[t] >> 24);
            const int t0 = bl[y0 * bp + x0];
            const int t1 = bl[y1 * bp + x1];
            val |= (t0 < t1) << t;
        }
    } else {
        const float ang = __fmul_rn(angle, (float)(3.14159265358979323846 / 180.f));  // factorPI (:105)
        const float a = (float)cos((double)ang), b = (float)sin((double)ang);
#pragma unroll
        for (int t = 0; t < 8; t++) {
            const float x0 = (float)(signed char)(pw[t]), y0 = (float)(signed char)(pw[t] >> 8);
            const float x1 = (float)(signed char)(pw[t] >> 16), y1 = (float)(signed char)(pw[t] >> 24);
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

// ------------------------------------------------------------------------------------------------------------
// K2+K5 fused (shipped mode, angle 0): the un-steered pattern only reads blurred pixels at offsets -13..12 around a keypoint, and
// keypoints sit >= 19 px inside the level, so the blurred 26 x 26 patch depends on the 32 x 32 input patch alone (no border
// handling).  One warp per keypoint stages that patch (pre-shifted so that patch column 0 is byte 0 of a shared row), runs the
// same bit-exact separable 7-tap filter as blur_kernel (DP4A rows, DP2A columns on vertical pairs) and samples the 256 tests
// from shared memory: the blurred pyramid is neither written nor read (blur_kernel remains for the steered mode and observers).
// ------------------------------------------------------------------------------------------------------------
#ifndef MCS_DESCF_WARPS
#define MCS_DESCF_WARPS 8
#endif
#ifndef MCS_DESCF_KPW
#define MCS_DESCF_KPW 4  // keypoints per warp
#endif
__global__ void __launch_bounds__(MCS_DESCF_WARPS * 32, 48 / MCS_DESCF_WARPS) describe_fused_kernel(const __grid_constant__ McsPlan P, ImageSet S,
                                                                              const uint32_t* __restrict__ sel,
                                                                              const int* __restrict__ sel_count,
                                                                              KeypointPod* __restrict__ kps, uint8_t* __restrict__ desc,
                                                                              int* __restrict__ kp_count) {
    // input patch rows -16..15, columns -16..15; row pitch 12 words so that the 4 row pairs x 8 word columns one H-pass step reads
    // fall into 32 different banks.  The blurred patch (26 x 28 bytes, rows/columns -13..12) reuses the space: the input is dead
    // once the H pass has finished.
    __shared__ __align__(16) uint32_t s_in[MCS_DESCF_WARPS][32][12];
    __shared__ __align__(16) uint32_t s_pair[MCS_DESCF_WARPS][16][28];  // row-filtered patch as vertical pairs (row 2p | row 2p+1 << 16)
    uint8_t* const s_blw = reinterpret_cast<uint8_t*>(&s_in[threadIdx.x >> 5][0][0]);
    const int img = blockIdx.y;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // lane l holds level l's selected count and the running total up to it: loaded and scanned once per warp, then every one of the
    // warp's MCS_DESCF_KPW keypoints finds its level with one ballot (the load + scan used to be paid per keypoint: 6 % of the
    // instructions and 30 % of the stall samples of the one-keypoint-per-warp form)
    const int cnt_l = lane < P.nlevels ? max(__ldg(&sel_count[img * P.nlevels + lane]), 0) : 0;
    int incl = cnt_l;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xFFFFFFFFu, incl, o);
        if (lane >= o) incl += v;
    }
    {
        const int total = __shfl_sync(0xFFFFFFFFu, incl, 31);
        if (blockIdx.x == 0 && threadIdx.x == 0) kp_count[img] = min(total, P.kp_cap);
    }
    const int excl = incl - cnt_l;
#pragma unroll 1
    for (int it = 0; it < MCS_DESCF_KPW; it++) {
    const int k = (blockIdx.x * MCS_DESCF_KPW + it) * MCS_DESCF_WARPS + warp;  // output slot: the block's warps take consecutive slots
    const unsigned lm = __ballot_sync(0xFFFFFFFFu, lane < P.nlevels && k < incl);  // first level whose running total exceeds k
    if (k >= P.kp_cap || lm == 0) break;  // past the last keypoint; warp-uniform, and slots only grow with `it`
    const int level = __ffs(lm) - 1;
    const int before = __shfl_sync(0xFFFFFFFFu, excl, level);
    __syncwarp();  // the previous keypoint's descriptor reads of s_blw are done
    const McsLevel& Lv = P.L[level];
    const uint32_t rec = sel[(size_t)img * P.sel_stride + Lv.sel_ofs + (k - before)];
    const int px = mcs_oct::cand_x(rec) + MCS_MIN_BORDER, py = mcs_oct::cand_y(rec) + MCS_MIN_BORDER;  // :883-884
    int pitch;
    const uint8_t* im = level_ptr(P, S, img, level, pitch);
    // stage: shared word (r, c) = image bytes (px - 16 + 4c) .. +3 of row py - 16 + r
    {
        const int gb = px - 16, aw = gb >> 2, sh = (gb & 3) * 8;
        const uint32_t* g0 = reinterpret_cast<const uint32_t*>(im + (size_t)(py - 16) * pitch) + aw;
        const int pw = pitch >> 2;
        uint32_t lo[8], hi[8];
#pragma unroll
        for (int i = 0; i < 8; i++) {
            const int idx = lane + 32 * i, r = idx >> 3, c = idx & 7;
            const uint32_t* g = g0 + r * pw + c;
            lo[i] = __ldg(g);
            hi[i] = __ldg(g + 1);
        }
#pragma unroll
        for (int i = 0; i < 8; i++) {
            const int idx = lane + 32 * i;
            s_in[warp][idx >> 3][idx & 7] = __funnelshift_r(lo[i], hi[i], sh);
        }
    }
    __syncwarp();
    const uint32_t wA = (uint32_t)c_gauss7[0] | ((uint32_t)c_gauss7[1] << 8) | ((uint32_t)c_gauss7[2] << 16) | ((uint32_t)c_gauss7[3] << 24);
    const uint32_t wB = (uint32_t)c_gauss7[4] | ((uint32_t)c_gauss7[5] << 8) | ((uint32_t)c_gauss7[6] << 16);
    // rows: output column j (0..25, patch column j - 13) = taps on bytes j .. j+6; item = (row pair p, 4 columns g)
    for (int pb = 0; pb < 16; pb += 4) {
        const int p = pb + (lane >> 3), g = lane & 7;  // lane = (row pair, word column); column 7 idles
        if (g == 7) continue;
        uint32_t hv[2][4];
#pragma unroll
        for (int rr = 0; rr < 2; rr++) {
            const uint32_t* row = s_in[warp][2 * p + rr];
            const uint32_t W0 = row[g], W1 = row[g + 1], W2 = g < 6 ? row[g + 2] : 0u;  // columns 26, 27 are padding
            hv[rr][0] = __dp4a(W1, wB, __dp4a(W0, wA, 0u));
            hv[rr][1] = __dp4a(__funnelshift_r(W1, W2, 8), wB, __dp4a(__funnelshift_r(W0, W1, 8), wA, 0u));
            hv[rr][2] = __dp4a(__funnelshift_r(W1, W2, 16), wB, __dp4a(__funnelshift_r(W0, W1, 16), wA, 0u));
            hv[rr][3] = __dp4a(__funnelshift_r(W1, W2, 24), wB, __dp4a(__funnelshift_r(W0, W1, 24), wA, 0u));
        }
        uint4 o;
        o.x = hv[0][0] | (hv[1][0] << 16);
        o.y = hv[0][1] | (hv[1][1] << 16);
        o.z = hv[0][2] | (hv[1][2] << 16);
        o.w = hv[0][3] | (hv[1][3] << 16);
        *reinterpret_cast<uint4*>(&s_pair[warp][p][4 * g]) = o;
    }
    __syncwarp();
    {
        const uint32_t k0 = c_gauss7[0], k1 = c_gauss7[1], k2 = c_gauss7[2], k3 = c_gauss7[3], k4 = c_gauss7[4], k5 = c_gauss7[5], k6 = c_gauss7[6];
        const uint32_t e0 = k0 | (k1 << 8), e1 = k2 | (k3 << 8), e2 = k4 | (k5 << 8), e3 = k6;       // even output row 2q
        const uint32_t o0 = k0 << 8, o1 = k1 | (k2 << 8), o2 = k3 | (k4 << 8), o3 = k5 | (k6 << 8);  // odd output row 2q + 1
        // columns: output row y (0..25, patch row y - 13) = taps on filtered rows y .. y+6; item = (output row pair q, 4 columns g)
        // lane = (segment of output row pairs, 4 columns g): segments {0-3, 4-6, 7-9, 10-12} walk down with a 4-row-pair window in
        // registers, so every filtered row pair is loaded once or twice instead of four times
        const int seg = lane >> 3, g = lane & 7;
        const int q0 = seg == 0 ? 0 : 1 + 3 * seg, q1 = 4 + 3 * seg;
        if (g < 7) {
            uint4 a = *reinterpret_cast<const uint4*>(&s_pair[warp][q0][4 * g]);
            uint4 b = *reinterpret_cast<const uint4*>(&s_pair[warp][q0 + 1][4 * g]);
            uint4 c = *reinterpret_cast<const uint4*>(&s_pair[warp][q0 + 2][4 * g]);
#pragma unroll
            for (int it = 0; it < 4; it++) {  // 3 or 4 row pairs per segment; unrolled so that the window rotates by renaming
                const int q = q0 + it;
                if (q >= q1) break;
                const uint4 d = *reinterpret_cast<const uint4*>(&s_pair[warp][q + 3][4 * g]);
                const uint32_t pa[4] = {a.x, a.y, a.z, a.w}, pb[4] = {b.x, b.y, b.z, b.w}, pc[4] = {c.x, c.y, c.z, c.w}, pd[4] = {d.x, d.y, d.z, d.w};
                uint32_t oe = 0, oo = 0;
#pragma unroll
                for (int kk = 0; kk < 4; kk++) {
                    const uint32_t ve = __dp2a_lo(pd[kk], e3, __dp2a_lo(pc[kk], e2, __dp2a_lo(pb[kk], e1, __dp2a_lo(pa[kk], e0, 32768u))));
                    const uint32_t vo = __dp2a_lo(pd[kk], o3, __dp2a_lo(pc[kk], o2, __dp2a_lo(pb[kk], o1, __dp2a_lo(pa[kk], o0, 32768u))));
                    oe |= (ve >> 16) << (8 * kk);
                    oo |= (vo >> 16) << (8 * kk);
                }
                *reinterpret_cast<uint32_t*>(s_blw + (2 * q) * 28 + 4 * g) = oe;
                *reinterpret_cast<uint32_t*>(s_blw + (2 * q + 1) * 28 + 4 * g) = oo;
                a = b; b = c; c = d;
            }
        }
    }
    __syncwarp();
    // 16 points = 8 tests = 1 descriptor byte per lane (computeOrbDescriptor :106-145 with a = 1, b = 0).  Slot s of lane i is one of
    // the tests 8i .. 8i+7, in an order the host chose so that the 32 lanes of a slot hit few common banks; each word holds the two
    // byte offsets into the blurred patch and the bit the result goes to.
    const uint4 pw0 = __ldg(reinterpret_cast<const uint4*>(g_pattern_ofs28) + lane * 2);  // (L1-resident; kept out of the loop-carried registers)
    const uint4 pw1 = __ldg(reinterpret_cast<const uint4*>(g_pattern_ofs28) + lane * 2 + 1);
    const uint32_t pw[8] = {pw0.x, pw0.y, pw0.z, pw0.w, pw1.x, pw1.y, pw1.z, pw1.w};
    int val = 0;
#pragma unroll
    for (int t = 0; t < 8; t++) val |= ((int)s_blw[pw[t] & 0x3FFu] < (int)s_blw[(pw[t] >> 10) & 0x3FFu]) << (pw[t] >> 20);
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
        kp.angle = 0.0f;  // computeOrientation writes 0.0 (:475)
        kp.response = (float)mcs_oct::cand_score(rec);
        kp.octave = level;
        kp.class_id = -1;
        kps[slot] = kp;
    }
    }  // keypoints of this warp
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
    CudaBlockCtx<false> ctx;
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
