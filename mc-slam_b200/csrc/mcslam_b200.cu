// mcslam_b200.cu — host engine + C ABI (include/mcslam_b200.h) over the sm_100a kernels in extract_kernels.cuh and
// match_kernels.cuh.  No CPU fallback: every stage is a kernel launch; creation fails without a CUDA device.
#include <sched.h>
#include <cuda.h>
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/mcs_orb_pattern.inc"
#include "../../include/mcslam_b200.h"
#include "extract_kernels.cuh"
#include "match_kernels.cuh"

namespace {

thread_local std::string g_err = "";
void set_err(const std::string& s) { g_err = s; }

#define CK(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) {                                                                   \
            set_err(std::string(#call) + ": " + cudaGetErrorString(e_) + " (" + __FILE__ + ":" + std::to_string(__LINE__) + ")"); \
            return MCS_ERR_CUDA;                                                                   \
        }                                                                                          \
    } while (0)

inline int cv_round_f(float v) { return (int)std::nearbyintf(v); }
inline int cv_floor_f(float v) {
    int i = (int)v;
    return i - (i > v);
}
inline short sat_short(float v) {
    int i = cv_round_f(v);
    return (short)std::min(std::max(i, -32768), 32767);
}
inline int align_up(int v, int a) { return (v + a - 1) / a * a; }

// ORBextractor constructor tables (ORBextractor.cpp:408-468)
struct OrbTables {
    int nfeatures = 0, nlevels = 0, ini_th = 0, min_th = 0;
    float scale_factor_f = 0;
    std::vector<float> scale, inv_scale, sigma2, inv_sigma2;
    std::vector<int> quota;
    int umax[16];
    void build(int nf, float sf, int nl, int ini, int mn) {
        nfeatures = nf; nlevels = nl; ini_th = ini; min_th = mn; scale_factor_f = sf;
        const double scaleFactor = sf;  // the member is a double initialised from the float argument (ORBextractor.h:103)
        scale.assign(nl, 1.f); sigma2.assign(nl, 1.f); inv_scale.assign(nl, 1.f); inv_sigma2.assign(nl, 1.f); quota.assign(nl, 0);
        for (int i = 1; i < nl; i++) {
            scale[i] = (float)(scale[i - 1] * scaleFactor);
            sigma2[i] = scale[i] * scale[i];
        }
        for (int i = 0; i < nl; i++) {
            inv_scale[i] = 1.0f / scale[i];
            inv_sigma2[i] = 1.0f / sigma2[i];
        }
        const float factor = (float)(1.0f / scaleFactor);
        float nDesired = nf * (1 - factor) / (1 - (float)std::pow((double)factor, (double)nl));
        int sum = 0;
        for (int l = 0; l < nl - 1; l++) {
            quota[l] = cv_round_f(nDesired);
            sum += quota[l];
            nDesired *= factor;
        }
        quota[nl - 1] = std::max(nf - sum, 0);
        const int HP = 15;
        int v, v0, vmax = cv_floor_f(HP * std::sqrt(2.f) / 2 + 1);
        int vmin = (int)std::ceil(HP * std::sqrt(2.f) / 2);
        const double hp2 = HP * HP;
        for (v = 0; v <= vmax; ++v) umax[v] = (int)std::nearbyint(std::sqrt(hp2 - v * v));
        for (v = HP, v0 = 0; v >= vmin; --v) {
            while (umax[v0] == umax[v0 + 1]) ++v0;
            umax[v] = v0;
            ++v0;
        }
    }
};

// cv::resize INTER_LINEAR 8U coefficient tables for one axis (OpenCV imgproc/resize.cpp semantics)
void build_resize_axis(int ssize, int dsize, std::vector<int2>& tab, bool is_x) {
    const double inv_scale = (double)dsize / ssize, scl = 1. / inv_scale;
    for (int d = 0; d < dsize; d++) {
        float f = (float)((d + 0.5) * scl - 0.5);
        int s = cv_floor_f(f);
        f -= s;
        if (is_x) {
            if (s < 0) { f = 0; s = 0; }
            if (s >= ssize - 1) { f = 0; s = ssize - 1; }
        }
        const short c0 = sat_short((1.f - f) * 2048.f), c1 = sat_short(f * 2048.f);
        int2 e;
        e.x = s;
        e.y = (int)((uint32_t)(uint16_t)c0 | ((uint32_t)(uint16_t)c1 << 16));
        tab.push_back(e);
    }
}

// getGaussianKernelBitExact + getGaussianKernelFixedPoint_ED (OpenCV imgproc/smooth.dispatch.cpp), 8 fractional bits
void gauss_kernel_fixed(int n, double sigma, int* out) {
    const int n2 = (n - 1) / 2;
    std::vector<double> v(n2 + 1), k(n);
    const double scale2x = -0.125 / (sigma * sigma);
    double sum = 0;
    for (int i = 0, x = 1 - n; i < n2; i++, x += 2) {
        v[i] = std::exp((double)(x * x) * scale2x);
        sum += v[i];
    }
    sum = sum * 2 + 1.0;
    const double mul1 = 1.0 / sum;
    for (int i = 0; i < n2; i++) k[i] = k[n - 1 - i] = v[i] * mul1;
    k[n2] = mul1;
    double err = 0;
    long long s = 0;
    for (int i = 0; i < n2; i++) {
        const double adj = k[i] * 256.0 + err;
        const long long v0 = (long long)std::nearbyint(adj);
        err = adj - (double)v0;
        out[i] = out[n - 1 - i] = (int)v0;
        s += v0;
    }
    out[n2] = (int)(256 - 2 * s);
}

// Builds the plan for (width x height); returns false (with g_err) when unsupported.
bool build_plan(const OrbTables& T, int width, int height, int use_angle, McsPlan& P, std::vector<int2>& xtab, std::vector<int2>& ytab) {
    memset(&P, 0, sizeof(P));
    xtab.clear();
    ytab.clear();
    if (T.nlevels < 1 || T.nlevels > MCS_PLAN_MAX_LEVELS) { set_err("nlevels out of range [1,16]"); return false; }
    if (width < 1 || height < 1 || width > 4096 || height > 4096) { set_err("image size out of range [1,4096]"); return false; }
    if (T.ini_th < 1 || T.min_th < 1 || T.ini_th > 254 || T.min_th > 254) { set_err("FAST thresholds must be in [1,254]"); return false; }
    P.nlevels = T.nlevels; P.width = width; P.height = height; P.nfeatures = T.nfeatures;
    P.ini_th = T.ini_th; P.min_th = T.min_th; P.use_angle = use_angle;
    for (int i = 0; i < 16; i++) P.umax[i] = T.umax[i];
    int plane = 0, cells = 0, cand = 0, sel = 0, tiles = 0, octmax = 0;
    for (int l = 0; l < T.nlevels; l++) {
        McsLevel& L = P.L[l];
        L.w = cv_round_f((float)width * T.inv_scale[l]);   // ORBextractor.cpp:1178
        L.h = cv_round_f((float)height * T.inv_scale[l]);
        if (L.w < 1 || L.h < 1) { set_err("pyramid level collapses to zero size"); return false; }
        L.pitch = align_up(L.w, 64);
        L.plane_ofs = plane;
        plane += align_up(L.pitch * L.h, 256);
        L.scale = T.scale[l];
        L.quota = T.quota[l];
        L.scaled_patch = (int)(31 * T.scale[l]);            // :879
        L.roi_w = L.w - 2 * MCS_MIN_BORDER;                 // :787-796
        L.roi_h = L.h - 2 * MCS_MIN_BORDER;
        const float W = 35;
        L.n_cols = L.roi_w > 0 ? (int)((float)L.roi_w / W) : 0;
        L.n_rows = L.roi_h > 0 ? (int)((float)L.roi_h / W) : 0;
        if (L.n_cols < 1 || L.n_rows < 1) {
            // the reference divides by zero here (:801-802); a level too small for one cell yields no keypoints
            L.n_cols = L.n_rows = 0; L.w_cell = L.h_cell = 1;
        } else {
            L.w_cell = (int)std::ceil((float)L.roi_w / L.n_cols);
            L.h_cell = (int)std::ceil((float)L.roi_h / L.n_rows);
            if (L.w_cell > 69 || L.h_cell > 69) { set_err("internal: FAST cell larger than 69 px"); return false; }
        }
        L.cell_base = cells;
        cells += L.n_cols * L.n_rows;
        L.cand_ofs = cand;
        L.cand_cap = L.n_cols * L.n_rows * ((L.w_cell + 1) / 2) * ((L.h_cell + 1) / 2);  // strict 8-neighbour maxima per cell
        cand += align_up(std::max(L.cand_cap, 4), 4);
        L.n_ini = L.n_cols ? std::max(1, (int)std::round((float)L.roi_w / L.roi_h)) : 1;  // :558
        L.hX = (float)L.roi_w / L.n_ini;
        L.oct_cap = std::max(L.quota, 4 * L.n_ini) + 4;
        if (L.oct_cap > 60000) { set_err("features per level too large"); return false; }
        octmax = std::max(octmax, L.oct_cap);
        L.sel_ofs = sel;
        sel += L.oct_cap;
        if (l > 0) {
            if (xtab.size() & 1) xtab.push_back(make_int2(0, 0));  // 16-byte aligned rows of the table (128-bit loads)
            L.xtab_ofs = (int)xtab.size();
            build_resize_axis(P.L[l - 1].w, L.w, xtab, true);
            for (int k = 0; k < 4; k++) xtab.push_back(xtab.back());  // padding read by the last thread of a row
            L.ytab_ofs = (int)ytab.size();
            build_resize_axis(P.L[l - 1].h, L.h, ytab, false);
        }
        L.tiles_x = (L.w + MCS_BLUR_TW - 1) / MCS_BLUR_TW;
        L.tiles_y = (L.h + MCS_BLUR_TH - 1) / MCS_BLUR_TH;
        L.tile_base = tiles;
        tiles += L.tiles_x * L.tiles_y;
    }
    {   // FAST shared-memory carve-up for the largest cell
        int cwm = 7, chm = 7;
        for (int l = 0; l < T.nlevels; l++)
            if (P.L[l].n_cols > 0) { cwm = std::max(cwm, P.L[l].w_cell + 6); chm = std::max(chm, P.L[l].h_cell + 6); }
        const int iwm = cwm - 6, ihm = chm - 6, G = (iwm + 3) / 4;
        P.fast_tpw = MCS_FAST_TPW;  // compile-time pitches; only the row counts follow the plan
        P.fast_sp = MCS_FAST_SP;
        if (G + 2 > MCS_FAST_TPW || iwm + 2 > MCS_FAST_SP) { set_err("internal: FAST cell wider than the tile pitch"); return false; }
        int o = align_up(P.fast_tpw * 4 * chm, 16);
        P.fast_ofs_score = o; o += align_up(P.fast_sp * (ihm + 2), 16);
        P.fast_ofs_flags = o; o += align_up(G * ihm, 16);
        P.fast_ofs_queue = o; o += align_up(2 * iwm * ihm * 2, 16);
        P.fast_maxsurv = ((iwm + 1) / 2) * ((ihm + 1) / 2);
        P.fast_ofs_surv = o; o += P.fast_maxsurv * 4;
        P.fast_smem_bytes = o;
    }
    P.n_cells = cells; P.n_tiles = tiles; P.pyr_stride = align_up(plane, 256); P.cand_stride = cand; P.sel_stride = sel;
    P.kp_cap = align_up(sel, 8);
    P.oct_cap_max = octmax;
    if (xtab.empty()) { xtab.push_back(make_int2(0, 0)); ytab.push_back(make_int2(0, 0)); }
    return true;
}


// Selector tables of resize_level_sel_kernel for one level (from the cv::resize tables of build_resize_axis).  Returns false when
// the level cannot use that kernel (4 output pixels spanning more than 7 source bytes, or coefficients outside the range that keeps
// the result in 0..255 without saturation).
bool build_resize_word_tables(const McsLevel& D, const McsLevel& Sv, const std::vector<int2>& xtab, const std::vector<int2>& ytab,
                              std::vector<int4>& xw, std::vector<int4>& yw) {
    auto coef_ok = [](int packed) {
        const int a0 = (short)(packed & 0xFFFF), a1 = packed >> 16;
        return a0 >= 0 && a1 >= 0 && a0 + a1 <= 2049;
    };
    const int dwords = (D.w + 3) / 4;
    for (int g = 0; g < dwords; g++) {
        const int2* e = &xtab[D.xtab_ofs + 4 * g];  // padded by 4 entries past the row end
        const int sx0 = e[0].x;
        uint32_t sel = 0;
        for (int k = 0; k < 4; k++) {
            const int o = e[k].x - sx0;
            if (o < 0 || o > 6 || e[k].x < 0 || !coef_ok(e[k].y)) return false;
            sel |= (uint32_t)(o | ((o + 1) << 4)) << (8 * k);
        }
        xw.push_back(make_int4(sx0 >> 2, (int)sel, e[0].y, e[1].y));
        xw.push_back(make_int4(e[2].y, e[3].y, (sx0 & 3) * 8, 0));
    }
    for (int dy = 0; dy < D.h; dy++) {
        const int2 y = ytab[D.ytab_ofs + dy];
        if (!coef_ok(y.y)) return false;
        const int sy0 = std::min(std::max(y.x, 0), Sv.h - 1), sy1 = std::min(std::max(y.x + 1, 0), Sv.h - 1);
        yw.push_back(make_int4(sy0, sy1, (short)(y.y & 0xFFFF), y.y >> 16));
    }
    return true;
}

// cv::initUndistortRectifyMap(K, dist, I, K', CV_16SC2) as cv::undistort runs it per stripe of max(1, 4096 / cols) rows with
// K'.cy = cy - y0 (OpenCV imgproc undistort): f64 arithmetic in OpenCV's operation order (3x3 inverse by cofactors, _x / _y / _w
// accumulated along the row), coordinates * 32 rounded half to even.  out[i] = {sx | sy << 16 (int16 each), fy * 32 + fx}.
void build_undistort_map(const mcs_calib& c, int w, int h, uint2* out) {
    const double k1 = c.dist[0], k2 = c.dist[1], p1 = c.dist[2], p2 = c.dist[3], k3 = c.dist[4];
    const double fx = c.K[0], fy = c.K[4], u0 = c.K[2], v0 = c.K[5];
    const int stripe0 = std::min(std::max(1, (1 << 12) / std::max(w, 1)), h);
    for (int y0 = 0; y0 < h; y0 += stripe0) {
        const int ss = std::min(stripe0, h - y0);
        double A[9];
        for (int i = 0; i < 9; i++) A[i] = c.K[i];
        A[5] = v0 - y0;
        const double det = A[0] * (A[4] * A[8] - A[5] * A[7]) - A[1] * (A[3] * A[8] - A[5] * A[6]) + A[2] * (A[3] * A[7] - A[4] * A[6]);
        const double d = 1. / det;
        const double ir[9] = {(A[4] * A[8] - A[5] * A[7]) * d, (A[2] * A[7] - A[1] * A[8]) * d, (A[1] * A[5] - A[2] * A[4]) * d,
                              (A[5] * A[6] - A[3] * A[8]) * d, (A[0] * A[8] - A[2] * A[6]) * d, (A[2] * A[3] - A[0] * A[5]) * d,
                              (A[3] * A[7] - A[4] * A[6]) * d, (A[1] * A[6] - A[0] * A[7]) * d, (A[0] * A[4] - A[1] * A[3]) * d};
        for (int i = 0; i < ss; i++) {
            double _x = i * ir[1] + ir[2], _y = i * ir[4] + ir[5], _w = i * ir[7] + ir[8];
            for (int j = 0; j < w; j++, _x += ir[0], _y += ir[3], _w += ir[6]) {
                const double iw = 1. / _w, x = _x * iw, y = _y * iw;
                const double x2 = x * x, y2 = y * y, r2 = x2 + y2, _2xy = 2 * x * y;
                const double kr = (1 + ((k3 * r2 + k2) * r2 + k1) * r2) / (1 + ((0 * r2 + 0) * r2 + 0) * r2);
                const double xd = x * kr + p1 * _2xy + p2 * (r2 + 2 * x2);
                const double yd = y * kr + p1 * (r2 + 2 * y2) + p2 * _2xy;
                const double u = fx * xd + u0, v = fy * yd + v0;
                auto sat = [](double t) { return t < -2147483648.0 ? (long)INT32_MIN : (t > 2147483647.0 ? (long)INT32_MAX : std::lrint(t)); };
                const long iu = sat(u * 32), iv = sat(v * 32);
                out[(size_t)(y0 + i) * w + j] =
                    make_uint2((uint32_t)(uint16_t)(short)(iu >> 5) | ((uint32_t)(uint16_t)(short)(iv >> 5) << 16), (uint32_t)((iv & 31) * 32 + (iu & 31)));
            }
        }
    }
}

// One record per FAST cell the reference actually runs (ComputeKeyPointsOctTree :804-821; cells with iniY >= maxBorderY - 3 or
// iniX >= maxBorderX - 6 are skipped there and are not listed): x = level, y = iniX | iniY << 16, z = sub-image width | height << 16,
// w = candidate coordinate offset (3 + j * wCell) | (3 + i * hCell) << 16.
void build_cell_table(const McsPlan& P, std::vector<uint4>& tab) {
    tab.clear();
    for (int l = 0; l < P.nlevels; l++) {
        const McsLevel& L = P.L[l];
        const int maxBX = L.w - MCS_MIN_BORDER, maxBY = L.h - MCS_MIN_BORDER;
        for (int i = 0; i < L.n_rows; i++)
            for (int j = 0; j < L.n_cols; j++) {
                const int iniX = MCS_MIN_BORDER + j * L.w_cell, iniY = MCS_MIN_BORDER + i * L.h_cell;
                if (iniY >= maxBY - 3 || iniX >= maxBX - 6) continue;
                const int cw = std::min(iniX + L.w_cell + 6, maxBX) - iniX, ch = std::min(iniY + L.h_cell + 6, maxBY) - iniY;
                if (cw - 6 <= 0 || ch - 6 <= 0) continue;
                tab.push_back(make_uint4((unsigned)l, (unsigned)iniX | ((unsigned)iniY << 16), (unsigned)cw | ((unsigned)ch << 16),
                                         (unsigned)(3 + j * L.w_cell) | ((unsigned)(3 + i * L.h_cell) << 16)));
            }
    }
}

// Strips of the FAST cell grid for fast_strips_kernel: up to 187 interior columns (<= 5 cells) of one cell row, the same cells
// build_cell_table lists (skip rules of :809, :818).  Two uint4 per strip:
//   a.x = level | cells << 8 | wCell << 16, a.y = iniX | iniY << 16 (first cell), a.z = interior width | height << 16,
//   a.w = candidate coordinate offset (3 + j0 * wCell) | (3 + i * hCell) << 16
//   b.x / b.y / b.z = ceil(2^20 / n) for n = pixel groups per row, 16-byte words per tile row, wCell (index divisions in the kernel;
//   exact for the operand ranges of one strip, verified here).
// Also sizes the kernel's shared memory (tile rows, score rows, corner list) for the largest strip of the plan.
bool build_strip_table(McsPlan& P, std::vector<uint4>& tab) {
    tab.clear();
    int chm = 7, area = 1;
    for (int& r : P.ft_box_rows) r = 0;
    auto magic = [](int n, int tmax) -> uint32_t {
        const uint32_t m = ((1u << 20) + n - 1) / n;
        for (int t = 0; t <= tmax; t++)
            if ((int)(((uint64_t)t * m) >> 20) != t / n || (uint64_t)t * m >= (1ull << 32)) return 0;
        return m;
    };
    for (int l = 0; l < P.nlevels; l++) {
        const McsLevel& L = P.L[l];
        const int maxBX = L.w - MCS_MIN_BORDER, maxBY = L.h - MCS_MIN_BORDER;
        if (L.n_cols < 1) continue;
        int live_cols = 0;  // cells of a row the reference runs
        for (int j = 0; j < L.n_cols; j++)
            if (MCS_MIN_BORDER + j * L.w_cell < maxBX - 6) live_cols = j + 1;
        if (live_cols < 1) continue;
        const int kmax = std::max(1, MCS_FS_MAXW / L.w_cell);
        const int nstrips = (live_cols + kmax - 1) / kmax;
        for (int i = 0; i < L.n_rows; i++) {
            const int iniY = MCS_MIN_BORDER + i * L.h_cell;
            if (iniY >= maxBY - 3) continue;
            const int ch = std::min(iniY + L.h_cell + 6, maxBY) - iniY;
            if (ch - 6 <= 0) continue;
            for (int s = 0; s < nstrips; s++) {
                const int j0 = (int)((long long)s * live_cols / nstrips), j1 = (int)((long long)(s + 1) * live_cols / nstrips);
                const int iniX = MCS_MIN_BORDER + j0 * L.w_cell;
                const int endX = std::min(MCS_MIN_BORDER + (j1 - 1) * L.w_cell + L.w_cell + 6, maxBX);
                const int wint = endX - iniX - 6, ih = ch - 6;
                if (wint < 1 || wint > MCS_FS_MAXW || j1 - j0 > 8) { set_err("internal: FAST strip geometry"); return false; }
                const int x0 = (iniX & 15) + 3;
                const int nG = ((x0 + wint - 1) >> 2) - (x0 >> 2) + 1, nq = (x0 + wint + 3 + 15) >> 4;
                const int nU = mcs::fs_units_per_row(x0, x0 + wint);  // pre-test units of two words from an even word (fs_scan8)
                const uint32_t m1 = magic(nG, nG * ih), m2 = magic(nq, nq * ch), m3 = magic(L.w_cell, 255), m4 = magic(nU, nU * ih);
                if (!m1 || !m2 || !m3 || !m4) { set_err("internal: FAST strip index division"); return false; }
                tab.push_back(make_uint4((unsigned)l | ((unsigned)(j1 - j0) << 8) | ((unsigned)L.w_cell << 16), (unsigned)iniX | ((unsigned)iniY << 16),
                                         (unsigned)wint | ((unsigned)ih << 16),
                                         (unsigned)(3 + j0 * L.w_cell) | ((unsigned)(3 + i * L.h_cell) << 16)));
                tab.push_back(make_uint4(m1, m2, m3, m4));
                chm = std::max(chm, ch);
                area = std::max(area, wint * ih);
                P.ft_box_rows[l] = std::max(P.ft_box_rows[l], ch);
            }
        }
    }
    int o = align_up(MCS_FS_TP * chm + 16, 16);
    P.fs_ofs_score = o; o += align_up(MCS_FS_TP * (chm - 6 + 2), 16);
    P.fs_ofs_wq = o; o += (MCS_FS_NT / 32) * MCS_FS_WQ8 * 2;
    P.fs_ofs_corner = o; o += align_up(area * 2, 16);
    P.fs_ofs_ctl = o; o += 16;  // mbarrier, corner count, cell mask
    P.fs_smem_bytes = o;
    // the persistent TMA kernel: [tile 0][tile 1][score][entry stacks][corner list], tiles 128-byte aligned (TMA destination)
    P.ft_tile_bytes = align_up(MCS_FS_TP * chm, 128);
    o = 2 * P.ft_tile_bytes;
    P.ft_ofs_score = o; P.ft_score_bytes = align_up(MCS_FS_TP * (chm - 6 + 2), 16); o += P.ft_score_bytes;
    P.ft_ofs_wq = o; o += (MCS_FS_NT / 32) * MCS_FS_WQ * 2;
    P.ft_ofs_corner = o; o += MCS_FT_CORNERS * 2;
    P.ft_ofs_ctl = o; o += 64;  // two mbarriers, corner count, work items, cell masks
    P.ft_smem_bytes = o;
    return true;
}

}  // namespace

// =================================================================================================================
// Engine: device workspace + launch sequence of ORBextractor::operator() for a batch of same-sized images.
// =================================================================================================================
namespace {

enum Stage { ST_UPLOAD = 0, ST_PYRAMID, ST_FAST, ST_OCTREE, ST_BLUR, ST_DESCRIBE, ST_MATCH, ST_MERGE, ST_TRIANG, ST_DOWNLOAD, ST_COUNT };

struct Engine {
    int device = 0;
    cudaStream_t stream = nullptr;
    OrbTables T;
    McsPlan plan;
    int max_images = 0, planned_w = 0, planned_h = 0, alloc_w = 0, alloc_h = 0;
    int use_angle = 0;
    long long launches = 0;
    // device buffers
    uint8_t *d_input = nullptr, *d_pyr = nullptr, *d_blur = nullptr;
    int2 *d_xtab = nullptr, *d_ytab = nullptr;
    int sm_count_ = 148;
    uint4* d_cell_tab = nullptr;  // FAST cell records (build_cell_table)
    uint4* d_strip_tab = nullptr;  // FAST strip records (build_strip_table)
    // optional high-priority side stream for the latency-bound kernels of a launch sequence (octree; merge and triangulation in
    // mcs_rig): their few long-lived blocks are placed ahead of the queued blocks of the issue-bound kernels of other chunks
    struct HiLane {
        cudaStream_t st = nullptr;
        cudaEvent_t e0 = nullptr, e1 = nullptr;
    };
    int fs_smem_launch = 0;        // dynamic shared memory of the strip kernel (>= plan.fs_smem_bytes; MCS_FS_BLOCKS pads it)
    int n_fast_strips = 0;
    // MCS_FAST_TMA: 1 (default) = fast_strips_kernel<true>, the tile arrives by TMA; 0 = fast_strips_kernel<false>, per-thread staging loop;
    // 2 = fast_strips_tma_kernel, persistent blocks with two tile buffers (measured slower, kept for A/B runs: profiles/README.md)
    int fast_tma = 1;
    bool fast_px8 = true;           // MCS_FAST_PX8=0: four pixels per thread in the pre-test (fs_scan) instead of eight (fs_scan8)
    mcs::FastTmaMaps ft_maps{};     // one 3-D tensor map (x, y, image) per level; level 0 follows the image set of the call
    const uint8_t* ft_l0_base = nullptr;
    int ft_l0_pitch = 0;
    size_t ft_l0_stride = 0;
    int ft_blocks_per_sm = MCS_FT_MINB;
    int ft_grid_mult = 1;           // MCS_FT_GRID_MULT: blocks per resident slot (large = one item per block, tuning only)
    int* d_ft_ctr = nullptr;        // work counters of the persistent kernel, one per first image of a launch
    bool fast_per_cell = false;  // MCS_FAST_CELLS=1: the one-block-per-cell kernel (A/B runs)
    int4* d_rs_tab = nullptr;     // resize_level_sel_kernel tables, all levels
    int oct_nt = MCS_OCT_NT;  // octree block size (MCS_OCT_THREADS overrides)
    bool oct_nt_forced = false;
    int rs_nr[MCS_PLAN_MAX_LEVELS] = {};  // source rows one group of MCS_RS_ROWS output rows spans (resize_level_sel_kernel<NR>)
    int rs_xw_ofs[MCS_PLAN_MAX_LEVELS] = {}, rs_yw_ofs[MCS_PLAN_MAX_LEVELS] = {};
    bool rs_sel_ok = false;
    int n_fast_cells = 0;
    uint32_t *d_cand = nullptr, *d_sel = nullptr;
    uint16_t* d_knode = nullptr;
    unsigned char* d_oct_tab = nullptr;  // octree node tables in global memory, only when a level's tables exceed shared memory
    size_t oct_tab_bytes = 0;
    bool oct_gtab = false;
    int *d_cand_count = nullptr, *d_sel_count = nullptr, *d_kp_count = nullptr, *d_mono = nullptr;
    mcs::KeypointPod *d_kps = nullptr, *d_kps_tmp = nullptr;
    uint8_t *d_desc = nullptr, *d_desc_tmp = nullptr;
    // pinned host mirrors
    mcs::KeypointPod* h_kps = nullptr;
    uint8_t* h_desc = nullptr;
    int *h_kp_count = nullptr, *h_mono = nullptr, *h_sel_count = nullptr, *h_cand_count = nullptr;
    size_t in_bytes = 0;
    mcs::ImageSet last_set{};
    int last_n = 0;
    bool profiling = false;
    bool blur_valid = false;  // d_blur holds the blurred pyramid of the last run (only the steered mode needs it on the path)
    cudaEvent_t ev[ST_COUNT + 1] = {};
    bool ev_rec[ST_COUNT + 1] = {};
    float stage_ms[ST_COUNT] = {};
    size_t oct_smem = 0;

    ~Engine() { release(); }

    void free_all() {
        cudaFree(d_input); cudaFree(d_pyr); cudaFree(d_blur); cudaFree(d_xtab); cudaFree(d_ytab); cudaFree(d_cell_tab); cudaFree(d_strip_tab); cudaFree(d_rs_tab); cudaFree(d_cand); cudaFree(d_sel);
        cudaFree(d_knode); cudaFree(d_ft_ctr); cudaFree(d_cand_count); cudaFree(d_sel_count); cudaFree(d_kp_count); cudaFree(d_mono); cudaFree(d_kps);
        cudaFree(d_kps_tmp); cudaFree(d_desc); cudaFree(d_desc_tmp); cudaFree(d_oct_tab);
        d_oct_tab = nullptr; oct_tab_bytes = 0;
        cudaFreeHost(h_kps); cudaFreeHost(h_desc); cudaFreeHost(h_kp_count); cudaFreeHost(h_mono); cudaFreeHost(h_sel_count);
        cudaFreeHost(h_cand_count);
        d_input = d_pyr = d_blur = nullptr; d_xtab = d_ytab = nullptr; d_cell_tab = nullptr; d_strip_tab = nullptr; d_rs_tab = nullptr; d_cand = d_sel = nullptr; d_knode = nullptr;
        d_ft_ctr = nullptr; d_cand_count = d_sel_count = d_kp_count = d_mono = nullptr; d_kps = d_kps_tmp = nullptr; d_desc = d_desc_tmp = nullptr;
        h_kps = nullptr; h_desc = nullptr; h_kp_count = h_mono = h_sel_count = h_cand_count = nullptr;
    }
    void release() {
        if (!stream) return;
        cudaSetDevice(device);
        cudaStreamSynchronize(stream);
        free_all();
        for (auto& e : ev) if (e) cudaEventDestroy(e);
        cudaStreamDestroy(stream);
        stream = nullptr;
    }

    int init(int nf, float sf, int nl, int ini, int mn, int dev, int max_w, int max_h, int max_img) {
        if (nf < 1 || nl < 1 || nl > MCS_PLAN_MAX_LEVELS || !(sf > 1.0f) || max_img < 1) { set_err("invalid extractor parameters"); return MCS_ERR_ARG; }
        int ndev = 0;
        if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1) { set_err("no CUDA device: this library has no CPU path"); return MCS_ERR_CUDA; }
        if (dev < 0 || dev >= ndev) { set_err("device ordinal out of range"); return MCS_ERR_ARG; }
        device = dev;
        CK(cudaSetDevice(device));
        T.build(nf, sf, nl, ini, mn);
        max_images = max_img;
        CK(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
        for (auto& e : ev) CK(cudaEventCreate(&e));
        int k7[7];
        gauss_kernel_fixed(7, 2.0, k7);
        CK(cudaMemcpyToSymbol(mcs::c_gauss7, k7, sizeof(k7)));
        CK(cudaMemcpyToSymbol(mcs::g_pattern, MCS_ORB_PATTERN_XY, 1024));
        {
            // slot order per lane: greedy over slots, each lane takes the remaining test whose two sample words add least to the fullest
            // banks of the slot's two loads (same word = broadcast, no conflict)
            uint32_t ofs[256];
            bool used[256] = {};
            for (int slot = 0; slot < 8; slot++) {
                int load[2][32] = {};
                std::vector<int> words[2];
                for (int lane = 0; lane < 32; lane++) {
                    int best = -1, best_cost = 1 << 30;
                    uint32_t best_o[2] = {0, 0};
                    for (int b = 0; b < 8; b++) {
                        const int t = lane * 8 + b;
                        if (used[t]) continue;
                        const signed char* q = MCS_ORB_PATTERN_XY + 4 * t;  // x0, y0, x1, y1
                        const uint32_t o[2] = {(uint32_t)(q[1] * 28 + q[0] + 377), (uint32_t)(q[3] * 28 + q[2] + 377)};
                        int cost = 0;
                        for (int e = 0; e < 2; e++) {
                            const int w = (int)(o[e] >> 2);
                            const bool bcast = std::find(words[e].begin(), words[e].end(), w) != words[e].end();
                            cost += bcast ? 0 : (load[e][w & 31] + 1) * (load[e][w & 31] + 1);
                        }
                        if (cost < best_cost) { best_cost = cost; best = b; best_o[0] = o[0]; best_o[1] = o[1]; }
                    }
                    used[lane * 8 + best] = true;
                    for (int e = 0; e < 2; e++) {
                        const int w = (int)(best_o[e] >> 2);
                        if (std::find(words[e].begin(), words[e].end(), w) == words[e].end()) { words[e].push_back(w); load[e][w & 31]++; }
                    }
                    ofs[lane * 8 + slot] = best_o[0] | (best_o[1] << 10) | ((uint32_t)best << 20);
                }
            }
            CK(cudaMemcpyToSymbol(mcs::g_pattern_ofs28, ofs, sizeof(ofs)));
        }
        return replan(max_w, max_h);
    }

    void launch_octree(const McsPlan& P, int n, int img0, int nt, cudaStream_t s, const uint32_t* cand, const int* cand_count, uint16_t* knode,
                       uint32_t* sel, int* sel_count) {
        if (oct_gtab)
            mcs::octree_kernel<true><<<dim3(P.nlevels, n), nt, 0, s>>>(P, cand, cand_count, knode, sel, sel_count,
                                                                     d_oct_tab + (size_t)img0 * P.nlevels * oct_smem, oct_smem);
        else
            mcs::octree_kernel<false><<<dim3(P.nlevels, n), nt, oct_smem, s>>>(P, cand, cand_count, knode, sel, sel_count, nullptr, 0);
    }

    // (re)build the plan for an image size; reallocates when the size grows
    int replan(int w, int h) {
        if (w == planned_w && h == planned_h && plan.use_angle == use_angle) return 0;
        std::vector<int2> xt, yt;
        McsPlan np;
        if (!build_plan(T, w, h, use_angle, np, xt, yt)) return MCS_ERR_ARG;
        CK(cudaSetDevice(device));
        const bool grow = (size_t)w * h > (size_t)alloc_w * alloc_h || np.pyr_stride > plan.pyr_stride || np.cand_stride > plan.cand_stride ||
                          np.sel_stride > plan.sel_stride || np.kp_cap > plan.kp_cap || !d_pyr;
        if (grow) {
            CK(cudaStreamSynchronize(stream));
            free_all();
            const size_t n = max_images;
            in_bytes = (size_t)w * h;
            CK(cudaMalloc(&d_input, n * in_bytes + 256));
            CK(cudaMalloc(&d_pyr, n * np.pyr_stride + 256));
            CK(cudaMalloc(&d_blur, n * np.pyr_stride + 256));
            CK(cudaMalloc(&d_cand, n * np.cand_stride * sizeof(uint32_t)));
            CK(cudaMalloc(&d_knode, n * np.cand_stride * sizeof(uint16_t)));
            CK(cudaMalloc(&d_sel, n * np.sel_stride * sizeof(uint32_t)));
            CK(cudaMalloc(&d_cand_count, n * MCS_PLAN_MAX_LEVELS * sizeof(int)));
            CK(cudaMalloc(&d_ft_ctr, n * sizeof(int)));
            CK(cudaMalloc(&d_sel_count, n * MCS_PLAN_MAX_LEVELS * sizeof(int)));
            CK(cudaMalloc(&d_kp_count, n * sizeof(int)));
            CK(cudaMalloc(&d_mono, n * sizeof(int)));
            CK(cudaMalloc(&d_kps, n * np.kp_cap * sizeof(mcs::KeypointPod)));
            CK(cudaMalloc(&d_kps_tmp, n * np.kp_cap * sizeof(mcs::KeypointPod)));
            CK(cudaMalloc(&d_desc, n * np.kp_cap * 32));
            CK(cudaMalloc(&d_desc_tmp, n * np.kp_cap * 32));
            CK(cudaMallocHost(&h_kps, n * np.kp_cap * sizeof(mcs::KeypointPod)));
            CK(cudaMallocHost(&h_desc, n * np.kp_cap * 32));
            CK(cudaMallocHost(&h_kp_count, n * sizeof(int)));
            CK(cudaMallocHost(&h_mono, n * sizeof(int)));
            CK(cudaMallocHost(&h_sel_count, n * MCS_PLAN_MAX_LEVELS * sizeof(int)));
            CK(cudaMallocHost(&h_cand_count, n * MCS_PLAN_MAX_LEVELS * sizeof(int)));
            CK(cudaMemset(d_kp_count, 0, n * sizeof(int)));
            alloc_w = w; alloc_h = h;
        } else {
            CK(cudaStreamSynchronize(stream));
        }
        cudaFree(d_xtab); cudaFree(d_ytab); cudaFree(d_cell_tab);
        d_xtab = d_ytab = nullptr; d_cell_tab = nullptr;
        cudaFree(d_strip_tab);
        d_strip_tab = nullptr;
        std::vector<uint4> strips;
        if (!build_strip_table(np, strips)) return MCS_ERR_ARG;
        n_fast_strips = (int)strips.size() / 2;
        CK(cudaMalloc(&d_strip_tab, std::max<size_t>(strips.size(), 1) * sizeof(uint4)));
        if (!strips.empty()) CK(cudaMemcpy(d_strip_tab, strips.data(), strips.size() * sizeof(uint4), cudaMemcpyHostToDevice));
        if (const char* e = getenv("MCS_FAST_CELLS")) fast_per_cell = atoi(e) != 0;
        // octree block size by quota: the per-round scans over a level's candidates want wider blocks when levels hold thousands of
        // features (config 4: 2.28 -> 1.02 ms per 8 frames at 512), narrow ones otherwise (config 2: 128 is 1 % faster than 256 in the pipeline)
        oct_nt = T.nfeatures > 3000 ? 512 : (T.nfeatures > 1500 ? 256 : MCS_OCT_NT);
        if (const char* e = getenv("MCS_OCT_THREADS")) {
            const int v = atoi(e);
            if (v >= 32 && v <= MCS_OCT_NT_MAX && v % 32 == 0) { oct_nt = v; oct_nt_forced = true; }
        }
        std::vector<uint4> cells;
        build_cell_table(np, cells);
        n_fast_cells = (int)cells.size();
        CK(cudaMalloc(&d_cell_tab, std::max<size_t>(cells.size(), 1) * sizeof(uint4)));
        if (!cells.empty()) CK(cudaMemcpy(d_cell_tab, cells.data(), cells.size() * sizeof(uint4), cudaMemcpyHostToDevice));
        {
            std::vector<int4> rs;
            rs_sel_ok = true;
            for (int l = 1; l < np.nlevels && rs_sel_ok; l++) {
                std::vector<int4> xw, yw;
                rs_sel_ok = build_resize_word_tables(np.L[l], np.L[l - 1], xt, yt, xw, yw);
                rs_nr[l] = 0;
                for (size_t r0 = 0; rs_sel_ok && r0 < yw.size(); r0 += MCS_RS_ROWS) {
                    const size_t r1 = std::min(yw.size(), r0 + MCS_RS_ROWS) - 1;
                    rs_nr[l] = std::max(rs_nr[l], yw[r1].y - yw[r0].x + 1);
                    for (size_t r = r0; r < r1; r++)
                        if (yw[r + 1].x < yw[r].x || yw[r].y < yw[r].x) rs_sel_ok = false;  // rows must not go backwards
                }
                if (rs_nr[l] > 8) rs_sel_ok = false;
                rs_xw_ofs[l] = (int)rs.size();
                rs.insert(rs.end(), xw.begin(), xw.end());
                rs_yw_ofs[l] = (int)rs.size();
                rs.insert(rs.end(), yw.begin(), yw.end());
            }
            cudaFree(d_rs_tab);
            d_rs_tab = nullptr;
            if (rs_sel_ok && !rs.empty()) {
                CK(cudaMalloc(&d_rs_tab, rs.size() * sizeof(int4)));
                CK(cudaMemcpy(d_rs_tab, rs.data(), rs.size() * sizeof(int4), cudaMemcpyHostToDevice));
            }
        }
        CK(cudaMalloc(&d_xtab, xt.size() * sizeof(int2)));
        CK(cudaMalloc(&d_ytab, yt.size() * sizeof(int2)));
        CK(cudaMemcpy(d_xtab, xt.data(), xt.size() * sizeof(int2), cudaMemcpyHostToDevice));
        CK(cudaMemcpy(d_ytab, yt.data(), yt.size() * sizeof(int2), cudaMemcpyHostToDevice));
        plan = np;
        planned_w = w; planned_h = h;
        oct_smem = mcs_oct::Shared::bytes(plan.oct_cap_max);
        // node tables of a level that do not fit a block's shared memory (a single level with ~2,150+ features: nlevels 1 or 2 with
        // thousands of features) go to a per-block slice of global memory, octree_kernel<true>
        oct_gtab = oct_smem > 200 * 1024;
        if (oct_gtab) {
            const size_t need = (size_t)max_images * plan.nlevels * oct_smem;
            if (need > ((size_t)8 << 30)) { set_err("features per level x images too large for the octree workspace"); return MCS_ERR_ARG; }
            if (need > oct_tab_bytes) {
                CK(cudaStreamSynchronize(stream));
                cudaFree(d_oct_tab);
                d_oct_tab = nullptr;
                oct_tab_bytes = 0;
                CK(cudaMalloc(&d_oct_tab, need));
                oct_tab_bytes = need;
            }
        }
        if (plan.fast_smem_bytes > 200 * 1024) { set_err("internal: FAST cell too large for shared memory"); return MCS_ERR_ARG; }
        CK(cudaFuncSetAttribute(mcs::fast_cells_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, std::max(plan.fast_smem_bytes, 1024)));
        if (plan.fs_smem_bytes > 200 * 1024) { set_err("internal: FAST strip too large for shared memory"); return MCS_ERR_ARG; }
        fs_smem_launch = plan.fs_smem_bytes;
        if (const char* e = getenv("MCS_FS_BLOCKS")) {  // tuning: cap the strip kernel's resident blocks per SM through its shared memory
            const int nb = atoi(e);
            if (nb >= 1 && nb <= 8) fs_smem_launch = std::max(fs_smem_launch, std::min(200 * 1024, (227 * 1024) / nb - 1024 - 64));
        }
        CK(cudaFuncSetAttribute(mcs::fast_strips_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, std::max(fs_smem_launch, 1024)));
        if (const char* e = getenv("MCS_FAST_TMA")) fast_tma = std::min(std::max(atoi(e), 0), 2);
        if (plan.ft_smem_bytes > 200 * 1024 && fast_tma == 2) fast_tma = 1;
        CK(cudaFuncSetAttribute(mcs::fast_strips_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, std::max(fs_smem_launch, 1024)));
        CK(cudaFuncSetAttribute(mcs::fast_strips_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, std::max(fs_smem_launch, 1024)));
        if (const char* e = getenv("MCS_FAST_PX8")) fast_px8 = atoi(e) != 0;
        if (fast_tma) {
            CK(cudaFuncSetAttribute(mcs::fast_strips_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, std::max(plan.ft_smem_bytes, 1024)));
            int nb = 0;
            CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, mcs::fast_strips_tma_kernel, MCS_FS_NT, plan.ft_smem_bytes));
            ft_blocks_per_sm = std::max(nb, 1);
            if (const char* e = getenv("MCS_FT_BLOCKS")) {
                const int v = atoi(e);
                if (v >= 1 && v <= ft_blocks_per_sm) ft_blocks_per_sm = v;
            }
            if (const char* e = getenv("MCS_FT_GRID_MULT")) ft_grid_mult = std::max(1, atoi(e));
            ft_l0_base = nullptr;
            for (int l = 1; l < plan.nlevels; l++) {
                if (plan.ft_box_rows[l] < 1) continue;
                int rc = encode_fast_map(l, d_pyr + plan.L[l].plane_ofs, plan.L[l].pitch, (size_t)plan.pyr_stride);
                if (rc) return rc;
            }
        }
        if (!oct_gtab) CK(cudaFuncSetAttribute(mcs::octree_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)std::max<size_t>(oct_smem, 1024)));
        last_n = 0;
        return 0;
    }

    // tensor map of level l for fast_strips_tma_kernel: u8 [max_images][h][w] with the given row pitch and image stride, box 208 x
    // box_rows x 1, no swizzle, out-of-bounds elements read as zero
    int encode_fast_map(int l, const uint8_t* base, int pitch, size_t image_stride) {
        typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                     const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
        static EncodeFn enc = nullptr;
        if (!enc) {
            void* fp = nullptr;
            cudaDriverEntryPointQueryResult qr;
            if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &qr) != cudaSuccess || qr != cudaDriverEntryPointSuccess || !fp) {
                set_err("cuTensorMapEncodeTiled is not available from this driver");
                return MCS_ERR_CUDA;
            }
            enc = (EncodeFn)fp;
        }
        if (((uintptr_t)base & 15) || (pitch & 15) || (image_stride & 15)) { set_err("internal: FAST tensor map alignment"); return MCS_ERR_ARG; }
        const cuuint64_t dims[3] = {(cuuint64_t)plan.L[l].w, (cuuint64_t)plan.L[l].h, (cuuint64_t)max_images};
        const cuuint64_t strides[2] = {(cuuint64_t)pitch, (cuuint64_t)image_stride};
        const cuuint32_t box[3] = {MCS_FS_TP, (cuuint32_t)plan.ft_box_rows[l], 1};
        const cuuint32_t es[3] = {1, 1, 1};
        const CUresult r = enc(&ft_maps.map[l], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, const_cast<uint8_t*>(base), dims, strides, box, es,
                               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) { set_err("cuTensorMapEncodeTiled failed (" + std::to_string((int)r) + ")"); return MCS_ERR_CUDA; }
        ft_maps.box_rows[l] = plan.ft_box_rows[l];
        return 0;
    }

    // level 0 follows the image set of the call (a caller's device buffer, d_input or the re-pitched plane in d_pyr)
    int encode_level0_map(const mcs::ImageSet& set0) {
        if (plan.ft_box_rows[0] < 1 || (set0.l0 == ft_l0_base && set0.l0_pitch == ft_l0_pitch && set0.l0_stride == ft_l0_stride)) return 0;
        int rc = encode_fast_map(0, set0.l0, set0.l0_pitch, set0.l0_stride);
        if (rc) return rc;
        ft_l0_base = set0.l0; ft_l0_pitch = set0.l0_pitch; ft_l0_stride = set0.l0_stride;
        return 0;
    }

    void mark(int st) {
        if (profiling) { cudaEventRecord(ev[st], stream); ev_rec[st] = true; }
    }
    void begin_profile() { for (auto& r : ev_rec) r = false; for (auto& m : stage_ms) m = 0; }
    // elapsed between consecutive recorded marks -> stage_ms[stage that started at the earlier mark]
    void collect_profile() {
        if (!profiling) return;
        int prev = -1;
        for (int s = 0; s <= ST_COUNT; s++) {
            if (!ev_rec[s]) continue;
            if (prev >= 0) { float ms = 0; cudaEventElapsedTime(&ms, ev[prev], ev[s]); stage_ms[prev] = ms; }
            prev = s;
        }
    }

    // Stage A on images [img0, img0 + n) of the workspace, already described by `set0` (the set of image 0), on stream `st`
    // (0 = the handle's main stream).  Leaves keypoints/descriptors/counts on the device.
    int run(const mcs::ImageSet& set0, int n, const int lap[2], int img0 = 0, cudaStream_t st = nullptr, const HiLane* hi = nullptr) {
        if (n < 1 || img0 < 0 || img0 + n > max_images) { set_err("n_images exceeds the handle's capacity"); return MCS_ERR_CAPACITY; }
        if (!st) st = stream;
        last_set = set0; last_n = img0 == 0 ? n : std::max(last_n, img0 + n);
        const McsPlan& P = plan;
        mcs::ImageSet set = set0;
        set.l0 = set0.l0 + (size_t)img0 * set0.l0_stride;
        set.pyr = set0.pyr + (size_t)img0 * P.pyr_stride;
        uint32_t* cand = d_cand + (size_t)img0 * P.cand_stride;
        uint16_t* knode = d_knode + (size_t)img0 * P.cand_stride;
        uint32_t* sel = d_sel + (size_t)img0 * P.sel_stride;
        int *cand_count = d_cand_count + (size_t)img0 * P.nlevels, *sel_count = d_sel_count + (size_t)img0 * P.nlevels;
        uint8_t* blur = d_blur + (size_t)img0 * P.pyr_stride;
        mcs::KeypointPod *kps = d_kps + (size_t)img0 * P.kp_cap, *kps_tmp = d_kps_tmp + (size_t)img0 * P.kp_cap;
        uint8_t *desc = d_desc + (size_t)img0 * P.kp_cap * 32, *desc_tmp = d_desc_tmp + (size_t)img0 * P.kp_cap * 32;
        int *kp_count = d_kp_count + img0, *mono = d_mono + img0;
        mark(ST_PYRAMID);
        CK(cudaMemsetAsync(cand_count, 0, (size_t)n * P.nlevels * sizeof(int), st));
        for (int l = 1; l < P.nlevels; l++) {
            dim3 b(32, 8), g((P.L[l].w + 127) / 128, (P.L[l].h + 7) / 8, n);
            if (rs_sel_ok && d_rs_tab) {
                g.y = (P.L[l].h + 8 * MCS_RS_ROWS - 1) / (8 * MCS_RS_ROWS);
                mcs::ResizeSelArgs A;
                A.src = l == 1 ? set.l0 : set.pyr + P.L[l - 1].plane_ofs;
                A.src_stride = l == 1 ? set.l0_stride : (size_t)P.pyr_stride;
                A.spitch = l == 1 ? set.l0_pitch : P.L[l - 1].pitch;
                A.dst = set.pyr + P.L[l].plane_ofs;
                A.dst_stride = (size_t)P.pyr_stride;
                A.dpitch = P.L[l].pitch;
                A.lastw = ((P.L[l - 1].w + 3) >> 2) - 1;
                A.sh = P.L[l - 1].h;
                A.dwords = (P.L[l].w + 3) / 4;
                A.dh = P.L[l].h;
                A.xw = d_rs_tab + rs_xw_ofs[l];
                A.yw = d_rs_tab + rs_yw_ofs[l];
                // level 0 aliasing a caller's buffer is the only source plane without readable padding behind its last row
                const uint8_t* l0b = set0.l0;
                const bool padded = l > 1 || (l0b >= d_pyr && l0b < d_pyr + (size_t)max_images * P.pyr_stride) ||
                                    (l0b >= d_input && l0b < d_input + (size_t)max_images * in_bytes);
                if (rs_nr[l] <= 6) {
                    if (padded) mcs::resize_level_sel_kernel<6, false><<<g, b, 0, st>>>(A);
                    else mcs::resize_level_sel_kernel<6, true><<<g, b, 0, st>>>(A);
                } else {
                    if (padded) mcs::resize_level_sel_kernel<8, false><<<g, b, 0, st>>>(A);
                    else mcs::resize_level_sel_kernel<8, true><<<g, b, 0, st>>>(A);
                }
            } else if (T.scale_factor_f <= 2.0f) mcs::resize_level_words_kernel<<<g, b, 0, st>>>(P, set, l, d_xtab, d_ytab);
            else mcs::resize_level_kernel<<<g, b, 0, st>>>(P, set, l, d_xtab, d_ytab);
            launches++;
        }
        mark(ST_FAST);
        if (fast_per_cell && n_fast_cells > 0) {
            mcs::fast_cells_kernel<<<dim3(n_fast_cells, n), MCS_FAST_NT, P.fast_smem_bytes, st>>>(P, set, d_cell_tab, cand, cand_count);
            launches++;
        } else if (n_fast_strips > 0 && fast_tma == 2) {
            int rc = encode_level0_map(set0);
            if (rc) return rc;
            const long long items = (long long)n_fast_strips * n;
            const int grid = (int)std::min<long long>(items, (long long)sm_count_ * ft_blocks_per_sm * ft_grid_mult);
            CK(cudaMemsetAsync(d_ft_ctr + img0, 0, sizeof(int), st));  // launches in flight together start at different images
            // tensor maps are encoded on image 0 of the workspace; cand / cand_count are already offset to image img0
            mcs::fast_strips_tma_kernel<<<grid, MCS_FS_NT, P.ft_smem_bytes, st>>>(P, ft_maps, d_strip_tab, n_fast_strips, (int)items, img0, cand, cand_count,
                                                                                   d_ft_ctr + img0);
            launches++;
        } else if (n_fast_strips > 0 && fast_tma == 1) {
            int rc = encode_level0_map(set0);
            if (rc) return rc;
            if (fast_px8) mcs::fast_strips_kernel<true, true><<<dim3(n_fast_strips, n), MCS_FS_NT, fs_smem_launch, st>>>(P, set, ft_maps, img0, d_strip_tab, cand, cand_count);
            else mcs::fast_strips_kernel<true, false><<<dim3(n_fast_strips, n), MCS_FS_NT, fs_smem_launch, st>>>(P, set, ft_maps, img0, d_strip_tab, cand, cand_count);
            launches++;
        } else if (n_fast_strips > 0) {
            mcs::fast_strips_kernel<false, false><<<dim3(n_fast_strips, n), MCS_FS_NT, fs_smem_launch, st>>>(P, set, ft_maps, img0, d_strip_tab, cand, cand_count);
            launches++;
        }
        mark(ST_OCTREE);
        // Few images (the per-frame latency mode): the launch cannot fill the SMs with blocks, so each block gets more threads and
        // its per-round passes over the candidates take fewer steps (one rig frame 0.49 -> 0.45 ms, eight 0.80 -> 0.76 ms); large
        // batches keep the narrow blocks, which interleave better with the other chunks' kernels.
        int oct_nt = this->oct_nt;
        if (!oct_nt_forced) {
            const int blocks = P.nlevels * n;
            if (blocks * 2 <= sm_count_) oct_nt = std::max(oct_nt, 512);
            else if (blocks <= 4 * sm_count_) oct_nt = std::max(oct_nt, 256);
        }
        if (hi) {
            CK(cudaEventRecord(hi->e0, st));
            CK(cudaStreamWaitEvent(hi->st, hi->e0, 0));
            launch_octree(P, n, img0, oct_nt, hi->st, cand, cand_count, knode, sel, sel_count);
            CK(cudaEventRecord(hi->e1, hi->st));
            CK(cudaStreamWaitEvent(st, hi->e1, 0));
        } else {
            launch_octree(P, n, img0, oct_nt, st, cand, cand_count, knode, sel, sel_count);
        }
        launches++;
        mark(ST_BLUR);
        if (P.use_angle) {
            // steered mode: rotated pattern offsets reach +-18 px, so the whole level is blurred first
            mcs::blur_kernel<<<dim3(P.n_tiles, n), 256, 0, st>>>(P, set, blur);
            launches++;
            mark(ST_DESCRIBE);
            mcs::describe_kernel<<<dim3((P.kp_cap + MCS_DESC_WARPS - 1) / MCS_DESC_WARPS, n), MCS_DESC_WARPS * 32, 0, st>>>(
                P, set, blur, sel, sel_count, kps, desc, kp_count);
            launches++;
            blur_valid = true;
        } else {
            // shipped mode (angle 0): blur only the 32 x 32 patch under every keypoint, inside the descriptor kernel
            mark(ST_DESCRIBE);
            mcs::describe_fused_kernel<<<dim3((P.kp_cap + MCS_DESCF_WARPS * MCS_DESCF_KPW - 1) / (MCS_DESCF_WARPS * MCS_DESCF_KPW), n), MCS_DESCF_WARPS * 32, 0, st>>>(
                P, set, sel, sel_count, kps, desc, kp_count);
            launches++;
            blur_valid = false;
        }
        // every keypoint has pt.x >= 19, so a lapping area ending before 19 (the callers' {0,0}) selects nothing (:1153)
        const bool lap_active = lap && lap[1] >= lap[0] && lap[1] >= MCS_EDGE_THRESHOLD;
        if (lap_active) {
            mcs::lapping_permute_kernel<<<n, 256, 0, st>>>(P, lap[0], lap[1], kps, desc, kps_tmp, desc_tmp, kp_count, mono);
            launches++;
        } else {
            CK(cudaMemcpyAsync(mono, kp_count, (size_t)n * sizeof(int), cudaMemcpyDeviceToDevice, st));
        }
        mark(ST_MATCH);
        CK(cudaGetLastError());
        return 0;
    }

    // host images [img0, img0 + n) -> d_input (tight); `set` describes image 0 of the workspace
    int upload(const uint8_t* images, int n, int w, int h, int stride, size_t image_stride, mcs::ImageSet& set, int img0 = 0,
               cudaStream_t st = nullptr) {
        if (n < 1 || img0 + n > max_images) { set_err("n_images exceeds the handle's capacity"); return MCS_ERR_CAPACITY; }
        if (!st) st = stream;
        mark(ST_UPLOAD);
        uint8_t* dst = d_input + (size_t)img0 * w * h;
        if (stride == w && image_stride == (size_t)w * h) {
            CK(cudaMemcpyAsync(dst, images, (size_t)n * w * h, cudaMemcpyHostToDevice, st));
        } else {
            for (int i = 0; i < n; i++)
                CK(cudaMemcpy2DAsync(dst + (size_t)i * w * h, w, images + i * image_stride, stride, w, h, cudaMemcpyHostToDevice, st));
        }
        return make_set(d_input, n, w, h, set, img0, st);
    }
    // setData conversion (K0): n tight f32 images (1 or 3 interleaved channels) in device memory -> tight u8 gray images at
    // d_input[img0 ...], then the usual image set.  w * h must be a multiple of 4 (checked by the callers).
    int stage_f32(const float* d_src, int channels, int n, int w, int h, mcs::ImageSet& set, int img0, cudaStream_t st) {
        if (n < 1 || img0 + n > max_images) { set_err("n_images exceeds the handle's capacity"); return MCS_ERR_CAPACITY; }
        const size_t quads = (size_t)n * w * h / 4;
        const int blocks = (int)std::min<size_t>((quads + 255) / 256, (size_t)sm_count_ * 16);
        uint8_t* dst = d_input + (size_t)img0 * w * h;
        if (channels == 1) mcs::stage_f32_kernel<1><<<blocks, 256, 0, st>>>(d_src, dst, quads);
        else mcs::stage_f32_kernel<3><<<blocks, 256, 0, st>>>(d_src, dst, quads);
        CK(cudaGetLastError());
        launches++;
        return make_set(d_input, n, w, h, set, img0, st);
    }
    // RECTIFY branch of setData: n tight images starting at image img0 of d_images are undistorted (K0b) into the pyramid's level-0
    // planes, which become the image set's level 0.
    int make_set_undistorted(const uint8_t* d_images, int n, int w, int h, const uint2* maps, int num_cams, mcs::ImageSet& set, int img0,
                             cudaStream_t st) {
        if (n < 1 || img0 + n > max_images) { set_err("n_images exceeds the handle's capacity"); return MCS_ERR_CAPACITY; }
        dim3 b(32, 8), g((w + 127) / 128, (h + 7) / 8, n);
        mcs::undistort_u8_kernel<<<g, b, 0, st>>>(d_images + (size_t)img0 * w * h, (size_t)w * h,
                                                  d_pyr + (size_t)img0 * plan.pyr_stride + plan.L[0].plane_ofs, plan.pyr_stride, plan.L[0].pitch,
                                                  maps, num_cams, img0, w, h);
        CK(cudaGetLastError());
        launches++;
        set.pyr = d_pyr;
        set.l0 = d_pyr + plan.L[0].plane_ofs; set.l0_pitch = plan.L[0].pitch; set.l0_stride = plan.pyr_stride;
        return 0;
    }
    // device-resident tight images (d_images = image 0); re-pitched only when 128-bit row loads would be misaligned
    int make_set(const uint8_t* d_images, int n, int w, int h, mcs::ImageSet& set, int img0 = 0, cudaStream_t st = nullptr) {
        if (!st) st = stream;
        set.pyr = d_pyr;
        if ((w & 15) == 0 && ((uintptr_t)d_images & 15) == 0) {  // rows start 16-byte aligned (FAST tile staging)
            set.l0 = d_images; set.l0_pitch = w; set.l0_stride = (size_t)w * h;
        } else {
            for (int i = img0; i < img0 + n; i++)
                CK(cudaMemcpy2DAsync(d_pyr + (size_t)i * plan.pyr_stride + plan.L[0].plane_ofs, plan.L[0].pitch, d_images + (size_t)i * w * h, w, w,
                                     h, cudaMemcpyDeviceToDevice, st));
            set.l0 = d_pyr + plan.L[0].plane_ofs; set.l0_pitch = plan.L[0].pitch; set.l0_stride = plan.pyr_stride;
        }
        return 0;
    }

    // device -> pinned host mirrors for images [img0, img0 + n); asynchronous on `st`
    int download_features_async(int img0, int n, cudaStream_t st) {
        const size_t nl = plan.nlevels, cap = plan.kp_cap;
        CK(cudaMemcpyAsync(h_kp_count + img0, d_kp_count + img0, (size_t)n * sizeof(int), cudaMemcpyDeviceToHost, st));
        CK(cudaMemcpyAsync(h_mono + img0, d_mono + img0, (size_t)n * sizeof(int), cudaMemcpyDeviceToHost, st));
        CK(cudaMemcpyAsync(h_sel_count + img0 * nl, d_sel_count + img0 * nl, (size_t)n * nl * sizeof(int), cudaMemcpyDeviceToHost, st));
        CK(cudaMemcpyAsync(h_cand_count + img0 * nl, d_cand_count + img0 * nl, (size_t)n * nl * sizeof(int), cudaMemcpyDeviceToHost, st));
        CK(cudaMemcpyAsync(h_kps + img0 * cap, d_kps + img0 * cap, (size_t)n * cap * sizeof(mcs::KeypointPod), cudaMemcpyDeviceToHost, st));
        CK(cudaMemcpyAsync(h_desc + img0 * cap * 32, d_desc + img0 * cap * 32, (size_t)n * cap * 32, cudaMemcpyDeviceToHost, st));
        return 0;
    }
    int download_features(int n) {
        mark(ST_DOWNLOAD);
        int rc = download_features_async(0, n, stream);
        if (rc) return rc;
        mark(ST_COUNT);
        CK(cudaStreamSynchronize(stream));
        collect_profile();
        return check_counts(n);
    }
    int check_counts(int n) {
        for (int i = 0; i < n; i++)
            for (int l = 0; l < plan.nlevels; l++) {
                if (h_sel_count[i * plan.nlevels + l] < 0) { set_err("octree node table overflow"); return MCS_ERR_CAPACITY; }
                if (h_cand_count[i * plan.nlevels + l] > plan.L[l].cand_cap) { set_err("FAST candidate list overflow"); return MCS_ERR_CAPACITY; }
            }
        return 0;
    }
};

}  // namespace

// =================================================================================================================
// C ABI — ORBextractor
// =================================================================================================================
struct mcs_extractor {
    Engine eng;
};

extern "C" {

const char* mcs_last_error(void) { return g_err.c_str(); }
const char* mcs_version(void) { return "mcslam_b200 0.1 (sm_100a)"; }
// Host-only: the FAST work decomposition of one extractor configuration (no device needed).  cells[i] = {level, iniX | iniY << 16,
// sub-image width | height << 16, offset}; strips[i] = the two uint4 of build_strip_table.  Returns 0, or a negative error code.
int mcs_plan_fast_tables(int width, int height, int nfeatures, float scale_factor, int nlevels, uint32_t* cells, int* n_cells,
                         uint32_t* strips, int* n_strips, int* strip_smem_bytes, int capacity) {
    if (width < 1 || height < 1 || nfeatures < 1 || nlevels < 1 || nlevels > MCS_PLAN_MAX_LEVELS || !(scale_factor > 1.f) || !n_cells || !n_strips) {
        set_err("mcs_plan_fast_tables: bad arguments");
        return MCS_ERR_ARG;
    }
    OrbTables T;
    T.build(nfeatures, scale_factor, nlevels, 20, 7);
    McsPlan P;
    std::vector<int2> xt, yt;
    if (!build_plan(T, width, height, 0, P, xt, yt)) return MCS_ERR_ARG;
    std::vector<uint4> ct, st;
    build_cell_table(P, ct);
    if (!build_strip_table(P, st)) return MCS_ERR_ARG;
    *n_cells = (int)ct.size();
    *n_strips = (int)st.size() / 2;
    if (strip_smem_bytes) *strip_smem_bytes = P.fs_smem_bytes;
    if ((int)ct.size() > capacity || (int)st.size() / 2 > capacity) { set_err("mcs_plan_fast_tables: capacity"); return MCS_ERR_CAPACITY; }
    if (cells) memcpy(cells, ct.data(), ct.size() * sizeof(uint4));
    if (strips) memcpy(strips, st.data(), st.size() * sizeof(uint4));
    return 0;
}
int mcs_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
    return n;
}

// Pin the calling host thread to the CPUs next to GPU `device` (its PCIe root's NUMA node, /sys/bus/pci/devices/<bdf>/local_cpulist):
// pinned staging buffers allocated afterwards are first-touched on that node, so the H2D / D2H copies of a multi-GPU host do not
// cross the socket interconnect.  Returns the number of CPUs in the set, 0 when the topology is not exposed (nothing changed).
int mcs_bind_host_thread(int device) {
    char bdf[32] = {0};
    if (cudaDeviceGetPCIBusId(bdf, sizeof bdf, device) != cudaSuccess) { set_err("mcs_bind_host_thread: bad device"); return MCS_ERR_ARG; }
    for (char* c = bdf; *c; c++) *c = (char)tolower(*c);
    const std::string path = std::string("/sys/bus/pci/devices/") + bdf + "/local_cpulist";
    FILE* f = fopen(path.c_str(), "r");
    if (!f) return 0;
    char line[4096] = {0};
    const bool got = fgets(line, sizeof line, f) != nullptr;
    fclose(f);
    if (!got) return 0;
    cpu_set_t set;
    CPU_ZERO(&set);
    int n = 0;
    for (char* p = line; *p;) {  // "0-31,64-95"
        if (*p < '0' || *p > '9') { p++; continue; }
        const long a = strtol(p, &p, 10);
        long b = a;
        if (*p == '-') b = strtol(p + 1, &p, 10);
        for (long c = a; c <= b && c < CPU_SETSIZE; c++) { CPU_SET((int)c, &set); n++; }
    }
    if (n == 0 || sched_setaffinity(0, sizeof set, &set) != 0) return 0;
    return n;
}

mcs_extractor* mcs_extractor_create(int nfeatures, float scale_factor, int nlevels, int ini_th_fast, int min_th_fast, int device,
                                    int max_width, int max_height, int max_images) {
    mcs_extractor* ex = new mcs_extractor();
    if (ex->eng.init(nfeatures, scale_factor, nlevels, ini_th_fast, min_th_fast, device, max_width, max_height, max_images) != 0) {
        delete ex;
        return nullptr;
    }
    return ex;
}
void mcs_extractor_destroy(mcs_extractor* ex) { delete ex; }

int mcs_extractor_set_orientation(mcs_extractor* ex, int enable) {
    if (!ex) return MCS_ERR_ARG;
    ex->eng.use_angle = enable ? 1 : 0;
    return ex->eng.replan(ex->eng.planned_w, ex->eng.planned_h);
}
int mcs_extractor_capacity(const mcs_extractor* ex) { return ex ? ex->eng.plan.kp_cap : MCS_ERR_ARG; }

int mcs_extract_batch(mcs_extractor* ex, const uint8_t* images, int n_images, int width, int height, int stride, size_t image_stride,
                      const int lapping[2], mcs_keypoint* keypoints, uint8_t* descriptors, int capacity, int* mono_index, int* n_total) {
    if (!ex) return MCS_ERR_ARG;
    if (!images || width <= 0 || height <= 0 || n_images <= 0) return MCS_ERR_EMPTY;  // `if(_image.empty()) return -1` (:1090)
    Engine& E = ex->eng;
    if (cudaSetDevice(E.device) != cudaSuccess) return MCS_ERR_CUDA;
    int rc = E.replan(width, height);
    if (rc) return rc;
    E.begin_profile();
    mcs::ImageSet set;
    if ((rc = E.upload(images, n_images, width, height, stride, image_stride, set))) return rc;
    if ((rc = E.run(set, n_images, lapping))) return rc;
    if ((rc = E.download_features(n_images))) return rc;
    const int cap = E.plan.kp_cap;
    for (int i = 0; i < n_images; i++) {
        const int n = E.h_kp_count[i];
        if (n > capacity) { set_err("keypoint capacity too small"); return MCS_ERR_CAPACITY; }
        if (keypoints) memcpy(keypoints + (size_t)i * capacity, E.h_kps + (size_t)i * cap, (size_t)n * sizeof(mcs_keypoint));
        if (descriptors) memcpy(descriptors + (size_t)i * capacity * 32, E.h_desc + (size_t)i * cap * 32, (size_t)n * 32);
        if (mono_index) mono_index[i] = E.h_mono[i];
        if (n_total) n_total[i] = n;
    }
    return 0;
}

int mcs_extract(mcs_extractor* ex, const uint8_t* image, int width, int height, int stride, const uint8_t* mask_ignored,
                const int lapping[2], mcs_keypoint* keypoints, uint8_t* descriptors, int capacity, int* n_total) {
    (void)mask_ignored;  // "Mask is ignored in the current implementation." (ORBextractor.h:56)
    int mono = 0, total = 0;
    const int rc = mcs_extract_batch(ex, image, 1, width, height, stride, (size_t)stride * (height > 0 ? height : 0), lapping, keypoints,
                                     descriptors, capacity, &mono, &total);
    if (rc) return rc;
    if (n_total) *n_total = total;
    return mono;
}

int mcs_extractor_get_tables(const mcs_extractor* ex, int* nlevels, float* scale_factors, float* inv_scale_factors, float* level_sigma2,
                             float* inv_level_sigma2, int* features_per_level) {
    if (!ex) return MCS_ERR_ARG;
    const OrbTables& T = ex->eng.T;
    if (nlevels) *nlevels = T.nlevels;
    for (int i = 0; i < T.nlevels; i++) {
        if (scale_factors) scale_factors[i] = T.scale[i];
        if (inv_scale_factors) inv_scale_factors[i] = T.inv_scale[i];
        if (level_sigma2) level_sigma2[i] = T.sigma2[i];
        if (inv_level_sigma2) inv_level_sigma2[i] = T.inv_sigma2[i];
        if (features_per_level) features_per_level[i] = T.quota[i];
    }
    return 0;
}

int mcs_extractor_level_size(const mcs_extractor* ex, int level, int* width, int* height) {
    if (!ex || level < 0 || level >= ex->eng.plan.nlevels) return MCS_ERR_ARG;
    if (width) *width = ex->eng.plan.L[level].w;
    if (height) *height = ex->eng.plan.L[level].h;
    return 0;
}

static int engine_download_level(Engine& E, int image_index, int level, int kind, uint8_t* dst, int dst_stride) {
    if (image_index < 0 || image_index >= E.last_n || level < 0 || level >= E.plan.nlevels || !dst) return MCS_ERR_ARG;
    if (cudaSetDevice(E.device) != cudaSuccess) return MCS_ERR_CUDA;
    const McsLevel& L = E.plan.L[level];
    const uint8_t* src;
    int pitch;
    if (kind == 2) {
        if (!E.blur_valid) {  // observer only: the shipped-mode path never materialises the blurred pyramid
            CK(cudaStreamSynchronize(E.stream));
            mcs::ImageSet set = E.last_set;
            mcs::blur_kernel<<<dim3(E.plan.n_tiles, E.last_n), 256, 0, E.stream>>>(E.plan, set, E.d_blur);
            CK(cudaStreamSynchronize(E.stream));
            E.blur_valid = true;
        }
        src = E.d_blur + (size_t)image_index * E.plan.pyr_stride + L.plane_ofs;
        pitch = L.pitch;
    } else if (level == 0) {
        src = E.last_set.l0 + (size_t)image_index * E.last_set.l0_stride;
        pitch = E.last_set.l0_pitch;
    } else {
        src = E.d_pyr + (size_t)image_index * E.plan.pyr_stride + L.plane_ofs;
        pitch = L.pitch;
    }
    if (kind == 1) {
        // mvImagePyramid's 19 px REFLECT_101 frame (ORBextractor.cpp:1188/:1193) is never read on the device path; it is
        // re-created here for observers only.
        std::vector<uint8_t> tmp((size_t)L.w * L.h);
        CK(cudaMemcpy2D(tmp.data(), L.w, src, pitch, L.w, L.h, cudaMemcpyDeviceToHost));
        const int B = MCS_EDGE_THRESHOLD;
        auto refl = [](int p, int len) { if (len == 1) return 0; while (p < 0 || p >= len) p = p < 0 ? -p : 2 * len - 2 - p; return p; };
        for (int y = 0; y < L.h + 2 * B; y++)
            for (int x = 0; x < L.w + 2 * B; x++) dst[(size_t)y * dst_stride + x] = tmp[(size_t)refl(y - B, L.h) * L.w + refl(x - B, L.w)];
        return 0;
    }
    CK(cudaMemcpy2D(dst, dst_stride, src, pitch, L.w, L.h, cudaMemcpyDeviceToHost));
    return 0;
}
int mcs_extractor_download_level(mcs_extractor* ex, int image_index, int level, int kind, uint8_t* dst, int dst_stride) {
    if (!ex) return MCS_ERR_ARG;
    return engine_download_level(ex->eng, image_index, level, kind, dst, dst_stride);
}

static int engine_download_candidates(Engine& E, int image_index, int level, int32_t* xys, int capacity) {
    if (image_index < 0 || image_index >= E.last_n || level < 0 || level >= E.plan.nlevels) return MCS_ERR_ARG;
    if (cudaSetDevice(E.device) != cudaSuccess) return MCS_ERR_CUDA;
    int n = 0;
    CK(cudaMemcpy(&n, E.d_cand_count + image_index * E.plan.nlevels + level, sizeof(int), cudaMemcpyDeviceToHost));
    n = std::min(n, E.plan.L[level].cand_cap);
    if (n > capacity) { set_err("candidate capacity too small"); return MCS_ERR_CAPACITY; }
    std::vector<uint32_t> rec(std::max(n, 1));
    CK(cudaMemcpy(rec.data(), E.d_cand + (size_t)image_index * E.plan.cand_stride + E.plan.L[level].cand_ofs, (size_t)n * 4, cudaMemcpyDeviceToHost));
    for (int i = 0; i < n; i++) {
        xys[i * 3] = mcs_oct::cand_x(rec[i]);
        xys[i * 3 + 1] = mcs_oct::cand_y(rec[i]);
        xys[i * 3 + 2] = mcs_oct::cand_score(rec[i]);
    }
    return n;
}
int mcs_extractor_download_candidates(mcs_extractor* ex, int image_index, int level, int32_t* xys, int capacity) {
    if (!ex) return MCS_ERR_ARG;
    return engine_download_candidates(ex->eng, image_index, level, xys, capacity);
}

void* mcs_device_malloc(int device, size_t bytes) {
    void* p = nullptr;
    if (cudaSetDevice(device) != cudaSuccess || cudaMalloc(&p, bytes) != cudaSuccess) { set_err("cudaMalloc failed"); return nullptr; }
    return p;
}
void mcs_device_free(int device, void* p) { if (cudaSetDevice(device) == cudaSuccess) cudaFree(p); }
int mcs_device_upload(int device, void* dst, const void* src, size_t bytes) {
    CK(cudaSetDevice(device));
    CK(cudaMemcpy(dst, src, bytes, cudaMemcpyHostToDevice));
    return 0;
}
int mcs_device_download(int device, void* dst, const void* src, size_t bytes) {
    CK(cudaSetDevice(device));
    CK(cudaMemcpy(dst, src, bytes, cudaMemcpyDeviceToHost));
    return 0;
}
void* mcs_host_alloc_pinned(size_t bytes) {
    void* p = nullptr;
    if (cudaMallocHost(&p, bytes) != cudaSuccess) { set_err("cudaMallocHost failed"); return nullptr; }
    return p;
}
void mcs_host_free_pinned(void* p) { cudaFreeHost(p); }

}  // extern "C"

// =================================================================================================================
// C ABI — rig frames (MultiCameraFrame extract / intra-match / triangulate)
// =================================================================================================================
namespace {

void mat3_mul(const double* A, const double* B, double* C) {
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) {
            double s = 0;
            for (int k = 0; k < 3; k++) s += A[i * 3 + k] * B[k * 3 + j];
            C[i * 3 + j] = s;
        }
}
bool mat3_inv(const double* m, double* o) {
    const double det = m[0] * (m[4] * m[8] - m[5] * m[7]) - m[1] * (m[3] * m[8] - m[5] * m[6]) + m[2] * (m[3] * m[7] - m[4] * m[6]);
    if (det == 0) return false;
    const double id = 1.0 / det;
    o[0] = (m[4] * m[8] - m[5] * m[7]) * id; o[1] = (m[2] * m[7] - m[1] * m[8]) * id; o[2] = (m[1] * m[5] - m[2] * m[4]) * id;
    o[3] = (m[5] * m[6] - m[3] * m[8]) * id; o[4] = (m[0] * m[8] - m[2] * m[6]) * id; o[5] = (m[2] * m[3] - m[0] * m[5]) * id;
    o[6] = (m[3] * m[7] - m[4] * m[6]) * id; o[7] = (m[1] * m[6] - m[0] * m[7]) * id; o[8] = (m[0] * m[4] - m[1] * m[3]) * id;
    return true;
}
// F = K_j^-T [t_ji]x R_ji K_i^-1, T_ji = T_j0 * T_i0^-1 (MultiCameraFrame.cpp:1126-1141); rigid inverse through R^T.
void fundamental_from_extrinsics(const mcs_calib& ci, const mcs_calib& cj, double* F) {
    double RiT[9], Rji[9], tji[3];
    for (int a = 0; a < 3; a++)
        for (int b = 0; b < 3; b++) RiT[a * 3 + b] = ci.R[b * 3 + a];
    mat3_mul(cj.R, RiT, Rji);
    for (int a = 0; a < 3; a++) {
        double s = 0;
        for (int b = 0; b < 3; b++) s += Rji[a * 3 + b] * ci.t[b];
        tji[a] = cj.t[a] - s;
    }
    const double tx[9] = {0, -tji[2], tji[1], tji[2], 0, -tji[0], -tji[1], tji[0], 0};
    double KjInv[9], KiInv[9], KjInvT[9], A[9], B[9];
    mat3_inv(cj.K, KjInv);
    mat3_inv(ci.K, KiInv);
    for (int a = 0; a < 3; a++)
        for (int b = 0; b < 3; b++) KjInvT[a * 3 + b] = KjInv[b * 3 + a];
    mat3_mul(KjInvT, tx, A);
    mat3_mul(A, Rji, B);
    mat3_mul(B, KiInv, F);
}

}  // namespace

struct mcs_rig {
    Engine eng;
    mcs_rig_config cfg{};
    std::vector<mcs_calib> calib;
    int C = 0, F = 0, kp_cap = 0, row_cap = 0;
    mcs::PairList pairs{};
    mcs::CalibDev* d_calib = nullptr;
    double* d_F = nullptr;
    float* d_sigma2 = nullptr;
    int *d_match_train = nullptr, *d_match_tmp = nullptr, *d_rows = nullptr, *d_row_count = nullptr;
    // A slot = what one in-flight chunk of frames needs for itself: a stream, the forked stream of the 2-view solver with its
    // fork/join events, and the triangulation work lists.  Slot 0 runs on the handle's main stream.
    struct Slot {
        cudaStream_t st = nullptr, st2 = nullptr;
        cudaEvent_t ev_fork = nullptr, ev_join = nullptr, ev_done = nullptr;
        int *d_tri_lists = nullptr, *d_tri_counts = nullptr;
        float* d_f32 = nullptr;  // staging of one chunk of f32 input images (process_f32 with host buffers)
        Engine::HiLane hi;       // high-priority side stream (use_hi)
        size_t f32_bytes = 0;
    };
#ifndef MCS_NSLOT
#define MCS_NSLOT 6
#endif
    static constexpr int NSLOT = MCS_NSLOT;
    Slot slots[NSLOT];
    cudaEvent_t ev_begin = nullptr;
    int tri_list_stride = 0, sm_count = 148;
    int chunk_frames = 64;  // frames per pipelined chunk (0 = no chunking)
    int host_chunk_div = 4; // host-buffer calls use chunks of chunk_frames / host_chunk_div (MCS_HOST_CHUNK_DIV): 6 slots x quarter chunks measured best
    int tri_lanes = -1;     // MCS_TRI_LANES: 4 = tri_solve4_kernel (four lanes per row), 1 = tri_solve_kernel (one thread per row), unset = by call size
    int tri4_per_sm = 16;   // MCS_TRI4_PER_SM: resident tri_solve4 blocks per SM
    int tri_per_sm = 3;     // resident tri_solve blocks per SM (MCS_TRI_PER_SM; 0 = as many as shared memory holds): 3 measured best — the
                            // one-warp blocks hold 44 KB of shared memory each for the length of a Jacobi chain, which other chunks' kernels need
    bool match_imma = false; // MCS_MATCH_IMMA=1: experiment, the pair matcher on the u8 integer MMA (hamming_match_imma_kernel); not the product path
    bool use_hi = true;     // octree, merge and triangulation on the slots' high-priority side streams (MCS_HIPRIO=0 switches it off): +0.7 %
    mcs::LandmarkOut lm{};
    float* d_masks = nullptr;
    size_t masks_bytes = 0;
    // RECTIFY (setData :123-136): per-camera undistortion maps and the buffer the undistorted masks go to
    uint2* d_und_maps = nullptr;
    float* d_masks_und = nullptr;
    size_t masks_und_bytes = 0;
    bool undistort_on = false;
    // RECTIFY = false with distorted lenses: image_kps_undist on the device (UndistortKeyPoints, MultiCameraFrame.cpp:300-347)
    bool undist_pts_on = false;
    mcs::KeypointPod *d_kps_undist = nullptr, *h_kps_undist = nullptr;
    mcs::UndistPtsCam* d_undist_cams = nullptr;
    bool kps_undist_host = false;
    std::vector<mcs_calib> calib_host;
    // pinned host mirrors
    int *h_rows = nullptr, *h_row_count = nullptr, *h_lm_rows = nullptr, *h_n_views = nullptr, *h_rep = nullptr, *h_n_acc = nullptr,
        *h_match_tmp = nullptr;
    uint8_t* h_accepted = nullptr;
    double* h_xyz = nullptr;
    float* h_uv = nullptr;
    int n_extracted = 0, n_matched = 0, n_triangulated = 0;
    bool feat_host = false, rows_host = false, lm_host = false;
    size_t merge_smem = 0;
    int merge_wide = -1;  // MCS_MERGE_WIDE: -1 by launch size, 0 / 1 forced
    bool merge_rows_smem = false;
    cudaEvent_t t0 = nullptr, t1 = nullptr;
    // Latency mode (mcs_rig_set_graph): the whole launch sequence of an unchunked device-resident call — memsets, every kernel, the
    // forks to the side streams and their joins — captured once per (image pointer, mask pointer, frame count, epipolar flag) into a
    // CUDA graph and replayed with one cudaGraphLaunch.  FrontEnd::processFrame (FrontEnd.cpp:991-1028) runs once per rig frame, so
    // a serving loop calls with the same buffers every time.
    struct GraphEntry {
        const void* d_images = nullptr;
        const void* d_masks = nullptr;
        int n_frames = 0, old = 0;
        cudaGraphExec_t exec = nullptr;
        long long launches = 0;        // kernels in the graph
        mcs::ImageSet last_set{};
        int last_n = 0;
        bool blur_valid = false;
    };
    std::vector<GraphEntry> graphs;
    bool use_graph = false;
    long long graph_replays = 0;

    void drop_graphs() {
        for (GraphEntry& g : graphs)
            if (g.exec) cudaGraphExecDestroy(g.exec);
        graphs.clear();
    }

    ~mcs_rig() {
        if (!eng.stream) return;
        cudaSetDevice(eng.device);
        cudaStreamSynchronize(eng.stream);
        drop_graphs();
        if (t0) { cudaEventDestroy(t0); cudaEventDestroy(t1); }
        for (int k = 0; k < NSLOT; k++) {
            Slot& S = slots[k];
            if (S.st && k > 0) { cudaStreamSynchronize(S.st); cudaStreamDestroy(S.st); }
            if (S.st2) { cudaStreamSynchronize(S.st2); cudaStreamDestroy(S.st2); }
            if (S.ev_fork) { cudaEventDestroy(S.ev_fork); cudaEventDestroy(S.ev_join); cudaEventDestroy(S.ev_done); }
            if (S.hi.st) { cudaStreamSynchronize(S.hi.st); cudaStreamDestroy(S.hi.st); cudaEventDestroy(S.hi.e0); cudaEventDestroy(S.hi.e1); }
            cudaFree(S.d_tri_lists); cudaFree(S.d_tri_counts); cudaFree(S.d_f32);
        }
        if (ev_begin) cudaEventDestroy(ev_begin);
        cudaFree(d_calib); cudaFree(d_F); cudaFree(d_sigma2); cudaFree(d_match_train); cudaFree(d_match_tmp); cudaFree(d_rows);
        cudaFree(d_row_count); cudaFree(lm.rows); cudaFree(lm.accepted); cudaFree(lm.n_views); cudaFree(lm.xyz); cudaFree(lm.uv_ref);
        cudaFree(lm.rep_cam); cudaFree(lm.n_accepted); cudaFree(d_masks); cudaFree(d_und_maps); cudaFree(d_masks_und); cudaFree(d_kps_undist); cudaFree(d_undist_cams); cudaFreeHost(h_kps_undist);
        cudaFreeHost(h_rows); cudaFreeHost(h_row_count); cudaFreeHost(h_lm_rows); cudaFreeHost(h_n_views); cudaFreeHost(h_rep);
        cudaFreeHost(h_n_acc); cudaFreeHost(h_match_tmp); cudaFreeHost(h_accepted); cudaFreeHost(h_xyz); cudaFreeHost(h_uv);
    }

    int init(const mcs_rig_config* c, const mcs_calib* cal) {
        cfg = *c;
        C = c->num_cams; F = c->max_frames;
        if (C < 1 || C > MCS_MAX_CAMS || F < 1 || !cal) { set_err("num_cams must be in [1,8], max_frames >= 1, calibration non-null"); return MCS_ERR_ARG; }
        calib_host.assign(cal, cal + C);
        int rc = eng.init(c->nfeatures, c->scale_factor, c->nlevels, c->ini_th_fast, c->min_th_fast, c->device, c->width, c->height, C * F);
        if (rc) return rc;
        kp_cap = eng.plan.kp_cap;
        row_cap = (C * kp_cap + 1) / 2;  // every new IntraMatch row consumes two unmatched features
        if (row_cap >= 65535) { set_err("too many features per frame for the merge table"); return MCS_ERR_ARG; }
        calib.assign(cal, cal + C);
        pairs.n_pairs = 0;
        for (int i = 0; i < C - 1; i++)
            for (int j = i + 1; j < C; j++) {
                pairs.a[pairs.n_pairs] = (unsigned char)i;
                pairs.b[pairs.n_pairs] = (unsigned char)j;
                pairs.n_pairs++;
            }
        const int NP = std::max(pairs.n_pairs, 1);
        std::vector<mcs::CalibDev> cd(C);
        for (int k = 0; k < C; k++) {
            memcpy(cd[k].K, cal[k].K, sizeof(double) * 9);
            for (int i = 0; i < 3; i++) {  // build_Rt (common_utils/src/tools.cpp:136-148)
                for (int j = 0; j < 3; j++) cd[k].Rt[i * 4 + j] = cal[k].R[i * 3 + j];
                cd[k].Rt[i * 4 + 3] = cal[k].t[i];
            }
            for (int r = 0; r < 3; r++)
                for (int cc = 0; cc < 4; cc++) {
                    double s = 0;
                    for (int q = 0; q < 3; q++) s += cal[k].K[r * 3 + q] * cd[k].Rt[q * 4 + cc];
                    cd[k].P[r * 4 + cc] = s;
                }
        }
        std::vector<double> Fm((size_t)NP * 9, 0.0);
        for (int p = 0; p < pairs.n_pairs; p++) fundamental_from_extrinsics(cal[pairs.a[p]], cal[pairs.b[p]], &Fm[(size_t)p * 9]);
        CK(cudaMalloc(&d_calib, sizeof(mcs::CalibDev) * C));
        CK(cudaMemcpy(d_calib, cd.data(), sizeof(mcs::CalibDev) * C, cudaMemcpyHostToDevice));
        CK(cudaMalloc(&d_F, Fm.size() * sizeof(double)));
        CK(cudaMemcpy(d_F, Fm.data(), Fm.size() * sizeof(double), cudaMemcpyHostToDevice));
        CK(cudaMalloc(&d_sigma2, sizeof(float) * eng.T.nlevels));
        CK(cudaMemcpy(d_sigma2, eng.T.sigma2.data(), sizeof(float) * eng.T.nlevels, cudaMemcpyHostToDevice));
        const size_t mt = (size_t)F * NP * kp_cap, rw = (size_t)F * row_cap;
        CK(cudaMalloc(&d_match_train, mt * sizeof(int)));
        CK(cudaMalloc(&d_match_tmp, (size_t)kp_cap * 3 * sizeof(int)));
        CK(cudaMalloc(&d_rows, rw * C * sizeof(int)));
        CK(cudaMalloc(&d_row_count, F * sizeof(int)));
        tri_list_stride = (int)rw;
        CK(cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, eng.device));
        eng.sm_count_ = sm_count;
        if ((rc = set_tri_smem())) return rc;
        CK(cudaEventCreateWithFlags(&ev_begin, cudaEventDisableTiming));
        for (int k = 0; k < NSLOT; k++) {
            Slot& S = slots[k];
            if (k == 0) S.st = eng.stream;
            else CK(cudaStreamCreateWithFlags(&S.st, cudaStreamNonBlocking));
            CK(cudaStreamCreateWithFlags(&S.st2, cudaStreamNonBlocking));
            CK(cudaEventCreateWithFlags(&S.ev_fork, cudaEventDisableTiming));
            CK(cudaEventCreateWithFlags(&S.ev_join, cudaEventDisableTiming));
            CK(cudaEventCreateWithFlags(&S.ev_done, cudaEventDisableTiming));
            CK(cudaMalloc(&S.d_tri_lists, (size_t)(MCS_MAX_CAMS + 1) * rw * sizeof(int)));
            CK(cudaMalloc(&S.d_tri_counts, (MCS_MAX_CAMS + 2) * sizeof(int)));  // per view count + the tile queue head
        }
        if (const char* e = getenv("MCS_CHUNK_FRAMES")) chunk_frames = atoi(e);
        if (const char* e = getenv("MCS_HIPRIO")) use_hi = atoi(e) != 0;
        if (const char* e = getenv("MCS_MATCH_IMMA")) match_imma = atoi(e) != 0;
        if (const char* e = getenv("MCS_MERGE_WIDE")) merge_wide = atoi(e);
        if (const char* e = getenv("MCS_HOST_CHUNK_DIV")) host_chunk_div = std::max(1, std::min(16, atoi(e)));
        if (const char* e = getenv("MCS_TRI_PER_SM")) tri_per_sm = std::max(0, atoi(e));
        if (const char* e = getenv("MCS_TRI_LANES")) tri_lanes = atoi(e) == 1 ? 1 : (atoi(e) == 4 ? 4 : -1);
        if (const char* e = getenv("MCS_TRI4_PER_SM")) tri4_per_sm = std::max(1, std::min(32, atoi(e)));
        if (use_hi) {
            int lo = 0, hi = 0;
            CK(cudaDeviceGetStreamPriorityRange(&lo, &hi));
            for (int k = 0; k < NSLOT; k++) {
                CK(cudaStreamCreateWithPriority(&slots[k].hi.st, cudaStreamNonBlocking, hi));
                CK(cudaEventCreateWithFlags(&slots[k].hi.e0, cudaEventDisableTiming));
                CK(cudaEventCreateWithFlags(&slots[k].hi.e1, cudaEventDisableTiming));
            }
        }
        CK(cudaMalloc(&lm.rows, rw * C * sizeof(int)));
        CK(cudaMalloc(&lm.accepted, rw));
        CK(cudaMalloc(&lm.n_views, rw * sizeof(int)));
        CK(cudaMalloc(&lm.xyz, rw * 3 * sizeof(double)));
        CK(cudaMalloc(&lm.uv_ref, rw * 2 * sizeof(float)));
        CK(cudaMalloc(&lm.rep_cam, rw * sizeof(int)));
        CK(cudaMalloc(&lm.n_accepted, F * sizeof(int)));
        CK(cudaMallocHost(&h_rows, rw * C * sizeof(int)));
        CK(cudaMallocHost(&h_row_count, F * sizeof(int)));
        CK(cudaMallocHost(&h_lm_rows, rw * C * sizeof(int)));
        CK(cudaMallocHost(&h_n_views, rw * sizeof(int)));
        CK(cudaMallocHost(&h_rep, rw * sizeof(int)));
        CK(cudaMallocHost(&h_n_acc, F * sizeof(int)));
        CK(cudaMallocHost(&h_match_tmp, (size_t)kp_cap * 3 * sizeof(int)));
        CK(cudaMallocHost(&h_accepted, rw));
        CK(cudaMallocHost(&h_xyz, rw * 3 * sizeof(double)));
        CK(cudaMallocHost(&h_uv, rw * 2 * sizeof(float)));
        // match_inv_idx (u16) + claim tables + (when it fits) the IntraMatch row table itself
        if (kp_cap >= 65535 || row_cap >= 65535) { set_err("too many features per camera for the merge kernel"); return MCS_ERR_ARG; }
        merge_smem = mcs::merge_smem_layout(C, kp_cap, row_cap, false).total;
        if (merge_smem > 200 * 1024) { set_err("too many features per frame for the merge kernel's shared table"); return MCS_ERR_ARG; }
        merge_rows_smem = mcs::merge_smem_layout(C, kp_cap, row_cap, true).total <= 100 * 1024;
        if (merge_rows_smem) merge_smem = mcs::merge_smem_layout(C, kp_cap, row_cap, true).total;
        CK(cudaFuncSetAttribute(mcs::merge_kernel<MCS_MERGE_NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)std::max<size_t>(merge_smem, 1024)));
        CK(cudaFuncSetAttribute(mcs::merge_kernel<MCS_MERGE_NT_WIDE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)std::max<size_t>(merge_smem, 1024)));
        return 0;
    }

    static size_t tri_smem(int nv) { return nv <= 2 ? 0 : (size_t)(3 * nv * (4 + nv) + 4 * (4 + nv)) * sizeof(double) * MCS_TRI_NT; }
    int set_tri_smem() {
        CK(cudaFuncSetAttribute(mcs::tri_solve_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tri_smem(MCS_MAX_CAMS)));
        CK(cudaFuncSetAttribute(mcs::tri_solve4_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)mcs::tri4_smem_bytes(MCS_MAX_CAMS)));
        return 0;
    }
    void invalidate() { feat_host = rows_host = lm_host = kps_undist_host = false; }

    // The three stages for frames [f0, f0 + nf) of the workspace on slot `k` (0 = main stream).  Every kernel indexes frames
    // relative to the pointers it is given, so a chunk is the same launch with offset pointers.
    int stage_extract(const mcs::ImageSet& set, int n_frames, int f0 = 0, int k = 0) {
        static const int lap[2] = {0, 0};  // vLapping = {0,0} (MultiCameraFrame.cpp:235, :274)
        int rc = eng.run(set, n_frames * C, lap, f0 * C, slots[k].st, use_hi ? &slots[k].hi : nullptr);
        if (rc) return rc;
        n_extracted = f0 == 0 ? n_frames : std::max(n_extracted, f0 + n_frames);
        if (f0 == 0) n_matched = n_triangulated = 0;
        invalidate();
        return undistort_keypoints(n_frames, f0, slots[k].st);
    }
    // image_kps_undist of the frames [f0, f0 + n_frames)
    int undistort_keypoints(int n_frames, int f0, cudaStream_t st) {
        if (!undist_pts_on) return 0;
        dim3 g((kp_cap + 127) / 128, n_frames * C);
        mcs::undistort_points_kernel<<<g, 128, 0, st>>>(eng.d_kps + (size_t)f0 * C * kp_cap, d_kps_undist + (size_t)f0 * C * kp_cap,
                                                        eng.d_kp_count + (size_t)f0 * C, kp_cap, d_undist_cams, C, f0 * C);
        CK(cudaGetLastError());
        eng.launches++;
        kps_undist_host = false;
        return 0;
    }
    const mcs::KeypointPod* kps_undist_dev() const { return undist_pts_on ? d_kps_undist : eng.d_kps; }
    int set_undistort_keypoints(bool on) {
        if (on && !d_kps_undist) {
            std::vector<mcs::UndistPtsCam> cams(C);
            for (int c = 0; c < C; c++) {
                const mcs_calib& k = calib_host[c];
                mcs::UndistPtsCam u{};
                u.fx = (double)(float)k.K[0]; u.fy = (double)(float)k.K[4]; u.cx = (double)(float)k.K[2]; u.cy = (double)(float)k.K[5];
                for (int i = 0; i < 5; i++) u.k[i] = (double)(float)k.dist[i];
                uint64_t bits;
                memcpy(&bits, &k.dist[0], 8);
                float lo;
                const uint32_t lo_bits = (uint32_t)bits;  // `.at<float>(0)` of a CV_64F matrix (:302)
                memcpy(&lo, &lo_bits, 4);
                u.copy_cam = lo == 0.0f ? 1 : 0;
                cams[c] = u;
            }
            CK(cudaMalloc(&d_undist_cams, sizeof(mcs::UndistPtsCam) * C));
            CK(cudaMemcpy(d_undist_cams, cams.data(), sizeof(mcs::UndistPtsCam) * C, cudaMemcpyHostToDevice));
            CK(cudaMalloc(&d_kps_undist, (size_t)F * C * kp_cap * sizeof(mcs::KeypointPod)));
            CK(cudaMallocHost(&h_kps_undist, (size_t)F * C * kp_cap * sizeof(mcs::KeypointPod)));
        }
        undist_pts_on = on;
        kps_undist_host = false;
        return 0;
    }
    int ensure_kps_undist() {
        if (!undist_pts_on || kps_undist_host) return 0;
        if (n_extracted < 1) { set_err("no extracted frames"); return MCS_ERR_STATE; }
        CK(cudaMemcpyAsync(h_kps_undist, d_kps_undist, (size_t)n_extracted * C * kp_cap * sizeof(mcs::KeypointPod), cudaMemcpyDeviceToHost, eng.stream));
        CK(cudaStreamSynchronize(eng.stream));
        kps_undist_host = true;
        return 0;
    }
    int stage_match(int n_frames, int old, float dist_thresh, float ratio, int f0 = 0, int k = 0) {
        if (n_frames < 1 || f0 + n_frames > n_extracted) { set_err("intra-match before extract"); return MCS_ERR_STATE; }
        cudaStream_t st = slots[k].st;
        const int NP = std::max(pairs.n_pairs, 1);
        int* match_train = d_match_train + (size_t)f0 * NP * kp_cap;
        eng.mark(ST_MATCH);
        if (pairs.n_pairs > 0) {
            dim3 g((kp_cap + MCS_MATCH_NT - 1) / MCS_MATCH_NT, pairs.n_pairs, n_frames);
            if (match_imma) {  // experiment (MCS_MATCH_IMMA=1): distances from the u8 integer MMA, same outputs
                dim3 gi((kp_cap + MCS_IMMA_NT / 2 - 1) / (MCS_IMMA_NT / 2), pairs.n_pairs, n_frames);
                mcs::hamming_match_imma_kernel<<<gi, MCS_IMMA_NT, 0, st>>>(C, kp_cap, pairs, f0, eng.d_desc, eng.d_kp_count, dist_thresh, ratio,
                                                                           match_train, nullptr, nullptr);
            } else
            mcs::hamming_match_kernel<<<g, MCS_MATCH_NT, MCS_MATCH_CHUNK * 32, st>>>(C, kp_cap, pairs, f0, eng.d_desc, eng.d_kp_count, dist_thresh,
                                                                                    ratio, match_train, nullptr, nullptr);
            eng.launches++;
        }
        eng.mark(ST_MERGE);
        cudaStream_t ml = st;
        if (use_hi) {
            CK(cudaEventRecord(slots[k].hi.e0, st));
            CK(cudaStreamWaitEvent(slots[k].hi.st, slots[k].hi.e0, 0));
            ml = slots[k].hi.st;
        }
        // a launch of few frames cannot fill the SMs: wide blocks (one query per thread) shorten its latency instead
        const bool wide = merge_wide < 0 ? n_frames * 4 <= sm_count : merge_wide != 0;
        const auto mk = wide ? &mcs::merge_kernel<MCS_MERGE_NT_WIDE> : &mcs::merge_kernel<MCS_MERGE_NT>;
        mk<<<n_frames, wide ? MCS_MERGE_NT_WIDE : MCS_MERGE_NT, merge_smem, ml>>>(C, kp_cap, row_cap, pairs, eng.d_kp_count + (size_t)f0 * C, match_train,
                                                                              kps_undist_dev() + (size_t)f0 * C * kp_cap, old, d_F, d_sigma2,
                                                                              d_rows + (size_t)f0 * row_cap * C, d_row_count + f0, merge_rows_smem ? 1 : 0);
        eng.launches++;
        if (use_hi) {
            CK(cudaEventRecord(slots[k].hi.e1, ml));
            CK(cudaStreamWaitEvent(st, slots[k].hi.e1, 0));
        }
        eng.mark(ST_TRIANG);
        CK(cudaGetLastError());
        n_matched = f0 == 0 ? n_frames : std::max(n_matched, f0 + n_frames);
        if (f0 == 0) n_triangulated = 0;
        rows_host = lm_host = false;
        return 0;
    }
    mcs::LandmarkOut lm_at(int f0) const {
        mcs::LandmarkOut o = lm;
        const size_t r = (size_t)f0 * row_cap;
        o.rows += r * C; o.accepted += r; o.n_views += r; o.xyz += r * 3; o.uv_ref += r * 2; o.rep_cam += r; o.n_accepted += f0;
        return o;
    }
    // d_mask_ptr: masks of frame 0 of the workspace (f32, or u8 decisions when mask_u8), or null
    int stage_triangulate(int n_frames, const float* d_mask_ptr, int f0 = 0, int k = 0, bool mask_u8 = false) {
        if (n_frames < 1 || f0 + n_frames > n_matched) { set_err("triangulate before intra-match"); return MCS_ERR_STATE; }
        Slot& S = slots[k];
        const mcs::LandmarkOut out = lm_at(f0);
        const mcs::KeypointPod* kps = eng.d_kps + (size_t)f0 * C * kp_cap;
        const uint8_t* desc = eng.d_desc + (size_t)f0 * C * kp_cap * 32;
        const float* mask = nullptr;
        if (d_mask_ptr) {
            const size_t ofs = (size_t)f0 * C * cfg.width * cfg.height;
            mask = mask_u8 ? reinterpret_cast<const float*>(reinterpret_cast<const uint8_t*>(d_mask_ptr) + ofs) : d_mask_ptr + ofs;
        }
        eng.mark(ST_TRIANG);
        cudaStream_t tl = S.st;  // the stream the triangulation kernels run on
        if (use_hi) {
            CK(cudaEventRecord(S.hi.e0, S.st));
            CK(cudaStreamWaitEvent(S.hi.st, S.hi.e0, 0));
            tl = S.hi.st;
        }
        CK(cudaMemsetAsync(out.n_accepted, 0, n_frames * sizeof(int), tl));
        CK(cudaMemsetAsync(S.d_tri_counts, 0, (MCS_MAX_CAMS + 2) * sizeof(int), tl));
        dim3 g((row_cap + 127) / 128, n_frames);
        mcs::tri_classify_kernel<MCS_MAX_CAMS><<<g, 128, 0, tl>>>(C, kp_cap, row_cap, cfg.width, cfg.height, d_rows + (size_t)f0 * row_cap * C,
                                                                    d_row_count + f0, kps, mask, mask_u8 ? 1 : 0, out, S.d_tri_lists, S.d_tri_counts,
                                                                    tri_list_stride);
        eng.launches++;
        // 2-view rows on a forked stream, every larger view count in one launch (blockIdx.y = class) on the slot's stream;
        // persistent grid-stride blocks because the list lengths live on the device
        if (C >= 2) {
            CK(cudaEventRecord(S.ev_fork, tl));
            CK(cudaStreamWaitEvent(S.st2, S.ev_fork, 0));
            const int nb2 = std::max(1, std::min(sm_count * 8, (n_frames * row_cap + 127) / 128));
            mcs::tri_solve2_kernel<<<nb2, 128, 0, S.st2>>>(C, kp_cap, row_cap, d_calib, kps, desc, out, S.d_tri_lists, S.d_tri_counts, tri_list_stride);
            eng.launches++;
            CK(cudaEventRecord(S.ev_join, S.st2));
            if (C >= 3) {
                // Four lanes per row shorten the chain of one solve (one rig frame: triangulation stage 0.141 -> 0.102 ms) but repeat the
                // scalar part of every rotation in four lanes: with more tiles than SMs that costs more than it saves (8 frames 0.148 ->
                // 0.194 ms, 64-frame chunks -2 % throughput).  So: calls of one or two frames only, unless MCS_TRI_LANES says otherwise.
                if (tri_lanes == 4 || (tri_lanes < 0 && n_frames <= 2)) {
                    // 8-row tiles, one warp per block: as many blocks as there can be tiles, up to 16 per SM
                    const size_t sm4 = mcs::tri4_smem_bytes(C);
                    const int per_sm = std::max(1, std::min(tri4_per_sm, (int)((200 * 1024) / (sm4 + 1024))));
                    const int tiles_max = (n_frames * row_cap + MCS_TRI4_ROWS - 1) / MCS_TRI4_ROWS + MCS_MAX_CAMS;
                    mcs::tri_solve4_kernel<<<std::min(sm_count * per_sm, tiles_max), 32, sm4, tl>>>(C, kp_cap, row_cap, d_calib, kps, desc, out, S.d_tri_lists,
                                                                                                   S.d_tri_counts, tri_list_stride,
                                                                                                   S.d_tri_counts + MCS_MAX_CAMS + 1);
                } else {
                    const int per_sm = std::max(1, std::min(tri_per_sm > 0 ? tri_per_sm : 16, (int)((227 * 1024) / (tri_smem(C) + 1024))));
                    mcs::tri_solve_kernel<<<sm_count * per_sm, MCS_TRI_NT, tri_smem(C), tl>>>(C, kp_cap, row_cap, d_calib, kps, desc, out, S.d_tri_lists,
                                                                                              S.d_tri_counts, tri_list_stride,
                                                                                              S.d_tri_counts + MCS_MAX_CAMS + 1);
                }
                eng.launches++;
            }
            CK(cudaStreamWaitEvent(tl, S.ev_join, 0));
        }
        if (use_hi) {
            CK(cudaEventRecord(S.hi.e1, tl));
            CK(cudaStreamWaitEvent(S.st, S.hi.e1, 0));
        }
        eng.mark(ST_DOWNLOAD);
        CK(cudaGetLastError());
        n_triangulated = f0 == 0 ? n_frames : std::max(n_triangulated, f0 + n_frames);
        lm_host = false;
        return 0;
    }
    // Result downloads are sized by counts, not capacities: the row tables hold up to row_cap = (num_cams * kp_cap + 1) / 2 rows per frame,
    // a frame fills well under half of that.  The copies are queued before the counts are known on the host, so each array goes down as
    // ONE strided copy per chunk (cudaMemcpy2DAsync: nf frames x the first `rows` rows of each frame's slab) with rows = rows_est, an
    // estimate kept from the previous call (largest row count seen + 25 % + 64; the full capacity until a call has finished).  A frame
    // that exceeds the estimate gets its remaining rows with a second, synchronous copy after the counts have arrived (top_up_rows).
    long long d2h_bytes = 0;  // bytes queued for download by the last process call (bench.py's e2e accounting)
    int rows_est = 0;         // 0 = not known yet: copy whole slabs
    int rows_sent = 0;        // rows per frame the last process call downloaded
    template <typename T>
    int copy_rows_2d(T* h, const T* d, size_t per_row, int f0, int nf, int rows, cudaStream_t st) {
        const size_t pitch = (size_t)row_cap * per_row * sizeof(T), o = (size_t)f0 * row_cap * per_row;
        CK(cudaMemcpy2DAsync(h + o, pitch, d + o, pitch, (size_t)rows * per_row * sizeof(T), nf, cudaMemcpyDeviceToHost, st));
        d2h_bytes += (long long)rows * per_row * sizeof(T) * nf;
        return 0;
    }
    int download_rows_async(int f0, int nf, cudaStream_t st, int rows = -1) {
        if (rows < 0) rows = row_cap;
        CK(cudaMemcpyAsync(h_row_count + f0, d_row_count + f0, nf * sizeof(int), cudaMemcpyDeviceToHost, st));
        d2h_bytes += nf * sizeof(int);
        return copy_rows_2d(h_rows, d_rows, C, f0, nf, rows, st);
    }
    int download_landmarks_async(int f0, int nf, cudaStream_t st, int rows = -1) {
        if (rows < 0) rows = row_cap;
        int rc;
        CK(cudaMemcpyAsync(h_row_count + f0, d_row_count + f0, nf * sizeof(int), cudaMemcpyDeviceToHost, st));
        CK(cudaMemcpyAsync(h_n_acc + f0, lm.n_accepted + f0, nf * sizeof(int), cudaMemcpyDeviceToHost, st));
        d2h_bytes += 2 * nf * sizeof(int);
        if ((rc = copy_rows_2d(h_lm_rows, lm.rows, C, f0, nf, rows, st))) return rc;
        if ((rc = copy_rows_2d(h_accepted, lm.accepted, 1, f0, nf, rows, st))) return rc;
        if ((rc = copy_rows_2d(h_n_views, lm.n_views, 1, f0, nf, rows, st))) return rc;
        if ((rc = copy_rows_2d(h_xyz, lm.xyz, 3, f0, nf, rows, st))) return rc;
        if ((rc = copy_rows_2d(h_uv, lm.uv_ref, 2, f0, nf, rows, st))) return rc;
        return copy_rows_2d(h_rep, lm.rep_cam, 1, f0, nf, rows, st);
    }
    // after the counts are on the host: fetch what a too-small estimate left behind, then re-estimate for the next call
    int top_up_rows(int n_frames) {
        int max_rows = 0;
        for (int f = 0; f < n_frames; f++) max_rows = std::max(max_rows, std::min(h_row_count[f], row_cap));
        if (max_rows > rows_sent) {
            for (int f = 0; f < n_frames; f++) {
                const int M = std::min(h_row_count[f], row_cap);
                if (M <= rows_sent) continue;
                const size_t o = (size_t)f * row_cap + rows_sent, n = (size_t)(M - rows_sent);
                CK(cudaMemcpyAsync(h_rows + o * C, d_rows + o * C, n * C * sizeof(int), cudaMemcpyDeviceToHost, eng.stream));
                CK(cudaMemcpyAsync(h_lm_rows + o * C, lm.rows + o * C, n * C * sizeof(int), cudaMemcpyDeviceToHost, eng.stream));
                CK(cudaMemcpyAsync(h_accepted + o, lm.accepted + o, n, cudaMemcpyDeviceToHost, eng.stream));
                CK(cudaMemcpyAsync(h_n_views + o, lm.n_views + o, n * sizeof(int), cudaMemcpyDeviceToHost, eng.stream));
                CK(cudaMemcpyAsync(h_xyz + o * 3, lm.xyz + o * 3, n * 3 * sizeof(double), cudaMemcpyDeviceToHost, eng.stream));
                CK(cudaMemcpyAsync(h_uv + o * 2, lm.uv_ref + o * 2, n * 2 * sizeof(float), cudaMemcpyDeviceToHost, eng.stream));
                CK(cudaMemcpyAsync(h_rep + o, lm.rep_cam + o, n * sizeof(int), cudaMemcpyDeviceToHost, eng.stream));
                d2h_bytes += (long long)n * (2 * C * 4 + 1 + 4 + 24 + 8 + 4);
            }
            CK(cudaStreamSynchronize(eng.stream));
        }
        rows_est = std::min(row_cap, max_rows + max_rows / 4 + 64);
        return 0;
    }
    // RECTIFY: the segmentation masks are undistorted with the same maps (MultiCameraFrame.cpp:131-134); *dm is replaced
    int undistort_masks(const float** dm, int n_frames) {
        if (!undistort_on || !*dm) return 0;
        if (n_frames < 1 || n_frames > F) { set_err("n_frames exceeds the handle's capacity"); return MCS_ERR_CAPACITY; }
        const int w = cfg.width, h = cfg.height;
        const size_t need = (size_t)n_frames * C * w * h * sizeof(float);
        if (need > masks_und_bytes) {
            CK(cudaStreamSynchronize(eng.stream));
            cudaFree(d_masks_und);
            d_masks_und = nullptr; masks_und_bytes = 0;
            CK(cudaMalloc(&d_masks_und, (size_t)F * C * w * h * sizeof(float)));
            masks_und_bytes = (size_t)F * C * w * h * sizeof(float);
        }
        dim3 b(32, 8), g((w + 31) / 32, (h + 7) / 8, n_frames * C);
        mcs::undistort_f32_kernel<<<g, b, 0, eng.stream>>>(*dm, d_masks_und, d_und_maps, C, w, h);
        CK(cudaGetLastError());
        eng.launches++;
        *dm = d_masks_und;
        return 0;
    }
    // image set of the tight device images [img0, img0 + n) of `d_images` (image 0 = base), undistorted first when RECTIFY is on
    int make_set(const uint8_t* d_images, int n, mcs::ImageSet& set, int img0, cudaStream_t st) {
        if (undistort_on) return eng.make_set_undistorted(d_images, n, cfg.width, cfg.height, d_und_maps, C, set, img0, st);
        return eng.make_set(d_images, n, cfg.width, cfg.height, set, img0, st);
    }
    int set_undistort(bool on) {
        if (on && !d_und_maps) {
            const size_t px = (size_t)cfg.width * cfg.height;
            std::vector<uint2> maps(px * C);
            for (int c = 0; c < C; c++) build_undistort_map(calib_host[c], cfg.width, cfg.height, maps.data() + px * c);
            CK(cudaMalloc(&d_und_maps, maps.size() * sizeof(uint2)));
            CK(cudaMemcpy(d_und_maps, maps.data(), maps.size() * sizeof(uint2), cudaMemcpyHostToDevice));
        }
        undistort_on = on;
        invalidate();
        return 0;
    }
    int upload_masks(const float* masks, int n_frames, const float** out) {
        *out = nullptr;
        if (!masks) return 0;
        if (n_frames < 1 || n_frames > F) { set_err("n_frames exceeds the handle's capacity"); return MCS_ERR_CAPACITY; }
        const size_t need = (size_t)n_frames * C * cfg.width * cfg.height * sizeof(float);
        if (need > masks_bytes) {
            CK(cudaStreamSynchronize(eng.stream));
            cudaFree(d_masks);
            d_masks = nullptr;
            CK(cudaMalloc(&d_masks, (size_t)F * C * cfg.width * cfg.height * sizeof(float)));
            masks_bytes = (size_t)F * C * cfg.width * cfg.height * sizeof(float);
        }
        CK(cudaMemcpyAsync(d_masks, masks, need, cudaMemcpyHostToDevice, eng.stream));
        *out = d_masks;
        return 0;
    }
    int ensure_features() {
        if (feat_host) return 0;
        if (n_extracted < 1) { set_err("no extracted frames"); return MCS_ERR_STATE; }
        int rc = eng.download_features(n_extracted * C);
        if (rc) return rc;
        feat_host = true;
        return 0;
    }
    int ensure_rows() {
        if (rows_host) return 0;
        if (n_matched < 1) { set_err("no intra-matches computed"); return MCS_ERR_STATE; }
        int rc = download_rows_async(0, n_matched, eng.stream);
        if (rc) return rc;
        CK(cudaStreamSynchronize(eng.stream));
        return check_rows();
    }
    int check_rows() {
        for (int f = 0; f < n_matched; f++)
            if (h_row_count[f] > row_cap) { set_err("IntraMatch row table overflow"); return MCS_ERR_CAPACITY; }
        rows_host = true;
        return 0;
    }
    int ensure_landmarks() {
        if (lm_host) return 0;
        if (n_triangulated < 1) { set_err("no triangulation computed"); return MCS_ERR_STATE; }
        int rc = download_landmarks_async(0, n_triangulated, eng.stream);
        if (rc) return rc;
        CK(cudaStreamSynchronize(eng.stream));
        lm_host = true;
        return 0;
    }

    // Whole path for n_frames frames.  Large batches are cut into chunks of chunk_frames that rotate over NSLOT streams, so the
    // upload of one chunk, the kernels of another and the download of a third overlap, and the latency-bound kernels (octree,
    // merge, Jacobi) of one chunk hide behind the throughput-bound ones of its neighbours.  Everything is ordered against the
    // main stream (fork at ev_begin, join at ev_done), so timers and synchronisation on the main stream cover all of it.
    // Masks: f32 segmentation masks (the reference's segMasks) or, with mask_u8, one byte per pixel holding the caller's decision
    // (0 = static, keep; non-zero = dynamic, i.e. mask >= 0.7) — a quarter of the PCIe traffic for the same 1-bit answer.  Host masks
    // are uploaded (and, with RECTIFY, undistorted) chunk by chunk on the chunk's own stream, like the images.
    int process(const uint8_t* h_images, const uint8_t* d_images, const float* h_masks, const float* d_masks_in, int n_frames, int old,
                bool download, const float* h_f32 = nullptr, const float* d_f32 = nullptr, int channels = 1, bool mask_u8 = false) {
        const int w = cfg.width, h = cfg.height;
        const size_t f32_frame = (size_t)C * w * h * channels;  // floats per rig frame
        const size_t mask_frame = (size_t)C * w * h * (mask_u8 ? 1 : sizeof(float));  // mask bytes per rig frame
        const float* dm = d_masks_in;
        int rc;
        if (mask_u8 && undistort_on && (h_masks || d_masks_in)) { set_err("u8 mask decisions cannot be undistorted: pass f32 masks with RECTIFY"); return MCS_ERR_ARG; }
        if (h_masks) {  // device staging for the host masks (sized for f32, the larger form)
            const size_t need = (size_t)F * C * w * h * sizeof(float);
            if (need > masks_bytes) {
                CK(cudaStreamSynchronize(eng.stream));
                cudaFree(d_masks);
                d_masks = nullptr; masks_bytes = 0;
                CK(cudaMalloc(&d_masks, need));
                masks_bytes = need;
            }
            dm = d_masks;
        }
        if (undistort_on && dm) {
            const size_t need = (size_t)F * C * w * h * sizeof(float);
            if (need > masks_und_bytes) {
                CK(cudaStreamSynchronize(eng.stream));
                cudaFree(d_masks_und);
                d_masks_und = nullptr; masks_und_bytes = 0;
                CK(cudaMalloc(&d_masks_und, need));
                masks_und_bytes = need;
            }
        }
        // host-buffer calls use smaller chunks (chunk_frames / host_chunk_div): more copies in flight hide more of the PCIe time
        const int cf = (h_images || h_f32) ? std::max(1, chunk_frames / host_chunk_div) : chunk_frames;
        const bool chunked = chunk_frames > 0 && n_frames > cf && !eng.profiling;
        const int step = chunked ? cf : n_frames;
        if (download) {
            d2h_bytes = 0;
            rows_sent = rows_est > 0 ? rows_est : row_cap;
        }
        mcs::ImageSet set{};
        // latency mode: replay (or capture, the first time) the launch sequence of this exact call
        const bool graphable = use_graph && d_images && !h_masks && !undistort_on && !chunked && !eng.profiling && !download;
        GraphEntry* ge = nullptr;
        long long launches_before = 0;
        if (graphable) {
            for (GraphEntry& g : graphs)
                if (g.d_images == d_images && g.d_masks == dm && g.n_frames == n_frames && g.old == (old | (mask_u8 ? 2 : 0))) ge = &g;
            if (ge) {
                CK(cudaGraphLaunch(ge->exec, eng.stream));
                eng.launches += ge->launches;
                eng.last_set = ge->last_set; eng.last_n = ge->last_n; eng.blur_valid = ge->blur_valid;
                n_extracted = n_matched = n_triangulated = n_frames;
                invalidate();
                graph_replays++;
                return 0;
            }
            if (graphs.size() >= 16) drop_graphs();  // a caller that keeps changing buffers gets no benefit; bound the cache
            launches_before = eng.launches;
            CK(cudaStreamBeginCapture(eng.stream, cudaStreamCaptureModeThreadLocal));
        }
        auto end_capture = [&](int rc_in) -> int {  // always leaves capture mode; instantiates and launches on success
            cudaGraph_t graph = nullptr;
            const cudaError_t e = cudaStreamEndCapture(eng.stream, &graph);
            if (rc_in) { if (graph) cudaGraphDestroy(graph); return rc_in; }
            if (e != cudaSuccess || !graph) { set_err(std::string("cudaStreamEndCapture: ") + cudaGetErrorString(e)); cudaGetLastError(); return MCS_ERR_CUDA; }
            GraphEntry g;
            g.d_images = d_images; g.d_masks = dm; g.n_frames = n_frames; g.old = old | (mask_u8 ? 2 : 0);
            const cudaError_t ei = cudaGraphInstantiate(&g.exec, graph, 0);
            cudaGraphDestroy(graph);
            if (ei != cudaSuccess) { set_err(std::string("cudaGraphInstantiate: ") + cudaGetErrorString(ei)); cudaGetLastError(); return MCS_ERR_CUDA; }
            g.launches = eng.launches - launches_before;
            g.last_set = eng.last_set; g.last_n = eng.last_n; g.blur_valid = eng.blur_valid;
            graphs.push_back(g);
            CK(cudaGraphLaunch(g.exec, eng.stream));
            return 0;
        };
        if (chunked) {
            CK(cudaEventRecord(ev_begin, eng.stream));
            for (int k = 1; k < NSLOT; k++) CK(cudaStreamWaitEvent(slots[k].st, ev_begin, 0));
        }
        // Host-buffer calls ramp the chunk size up (step/4, step/2, step, ...): the kernels of the first chunk start after a quarter
        // of the upload time, and that chunk's latency-bound kernels overlap the upload of the next ones.
        int ci = 0, nf = 0;
        const bool ramp = chunked && (h_images || h_f32) && step >= 8;
        for (int f0 = 0; f0 < n_frames; f0 += nf, ci++) {
            const int want = !ramp ? step : (ci == 0 ? step / 4 : (ci == 1 ? step / 2 : step));
            nf = std::min(want, n_frames - f0);
            const int k = chunked ? ci % NSLOT : 0;
            cudaStream_t st = slots[k].st;
            if (h_f32) {
                Slot& S = slots[k];
                const size_t bytes = (size_t)nf * f32_frame * sizeof(float);
                if (bytes > S.f32_bytes) {
                    CK(cudaStreamSynchronize(st));
                    cudaFree(S.d_f32);
                    S.d_f32 = nullptr; S.f32_bytes = 0;
                    CK(cudaMalloc(&S.d_f32, (size_t)step * f32_frame * sizeof(float)));
                    S.f32_bytes = (size_t)step * f32_frame * sizeof(float);
                }
                eng.mark(ST_UPLOAD);
                CK(cudaMemcpyAsync(S.d_f32, h_f32 + (size_t)f0 * f32_frame, bytes, cudaMemcpyHostToDevice, st));
                rc = eng.stage_f32(S.d_f32, channels, nf * C, w, h, set, f0 * C, st);
            } else if (d_f32) rc = eng.stage_f32(d_f32 + (size_t)f0 * f32_frame, channels, nf * C, w, h, set, f0 * C, st);
            else if (h_images) rc = eng.upload(h_images + (size_t)f0 * C * w * h, nf * C, w, h, w, (size_t)w * h, set, f0 * C, st);
            else rc = make_set(d_images, nf * C, set, f0 * C, st);
            if (!rc && undistort_on && (h_f32 || d_f32 || h_images))  // staged tight in d_input: undistort from there
                rc = eng.make_set_undistorted(eng.d_input, nf * C, w, h, d_und_maps, C, set, f0 * C, st);
            const float* dmk = dm;  // masks of frame 0 as this chunk's triangulation reads them
            if (!rc && h_masks)
                CK(cudaMemcpyAsync(reinterpret_cast<uint8_t*>(d_masks) + (size_t)f0 * mask_frame, reinterpret_cast<const uint8_t*>(h_masks) + (size_t)f0 * mask_frame,
                                   (size_t)nf * mask_frame, cudaMemcpyHostToDevice, st));
            if (!rc && undistort_on && dm) {  // RECTIFY: the masks go through the same maps (MultiCameraFrame.cpp:131-134)
                const size_t ofs = (size_t)f0 * C * w * h;
                dim3 b(32, 8), g((w + 31) / 32, (h + 7) / 8, nf * C);
                mcs::undistort_f32_kernel<<<g, b, 0, st>>>(dm + ofs, d_masks_und + ofs, d_und_maps, C, w, h);
                eng.launches++;
                dmk = d_masks_und;
            }
            if (!rc) rc = stage_extract(set, nf, f0, k);
            // computeIntraMatches(matches_map, old): BruteForceMatch(i, j, 75, 0.85, ...) (MultiCameraFrame.cpp:1118)
            if (!rc) rc = stage_match(nf, old, 75.f, 0.85f, f0, k);
            if (!rc) rc = stage_triangulate(nf, dmk, f0, k, mask_u8);
            if (graphable) return end_capture(rc);  // one unchunked pass; nothing is downloaded in this mode
            if (rc) return rc;
            if (download) {
                if ((rc = eng.download_features_async(f0 * C, nf * C, st))) return rc;
                d2h_bytes += (long long)nf * C * (8 + 8 * eng.plan.nlevels + (long long)kp_cap * (sizeof(mcs::KeypointPod) + 32));
                if ((rc = download_rows_async(f0, nf, st, rows_sent))) return rc;
                if ((rc = download_landmarks_async(f0, nf, st, rows_sent))) return rc;
            }
        }
        if (chunked) {
            for (int k = 1; k < NSLOT; k++) {
                CK(cudaEventRecord(slots[k].ev_done, slots[k].st));
                CK(cudaStreamWaitEvent(eng.stream, slots[k].ev_done, 0));
            }
        }
        if (download) {
            eng.mark(ST_COUNT);
            CK(cudaStreamSynchronize(eng.stream));
            eng.collect_profile();
            if ((rc = eng.check_counts(n_frames * C))) return rc;
            feat_host = true;
            if ((rc = check_rows())) return rc;
            if ((rc = top_up_rows(n_frames))) return rc;
            lm_host = true;
        }
        return 0;
    }
};

extern "C" {

mcs_rig* mcs_rig_create(const mcs_rig_config* cfg, const mcs_calib* calib) {
    if (!cfg || !calib) { set_err("null config/calib"); return nullptr; }
    mcs_rig* r = new mcs_rig();
    if (r->init(cfg, calib) != 0) {
        delete r;
        return nullptr;
    }
    return r;
}
void mcs_rig_destroy(mcs_rig* rig) { delete rig; }
int mcs_rig_feature_capacity(const mcs_rig* rig) { return rig ? rig->kp_cap : MCS_ERR_ARG; }
int mcs_rig_row_capacity(const mcs_rig* rig) { return rig ? rig->row_cap : MCS_ERR_ARG; }
int mcs_rig_set_orientation(mcs_rig* rig, int enable) {
    if (!rig) return MCS_ERR_ARG;
    cudaStreamSynchronize(rig->eng.stream);
    rig->drop_graphs();
    rig->eng.use_angle = enable ? 1 : 0;
    return rig->eng.replan(rig->cfg.width, rig->cfg.height);
}

#define RIG_ENTER(rig)                                                  \
    if (!(rig)) return MCS_ERR_ARG;                                     \
    if (cudaSetDevice((rig)->eng.device) != cudaSuccess) return MCS_ERR_CUDA;

int mcs_rig_extract(mcs_rig* rig, const uint8_t* images, int n_frames) {
    RIG_ENTER(rig);
    if (!images || n_frames < 1) return MCS_ERR_EMPTY;
    if (n_frames > rig->F) { set_err("n_frames exceeds max_frames"); return MCS_ERR_CAPACITY; }
    rig->eng.begin_profile();
    mcs::ImageSet set;
    const int w = rig->cfg.width, h = rig->cfg.height;
    int rc = rig->eng.upload(images, n_frames * rig->C, w, h, w, (size_t)w * h, set);
    if (!rc && rig->undistort_on) rc = rig->eng.make_set_undistorted(rig->eng.d_input, n_frames * rig->C, w, h, rig->d_und_maps, rig->C, set, 0, rig->eng.stream);
    if (rc) return rc;
    return rig->stage_extract(set, n_frames);
}
int mcs_rig_extract_device(mcs_rig* rig, const void* d_images, int n_frames) {
    RIG_ENTER(rig);
    if (!d_images || n_frames < 1) return MCS_ERR_EMPTY;
    if (n_frames > rig->F) { set_err("n_frames exceeds max_frames"); return MCS_ERR_CAPACITY; }
    rig->eng.begin_profile();
    mcs::ImageSet set;
    int rc = rig->make_set((const uint8_t*)d_images, n_frames * rig->C, set, 0, rig->eng.stream);
    if (rc) return rc;
    return rig->stage_extract(set, n_frames);
}

int mcs_rig_get_features(mcs_rig* rig, int frame, int cam, mcs_keypoint* keypoints, uint8_t* descriptors, int capacity) {
    RIG_ENTER(rig);
    if (frame < 0 || frame >= rig->n_extracted || cam < 0 || cam >= rig->C) return MCS_ERR_ARG;
    int rc = rig->ensure_features();
    if (rc) return rc;
    const int img = frame * rig->C + cam, n = rig->eng.h_kp_count[img];
    if (n > capacity) { set_err("keypoint capacity too small"); return MCS_ERR_CAPACITY; }
    if (keypoints) memcpy(keypoints, rig->eng.h_kps + (size_t)img * rig->kp_cap, (size_t)n * sizeof(mcs_keypoint));
    if (descriptors) memcpy(descriptors, rig->eng.h_desc + (size_t)img * rig->kp_cap * 32, (size_t)n * 32);
    return n;
}

// Features computed elsewhere (another extractor, a test) for the rig frames [0, n_frames): replaces stage A for the stages that
// follow.  counts[n_frames * num_cams]; keypoints / descriptors: the cameras' features back to back in image order.
int mcs_rig_set_features(mcs_rig* rig, int n_frames, const int* counts, const mcs_keypoint* keypoints, const uint8_t* descriptors) {
    RIG_ENTER(rig);
    if (!counts || !keypoints || !descriptors || n_frames < 1) return MCS_ERR_ARG;
    if (n_frames > rig->F) { set_err("n_frames exceeds max_frames"); return MCS_ERR_CAPACITY; }
    const int n_img = n_frames * rig->C, cap = rig->kp_cap;
    cudaStream_t st = rig->eng.stream;
    size_t ofs = 0;
    for (int i = 0; i < n_img; i++) {
        const int n = counts[i];
        if (n < 0 || n > cap) { set_err("set_features: more features than the handle's capacity"); return MCS_ERR_CAPACITY; }
        if (n) {
            CK(cudaMemcpyAsync(rig->eng.d_kps + (size_t)i * cap, keypoints + ofs, (size_t)n * sizeof(mcs_keypoint), cudaMemcpyHostToDevice, st));
            CK(cudaMemcpyAsync(rig->eng.d_desc + (size_t)i * cap * 32, descriptors + ofs * 32, (size_t)n * 32, cudaMemcpyHostToDevice, st));
        }
        ofs += n;
    }
    CK(cudaMemcpyAsync(rig->eng.d_kp_count, counts, (size_t)n_img * sizeof(int), cudaMemcpyHostToDevice, st));
    CK(cudaStreamSynchronize(st));
    rig->n_extracted = n_frames;
    rig->n_matched = rig->n_triangulated = 0;
    rig->eng.last_n = n_img;
    rig->invalidate();
    return rig->undistort_keypoints(n_frames, 0, st);
}

int mcs_rig_brute_force_match(mcs_rig* rig, int frame, int cam1, int cam2, float dist_thresh, float neigh_ratio, uint32_t* indices_1,
                              uint32_t* indices_2, mcs_keypoint* kps1, mcs_keypoint* kps2, int capacity) {
    RIG_ENTER(rig);
    if (frame < 0 || frame >= rig->n_extracted || cam1 < 0 || cam1 >= rig->C || cam2 < 0 || cam2 >= rig->C) return MCS_ERR_ARG;
    mcs::PairList one{};
    one.n_pairs = 1;
    one.a[0] = (unsigned char)cam1;
    one.b[0] = (unsigned char)cam2;
    const int cap = rig->kp_cap;
    dim3 g((cap + MCS_MATCH_NT - 1) / MCS_MATCH_NT, 1, 1);
    CK(cudaMemsetAsync(rig->d_match_tmp, 0xFF, (size_t)cap * 3 * sizeof(int), rig->eng.stream));
    if (rig->match_imma) {  // experiment (MCS_MATCH_IMMA=1)
        dim3 gi((cap + MCS_IMMA_NT / 2 - 1) / (MCS_IMMA_NT / 2), 1, 1);
        mcs::hamming_match_imma_kernel<<<gi, MCS_IMMA_NT, 0, rig->eng.stream>>>(rig->C, cap, one, frame, rig->eng.d_desc, rig->eng.d_kp_count, dist_thresh,
                                                                                neigh_ratio, rig->d_match_tmp, rig->d_match_tmp + cap,
                                                                                rig->d_match_tmp + 2 * cap);
    } else
    mcs::hamming_match_kernel<<<g, MCS_MATCH_NT, MCS_MATCH_CHUNK * 32, rig->eng.stream>>>(rig->C, cap, one, frame, rig->eng.d_desc,
                                                                                          rig->eng.d_kp_count, dist_thresh, neigh_ratio,
                                                                                          rig->d_match_tmp, rig->d_match_tmp + cap,
                                                                                          rig->d_match_tmp + 2 * cap);
    rig->eng.launches++;
    CK(cudaMemcpyAsync(rig->h_match_tmp, rig->d_match_tmp, (size_t)cap * 3 * sizeof(int), cudaMemcpyDeviceToHost, rig->eng.stream));
    CK(cudaStreamSynchronize(rig->eng.stream));
    int rc = rig->ensure_features();
    if (!rc) rc = rig->ensure_kps_undist();
    if (rc) return rc;
    const mcs::KeypointPod* hk = rig->undist_pts_on ? rig->h_kps_undist : rig->eng.h_kps;
    const int ia = frame * rig->C + cam1, ib = frame * rig->C + cam2;
    const int nq = rig->eng.h_kp_count[ia];
    int n = 0;
    for (int q = 0; q < nq; q++) {  // query order is preserved (:1060-1076)
        const int t = rig->h_match_tmp[q];
        if (t < 0) continue;
        if (n >= capacity) { set_err("match capacity too small"); return MCS_ERR_CAPACITY; }
        if (indices_1) indices_1[n] = (uint32_t)q;
        if (indices_2) indices_2[n] = (uint32_t)t;
        // kps come from image_kps_undist; with zero distortion that is a copy of image_kps (MultiCameraFrame.cpp:302-305)
        if (kps1) memcpy(&kps1[n], hk + (size_t)ia * cap + q, sizeof(mcs_keypoint));
        if (kps2) memcpy(&kps2[n], hk + (size_t)ib * cap + t, sizeof(mcs_keypoint));
        n++;
    }
    return n;
}

int mcs_rig_compute_intra_matches(mcs_rig* rig, int n_frames, int old_epipolar, float dist_thresh, float neigh_ratio) {
    RIG_ENTER(rig);
    return rig->stage_match(n_frames, old_epipolar ? 1 : 0, dist_thresh, neigh_ratio);
}
int mcs_rig_get_intra_matches(mcs_rig* rig, int frame, int32_t* rows, int capacity) {
    RIG_ENTER(rig);
    if (frame < 0 || frame >= rig->n_matched) return MCS_ERR_ARG;
    int rc = rig->ensure_rows();
    if (rc) return rc;
    const int M = rig->h_row_count[frame];
    if (M > capacity) { set_err("row capacity too small"); return MCS_ERR_CAPACITY; }
    if (rows) memcpy(rows, rig->h_rows + (size_t)frame * rig->row_cap * rig->C, (size_t)M * rig->C * sizeof(int));
    return M;
}

int mcs_rig_triangulate(mcs_rig* rig, int n_frames, const float* masks) {
    RIG_ENTER(rig);
    if (n_frames < 1 || n_frames > rig->F || n_frames > rig->n_matched) { set_err("triangulate: n_frames out of range (1 .. frames intra-matched)"); return n_frames > rig->F ? MCS_ERR_CAPACITY : MCS_ERR_STATE; }
    const float* dm = nullptr;
    int rc = rig->upload_masks(masks, n_frames, &dm);
    if (!rc) rc = rig->undistort_masks(&dm, n_frames);
    if (rc) return rc;
    return rig->stage_triangulate(n_frames, dm);
}
int mcs_rig_triangulate_device_masks(mcs_rig* rig, int n_frames, const void* d_masks) {
    RIG_ENTER(rig);
    if (n_frames < 1 || n_frames > rig->F || n_frames > rig->n_matched) { set_err("triangulate: n_frames out of range (1 .. frames intra-matched)"); return n_frames > rig->F ? MCS_ERR_CAPACITY : MCS_ERR_STATE; }
    const float* dm = (const float*)d_masks;
    int rc = rig->undistort_masks(&dm, n_frames);
    if (rc) return rc;
    return rig->stage_triangulate(n_frames, dm);
}
int mcs_rig_get_landmarks(mcs_rig* rig, int frame, mcs_landmarks* out, int capacity) {
    RIG_ENTER(rig);
    if (!out || frame < 0 || frame >= rig->n_triangulated) return MCS_ERR_ARG;
    int rc = rig->ensure_landmarks();
    if (rc) return rc;
    const int M = std::min(rig->h_row_count[frame], rig->row_cap);
    if (M > capacity) { set_err("landmark capacity too small"); return MCS_ERR_CAPACITY; }
    const size_t o = (size_t)frame * rig->row_cap;
    if (out->rows) memcpy(out->rows, rig->h_lm_rows + o * rig->C, (size_t)M * rig->C * sizeof(int));
    if (out->accepted) memcpy(out->accepted, rig->h_accepted + o, M);
    if (out->n_views) memcpy(out->n_views, rig->h_n_views + o, (size_t)M * sizeof(int));
    if (out->xyz) memcpy(out->xyz, rig->h_xyz + o * 3, (size_t)M * 3 * sizeof(double));
    if (out->uv_ref) memcpy(out->uv_ref, rig->h_uv + o * 2, (size_t)M * 2 * sizeof(float));
    if (out->rep_cam) memcpy(out->rep_cam, rig->h_rep + o, (size_t)M * sizeof(int));
    return M;
}

int mcs_rig_triangulate_nview(mcs_rig* rig, const int* view_inds, int n_views, const double* xs, const double* ys, int n_points, double* xyz) {
    RIG_ENTER(rig);
    if (n_views < 2 || n_views > MCS_MAX_CAMS || n_points < 1 || !view_inds || !xs || !ys || !xyz) return MCS_ERR_ARG;
    for (int i = 0; i < n_views; i++)
        if (view_inds[i] < 0 || view_inds[i] >= rig->C) return MCS_ERR_ARG;
    int* d_vi = nullptr;
    double *d_x = nullptr, *d_y = nullptr, *d_o = nullptr;
    const size_t nb = (size_t)n_points * n_views * sizeof(double);
    CK(cudaMalloc(&d_vi, n_views * sizeof(int)));
    CK(cudaMalloc(&d_x, nb));
    CK(cudaMalloc(&d_y, nb));
    CK(cudaMalloc(&d_o, (size_t)n_points * 3 * sizeof(double)));
    cudaStream_t s = rig->eng.stream;
    CK(cudaMemcpyAsync(d_vi, view_inds, n_views * sizeof(int), cudaMemcpyHostToDevice, s));
    CK(cudaMemcpyAsync(d_x, xs, nb, cudaMemcpyHostToDevice, s));
    CK(cudaMemcpyAsync(d_y, ys, nb, cudaMemcpyHostToDevice, s));
    mcs::triangulate_nview_kernel<MCS_MAX_CAMS><<<(n_points + 63) / 64, 64, 0, s>>>(rig->d_calib, d_vi, n_views, d_x, d_y, n_points, d_o);
    rig->eng.launches++;
    CK(cudaMemcpyAsync(xyz, d_o, (size_t)n_points * 3 * sizeof(double), cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    cudaFree(d_vi); cudaFree(d_x); cudaFree(d_y); cudaFree(d_o);
    return 0;
}
int mcs_rig_triangulate_pair(mcs_rig* rig, int cam1, int cam2, const float* pts1, const float* pts2, int n_points, float* xyz) {
    RIG_ENTER(rig);
    if (cam1 < 0 || cam1 >= rig->C || cam2 < 0 || cam2 >= rig->C || n_points < 1 || !pts1 || !pts2 || !xyz) return MCS_ERR_ARG;
    float *d_a = nullptr, *d_b = nullptr, *d_o = nullptr;
    const size_t nb = (size_t)n_points * 2 * sizeof(float);
    CK(cudaMalloc(&d_a, nb));
    CK(cudaMalloc(&d_b, nb));
    CK(cudaMalloc(&d_o, (size_t)n_points * 3 * sizeof(float)));
    cudaStream_t s = rig->eng.stream;
    CK(cudaMemcpyAsync(d_a, pts1, nb, cudaMemcpyHostToDevice, s));
    CK(cudaMemcpyAsync(d_b, pts2, nb, cudaMemcpyHostToDevice, s));
    mcs::triangulate_pair_f32_kernel<<<(n_points + 63) / 64, 64, 0, s>>>(rig->d_calib, cam1, cam2, d_a, d_b, n_points, d_o);
    rig->eng.launches++;
    CK(cudaMemcpyAsync(xyz, d_o, (size_t)n_points * 3 * sizeof(float), cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    cudaFree(d_a); cudaFree(d_b); cudaFree(d_o);
    return 0;
}

int mcs_rig_process(mcs_rig* rig, const uint8_t* images, const float* masks, int n_frames, int old_epipolar) {
    RIG_ENTER(rig);
    if (!images || n_frames < 1) return MCS_ERR_EMPTY;
    if (n_frames > rig->F) { set_err("n_frames exceeds max_frames"); return MCS_ERR_CAPACITY; }
    rig->eng.begin_profile();
    return rig->process(images, nullptr, masks, nullptr, n_frames, old_epipolar ? 1 : 0, true);
}
int mcs_rig_process_device(mcs_rig* rig, const void* d_images, const void* d_masks, int n_frames, int old_epipolar) {
    RIG_ENTER(rig);
    if (!d_images || n_frames < 1) return MCS_ERR_EMPTY;
    if (n_frames > rig->F) { set_err("n_frames exceeds max_frames"); return MCS_ERR_CAPACITY; }
    rig->eng.begin_profile();
    return rig->process(nullptr, (const uint8_t*)d_images, nullptr, (const float*)d_masks, n_frames, old_epipolar ? 1 : 0, false);
}
int mcs_rig_process_masks_u8(mcs_rig* rig, const uint8_t* images, const uint8_t* masks_u8, int n_frames, int old_epipolar) {
    RIG_ENTER(rig);
    if (!images || n_frames < 1) return MCS_ERR_EMPTY;
    if (n_frames > rig->F) { set_err("n_frames exceeds max_frames"); return MCS_ERR_CAPACITY; }
    rig->eng.begin_profile();
    return rig->process(images, nullptr, reinterpret_cast<const float*>(masks_u8), nullptr, n_frames, old_epipolar ? 1 : 0, true, nullptr, nullptr, 1, true);
}
int mcs_rig_process_device_masks_u8(mcs_rig* rig, const void* d_images, const void* d_masks_u8, int n_frames, int old_epipolar) {
    RIG_ENTER(rig);
    if (!d_images || n_frames < 1) return MCS_ERR_EMPTY;
    if (n_frames > rig->F) { set_err("n_frames exceeds max_frames"); return MCS_ERR_CAPACITY; }
    rig->eng.begin_profile();
    return rig->process(nullptr, (const uint8_t*)d_images, nullptr, (const float*)d_masks_u8, n_frames, old_epipolar ? 1 : 0, false, nullptr, nullptr, 1, true);
}
int mcs_rig_set_graph(mcs_rig* rig, int enable) {
    RIG_ENTER(rig);
    rig->use_graph = enable != 0;
    if (!rig->use_graph) {
        CK(cudaStreamSynchronize(rig->eng.stream));
        rig->drop_graphs();
    }
    return 0;
}
long long mcs_rig_graph_replays(const mcs_rig* rig) { return rig ? rig->graph_replays : 0; }
long long mcs_rig_last_d2h_bytes(const mcs_rig* rig) { return rig ? rig->d2h_bytes : 0; }
static int check_f32_args(mcs_rig* rig, const void* images, int channels, int n_frames) {
    if (!images || n_frames < 1) return MCS_ERR_EMPTY;
    if (channels != 1 && channels != 3) { set_err("f32 input: channels must be 1 (gray) or 3 (BGR)"); return MCS_ERR_ARG; }
    if (n_frames > rig->F) { set_err("n_frames exceeds max_frames"); return MCS_ERR_CAPACITY; }
    if (((size_t)rig->cfg.width * rig->cfg.height) % 4 != 0) { set_err("f32 input: width * height must be a multiple of 4"); return MCS_ERR_ARG; }
    if (((uintptr_t)images & 15) != 0) { set_err("f32 input: buffer must be 16-byte aligned"); return MCS_ERR_ARG; }
    return 0;
}
int mcs_rig_process_f32(mcs_rig* rig, const float* images, int channels, const float* masks, int n_frames, int old_epipolar) {
    RIG_ENTER(rig);
    int rc = check_f32_args(rig, images, channels, n_frames);
    if (rc) return rc;
    rig->eng.begin_profile();
    return rig->process(nullptr, nullptr, masks, nullptr, n_frames, old_epipolar ? 1 : 0, true, images, nullptr, channels);
}
int mcs_rig_process_device_f32(mcs_rig* rig, const void* d_images, int channels, const void* d_masks, int n_frames, int old_epipolar) {
    RIG_ENTER(rig);
    int rc = check_f32_args(rig, d_images, channels, n_frames);
    if (rc) return rc;
    rig->eng.begin_profile();
    return rig->process(nullptr, nullptr, nullptr, (const float*)d_masks, n_frames, old_epipolar ? 1 : 0, false, nullptr, (const float*)d_images,
                        channels);
}
int mcs_rig_set_undistort(mcs_rig* rig, int enable) {
    RIG_ENTER(rig);
    cudaStreamSynchronize(rig->eng.stream);
    rig->drop_graphs();
    return rig->set_undistort(enable != 0);
}
int mcs_rig_set_undistort_keypoints(mcs_rig* rig, int enable) {
    RIG_ENTER(rig);
    cudaStreamSynchronize(rig->eng.stream);
    rig->drop_graphs();
    return rig->set_undistort_keypoints(enable != 0);
}
int mcs_rig_get_keypoints_undist(mcs_rig* rig, int frame, int cam, mcs_keypoint* keypoints, int capacity) {
    RIG_ENTER(rig);
    if (frame < 0 || frame >= rig->n_extracted || cam < 0 || cam >= rig->C || !keypoints) return MCS_ERR_ARG;
    int rc = rig->ensure_features();
    if (!rc) rc = rig->ensure_kps_undist();
    if (rc) return rc;
    const int img = frame * rig->C + cam, n = rig->eng.h_kp_count[img];
    if (n > capacity) { set_err("keypoint capacity too small"); return MCS_ERR_CAPACITY; }
    const mcs::KeypointPod* hk = rig->undist_pts_on ? rig->h_kps_undist : rig->eng.h_kps;
    memcpy(keypoints, hk + (size_t)img * rig->kp_cap, (size_t)n * sizeof(mcs_keypoint));
    return n;
}
int mcs_rig_set_chunk_frames(mcs_rig* rig, int chunk_frames) {
    if (!rig || chunk_frames < 0) return MCS_ERR_ARG;
    rig->chunk_frames = chunk_frames;
    return 0;
}

int mcs_rig_get_counts(mcs_rig* rig, int frame, int* kp_counts, int* n_rows, int* n_accepted) {
    RIG_ENTER(rig);
    if (frame < 0 || frame >= rig->n_extracted) return MCS_ERR_ARG;
    int rc;
    if (kp_counts) {
        if ((rc = rig->ensure_features())) return rc;
        for (int c = 0; c < rig->C; c++) kp_counts[c] = rig->eng.h_kp_count[frame * rig->C + c];
    }
    if (n_rows) {
        if (frame >= rig->n_matched) return MCS_ERR_STATE;
        if ((rc = rig->ensure_rows())) return rc;
        *n_rows = rig->h_row_count[frame];
    }
    if (n_accepted) {
        if (frame >= rig->n_triangulated) return MCS_ERR_STATE;
        if ((rc = rig->ensure_landmarks())) return rc;
        *n_accepted = rig->h_n_acc[frame];
    }
    return 0;
}

void* mcs_rig_stream(mcs_rig* rig) { return rig ? (void*)rig->eng.stream : nullptr; }
int mcs_rig_synchronize(mcs_rig* rig) {
    RIG_ENTER(rig);
    CK(cudaStreamSynchronize(rig->eng.stream));
    CK(cudaGetLastError());
    return 0;
}
long long mcs_rig_kernel_launches(const mcs_rig* rig) { return rig ? rig->eng.launches : 0; }
int mcs_rig_set_profiling(mcs_rig* rig, int enable) {
    if (!rig) return MCS_ERR_ARG;
    rig->eng.profiling = enable != 0;
    return 0;
}
int mcs_rig_stage_times(mcs_rig* rig, float* ms10) {
    RIG_ENTER(rig);
    if (!ms10) return MCS_ERR_ARG;
    if (rig->eng.profiling) {
        cudaEventRecord(rig->eng.ev[ST_COUNT], rig->eng.stream);
        rig->eng.ev_rec[ST_COUNT] = true;
        CK(cudaStreamSynchronize(rig->eng.stream));
        rig->eng.collect_profile();
    }
    for (int i = 0; i < ST_COUNT; i++) ms10[i] = rig->eng.stage_ms[i];
    return 0;
}

int mcs_rig_timer_start(mcs_rig* rig) {
    RIG_ENTER(rig);
    if (!rig->t0) { CK(cudaEventCreate(&rig->t0)); CK(cudaEventCreate(&rig->t1)); }
    CK(cudaEventRecord(rig->t0, rig->eng.stream));
    return 0;
}
int mcs_rig_timer_stop(mcs_rig* rig, float* ms) {
    RIG_ENTER(rig);
    if (!rig->t0 || !ms) return MCS_ERR_STATE;
    CK(cudaEventRecord(rig->t1, rig->eng.stream));
    CK(cudaEventSynchronize(rig->t1));
    CK(cudaEventElapsedTime(ms, rig->t0, rig->t1));
    return 0;
}

// Rig-level observers mirroring the extractor's (parity tests)
MCS_API int mcs_rig_download_level(mcs_rig* rig, int frame, int cam, int level, int kind, uint8_t* dst, int dst_stride) {
    RIG_ENTER(rig);
    return engine_download_level(rig->eng, frame * rig->C + cam, level, kind, dst, dst_stride);
}
MCS_API int mcs_rig_level_size(const mcs_rig* rig, int level, int* width, int* height) {
    if (!rig || level < 0 || level >= rig->eng.plan.nlevels) return MCS_ERR_ARG;
    if (width) *width = rig->eng.plan.L[level].w;
    if (height) *height = rig->eng.plan.L[level].h;
    return 0;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------------------
// Inter-frame matcher primitive: batched knnMatch(k = 2).
// ------------------------------------------------------------------------------------------------------------
struct mcs_matcher {
    int device = 0, max_sets = 0, max_q = 0, max_t = 0;
    cudaStream_t stream = nullptr;
    uint8_t *d_q = nullptr, *d_t = nullptr;
    mcs::Knn2Set *d_sets = nullptr, *h_sets = nullptr;
    int4* d_out = nullptr;
    long long launches = 0;
    // h_sets is one pinned staging buffer read by an asynchronous H2D copy: ev_sets marks the end of the last copy out of it, and
    // layout() waits for that copy (not for the kernel behind it) before it overwrites the table for the next call
    cudaEvent_t ev_sets = nullptr;
    bool sets_in_flight = false;
    // subset matcher workspace (grown on demand): CSR offsets, index lists, per-query scratch, outputs
    int *d_sub_i32 = nullptr;
    size_t sub_cap = 0;
    ~mcs_matcher() {
        if (cudaSetDevice(device) != cudaSuccess) return;
        if (stream) { cudaStreamSynchronize(stream); cudaStreamDestroy(stream); }
        if (ev_sets) cudaEventDestroy(ev_sets);
        cudaFree(d_q); cudaFree(d_t); cudaFree(d_sets); cudaFree(d_out); cudaFree(d_sub_i32);
        if (h_sets) cudaFreeHost(h_sets);
    }
    int init() {
        CK(cudaSetDevice(device));
        CK(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
        CK(cudaMalloc(&d_q, (size_t)std::max(max_q, 1) * 32));
        CK(cudaMalloc(&d_t, (size_t)std::max(max_t, 1) * 32));
        CK(cudaMalloc(&d_sets, (size_t)max_sets * sizeof(mcs::Knn2Set)));
        CK(cudaMallocHost(&h_sets, (size_t)max_sets * sizeof(mcs::Knn2Set)));
        CK(cudaMalloc(&d_out, (size_t)std::max(max_q, 1) * sizeof(int4)));
        CK(cudaEventCreateWithFlags(&ev_sets, cudaEventDisableTiming));
        return 0;
    }
    // fills h_sets from the per-set counts; returns the largest query count or a negative error
    int layout(const int* q_count, const int* t_count, int n_sets, int& tot_q, int& tot_t) {
        if (n_sets < 1 || n_sets > max_sets || !q_count || !t_count) { set_err("knn2: bad n_sets"); return MCS_ERR_ARG; }
        tot_q = tot_t = 0;
        int max_nq = 0;
        if (sets_in_flight) {  // the previous call's copy of the set table may still be reading h_sets
            CK(cudaEventSynchronize(ev_sets));
            sets_in_flight = false;
        }
        for (int i = 0; i < n_sets; i++) {
            if (q_count[i] < 0 || t_count[i] < 0) { set_err("knn2: negative count"); return MCS_ERR_ARG; }
            if (t_count[i] > 65535) { set_err("knn2: more than 65535 train descriptors in one set"); return MCS_ERR_CAPACITY; }
            h_sets[i] = mcs::Knn2Set{tot_q, q_count[i], tot_t, t_count[i], tot_q, 0, 0, 0};
            tot_q += q_count[i];
            tot_t += t_count[i];
            max_nq = std::max(max_nq, q_count[i]);
        }
        return max_nq;
    }
    int launch(const uint8_t* dq, const uint8_t* dt, const int* counts, int n_sets, int max_nq, int4* out) {
        CK(cudaMemcpyAsync(d_sets, h_sets, (size_t)n_sets * sizeof(mcs::Knn2Set), cudaMemcpyHostToDevice, stream));
        CK(cudaEventRecord(ev_sets, stream));
        sets_in_flight = true;
        if (max_nq > 0) {
            dim3 g((max_nq + MCS_MATCH_NT - 1) / MCS_MATCH_NT, n_sets);
            mcs::knn2_kernel<<<g, MCS_MATCH_NT, MCS_MATCH_CHUNK * 32, stream>>>(dq, dt, d_sets, counts, out);
            CK(cudaGetLastError());
            launches++;
        }
        return 0;
    }
};

extern "C" {

mcs_matcher* mcs_matcher_create(int device, int max_sets, int max_query_rows, int max_train_rows) {
    if (max_sets < 1 || max_query_rows < 0 || max_train_rows < 0) { set_err("mcs_matcher_create: bad sizes"); return nullptr; }
    mcs_matcher* m = new mcs_matcher();
    m->device = device; m->max_sets = max_sets; m->max_q = max_query_rows; m->max_t = max_train_rows;
    if (m->init()) { delete m; return nullptr; }
    return m;
}
void mcs_matcher_destroy(mcs_matcher* m) { delete m; }
long long mcs_matcher_kernel_launches(const mcs_matcher* m) { return m ? m->launches : 0; }

int mcs_knn2(mcs_matcher* m, const uint8_t* q_desc, const int* q_count, const uint8_t* t_desc, const int* t_count, int n_sets,
             int32_t* out_knn) {
    if (!m || !out_knn) return MCS_ERR_ARG;
    CK(cudaSetDevice(m->device));
    int tot_q, tot_t;
    const int max_nq = m->layout(q_count, t_count, n_sets, tot_q, tot_t);
    if (max_nq < 0) return max_nq;
    if (tot_q > m->max_q || tot_t > m->max_t) { set_err("knn2: descriptor rows exceed the matcher's capacity"); return MCS_ERR_CAPACITY; }
    if (tot_q == 0) return 0;
    if ((!q_desc) || (tot_t > 0 && !t_desc)) return MCS_ERR_ARG;
    CK(cudaMemcpyAsync(m->d_q, q_desc, (size_t)tot_q * 32, cudaMemcpyHostToDevice, m->stream));
    if (tot_t) CK(cudaMemcpyAsync(m->d_t, t_desc, (size_t)tot_t * 32, cudaMemcpyHostToDevice, m->stream));
    int rc = m->launch(m->d_q, m->d_t, nullptr, n_sets, max_nq, m->d_out);
    if (rc) return rc;
    CK(cudaMemcpyAsync(out_knn, m->d_out, (size_t)tot_q * sizeof(int4), cudaMemcpyDeviceToHost, m->stream));
    CK(cudaStreamSynchronize(m->stream));
    return tot_q;
}

int mcs_knn2_device(mcs_matcher* m, const void* d_q_desc, const int* q_count, const void* d_t_desc, const int* t_count, int n_sets,
                    void* d_out_knn) {
    if (!m || !d_out_knn || !d_q_desc || !d_t_desc) return MCS_ERR_ARG;
    CK(cudaSetDevice(m->device));
    int tot_q, tot_t;
    const int max_nq = m->layout(q_count, t_count, n_sets, tot_q, tot_t);
    if (max_nq < 0) return max_nq;
    int rc = m->launch((const uint8_t*)d_q_desc, (const uint8_t*)d_t_desc, nullptr, n_sets, max_nq, (int4*)d_out_knn);
    if (rc) return rc;
    return tot_q;
}
// ------------------------------------------------------------------------------------------------------------
// mcs_knn2_device_static: knnMatch(k = 2) on device-resident sets WITHOUT any allocation (
// no counterpart in the reference).  The set table that knn2_kernel reads is the kernel's scratch: here it lives in a file-scope
// __device__ array instead of a cudaMalloc'ed buffer, and a one-block kernel builds it on the device from the counts (passed by value
// as a kernel argument), so the call needs no handle, no workspace, no host-to-device copy.  Every GPU that loads the library holds
// the 8 KB array until exit.  In host code `g_knn2_static_sets` is only a stand-in for the symbol: its device address is taken once
// per device with cudaGetSymbolAddress (that device current) and passed to the kernels.
// ------------------------------------------------------------------------------------------------------------
}  // extern "C"

namespace {
static __device__ __align__(256) mcs::Knn2Set g_knn2_static_sets[MCS_KNN2_STATIC_MAX_SETS];

struct Knn2StaticCounts {  // by-value kernel argument: 2 KB
    int q[MCS_KNN2_STATIC_MAX_SETS], t[MCS_KNN2_STATIC_MAX_SETS];
};

// entries [0, n_sets) are rewritten by every call (nothing in the scratch has to start at zero); one thread: n_sets <= 256
__global__ void knn2_static_layout_kernel(const __grid_constant__ Knn2StaticCounts c, int n_sets, mcs::Knn2Set* __restrict__ sets) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    int tq = 0, tt = 0;
    for (int i = 0; i < n_sets; i++) {
        sets[i] = mcs::Knn2Set{tq, c.q[i], tt, c.t[i], tq, 0, 0, 0};
        tq += c.q[i];
        tt += c.t[i];
    }
}

constexpr int kMaxDevices = 64;
std::mutex g_knn2_static_mu;
mcs::Knn2Set* g_knn2_static_ptr[kMaxDevices] = {};

// device address of the scratch array on `device` (which must be current); looked up once per device
int knn2_static_scratch(int device, mcs::Knn2Set** out) {
    std::lock_guard<std::mutex> lk(g_knn2_static_mu);
    if (!g_knn2_static_ptr[device]) {
        void* p = nullptr;
        CK(cudaGetSymbolAddress(&p, g_knn2_static_sets));
        if (!p) { set_err("knn2_device_static: cudaGetSymbolAddress returned null"); return MCS_ERR_CUDA; }
        g_knn2_static_ptr[device] = (mcs::Knn2Set*)p;
    }
    *out = g_knn2_static_ptr[device];
    return 0;
}
}  // namespace

extern "C" {

int mcs_knn2_device_static(int device, void* stream, const void* d_q_desc, const int* q_count, const void* d_t_desc, const int* t_count,
                           int n_sets, void* d_out_knn) {
    if (!d_q_desc || !d_t_desc || !d_out_knn || !q_count || !t_count) { set_err("knn2_device_static: null argument"); return MCS_ERR_ARG; }
    if (n_sets < 1 || n_sets > MCS_KNN2_STATIC_MAX_SETS) {
        set_err("knn2_device_static: n_sets must be in [1, " + std::to_string(MCS_KNN2_STATIC_MAX_SETS) + "] (size of the static set table)");
        return MCS_ERR_CAPACITY;
    }
    if (device < 0 || device >= kMaxDevices) { set_err("knn2_device_static: device ordinal out of range"); return MCS_ERR_ARG; }
    Knn2StaticCounts c;
    long long tot_q = 0, tot_t = 0;
    int max_nq = 0;
    for (int i = 0; i < n_sets; i++) {
        if (q_count[i] < 0 || t_count[i] < 0) { set_err("knn2_device_static: negative count"); return MCS_ERR_ARG; }
        if (t_count[i] > 65535) { set_err("knn2_device_static: more than 65535 train descriptors in one set"); return MCS_ERR_CAPACITY; }
        c.q[i] = q_count[i];
        c.t[i] = t_count[i];
        tot_q += q_count[i];
        tot_t += t_count[i];
        max_nq = std::max(max_nq, q_count[i]);
    }
    if (tot_q > MCS_KNN2_STATIC_MAX_ROWS || tot_t > MCS_KNN2_STATIC_MAX_ROWS) {
        set_err("knn2_device_static: more than " + std::to_string(MCS_KNN2_STATIC_MAX_ROWS) + " query or train rows in one call");
        return MCS_ERR_CAPACITY;
    }
    CK(cudaSetDevice(device));
    mcs::Knn2Set* d_sets = nullptr;
    int rc = knn2_static_scratch(device, &d_sets);
    if (rc) return rc;
    if (max_nq == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    knn2_static_layout_kernel<<<1, 32, 0, st>>>(c, n_sets, d_sets);
    dim3 g((max_nq + MCS_MATCH_NT - 1) / MCS_MATCH_NT, n_sets);
    mcs::knn2_kernel<<<g, MCS_MATCH_NT, MCS_MATCH_CHUNK * 32, st>>>((const uint8_t*)d_q_desc, (const uint8_t*)d_t_desc, d_sets, nullptr,
                                                                    (int4*)d_out_knn);
    CK(cudaGetLastError());
    return (int)tot_q;
}

void* mcs_matcher_stream(mcs_matcher* m) { return m ? (void*)m->stream : nullptr; }
int mcs_matcher_synchronize(mcs_matcher* m) {
    if (!m) return MCS_ERR_ARG;
    CK(cudaSetDevice(m->device));
    CK(cudaStreamSynchronize(m->stream));
    return 0;
}

// ORBextractor::getMatches_distRatio for n_groups candidate-list pairs (see K6c).  Host buffers; blocking.
int mcs_match_subsets(mcs_matcher* m, const uint8_t* descA, int nA, const uint8_t* descB, int nB, int n_groups, const int* a_ofs,
                      const uint32_t* a_idx, const int* b_ofs, const uint32_t* b_idx, int th_low, double max_neighbor_ratio,
                      uint32_t* match_a, uint32_t* match_b, int* group_count, long long* book_k) {
    if (!m || n_groups < 0 || nA < 0 || nB < 0) return MCS_ERR_ARG;
    if (book_k) *book_k = 0;
    if (n_groups == 0) return 0;
    if (!a_ofs || !b_ofs || !match_a || !match_b) return MCS_ERR_ARG;
    CK(cudaSetDevice(m->device));
    const int ta = a_ofs[n_groups], tb = b_ofs[n_groups];
    if (a_ofs[0] != 0 || b_ofs[0] != 0 || ta < 0 || tb < 0) { set_err("match_subsets: offsets must start at 0"); return MCS_ERR_ARG; }
    for (int g = 0; g < n_groups; g++)
        if (a_ofs[g + 1] < a_ofs[g] || b_ofs[g + 1] < b_ofs[g]) { set_err("match_subsets: offsets must not decrease"); return MCS_ERR_ARG; }
    for (int i = 0; i < ta; i++) if (a_idx[i] >= (uint32_t)nA) { set_err("match_subsets: index into A out of range"); return MCS_ERR_ARG; }
    for (int i = 0; i < tb; i++) if (b_idx[i] >= (uint32_t)nB) { set_err("match_subsets: index into B out of range"); return MCS_ERR_ARG; }
    if (nA > m->max_q || nB > m->max_t) { set_err("match_subsets: descriptor rows exceed the matcher's capacity"); return MCS_ERR_CAPACITY; }
    if (ta == 0) { if (group_count) memset(group_count, 0, n_groups * sizeof(int)); return 0; }
    // workspace layout (int32 units): a_ofs | b_ofs | a_idx | b_idx | tmp_b | tmp_d | match_a | match_b | match_d | count | dups
    const size_t G1 = (size_t)n_groups + 1;
    const size_t need = 2 * G1 + (size_t)ta + tb + 5 * (size_t)ta + 2 * (size_t)n_groups;
    if (need > m->sub_cap) {
        cudaFree(m->d_sub_i32);
        m->d_sub_i32 = nullptr;
        m->sub_cap = 0;
        CK(cudaMalloc(&m->d_sub_i32, need * 2 * sizeof(int)));
        m->sub_cap = need * 2;
    }
    int* w = m->d_sub_i32;
    int* d_aofs = w; w += G1;
    int* d_bofs = w; w += G1;
    uint32_t* d_aidx = (uint32_t*)w; w += ta;
    uint32_t* d_bidx = (uint32_t*)w; w += tb;
    uint32_t* d_tmpb = (uint32_t*)w; w += ta;
    int* d_tmpd = w; w += ta;
    uint32_t* d_ma = (uint32_t*)w; w += ta;
    uint32_t* d_mb = (uint32_t*)w; w += ta;
    int* d_md = w; w += ta;
    int* d_cnt = w; w += n_groups;
    int* d_dup = w;
    cudaStream_t st = m->stream;
    CK(cudaMemcpyAsync(m->d_q, descA, (size_t)nA * 32, cudaMemcpyHostToDevice, st));
    if (nB) CK(cudaMemcpyAsync(m->d_t, descB, (size_t)nB * 32, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(d_aofs, a_ofs, G1 * sizeof(int), cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(d_bofs, b_ofs, G1 * sizeof(int), cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(d_aidx, a_idx, (size_t)ta * sizeof(int), cudaMemcpyHostToDevice, st));
    if (tb) CK(cudaMemcpyAsync(d_bidx, b_idx, (size_t)tb * sizeof(int), cudaMemcpyHostToDevice, st));
    mcs::subset_match_kernel<<<n_groups, 128, 0, st>>>(m->d_q, m->d_t, d_aofs, d_aidx, d_bofs, d_bidx, th_low, max_neighbor_ratio, d_tmpb,
                                                        d_tmpd, d_ma, d_mb, d_md, d_cnt, d_dup);
    CK(cudaGetLastError());
    m->launches++;
    std::vector<uint32_t> ha(ta), hb(ta);
    std::vector<int> hc(2 * (size_t)n_groups);
    CK(cudaMemcpyAsync(ha.data(), d_ma, (size_t)ta * sizeof(int), cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(hb.data(), d_mb, (size_t)ta * sizeof(int), cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(hc.data(), d_cnt, 2 * (size_t)n_groups * sizeof(int), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    int n = 0;
    long long book = 0;
    for (int g = 0; g < n_groups; g++) {  // concatenate the groups' match lists in group order
        const int c = hc[g];
        memcpy(match_a + n, ha.data() + a_ofs[g], c * sizeof(uint32_t));
        memcpy(match_b + n, hb.data() + a_ofs[g], c * sizeof(uint32_t));
        if (group_count) group_count[g] = c;
        n += c;
        book += (long long)(a_ofs[g + 1] - a_ofs[g]) * (b_ofs[g + 1] - b_ofs[g]) + hc[n_groups + g];
    }
    if (book_k) *book_k = book;
    return n;
}

// Inner loop of the BoW-guided computeIntraMatches(matches, words_) for n_groups (node, cam1, cam2) groups of frame `frame` (see K6d)
int mcs_rig_bow_best2(mcs_rig* rig, int frame, int n_groups, const int* cam_a, const int* cam_b, const int* a_ofs, const uint32_t* a_idx,
                      const int* b_ofs, const uint32_t* b_idx, float y_gate, int32_t* best_j, int32_t* d1, int32_t* d2) {
    RIG_ENTER(rig);
    if (frame < 0 || frame >= rig->n_extracted || n_groups < 0) return MCS_ERR_ARG;
    if (n_groups == 0) return 0;
    if (!cam_a || !cam_b || !a_ofs || !b_ofs || !best_j || !d1 || !d2) return MCS_ERR_ARG;
    int rc = rig->ensure_features();
    if (rc) return rc;
    const int ta = a_ofs[n_groups], tb = b_ofs[n_groups];
    std::vector<int> img(2 * (size_t)n_groups);
    for (int g = 0; g < n_groups; g++) {
        if (cam_a[g] < 0 || cam_a[g] >= rig->C || cam_b[g] < 0 || cam_b[g] >= rig->C || a_ofs[g + 1] < a_ofs[g] || b_ofs[g + 1] < b_ofs[g]) {
            set_err("bow_best2: bad group");
            return MCS_ERR_ARG;
        }
        img[g] = frame * rig->C + cam_a[g];
        img[n_groups + g] = frame * rig->C + cam_b[g];
        const int na = rig->eng.h_kp_count[img[g]], nb = rig->eng.h_kp_count[img[n_groups + g]];
        for (int i = a_ofs[g]; i < a_ofs[g + 1]; i++) if ((int)a_idx[i] >= na) { set_err("bow_best2: feature index out of range"); return MCS_ERR_ARG; }
        for (int i = b_ofs[g]; i < b_ofs[g + 1]; i++) if ((int)b_idx[i] >= nb) { set_err("bow_best2: feature index out of range"); return MCS_ERR_ARG; }
    }
    if (ta == 0) return 0;
    cudaStream_t st = rig->eng.stream;
    const size_t G1 = (size_t)n_groups + 1, words = 2 * (size_t)n_groups + 2 * G1 + ta + tb + 3 * (size_t)ta;
    int* w = nullptr;
    CK(cudaMallocAsync(&w, words * sizeof(int), st));
    int *d_img = w, *d_aofs = d_img + 2 * n_groups, *d_bofs = d_aofs + G1;
    uint32_t *d_aidx = (uint32_t*)(d_bofs + G1), *d_bidx = d_aidx + ta;
    int *d_bj = (int*)(d_bidx + tb), *d_d1 = d_bj + ta, *d_d2 = d_d1 + ta;
    CK(cudaMemcpyAsync(d_img, img.data(), img.size() * sizeof(int), cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(d_aofs, a_ofs, G1 * sizeof(int), cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(d_bofs, b_ofs, G1 * sizeof(int), cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(d_aidx, a_idx, (size_t)ta * sizeof(int), cudaMemcpyHostToDevice, st));
    if (tb) CK(cudaMemcpyAsync(d_bidx, b_idx, (size_t)tb * sizeof(int), cudaMemcpyHostToDevice, st));
    mcs::bow_best2_kernel<<<n_groups, 64, 0, st>>>(rig->eng.d_desc, rig->kps_undist_dev(), rig->kp_cap, d_img, d_img + n_groups, d_aofs, d_aidx,
                                                   d_bofs, d_bidx, y_gate, d_bj, d_d1, d_d2);
    CK(cudaGetLastError());
    rig->eng.launches++;
    CK(cudaMemcpyAsync(best_j, d_bj, (size_t)ta * sizeof(int), cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(d1, d_d1, (size_t)ta * sizeof(int), cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(d2, d_d2, (size_t)ta * sizeof(int), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    CK(cudaFreeAsync(w, st));
    return ta;
}

// knnMatch(k = 2) between feature sets the rig already holds on the device: set s matches image q_image[s] (= frame * num_cams + cam)
// against image t_image[s].  out_knn: [n_sets][feature_capacity][4]; rows >= the query image's keypoint count are left at -1.
int mcs_rig_knn2_images(mcs_rig* rig, const int* q_image, const int* t_image, int n_sets, int32_t* out_knn) {
    RIG_ENTER(rig);
    if (!q_image || !t_image || !out_knn || n_sets < 1) return MCS_ERR_ARG;
    const int n_img = rig->n_extracted * rig->C, cap = rig->kp_cap;
    if (rig->n_extracted < 1) { set_err("knn2_images before extract"); return MCS_ERR_STATE; }
    if (cap > 65535) return MCS_ERR_CAPACITY;
    std::vector<mcs::Knn2Set> sets(n_sets);
    for (int i = 0; i < n_sets; i++) {
        if (q_image[i] < 0 || q_image[i] >= n_img || t_image[i] < 0 || t_image[i] >= n_img) { set_err("knn2_images: image index out of range"); return MCS_ERR_ARG; }
        sets[i] = mcs::Knn2Set{q_image[i] * cap, q_image[i], t_image[i] * cap, t_image[i], i * cap, 0, 0, 0};
    }
    mcs::Knn2Set* d_sets = nullptr;
    int4* d_out = nullptr;
    const size_t out_bytes = (size_t)n_sets * cap * sizeof(int4);
    cudaStream_t st = rig->eng.stream;
    CK(cudaMallocAsync(&d_sets, sets.size() * sizeof(mcs::Knn2Set), st));
    CK(cudaMallocAsync(&d_out, out_bytes, st));
    CK(cudaMemcpyAsync(d_sets, sets.data(), sets.size() * sizeof(mcs::Knn2Set), cudaMemcpyHostToDevice, st));
    CK(cudaMemsetAsync(d_out, 0xFF, out_bytes, st));
    dim3 g((cap + MCS_MATCH_NT - 1) / MCS_MATCH_NT, n_sets);
    mcs::knn2_kernel<<<g, MCS_MATCH_NT, MCS_MATCH_CHUNK * 32, st>>>(rig->eng.d_desc, rig->eng.d_desc, d_sets, rig->eng.d_kp_count, d_out);
    CK(cudaGetLastError());
    rig->eng.launches++;
    CK(cudaMemcpyAsync(out_knn, d_out, out_bytes, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    CK(cudaFreeAsync(d_sets, st));
    CK(cudaFreeAsync(d_out, st));
    return 0;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------------------
// DBoW2 vocabulary (text format of the ORB-SLAM2 fork's loadFromTextFile, FrontEnd.h:138) + transform.
// ------------------------------------------------------------------------------------------------------------
struct mcs_vocab {
    int device = 0, k = 0, L = 0, scoring = 0, weighting = 0, n_nodes = 0, n_words = 0;
    cudaStream_t stream = nullptr;
    mcs::VocabNode* d_nodes = nullptr;
    uint4* d_desc = nullptr;
    double* d_weight = nullptr;
    // workspace of mcs_vocab_transform
    uint8_t* d_in = nullptr;
    int *d_word = nullptr, *d_node = nullptr;
    double* d_w = nullptr;
    size_t cap = 0;
    ~mcs_vocab() {
        if (cudaSetDevice(device) != cudaSuccess) return;
        if (stream) cudaStreamDestroy(stream);
        cudaFree(d_nodes); cudaFree(d_desc); cudaFree(d_weight); cudaFree(d_in); cudaFree(d_word); cudaFree(d_node); cudaFree(d_w);
    }
    // parent[n], is_leaf[n], desc[n][32], weight[n] for the nodes 1..n of the file (node 0 is the root)
    int build(int n, const int* parent, const int* is_leaf, const uint8_t* desc, const double* weight) {
        n_nodes = n + 1;
        std::vector<std::vector<int>> children(n_nodes);
        std::vector<int> word(n_nodes, -1);
        n_words = 0;
        for (int i = 1; i <= n; i++) {
            const int p = parent[i - 1];
            if (p < 0 || p >= i) { set_err("vocabulary: parent id must precede the node"); return MCS_ERR_ARG; }
            children[p].push_back(i);  // file order = DBoW2's children order
            if (is_leaf[i - 1] > 0) word[i] = n_words++;
        }
        // breadth-first renumbering so that every node's children are contiguous
        std::vector<int> order(1, 0), new_id(n_nodes, -1);
        new_id[0] = 0;
        for (size_t h = 0; h < order.size(); h++)
            for (int c : children[order[h]]) { new_id[c] = (int)order.size(); order.push_back(c); }
        if ((int)order.size() != n_nodes) { set_err("vocabulary: unreachable nodes"); return MCS_ERR_ARG; }
        std::vector<mcs::VocabNode> nodes(n_nodes);
        std::vector<uint8_t> nd((size_t)n_nodes * 32, 0);
        std::vector<double> nw(n_nodes, 0.0);
        for (int t = 0; t < n_nodes; t++) {
            const int o = order[t];
            nodes[t].child_count = (int)children[o].size();
            nodes[t].child_begin = nodes[t].child_count ? new_id[children[o][0]] : 0;
            nodes[t].word_id = word[o];
            nodes[t].orig_id = o;
            if (o > 0) { memcpy(&nd[(size_t)t * 32], desc + (size_t)(o - 1) * 32, 32); nw[t] = weight[o - 1]; }
            if (nodes[t].child_count == 0 && word[o] < 0 && n > 0) { set_err("vocabulary: inner node without children"); return MCS_ERR_ARG; }
        }
        CK(cudaSetDevice(device));
        CK(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
        CK(cudaMalloc(&d_nodes, nodes.size() * sizeof(mcs::VocabNode)));
        CK(cudaMalloc(&d_desc, nd.size()));
        CK(cudaMalloc(&d_weight, nw.size() * sizeof(double)));
        CK(cudaMemcpy(d_nodes, nodes.data(), nodes.size() * sizeof(mcs::VocabNode), cudaMemcpyHostToDevice));
        CK(cudaMemcpy(d_desc, nd.data(), nd.size(), cudaMemcpyHostToDevice));
        CK(cudaMemcpy(d_weight, nw.data(), nw.size() * sizeof(double), cudaMemcpyHostToDevice));
        return 0;
    }
};

extern "C" {

mcs_vocab* mcs_vocab_create(int device, int k, int L, int scoring, int weighting, int n_nodes, const int* parent, const int* is_leaf,
                            const uint8_t* descriptors, const double* weights) {
    if (n_nodes < 1 || !parent || !is_leaf || !descriptors || !weights || L < 1) { set_err("mcs_vocab_create: bad arguments"); return nullptr; }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1) { set_err("no CUDA device: this library has no CPU path"); return nullptr; }
    mcs_vocab* v = new mcs_vocab();
    v->device = device; v->k = k; v->L = L; v->scoring = scoring; v->weighting = weighting;
    if (v->build(n_nodes, parent, is_leaf, descriptors, weights)) { delete v; return nullptr; }
    return v;
}

// loadFromTextFile: "k L scoring weighting", then one line per node: parent is_leaf d0 .. d31 weight
mcs_vocab* mcs_vocab_load_text(const char* path, int device) {
    FILE* f = path ? fopen(path, "r") : nullptr;
    if (!f) { set_err("mcs_vocab_load_text: cannot open file"); return nullptr; }
    int k = 0, L = 0, n1 = 0, n2 = 0;
    if (fscanf(f, "%d %d %d %d", &k, &L, &n1, &n2) != 4 || k < 0 || k > 20 || L < 1 || L > 10 || n1 < 0 || n1 > 5 || n2 < 0 || n2 > 3) {
        fclose(f);
        set_err("mcs_vocab_load_text: vocabulary header out of range");
        return nullptr;
    }
    std::vector<int> parent, leaf;
    std::vector<uint8_t> desc;
    std::vector<double> weight;
    for (;;) {
        int pid, il;
        if (fscanf(f, "%d %d", &pid, &il) != 2) break;
        uint8_t d[32];
        bool ok = true;
        for (int i = 0; i < 32 && ok; i++) {
            int b;
            ok = fscanf(f, "%d", &b) == 1;
            d[i] = (uint8_t)b;
        }
        double w;
        if (!ok || fscanf(f, "%lf", &w) != 1) { fclose(f); set_err("mcs_vocab_load_text: truncated node line"); return nullptr; }
        parent.push_back(pid); leaf.push_back(il); weight.push_back(w);
        desc.insert(desc.end(), d, d + 32);
    }
    fclose(f);
    return mcs_vocab_create(device, k, L, n1, n2, (int)parent.size(), parent.data(), leaf.data(), desc.data(), weight.data());
}
void mcs_vocab_destroy(mcs_vocab* v) { delete v; }
int mcs_vocab_info(const mcs_vocab* v, int* k, int* L, int* n_nodes, int* n_words) {
    if (!v) return MCS_ERR_ARG;
    if (k) *k = v->k;
    if (L) *L = v->L;
    if (n_nodes) *n_nodes = v->n_nodes;
    if (n_words) *n_words = v->n_words;
    return 0;
}

// TemplatedVocabulary::transform(feature, word_id, weight, nid, levelsup) for n descriptors (host buffers)
int mcs_vocab_transform(mcs_vocab* v, const uint8_t* descriptors, int n, int levelsup, int32_t* word_id, int32_t* node_id, double* weight) {
    if (!v || n < 0 || (n && (!descriptors || !word_id || !node_id))) return MCS_ERR_ARG;
    if (n == 0) return 0;
    CK(cudaSetDevice(v->device));
    if ((size_t)n > v->cap) {
        cudaFree(v->d_in); cudaFree(v->d_word); cudaFree(v->d_node); cudaFree(v->d_w);
        v->d_in = nullptr; v->d_word = v->d_node = nullptr; v->d_w = nullptr; v->cap = 0;
        const size_t c = (size_t)n * 2;
        CK(cudaMalloc(&v->d_in, c * 32));
        CK(cudaMalloc(&v->d_word, c * sizeof(int)));
        CK(cudaMalloc(&v->d_node, c * sizeof(int)));
        CK(cudaMalloc(&v->d_w, c * sizeof(double)));
        v->cap = c;
    }
    cudaStream_t st = v->stream;
    CK(cudaMemcpyAsync(v->d_in, descriptors, (size_t)n * 32, cudaMemcpyHostToDevice, st));
    mcs::vocab_transform_kernel<<<dim3((n + 127) / 128, 1), 128, 0, st>>>(v->d_nodes, v->d_desc, v->d_weight, v->L, levelsup, v->d_in, nullptr, n,
                                                                         v->d_word, v->d_node, v->d_w);
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(word_id, v->d_word, (size_t)n * sizeof(int), cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(node_id, v->d_node, (size_t)n * sizeof(int), cudaMemcpyDeviceToHost, st));
    if (weight) CK(cudaMemcpyAsync(weight, v->d_w, (size_t)n * sizeof(double), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    return n;
}

// The same for every image a rig handle holds (descriptors stay on the device): outputs [n_images][feature_capacity]
int mcs_rig_vocab_transform(mcs_rig* rig, mcs_vocab* v, int levelsup, int32_t* word_id, int32_t* node_id, double* weight) {
    RIG_ENTER(rig);
    if (!v || !word_id || !node_id) return MCS_ERR_ARG;
    if (rig->n_extracted < 1) { set_err("vocab transform before extract"); return MCS_ERR_STATE; }
    if (v->device != rig->eng.device) { set_err("vocabulary lives on another device"); return MCS_ERR_ARG; }
    const int n_img = rig->n_extracted * rig->C, cap = rig->kp_cap;
    const size_t tot = (size_t)n_img * cap;
    int *d_word = nullptr, *d_node = nullptr;
    double* d_w = nullptr;
    cudaStream_t st = rig->eng.stream;
    CK(cudaMallocAsync(&d_word, tot * sizeof(int), st));
    CK(cudaMallocAsync(&d_node, tot * sizeof(int), st));
    CK(cudaMallocAsync(&d_w, tot * sizeof(double), st));
    CK(cudaMemsetAsync(d_word, 0xFF, tot * sizeof(int), st));
    CK(cudaMemsetAsync(d_node, 0xFF, tot * sizeof(int), st));
    CK(cudaMemsetAsync(d_w, 0, tot * sizeof(double), st));
    mcs::vocab_transform_kernel<<<dim3((cap + 127) / 128, n_img), 128, 0, st>>>(v->d_nodes, v->d_desc, v->d_weight, v->L, levelsup, rig->eng.d_desc,
                                                                               rig->eng.d_kp_count, cap, d_word, d_node, d_w);
    CK(cudaGetLastError());
    rig->eng.launches++;
    CK(cudaMemcpyAsync(word_id, d_word, tot * sizeof(int), cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(node_id, d_node, tot * sizeof(int), cudaMemcpyDeviceToHost, st));
    if (weight) CK(cudaMemcpyAsync(weight, d_w, tot * sizeof(double), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    CK(cudaFreeAsync(d_word, st));
    CK(cudaFreeAsync(d_node, st));
    CK(cudaFreeAsync(d_w, st));
    return 0;
}

}  // extern "C"
