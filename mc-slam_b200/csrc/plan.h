// plan.h — host-built description of one extractor configuration (image size x ORB parameters), passed BY VALUE to
// every kernel as a __grid_constant__ parameter.  Restates the per-level tables of the reference constructor
// (/root/reference/MCSlam/src/ORBextractor.cpp:408-468), the level sizes of ComputePyramid (:1173-1198) and the FAST
// cell grid of ComputeKeyPointsOctTree (:780-803).
#ifndef MCS_PLAN_H
#define MCS_PLAN_H

#include <stdint.h>

#define MCS_PLAN_MAX_LEVELS 16
#define MCS_EDGE_THRESHOLD 19
#define MCS_MIN_BORDER 16  // EDGE_THRESHOLD - 3
#define MCS_BLUR_TW 128    // blur output tile
#define MCS_BLUR_TH 32

struct McsLevel {
    int w, h, pitch;     // level image; pitch is a multiple of 64 bytes
    int plane_ofs;       // byte offset of the plane inside one image's pyramid block (256-aligned)
    float scale;         // mvScaleFactor[level]
    int quota;           // mnFeaturesPerLevel[level]
    int roi_w, roi_h;    // maxBorderX - minBorderX, maxBorderY - minBorderY
    int n_cols, n_rows;  // FAST cell grid (0 when the level is too small for one cell)
    int w_cell, h_cell;
    int cell_base;       // index of this level's first cell in the per-image cell list
    int cand_ofs;        // offset (records) of this level's candidate list inside one image's candidate block
    int cand_cap;
    int n_ini;           // octree roots
    float hX;
    int oct_cap;         // node-table capacity == max selected keypoints of this level
    int sel_ofs;         // offset (records) of this level's selected list inside one image's block
    int xtab_ofs;        // resize tables (entries), level >= 1
    int ytab_ofs;
    int tile_base;       // blur tiles
    int tiles_x, tiles_y;
    int scaled_patch;    // (int)(PATCH_SIZE * scale)
};

struct McsPlan {
    int nlevels;
    int width, height;
    int nfeatures;
    int ini_th, min_th;
    int n_cells;        // cells per image over all levels
    int n_tiles;        // blur tiles per image over all levels
    int pyr_stride;     // bytes per image pyramid block
    int cand_stride;    // candidate records per image
    int sel_stride;     // selected records per image
    int kp_cap;         // output keypoints per image (>= sum of oct_cap)
    int oct_cap_max;
    int use_angle;      // 0 = shipped behaviour (angle 0)
    // FAST kernel shared-memory layout, sized for the largest cell of the plan (bytes unless noted)
    int fast_tpw;       // tile pitch in 32-bit words
    int fast_sp;        // score tile pitch
    int fast_ofs_score, fast_ofs_flags, fast_ofs_queue, fast_ofs_surv, fast_smem_bytes;
    int fast_maxsurv;   // survivor capacity (records)
    // strip FAST kernel (fast_strips_kernel): score tile, per-warp entry stacks, corner list (bytes)
    int fs_ofs_score, fs_ofs_wq, fs_ofs_corner, fs_ofs_ctl, fs_smem_bytes;
    // persistent TMA strip kernel (fast_strips_tma_kernel): two tile buffers, score tile, per-warp entry stacks, capped corner list
    int ft_tile_bytes, ft_ofs_score, ft_score_bytes, ft_ofs_wq, ft_ofs_corner, ft_ofs_ctl, ft_smem_bytes;
    int ft_box_rows[MCS_PLAN_MAX_LEVELS];  // TMA box height per level = tallest strip tile of the level
    int umax[16];       // circular patch row half-widths for IC_Angle (ORBextractor.cpp:452-467)
    McsLevel L[MCS_PLAN_MAX_LEVELS];
};

#endif
