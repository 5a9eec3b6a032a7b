/* mcslam_b200.h — C ABI of the B200-native MC-SLAM front-end hot path (ORB extract -> intra-rig match -> triangulate).
 *
 * Drop-in boundary for the two C++ classes the reference front-end calls directly (it has no plugin/FFI layer):
 *   ORBextractor        /root/reference/MCSlam/include/MCSlam/ORBextractor.h:44-112
 *   MultiCameraFrame    /root/reference/MCSlam/src/MultiCameraFrame.cpp:203 (extractFeaturesParallel), :1024 (BruteForceMatch),
 *                       :1100 (computeIntraMatches(vector<IntraMatch>&, bool)), :411/:481 (triangulate*), and the
 *                       triangulation block of FrontEnd::obtainLfFeatures, MCSlam/src/FrontEnd.cpp:249-370.
 * Plain pointers and sizes only; no torch / OpenCV types.  The C++ adaptor with the reference's method signatures is
 * mc-slam_b200/host/mcslam_adaptor.hpp; the binding a maintainer adds is shown in INTEGRATION.md.
 *
 * Every entry point runs hand-written sm_100a CUDA kernels.  There is no CPU fallback: without a CUDA device the
 * create calls fail (NULL + mcs_last_error()).
 *
 * Error convention: functions returning int give >= 0 on success and a negative MCS_ERR_* code on failure; the reference's
 * own conventions are kept where it has them (-1 for an empty image, ORBextractor.cpp:1090-1091).
 */
#ifndef MCSLAM_B200_H
#define MCSLAM_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MCS_API __attribute__((visibility("default")))

#define MCS_ERR_EMPTY -1    /* empty image: same value ORBextractor::operator() returns (ORBextractor.cpp:1091) */
#define MCS_ERR_CAPACITY -2 /* caller buffer / configured capacity too small */
#define MCS_ERR_ARG -3      /* invalid argument */
#define MCS_ERR_CUDA -4     /* CUDA runtime error; text in mcs_last_error() */
#define MCS_ERR_STATE -5    /* stage called before its producer stage */

#define MCS_MAX_LEVELS 16
#define MCS_MAX_CAMS 8

/* POD mirror of cv::KeyPoint, same field order and size (28 bytes). */
typedef struct mcs_keypoint {
    float x, y;       /* pt, in level-0 pixel coordinates (pt *= mvScaleFactor[level], ORBextractor.cpp:1149-1151) */
    float size;       /* (int)(31 * scale[level])  (ORBextractor.cpp:879, :889) */
    float angle;      /* 0 in parity mode (ORBextractor.cpp:475); IC_Angle degrees when orientation is enabled */
    float response;   /* FAST corner score */
    int32_t octave;   /* pyramid level */
    int32_t class_id; /* -1 */
} mcs_keypoint;

/* Per-camera calibration, the fields of CamArrayConfig (MCDataUtils/include/MCDataUtils/CamArrayConfig.h:24), all f64,
 * row-major.  X_cam = R * X_ref + t. */
typedef struct mcs_calib {
    double K[9];
    double R[9];
    double t[3];
    double dist[5]; /* dist[0] == 0 => keypoint undistortion is a copy (MultiCameraFrame.cpp:302-305) */
} mcs_calib;

MCS_API const char* mcs_last_error(void); /* thread-local, never NULL */
MCS_API int mcs_device_count(void);
/* Pin the calling host thread to the CPUs local to GPU `device` (NUMA node of its PCIe root); call before creating handles and pinned
 * buffers in a one-process-per-GPU host.  Returns the CPU count of the set, 0 if the topology is not exposed, < 0 on a bad device. */
MCS_API int mcs_bind_host_thread(int device);
MCS_API const char* mcs_version(void);

/* ------------------------------------------------------------------------------------------------------------------
 * ORBextractor  (replaces `new ORBextractor(nFeatures, fScaleFactor, nLevels, fIniThFAST, fMinThFAST)`, FrontEnd.h:190)
 * One handle per concurrent caller, like the reference's one object per camera thread (MultiCameraFrame.cpp:234).
 * max_images = how many same-sized images one batched call may carry (1 for the plain operator()).
 * ------------------------------------------------------------------------------------------------------------------ */
typedef struct mcs_extractor mcs_extractor;

MCS_API mcs_extractor* mcs_extractor_create(int nfeatures, float scale_factor, int nlevels, int ini_th_fast, int min_th_fast,
                                            int device, int max_width, int max_height, int max_images);
MCS_API void mcs_extractor_destroy(mcs_extractor* ex);

/* 0 = shipped behaviour (angle 0, un-steered BRIEF); 1 = IC_Angle + steered rBRIEF (ORBextractor.cpp:75-102, :106-145). */
MCS_API int mcs_extractor_set_orientation(mcs_extractor* ex, int enable);

/* Output capacity (keypoints per image) a caller must provide: nfeatures + 3 * nlevels rounded up. */
MCS_API int mcs_extractor_capacity(const mcs_extractor* ex);

/* ORBextractor::operator()(image, mask, keypoints, descriptors, vLappingArea)  (ORBextractor.cpp:1085-1171).
 * image: CV_8UC1 host pixels, `stride` bytes per row.  mask is accepted and ignored exactly as the reference does
 * (ORBextractor.h:56).  lapping = {x0, x1}: keypoints with x0 <= pt.x <= x1 are written from the back (:1153-1157).
 * keypoints[capacity], descriptors[capacity*32].  *n_total = number of keypoints written (mono + stereo).
 * Returns monoIndex (:1170), MCS_ERR_EMPTY for an empty image, or another negative code. */
MCS_API int mcs_extract(mcs_extractor* ex, const uint8_t* image, int width, int height, int stride, const uint8_t* mask_ignored,
                        const int lapping[2], mcs_keypoint* keypoints, uint8_t* descriptors, int capacity, int* n_total);

/* The same for n_images same-sized images in one launch sequence (replaces the thread-per-camera fan-out,
 * MultiCameraFrame.cpp:212-227).  images: n_images planes, `image_stride` bytes apart.  keypoints[n_images*capacity],
 * descriptors[n_images*capacity*32], mono_index[n_images], n_total[n_images]. */
MCS_API int mcs_extract_batch(mcs_extractor* ex, const uint8_t* images, int n_images, int width, int height, int stride,
                              size_t image_stride, const int lapping[2], mcs_keypoint* keypoints, uint8_t* descriptors, int capacity,
                              int* mono_index, int* n_total);

/* GetLevels / GetScaleFactor(s) / GetInverseScaleFactors / GetScaleSigmaSquares / GetInverseScaleSigmaSquares
 * (ORBextractor.h:61-81) and mnFeaturesPerLevel.  Each array has nlevels entries; NULL pointers are skipped. */
/* Host-only inspection of the FAST work decomposition of one configuration (the cell grid of ComputeKeyPointsOctTree,
 * ORBextractor.cpp:780-821, and the strips of cells one block of the FAST kernel processes).  Needs no device.  cells: capacity x 4
 * words {level, iniX | iniY << 16, width | height << 16 of the sub-image handed to cv::FAST, candidate offset}; strips: capacity x 8
 * words (a.x = level | cells << 8 | wCell << 16, a.y = iniX | iniY << 16, a.z = interior width | height << 16, a.w = offset,
 * b = index-division constants).  Either array may be NULL. */
MCS_API int mcs_plan_fast_tables(int width, int height, int nfeatures, float scale_factor, int nlevels, uint32_t* cells, int* n_cells,
                                 uint32_t* strips, int* n_strips, int* strip_smem_bytes, int capacity);

MCS_API int mcs_extractor_get_tables(const mcs_extractor* ex, int* nlevels, float* scale_factors, float* inv_scale_factors,
                                     float* level_sigma2, float* inv_level_sigma2, int* features_per_level);

/* Observers of mvImagePyramid (ORBextractor.h:89) for image `image_index` of the last call.
 * kind 0 = level pixels (w x h), 1 = level with its 19 px REFLECT_101 frame ((w+38) x (h+38)), 2 = 7x7 Gaussian-blurred level. */
MCS_API int mcs_extractor_level_size(const mcs_extractor* ex, int level, int* width, int* height);
MCS_API int mcs_extractor_download_level(mcs_extractor* ex, int image_index, int level, int kind, uint8_t* dst, int dst_stride);

/* Stage observers for parity tests: FAST candidates of (image, level) before the octree (vToDistributeKeys,
 * ORBextractor.cpp:864-867) as int32 triples (x, y, score) in ROI coordinates, in unspecified order. */
MCS_API int mcs_extractor_download_candidates(mcs_extractor* ex, int image_index, int level, int32_t* xys, int capacity);

/* ------------------------------------------------------------------------------------------------------------------
 * Multi-camera rig frame(s)  (replaces MultiCameraFrame's extract / intra-match / triangulate methods)
 * A rig handle owns device workspace for up to max_frames rig frames of num_cams cameras and one CUDA stream.
 * ------------------------------------------------------------------------------------------------------------------ */
typedef struct mcs_rig mcs_rig;

typedef struct mcs_rig_config {
    int num_cams;       /* <= MCS_MAX_CAMS (the reference hard-codes 5 in IntraMatch, MultiCameraFrame.h:44) */
    int width, height;  /* every camera of the rig has the same size */
    int nfeatures;      /* ORBextractor.nFeatures */
    float scale_factor; /* ORBextractor.scaleFactor */
    int nlevels;        /* ORBextractor.nLevels */
    int ini_th_fast;    /* ORBextractor.iniThFAST */
    int min_th_fast;    /* ORBextractor.minThFAST */
    int max_frames;     /* rig frames per call (batch capacity) */
    int device;         /* CUDA device ordinal */
} mcs_rig_config;

MCS_API mcs_rig* mcs_rig_create(const mcs_rig_config* cfg, const mcs_calib* calib /* [num_cams] */);
MCS_API void mcs_rig_destroy(mcs_rig* rig);
MCS_API int mcs_rig_feature_capacity(const mcs_rig* rig);  /* keypoints per camera */
MCS_API int mcs_rig_row_capacity(const mcs_rig* rig);      /* IntraMatch rows per frame */
MCS_API int mcs_rig_set_orientation(mcs_rig* rig, int enable);

/* Stage A — extractFeaturesParallel()/extractFeatures() for n_frames rig frames (MultiCameraFrame.cpp:203, :266).
 * images: host, [n_frames][num_cams][height][width] u8 tightly packed.  Results stay on the device. */
MCS_API int mcs_rig_extract(mcs_rig* rig, const uint8_t* images, int n_frames);
/* The same with images already resident in device memory (bench `value` leg); layout as above. */
MCS_API int mcs_rig_extract_device(mcs_rig* rig, const void* d_images, int n_frames);

/* Features computed elsewhere for the rig frames [0, n_frames) instead of stage A (image_kps / image_descriptors filled by another
 * extractor): counts[n_frames * num_cams], then the cameras' keypoints and 32-byte descriptors back to back in image order. */
MCS_API int mcs_rig_set_features(mcs_rig* rig, int n_frames, const int* counts, const mcs_keypoint* keypoints, const uint8_t* descriptors);
/* image_kps[cam] / image_descriptors[cam] of frame `frame`.  Returns the keypoint count. */
MCS_API int mcs_rig_get_features(mcs_rig* rig, int frame, int cam, mcs_keypoint* keypoints, uint8_t* descriptors, int capacity);

/* Stage B1 — BruteForceMatch(img1_ind, img2_ind, dist_thresh, neigh_ratio, indices_1, indices_2, kps1, kps2)
 * (MultiCameraFrame.cpp:1024-1086): knn k=2 Hamming, keep iff d0 < ratio*d1 and not d0 > dist_thresh, query order.
 * kps1/kps2 may be NULL.  Returns the number of matches. */
MCS_API int mcs_rig_brute_force_match(mcs_rig* rig, int frame, int cam1, int cam2, float dist_thresh, float neigh_ratio,
                                      uint32_t* indices_1, uint32_t* indices_2, mcs_keypoint* kps1, mcs_keypoint* kps2, int capacity);

/* Stage B2 — computeIntraMatches(matches, old) (MultiCameraFrame.cpp:1100-1288) for the first n_frames frames:
 * all camera pairs i<j through BruteForceMatch(i, j, dist_thresh, neigh_ratio) (the reference passes 75, 0.85),
 * optional epipolar gate (old = true), merge into IntraMatch rows.  Results stay on the device. */
MCS_API int mcs_rig_compute_intra_matches(mcs_rig* rig, int n_frames, int old_epipolar, float dist_thresh, float neigh_ratio);
/* rows: int32 [capacity][num_cams] = IntraMatch::matchIndex (-1 = not seen).  Returns the row count M. */
MCS_API int mcs_rig_get_intra_matches(mcs_rig* rig, int frame, int32_t* rows, int capacity);

/* Stage C — triangulation block of FrontEnd::obtainLfFeatures (FrontEnd.cpp:249-370) for the first n_frames frames:
 * mask filter (keep iff mask(int y, int x) < 0.7; masks = host float [n_frames][num_cams][height][width] or NULL for the
 * all-zero masks the app ships), normalised coordinates, N-view DLT (cv::sfm::triangulatePoints), depth gate 0.5 < z < 40,
 * uv_ref through K0, representative descriptor. */
MCS_API int mcs_rig_triangulate(mcs_rig* rig, int n_frames, const float* masks);
MCS_API int mcs_rig_triangulate_device_masks(mcs_rig* rig, int n_frames, const void* d_masks);
typedef struct mcs_landmarks {
    int32_t* rows;      /* [capacity][num_cams] matchIndex after the mask filter */
    uint8_t* accepted;  /* [capacity] 1 = pushed to intraMatches (passed the depth gate) */
    int32_t* n_views;   /* [capacity] views left after the mask filter (n_rays) */
    double* xyz;        /* [capacity][3] point3D (f64) — valid when n_views >= 2 */
    float* uv_ref;      /* [capacity][2] */
    int32_t* rep_cam;   /* [capacity] camera whose descriptor is the representative one (matchDesc), -1 if none */
} mcs_landmarks;
MCS_API int mcs_rig_get_landmarks(mcs_rig* rig, int frame, mcs_landmarks* out, int capacity); /* returns M */

/* MultiCameraFrame::triangulateIntraMatches(x, view_inds, pt) (MultiCameraFrame.cpp:411-441): N-view DLT of pixel
 * coordinates through K[R|t].  xs/ys: [n_points][n_views]; xyz: [n_points][3]. */
MCS_API int mcs_rig_triangulate_nview(mcs_rig* rig, const int* view_inds, int n_views, const double* xs, const double* ys, int n_points,
                                      double* xyz);
/* MultiCameraFrame::triangulate(kps_1, kps_2, pts3D, cam1, cam2) (MultiCameraFrame.cpp:481-507): 2-view f32 DLT. */
MCS_API int mcs_rig_triangulate_pair(mcs_rig* rig, int cam1, int cam2, const float* pts1 /*[n][2]*/, const float* pts2, int n_points,
                                     float* xyz /*[n][3]*/);

/* Whole path in one call with HOST buffers (upload -> A -> B -> C -> download of counts, rows and landmarks): the call
 * FrontEnd::processFrame makes, FrontEnd.cpp:991-1028.  Returns 0; per-frame results are read with the getters above. */
MCS_API int mcs_rig_process(mcs_rig* rig, const uint8_t* images, const float* masks, int n_frames, int old_epipolar);
/* Same, inputs resident in device memory, nothing copied back (results are read with the getters). */
MCS_API int mcs_rig_process_device(mcs_rig* rig, const void* d_images, const void* d_masks, int n_frames, int old_epipolar);

/* The same two calls with the segmentation decision already taken: masks_u8 holds one byte per pixel, 0 where the reference's
 * `segMasks[i].at<float>(p.y, p.x) < 0.7` (FrontEnd.cpp:266) is true (static scene, the view is kept) and non-zero where it is false
 * (dynamic object, the view is dropped).  [n_frames][num_cams][height][width]; a quarter of the f32 masks' PCIe traffic for the same
 * results.  Not available together with mcs_rig_set_undistort (a decision cannot be interpolated; pass the f32 masks there). */
MCS_API int mcs_rig_process_masks_u8(mcs_rig* rig, const uint8_t* images, const uint8_t* masks_u8, int n_frames, int old_epipolar);
MCS_API int mcs_rig_process_device_masks_u8(mcs_rig* rig, const void* d_images, const void* d_masks_u8, int n_frames, int old_epipolar);

/* Latency mode for the per-frame call of FrontEnd::processFrame (FrontEnd.cpp:991-1028).  When enabled, an unchunked
 * mcs_rig_process_device / mcs_rig_process_device_masks_u8 call (n_frames <= chunk_frames, no RECTIFY) is captured into a CUDA graph the
 * first time it is seen with a given (image pointer, mask pointer, n_frames, old_epipolar) and replayed with one cudaGraphLaunch
 * afterwards; results are identical to the stream path.  Up to 16 shapes are cached per handle.  mcs_rig_graph_replays counts the
 * replays (tests). */
MCS_API int mcs_rig_set_graph(mcs_rig* rig, int enable);
MCS_API long long mcs_rig_graph_replays(const mcs_rig* rig);
/* Bytes the last host-buffer process call (mcs_rig_process, _masks_u8, _f32) copied back to the host: keypoints and descriptors by
 * capacity, IntraMatch rows and landmark arrays by the row count of the fullest frame of the previous call + 25 % (whole slabs on the
 * first call; frames that outgrow the estimate are completed by a second copy).  For harnesses that account PCIe traffic. */
MCS_API long long mcs_rig_last_d2h_bytes(const mcs_rig* rig);

/* MultiCameraFrame::setData (MultiCameraFrame.cpp:95-152) fused in front of the path: images as the application holds them — CV_32F in
 * [0, 1] (imread / 255, MCDataUtils/src/DatasetReader.cpp:699-711), `channels` = 1 or 3 (interleaved BGR), tightly packed
 * [n_frames][num_cams][height][width][channels] — are multiplied by 255, converted to CV_8U (round half to even, saturate; :109-110)
 * and reduced to gray (cvtColor BGR2GRAY on the 8-bit values, :112-116) on the device, then processed like mcs_rig_process /
 * mcs_rig_process_device.  width * height must be a multiple of 4 and the buffer 16-byte aligned.  cv::undistort (RECTIFY,
 * :123-136) is not applied. */
MCS_API int mcs_rig_process_f32(mcs_rig* rig, const float* images, int channels, const float* masks, int n_frames, int old_epipolar);
MCS_API int mcs_rig_process_device_f32(mcs_rig* rig, const void* d_images, int channels, const void* d_masks, int n_frames,
                                       int old_epipolar);

/* RECTIFY branch of MultiCameraFrame::setData (MultiCameraFrame.cpp:123-136; CamArrayConfig::RECTIFY defaults to true): when enabled,
 * every image entering the handle (all extract / process entry points) goes through cv::undistort(img, K, dist_coeffs) with
 * dist = (k1, k2, p1, p2, k3) of mcs_calib, and every mask through the same maps (float weights), on the device.  Keypoints are then in
 * undistorted coordinates (image_kps_undist = image_kps, :241-243).  Off by default (the u8 entry points expect what setData hands
 * to the extractor). */
MCS_API int mcs_rig_set_undistort(mcs_rig* rig, int enable);

/* RECTIFY = false branch of extractFeatureSingle (MultiCameraFrame.cpp:244-245): image_kps_undist = UndistortKeyPoints (:300-347) =
 * cv::undistortPoints(image_kps, K, dist, noArray(), K) with K and dist narrowed to f32; cameras whose first coefficient reads as 0
 * through the reference's `.at<float>(0)` on the CV_64F matrix (:302, i.e. the low 32 bits of the double are zero) are copied.  When
 * enabled the undistorted keypoints feed what the reference feeds with image_kps_undist: the keypoints BruteForceMatch returns and the
 * epipolar gate of computeIntraMatches(matches, old); triangulation keeps using image_kps (FrontEnd.cpp:265, :292-298). */
MCS_API int mcs_rig_set_undistort_keypoints(mcs_rig* rig, int enable);
/* image_kps_undist[cam] of frame `frame` (a copy of image_kps while the option above is off).  Returns the keypoint count. */
MCS_API int mcs_rig_get_keypoints_undist(mcs_rig* rig, int frame, int cam, mcs_keypoint* keypoints, int capacity);

/* mcs_rig_process / mcs_rig_process_device cut batches larger than `chunk_frames` rig frames into chunks that rotate over six
 * CUDA streams (upload, kernels and download of neighbouring chunks overlap; host-buffer calls use chunks of a quarter of that).
 * 0 = one chunk; default 64 (env MCS_CHUNK_FRAMES). */
MCS_API int mcs_rig_set_chunk_frames(mcs_rig* rig, int chunk_frames);

/* Compact per-frame summary copied back by mcs_rig_process: n keypoints per camera, M rows, accepted landmarks. */
MCS_API int mcs_rig_get_counts(mcs_rig* rig, int frame, int* kp_counts /*[num_cams]*/, int* n_rows, int* n_accepted);

/* Observers for tests: pyramid level / blurred level of one camera image of the last batch (same kinds as
 * mcs_extractor_download_level). */
MCS_API int mcs_rig_download_level(mcs_rig* rig, int frame, int cam, int level, int kind, uint8_t* dst, int dst_stride);
MCS_API int mcs_rig_level_size(const mcs_rig* rig, int level, int* width, int* height);

/* Stream / timing plumbing for harnesses. */
MCS_API void* mcs_rig_stream(mcs_rig* rig);                 /* cudaStream_t the rig launches on */
MCS_API int mcs_rig_synchronize(mcs_rig* rig);
MCS_API long long mcs_rig_kernel_launches(const mcs_rig* rig); /* kernels launched by this handle since creation */
/* Device-time (ms, CUDA events on the rig's stream) of each stage of the last process/stage call:
 * [0] upload, [1] pyramid, [2] FAST, [3] octree, [4] blur, [5] describe, [6] match, [7] merge, [8] triangulate, [9] download */
MCS_API int mcs_rig_stage_times(mcs_rig* rig, float* ms10);
MCS_API int mcs_rig_set_profiling(mcs_rig* rig, int enable); /* record the per-stage events (adds event overhead) */
/* Stopwatch on the rig's own stream (torch.cuda.Event only sees torch's current stream): start records an event, stop records a
 * second one, waits for it and returns the device time between the two in milliseconds. */
MCS_API int mcs_rig_timer_start(mcs_rig* rig);
MCS_API int mcs_rig_timer_stop(mcs_rig* rig, float* ms);

/* ---- Inter-frame matcher primitive ---------------------------------------------------
 * Replaces  Ptr<DescriptorMatcher> matcher = DescriptorMatcher::create("BruteForce-Hamming"); matcher->knnMatch(q, t, matches, 2);
 * in FrontEnd::findMatchesMono (MCSlam/src/FrontEnd.cpp:3134-3135) and FrontEnd::findInterMatches (:3199-3201, :3373-3375):
 * n_sets independent (query, train) problems per launch.  Result rows are int32 (trainIdx0, distance0, trainIdx1, distance1) per
 * query in query order, -1 where the train set has fewer than 2 rows; ties are ordered by train index like OpenCV's.  The
 * ratio / distance / mask / 3-D gates differ between the two call sites and stay with the caller (mcslam_adaptor.hpp). */
typedef struct mcs_matcher mcs_matcher;
MCS_API mcs_matcher* mcs_matcher_create(int device, int max_sets, int max_query_rows, int max_train_rows);
MCS_API void mcs_matcher_destroy(mcs_matcher* m);
/* Host buffers: q_desc = the sets' query descriptors concatenated [sum(q_count)][32], t_desc likewise; out_knn [sum(q_count)][4].
 * Returns the number of query rows written.  At most 65535 train rows per set. */
MCS_API int mcs_knn2(mcs_matcher* m, const uint8_t* q_desc, const int* q_count, const uint8_t* t_desc, const int* t_count, int n_sets,
                     int32_t* out_knn);
/* Same with descriptors and output resident in device memory; asynchronous on mcs_matcher_stream(). */
MCS_API int mcs_knn2_device(mcs_matcher* m, const void* d_q_desc, const int* q_count, const void* d_t_desc, const int* t_count,
                            int n_sets, void* d_out_knn);
/* mcs_knn2_device with NO allocation anywhere: no handle, no cudaMalloc / cudaFree / pinned memory, no host-to-device copy.  It has
 * no counterpart in the reference; it was added for callers that cannot allocate (code under
 * CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  The kernel's scratch - the per-set offset table - is a
 * file-scope __device__ array of MCS_KNN2_STATIC_MAX_SETS entries (8 KB, held by every GPU that loads the library until exit) that a
 * one-block kernel fills on the device from the counts.  Inputs and outputs are the caller's device buffers as for mcs_knn2_device;
 * `stream` is a cudaStream_t (NULL = the default stream); the call is asynchronous on it.  Limits: n_sets <= MCS_KNN2_STATIC_MAX_SETS,
 * <= 65535 train rows per set, <= MCS_KNN2_STATIC_MAX_ROWS query rows and train rows per call; beyond them MCS_ERR_CAPACITY with a
 * text in mcs_last_error().  All calls on one GPU share the one table, which keeps the last call's contents (entries [0, n_sets)
 * are rewritten by each call; nothing has to start at zero): calls that may OVERLAP on the GPU (different streams, no ordering
 * between them) race on it - issue them on one stream or order them with events.  Returns the number of query rows. */
#define MCS_KNN2_STATIC_MAX_SETS 256
#define MCS_KNN2_STATIC_MAX_ROWS (1 << 24)
MCS_API int mcs_knn2_device_static(int device, void* stream, const void* d_q_desc, const int* q_count, const void* d_t_desc,
                                   const int* t_count, int n_sets, void* d_out_knn);
MCS_API void* mcs_matcher_stream(mcs_matcher* m);
MCS_API int mcs_matcher_synchronize(mcs_matcher* m);
MCS_API long long mcs_matcher_kernel_launches(const mcs_matcher* m);
/* ORBextractor::getMatches_distRatio (MCSlam/src/ORBextractor.cpp:1228-1290) for n_groups candidate-list pairs in one launch: the
 * per-node calls of MultiCameraFrame::BowMatching (MultiCameraFrame.cpp:946-1022) and FrontEnd::InterMatchingBow (FrontEnd.cpp:3791,
 * :3905).  Group g matches the features a_idx[a_ofs[g] .. a_ofs[g+1]) of descA against b_idx[b_ofs[g] .. b_ofs[g+1]) of descB
 * (CSR lists; the caller intersects the two DBoW2 FeatureVectors).  Accept: best <= th_low (TH_LOW = 75, ORBextractor.h:27) and
 * best / second <= max_neighbor_ratio (0.85, :90) in double; a B feature claimed twice keeps the closer A feature.  match_a /
 * match_b (capacity a_ofs[n_groups]) receive the groups' i_match_A / i_match_B concatenated in group order; group_count[g] (optional)
 * the matches of group g; *book_k (optional) the BookK distance-evaluation count.  Returns the total number of matches. */
MCS_API int mcs_match_subsets(mcs_matcher* m, const uint8_t* descA, int nA, const uint8_t* descB, int nB, int n_groups, const int* a_ofs,
                              const uint32_t* a_idx, const int* b_ofs, const uint32_t* b_idx, int th_low, double max_neighbor_ratio,
                              uint32_t* match_a, uint32_t* match_b, int* group_count, long long* book_k);
/* Inner loop of the BoW-guided MultiCameraFrame::computeIntraMatches(matches, words_) (MultiCameraFrame.cpp:716-751), for the
 * resident features of rig frame `frame`: group g = one vocabulary node seen by cameras cam_a[g] < cam_b[g] with the node's feature
 * lists a_idx[a_ofs[g]..) / b_idx[b_ofs[g]..).  Per cam_a feature: position of the best cam_b feature inside the group's list (or -1),
 * best and second-best Hamming distance (1000000000 = the reference's 1e9) over the cam_b features whose undistorted keypoint is
 * less than y_gate (50, :731) rows away.  The row bookkeeping (:752-816) runs on the host (mcslam_adaptor.hpp). */
MCS_API int mcs_rig_bow_best2(mcs_rig* rig, int frame, int n_groups, const int* cam_a, const int* cam_b, const int* a_ofs,
                              const uint32_t* a_idx, const int* b_ofs, const uint32_t* b_idx, float y_gate, int32_t* best_j, int32_t* d1,
                              int32_t* d2);
/* knnMatch(k = 2) between feature sets a rig handle already holds on the device (no descriptor traffic over PCIe): set s matches
 * image q_image[s] = frame * num_cams + cam against image t_image[s], e.g. findMatchesMono for every consecutive frame pair of a
 * batch.  out_knn: [n_sets][feature_capacity][4]; rows past the query image's keypoint count stay -1. */
MCS_API int mcs_rig_knn2_images(mcs_rig* rig, const int* q_image, const int* t_image, int n_sets, int32_t* out_knn);

/* ---- DBoW2 vocabulary transform ---------------------------------------------------------------------------------
 * DBoW2 (ORB-SLAM2 fork with loadFromTextFile) is an external, un-vendored dependency of the reference; the front-end calls
 * orb_vocabulary->loadFromTextFile(voc_file) (MCSlam/include/MCSlam/FrontEnd.h:137-138) and orb_vocabulary->transform(vec_desc, bowVec,
 * featVec, 4) (MCSlam/src/MultiCameraFrame.cpp:257, MCSlam/src/FrontEnd.cpp:533).  TemplatedVocabulary::transform per feature: descend
 * from the root to the Hamming-nearest child (first child wins ties) down to a leaf; word id and weight are the leaf's, node id the
 * node passed at level L - levelsup.  A feature enters the FeatureVector under its node id when its weight is > 0. */
typedef struct mcs_vocab mcs_vocab;
/* Text format: first line "k L scoring weighting", then one line per node 1..n: "parent is_leaf d0 ... d31 weight". */
MCS_API mcs_vocab* mcs_vocab_load_text(const char* path, int device);
/* The same from arrays (node i + 1 of the file = entry i). */
MCS_API mcs_vocab* mcs_vocab_create(int device, int k, int L, int scoring, int weighting, int n_nodes, const int* parent, const int* is_leaf,
                                    const uint8_t* descriptors /*[n_nodes][32]*/, const double* weights);
MCS_API void mcs_vocab_destroy(mcs_vocab* v);
MCS_API int mcs_vocab_info(const mcs_vocab* v, int* k, int* L, int* n_nodes /* incl. root */, int* n_words);
MCS_API int mcs_vocab_transform(mcs_vocab* v, const uint8_t* descriptors, int n, int levelsup, int32_t* word_id, int32_t* node_id,
                                double* weight /* may be NULL */);
/* For every image of the last batch of a rig handle, on the resident descriptors: outputs [n_frames * num_cams][feature_capacity]
 * (-1 / 0 past an image's keypoint count). */
MCS_API int mcs_rig_vocab_transform(mcs_rig* rig, mcs_vocab* v, int levelsup, int32_t* word_id, int32_t* node_id, double* weight);

/* Plain device-memory helpers so harnesses written without a CUDA binding can stage resident inputs. */
MCS_API void* mcs_device_malloc(int device, size_t bytes);
MCS_API void mcs_device_free(int device, void* p);
MCS_API int mcs_device_upload(int device, void* dst, const void* src, size_t bytes);
MCS_API int mcs_device_download(int device, void* dst, const void* src, size_t bytes);
MCS_API void* mcs_host_alloc_pinned(size_t bytes);
MCS_API void mcs_host_free_pinned(void* p);

#ifdef __cplusplus
}
#endif
#endif /* MCSLAM_B200_H */
