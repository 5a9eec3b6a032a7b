// mcslam_adaptor.hpp — header-only C++17 host side over the C ABI (include/mcslam_b200.h) that keeps the call shapes of the two
// reference classes the front-end uses, so FrontEnd code compiles against it unchanged:
//   ORBextractor       /root/reference/MCSlam/include/MCSlam/ORBextractor.h:44-112
//   MultiCameraFrame   /root/reference/MCSlam/include/MCSlam/MultiCameraFrame.h (extract / intra-match / triangulate subset)
//   IntraMatch         /root/reference/MCSlam/include/MCSlam/MultiCameraFrame.h:42-57
//   ORBVocabulary      DBoW2 vocabulary: loadFromTextFile (FrontEnd.h:137-138), transform(features, bowVec, featVec, 4) (MultiCameraFrame.cpp:257)
//   BowMatcher         ORBextractor::getMatches_distRatio (ORBextractor.cpp:1228-1290) + the BowMatching / InterMatchingBow node walk
//   InterFrameMatcher  FrontEnd::findMatchesMono / findInterMatches, /root/reference/MCSlam/src/FrontEnd.cpp:3114-3470
// Without OpenCV headers the containers are PODs (mcs_keypoint, byte vectors).  Define MCS_WITH_OPENCV before including this
// file (in a build that has OpenCV) to get the cv::InputArray / cv::KeyPoint / cv::Mat overloads with the reference's exact
// signatures.  There is no CPU path: every method launches the sm_100a kernels and throws std::runtime_error when the library
// reports an error (the reference's own conventions are kept where it has them: operator() returns -1 for an empty image).
#ifndef MCSLAM_ADAPTOR_HPP
#define MCSLAM_ADAPTOR_HPP

#include <algorithm>
#include <array>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <numeric>
#include <set>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/mcslam_b200.h"

#ifdef MCS_WITH_OPENCV
#include <opencv2/core/core.hpp>
#include <opencv2/features2d/features2d.hpp>
#endif

namespace mcs_b200 {

inline void check(int rc, const char* what) {
    if (rc < 0) throw std::runtime_error(std::string(what) + ": " + mcs_last_error());
}

// ORBextractor.h:44-112.  One object per concurrent caller (the reference keeps one per camera thread, FrontEnd.h:190-193).
class ORBextractor {
public:
    enum { HARRIS_SCORE = 0, FAST_SCORE = 1 };

    ORBextractor(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST, int device = 0, int max_width = 720,
                 int max_height = 540)
        : nlevels_(nlevels), scaleFactor_(scaleFactor) {
        h_ = mcs_extractor_create(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, device, max_width, max_height, 1);
        if (!h_) throw std::runtime_error(std::string("mcs_extractor_create: ") + mcs_last_error());
        mvScaleFactor.resize(nlevels); mvInvScaleFactor.resize(nlevels); mvLevelSigma2.resize(nlevels); mvInvLevelSigma2.resize(nlevels);
        mnFeaturesPerLevel.resize(nlevels);
        int nl = 0;
        check(mcs_extractor_get_tables(h_, &nl, mvScaleFactor.data(), mvInvScaleFactor.data(), mvLevelSigma2.data(),
                                       mvInvLevelSigma2.data(), mnFeaturesPerLevel.data()), "mcs_extractor_get_tables");
    }
    ~ORBextractor() { mcs_extractor_destroy(h_); }
    ORBextractor(const ORBextractor&) = delete;
    ORBextractor& operator=(const ORBextractor&) = delete;

    // operator()(image, mask, keypoints, descriptors, vLappingArea) on raw CV_8UC1 pixels.  descriptors: n x 32 bytes, row-major.
    // Returns monoIndex (ORBextractor.cpp:1170) or -1 for an empty image (:1090-1091).
    int operator()(const uint8_t* image, int width, int height, int stride, const uint8_t* mask /*ignored, ORBextractor.h:56*/,
                   std::vector<mcs_keypoint>& keypoints, std::vector<uint8_t>& descriptors, std::vector<int>& vLappingArea) {
        if (!image || width <= 0 || height <= 0) return -1;
        const int lap[2] = {vLappingArea.size() > 0 ? vLappingArea[0] : 0, vLappingArea.size() > 1 ? vLappingArea[1] : 0};
        int rc = MCS_ERR_CAPACITY, total = 0;
        for (int attempt = 0; attempt < 2 && rc == MCS_ERR_CAPACITY; attempt++) {  // capacity follows the image-size-dependent plan
            const int cap = mcs_extractor_capacity(h_) > 8 ? mcs_extractor_capacity(h_) : 8;
            keypoints.resize(cap);
            descriptors.resize((size_t)cap * 32);
            rc = mcs_extract(h_, image, width, height, stride, mask, lap, keypoints.data(), descriptors.data(), cap, &total);
        }
        if (rc == MCS_ERR_EMPTY) return -1;
        check(rc, "mcs_extract");
        keypoints.resize(total);              // _keypoints = vector<KeyPoint>(nkeypoints)  (:1118)
        descriptors.resize((size_t)total * 32);  // _descriptors.create(nkeypoints, 32, CV_8U) (:1112)
        return rc;
    }

#ifdef MCS_WITH_OPENCV
    int operator()(cv::InputArray _image, cv::InputArray _mask, std::vector<cv::KeyPoint>& _keypoints, cv::OutputArray _descriptors,
                   std::vector<int>& vLappingArea) {
        (void)_mask;
        if (_image.empty()) return -1;
        cv::Mat image = _image.getMat();
        CV_Assert(image.type() == CV_8UC1);  // ORBextractor.cpp:1094
        std::vector<mcs_keypoint> k;
        std::vector<uint8_t> d;
        const int mono = (*this)(image.data, image.cols, image.rows, (int)image.step, nullptr, k, d, vLappingArea);
        if (k.empty()) _descriptors.release();
        else {
            _descriptors.create((int)k.size(), 32, CV_8U);
            std::memcpy(_descriptors.getMat().data, d.data(), d.size());
        }
        _keypoints.resize(k.size());
        static_assert(sizeof(cv::KeyPoint) == sizeof(mcs_keypoint), "mcs_keypoint mirrors cv::KeyPoint");
        if (!k.empty()) std::memcpy((void*)_keypoints.data(), k.data(), k.size() * sizeof(mcs_keypoint));
        return mono;
    }
#endif

    int GetLevels() { return nlevels_; }
    float GetScaleFactor() { return scaleFactor_; }
    std::vector<float> GetScaleFactors() { return mvScaleFactor; }
    std::vector<float> GetInverseScaleFactors() { return mvInvScaleFactor; }
    std::vector<float> GetScaleSigmaSquares() { return mvLevelSigma2; }
    std::vector<float> GetInverseScaleSigmaSquares() { return mvInvLevelSigma2; }

    // ORBextractor::DescriptorDistance (ORBextractor.cpp:1202-1218) on two 32-byte descriptors
    static int DescriptorDistance(const uint8_t* a, const uint8_t* b) {
        int d = 0;
        for (int i = 0; i < 8; i++) {
            uint32_t x, y;
            std::memcpy(&x, a + 4 * i, 4);
            std::memcpy(&y, b + 4 * i, 4);
            d += __builtin_popcount(x ^ y);
        }
        return d;
    }

    // mvImagePyramid[level] (ORBextractor.h:89) of the last call, w x h tightly packed
    std::vector<uint8_t> imagePyramidLevel(int level, int* w = nullptr, int* h = nullptr) {
        int lw = 0, lh = 0;
        check(mcs_extractor_level_size(h_, level, &lw, &lh), "mcs_extractor_level_size");
        std::vector<uint8_t> out((size_t)lw * lh);
        check(mcs_extractor_download_level(h_, 0, level, 0, out.data(), lw), "mcs_extractor_download_level");
        if (w) *w = lw;
        if (h) *h = lh;
        return out;
    }

    double max_neighbor_ratio = 0.85;  // ORBextractor.h:90
    std::vector<float> mvScaleFactor, mvInvScaleFactor, mvLevelSigma2, mvInvLevelSigma2;
    std::vector<int> mnFeaturesPerLevel;
    mcs_extractor* handle() { return h_; }

private:
    mcs_extractor* h_ = nullptr;
    int nlevels_;
    float scaleFactor_;
};

// MultiCameraFrame.h:42-57.  The reference hard-codes 5 cameras; rigs with more cameras use matchIndexN.
struct IntraMatch {
    std::array<int, 5> matchIndex{{-1, -1, -1, -1, -1}};
    std::vector<int> matchIndexN;  // [num_cams] (all cameras; equals matchIndex for num_cams <= 5)
    float uv_ref[2] = {0, 0};
    bool mono = true;
    int n_rays = 0;
    std::array<uint8_t, 32> matchDesc{};  // representative descriptor (MultiCameraFrame.cpp:530-567)
    double point3D[3] = {0, 0, 0};
};

// DBoW2::FeatureVector as the matchers read it: node id -> indices of the features that fell into that vocabulary node, ordered by
// node id.  ORBVocabulary::transform (below) builds it on the GPU.
typedef std::map<unsigned int, std::vector<unsigned int>> FeatureVector;

// Kalibr camchain calibration file as the data readers parse it (MCDataUtils/src/RosDataReader.cpp:258-380, DatasetReader.cpp:169-268):
// top-level nodes in file order, stopping at the first whose name does not contain "cam" (RosDataReader.cpp:266); per camera
// `distortion_coeffs` (radtan: copied in order; otherwise equidistant-style k1 k2 -> slots 0, 1 and the third value -> slot 4, :293-304),
// `resolution` (must be the same for all cameras), `intrinsics` fx fy cx cy, `rostopic`, and `T_cn_cnm1` (rows 0-2 -> R | t; absent ->
// identity).  With camchain = true the transforms are composed towards camera 0: R_i = R * R_{i-1}, t_i = R * t_{i-1} + t (:343-350).
// The file is OpenCV-FileStorage YAML; the subset Kalibr writes (block maps, flow sequences, `- [..]` rows, # comments) is parsed here.
struct CamArrayConfig {
    std::vector<mcs_calib> calib;          // K_mats_, R_mats_, t_vecs_, dist_coeffs_ per camera
    std::vector<mcs_calib> calib_kalibr;   // the un-chained R_mats_kalibr / t_vecs_kalibr
    std::vector<std::string> cam_topics;
    int width = 0, height = 0;
    int num_cams() const { return (int)calib.size(); }
};

inline CamArrayConfig read_kalibr_data(const std::string& path, bool radtan = true, bool camchain = true) {
    FILE* fp = std::fopen(path.c_str(), "r");
    if (!fp) throw std::runtime_error("read_kalibr_data: cannot open " + path);
    struct Node { std::string name; std::map<std::string, std::string> scalars; std::map<std::string, std::vector<double>> lists; };
    std::vector<Node> nodes;
    std::string pending_key;  // key whose value is a block sequence on the following lines
    int depth2_indent = -1;
    auto strip = [](std::string v) {
        size_t a = v.find_first_not_of(" \t\r\n"), b = v.find_last_not_of(" \t\r\n");
        return a == std::string::npos ? std::string() : v.substr(a, b - a + 1);
    };
    auto numbers = [&](const std::string& v, std::vector<double>& out) {
        const char* p = v.c_str();
        while (*p) {
            if ((*p >= '0' && *p <= '9') || *p == '-' || *p == '+' || *p == '.') {
                char* e = nullptr;
                const double d = std::strtod(p, &e);
                if (e == p) { p++; continue; }
                out.push_back(d);
                p = e;
            } else p++;
        }
    };
    char buf[4096];
    while (std::fgets(buf, sizeof buf, fp)) {
        std::string line(buf);
        for (size_t i = 0; i < line.size(); i++)  // comments: '#' at line start or after whitespace
            if (line[i] == '#' && (i == 0 || line[i - 1] == ' ' || line[i - 1] == '\t')) { line.resize(i); break; }
        if (strip(line).empty() || line[0] == '%' || line.compare(0, 3, "---") == 0) continue;
        const int indent = (int)line.find_first_not_of(" \t");
        const std::string body = strip(line);
        if (indent == 0) {
            const size_t c = body.find(':');
            if (c == std::string::npos) continue;
            Node n;
            n.name = strip(body.substr(0, c));
            nodes.push_back(n);
            pending_key.clear();
            depth2_indent = -1;
            continue;
        }
        if (nodes.empty()) continue;
        Node& n = nodes.back();
        if (body[0] == '-') {  // a row of the pending block sequence
            if (!pending_key.empty()) numbers(body.substr(1), n.lists[pending_key]);
            continue;
        }
        if (depth2_indent < 0) depth2_indent = indent;
        if (indent > depth2_indent) continue;  // deeper maps (imu / gps blocks) are not camera data
        const size_t c = body.find(':');
        if (c == std::string::npos) continue;
        const std::string key = strip(body.substr(0, c));
        std::string val = strip(body.substr(c + 1));
        pending_key.clear();
        if (val.empty()) { pending_key = key; n.lists[key]; continue; }
        if (val[0] == '[') {
            while (val.find(']') == std::string::npos && std::fgets(buf, sizeof buf, fp)) val += buf;  // flow sequence over several lines
            numbers(val, n.lists[key]);
        } else {
            if (val.size() >= 2 && (val[0] == '"' || val[0] == '\'')) val = val.substr(1, val.size() - 2);
            n.scalars[key] = val;
        }
    }
    std::fclose(fp);
    CamArrayConfig cfg;
    for (const Node& n : nodes) {
        if (n.name.find("cam") == std::string::npos) break;
        mcs_calib c;
        std::memset(&c, 0, sizeof c);
        const int i = (int)cfg.calib.size();
        auto need = [&](const char* key, size_t count) -> const std::vector<double>& {
            auto it = n.lists.find(key);
            if (it == n.lists.end() || it->second.size() < count) throw std::runtime_error("read_kalibr_data: " + n.name + " lacks " + key);
            return it->second;
        };
        const std::vector<double>& dc = need("distortion_coeffs", radtan ? 0 : 3);
        if (radtan) {
            for (size_t j = 0; j < dc.size() && j < 5; j++) c.dist[j] = dc[j];
        } else {
            c.dist[0] = dc[0]; c.dist[1] = dc[1]; c.dist[4] = dc[2];
        }
        const std::vector<double>& res = need("resolution", 2);
        if (i > 0) {
            if ((int)res[0] != cfg.width || (int)res[1] != cfg.height) throw std::runtime_error("Resolution of all images is not the same!");
        } else {
            cfg.width = (int)res[0]; cfg.height = (int)res[1];
        }
        const std::vector<double>& in = need("intrinsics", 4);
        c.K[0] = in[0]; c.K[4] = in[1]; c.K[2] = in[2]; c.K[5] = in[3]; c.K[8] = 1.0;
        auto tn = n.lists.find("T_cn_cnm1");
        if (tn == n.lists.end() || tn->second.empty()) {
            c.R[0] = c.R[4] = c.R[8] = 1.0;
        } else {
            if (tn->second.size() < 12) throw std::runtime_error("read_kalibr_data: " + n.name + " has a short T_cn_cnm1");
            for (int r = 0; r < 3; r++) {
                for (int k = 0; k < 3; k++) c.R[r * 3 + k] = tn->second[r * 4 + k];
                c.t[r] = tn->second[r * 4 + 3];
            }
        }
        cfg.calib_kalibr.push_back(c);
        if (camchain && i > 0) {
            const mcs_calib& p = cfg.calib[i - 1];
            mcs_calib o = c;
            for (int r = 0; r < 3; r++) {
                for (int k = 0; k < 3; k++) {
                    double acc = 0;
                    for (int m = 0; m < 3; m++) acc += c.R[r * 3 + m] * p.R[m * 3 + k];
                    o.R[r * 3 + k] = acc;
                }
                double acc = 0;
                for (int m = 0; m < 3; m++) acc += c.R[r * 3 + m] * p.t[m];
                o.t[r] = acc + c.t[r];
            }
            c = o;
        }
        cfg.calib.push_back(c);
        auto topic = n.scalars.find("rostopic");
        cfg.cam_topics.push_back(topic == n.scalars.end() ? std::string() : topic->second);
    }
    return cfg;
}

// FrontEnd::compute_overlap (FrontEnd.cpp:43-198): per camera the 40 x 30 px grid cells whose centre, unprojected at depth dmin = 1 and
// moved into another camera (T_ji = T_j0 * T_i0^-1), projects inside that camera's image (x in (10, width), y in (10, height)).  Outputs
// what the reference keeps: overlap_masks[c] = CV_8U image, 50 x (number of other cameras that see the cell) over the cell's pixels
// (rows / columns beyond the last whole cell stay 0), and non_overlap_fraction[c] = cells of camera c nobody else sees / the same count
// summed over the cameras (float; the reference divides by zero when every cell is seen).  Calibration-only set-up work that runs once
// per rig (a few thousand 3 x 4 products), so it is host code; the shipped front-end branch reads the masks only when `tiling` is on
// (FrontEnd.cpp:493).  The epipolar-line part of the loop and the imshow block compute nothing that is kept.
struct OverlapInfo {
    std::vector<std::vector<uint8_t>> overlap_masks;  // [num_cams][height * width]
    std::vector<float> non_overlap_fraction;          // [num_cams]
};
inline OverlapInfo compute_overlap(const std::vector<mcs_calib>& calib, int width, int height) {
    const int cell_size_x = 40, cell_size_y = 30;
    const int grid_rows = height / cell_size_y, grid_cols = width / cell_size_x;
    const int C = (int)calib.size();
    OverlapInfo out;
    int total_non_overlap_cells = 0;
    auto T4 = [](const mcs_calib& c, double T[16]) {
        for (int r = 0; r < 3; r++) {
            for (int k = 0; k < 3; k++) T[r * 4 + k] = c.R[r * 3 + k];
            T[r * 4 + 3] = c.t[r];
        }
        T[12] = T[13] = T[14] = 0; T[15] = 1;
    };
    auto inv4 = [](const double A[16], double X[16]) {  // Gauss-Jordan with partial pivoting (cv::Mat::inv(), DECOMP_LU)
        double a[4][8];
        for (int r = 0; r < 4; r++)
            for (int k = 0; k < 4; k++) { a[r][k] = A[r * 4 + k]; a[r][4 + k] = r == k ? 1.0 : 0.0; }
        for (int i = 0; i < 4; i++) {
            int p = i;
            for (int r = i + 1; r < 4; r++)
                if (std::fabs(a[r][i]) > std::fabs(a[p][i])) p = r;
            if (p != i) for (int k = 0; k < 8; k++) std::swap(a[i][k], a[p][k]);
            const double d = 1.0 / a[i][i];
            for (int k = 0; k < 8; k++) a[i][k] *= d;
            for (int r = 0; r < 4; r++) {
                if (r == i) continue;
                const double f = a[r][i];
                for (int k = 0; k < 8; k++) a[r][k] -= f * a[i][k];
            }
        }
        for (int r = 0; r < 4; r++)
            for (int k = 0; k < 4; k++) X[r * 4 + k] = a[r][4 + k];
    };
    for (int c1 = 0; c1 < C; c1++) {
        std::vector<double> overlap_grid2((size_t)grid_rows * grid_cols, 0.0);
        for (int c2 = 0; c2 < C; c2++) {
            if (c1 == c2) continue;
            double Ti[16], Tj[16], Tinv[16], Tji[16];
            T4(calib[c1], Ti);
            T4(calib[c2], Tj);
            inv4(Ti, Tinv);
            for (int r = 0; r < 4; r++)
                for (int k = 0; k < 4; k++) {
                    double s = 0;
                    for (int q = 0; q < 4; q++) s += Tj[r * 4 + q] * Tinv[q * 4 + k];
                    Tji[r * 4 + k] = s;
                }
            const double dmin = 1.0;
            const double* K1 = calib[c1].K;
            const double* K2 = calib[c2].K;
            for (int i = 0; i < grid_cols; i++)
                for (int j = 0; j < grid_rows; j++) {
                    const double gx = cell_size_x * i + (cell_size_x / 2.0), gy = cell_size_y * j + (cell_size_y / 2.0);
                    const double p[4] = {(gx - K1[2]) / K1[0] * dmin, (gy - K1[5]) / K1[4] * dmin, dmin, 1.0};
                    double q[3], pr[3];
                    for (int r = 0; r < 3; r++) {
                        double s = 0;
                        for (int k = 0; k < 4; k++) s += Tji[r * 4 + k] * p[k];
                        q[r] = s;
                    }
                    for (int r = 0; r < 3; r++) {
                        double s = 0;
                        for (int k = 0; k < 3; k++) s += K2[r * 3 + k] * q[k];
                        pr[r] = s;
                    }
                    const double ex = pr[0] / pr[2], ey = pr[1] / pr[2];
                    if (ex > 10 && ex < width && ey < height && ey > 10) overlap_grid2[(size_t)j * grid_cols + i] += 1;
                }
        }
        int nz = 0;
        for (double v : overlap_grid2) nz += v != 0;
        const float no_overlap_cells = (float)(grid_rows * grid_cols - nz);
        out.non_overlap_fraction.push_back(no_overlap_cells);
        total_non_overlap_cells += (int)no_overlap_cells;
        std::vector<uint8_t> mask((size_t)width * height, 0);
        for (int j = 0; j < grid_rows; j++)
            for (int i = 0; i < grid_cols; i++) {
                const double v = overlap_grid2[(size_t)j * grid_cols + i] * 50;  // overlap_grid * 50, convertTo(CV_8U)
                const uint8_t b = (uint8_t)std::min(255.0, std::max(0.0, std::nearbyint(v)));
                if (!b) continue;
                for (int y = cell_size_y * j; y < cell_size_y * (j + 1); y++)
                    std::memset(&mask[(size_t)y * width + cell_size_x * i], b, cell_size_x);
            }
        out.overlap_masks.push_back(std::move(mask));
    }
    for (int i = 0; i < C; i++) out.non_overlap_fraction[i] = out.non_overlap_fraction[i] / total_non_overlap_cells;
    return out;
}

class BowMatcher;
struct DMatch;

// The extract / intra-match / triangulate entry points of MultiCameraFrame for ONE rig frame (the class also batches frames
// through process_frames for throughput).  Image data: num_cams CV_8UC1 planes of the same size, tightly packed.
class MultiCameraFrame {
public:
    MultiCameraFrame(const std::vector<mcs_calib>& calib, int width, int height, int nfeatures, float scaleFactor, int nlevels,
                     int iniThFAST, int minThFAST, int device = 0, int max_frames = 1)
        : num_cams_((int)calib.size()), calib_(calib), w_(width), h_img_(height) {
        mcs_rig_config cfg{};
        cfg.num_cams = num_cams_; cfg.width = width; cfg.height = height; cfg.nfeatures = nfeatures; cfg.scale_factor = scaleFactor;
        cfg.nlevels = nlevels; cfg.ini_th_fast = iniThFAST; cfg.min_th_fast = minThFAST; cfg.max_frames = max_frames; cfg.device = device;
        h_ = mcs_rig_create(&cfg, calib.data());
        if (!h_) throw std::runtime_error(std::string("mcs_rig_create: ") + mcs_last_error());
        kp_cap_ = mcs_rig_feature_capacity(h_);
        row_cap_ = mcs_rig_row_capacity(h_);
        image_kps.resize(num_cams_); image_kps_undist.resize(num_cams_); image_descriptors.resize(num_cams_); BoW_feats.resize(num_cams_);
    }
    ~MultiCameraFrame() { mcs_rig_destroy(h_); }
    MultiCameraFrame(const MultiCameraFrame&) = delete;
    MultiCameraFrame& operator=(const MultiCameraFrame&) = delete;

    // setData (MultiCameraFrame.cpp:95-152) for already-gray 8-bit images: [num_cams][h][w]
    void setData(const uint8_t* images) { images_ = images; }

    // extractFeaturesParallel() (MultiCameraFrame.cpp:203-228): fills image_kps / image_kps_undist / image_descriptors.
    // With zero distortion image_kps_undist is a copy (:302-305); distorted rigs are out of scope.
    void extractFeaturesParallel() {
        if (!images_) throw std::runtime_error("extractFeaturesParallel: setData first");
        check(mcs_rig_extract(h_, images_, 1), "mcs_rig_extract");
        for (int c = 0; c < num_cams_; c++) {
            image_kps[c].resize(kp_cap_);
            image_descriptors[c].resize((size_t)kp_cap_ * 32);
            const int n = mcs_rig_get_features(h_, 0, c, image_kps[c].data(), image_descriptors[c].data(), kp_cap_);
            check(n, "mcs_rig_get_features");
            image_kps[c].resize(n);
            image_descriptors[c].resize((size_t)n * 32);
            image_kps_undist[c].resize(n);  // UndistortKeyPoints (:300-347) when setUndistortKeypoints(true), else the copy of :241-243 / :302-305
            if (n) check(mcs_rig_get_keypoints_undist(h_, 0, c, image_kps_undist[c].data(), n), "mcs_rig_get_keypoints_undist");
        }
    }
    void extractFeatures() { extractFeaturesParallel(); }  // :266-298, same results without the thread fan-out

    // BruteForceMatch(img1_ind, img2_ind, dist_thresh, neigh_ratio, indices_1, indices_2, kps_1, kps_2) (:1024-1086)
    void BruteForceMatch(int img1_ind, int img2_ind, float dist_thresh, float neigh_ratio, std::vector<unsigned int>& indices_1,
                         std::vector<unsigned int>& indices_2, std::vector<mcs_keypoint>& kps_1, std::vector<mcs_keypoint>& kps_2) {
        indices_1.resize(kp_cap_); indices_2.resize(kp_cap_); kps_1.resize(kp_cap_); kps_2.resize(kp_cap_);
        const int n = mcs_rig_brute_force_match(h_, 0, img1_ind, img2_ind, dist_thresh, neigh_ratio, indices_1.data(), indices_2.data(),
                                                kps_1.data(), kps_2.data(), kp_cap_);
        check(n, "mcs_rig_brute_force_match");
        indices_1.resize(n); indices_2.resize(n); kps_1.resize(n); kps_2.resize(n);
    }

    // computeIntraMatches(matches, old) (:1100-1288)
    void computeIntraMatches(std::vector<IntraMatch>& matches, bool old) {
        check(mcs_rig_compute_intra_matches(h_, 1, old ? 1 : 0, 75.f, 0.85f), "mcs_rig_compute_intra_matches");
        std::vector<int32_t> rows((size_t)row_cap_ * num_cams_);
        const int M = mcs_rig_get_intra_matches(h_, 0, rows.data(), row_cap_);
        check(M, "mcs_rig_get_intra_matches");
        matches.assign(M, IntraMatch());
        for (int r = 0; r < M; r++) fill_row(matches[r], &rows[(size_t)r * num_cams_]);
    }

    // computeIntraMatches(matches, words_) (:586-943) — the BoW-guided intra-rig matcher FrontEnd runs by default (FrontEnd.cpp:1009),
    // over BoW_feats[cam] (the FeatureVectors of :257).  The walk over the vocabulary nodes depends on node ids only, so it is done
    // twice: once to list every (node, cam1 < cam2) group, whose best / second-best search (:716-751) then runs on the GPU in ONE
    // launch over the features the rig handle already holds, and once more to replay the reference's order-dependent row bookkeeping
    // (:752-816) on those tables.  Reference quirks kept: the last node of every FeatureVector is never visited (:627), words_ gets one
    // entry per successful cam1 feature even when its row was merged into an older one, stolen features leave rows with n_rays == 1.
    std::vector<FeatureVector> BoW_feats;
    void computeIntraMatches(std::vector<IntraMatch>& matches, std::vector<unsigned int>& words_, int TH_LOW = 75,
                             double max_neighbor_ratio = 0.85) {
        if ((int)BoW_feats.size() != num_cams_) throw std::invalid_argument("computeIntraMatches: BoW_feats must hold one FeatureVector per camera");
        typedef FeatureVector::const_iterator It;
        std::vector<It> begin(num_cams_), last(num_cams_);
        for (int c = 0; c < num_cams_; c++) {
            if (BoW_feats[c].empty()) return;
            begin[c] = BoW_feats[c].begin();
            last[c] = std::prev(BoW_feats[c].end());
        }
        struct Step { unsigned int node; std::vector<int> cams; std::vector<const std::vector<unsigned int>*> feats; int group0; };
        std::vector<Step> steps;
        std::vector<int> cam_a, cam_b, a_ofs(1, 0), b_ofs(1, 0);
        std::vector<uint32_t> a_idx, b_idx;
        {
            std::vector<It> it = begin;
            auto done = [&]() {
                bool res = true;
                for (int c = 0; c < num_cams_; c++) res = res && it[c]->first >= last[c]->first;
                return res;
            };
            while (!done()) {
                long lo = (long)INT32_MAX - 1;
                Step st;
                for (int c = 0; c < num_cams_; c++) {
                    const long w = it[c] == last[c] ? (long)INT32_MAX : (long)it[c]->first;
                    if (w < lo) { lo = w; st.cams.assign(1, c); }
                    else if (w == lo) st.cams.push_back(c);
                }
                st.node = (unsigned int)lo;
                for (int c : st.cams) st.feats.push_back(&it[c]->second);
                st.group0 = (int)cam_a.size();
                const int S = (int)st.cams.size();
                if (S >= 2) {
                    for (int i = 0; i + 1 < S; i++)
                        for (int j = i + 1; j < S; j++) {
                            cam_a.push_back(st.cams[i]); cam_b.push_back(st.cams[j]);
                            a_idx.insert(a_idx.end(), st.feats[i]->begin(), st.feats[i]->end());
                            b_idx.insert(b_idx.end(), st.feats[j]->begin(), st.feats[j]->end());
                            a_ofs.push_back((int)a_idx.size()); b_ofs.push_back((int)b_idx.size());
                        }
                    steps.push_back(st);
                }
                for (int c : st.cams) ++it[c];
            }
        }
        std::vector<int32_t> best_j(a_idx.size()), d1(a_idx.size()), d2(a_idx.size());
        check(mcs_rig_bow_best2(h_, 0, (int)cam_a.size(), cam_a.data(), cam_b.data(), a_ofs.data(), a_idx.data(), b_ofs.data(), b_idx.data(),
                                50.f, best_j.data(), d1.data(), d2.data()), "mcs_rig_bow_best2");
        auto popcount256 = [&](int camx, int fx, int camy, int fy) {
            const uint8_t *x = &image_descriptors[camx][(size_t)fx * 32], *y = &image_descriptors[camy][(size_t)fy * 32];
            int d = 0;
            for (int k = 0; k < 32; k++) d += __builtin_popcount((unsigned)(x[k] ^ y[k]));
            return d;
        };
        matches.reserve(10000);
        std::vector<std::vector<int>> rows;  // rows[r][cam], mirrored into `matches` at the end
        std::vector<int> rays;
        int cur = 0;  // intraMatchInd
        for (const Step& st : steps) {
            const int S = (int)st.cams.size();
            std::vector<std::vector<int>> flags(S);
            for (int i = 0; i < S; i++) flags[i].assign(st.feats[i]->size(), -1);
            int g = st.group0;
            for (int i = 0; i + 1 < S; g += S - 1 - i, i++) {
                const int cam1 = st.cams[i];
                const std::vector<unsigned int>& f1 = *st.feats[i];
                for (size_t c1 = 0; c1 < f1.size(); c1++) {
                    if (flags[i][c1] != -1) continue;
                    bool found = false, updateOnce = true;
                    rows.emplace_back(num_cams_, -1);
                    rows.back()[cam1] = (int)f1[c1];
                    rays.push_back(1);
                    flags[i][c1] = cur;
                    for (int j = i + 1; j < S; j++) {
                        const int cam2 = st.cams[j], gg = g + (j - i - 1);
                        const std::vector<unsigned int>& f2 = *st.feats[j];
                        const int bj = best_j[a_ofs[gg] + c1];
                        const double b1 = d1[a_ofs[gg] + c1] >= 1000000000 ? 1e9 : (double)d1[a_ofs[gg] + c1];
                        const double b2 = d2[a_ofs[gg] + c1] >= 1000000000 ? 1e9 : (double)d2[a_ofs[gg] + c1];
                        if (!(b1 <= TH_LOW && b1 / b2 <= max_neighbor_ratio)) continue;
                        const int owner = flags[j][bj];
                        if (owner == cur) continue;
                        if (owner == -1) {
                            rows[cur][cam2] = (int)f2[bj]; rays[cur]++;
                            flags[j][bj] = cur;
                            found = true;
                            continue;
                        }
                        const int old_cam1 = rows[owner][cam1];
                        if (old_cam1 == -1) {  // the owning row has no cam1 feature yet: fold this row into it, once per cam1 feature (:762-795)
                            if (!updateOnce) continue;
                            updateOnce = false;
                            bool disjoint = true;
                            int inc = 0;
                            for (int t = 0; t < num_cams_ && disjoint; t++)
                                if (rows[cur][t] != -1) { if (rows[owner][t] != -1) disjoint = false; else inc++; }
                            if (disjoint) {
                                for (int t = 0; t < num_cams_; t++) if (rows[cur][t] != -1) rows[owner][t] = rows[cur][t];
                                rays[owner] += inc;
                                flags[i][c1] = owner;
                                cur = owner;
                                found = true;
                                rows.pop_back(); rays.pop_back();
                            }
                            continue;
                        }
                        if (b1 < (double)popcount256(cam1, old_cam1, cam2, (int)f2[bj])) {  // closer than the current owner: steal (:797-806)
                            rows[owner][cam2] = -1; rays[owner]--;
                            rows[cur][cam2] = (int)f2[bj]; rays[cur]++;
                            flags[j][bj] = cur;
                            found = true;
                        }
                    }
                    if (found) {
                        words_.push_back(st.node);
                        cur = (int)rows.size();
                    } else {
                        rows.pop_back(); rays.pop_back();
                        flags[i][c1] = -1;
                    }
                }
            }
        }
        for (size_t r = 0; r < rows.size(); r++) {
            IntraMatch m;
            fill_row(m, rows[r].data());
            m.n_rays = rays[r];
            matches.push_back(m);
        }
    }

    // BowMatching(cam1_ind, cam2_ind, indices_1, indices_2, kps1, kps2, words) (:946-1022): the per-node ratio-test matcher between two
    // cameras' FeatureVectors (BoW_feats); like the reference it clears the index lists but APPENDS to kps1 / kps2.
    inline void BowMatching(int cam1_ind, int cam2_ind, std::vector<unsigned int>& indices_1, std::vector<unsigned int>& indices_2,
                            std::vector<mcs_keypoint>& kps1, std::vector<mcs_keypoint>& kps2, std::set<unsigned int>& words);
    // computeIntraMatches_chained(matches, old) (:1290-1355): BowMatching between neighbouring cameras (j-1, j), |dy| < 50 gate on the
    // undistorted keypoints, rows chained through the shared camera.  n_rays is left at its default, as in the reference.
    inline void computeIntraMatches_chained(std::vector<IntraMatch>& matches, bool old);
    // computeIntraMatches(matches, kps_1, kps_2) (:1365-1466): per camera pair (i < j) the BowMatching matches as DMatch(i_idx, j_idx, 0)
    // with the (distorted) image_kps points.  x/y pairs in kps_1 / kps_2.
    inline void computeIntraMatches(std::vector<std::vector<std::vector<DMatch>>>& matches,
                                    std::vector<std::vector<std::vector<std::array<float, 2>>>>& kps_1,
                                    std::vector<std::vector<std::vector<std::array<float, 2>>>>& kps_2);

    // Triangulation block of FrontEnd::obtainLfFeatures (FrontEnd.cpp:249-370): mask filter (masks: [num_cams][h][w] f32 or null),
    // N-view DLT, depth gate.  Returns the accepted rows in order (what the reference pushes to currentFrame->intraMatches);
    // single-view rows (mono candidates, :371-384) are returned in `mono_rows`.
    void triangulateIntraMatches(const float* segMasks, std::vector<IntraMatch>& accepted, std::vector<IntraMatch>* mono_rows = nullptr) {
        check(mcs_rig_triangulate(h_, 1, segMasks), "mcs_rig_triangulate");
        const int cap = row_cap_;
        std::vector<int32_t> rows((size_t)cap * num_cams_), nv(cap), rep(cap);
        std::vector<uint8_t> acc(cap);
        std::vector<double> xyz((size_t)cap * 3);
        std::vector<float> uv((size_t)cap * 2);
        mcs_landmarks lm{rows.data(), acc.data(), nv.data(), xyz.data(), uv.data(), rep.data()};
        const int M = mcs_rig_get_landmarks(h_, 0, &lm, cap);
        check(M, "mcs_rig_get_landmarks");
        accepted.clear();
        if (mono_rows) mono_rows->clear();
        for (int r = 0; r < M; r++) {
            if (!acc[r] && !(mono_rows && nv[r] == 1)) continue;
            IntraMatch m;
            fill_row(m, &rows[(size_t)r * num_cams_]);
            m.n_rays = nv[r];
            if (rep[r] >= 0) {
                const int f = rows[(size_t)r * num_cams_ + rep[r]];
                std::memcpy(m.matchDesc.data(), &image_descriptors[rep[r]][(size_t)f * 32], 32);
            }
            if (acc[r]) {
                m.mono = false;
                std::memcpy(m.point3D, &xyz[(size_t)r * 3], 24);
                m.uv_ref[0] = uv[(size_t)r * 2]; m.uv_ref[1] = uv[(size_t)r * 2 + 1];
                accepted.push_back(m);
            } else {
                mono_rows->push_back(m);
            }
        }
    }

    // The rest of FrontEnd::obtainLfFeatures around the triangulation block (FrontEnd.cpp:213-533, shipped branch: tiling and refview
    // off, no vocabulary transform): intraMatches = accepted rows in row order, then the mono features — single-view rows (:371-384)
    // followed by every keypoint no accepted row uses, camera-major (:487-505) — ordered by descending response with the reference's
    // argsorte (utils.h:20-30: std::sort, unstable) and cut to total_feats - intramatch_size (:513-521).  lIds all -1.
    struct LfFeatures {
        std::vector<IntraMatch> intraMatches;
        std::vector<int> lIds;
        int intramatch_size = 0, mono_size = 0;
    };
    void obtainLfFeatures(const float* segMasks, LfFeatures& out, int total_feats = 3000) {
        std::vector<IntraMatch> accepted, mono_keypoints;
        triangulateIntraMatches(segMasks, accepted, &mono_keypoints);
        std::vector<std::vector<char>> keypoint_mask(num_cams_);
        for (int c = 0; c < num_cams_; c++) keypoint_mask[c].assign(image_kps_undist[c].size(), 1);
        std::vector<float> responses;
        for (const IntraMatch& m : accepted)
            for (int c = 0; c < num_cams_; c++)
                if (m.matchIndexN[c] >= 0) keypoint_mask[c][m.matchIndexN[c]] = 0;
        for (IntraMatch& m : mono_keypoints) {
            int c = 0;
            while (m.matchIndexN[c] < 0) c++;
            const int j = m.matchIndexN[c];
            std::memcpy(m.matchDesc.data(), &image_descriptors[c][(size_t)j * 32], 32);
            m.uv_ref[0] = image_kps_undist[c][j].x; m.uv_ref[1] = image_kps_undist[c][j].y;
            m.mono = true;
            responses.push_back(image_kps_undist[c][j].response);
            keypoint_mask[c][j] = 0;
        }
        for (int c = 0; c < num_cams_; c++)
            for (size_t j = 0; j < image_kps_undist[c].size(); j++) {
                if (!keypoint_mask[c][j]) continue;
                IntraMatch m;
                m.matchIndexN.assign(num_cams_, -1);
                m.matchIndexN[c] = (int)j;
                if (c < 5) m.matchIndex[c] = (int)j;
                std::memcpy(m.matchDesc.data(), &image_descriptors[c][j * 32], 32);
                m.uv_ref[0] = image_kps_undist[c][j].x; m.uv_ref[1] = image_kps_undist[c][j].y;
                m.n_rays = 1;
                m.mono = true;
                mono_keypoints.push_back(m);
                responses.push_back(image_kps_undist[c][j].response);
                keypoint_mask[c][j] = 0;
            }
        std::vector<int> sorted_kp_ind(responses.size());
        std::iota(sorted_kp_ind.begin(), sorted_kp_ind.end(), 0);
        std::sort(sorted_kp_ind.begin(), sorted_kp_ind.end(), [&responses](int i, int j) -> bool { return responses[i] > responses[j]; });
        out.intraMatches = accepted;
        out.intramatch_size = (int)accepted.size();
        out.mono_size = 0;
        for (size_t i = 0; i < sorted_kp_ind.size() && (int)i < total_feats - out.intramatch_size; i++) {
            out.intraMatches.push_back(mono_keypoints[sorted_kp_ind[i]]);
            out.mono_size++;
        }
        out.lIds.assign(out.intraMatches.size(), -1);
    }

    // triangulateIntraMatches(x, view_inds, pt) (:411-441): x = 2 x n pixel coordinates (row 0 = u, row 1 = v)
    void triangulateIntraMatches(const double* x_2xn, std::vector<int> view_inds, double pt[3]) {
        const int n = (int)view_inds.size();
        check(mcs_rig_triangulate_nview(h_, view_inds.data(), n, x_2xn, x_2xn + n, 1, pt), "mcs_rig_triangulate_nview");
    }
    // triangulate(kps_1, kps_2, pts3D, cam1, cam2) (:481-507): points as (x, y) float pairs
    void triangulate(const std::vector<std::array<float, 2>>& kps_1, const std::vector<std::array<float, 2>>& kps_2,
                     std::vector<std::array<float, 3>>& pts3D, int cam1, int cam2) {
        pts3D.resize(kps_1.size());
        if (kps_1.empty()) return;
        check(mcs_rig_triangulate_pair(h_, cam1, cam2, &kps_1[0][0], &kps_2[0][0], (int)kps_1.size(), &pts3D[0][0]), "mcs_rig_triangulate_pair");
    }

    // triangulateIntraMatches(kps_1, kps_2) (:446-472): kps_1[i][ind] / kps_2[i][ind] hold the matched pixel pairs of camera i and
    // camera j = i + 1 + ind; every pair is triangulated with the 2-view DLT above and moved into camera 0's frame
    // (R_i^T, -R_i^T t_i converted to f32, cv::Affine3f product).  Fills pts3D_ in the reference's pair order.
    void triangulateIntraMatches(const std::vector<std::vector<std::vector<std::array<float, 2>>>>& kps_1,
                                 const std::vector<std::vector<std::vector<std::array<float, 2>>>>& kps_2) {
        pts3D_.clear();
        for (int i = 0; i < num_cams_ - 1; i++) {
            int ind = 0;
            for (int j = i + 1; j < num_cams_; j++, ind++) {
                std::vector<std::array<float, 3>> pts;
                triangulate(kps_1[i][ind], kps_2[i][ind], pts, i, j);
                if (i != 0) {
                    float R[9], t[3];
                    const mcs_calib& c = calib_[i];
                    for (int r = 0; r < 3; r++) {
                        for (int k = 0; k < 3; k++) R[r * 3 + k] = (float)c.R[k * 3 + r];
                        t[r] = (float)((-c.R[0 * 3 + r]) * c.t[0] + (-c.R[1 * 3 + r]) * c.t[1] + (-c.R[2 * 3 + r]) * c.t[2]);
                    }
                    for (auto& p : pts) {
                        const float x = p[0], y = p[1], z = p[2];
                        p[0] = R[0] * x + R[1] * y + R[2] * z + t[0];
                        p[1] = R[3] * x + R[4] * y + R[5] * z + t[1];
                        p[2] = R[6] * x + R[7] * y + R[8] * z + t[2];
                    }
                }
                pts3D_.insert(pts3D_.end(), pts.begin(), pts.end());
            }
        }
    }
    std::vector<std::array<float, 3>> pts3D_;  // MultiCameraFrame.h pts3D_ (get3DPoints, :510-514)

    // Throughput form of FrontEnd::processFrame (FrontEnd.cpp:991-1028) for n_frames rig frames in one launch sequence.
    void process_frames(const uint8_t* images, const float* masks, int n_frames, bool old = false) {
        check(mcs_rig_process(h_, images, masks, n_frames, old ? 1 : 0), "mcs_rig_process");
    }

    // camconfig_.RECTIFY (MultiCameraFrame.cpp:123-136, :241-243): cv::undistort of every image and segmentation mask entering the
    // handle with the calibration's distortion coefficients; keypoints are then already undistorted (image_kps_undist = image_kps).
    void setRectify(bool on) { check(mcs_rig_set_undistort(h_, on ? 1 : 0), "mcs_rig_set_undistort"); }
    // RECTIFY = false with distorted lenses: image_kps_undist = cv::undistortPoints(image_kps) (UndistortKeyPoints, :300-347)
    void setUndistortKeypoints(bool on) { check(mcs_rig_set_undistort_keypoints(h_, on ? 1 : 0), "mcs_rig_set_undistort_keypoints"); }

    // setData (MultiCameraFrame.cpp:95-152) + processFrame for the application's native CV_32F [0, 1] images (1 or 3 BGR channels,
    // [n_frames][num_cams][h][w][channels]): multiply by 255, convertTo(CV_8U), BGR2GRAY run on the device in front of the path.
    void process_frames_f32(const float* images, int channels, const float* masks, int n_frames, bool old = false) {
        check(mcs_rig_process_f32(h_, images, channels, masks, n_frames, old ? 1 : 0), "mcs_rig_process_f32");
    }

    int num_cams_;
    std::vector<std::vector<mcs_keypoint>> image_kps, image_kps_undist;
    std::vector<std::vector<uint8_t>> image_descriptors;  // per camera n x 32 bytes
    mcs_rig* handle() { return h_; }

private:
    void fill_row(IntraMatch& m, const int32_t* row) const {
        m.matchIndexN.assign(row, row + num_cams_);
        for (int c = 0; c < 5 && c < num_cams_; c++) m.matchIndex[c] = row[c];
    }
    std::shared_ptr<BowMatcher> bow_;  // created on first use by the BoW pair matchers
    std::vector<mcs_calib> calib_;
    mcs_rig* h_ = nullptr;
    int w_, h_img_, kp_cap_ = 0, row_cap_ = 0;
    const uint8_t* images_ = nullptr;
};

// DBoW2::TemplatedVocabulary as the front-end uses it (MCSlam/include/MCSlam/ORBVocabulary.h; DBoW2 itself is an external, un-vendored
// dependency): loadFromTextFile (FrontEnd.h:137-138) and transform(features, bowVec, featVec, levelsup) (MultiCameraFrame.cpp:257,
// FrontEnd.cpp:533).  The tree descent runs on the GPU for all features of a call; word ids / weights are returned alongside the
// FeatureVector for callers that build a BowVector.
class ORBVocabulary {
public:
    explicit ORBVocabulary(int device = 0) : device_(device) {}
    ~ORBVocabulary() { mcs_vocab_destroy(h_); }
    ORBVocabulary(const ORBVocabulary&) = delete;
    ORBVocabulary& operator=(const ORBVocabulary&) = delete;
    bool loadFromTextFile(const std::string& filename) {
        mcs_vocab_destroy(h_);
        h_ = mcs_vocab_load_text(filename.c_str(), device_);
        return h_ != nullptr;
    }
    // features: n x 32 bytes.  featVec: node id (at level L - levelsup) -> feature indices in ascending order, features whose word has
    // weight 0 left out (DBoW2: `if(w > 0)`).
    void transform(const uint8_t* features, int n, FeatureVector& featVec, int levelsup, std::vector<int>* word_ids = nullptr,
                   std::vector<double>* weights = nullptr) const {
        if (!h_) throw std::runtime_error("ORBVocabulary::transform: no vocabulary loaded");
        std::vector<int32_t> wid(n > 0 ? n : 1), nid(n > 0 ? n : 1);
        std::vector<double> w(n > 0 ? n : 1);
        check(mcs_vocab_transform(h_, features, n, levelsup, wid.data(), nid.data(), w.data()), "mcs_vocab_transform");
        featVec.clear();
        for (int i = 0; i < n; i++)
            if (w[i] > 0) featVec[(unsigned int)nid[i]].push_back((unsigned int)i);
        if (word_ids) word_ids->assign(wid.begin(), wid.begin() + n);
        if (weights) weights->assign(w.begin(), w.begin() + n);
    }
    mcs_vocab* handle() const { return h_; }

private:
    mcs_vocab* h_ = nullptr;
    int device_ = 0;
};

// The BoW-guided matchers: ORBextractor::getMatches_distRatio (ORBextractor.cpp:1228-1290) and the node walk around it in
// MultiCameraFrame::BowMatching (MultiCameraFrame.cpp:946-1022) and FrontEnd::InterMatchingBow (FrontEnd.cpp:3791-3840, :3905-3965).
// All common nodes of the two feature vectors are matched in ONE kernel launch (mcs_match_subsets) instead of one call per node.
class BowMatcher {
public:
    explicit BowMatcher(int device = 0, int max_rows = 16384) {
        h_ = mcs_matcher_create(device, 1, max_rows, max_rows);
        if (!h_) throw std::runtime_error(std::string("mcs_matcher_create: ") + mcs_last_error());
    }
    ~BowMatcher() { mcs_matcher_destroy(h_); }
    BowMatcher(const BowMatcher&) = delete;
    BowMatcher& operator=(const BowMatcher&) = delete;

    double max_neighbor_ratio = 0.85;  // ORBextractor.h:90
    int TH_LOW = 75;                   // ORBextractor.h:27

    // getMatches_distRatio(A, i_A, B, i_B, i_match_A, i_match_B, BookK): A, B = descriptor rows (n x 32 bytes)
    void getMatches_distRatio(const uint8_t* A, int nA, const std::vector<unsigned int>& i_A, const uint8_t* B, int nB,
                              const std::vector<unsigned int>& i_B, std::vector<unsigned int>& i_match_A,
                              std::vector<unsigned int>& i_match_B, int& BookK) {
        const int a_ofs[2] = {0, (int)i_A.size()}, b_ofs[2] = {0, (int)i_B.size()};
        i_match_A.assign(i_A.size(), 0);
        i_match_B.assign(i_A.size(), 0);
        long long book = 0;
        const int n = mcs_match_subsets(h_, A, nA, B, nB, 1, a_ofs, i_A.data(), b_ofs, i_B.data(), TH_LOW, max_neighbor_ratio,
                                        i_match_A.data(), i_match_B.data(), nullptr, &book);
        check(n, "mcs_match_subsets");
        i_match_A.resize(n);
        i_match_B.resize(n);
        BookK += (int)book;
    }

    // BowMatching / InterMatchingBow: walk both feature vectors in node order (the lower_bound walk of the reference visits exactly
    // the node ids present in both), match every common node, concatenate in node order; `words` receives the nodes with >= 1 match.
    void BowMatching(const FeatureVector& featVec1, const FeatureVector& featVec2, const uint8_t* descs1, int n1, const uint8_t* descs2,
                     int n2, std::vector<unsigned int>& indices_1, std::vector<unsigned int>& indices_2,
                     std::set<unsigned int>* words = nullptr) {
        indices_1.clear();
        indices_2.clear();
        std::vector<int> a_ofs(1, 0), b_ofs(1, 0);
        std::vector<unsigned int> a_idx, b_idx, nodes;
        FeatureVector::const_iterator it1 = featVec1.begin(), it2 = featVec2.begin();
        while (it1 != featVec1.end() && it2 != featVec2.end()) {
            if (it1->first == it2->first) {
                nodes.push_back(it1->first);
                a_idx.insert(a_idx.end(), it1->second.begin(), it1->second.end());
                b_idx.insert(b_idx.end(), it2->second.begin(), it2->second.end());
                a_ofs.push_back((int)a_idx.size());
                b_ofs.push_back((int)b_idx.size());
                ++it1; ++it2;
            } else if (it1->first < it2->first) {
                it1 = featVec1.lower_bound(it2->first);
            } else {
                it2 = featVec2.lower_bound(it1->first);
            }
        }
        if (nodes.empty()) return;
        indices_1.assign(a_idx.size(), 0);
        indices_2.assign(a_idx.size(), 0);
        std::vector<int> counts(nodes.size(), 0);
        const int n = mcs_match_subsets(h_, descs1, n1, descs2, n2, (int)nodes.size(), a_ofs.data(), a_idx.data(), b_ofs.data(), b_idx.data(),
                                        TH_LOW, max_neighbor_ratio, indices_1.data(), indices_2.data(), counts.data(), nullptr);
        check(n, "mcs_match_subsets");
        indices_1.resize(n);
        indices_2.resize(n);
        if (words)
            for (size_t g = 0; g < nodes.size(); g++)
                if (counts[g] >= 1) words->insert(nodes[g]);
    }

private:
    mcs_matcher* h_ = nullptr;
};

// cv::DMatch stand-in (queryIdx, trainIdx, distance); with MCS_WITH_OPENCV the methods below also fill std::vector<cv::DMatch>.
struct DMatch {
    int queryIdx = -1, trainIdx = -1;
    float distance = 0.f;
};

inline void MultiCameraFrame::BowMatching(int cam1_ind, int cam2_ind, std::vector<unsigned int>& indices_1, std::vector<unsigned int>& indices_2,
                                          std::vector<mcs_keypoint>& kps1, std::vector<mcs_keypoint>& kps2, std::set<unsigned int>& words) {
    if (!bow_) bow_ = std::make_shared<BowMatcher>();
    bow_->BowMatching(BoW_feats.at(cam1_ind), BoW_feats.at(cam2_ind), image_descriptors[cam1_ind].data(),
                      (int)image_descriptors[cam1_ind].size() / 32, image_descriptors[cam2_ind].data(),
                      (int)image_descriptors[cam2_ind].size() / 32, indices_1, indices_2, &words);
    for (size_t m = 0; m < indices_1.size(); m++) {
        kps1.push_back(image_kps_undist[cam1_ind][indices_1[m]]);
        kps2.push_back(image_kps_undist[cam2_ind][indices_2[m]]);
    }
}

inline void MultiCameraFrame::computeIntraMatches_chained(std::vector<IntraMatch>& matches, bool /*old*/) {
    matches.reserve(10000);
    std::vector<std::vector<int>> inv(num_cams_);
    for (int c = 0; c < num_cams_; c++) inv[c].assign(image_kps[c].size(), -1);
    std::set<unsigned int> wordIds;
    for (int j = 1; j < num_cams_; j++) {
        const int i = j - 1;
        std::vector<unsigned int> indices1, indices2;
        std::vector<mcs_keypoint> kp1, kp2;
        BowMatching(i, j, indices1, indices2, kp1, kp2, wordIds);
        for (size_t k = 0; k < indices1.size(); k++) {
            const int fi = (int)indices1[k], fj = (int)indices2[k];
            const int row_i = inv[i][fi], row_j = inv[j][fj];
            if (std::fabs(image_kps_undist[i][fi].y - image_kps_undist[j][fj].y) >= 50) continue;
            if (row_i == -1 && row_j == -1) {
                IntraMatch m;
                m.matchIndexN.assign(num_cams_, -1);
                m.matchIndexN[i] = fi; m.matchIndexN[j] = fj;
                if (i < 5) m.matchIndex[i] = fi;
                if (j < 5) m.matchIndex[j] = fj;
                inv[i][fi] = inv[j][fj] = (int)matches.size();
                matches.push_back(m);
            } else if (row_j == -1) {  // :1340-1343 — as in the reference, row_i may be -1 here only if row_j is set, so row_i is valid
                matches[row_i].matchIndexN[j] = fj;
                if (j < 5) matches[row_i].matchIndex[j] = fj;
                inv[j][fj] = row_i;
            }
        }
    }
}

inline void MultiCameraFrame::computeIntraMatches(std::vector<std::vector<std::vector<DMatch>>>& matches,
                                                  std::vector<std::vector<std::vector<std::array<float, 2>>>>& kps_1,
                                                  std::vector<std::vector<std::vector<std::array<float, 2>>>>& kps_2) {
    if ((int)BoW_feats.size() != num_cams_) throw std::invalid_argument("computeIntraMatches: BoW_feats must hold one FeatureVector per camera");
    kps_1.clear();
    kps_2.clear();
    std::set<unsigned int> unused;
    for (int i = 0; i + 1 < num_cams_; i++) {
        matches.emplace_back();
        kps_1.emplace_back();
        kps_2.emplace_back();
        for (int j = i + 1; j < num_cams_; j++) {
            std::vector<unsigned int> i_ind, j_ind;
            std::vector<mcs_keypoint> ku1, ku2;
            BowMatching(i, j, i_ind, j_ind, ku1, ku2, unused);
            std::vector<DMatch> m_vec(i_ind.size());
            std::vector<std::array<float, 2>> p1(i_ind.size()), p2(i_ind.size());
            for (size_t k = 0; k < i_ind.size(); k++) {
                m_vec[k].queryIdx = (int)i_ind[k]; m_vec[k].trainIdx = (int)j_ind[k]; m_vec[k].distance = 0.f;
                p1[k] = {image_kps[i][i_ind[k]].x, image_kps[i][i_ind[k]].y};
                p2[k] = {image_kps[j][j_ind[k]].x, image_kps[j][j_ind[k]].y};
            }
            matches.back().push_back(m_vec);
            kps_1.back().push_back(p1);
            kps_2.back().push_back(p2);
        }
    }
}

// What FrontEnd::findInterMatches reads of a MultiCameraFrame (MultiCameraFrame.h:138-151): the frame's IntraMatch list
// (triangulated rows first, mono features appended), the landmark ids and the number of triangulated rows.
struct InterFrameView {
    const std::vector<IntraMatch>* intraMatches = nullptr;
    const std::vector<int>* lIds = nullptr;
    int intramatch_size = 0;
};

// Inter-frame matching of FrontEnd (FrontEnd.cpp:3114-3470).  The brute-force knnMatch(k = 2) runs on the GPU (mcs_knn2); the
// accept / reject rules that differ between the call sites and the order-dependent one-to-one bookkeeping stay on the host, in
// the reference's arithmetic (float distances promoted to double against 0.7 * m[1].distance).
class InterFrameMatcher {
public:
    explicit InterFrameMatcher(int device = 0, int max_rows = 16384) : max_rows_(max_rows) {
        h_ = mcs_matcher_create(device, 1, max_rows, max_rows);
        if (!h_) throw std::runtime_error(std::string("mcs_matcher_create: ") + mcs_last_error());
    }
    ~InterFrameMatcher() { mcs_matcher_destroy(h_); }
    InterFrameMatcher(const InterFrameMatcher&) = delete;
    InterFrameMatcher& operator=(const InterFrameMatcher&) = delete;

    // matcher->knnMatch(descs1, descs2, matches, 2) (FrontEnd.cpp:3135): rows of (trainIdx0, dist0, trainIdx1, dist1)
    const std::vector<int32_t>& knnMatch(const uint8_t* descs1, int n1, const uint8_t* descs2, int n2) {
        knn_.assign((size_t)(n1 > 0 ? n1 : 1) * 4, -1);
        if (n1 > 0) check(mcs_knn2(h_, descs1, &n1, descs2, &n2, 1, knn_.data()), "mcs_knn2");
        return knn_;
    }

    // findMatchesMono(lf_prev, lf_cur, cam_ind, matches) (FrontEnd.cpp:3114-3154).  descs*: image_descriptors[cam_ind] rows,
    // kps*_undist: image_kps_undist[cam_ind], segMask*: segMasks[cam_ind] as CV_32F planes of `width` columns (nullptr = all static).
    void findMatchesMono(const uint8_t* descs1, const std::vector<mcs_keypoint>& kps1_undist, const float* segMask1, const uint8_t* descs2,
                         const std::vector<mcs_keypoint>& kps2_undist, const float* segMask2, int width, std::vector<DMatch>& matches) {
        matches.clear();
        const int n1 = (int)kps1_undist.size(), n2 = (int)kps2_undist.size();
        const std::vector<int32_t>& k = knnMatch(descs1, n1, descs2, n2);
        for (int q = 0; q < n1; q++) {
            const int32_t* m = &k[(size_t)q * 4];
            if (m[2] < 0) continue;  // fewer than two train descriptors: the reference would index m[1] out of range
            const float d0 = (float)m[1], d1 = (float)m[3];
            if (d0 < 0.7 * d1) {
                if (d0 > 50) continue;
                const mcs_keypoint &p = kps1_undist[q], &c = kps2_undist[m[0]];
                const float s1 = segMask1 ? segMask1[(size_t)(int)p.y * width + (int)p.x] : 0.f;
                const float s2 = segMask2 ? segMask2[(size_t)(int)c.y * width + (int)c.x] : 0.f;
                if (s1 < 0.7 && s2 < 0.7) matches.push_back(DMatch{q, m[0], d0});
            }
        }
    }

    // findInterMatches(lf_prev, lf_cur, matches_z_filtered, matches_mono_z_filtered, viz) (FrontEnd.cpp:3157-3343).  Like the
    // reference only matches_z_filtered is cleared on entry.
    void findInterMatches(const InterFrameView& prev, const InterFrameView& cur, std::vector<DMatch>& matches_z_filtered,
                          std::vector<DMatch>& matches_mono_z_filtered) {
        inter(prev, cur, matches_z_filtered, &matches_mono_z_filtered);
    }
    // findInterMatches(lf_prev, lf_cur, matches_z_filtered, viz) (FrontEnd.cpp:3344-3470)
    void findInterMatches(const InterFrameView& prev, const InterFrameView& cur, std::vector<DMatch>& matches_z_filtered) {
        inter(prev, cur, matches_z_filtered, nullptr);
    }

private:
    void inter(const InterFrameView& prev, const InterFrameView& cur, std::vector<DMatch>& lf, std::vector<DMatch>* mono) {
        const std::vector<IntraMatch>&a = *prev.intraMatches, &b = *cur.intraMatches;
        const int n1 = (int)a.size(), n2 = (int)b.size();
        d1_.resize((size_t)n1 * 32);
        d2_.resize((size_t)n2 * 32);
        for (int i = 0; i < n1; i++) std::memcpy(&d1_[(size_t)i * 32], a[i].matchDesc.data(), 32);
        for (int i = 0; i < n2; i++) std::memcpy(&d2_[(size_t)i * 32], b[i].matchDesc.data(), 32);
        const std::vector<int32_t>& k = knnMatch(d1_.data(), n1, d2_.data(), n2);
        lf.clear();
        std::vector<int> inds1, inds2, mono_inds1, mono_inds2;
        for (int q = 0; q < n1; q++) {
            const int32_t* m = &k[(size_t)q * 4];
            if (m[0] < 0) continue;
            const int t = m[0];
            const float d0 = (float)m[1];
            if ((*prev.lIds)[q] == -1) {  // :3212-3218: landmarks skip the ratio test
                if (m[2] < 0) continue;
                if (d0 > 0.7 * (float)m[3]) continue;
            }
            if (!a[q].mono && !b[t].mono) {  // :3221-3232
                const double dx = a[q].point3D[0] - b[t].point3D[0], dy = a[q].point3D[1] - b[t].point3D[1],
                             dz = a[q].point3D[2] - b[t].point3D[2];
                const float distance = (float)std::sqrt(dx * dx + dy * dy + dz * dz);
                if (!(distance <= 2.0)) continue;
            }
            int pos_lf = -1, pos_mono = -1;
            if (!mono || t < cur.intramatch_size) pos_lf = find(inds2, t);
            if (mono) pos_mono = find(mono_inds2, t);
            const DMatch dm{q, t, d0};
            if (pos_lf < 0 && pos_mono < 0) {
                if (!mono || (q < prev.intramatch_size && t < cur.intramatch_size)) {
                    inds2.push_back(t); inds1.push_back(q); lf.push_back(dm);
                } else {
                    mono_inds2.push_back(t); mono_inds1.push_back(q); mono->push_back(dm);
                }
            } else if (pos_mono < 0) {  // train feature already taken by an intra-match inter-match: keep the closer one
                if (d0 < lf[pos_lf].distance) { inds1[pos_lf] = q; lf[pos_lf] = dm; }
            } else {
                if (d0 < (*mono)[pos_mono].distance) { mono_inds1[pos_mono] = q; (*mono)[pos_mono] = dm; }
            }
        }
    }
    static int find(const std::vector<int>& v, int x) {
        for (size_t i = 0; i < v.size(); i++) if (v[i] == x) return (int)i;
        return -1;
    }
    mcs_matcher* h_ = nullptr;
    int max_rows_ = 0;
    std::vector<int32_t> knn_;
    std::vector<uint8_t> d1_, d2_;
};

}  // namespace mcs_b200
#endif
