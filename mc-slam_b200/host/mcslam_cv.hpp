// mcslam_cv.hpp — the cv::/Eigen-typed face of the adaptor: the reference's own member names and method signatures, so FrontEnd
// code that touches `lf_frame->image_kps[i][j].pt`, `image_descriptors[i][j]` (one cv::Mat row per feature), `im.matchDesc`,
// `im.point3D.at<double>(2, 0)` or calls BruteForceMatch / triangulate with cv containers compiles against it unchanged.
//   IntraMatch                     /root/reference/MCSlam/include/MCSlam/MultiCameraFrame.h:42-57
//   MultiCameraFrame (members)     /root/reference/MCSlam/include/MCSlam/MultiCameraFrame.h (image_kps, image_kps_undist, image_descriptors,
//                                  segMasks, intraMatches, lIds, intramatch_size, mono_size, num_cams_)
//   extractFeaturesParallel        /root/reference/MCSlam/src/MultiCameraFrame.cpp:203-262
//   BruteForceMatch                :1024-1086      computeIntraMatches(matches, old)   :1100-1288
//   triangulateIntraMatches(x, view_inds, pt)      :411-441
//   triangulate(kps_1, kps_2, pts3D, cam1, cam2)   :481-507
//   computeRepresentativeDesc      :530-567
//   obtainLfFeatures               /root/reference/MCSlam/src/FrontEnd.cpp:213-533 (the loop a maintainer replaces by one call)
// Needs OpenCV core + features2d and Eigen/Core headers (the reference's own dependencies).  It is a thin conversion layer over
// mcs_b200::MultiCameraFrame (mcslam_adaptor.hpp): every method launches the sm_100a kernels through the C ABI; nothing is computed
// on the host except container conversion and the 5 x 5 median pick of computeRepresentativeDesc when it is called on its own.
// tests/cpp/test_cv_api.cpp compiles and runs it against oracle/ref_shim's OpenCV stand-in together with call-site lines cut
// from the reference by line range.
#ifndef MCSLAM_CV_HPP
#define MCSLAM_CV_HPP

#ifndef MCS_WITH_OPENCV
#define MCS_WITH_OPENCV
#endif
#include <opencv2/core/core.hpp>
#include <opencv2/features2d/features2d.hpp>
#include <Eigen/Core>

#include <climits>

#include "mcslam_adaptor.hpp"

namespace mcs_b200 {
namespace cvapi {

// MultiCameraFrame.h:42-57, member for member
class IntraMatch {
public:
    std::array<int, 5> matchIndex;
    cv::Point2f uv_ref;
    bool mono;
    int n_rays;

    cv::Mat matchDesc;
    cv::Mat point3D;

    IntraMatch() : matchIndex{{-1, -1, -1, -1, -1}}, mono(true), n_rays(0) {}
};

static_assert(sizeof(cv::KeyPoint) == sizeof(mcs_keypoint), "mcs_keypoint mirrors cv::KeyPoint");

inline std::vector<cv::KeyPoint> to_cv(const std::vector<mcs_keypoint>& k) {
    std::vector<cv::KeyPoint> out(k.size());
    if (!k.empty()) std::memcpy((void*)out.data(), k.data(), k.size() * sizeof(mcs_keypoint));
    return out;
}

class MultiCameraFrame {
public:
    // calibration as K / R / t / dist per camera (CamArrayConfig::K_mats_, R_mats_, t_mats_, dist_coeffs_); <= 5 cameras, like IntraMatch
    MultiCameraFrame(const std::vector<mcs_calib>& calib, int width, int height, int nfeatures, float scaleFactor, int nlevels, int iniThFAST,
                     int minThFAST, int device = 0)
        : num_cams_((int)calib.size()), impl_(calib, width, height, nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, device, 1), w_(width),
          h_(height) {
        if (num_cams_ > 5) throw std::runtime_error("cvapi::MultiCameraFrame: IntraMatch::matchIndex holds 5 cameras (MultiCameraFrame.h:44)");
        image_kps.resize(num_cams_); image_kps_undist.resize(num_cams_); image_descriptors.resize(num_cams_);
    }

    // setData (MultiCameraFrame.cpp:95-152) for CV_8UC1 images (the application's CV_32F images go through
    // mcs_b200::MultiCameraFrame::process_frames_f32, which runs the conversion on the device); segmap_set: CV_32FC1 or empty
    void setData(const std::vector<cv::Mat>& img_set, const std::vector<cv::Mat>& segmap_set) {
        if ((int)img_set.size() != num_cams_) throw std::runtime_error("setData: one image per camera");
        imgs = img_set;
        segMasks = segmap_set;
        pixels_.resize((size_t)num_cams_ * w_ * h_);
        for (int c = 0; c < num_cams_; c++) {
            const cv::Mat& m = imgs[c];
            if (m.type() != CV_8UC1 || m.cols != w_ || m.rows != h_) throw std::runtime_error("setData: CV_8UC1 images of the configured size");
            for (int y = 0; y < h_; y++) std::memcpy(&pixels_[((size_t)c * h_ + y) * w_], m.ptr<uint8_t>(y), (size_t)w_);
        }
        impl_.setData(pixels_.data());
    }

    void extractFeaturesParallel() {
        impl_.extractFeaturesParallel();
        for (int c = 0; c < num_cams_; c++) {
            image_kps[c] = to_cv(impl_.image_kps[c]);
            image_kps_undist[c] = to_cv(impl_.image_kps_undist[c]);
            const int n = (int)image_kps[c].size();
            // one Mat per camera, one row header per feature: what `vec_desc.push_back(descs.row(j))` leaves (:252-254)
            cv::Mat descs(std::max(n, 1), 32, CV_8U);
            if (n) std::memcpy(descs.data, impl_.image_descriptors[c].data(), (size_t)n * 32);
            image_descriptors[c].clear();
            image_descriptors[c].reserve(n);
            for (int j = 0; j < n; j++) image_descriptors[c].push_back(descs.row(j));
        }
    }
    void extractFeatures() { extractFeaturesParallel(); }

    void BruteForceMatch(int img1_ind, int img2_ind, float dist_thresh, float neigh_ratio, std::vector<unsigned int>& indices_1,
                         std::vector<unsigned int>& indices_2, std::vector<cv::KeyPoint>& kps_1, std::vector<cv::KeyPoint>& kps_2) {
        std::vector<mcs_keypoint> k1, k2;
        impl_.BruteForceMatch(img1_ind, img2_ind, dist_thresh, neigh_ratio, indices_1, indices_2, k1, k2);
        kps_1 = to_cv(k1);
        kps_2 = to_cv(k2);
    }

    void computeIntraMatches(std::vector<IntraMatch>& matches, bool old) {
        std::vector<mcs_b200::IntraMatch> rows;
        impl_.computeIntraMatches(rows, old);
        matches.assign(rows.size(), IntraMatch());
        for (size_t r = 0; r < rows.size(); r++) {
            matches[r].matchIndex = rows[r].matchIndex;
            matches[r].n_rays = rows[r].n_rays;
        }
    }

    // x: 2 x n pixel coordinates, column k = view view_inds[k] (:411-441)
    void triangulateIntraMatches(const cv::Mat_<double>& x, std::vector<int> view_inds, Eigen::Vector3d& pt) {
        const int n = (int)view_inds.size();
        std::vector<double> xy((size_t)2 * n);
        for (int k = 0; k < n; k++) {
            xy[k] = x(0, k);
            xy[n + k] = x(1, k);
        }
        double p[3];
        impl_.triangulateIntraMatches(xy.data(), view_inds, p);
        pt[0] = p[0]; pt[1] = p[1]; pt[2] = p[2];
    }

    void triangulate(std::vector<cv::Point2f>& kps_1, std::vector<cv::Point2f>& kps_2, std::vector<cv::Point3f>& pts3D, int cam1, int cam2) {
        std::vector<std::array<float, 2>> a(kps_1.size()), b(kps_2.size());
        for (size_t i = 0; i < a.size(); i++) a[i] = {{kps_1[i].x, kps_1[i].y}};
        for (size_t i = 0; i < b.size(); i++) b[i] = {{kps_2[i].x, kps_2[i].y}};
        std::vector<std::array<float, 3>> p;
        impl_.triangulate(a, b, p, cam1, cam2);
        pts3D.clear();
        for (const auto& q : p) pts3D.push_back(cv::Point3f(q[0], q[1], q[2]));
    }

    // the descriptor with the least median Hamming distance to the others (:530-567); on the path the triangulation kernel makes this
    // choice (mcs_landmarks::rep_cam) — this form is for callers that hold the cv::Mat rows
    void computeRepresentativeDesc(std::vector<cv::Mat> descs, cv::Mat& desc_out) {
        const size_t N = descs.size();
        if (!N) return;
        std::vector<std::vector<int>> D(N, std::vector<int>(N, 0));
        for (size_t i = 0; i < N; i++)
            for (size_t j = i + 1; j < N; j++) D[i][j] = D[j][i] = ORBextractor::DescriptorDistance(descs[i].ptr<uint8_t>(0), descs[j].ptr<uint8_t>(0));
        int BestMedian = INT_MAX, BestIdx = 0;
        for (size_t i = 0; i < N; i++) {
            std::vector<int> v = D[i];
            std::sort(v.begin(), v.end());
            const int median = v[(size_t)(0.5 * (N - 1))];
            if (median < BestMedian) { BestMedian = median; BestIdx = (int)i; }
        }
        desc_out = descs[BestIdx].clone();
    }

    // FrontEnd::obtainLfFeatures (FrontEnd.cpp:213-533, shipped branch) on the intra-matches of the last computeIntraMatches call:
    // fills intraMatches (accepted rows, then the mono top-up), lIds, intramatch_size, mono_size exactly as the reference leaves them in
    // currentFrame / lf_frame.  segMasks (setData) are applied in the triangulation kernel.
    void obtainLfFeatures(int total_feats = 3000) {
        const float* masks = nullptr;
        if (!segMasks.empty()) {
            if ((int)segMasks.size() != num_cams_) throw std::runtime_error("obtainLfFeatures: one mask per camera");
            masks_.resize((size_t)num_cams_ * w_ * h_);
            for (int c = 0; c < num_cams_; c++) {
                const cv::Mat& m = segMasks[c];
                if (m.type() != CV_32FC1 || m.cols != w_ || m.rows != h_) throw std::runtime_error("obtainLfFeatures: CV_32FC1 masks of the image size");
                for (int y = 0; y < h_; y++) std::memcpy(&masks_[((size_t)c * h_ + y) * w_], m.ptr<float>(y), (size_t)w_ * sizeof(float));
            }
            masks = masks_.data();
        }
        mcs_b200::MultiCameraFrame::LfFeatures lf;
        impl_.obtainLfFeatures(masks, lf, total_feats);
        intraMatches.assign(lf.intraMatches.size(), IntraMatch());
        for (size_t r = 0; r < lf.intraMatches.size(); r++) {
            const mcs_b200::IntraMatch& s = lf.intraMatches[r];
            IntraMatch& d = intraMatches[r];
            d.matchIndex = s.matchIndex;
            d.uv_ref = cv::Point2f(s.uv_ref[0], s.uv_ref[1]);
            d.mono = s.mono;
            d.n_rays = s.n_rays;
            d.matchDesc = cv::Mat(1, 32, CV_8U);
            std::memcpy(d.matchDesc.data, s.matchDesc.data(), 32);
            if (!s.mono) {  // pt3d_sfm: 3 x 1 CV_64F (FrontEnd.cpp:351)
                d.point3D = cv::Mat(3, 1, CV_64F);
                for (int i = 0; i < 3; i++) d.point3D.at<double>(i, 0) = s.point3D[i];
            }
        }
        lIds = lf.lIds;
        intramatch_size = lf.intramatch_size;
        mono_size = lf.mono_size;
    }

    int num_cams_;
    std::vector<cv::Mat> imgs, segMasks;
    std::vector<std::vector<cv::KeyPoint>> image_kps, image_kps_undist;
    std::vector<std::vector<cv::Mat>> image_descriptors;
    std::vector<IntraMatch> intraMatches;
    std::vector<int> lIds;
    int intramatch_size = 0, mono_size = 0;
    mcs_b200::MultiCameraFrame& impl() { return impl_; }

private:
    mcs_b200::MultiCameraFrame impl_;
    int w_, h_;
    std::vector<uint8_t> pixels_;
    std::vector<float> masks_;
};

}  // namespace cvapi
}  // namespace mcs_b200

#endif  // MCSLAM_CV_HPP
