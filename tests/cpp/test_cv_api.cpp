// Drives mc-slam_b200/host/mcslam_cv.hpp — the cv::/Eigen-typed adaptor layer — the way MCSlam's front-end drives the reference classes.
// Built against an OpenCV stand-in (oracle/ref_shim/opencv2: functional Mat / KeyPoint / InputArray, test infrastructure) and
// tests/cpp/stub_eigen, because neither library is installed here; with the real headers the same file compiles unchanged.
//   argv: <raw u8 images [C][h][w]> C w h <out file> [masks f32 [C][h][w]]
// Output (text): kp / row / lm / mono lines in the format of test_adaptor.cpp plus the results of the consumer patterns below.
#include "../../mc-slam_b200/host/mcslam_cv.hpp"

#include <fstream>
#include <iostream>

using namespace std;
using namespace cv;
using mcs_b200::ORBextractor;
namespace api = mcs_b200::cvapi;

#ifdef MCS_REFERENCE_SNIPPETS
// Lines cut from the reference by line range at test time (tests/test_host_side.py::test_cv_api_compiles_reference_call_sites), compiled
// verbatim against the typed layer.  Not run: they only have to compile.
struct NullLog {
    template <class T> NullLog& operator<<(const T&) { return *this; }
    NullLog& operator<<(std::ostream& (*)(std::ostream&)) { return *this; }
};
#define VLOG(n) NullLog()
namespace DBoW2 { struct BowVector {}; struct FeatureVector {}; }
struct VocabShell { void transform(const std::vector<cv::Mat>&, DBoW2::BowVector&, DBoW2::FeatureVector&, int) {} };
struct ExtractShell {  // the members MultiCameraFrame::extractFeatureSingle touches (MultiCameraFrame.h)
    vector<ORBextractor*> orBextractors;
    vector<Mat> imgs;
    vector<vector<KeyPoint>> image_kps, image_kps_undist;
    vector<vector<Mat>> image_descriptors;
    vector<DBoW2::FeatureVector> BoW_feats;
    vector<DBoW2::BowVector> BoW_vecs;
    VocabShell* orb_vocabulary;
    struct { bool RECTIFY = true; } camconfig_;
    void UndistortKeyPoints(int) {}
    void extractFeatureSingle(int cam_ind) {
#include "ref_mcf_extractFeatureSingle_body.inc"  // MultiCameraFrame.cpp:231-262, the function body
    }
};
void inter_match_consumer(api::MultiCameraFrame* lf_prev, api::MultiCameraFrame* lf_cur) {
#include "ref_frontend_descs_from_intramatches.inc"  // FrontEnd.cpp:3165-3181
    (void)descs1; (void)descs2;
}
float depth_gate_consumer(api::MultiCameraFrame* lf_prev, api::MultiCameraFrame* lf_cur, std::vector<DMatch> m) {
#include "ref_frontend_point3d_diff.inc"  // FrontEnd.cpp:3224-3225
    return distance;
}
#endif

int main(int argc, char** argv) {
    if (argc < 6) { fprintf(stderr, "usage: %s images C w h out [masks]\n", argv[0]); return 2; }
    const int C = atoi(argv[2]), w = atoi(argv[3]), h = atoi(argv[4]);
    std::vector<uint8_t> img((size_t)C * w * h);
    std::ifstream(argv[1], std::ios::binary).read((char*)img.data(), img.size());
    std::vector<mcs_calib> calib(C);
    std::ifstream(std::string(argv[1]) + ".calib", std::ios::binary).read((char*)calib.data(), C * sizeof(mcs_calib));
    std::vector<float> maskbuf;
    if (argc >= 7) {
        maskbuf.resize((size_t)C * w * h);
        std::ifstream(argv[6], std::ios::binary).read((char*)maskbuf.data(), maskbuf.size() * sizeof(float));
    }
    try {
        FILE* f = fopen(argv[5], "w");
        // (1) ORBextractor::operator()(InputArray, InputArray, vector<KeyPoint>&, OutputArray, vector<int>&) as extractFeatureSingle calls it
        ORBextractor orb(400, 1.2f, 8, 20, 7, 0, w, h);
        vector<KeyPoint> kps;
        Mat descs;
        vector<int> vLapping = {0, 0};
        Mat im0(h, w, CV_8UC1, img.data());
        const int mono = orb(im0, Mat(), kps, descs, vLapping);
        fprintf(f, "orb %d %d %d %d\n", mono, (int)kps.size(), descs.rows, descs.cols);
        std::vector<Mat> vec_desc;
        vec_desc.reserve(descs.rows);
        for (int j = 0; j < descs.rows; j++) vec_desc.push_back(descs.row(j));

        // (2) the frame: setData / extractFeaturesParallel / computeIntraMatches / obtainLfFeatures
        api::MultiCameraFrame fr(calib, w, h, 400, 1.2f, 8, 20, 7);
        vector<Mat> imgs, segs;
        for (int c = 0; c < C; c++) imgs.push_back(Mat(h, w, CV_8UC1, img.data() + (size_t)c * w * h));
        for (int c = 0; c < C && !maskbuf.empty(); c++) segs.push_back(Mat(h, w, CV_32FC1, maskbuf.data() + (size_t)c * w * h));
        fr.setData(imgs, segs);
        fr.extractFeaturesParallel();
        for (int c = 0; c < C; c++) {
            fprintf(f, "cam %d %d %d\n", c, (int)fr.image_kps[c].size(), (int)fr.image_descriptors[c].size());
            for (size_t j = 0; j < fr.image_kps[c].size(); j++) {
                const KeyPoint& k = fr.image_kps[c][j];
                const Mat& d = fr.image_descriptors[c][j];  // 1 x 32 CV_8U row
                fprintf(f, "kp %d %.9g %.9g %d %.9g %d %d %d %d\n", c, k.pt.x, k.pt.y, k.octave, k.response, (int)d.at<uchar>(0, 0), (int)d.at<uchar>(0, 31),
                        d.rows, d.cols);
            }
        }
        // camera 0 through the typed ORBextractor equals camera 0 of the frame
        int same0 = (int)kps.size() == (int)fr.image_kps[0].size();
        for (size_t j = 0; same0 && j < kps.size(); j++)
            same0 = kps[j].pt.x == fr.image_kps[0][j].pt.x && kps[j].pt.y == fr.image_kps[0][j].pt.y &&
                    memcmp(vec_desc[j].ptr(0), fr.image_descriptors[0][j].ptr(0), 32) == 0;
        fprintf(f, "same0 %d\n", same0);

        vector<api::IntraMatch> matches;
        fr.computeIntraMatches(matches, false);
        for (const api::IntraMatch& m : matches) {
            fprintf(f, "row");
            for (int c = 0; c < C; c++) fprintf(f, " %d", m.matchIndex[c]);
            fprintf(f, "\n");
        }
        vector<unsigned int> i1, i2;
        vector<KeyPoint> k1, k2;
        fr.BruteForceMatch(0, 1, 75.f, 0.85f, i1, i2, k1, k2);
        fprintf(f, "bf %d", (int)i1.size());
        for (size_t k = 0; k < i1.size(); k++) fprintf(f, " %u:%u", i1[k], i2[k]);
        fprintf(f, "\n");
        int bf_ok = 1;
        for (size_t k = 0; k < i1.size(); k++)
            bf_ok &= k1[k].pt.x == fr.image_kps_undist[0][i1[k]].pt.x && k2[k].pt.y == fr.image_kps_undist[1][i2[k]].pt.y;
        fprintf(f, "bfkps %d\n", bf_ok);

        fr.obtainLfFeatures(700);
        fprintf(f, "lfsizes %d %d %d %d\n", fr.intramatch_size, fr.mono_size, (int)fr.intraMatches.size(), (int)fr.lIds.size());
        // consumer patterns of the front-end: im.point3D.at<double>(k, 0) (FrontEnd.cpp:3992), d.matchDesc.copyTo(descs1.row(ind)) (:3174),
        // KeyPoint(d.uv_ref, 3.0) (:3175), point3D differences (:3224)
        Mat descs1 = Mat((int)fr.intraMatches.size(), fr.intraMatches[0].matchDesc.cols, CV_8U);
        vector<KeyPoint> kps1;
        int ind = 0;
        for (auto& d : fr.intraMatches) {
            d.matchDesc.copyTo(descs1.row(ind));
            kps1.push_back(KeyPoint(d.uv_ref, 3.0));
            ind++;
        }
        for (int r = 0; r < (int)fr.intraMatches.size(); r++) {
            const api::IntraMatch& im = fr.intraMatches[r];
            if (!im.mono) {
                fprintf(f, "lm %d %.17g %.17g %.17g %.9g %.9g %d %d\n", im.n_rays, im.point3D.at<double>(0, 0), im.point3D.at<double>(1, 0),
                        im.point3D.at<double>(2, 0), kps1[r].pt.x, kps1[r].pt.y, (int)descs1.at<uchar>(r, 0), (int)descs1.at<uchar>(r, 31));
            } else {
                int c = 0;
                while (im.matchIndex[c] < 0) c++;
                fprintf(f, "lfmono %d %d %d %d\n", c, im.matchIndex[c], im.mono ? 1 : 0, (int)descs1.at<uchar>(r, 0));
            }
        }
        if (fr.intramatch_size >= 2) {
            cv::Mat dif = fr.intraMatches[0].point3D - fr.intraMatches[1].point3D;
            fprintf(f, "dif %.17g %.17g %.17g\n", dif.at<double>(0, 0), dif.at<double>(1, 0), dif.at<double>(2, 0));
        }
        // (3) triangulateIntraMatches(const Mat_<double>& x, vector<int> view_inds, Eigen::Vector3d& pt) on the first accepted row with >= 2 views
        for (const api::IntraMatch& im : fr.intraMatches) {
            if (im.mono) continue;
            vector<int> view_inds;
            for (int c = 0; c < C; c++)
                if (im.matchIndex[c] >= 0) view_inds.push_back(c);
            Mat_<double> x(2, (int)view_inds.size());
            for (size_t k = 0; k < view_inds.size(); k++) {
                const Point2f p = fr.image_kps_undist[view_inds[k]][im.matchIndex[view_inds[k]]].pt;
                x(0, (int)k) = p.x;
                x(1, (int)k) = p.y;
            }
            Eigen::Vector3d pt;
            fr.triangulateIntraMatches(x, view_inds, pt);
            fprintf(f, "tri %.17g %.17g %.17g\n", pt[0], pt[1], pt[2]);
            // computeRepresentativeDesc(vector<Mat> descs, Mat& desc_out) on the rows of the same match
            vector<Mat> dd;
            for (int c : view_inds) dd.push_back(fr.image_descriptors[c][im.matchIndex[c]]);
            Mat rep;
            fr.computeRepresentativeDesc(dd, rep);
            fprintf(f, "rep %d\n", memcmp(rep.ptr(0), im.matchDesc.ptr(0), 32) == 0);
            break;
        }
        // (4) triangulate(vector<Point2f>&, vector<Point2f>&, vector<Point3f>&, cam1, cam2) on the BruteForceMatch pairs
        vector<Point2f> p1, p2;
        for (size_t k = 0; k < k1.size(); k++) { p1.push_back(k1[k].pt); p2.push_back(k2[k].pt); }
        vector<Point3f> p3;
        fr.triangulate(p1, p2, p3, 0, 1);
        for (const Point3f& p : p3) fprintf(f, "p3 %.9g %.9g %.9g\n", p.x, p.y, p.z);
        fclose(f);
    } catch (const std::exception& e) {
        fprintf(stderr, "%s\n", e.what());
        return 3;
    }
    return 0;
}
