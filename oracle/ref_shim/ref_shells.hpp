// ref_shells.hpp — class shells around the reference method bodies that oracle/ref_build.py extracts by line range.
// TEST INFRASTRUCTURE ONLY.
//
// MultiCameraFrame.h / FrontEnd.h cannot be included (they pull in Eigen, OpenGV, GTSAM, ROS, DBoW2, glog), so the two classes
// are re-declared here with exactly the members the extracted bodies touch, under the reference's names and types
// (MultiCameraFrame.h:59-183, FrontEnd.h:644-683).  The method BODIES are the reference's own text, included verbatim from
// oracle/_ref/gen/*.inc; `class IntraMatch` and `argsorte` are extracted text too.  CamArrayConfig.h is the reference's header.
#pragma once
#include <array>
#include <chrono>
#include <list>
#include <map>
#include <mutex>
#include <numeric>
#include <set>
#include <thread>
#include <vector>

#include <opencv2/opencv.hpp>

#include "MCDataUtils/CamArrayConfig.h"  // /root/reference/MCDataUtils/include (compiled as is)
#include "MCSlam/ORBextractor.h"         // /root/reference/MCSlam/include (compiled as is)

using namespace std;
using namespace cv;

// glog's VLOG(n) << ... : swallowed
struct RefNullStream {
    template <typename T> RefNullStream& operator<<(const T&) { return *this; }
    RefNullStream& operator<<(std::ostream& (*)(std::ostream&)) { return *this; }
};
#define VLOG(n) RefNullStream()

namespace DBoW2 {  // only the names the extracted bodies mention; the vocabulary itself is out of this checker's scope
typedef unsigned int NodeId;
typedef unsigned int WordId;
struct BowVector : std::map<WordId, double> {};
struct FeatureVector : std::map<NodeId, std::vector<unsigned int>> {};
struct ORBVocabulary {
    // TemplatedVocabulary::transform(features, bowVec, featVec, levelsup): results unused by the checked outputs
    void transform(const std::vector<cv::Mat>&, BowVector&, FeatureVector&, int) const {}
};
}  // namespace DBoW2
typedef DBoW2::ORBVocabulary ORBVocabulary;

namespace Eigen {
struct Vector3d {
    double v[3] = {0, 0, 0};
    double& operator[](int i) { return v[i]; }
    const double& operator[](int i) const { return v[i]; }
};
}  // namespace Eigen

Mat build_Rt(Mat R, Mat t);  // common_utils/src/tools.cpp:136-148 (extracted)

#include "gen/utils_argsorte.inc"   // MCSlam/include/MCSlam/utils.h:20-30
#include "gen/IntraMatch.inc"       // MCSlam/include/MCSlam/MultiCameraFrame.h:41-57

class MultiCameraFrame {
public:
    explicit MultiCameraFrame(CamArrayConfig& camconfig) : num_cams_(camconfig.num_cams_), img_size(camconfig.im_size_), camconfig_(camconfig) {}

    // bodies: oracle/_ref/gen/mcf_*.inc
    void extractFeaturesParallel();
    void extractFeatureSingle(int cam_ind);
    void UndistortKeyPoints(int cam);
    void computeIntraMatches(std::vector<IntraMatch>& matches, bool old);
    void BruteForceMatch(int img1_ind, int img2_ind, float dist_thresh, float neigh_ratio, vector<unsigned int>& indices_1,
                         vector<unsigned int>& indices_2, vector<KeyPoint>& kps1, vector<KeyPoint>& kps2);
    void computeRepresentativeDesc(vector<cv::Mat> descs, cv::Mat& desc_out);
    void triangulate(vector<Point2f>& kps_1, vector<Point2f>& kps_2, vector<Point3f>& pts3D, int cam1, int cam2);
    void triangulateIntraMatches(const Mat_<double>& x, vector<int> view_inds, Eigen::Vector3d& pt);

    int num_cams_;
    Size img_size;
    vector<Mat> imgs;
    vector<Mat> segMasks;
    CamArrayConfig& camconfig_;
    ORBextractor* orBextractor = nullptr;
    vector<ORBextractor*> orBextractors;
    DBoW2::ORBVocabulary* orb_vocabulary = nullptr;
    vector<DBoW2::FeatureVector> BoW_feats;
    vector<DBoW2::BowVector> BoW_vecs;
    vector<vector<cv::KeyPoint>> image_kps, image_kps_undist;
    vector<vector<cv::Mat>> image_descriptors;
    int intramatch_size = 0, mono_size = 0;
    vector<int> lIds;
    std::vector<IntraMatch> intraMatches;
    DBoW2::BowVector lfBoW;
    DBoW2::FeatureVector lfFeatVec;
    vector<Point3f> pts3D_;
};

class FrontEnd {
public:
    explicit FrontEnd(CamArrayConfig& camconfig) : camconfig_(camconfig) {}

    // body: oracle/_ref/gen/frontend_obtainLfFeatures.inc (FrontEnd.cpp:213-593)
    void obtainLfFeatures(std::vector<IntraMatch>& matches_map, MultiCameraFrame* lf_frame, std::vector<IntraMatch>& filtered_intra_matches,
                          vector<DBoW2::NodeId>& words_, set<DBoW2::NodeId>& filtered_words);

    MultiCameraFrame* currentFrame = nullptr;
    CamArrayConfig& camconfig_;
    ORBextractor* orBextractor = nullptr;
    ORBVocabulary* orb_vocabulary = nullptr;
    vector<Mat> overlap_masks;
};
