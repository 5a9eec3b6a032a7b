// ref_abi.cpp — plain-C entry points over the compiled reference classes, for ctypes (oracle/ref_lib.py).
// TEST INFRASTRUCTURE ONLY.  This file only marshals arrays in and out; every decision is taken by the reference's own code
// (ORBextractor.cpp compiled whole; MultiCameraFrame / FrontEnd method bodies extracted by line range, see oracle/ref_build.py).
#include <cmath>
#include <cstring>

#include "ref_shells.hpp"

namespace {

struct RefCalib {  // same layout as mcs_calib / the oracle's Calib: K[9], R[9], t[3], dist[5]
    double K[9], R[9], t[3], dist[5];
};

thread_local std::string g_err;

// protected members of ORBextractor are reachable from a derived class; the object stays a plain ORBextractor for its callers
struct Probe : ORBextractor {
    using ORBextractor::ORBextractor;
    using ORBextractor::DistributeOctTree;
    using ORBextractor::mnFeaturesPerLevel;
    using ORBextractor::umax;
};

struct RefRig {
    CamArrayConfig cfg;
    DBoW2::ORBVocabulary voc;
    std::vector<ORBextractor*> orbs;
    ORBextractor* orb = nullptr;
    MultiCameraFrame* frame = nullptr;
    FrontEnd* fe = nullptr;
    std::vector<IntraMatch> matches;  // computeIntraMatches output, obtainLfFeatures input
    std::vector<std::vector<uint8_t>> desc_store;  // descriptors injected through ref_rig_set_features
    int w = 0, h = 0;
    ~RefRig() {
        delete fe;
        delete frame;
        for (auto* o : orbs) delete o;
        delete orb;
    }
};

Mat mat64(int r, int c, const double* v) {
    Mat m(r, c, CV_64F);
    for (int i = 0; i < r; i++) for (int j = 0; j < c; j++) m.at<double>(i, j) = v[i * c + j];
    return m;
}

template <typename F> int guarded(F f) {
    try {
        return f();
    } catch (const std::exception& e) {
        g_err = e.what();
        return -1000;
    }
}

}  // namespace

extern "C" {

const char* ref_last_error() { return g_err.c_str(); }

// ---- ORBextractor (ORBextractor.h:44-112) -----------------------------------------------------------------------------------
void* ref_extractor_create(int nfeatures, float scale, int nlevels, int ini_th, int min_th) { return new Probe(nfeatures, scale, nlevels, ini_th, min_th); }
void ref_extractor_destroy(void* e) { delete (Probe*)e; }

void ref_extractor_tables(void* e_, float* scale, float* inv, float* s2, float* is2, int* quota, int* umax16) {
    Probe* e = (Probe*)e_;
    const int n = e->GetLevels();
    std::vector<float> a = e->GetScaleFactors(), b = e->GetInverseScaleFactors(), c = e->GetScaleSigmaSquares(), d = e->GetInverseScaleSigmaSquares();
    for (int i = 0; i < n; i++) { scale[i] = a[i]; inv[i] = b[i]; s2[i] = c[i]; is2[i] = d[i]; quota[i] = e->mnFeaturesPerLevel[i]; }
    for (int i = 0; i < 16; i++) umax16[i] = e->umax[i];
}

// operator()(image, mask, keypoints, descriptors, vLappingArea): returns monoIndex (or -1), *total = keypoints.size()
int ref_extract(void* e_, const uint8_t* img, int w, int h, int stride, int lap0, int lap1, void* kps_out, uint8_t* desc_out, int cap, int* total) {
    return guarded([&]() {
        Probe* e = (Probe*)e_;
        Mat image = (img && w > 0 && h > 0) ? Mat(h, w, CV_8UC1, (void*)img, (size_t)stride) : Mat();
        std::vector<KeyPoint> kps;
        Mat desc;
        std::vector<int> lap = {lap0, lap1};
        const int mono = (*e)(image, Mat(), kps, desc, lap);
        if (total) *total = mono < 0 ? 0 : (int)kps.size();
        if (mono < 0) return mono;
        if ((int)kps.size() > cap) { g_err = "capacity"; return -2; }
        if (!kps.empty()) std::memcpy(kps_out, kps.data(), kps.size() * sizeof(KeyPoint));
        for (int i = 0; i < desc.rows; i++) std::memcpy(desc_out + (size_t)i * 32, desc.ptr(i), 32);
        return mono;
    });
}

int ref_level_size(void* e_, int level, int* w, int* h) {
    Probe* e = (Probe*)e_;
    if (level < 0 || level >= e->GetLevels() || e->mvImagePyramid[level].empty()) return -1;
    *w = e->mvImagePyramid[level].cols;
    *h = e->mvImagePyramid[level].rows;
    return 0;
}

// mvImagePyramid[level] (ORBextractor.h:89); padded = with the 19 px frame ComputePyramid writes around it (ORBextractor.cpp:1180-1193)
int ref_get_level(void* e_, int level, int padded, uint8_t* dst) {
    Probe* e = (Probe*)e_;
    int w, h;
    if (ref_level_size(e_, level, &w, &h)) return 0;
    const Mat& m = e->mvImagePyramid[level];
    const int pad = padded ? 19 : 0;
    for (int y = -pad; y < h + pad; y++) std::memcpy(dst + (size_t)(y + pad) * (w + 2 * pad), m.data + (ptrdiff_t)y * (ptrdiff_t)m.step - pad, (size_t)(w + 2 * pad));
    return 1;
}

// DistributeOctTree (ORBextractor.cpp:554-778) on caller-supplied candidates (x, y, response triples): retained keys in list order
int ref_distribute_octtree(void* e_, const float* xyr, int n, int minX, int maxX, int minY, int maxY, int N, int level, float* out_xyr) {
    return guarded([&]() {
        Probe* e = (Probe*)e_;
        std::vector<KeyPoint> in(n);
        for (int i = 0; i < n; i++) in[i] = KeyPoint(xyr[3 * i], xyr[3 * i + 1], 7.f, -1, xyr[3 * i + 2]);
        std::vector<KeyPoint> out = e->DistributeOctTree(in, minX, maxX, minY, maxY, N, level);
        for (size_t i = 0; i < out.size(); i++) { out_xyr[3 * i] = out[i].pt.x; out_xyr[3 * i + 1] = out[i].pt.y; out_xyr[3 * i + 2] = out[i].response; }
        return (int)out.size();
    });
}

int ref_descriptor_distance(void* e_, const uint8_t* a, const uint8_t* b) {
    Mat ma(1, 32, CV_8U, (void*)a), mb(1, 32, CV_8U, (void*)b);
    return ((Probe*)e_)->DescriptorDistance(ma, mb);
}

// getMatches_distRatio (ORBextractor.cpp:1228-1290); max_neighbor_ratio is the public member (ORBextractor.h:90)
int ref_matches_dist_ratio(void* e_, const uint8_t* A, int nA_rows, const uint32_t* iA, int nA, const uint8_t* B, int nB_rows, const uint32_t* iB, int nB,
                           double max_ratio, uint32_t* match_a, uint32_t* match_b, long long* book_k) {
    return guarded([&]() {
        Probe* e = (Probe*)e_;
        std::vector<Mat> vA(nA_rows), vB(nB_rows);
        for (int i = 0; i < nA_rows; i++) vA[i] = Mat(1, 32, CV_8U, (void*)(A + (size_t)i * 32));
        for (int i = 0; i < nB_rows; i++) vB[i] = Mat(1, 32, CV_8U, (void*)(B + (size_t)i * 32));
        std::vector<unsigned int> ia(iA, iA + nA), ib(iB, iB + nB), ma, mb;
        int bk = 0;
        const double keep = e->max_neighbor_ratio;
        e->max_neighbor_ratio = max_ratio;
        e->getMatches_distRatio(vA, ia, vB, ib, ma, mb, bk);
        e->max_neighbor_ratio = keep;
        for (size_t i = 0; i < ma.size(); i++) { match_a[i] = ma[i]; match_b[i] = mb[i]; }
        if (book_k) *book_k = bk;
        return (int)ma.size();
    });
}

// ---- MultiCameraFrame / FrontEnd --------------------------------------------------------------------------------------------
void* ref_rig_create(int num_cams, const void* calib_, int w, int h, int nfeatures, float scale, int nlevels, int ini_th, int min_th, int rectify) {
    const RefCalib* cal = (const RefCalib*)calib_;
    RefRig* r = new RefRig();
    r->w = w;
    r->h = h;
    r->cfg.num_cams_ = num_cams;
    r->cfg.im_size_ = Size(w, h);
    r->cfg.RECTIFY = rectify != 0;
    for (int c = 0; c < num_cams; c++) {
        r->cfg.K_mats_.push_back(mat64(3, 3, cal[c].K));
        r->cfg.R_mats_.push_back(mat64(3, 3, cal[c].R));
        r->cfg.t_mats_.push_back(mat64(3, 1, cal[c].t));
        r->cfg.dist_coeffs_.push_back(mat64(1, 5, cal[c].dist));  // Mat_<double>::zeros(1,5) in the readers (RosDataReader.cpp:291)
        r->orbs.push_back(new ORBextractor(nfeatures, scale, nlevels, ini_th, min_th));  // one per camera, FrontEnd.h:190-193
    }
    r->orb = new ORBextractor(nfeatures, scale, nlevels, ini_th, min_th);
    r->frame = new MultiCameraFrame(r->cfg);
    r->frame->orBextractor = r->orb;
    r->frame->orBextractors = r->orbs;
    r->frame->orb_vocabulary = &r->voc;
    r->frame->imgs.assign(num_cams, Mat());
    r->frame->image_kps.assign(num_cams, {});
    r->frame->image_kps_undist.assign(num_cams, {});
    r->frame->image_descriptors.assign(num_cams, {});
    r->fe = new FrontEnd(r->cfg);
    r->fe->currentFrame = r->frame;
    r->fe->orBextractor = r->orb;
    r->fe->orb_vocabulary = &r->voc;
    for (int c = 0; c < num_cams; c++) r->fe->overlap_masks.push_back(Mat(h, w, CV_8U));
    r->desc_store.resize(num_cams);
    return r;
}
void ref_rig_destroy(void* r) { delete (RefRig*)r; }

// images: num_cams tight w*h planes (copied); then extractFeaturesParallel() (MultiCameraFrame.cpp:203-228)
int ref_rig_extract(void* r_, const uint8_t* images) {
    return guarded([&]() {
        RefRig* r = (RefRig*)r_;
        for (int c = 0; c < r->cfg.num_cams_; c++) r->frame->imgs[c] = Mat(r->h, r->w, CV_8U, (void*)(images + (size_t)c * r->w * r->h)).clone();
        r->frame->extractFeaturesParallel();
        return 0;
    });
}

// inject features for one camera (image_kps, image_descriptors), then image_kps_undist the way extractFeatureSingle does (:241-245)
int ref_rig_set_features(void* r_, int cam, const void* kps, int n, const uint8_t* desc) {
    return guarded([&]() {
        RefRig* r = (RefRig*)r_;
        MultiCameraFrame* f = r->frame;
        f->image_kps[cam].assign((const KeyPoint*)kps, (const KeyPoint*)kps + n);
        r->desc_store[cam].assign(desc, desc + (size_t)n * 32);
        f->image_descriptors[cam].clear();
        for (int i = 0; i < n; i++) f->image_descriptors[cam].push_back(Mat(1, 32, CV_8U, r->desc_store[cam].data() + (size_t)i * 32));
        f->image_kps_undist[cam].clear();
        if (r->cfg.RECTIFY) f->image_kps_undist[cam] = f->image_kps[cam];
        else f->UndistortKeyPoints(cam);
        return 0;
    });
}

int ref_rig_num_features(void* r_, int cam) { return (int)((RefRig*)r_)->frame->image_kps[cam].size(); }

int ref_rig_get_features(void* r_, int cam, void* kps_out, void* kps_undist_out, uint8_t* desc_out) {
    MultiCameraFrame* f = ((RefRig*)r_)->frame;
    const size_t n = f->image_kps[cam].size();
    if (n && kps_out) std::memcpy(kps_out, f->image_kps[cam].data(), n * sizeof(KeyPoint));
    if (kps_undist_out && f->image_kps_undist[cam].size() == n && n) std::memcpy(kps_undist_out, f->image_kps_undist[cam].data(), n * sizeof(KeyPoint));
    if (desc_out) for (size_t i = 0; i < n; i++) std::memcpy(desc_out + i * 32, f->image_descriptors[cam][i].data, 32);
    return (int)n;
}

// BruteForceMatch (MultiCameraFrame.cpp:1024-1086).  The reference indexes m[1] of every knn row, so it needs >= 2 train descriptors
// and >= 1 query descriptor (image_descriptors[..][0].cols): refused here instead of running into undefined behaviour.
int ref_rig_bf_match(void* r_, int i, int j, float dist_thresh, float ratio, uint32_t* idx1, uint32_t* idx2, void* kps1_out, void* kps2_out) {
    return guarded([&]() {
        MultiCameraFrame* f = ((RefRig*)r_)->frame;
        if (f->image_descriptors[i].empty() || f->image_descriptors[j].size() < 2) { g_err = "undefined in the reference: fewer than 2 train or no query descriptors"; return -3; }
        std::vector<unsigned int> a, b;
        std::vector<KeyPoint> k1, k2;
        f->BruteForceMatch(i, j, dist_thresh, ratio, a, b, k1, k2);
        for (size_t k = 0; k < a.size(); k++) { idx1[k] = a[k]; idx2[k] = b[k]; }
        if (kps1_out && !k1.empty()) std::memcpy(kps1_out, k1.data(), k1.size() * sizeof(KeyPoint));
        if (kps2_out && !k2.empty()) std::memcpy(kps2_out, k2.data(), k2.size() * sizeof(KeyPoint));
        return (int)a.size();
    });
}

// computeIntraMatches(matches, old) (MultiCameraFrame.cpp:1100-1288): rows_out [cap][5] (IntraMatch::matchIndex is array<int,5>)
int ref_rig_intra_matches(void* r_, int old, int32_t* rows_out, int cap) {
    return guarded([&]() {
        RefRig* r = (RefRig*)r_;
        MultiCameraFrame* f = r->frame;
        if (r->cfg.num_cams_ > 5) { g_err = "IntraMatch::matchIndex holds 5 cameras (MultiCameraFrame.h:44)"; return -4; }
        for (int c = 0; c < r->cfg.num_cams_; c++)
            if (f->image_descriptors[c].size() < 2) { g_err = "undefined in the reference: a camera with fewer than 2 descriptors"; return -3; }
        r->matches.clear();
        f->computeIntraMatches(r->matches, old != 0);
        if ((int)r->matches.size() > cap) { g_err = "capacity"; return -2; }
        for (size_t m = 0; m < r->matches.size(); m++)
            for (int c = 0; c < 5; c++) rows_out[m * 5 + c] = r->matches[m].matchIndex[c];
        return (int)r->matches.size();
    });
}

// replace the stored IntraMatch rows (so that obtainLfFeatures can be checked on rows that did not come from computeIntraMatches)
int ref_rig_set_rows(void* r_, const int32_t* rows, int M) {
    RefRig* r = (RefRig*)r_;
    r->matches.assign(M, IntraMatch());
    for (int m = 0; m < M; m++)
        for (int c = 0; c < 5; c++) r->matches[m].matchIndex[c] = c < r->cfg.num_cams_ ? rows[(size_t)m * 5 + c] : -1;
    return M;
}

// FrontEnd::obtainLfFeatures (FrontEnd.cpp:213-593) on the stored rows.  masks: num_cams float planes h*w, or null = all zero
// (mc_slam_app.cpp:223-224).  Outputs describe currentFrame->intraMatches in order: matchIndex [n][5], uv_ref [n][2], mono [n],
// n_rays [n], xyz [n][3] (NaN when the row carries no point3D), desc [n][32].  Returns n; *intramatch_size / *mono_size as the
// reference counts them.
int ref_rig_obtain_lf_features(void* r_, const float* const* masks, int32_t* match_index, float* uv_ref, uint8_t* mono, int32_t* n_rays,
                               double* xyz, uint8_t* desc, int cap, int* intramatch_size, int* mono_size) {
    return guarded([&]() {
        RefRig* r = (RefRig*)r_;
        MultiCameraFrame* f = r->frame;
        f->segMasks.clear();
        for (int c = 0; c < r->cfg.num_cams_; c++) {
            Mat m(r->h, r->w, CV_32F);
            if (masks && masks[c]) std::memcpy(m.data, masks[c], (size_t)r->w * r->h * 4);
            f->segMasks.push_back(m);
        }
        std::vector<IntraMatch> filtered;
        std::vector<DBoW2::NodeId> words(r->matches.size(), 0);
        std::set<DBoW2::NodeId> fwords;
        r->fe->obtainLfFeatures(r->matches, f, filtered, words, fwords);
        const std::vector<IntraMatch>& out = f->intraMatches;
        if ((int)out.size() > cap) { g_err = "capacity"; return -2; }
        for (size_t i = 0; i < out.size(); i++) {
            for (int c = 0; c < 5; c++) match_index[i * 5 + c] = out[i].matchIndex[c];
            uv_ref[i * 2] = out[i].uv_ref.x;
            uv_ref[i * 2 + 1] = out[i].uv_ref.y;
            mono[i] = out[i].mono ? 1 : 0;
            n_rays[i] = out[i].n_rays;
            for (int k = 0; k < 3; k++) xyz[i * 3 + k] = out[i].point3D.empty() ? NAN : out[i].point3D.at<double>(k, 0);
            if (!out[i].matchDesc.empty()) std::memcpy(desc + i * 32, out[i].matchDesc.data, 32);
            else std::memset(desc + i * 32, 0, 32);
        }
        if (intramatch_size) *intramatch_size = f->intramatch_size;
        if (mono_size) *mono_size = f->mono_size;
        return (int)out.size();
    });
}

// triangulateIntraMatches(x, view_inds, pt) (MultiCameraFrame.cpp:411-441)
int ref_rig_triangulate_nview(void* r_, const int* view_inds, const double* xs, const double* ys, int nviews, double* X3) {
    return guarded([&]() {
        MultiCameraFrame* f = ((RefRig*)r_)->frame;
        Mat_<double> x(2, nviews);
        std::vector<int> vi(view_inds, view_inds + nviews);
        for (int i = 0; i < nviews; i++) { x(0, i) = xs[i]; x(1, i) = ys[i]; }
        Eigen::Vector3d pt;
        f->triangulateIntraMatches(x, vi, pt);
        for (int k = 0; k < 3; k++) X3[k] = pt[k];
        return 0;
    });
}

// triangulate(kps_1, kps_2, pts3D, cam1, cam2) (MultiCameraFrame.cpp:480-507)
int ref_rig_triangulate_pair(void* r_, int cam1, int cam2, const float* p1, const float* p2, int n, float* out_xyz) {
    return guarded([&]() {
        MultiCameraFrame* f = ((RefRig*)r_)->frame;
        std::vector<Point2f> a(n), b(n);
        for (int i = 0; i < n; i++) { a[i] = Point2f(p1[2 * i], p1[2 * i + 1]); b[i] = Point2f(p2[2 * i], p2[2 * i + 1]); }
        std::vector<Point3f> pts;
        f->triangulate(a, b, pts, cam1, cam2);
        for (int i = 0; i < n; i++) { out_xyz[3 * i] = pts[i].x; out_xyz[3 * i + 1] = pts[i].y; out_xyz[3 * i + 2] = pts[i].z; }
        return 0;
    });
}

// computeRepresentativeDesc (MultiCameraFrame.cpp:530-567)
int ref_rig_representative_desc(void* r_, const uint8_t* descs, int n, uint8_t* out32) {
    return guarded([&]() {
        MultiCameraFrame* f = ((RefRig*)r_)->frame;
        std::vector<Mat> v;
        for (int i = 0; i < n; i++) v.push_back(Mat(1, 32, CV_8U, (void*)(descs + (size_t)i * 32)));
        Mat out;
        f->computeRepresentativeDesc(v, out);
        std::memcpy(out32, out.data, 32);
        return 0;
    });
}

// One rig frame through the reference's own front-end calls, as FrontEnd::processFrame strings them together for the brute-force
// configuration (extractFeaturesParallel -> computeIntraMatches(matches, false) -> obtainLfFeatures): the timed CPU baseline.
int ref_rig_process(void* r_, const uint8_t* images, int do_match, int do_triangulate, int* kp_counts, int* accepted) {
    return guarded([&]() {
        RefRig* r = (RefRig*)r_;
        MultiCameraFrame* f = r->frame;
        for (int c = 0; c < r->cfg.num_cams_; c++) r->frame->imgs[c] = Mat(r->h, r->w, CV_8U, (void*)(images + (size_t)c * r->w * r->h));
        f->extractFeaturesParallel();
        if (kp_counts) for (int c = 0; c < r->cfg.num_cams_; c++) kp_counts[c] = (int)f->image_kps[c].size();
        if (!do_match) return 0;
        for (int c = 0; c < r->cfg.num_cams_; c++)
            if (f->image_descriptors[c].size() < 2) { g_err = "undefined in the reference: a camera with fewer than 2 descriptors"; return -3; }
        r->matches.clear();
        f->computeIntraMatches(r->matches, false);
        if (!do_triangulate) return (int)r->matches.size();
        if (f->segMasks.size() != (size_t)r->cfg.num_cams_) {
            f->segMasks.clear();
            for (int c = 0; c < r->cfg.num_cams_; c++) f->segMasks.push_back(Mat(r->h, r->w, CV_32F));
        }
        std::vector<IntraMatch> filtered;
        std::vector<DBoW2::NodeId> words(r->matches.size(), 0);
        std::set<DBoW2::NodeId> fwords;
        r->fe->obtainLfFeatures(r->matches, f, filtered, words, fwords);
        if (accepted) *accepted = f->intramatch_size;
        return (int)r->matches.size();
    });
}

}  // extern "C"
