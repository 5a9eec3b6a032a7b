// mini_cv.cpp — out-of-line part of the OpenCV stand-in (see mini_cv.hpp).  TEST INFRASTRUCTURE ONLY.
// The pixel / descriptor primitives forward to liboracle.so's cv2-pinned restatements; nothing here is product code.
#include "mini_cv.hpp"

#include <limits>

extern "C" {  // oracle/mcs_oracle.cpp (liboracle.so)
void orc_resize_linear_u8(const uint8_t* src, int sw, int sh, int sstride, uint8_t* dst, int dw, int dh, int dstride);
void orc_gauss7_u8(const uint8_t* src, int w, int h, int sstride, uint8_t* dst, int dstride);
int orc_fast9(const uint8_t* img, int w, int h, int stride, int threshold, int32_t* out_xys, int cap);
float orc_fast_atan2(float y, float x);
void orc_undistort_points(const float* xy_in, int n, const double* K, const double* dist5, float* xy_out);
void orc_knn2(const uint8_t* q, int nq, const uint8_t* t, int nt, int* i0, int* d0, int* i1, int* d1);
}

namespace cv {

float fastAtan2(float y, float x) { return orc_fast_atan2(y, x); }

std::ostream& operator<<(std::ostream& o, const Mat& m) {
    o << "[";
    for (int i = 0; i < m.rows; i++) {
        for (int j = 0; j < m.cols; j++) o << (j ? ", " : "") << m.getd(i, j);
        o << (i + 1 < m.rows ? ";\n " : "");
    }
    return o << "]";
}

// cv::FAST(img, kps, threshold, true): TYPE_9_16, raster order, KeyPoint(x, y, 7.f, -1, score)
void FAST(InputArray image, std::vector<KeyPoint>& keypoints, int threshold, bool nonmaxSuppression) {
    CV_Assert(nonmaxSuppression);
    Mat img = image.getMat();
    CV_Assert(img.type() == CV_8U);
    keypoints.clear();
    if (img.rows < 7 || img.cols < 7) return;
    const int cap = (img.rows * img.cols) / 4 + 16;
    std::vector<int32_t> out((size_t)cap * 3);
    const int n = orc_fast9(img.data, img.cols, img.rows, (int)img.step, threshold, out.data(), cap);
    keypoints.reserve(n);
    for (int i = 0; i < n; i++) keypoints.push_back(KeyPoint((float)out[3 * i], (float)out[3 * i + 1], 7.f, -1, (float)out[3 * i + 2]));
}

void resize(InputArray src_, OutputArray dst_, Size dsize, double fx, double fy, int interpolation) {
    Mat src = src_.getMat();
    CV_Assert(src.type() == CV_8U && interpolation == INTER_LINEAR && fx == 0 && fy == 0);
    dst_.create(dsize.height, dsize.width, CV_8U);
    Mat dst = dst_.getMat();
    orc_resize_linear_u8(src.data, src.cols, src.rows, (int)src.step, dst.data, dst.cols, dst.rows, (int)dst.step);
}

static int border101(int p, int len) {
    if (len == 1) return 0;
    while (p < 0 || p >= len) p = p < 0 ? -p : 2 * len - 2 - p;
    return p;
}

void copyMakeBorder(InputArray src_, OutputArray dst_, int top, int bottom, int left, int right, int borderType) {
    Mat src = src_.getMat();
    CV_Assert(src.type() == CV_8U && (borderType & ~BORDER_ISOLATED) == BORDER_REFLECT_101);
    const int w = src.cols, h = src.rows;
    dst_.create(h + top + bottom, w + left + right, CV_8U);
    Mat dst = dst_.getMat();
    // src may be the interior of dst (ORBextractor.cpp:1188): interior rows first (memmove), then borders from the interior
    for (int y = 0; y < h; y++) {
        uchar* d = dst.ptr(y + top) + left;
        if (d != src.ptr(y)) std::memmove(d, src.ptr(y), (size_t)w);
    }
    for (int y = 0; y < h; y++) {
        uchar* row = dst.ptr(y + top);
        for (int x = 0; x < left; x++) row[x] = row[left + border101(x - left, w)];
        for (int x = 0; x < right; x++) row[left + w + x] = row[left + border101(w + x, w)];
    }
    for (int y = 0; y < top; y++) std::memcpy(dst.ptr(y), dst.ptr(top + border101(y - top, h)), (size_t)dst.cols);
    for (int y = 0; y < bottom; y++) std::memcpy(dst.ptr(top + h + y), dst.ptr(top + border101(h + y, h)), (size_t)dst.cols);
}

void GaussianBlur(InputArray src_, OutputArray dst_, Size ksize, double sigmaX, double sigmaY, int borderType) {
    Mat src = src_.getMat().clone();  // in-place call at ORBextractor.cpp:1133
    CV_Assert(src.type() == CV_8U && ksize.width == 7 && ksize.height == 7 && sigmaX == 2 && sigmaY == 2 && borderType == BORDER_REFLECT_101);
    dst_.create(src.rows, src.cols, CV_8U);
    Mat dst = dst_.getMat();
    orc_gauss7_u8(src.data, src.cols, src.rows, (int)src.step, dst.data, (int)dst.step);
}

void undistortPoints(InputArray src_, OutputArray dst_, InputArray K_, InputArray dist_, InputArray R_, InputArray P_) {
    Mat src = src_.getMat(), K = K_.getMat(), dist = dist_.getMat(), P = P_.getMat();
    CV_Assert(src.type() == CV_32F && src.channels() == 2 && src.cols == 1 && R_.empty());
    CV_Assert(K.rows == 3 && K.cols == 3 && P.rows == 3 && P.cols == 3);
    double Kd[9], k[5] = {0, 0, 0, 0, 0};
    for (int i = 0; i < 9; i++) {
        Kd[i] = K.getd(i / 3, i % 3);
        CV_Assert(P.getd(i / 3, i % 3) == Kd[i]);  // the only call site re-projects through the same matrix
    }
    const int nd = (int)dist.total();
    CV_Assert(nd == 4 || nd == 5);
    for (int i = 0; i < nd; i++) k[i] = dist.rows == 1 ? dist.getd(0, i) : dist.getd(i, 0);
    const int n = src.rows;
    std::vector<float> in((size_t)n * 2), out((size_t)n * 2);
    for (int i = 0; i < n; i++) { in[2 * i] = src.ptr<float>(i)[0]; in[2 * i + 1] = src.ptr<float>(i)[1]; }
    orc_undistort_points(in.data(), n, Kd, k, out.data());
    Mat dst = dst_.getMat();
    CV_Assert(dst.rows == n && dst.channels() == 2 && dst.type() == CV_32F);
    for (int i = 0; i < n; i++) { dst.ptr<float>(i)[0] = out[2 * i]; dst.ptr<float>(i)[1] = out[2 * i + 1]; }
}

void KeyPointsFilter::retainBest(std::vector<KeyPoint>& keypoints, int npoints) {
    if (npoints >= 0 && (int)keypoints.size() > npoints) {
        std::stable_sort(keypoints.begin(), keypoints.end(), [](const KeyPoint& a, const KeyPoint& b) { return a.response > b.response; });
        keypoints.resize(npoints);
    }
}

Ptr<DescriptorMatcher> DescriptorMatcher::create(const std::string& name) {
    CV_Assert(name == "BruteForce-Hamming");
    return std::make_shared<DescriptorMatcher>();
}

void DescriptorMatcher::knnMatch(InputArray query_, InputArray train_, std::vector<std::vector<DMatch>>& matches, int k) {
    Mat q = query_.getMat(), t = train_.getMat();
    CV_Assert(k == 2 && q.type() == CV_8U && t.type() == CV_8U && q.cols == 32 && t.cols == 32 && q.isContinuous() && t.isContinuous());
    const int nq = q.rows, nt = t.rows;
    std::vector<int> i0(nq + 1), d0(nq + 1), i1(nq + 1), d1(nq + 1);
    orc_knn2(q.data, nq, t.data, nt, i0.data(), d0.data(), i1.data(), d1.data());
    matches.assign(nq, {});
    for (int i = 0; i < nq; i++) {
        if (i0[i] >= 0) matches[i].push_back(DMatch(i, i0[i], (float)d0[i]));
        if (i1[i] >= 0) matches[i].push_back(DMatch(i, i1[i], (float)d1[i]));
    }
}

// ---------------------------------------------------------------------------------------------------------------------
MatExpr shim_invert(const Mat& a) {
    CV_Assert(a.rows == a.cols && (a.type() == CV_32F || a.type() == CV_64F));
    const int n = a.rows;
    Mat r(n, n, a.type());
    if (n == 2) {
        double d = a.getd(0, 0) * a.getd(1, 1) - a.getd(0, 1) * a.getd(1, 0);
        if (d != 0.) {
            d = 1. / d;
            r.setd(0, 0, a.getd(1, 1) * d);
            r.setd(0, 1, -a.getd(0, 1) * d);
            r.setd(1, 0, -a.getd(1, 0) * d);
            r.setd(1, 1, a.getd(0, 0) * d);
        }
        return MatExpr(r);
    }
    if (n == 3) {
        auto A = [&](int i, int j) { return a.getd(i, j); };
        double d = A(0, 0) * (A(1, 1) * A(2, 2) - A(1, 2) * A(2, 1)) - A(0, 1) * (A(1, 0) * A(2, 2) - A(1, 2) * A(2, 0)) +
                   A(0, 2) * (A(1, 0) * A(2, 1) - A(1, 1) * A(2, 0));
        if (d != 0.) {
            d = 1. / d;
            double t[9];
            t[0] = (A(1, 1) * A(2, 2) - A(1, 2) * A(2, 1)) * d;
            t[1] = (A(0, 2) * A(2, 1) - A(0, 1) * A(2, 2)) * d;
            t[2] = (A(0, 1) * A(1, 2) - A(0, 2) * A(1, 1)) * d;
            t[3] = (A(1, 2) * A(2, 0) - A(1, 0) * A(2, 2)) * d;
            t[4] = (A(0, 0) * A(2, 2) - A(0, 2) * A(2, 0)) * d;
            t[5] = (A(0, 2) * A(1, 0) - A(0, 0) * A(1, 2)) * d;
            t[6] = (A(1, 0) * A(2, 1) - A(1, 1) * A(2, 0)) * d;
            t[7] = (A(0, 1) * A(2, 0) - A(0, 0) * A(2, 1)) * d;
            t[8] = (A(0, 0) * A(1, 1) - A(0, 1) * A(1, 0)) * d;
            for (int i = 0; i < 9; i++) r.setd(i / 3, i % 3, t[i]);
        }
        return MatExpr(r);
    }
    std::vector<double> M((size_t)n * 2 * n, 0.0);
    for (int i = 0; i < n; i++) {
        for (int j = 0; j < n; j++) M[(size_t)i * 2 * n + j] = a.getd(i, j);
        M[(size_t)i * 2 * n + n + i] = 1.0;
    }
    for (int c = 0; c < n; c++) {
        int piv = c;
        for (int i = c + 1; i < n; i++)
            if (std::abs(M[(size_t)i * 2 * n + c]) > std::abs(M[(size_t)piv * 2 * n + c])) piv = i;
        if (std::abs(M[(size_t)piv * 2 * n + c]) < std::numeric_limits<double>::epsilon()) return MatExpr(Mat(n, n, a.type()));
        if (piv != c)
            for (int j = 0; j < 2 * n; j++) std::swap(M[(size_t)piv * 2 * n + j], M[(size_t)c * 2 * n + j]);
        const double d = 1.0 / M[(size_t)c * 2 * n + c];
        for (int j = 0; j < 2 * n; j++) M[(size_t)c * 2 * n + j] *= d;
        for (int i = 0; i < n; i++) {
            if (i == c) continue;
            const double f = M[(size_t)i * 2 * n + c];
            if (f == 0) continue;
            for (int j = 0; j < 2 * n; j++) M[(size_t)i * 2 * n + j] -= f * M[(size_t)c * 2 * n + j];
        }
    }
    for (int i = 0; i < n; i++)
        for (int j = 0; j < n; j++) r.setd(i, j, M[(size_t)i * 2 * n + n + j]);
    return MatExpr(r);
}

// One-sided (Hestenes) Jacobi on the columns of A (m x n), V accumulated; singular values = column norms, sorted descending.
template <typename T> static void jacobi_svd(std::vector<T>& A, int m, int n, std::vector<T>& W, std::vector<T>& V) {
    V.assign((size_t)n * n, 0);
    for (int i = 0; i < n; i++) V[(size_t)i * n + i] = 1;
    const T eps = std::numeric_limits<T>::epsilon();
    for (int sweep = 0; sweep < 60; sweep++) {
        bool changed = false;
        for (int p = 0; p < n - 1; p++)
            for (int q = p + 1; q < n; q++) {
                T a = 0, b = 0, g = 0;
                for (int k = 0; k < m; k++) {
                    a += A[(size_t)k * n + p] * A[(size_t)k * n + p];
                    b += A[(size_t)k * n + q] * A[(size_t)k * n + q];
                    g += A[(size_t)k * n + p] * A[(size_t)k * n + q];
                }
                if (g == 0 || std::abs(g) <= eps * std::sqrt(a * b)) continue;
                changed = true;
                const T zeta = (b - a) / (2 * g);
                const T t = (zeta >= 0 ? 1 : -1) / (std::abs(zeta) + std::sqrt(1 + zeta * zeta));
                const T c = 1 / std::sqrt(1 + t * t), s = c * t;
                for (int k = 0; k < m; k++) {
                    const T x = A[(size_t)k * n + p], y = A[(size_t)k * n + q];
                    A[(size_t)k * n + p] = c * x - s * y;
                    A[(size_t)k * n + q] = s * x + c * y;
                }
                for (int k = 0; k < n; k++) {
                    const T x = V[(size_t)k * n + p], y = V[(size_t)k * n + q];
                    V[(size_t)k * n + p] = c * x - s * y;
                    V[(size_t)k * n + q] = s * x + c * y;
                }
            }
        if (!changed) break;
    }
    W.assign(n, 0);
    for (int j = 0; j < n; j++) {
        T s = 0;
        for (int k = 0; k < m; k++) s += A[(size_t)k * n + j] * A[(size_t)k * n + j];
        W[j] = std::sqrt(s);
    }
}

template <typename T> static void svd_impl(const Mat& A_, Mat* w, Mat* u, Mat* vt, int type) {
    const int m = A_.rows, n = A_.cols;
    CV_Assert(m >= n);
    std::vector<T> A((size_t)m * n), W, V;
    for (int i = 0; i < m; i++) for (int j = 0; j < n; j++) A[(size_t)i * n + j] = (T)A_.getd(i, j);
    jacobi_svd<T>(A, m, n, W, V);
    std::vector<int> order(n);
    for (int i = 0; i < n; i++) order[i] = i;
    std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return W[a] > W[b]; });
    if (w) { w->create(n, 1, type); for (int i = 0; i < n; i++) w->setd(i, 0, W[order[i]]); }
    if (vt) {
        vt->create(n, n, type);
        for (int i = 0; i < n; i++) for (int k = 0; k < n; k++) vt->setd(i, k, V[(size_t)k * n + order[i]]);
    }
    if (u) {  // thin U (m x n); FULL_UV is only ever requested for square matrices here
        u->create(m, n, type);
        for (int i = 0; i < n; i++) {
            const T s = W[order[i]];
            for (int k = 0; k < m; k++) u->setd(k, i, s > 0 ? A[(size_t)k * n + order[i]] / s : 0);
        }
    }
}

void SVD::compute(InputArray A_, OutputArray w_, OutputArray u_, OutputArray vt_, int) {
    Mat A = A_.getMat(), w, u, vt;
    CV_Assert(A.type() == CV_32F || A.type() == CV_64F);
    if (A.type() == CV_32F) svd_impl<float>(A, &w, &u, &vt, CV_32F);
    else svd_impl<double>(A, &w, &u, &vt, CV_64F);
    if (w_.needed()) w.copyTo(w_);
    if (u_.needed()) u.copyTo(u_);
    if (vt_.needed()) vt.copyTo(vt_);
}

void SVD::solveZ(InputArray A_, OutputArray z_) {
    Mat A = A_.getMat(), vt;
    CV_Assert(A.type() == CV_32F || A.type() == CV_64F);
    if (A.type() == CV_32F) svd_impl<float>(A, nullptr, nullptr, &vt, CV_32F);
    else svd_impl<double>(A, nullptr, nullptr, &vt, CV_64F);
    Mat(vt.row(vt.rows - 1).t()).copyTo(z_);
}

namespace sfm {

void homogeneousToEuclidean(InputArray src_, OutputArray dst_) {
    Mat X = src_.getMat();
    const int d = X.rows, n = X.cols;
    Mat x(d - 1, n, X.type());
    for (int i = 0; i < n; i++) {
        const double h = X.getd(d - 1, i);
        for (int r = 0; r < d - 1; r++) x.setd(r, i, X.getd(r, i) / h);
    }
    x.copyTo(dst_);
}

void homogeneousToEuclidean(InputArray src_, Vec3d& dst) {
    Mat x;
    homogeneousToEuclidean(src_, x);
    CV_Assert(x.rows == 3 && x.cols == 1);
    for (int i = 0; i < 3; i++) dst[i] = x.getd(i, 0);
}

static void triangulate_dlt(const double* xl, const double* xr, const Mat& Pl, const Mat& Pr, double* X3) {
    Mat_<double> design(4, 4);
    for (int i = 0; i < 4; i++) {
        design(0, i) = xl[0] * Pl.at<double>(2, i) - Pl.at<double>(0, i);
        design(1, i) = xl[1] * Pl.at<double>(2, i) - Pl.at<double>(1, i);
        design(2, i) = xr[0] * Pr.at<double>(2, i) - Pr.at<double>(0, i);
        design(3, i) = xr[1] * Pr.at<double>(2, i) - Pr.at<double>(1, i);
    }
    Mat XH;
    SVD::solveZ(design, XH);
    Vec3d X;
    homogeneousToEuclidean(XH, X);
    for (int i = 0; i < 3; i++) X3[i] = X[i];
}

static void triangulate_nviews(const Mat_<double>& x, const std::vector<Mat>& Ps, double* X3) {
    const int nviews = x.cols;
    Mat_<double> design = Mat_<double>::zeros(3 * nviews, 4 + nviews);
    for (int i = 0; i < nviews; i++) {
        for (int jj = 0; jj < 3; jj++)
            for (int ii = 0; ii < 4; ii++) design(3 * i + jj, ii) = -Ps[i].at<double>(jj, ii);
        design(3 * i + 0, 4 + i) = x(0, i);
        design(3 * i + 1, 4 + i) = x(1, i);
        design(3 * i + 2, 4 + i) = 1.0;
    }
    Mat X_and_alphas;
    SVD::solveZ(design, X_and_alphas);
    Vec3d X;
    homogeneousToEuclidean(X_and_alphas.rowRange(0, 4), X);
    for (int i = 0; i < 3; i++) X3[i] = X[i];
}

void triangulatePoints(const std::vector<Mat_<double>>& points2d, const std::vector<Mat>& Ps, OutputArray points3d) {
    const int nviews = (int)points2d.size();
    CV_Assert(nviews >= 2 && nviews == (int)Ps.size());
    const int npts = points2d[0].cols;
    Mat out(3, npts, CV_64F);
    for (int p = 0; p < npts; p++) {
        double X[3];
        if (nviews == 2) {
            const double xl[2] = {points2d[0](0, p), points2d[0](1, p)}, xr[2] = {points2d[1](0, p), points2d[1](1, p)};
            triangulate_dlt(xl, xr, Ps[0], Ps[1], X);
        } else {
            Mat_<double> x(2, nviews);
            for (int k = 0; k < nviews; k++) { x(0, k) = points2d[k](0, p); x(1, k) = points2d[k](1, p); }
            triangulate_nviews(x, Ps, X);
        }
        for (int i = 0; i < 3; i++) out.at<double>(i, p) = X[i];
    }
    out.copyTo(points3d);
}

}  // namespace sfm
}  // namespace cv
