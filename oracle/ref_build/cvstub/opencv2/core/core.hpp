// ORACLE reference build only (test infrastructure, never seen by the product): what the reference's
// src/{ORBmatcher,Frame,KeyFrame,MapPoint,Map,KeyFrameDatabase}.cc and Thirdparty/DBoW2 use of OpenCV beyond the
// type shim in include/minicv, so that those files compile UNMODIFIED where they lie:
//   * the CV_32F matrix algebra (gemm through MatExpr, transpose, scale, +/-, dot, norm) with the arithmetic of
//     OpenCV 4.13 written out (pinned against cv2 4.13 by tests/test_cvprim_golden.py::test_mat_algebra_golden):
//       A*B (+C), no transposition, inner length 2..4  -> gemm's small-matrix branch: float products and sums left to
//                                                         right, then (float)((double)t*alpha + (double)c*beta)
//       anything with a transposed operand              -> GEMMSingleMul<float,double>: double accumulation in k
//                                                         order, (float)(s*alpha + c*beta)
//       a.dot(b), norm(a) on CV_32F                     -> double accumulators
//       alpha*A, A/alpha materialised                   -> float product with (float)alpha
//   * cv::FileStorage / cv::FileNode names (DBoW2's unused yaml members), cv::undistortPoints (never reached: the
//     wrappers use undistorted key points): abort.
#ifndef ORACLE_CVSTUB_CORE_HPP
#define ORACLE_CVSTUB_CORE_HPP
#include <climits>
#include <cmath>
#include <cstdlib>
#include <iostream>
#include <sstream>
#include <string>

namespace cv { struct MatT; struct MatS; class Mat; }
#define MINICV_MAT_EXTRA_MEMBERS                                                             \
    MatT t() const;                                                                          \
    double dot(const Mat& m) const;                                                          \
    Mat reshape(int cn, int rows = 0) const;                                                 \
    void convertTo(Mat& m, int rtype) const;                                                 \
    static Mat ones(int r, int c, int t);                                                    \
    static Mat eye(int r, int c, int t);                                                     \
    void copyTo(Mat&& dst) const { Mat d(dst); copyTo(d); }                                  \
    Mat(const MatT& e);                                                                      \
    Mat& operator=(const MatT& e);                                                           \
    Mat(const MatS& e);                                                                      \
    Mat& operator=(const MatS& e);
#include_next <opencv2/core/core.hpp>

namespace cv {

enum { NORM_L1 = 2, NORM_L2 = 4 };

// Rcw.t() with a pending scale: MatExpr op 'T' (alpha * A^T)
struct MatT { Mat a; double alpha; };

inline Mat materialize(const MatT& e) {
    Mat r(e.a.cols, e.a.rows, CV_32F);
    for (int i = 0; i < r.rows; i++)
        for (int j = 0; j < r.cols; j++) {
            const float x = e.a.at<float>(j, i);
            r.at<float>(i, j) = e.alpha == 1.0 ? x : x * (float)e.alpha;
        }
    return r;
}
inline Mat::Mat(const MatT& e) : Mat() { *this = materialize(e); }
inline Mat& Mat::operator=(const MatT& e) { Mat r = materialize(e); *this = r; return *this; }
inline MatT Mat::t() const { return MatT{*this, 1.0}; }
inline MatT operator-(const MatT& e) { return MatT{e.a, -e.alpha}; }
inline MatT operator*(double s, const MatT& e) { return MatT{e.a, e.alpha * s}; }

// gemm(A, B, alpha, C, beta) for CV_32F, A not transposed
inline Mat gemm_nn(const Mat& A, const Mat& B, double alpha, const Mat* C, double beta) {
    const int len = A.cols, H = A.rows, W = B.cols;
    Mat D(H, W, CV_32F);
    if (len >= 2 && len <= 4 && (len == W || len == H)) {
        for (int i = 0; i < H; i++)
            for (int j = 0; j < W; j++) {
                float t = A.at<float>(i, 0) * B.at<float>(0, j);
                for (int k = 1; k < len; k++) t = t + A.at<float>(i, k) * B.at<float>(k, j);
                const float c = C ? C->at<float>(i, j) : 0.0f;
                D.at<float>(i, j) = (float)((double)t * alpha + (double)c * beta);
            }
    } else {
        for (int i = 0; i < H; i++)
            for (int j = 0; j < W; j++) {
                double s = 0;
                for (int k = 0; k < len; k++) s += (double)A.at<float>(i, k) * (double)B.at<float>(k, j);
                D.at<float>(i, j) = C ? (float)(s * alpha + (double)C->at<float>(i, j) * beta) : (float)(s * alpha);
            }
    }
    return D;
}
// gemm(A, B, alpha, C, beta, GEMM_1_T): A given untransposed
inline Mat gemm_tn(const Mat& A, const Mat& B, double alpha, const Mat* C, double beta) {
    const int len = A.rows, H = A.cols, W = B.cols;
    Mat D(H, W, CV_32F);
    for (int i = 0; i < H; i++)
        for (int j = 0; j < W; j++) {
            double s = 0;
            for (int k = 0; k < len; k++) s += (double)A.at<float>(k, i) * (double)B.at<float>(k, j);
            D.at<float>(i, j) = C ? (float)(s * alpha + (double)C->at<float>(i, j) * beta) : (float)(s * alpha);
        }
    return D;
}

// A*B kept lazy so that "A*B + C" becomes ONE gemm with beta = 1 (MatExpr op 'G')
struct MatMul {
    Mat a, b; double alpha; bool ta;
    Mat eval(const Mat* C = nullptr, double beta = 0) const { return ta ? gemm_tn(a, b, alpha, C, beta) : gemm_nn(a, b, alpha, C, beta); }
    operator Mat() const { return eval(); }
};
// -A and s*A kept lazy so that "(-A)*B" folds the scale into gemm's alpha (MatExpr op 'S')
struct MatS { Mat a; double alpha; };
inline Mat materialize(const MatS& e) {
    Mat r(e.a.rows, e.a.cols, CV_32F);
    for (int i = 0; i < r.rows; i++)
        for (int j = 0; j < r.cols; j++) r.at<float>(i, j) = e.a.at<float>(i, j) * (float)e.alpha;
    return r;
}

inline Mat::Mat(const MatS& e) : Mat() { *this = materialize(e); }
inline Mat& Mat::operator=(const MatS& e) { Mat r = materialize(e); *this = r; return *this; }
inline MatMul operator*(const Mat& a, const Mat& b) { return MatMul{a, b, 1.0, false}; }
inline MatMul operator*(const MatT& a, const Mat& b) { return MatMul{a.a, b, a.alpha, true}; }
inline MatMul operator*(const MatS& a, const Mat& b) { return MatMul{a.a, b, a.alpha, false}; }
inline Mat operator+(const MatMul& m, const Mat& c) { return m.eval(&c, 1.0); }
inline Mat operator+(const Mat& c, const MatMul& m) { return m.eval(&c, 1.0); }
inline MatS operator-(const Mat& a) { return MatS{a, -1.0}; }
inline MatS operator*(double s, const Mat& a) { return MatS{a, s}; }
inline MatS operator*(const Mat& a, double s) { return MatS{a, s}; }
inline MatS operator/(const Mat& a, double s) { return MatS{a, 1.0 / s}; }

inline Mat elementwise(const Mat& a, const Mat& b, float sb) {
    Mat r(a.rows, a.cols, CV_32F);
    for (int i = 0; i < a.rows; i++)
        for (int j = 0; j < a.cols; j++) r.at<float>(i, j) = a.at<float>(i, j) + sb * b.at<float>(i, j);
    return r;
}
inline Mat operator+(const Mat& a, const Mat& b) { return elementwise(a, b, 1.0f); }
inline Mat operator-(const Mat& a, const Mat& b) { return elementwise(a, b, -1.0f); }
inline Mat operator+(const Mat& a, const MatS& b) { return elementwise(a, materialize(b), 1.0f); }      // a + b/s: addWeighted path
inline Mat operator-(const Mat& a, const MatS& b) { return elementwise(a, materialize(b), -1.0f); }

inline double norm(const Mat& a) {                      // NORM_L2 on CV_32F
    double s = 0;
    for (int i = 0; i < a.rows; i++)
        for (int j = 0; j < a.cols; j++) s += (double)a.at<float>(i, j) * (double)a.at<float>(i, j);
    return std::sqrt(s);
}
inline double norm(const Mat& a, const Mat& b, int type) {  // NORM_L1 of the difference, CV_32F
    (void)type;
    double s = 0;
    for (int i = 0; i < a.rows; i++)
        for (int j = 0; j < a.cols; j++) s += std::fabs((double)a.at<float>(i, j) - (double)b.at<float>(i, j));
    return s;
}
inline double Mat::dot(const Mat& m) const {
    double s = 0;
    for (int i = 0; i < rows; i++)
        for (int j = 0; j < cols; j++) s += (double)at<float>(i, j) * (double)m.at<float>(i, j);
    return s;
}
inline Mat Mat::reshape(int cn, int r) const { (void)cn; (void)r; std::abort(); }
inline void Mat::convertTo(Mat& m, int rtype) const {
    if (rtype != CV_32F || depth() != CV_8U) std::abort();
    Mat r(rows, cols, CV_32F);
    for (int i = 0; i < rows; i++)
        for (int j = 0; j < cols; j++) r.at<float>(i, j) = (float)at<uchar>(i, j);
    m = r;
}
inline Mat Mat::ones(int r, int c, int t) {
    if (t != CV_32F) std::abort();
    Mat m(r, c, t);
    for (int i = 0; i < r; i++) for (int j = 0; j < c; j++) m.at<float>(i, j) = 1.0f;
    return m;
}
inline Mat Mat::eye(int r, int c, int t) {
    if (t != CV_32F) std::abort();
    Mat m(r, c, t);
    for (int i = 0; i < r; i++) for (int j = 0; j < c; j++) m.at<float>(i, j) = i == j ? 1.0f : 0.0f;
    return m;
}
template <typename A, typename B, typename C, typename D, typename E, typename F>
inline void undistortPoints(const A&, const B&, const C&, const D&, const E&, const F&) { std::abort(); }

// (cv::Mat_<float>(3,1) << x, y, z)
template <typename T> struct Mat_ : Mat { Mat_(int r, int c) : Mat(r, c, CV_32F) { static_assert(sizeof(T) == 4, "float only"); } };
template <typename T> struct MatCommaInit {
    Mat m; int idx;
    template <typename V> MatCommaInit& operator,(V v) { ((T*)m.data)[idx++] = (T)v; return *this; }
    operator Mat() const { return m; }
};
template <typename T, typename V> inline MatCommaInit<T> operator<<(const Mat_<T>& m, V v) {
    MatCommaInit<T> c{m, 0};
    return (c, v);
}

struct FileNode {
    FileNode operator[](const char*) const { std::abort(); }
    FileNode operator[](const std::string&) const { std::abort(); }
    FileNode operator[](int) const { std::abort(); }
    size_t size() const { std::abort(); }
    operator int() const { std::abort(); }
    operator double() const { std::abort(); }
    operator std::string() const { std::abort(); }
};
struct FileStorage {
    enum { READ = 0, WRITE = 1 };
    FileStorage(const char*, int) {}
    bool isOpened() const { return false; }
    FileNode operator[](const std::string&) const { std::abort(); }
    void release() {}
};
template <typename T> inline FileStorage& operator<<(FileStorage& fs, const T&) { std::abort(); return fs; }

}  // namespace cv
#endif
