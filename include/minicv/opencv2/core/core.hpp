// minicv — header-only stand-in for the handful of OpenCV *types* that the
// ORB_SLAM2::ORBextractor / ORBmatcher / Frame signatures mention
// (reference: include/ORBextractor.h:59-61,85; include/ORBmatcher.h:41-83;
// include/Frame.h:87-89).  Neither this container nor the GPU box has OpenCV
// C++ headers, so the drop-in classes are compiled against this file; with a
// real OpenCV on the include path this directory is simply left out (-I order)
// and the same sources compile against <opencv2/core/core.hpp> proper.
//
// Only data carriers live here (Mat, KeyPoint, Point_, Size, Rect,
// Input/OutputArray, cvRound & friends).  No image-processing primitive is
// implemented in this header: the product computes on the GPU, and the CPU
// restatements of resize / FAST / GaussianBlur / ... belong to oracle/.
#ifndef MINICV_CORE_HPP
#define MINICV_CORE_HPP

#include <algorithm>
#include <cassert>
#include <cfloat>
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <vector>

#define MINICV 1

typedef unsigned char uchar;
typedef unsigned short ushort;

#define CV_PI 3.1415926535897932384626433832795

#define CV_8U 0
#define CV_8S 1
#define CV_16U 2
#define CV_16S 3
#define CV_32S 4
#define CV_32F 5
#define CV_64F 6
#define CV_CN_SHIFT 3
#define CV_MAKETYPE(depth, cn) ((depth) + (((cn)-1) << CV_CN_SHIFT))
#define CV_8UC1 CV_MAKETYPE(CV_8U, 1)
#define CV_32FC1 CV_MAKETYPE(CV_32F, 1)
#define CV_32SC1 CV_MAKETYPE(CV_32S, 1)
#define CV_MAT_DEPTH(t) ((t) & 7)
#define CV_MAT_CN(t) ((((t) >> CV_CN_SHIFT) & 63) + 1)

// cvRound = round-half-to-even (SSE cvtsd2si / lrint under the default
// rounding mode), cvFloor / cvCeil as in opencv2/core/fast_math.hpp.
static inline int cvRound(double v) { return (int)lrint(v); }
static inline int cvRound(float v) { return (int)lrintf(v); }
static inline int cvRound(int v) { return v; }
static inline int cvFloor(double v) { int i = (int)v; return i - (i > v); }
static inline int cvFloor(float v) { int i = (int)v; return i - (i > v); }
static inline int cvCeil(double v) { int i = (int)v; return i + (i < v); }
static inline int cvCeil(float v) { int i = (int)v; return i + (i < v); }

namespace cv {

enum BorderTypes {
    BORDER_CONSTANT = 0, BORDER_REPLICATE = 1, BORDER_REFLECT = 2, BORDER_WRAP = 3,
    BORDER_REFLECT_101 = 4, BORDER_REFLECT101 = 4, BORDER_DEFAULT = 4, BORDER_ISOLATED = 16
};
enum InterpolationFlags { INTER_NEAREST = 0, INTER_LINEAR = 1, INTER_CUBIC = 2, INTER_AREA = 3 };

template <typename T> static inline T saturate_cast(int v);
template <> inline uchar saturate_cast<uchar>(int v) { return (uchar)((unsigned)v <= 255 ? v : v > 0 ? 255 : 0); }
template <> inline short saturate_cast<short>(int v) { return (short)((unsigned)(v + 32768) <= 65535 ? v : v > 0 ? 32767 : -32768); }

template <typename T> class Point_ {
public:
    T x, y;
    Point_() : x(0), y(0) {}
    Point_(T _x, T _y) : x(_x), y(_y) {}
    template <typename U> Point_(const Point_<U>& p) : x((T)p.x), y((T)p.y) {}
    Point_& operator*=(float s) { x = (T)(x * s); y = (T)(y * s); return *this; }
    Point_& operator+=(const Point_& p) { x += p.x; y += p.y; return *this; }
    bool operator==(const Point_& p) const { return x == p.x && y == p.y; }
};
// Point2i(float,float) with float args is used by the reference
// (ORBextractor.cc:667-668): the implicit float->int conversion truncates.
typedef Point_<int> Point2i;
typedef Point_<int> Point;
typedef Point_<float> Point2f;
typedef Point_<double> Point2d;

template <typename T> class Size_ {
public:
    T width, height;
    Size_() : width(0), height(0) {}
    Size_(T w, T h) : width(w), height(h) {}
};
typedef Size_<int> Size;

template <typename T> class Rect_ {
public:
    T x, y, width, height;
    Rect_() : x(0), y(0), width(0), height(0) {}
    Rect_(T _x, T _y, T w, T h) : x(_x), y(_y), width(w), height(h) {}
};
typedef Rect_<int> Rect;

// Layout-compatible with cv::KeyPoint of OpenCV 2.4/3.x/4.x: 7 x 4 bytes.
class KeyPoint {
public:
    Point2f pt;
    float size;
    float angle;
    float response;
    int octave;
    int class_id;
    KeyPoint() : pt(0, 0), size(0), angle(-1), response(0), octave(0), class_id(-1) {}
    KeyPoint(Point2f _pt, float _size, float _angle = -1, float _response = 0, int _octave = 0, int _class_id = -1)
        : pt(_pt), size(_size), angle(_angle), response(_response), octave(_octave), class_id(_class_id) {}
    KeyPoint(float x, float y, float _size, float _angle = -1, float _response = 0, int _octave = 0, int _class_id = -1)
        : pt(x, y), size(_size), angle(_angle), response(_response), octave(_octave), class_id(_class_id) {}
};
static_assert(sizeof(KeyPoint) == 28, "cv::KeyPoint must stay 28 bytes (C ABI orbx_keypoint aliases it)");

class Mat;

// Result of Mat::zeros(): assigning it to an existing Mat of the same shape
// writes zeros *in place* (OpenCV MatExpr semantics, relied upon by
// ORBextractor.cc:1228 where `descriptors` is a row range of the output).
struct MatZerosExpr { int rows, cols, type; };

class Mat {
public:
    int flags;          // type
    int rows, cols;
    uchar* data;
    size_t step;        // bytes per row

    Mat() : flags(0), rows(0), cols(0), data(nullptr), step(0) {}
    Mat(int r, int c, int t) : Mat() { create(r, c, t); }
    Mat(Size sz, int t) : Mat() { create(sz.height, sz.width, t); }
    // user-owned memory, no copy (OpenCV: Mat(rows, cols, type, void* data, size_t step))
    Mat(int r, int c, int t, void* d, size_t s = 0)
        : flags(t), rows(r), cols(c), data((uchar*)d), step(s ? s : (size_t)c * esz(t)) {}
    Mat(const MatZerosExpr& z) : Mat() { *this = z; }

    static size_t esz(int t) {
        static const int d[8] = {1, 1, 2, 2, 4, 4, 8, 2};
        return (size_t)d[CV_MAT_DEPTH(t)] * CV_MAT_CN(t);
    }
    void create(int r, int c, int t) {
        if (data && r == rows && c == cols && t == flags) return;
        flags = t; rows = r; cols = c;
        step = (size_t)c * esz(t);
        size_t n = step * (size_t)r;
        owner.reset(n ? new uchar[n] : nullptr, std::default_delete<uchar[]>());
        data = owner.get();
    }
    void create(Size sz, int t) { create(sz.height, sz.width, t); }
    void release() { owner.reset(); data = nullptr; rows = cols = 0; step = 0; }
    Mat& operator=(const MatZerosExpr& z) {
        create(z.rows, z.cols, z.type);
        for (int i = 0; i < rows; i++) std::memset(ptr(i), 0, (size_t)cols * elemSize());
        return *this;
    }
    static MatZerosExpr zeros(int r, int c, int t) { return MatZerosExpr{r, c, t}; }

    int type() const { return flags; }
    int depth() const { return CV_MAT_DEPTH(flags); }
    int channels() const { return CV_MAT_CN(flags); }
    size_t elemSize() const { return esz(flags); }
    size_t step1() const { return step / esz(CV_MAT_DEPTH(flags)); }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    bool isContinuous() const { return step == (size_t)cols * elemSize() || rows <= 1; }
    Size size() const { return Size(cols, rows); }
    size_t total() const { return (size_t)rows * cols; }

    uchar* ptr(int r = 0) { return data + step * (size_t)r; }
    const uchar* ptr(int r = 0) const { return data + step * (size_t)r; }
    template <typename T> T* ptr(int r = 0) { return (T*)(data + step * (size_t)r); }
    template <typename T> const T* ptr(int r = 0) const { return (const T*)(data + step * (size_t)r); }
    template <typename T> T& at(int r, int c) { return ((T*)(data + step * (size_t)r))[c]; }
    template <typename T> const T& at(int r, int c) const { return ((const T*)(data + step * (size_t)r))[c]; }
    template <typename T> T& at(int i) {
        return rows == 1 ? ((T*)data)[i] : *(T*)(data + step * (size_t)i);
    }
    template <typename T> const T& at(int i) const {
        return rows == 1 ? ((const T*)data)[i] : *(const T*)(data + step * (size_t)i);
    }

    Mat operator()(const Rect& r) const {
        Mat m(*this);
        m.data = data + step * (size_t)r.y + (size_t)r.x * elemSize();
        m.rows = r.height; m.cols = r.width;
        return m;
    }
    Mat rowRange(int a, int b) const { return (*this)(Rect(0, a, cols, b - a)); }
    Mat colRange(int a, int b) const { return (*this)(Rect(a, 0, b - a, rows)); }
    Mat row(int r) const { return rowRange(r, r + 1); }
    Mat col(int c) const { return colRange(c, c + 1); }
    Mat clone() const {
        Mat m;
        if (empty()) return m;
        m.create(rows, cols, flags);
        for (int i = 0; i < rows; i++) std::memcpy(m.ptr(i), ptr(i), (size_t)cols * elemSize());
        return m;
    }
    void copyTo(Mat& dst) const {
        dst.create(rows, cols, flags);
        for (int i = 0; i < rows; i++) std::memmove(dst.ptr(i), ptr(i), (size_t)cols * elemSize());
    }

#ifdef MINICV_MAT_EXTRA_MEMBERS     // hook for the oracle's reference build (oracle/ref_build/cvstub): float algebra members
    MINICV_MAT_EXTRA_MEMBERS
#endif

private:
    std::shared_ptr<uchar> owner;   // null for user-owned memory
};

class _InputArray {
public:
    _InputArray() : m_(nullptr) {}
    _InputArray(const Mat& m) : m_(const_cast<Mat*>(&m)) {}
    Mat getMat() const { return m_ ? *m_ : Mat(); }
    bool empty() const { return !m_ || m_->empty(); }
protected:
    Mat* m_;
};
class _OutputArray : public _InputArray {
public:
    _OutputArray() {}
    _OutputArray(Mat& m) : _InputArray(m) {}
    void create(int rows, int cols, int type) const { if (m_) m_->create(rows, cols, type); }
    void release() const { if (m_) m_->release(); }
};
typedef const _InputArray& InputArray;
typedef const _OutputArray& OutputArray;
static inline InputArray noArray() { static _InputArray none; return none; }

}  // namespace cv

#endif  // MINICV_CORE_HPP
