// TEST INFRASTRUCTURE — NOT PRODUCT CODE (see oracle/cv_exact.h header).
//
// Minimal stand-in for the slice of the cv:: API that /root/reference/src/ORBextractor.cc and
// /root/reference/include/ORBextractor.h touch, so that the reference file
// can be compiled *verbatim, by path* into oracle/_ref/ without OpenCV's C++ SDK (absent from this
// image).  Arithmetic is delegated to oracle/cv_exact.h, which is pinned bit-for-bit to cv2 4.13.0.
// Only 8-bit single-channel matrices exist here.
#pragma once
#include <cstddef>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <cassert>
#include <vector>
#include <algorithm>
#include "../../../cv_exact.h"

#define CV_8U 0
#define CV_8UC1 0
#define CV_PI 3.1415926535897932384626433832795

namespace cv {

typedef unsigned char uchar;

using cvx::cvRound;
using cvx::cvFloor;
using cvx::cvCeil;

template <typename T> struct Point_ {
    T x, y;
    Point_() : x(0), y(0) {}
    Point_(T _x, T _y) : x(_x), y(_y) {}
    template <typename U> Point_(const Point_<U>& p) : x((T)p.x), y((T)p.y) {}
    Point_& operator*=(float s) { x = (T)(x * s); y = (T)(y * s); return *this; }
};
typedef Point_<int> Point2i;
typedef Point2i Point;
typedef Point_<float> Point2f;

struct Size {
    int width, height;
    Size() : width(0), height(0) {}
    Size(int w, int h) : width(w), height(h) {}
};

struct Rect {
    int x, y, width, height;
    Rect() : x(0), y(0), width(0), height(0) {}
    Rect(int _x, int _y, int w, int h) : x(_x), y(_y), width(w), height(h) {}
};

struct KeyPoint {
    Point2f pt;
    float size, angle, response;
    int octave, class_id;
    KeyPoint() : pt(0, 0), size(0), angle(-1), response(0), octave(0), class_id(-1) {}
    KeyPoint(float x, float y, float _size, float _angle = -1, float _response = 0, int _octave = 0, int _class_id = -1)
        : pt(x, y), size(_size), angle(_angle), response(_response), octave(_octave), class_id(_class_id) {}
};

struct MatZeros { int rows, cols, type; };   // what Mat::zeros() hands to operator=

class Mat {
    // intrusive reference count placed in front of the pixels (malloc'ed: the canonical-order build of
    // the oracle swaps operator new for a bump arena, and pyramid buffers must outlive arena resets)
    struct Block { int refs; int pad[3]; };
    Block* blk;
    void retain() { if (blk) ++blk->refs; }
    void drop() { if (blk && --blk->refs == 0) free(blk); blk = nullptr; }
public:
    int rows, cols;
    size_t step;
    uchar* data;

    Mat() : blk(nullptr), rows(0), cols(0), step(0), data(nullptr) {}
    Mat(Size sz, int /*type*/) : blk(nullptr), rows(0), cols(0), step(0), data(nullptr) { create(sz.height, sz.width, CV_8U); }
    Mat(int r, int c, int /*type*/) : blk(nullptr), rows(0), cols(0), step(0), data(nullptr) { create(r, c, CV_8U); }
    // user-owned buffer
    Mat(int r, int c, int /*type*/, void* p, size_t st) : blk(nullptr), rows(r), cols(c), step(st), data((uchar*)p) {}
    Mat(const Mat& o) : blk(o.blk), rows(o.rows), cols(o.cols), step(o.step), data(o.data) { retain(); }
    Mat& operator=(const Mat& o)
    {
        if (this != &o) {
            Block* nb = o.blk; if (nb) ++nb->refs;
            drop();
            blk = nb; rows = o.rows; cols = o.cols; step = o.step; data = o.data;
        }
        return *this;
    }
    ~Mat() { drop(); }

    void create(int r, int c, int /*type*/)
    {
        if (data && rows == r && cols == c) return;
        drop();
        rows = r; cols = c; step = (size_t)c; data = nullptr;
        size_t n = (size_t)r * c;
        if (n) {
            blk = (Block*)malloc(sizeof(Block) + n);
            blk->refs = 1;
            data = (uchar*)(blk + 1);
        }
    }
    void release() { drop(); data = nullptr; rows = cols = 0; step = 0; }

    Mat operator()(const Rect& r) const
    {
        Mat m(*this);
        m.rows = r.height; m.cols = r.width;
        m.data = data + (size_t)r.y * step + r.x;
        return m;
    }
    Mat rowRange(int a, int b) const { return (*this)(Rect(0, a, cols, b - a)); }
    Mat colRange(int a, int b) const { return (*this)(Rect(a, 0, b - a, rows)); }

    template <typename T> T& at(int r, int c) { return *(T*)(data + (size_t)r * step + c); }
    template <typename T> const T& at(int r, int c) const { return *(const T*)(data + (size_t)r * step + c); }
    uchar* ptr(int r = 0) { return data + (size_t)r * step; }
    const uchar* ptr(int r = 0) const { return data + (size_t)r * step; }
    size_t step1() const { return step; }
    int type() const { return CV_8UC1; }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }

    Mat clone() const
    {
        Mat m; m.create(rows, cols, CV_8U);
        for (int y = 0; y < rows; y++) memcpy(m.data + (size_t)y * m.step, data + (size_t)y * step, cols);
        return m;
    }

    static MatZeros zeros(int r, int c, int t) { return MatZeros{ r, c, t }; }
    // cv::MatExpr assignment semantics: an existing header of the right size is filled IN PLACE
    // (ORBextractor.cc:901 relies on it: `descriptors` there aliases rows of the caller's output).
    Mat& operator=(const MatZeros& z)
    {
        create(z.rows, z.cols, z.type);
        for (int y = 0; y < rows; y++) memset(data + (size_t)y * step, 0, cols);
        return *this;
    }
};

// ---- Input/OutputArray proxies -------------------------------------------------------------
class _InputArray {
public:
    const Mat* m;
    _InputArray(const Mat& mm) : m(&mm) {}
    bool empty() const { return m->empty(); }
    Mat getMat() const { return *m; }
};
typedef const _InputArray& InputArray;

class _OutputArray {
public:
    Mat* m;
    _OutputArray(Mat& mm) : m(&mm) {}
    void release() const { m->release(); }
    void create(int r, int c, int t) const { m->create(r, c, t); }
    Mat getMat() const { return *m; }
};
typedef const _OutputArray& OutputArray;

// ---- functions -----------------------------------------------------------------------------
enum { INTER_LINEAR = 1 };
enum { BORDER_REFLECT_101 = 4, BORDER_ISOLATED = 16 };

inline float fastAtan2(float y, float x) { return cvx::fastAtan2(y, x); }

// hook: every cv::FAST result can be observed by the test wrapper (oracle/ref_wrapper.cpp)
typedef void (*FastHook)(const Mat& img, int threshold, const std::vector<KeyPoint>& kps);
inline FastHook& fast_hook() { static FastHook h = nullptr; return h; }

inline void FAST(const Mat& img, std::vector<KeyPoint>& keypoints, int threshold, bool nonmax = true)
{
    std::vector<cvx::FastPt> pts;
    cvx::fast9_16(img.data, img.cols, img.rows, img.step, threshold, nonmax, pts);
    keypoints.clear();
    keypoints.reserve(pts.size());
    for (size_t i = 0; i < pts.size(); i++)
        keypoints.push_back(KeyPoint((float)pts[i].x, (float)pts[i].y, 7.f, -1, (float)pts[i].score));
    if (fast_hook()) fast_hook()(img, threshold, keypoints);
}

inline void resize(const Mat& src, Mat& dst, Size dsize, double /*fx*/ = 0, double /*fy*/ = 0, int /*interp*/ = INTER_LINEAR)
{
    dst.create(dsize.height, dsize.width, CV_8U);   // no-op for the pre-sized ROI the reference passes
    // cv::resize switches INTER_LINEAR to INTER_AREA when both axes shrink by exactly 2;
    // os1 never configures that (scaleFactor 1.2 in every yaml) and this shim refuses it loudly.
    if (src.cols == 2 * dsize.width && src.rows == 2 * dsize.height) abort();
    cvx::resize_linear_u8(src.data, src.cols, src.rows, src.step, dst.data, dsize.width, dsize.height, dst.step);
}

inline void copyMakeBorder(const Mat& src, Mat& dst, int top, int bottom, int left, int right, int /*borderType*/)
{
    // the reference calls this with dst == the parent buffer of src's ROI (ORBextractor.cc:986) or a
    // fresh parent (:991); writing the centre is then the identity, so going through a copy is safe
    Mat s = src.clone();
    dst.create(src.rows + top + bottom, src.cols + left + right, CV_8U);
    cvx::copy_make_border_reflect101(s.data, s.cols, s.rows, s.step, dst.data, dst.step, top, bottom, left, right);
}

inline void GaussianBlur(const Mat& src, Mat& dst, Size ksize, double sx, double sy, int /*borderType*/)
{
    if (ksize.width != 7 || ksize.height != 7 || sx != 2 || sy != 2) abort();   // only the reference's call
    dst.create(src.rows, src.cols, CV_8U);
    cvx::gaussian7_s2_u8(src.data, src.cols, src.rows, src.step, dst.data, dst.step);
}

} // namespace cv
