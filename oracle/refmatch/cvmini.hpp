// TEST INFRASTRUCTURE — NOT PRODUCT CODE (see oracle/cv_exact.h header).
//
// The slice of the cv:: API that the sliced bodies of /root/reference/src/ORBmatcher.cc, Frame.cc and MapPoint.cc touch
// (oracle/slice_reference.py cuts them out BY PATH AND LINE RANGE at build time; nothing of the reference is copied into
// this repository).  Only what those bodies call exists here:
//   * Mat: 8-bit and 32-bit-float, two-dimensional, reference counted; row / col / rowRange / colRange headers,
//     at<T>, ptr<T>, reshape between 1 and 2 channels, clone;
//   * the three matrix expressions of ORBmatcher.cc:1304-1312,:1325 — `A*B+C`, `-A.t()*B`, `A*B` — evaluated the way
//     cv::gemm evaluates them for these shapes (pinned against cv2.gemm 4.13 by tests/test_oracle_primitives.py:
//     3x3 * 3x1 with flags == 0 takes gemm's small-matrix path = float products and float sums, left to right, then
//     (float)(t*alpha + c*beta) in double; a transposed operand takes the general path = double accumulation);
//   * undistortPoints(src, dst, K, dist, Mat(), K) through cvx::undistort_points (pinned against cv2 by the same test);
//   * Vec2f / Vec2d / Vec3d / Matx33f / Matx33d for Frame::antidistorsionarProyeccionEquidistante (Frame.cc:355-384).
//     ASSUMPTION, not a pin: `Matx33f * Vec3d` (Frame.cc:376) is evaluated here in double with the float matrix entries
//     widened — which overload real OpenCV selects for that mixed-type product cannot be checked without its headers.
#pragma once
#include <algorithm>
#include <cassert>
#include <climits>
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <vector>
#include "../cv_exact.h"

#define CV_8U 0
#define CV_8UC1 0
#define CV_32F 5
#define CV_32FC1 5
#define CV_32FC2 13

namespace cv {

typedef unsigned char uchar;

template <typename T> struct Point_ {
    T x, y;
    Point_() : x(0), y(0) {}
    Point_(T _x, T _y) : x(_x), y(_y) {}
};
typedef Point_<float> Point2f;

struct KeyPoint {
    Point2f pt;
    float size, angle, response;
    int octave, class_id;
    KeyPoint() : pt(0, 0), size(0), angle(-1), response(0), octave(0), class_id(-1) {}
};

template <typename T, int N> struct Vec {
    T val[N];
    Vec() { for (int i = 0; i < N; i++) val[i] = T(0); }
    Vec(T a, T b) { static_assert(N == 2, "two-element constructor"); val[0] = a; val[1] = b; }
    Vec(T a, T b, T c) { static_assert(N == 3, "three-element constructor"); val[0] = a; val[1] = b; val[2] = c; }
    template <typename U> Vec(const Vec<U, N>& o) { for (int i = 0; i < N; i++) val[i] = (T)o.val[i]; }      // saturate_cast between float types is a plain cast
    T& operator[](int i) { return val[i]; }
    const T& operator[](int i) const { return val[i]; }
};
typedef Vec<float, 2> Vec2f;
typedef Vec<double, 2> Vec2d;
typedef Vec<double, 3> Vec3d;
template <typename T, int N> inline Vec<T, N> operator*(const Vec<T, N>& a, double s)
{
    Vec<T, N> r;
    for (int i = 0; i < N; i++) r.val[i] = (T)(a.val[i] * s);
    return r;
}

template <typename T> struct Matx33 {
    T val[9];
    Matx33() { for (int i = 0; i < 9; i++) val[i] = T(0); }
    explicit Matx33(const T* p) { for (int i = 0; i < 9; i++) val[i] = p[i]; }
};
typedef Matx33<float> Matx33f;
typedef Matx33<double> Matx33d;
// see the ASSUMPTION in the file header
inline Vec3d operator*(const Matx33f& K, const Vec3d& v)
{
    Vec3d r;
    for (int i = 0; i < 3; i++) r.val[i] = K.val[3 * i] * v.val[0] + K.val[3 * i + 1] * v.val[1] + K.val[3 * i + 2] * v.val[2];
    return r;
}

class Mat;
struct MulExpr;
struct TransExpr;

class Mat {
    struct Block { int refs; int pad[3]; };
    Block* blk;
    void retain() { if (blk) ++blk->refs; }
    void drop() { if (blk && --blk->refs == 0) free(blk); blk = nullptr; }
    int tp;          // CV_8U or CV_32F
    int cn;          // channels (1 or 2)
public:
    int rows, cols;
    size_t step;
    uchar* data;

    Mat() : blk(nullptr), tp(CV_8U), cn(1), rows(0), cols(0), step(0), data(nullptr) {}
    Mat(int r, int c, int type) : blk(nullptr), tp(CV_8U), cn(1), rows(0), cols(0), step(0), data(nullptr) { create(r, c, type); }
    Mat(int r, int c, int type, void* p, size_t st = 0)
        : blk(nullptr), tp(type & 7), cn((type >> 3) + 1), rows(r), cols(c), step(st ? st : (size_t)c * elemSize()), data((uchar*)p) {}
    Mat(const Mat& o) : blk(o.blk), tp(o.tp), cn(o.cn), rows(o.rows), cols(o.cols), step(o.step), data(o.data) { retain(); }
    Mat(const MulExpr& e);
    Mat& operator=(const Mat& o)
    {
        if (this != &o) {
            Block* nb = o.blk; if (nb) ++nb->refs;
            drop();
            blk = nb; tp = o.tp; cn = o.cn; rows = o.rows; cols = o.cols; step = o.step; data = o.data;
        }
        return *this;
    }
    ~Mat() { drop(); }

    size_t elemSize1() const { return tp == CV_32F ? 4 : 1; }
    size_t elemSize() const { return elemSize1() * cn; }
    int type() const { return tp | ((cn - 1) << 3); }
    int channels() const { return cn; }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    bool isContinuous() const { return rows <= 1 || step == (size_t)cols * elemSize(); }

    void create(int r, int c, int type)
    {
        drop();
        tp = type & 7; cn = (type >> 3) + 1;
        rows = r; cols = c; step = (size_t)c * elemSize(); data = nullptr;
        const size_t n = (size_t)r * step;
        if (n) {
            blk = (Block*)malloc(sizeof(Block) + n);
            blk->refs = 1;
            data = (uchar*)(blk + 1);
            memset(data, 0, n);
        }
    }

    Mat sub(int r0, int r1, int c0, int c1) const
    {
        Mat m(*this);
        m.rows = r1 - r0; m.cols = c1 - c0;
        m.data = data + (size_t)r0 * step + (size_t)c0 * elemSize();
        return m;
    }
    Mat row(int r) const { return sub(r, r + 1, 0, cols); }
    Mat col(int c) const { return sub(0, rows, c, c + 1); }
    Mat rowRange(int a, int b) const { return sub(a, b, 0, cols); }
    Mat colRange(int a, int b) const { return sub(0, rows, a, b); }

    template <typename T> T& at(int r, int c) { return *(T*)(data + (size_t)r * step + (size_t)c * sizeof(T)); }
    template <typename T> const T& at(int r, int c) const { return *(const T*)(data + (size_t)r * step + (size_t)c * sizeof(T)); }
    // single index: element i0 of a row vector or of a column vector (cv::Mat::at(int))
    template <typename T> T& at(int i0) { return (rows == 1 || isContinuous()) ? ((T*)data)[i0] : *(T*)(data + (size_t)i0 * step); }
    template <typename T> const T& at(int i0) const { return (rows == 1 || isContinuous()) ? ((const T*)data)[i0] : *(const T*)(data + (size_t)i0 * step); }
    uchar* ptr(int r = 0) { return data + (size_t)r * step; }
    const uchar* ptr(int r = 0) const { return data + (size_t)r * step; }
    template <typename T> T* ptr(int r = 0) { return (T*)(data + (size_t)r * step); }
    template <typename T> const T* ptr(int r = 0) const { return (const T*)(data + (size_t)r * step); }

    // reshape(cn): same bytes, `cn` channels per element (only between 1 and 2 channels of a continuous matrix)
    Mat reshape(int newCn) const
    {
        assert(isContinuous());
        Mat m(*this);
        const int total = cols * cn;
        assert(total % newCn == 0);
        m.cn = newCn; m.cols = total / newCn;
        return m;
    }

    Mat clone() const
    {
        Mat m; m.create(rows, cols, type());
        for (int y = 0; y < rows; y++) memcpy(m.data + (size_t)y * m.step, data + (size_t)y * step, (size_t)cols * elemSize());
        return m;
    }

    TransExpr t() const;
};

// ---- the matrix expressions of the sliced bodies, evaluated as cv::gemm does for these shapes -------------------
struct TransExpr { Mat a; double alpha; };
struct MulExpr { Mat a, b; double alpha; bool aT; };
inline TransExpr Mat::t() const { return TransExpr{ *this, 1.0 }; }
inline TransExpr operator-(const TransExpr& e) { return TransExpr{ e.a, -e.alpha }; }
inline MulExpr operator*(const Mat& a, const Mat& b) { return MulExpr{ a, b, 1.0, false }; }
inline MulExpr operator*(const TransExpr& a, const Mat& b) { return MulExpr{ a.a, b, a.alpha, true }; }

// D = alpha * op(A) * B + beta * C for CV_32F matrices (C may be empty)
inline Mat gemm32f(const Mat& A, bool aT, const Mat& B, double alpha, const Mat* C, double beta)
{
    const int M = aT ? A.cols : A.rows, K = aT ? A.rows : A.cols, N = B.cols;
    assert(B.rows == K);
    Mat D(M, N, CV_32F);
    auto a = [&](int i, int k) -> float { return aT ? A.at<float>(k, i) : A.at<float>(i, k); };
    const bool small = !aT && K >= 2 && K <= 4 && (K == N || K == M);      // gemm's 2..4 special case needs flags == 0
    for (int i = 0; i < M; i++)
        for (int j = 0; j < N; j++) {
            const double c = C ? (double)C->at<float>(i, j) : 0.0;
            if (small) {
                volatile float t = a(i, 0) * B.at<float>(0, j);          // float products and sums; volatile forbids contraction
                for (int k = 1; k < K; k++) { volatile float p = a(i, k) * B.at<float>(k, j); t = t + p; }
                D.at<float>(i, j) = (float)((double)t * alpha + c * beta);
            } else {
                double s = 0;
                for (int k = 0; k < K; k++) s += (double)a(i, k) * (double)B.at<float>(k, j);
                D.at<float>(i, j) = (float)(s * alpha + c * beta);
            }
        }
    return D;
}
inline Mat::Mat(const MulExpr& e) : blk(nullptr), tp(CV_8U), cn(1), rows(0), cols(0), step(0), data(nullptr) { *this = gemm32f(e.a, e.aT, e.b, e.alpha, nullptr, 0.0); }
inline Mat operator+(const MulExpr& e, const Mat& c) { return gemm32f(e.a, e.aT, e.b, e.alpha, &c, 1.0); }

// cv::undistortPoints(src, dst, K, dist, R = Mat(), P = K) on an N x 1 two-channel float matrix, in place or not
inline void undistortPoints(const Mat& src, Mat& dst, const Mat& K, const Mat& dist, const Mat& /*R*/, const Mat& P)
{
    const int n = src.rows * src.cols;
    std::vector<float> in((size_t)n * 2), out((size_t)n * 2);
    memcpy(in.data(), src.data, (size_t)n * 8);
    const float K4[4] = { K.at<float>(0, 0), K.at<float>(1, 1), K.at<float>(0, 2), K.at<float>(1, 2) };
    assert(P.at<float>(0, 0) == K4[0] && P.at<float>(1, 1) == K4[1]);
    std::vector<float> d;
    const int nd = dist.rows * dist.cols;
    for (int i = 0; i < nd; i++) d.push_back(dist.rows == 1 ? dist.at<float>(0, i) : dist.at<float>(i, 0));
    cvx::undistort_points(in.data(), n, K4, d.data(), nd, out.data());
    if (dst.data != src.data || dst.rows != src.rows) dst.create(src.rows, src.cols, CV_32FC2);
    memcpy(dst.data, out.data(), (size_t)n * 8);
}

} // namespace cv
