// TEST INFRASTRUCTURE — NOT PRODUCT CODE.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may use
// anything under oracle/.  The product path (os1_b200/) never includes, links or calls this.
//
// cv_exact.h — CPU restatement of the OpenCV 4.x primitives that os1's ORBextractor.cc calls
// (the arithmetic lives in OpenCV, a system dependency of the reference that is not vendored and
// not version-pinned: /root/reference/.cproject:40,58-64).  Each routine restates the *published*
// OpenCV algorithm for exactly the call made by the reference and is pinned bit-for-bit against
// opencv-python-headless 4.13.0 by tests/test_oracle_primitives.py (+ tests/golden/ fixtures made by
// oracle/gen_golden.py).  Call sites restated:
//   resize INTER_LINEAR u8      ORBextractor.cc:984
//   GaussianBlur 7x7 s=2 R101   ORBextractor.cc:950
//   FAST(9_16, nonmax)          ORBextractor.cc:848,854
//   fastAtan2                   ORBextractor.cc:112
//   cvRound / cvFloor / cvCeil  ORBextractor.cc:90,139,143-144,474,488-492,976
//   copyMakeBorder REFLECT_101  ORBextractor.cc:986,991
#pragma once
#include <cmath>
#include <cfloat>
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <vector>
#include <algorithm>

namespace cvx {

typedef unsigned char uchar;
using std::ptrdiff_t;

// round-half-to-even, as cvRound does through cvtsd2si / cvtss2si on x86-64
static inline int cvRound(double v) { return (int)lrint(v); }
static inline int cvRound(float v) { return (int)lrintf(v); }
static inline int cvRound(int v) { return v; }
static inline int cvFloor(double v) { int i = (int)v; return i - (i > v); }
static inline int cvFloor(float v) { int i = (int)v; return i - (i > v); }
static inline int cvCeil(double v) { int i = (int)v; return i + (i < v); }
static inline int cvCeil(float v) { int i = (int)v; return i + (i < v); }

static inline short sat_short(float v)
{
    int i = cvRound(v);
    return (short)(i < -32768 ? -32768 : i > 32767 ? 32767 : i);
}

// ---- fastAtan2 (degrees, [0,360)) : 7th-order odd polynomial, f32, no contraction ------------
static inline float fastAtan2(float y, float x)
{
    // constants are products of two floats, rounded to float (as the OpenCV source spells them)
    static const float R = (float)(180.0 / 3.1415926535897932384626433832795);
    static const float p1 = 0.9997878412794807f * R;
    static const float p3 = -0.3258083974640975f * R;
    static const float p5 = 0.1555786518463281f * R;
    static const float p7 = -0.04432655554792128f * R;
    volatile float t;   // volatile stores forbid fma contraction whatever -march the oracle gets
    float ax = std::fabs(x), ay = std::fabs(y), a, c, c2;
    if (ax >= ay) {
        c = ay / (ax + (float)DBL_EPSILON);
        t = c * c; c2 = t;
        t = p7 * c2; t = t + p5; t = t * c2; t = t + p3; t = t * c2; t = t + p1; t = t * c;
        a = t;
    } else {
        c = ax / (ay + (float)DBL_EPSILON);
        t = c * c; c2 = t;
        t = p7 * c2; t = t + p5; t = t * c2; t = t + p3; t = t * c2; t = t + p1; t = t * c;
        a = 90.f - t;
    }
    if (x < 0) a = 180.f - a;
    if (y < 0) a = 360.f - a;
    return a;
}

// ---- resize, INTER_LINEAR, 8-bit single channel -------------------------------------------
// Fixed-point bilinear with 11-bit coefficients: horizontal pass into int32, vertical pass
// (((b0*(S0>>4))>>16) + ((b1*(S1>>4))>>16) + 2) >> 2.
struct LinearTab { std::vector<int> ofs; std::vector<short> c0, c1; };

static inline void linear_tab_x(int ssize, int dsize, LinearTab& t)
{
    double inv_scale = (double)dsize / ssize;
    double scale = 1. / inv_scale;
    t.ofs.resize(dsize); t.c0.resize(dsize); t.c1.resize(dsize);
    for (int d = 0; d < dsize; d++) {
        float f = (float)((d + 0.5) * scale - 0.5);
        int s = cvFloor(f);
        f -= s;
        if (s < 0) { f = 0; s = 0; }
        if (s >= ssize - 1) { f = 0; s = ssize - 1; }
        t.ofs[d] = s;
        t.c0[d] = sat_short((1.f - f) * 2048.f);
        t.c1[d] = sat_short(f * 2048.f);
    }
}
// the vertical table keeps the fractional weight and clamps the two ROW INDICES instead
static inline void linear_tab_y(int ssize, int dsize, LinearTab& t)
{
    double inv_scale = (double)dsize / ssize;
    double scale = 1. / inv_scale;
    t.ofs.resize(dsize); t.c0.resize(dsize); t.c1.resize(dsize);
    for (int d = 0; d < dsize; d++) {
        float f = (float)((d + 0.5) * scale - 0.5);
        int s = cvFloor(f);
        f -= s;
        t.ofs[d] = s;
        t.c0[d] = sat_short((1.f - f) * 2048.f);
        t.c1[d] = sat_short(f * 2048.f);
    }
}

static inline void resize_linear_u8(const uchar* src, int sw, int sh, size_t sstep,
                                    uchar* dst, int dw, int dh, size_t dstep)
{
    LinearTab tx, ty;
    linear_tab_x(sw, dw, tx);
    linear_tab_y(sh, dh, ty);
    // horizontally filtered source rows, computed on first use
    std::vector<int> H((size_t)sh * dw);
    std::vector<char> done(sh, 0);
    for (int dy = 0; dy < dh; dy++) {
        int sy[2];
        for (int k = 0; k < 2; k++) {
            sy[k] = std::min(std::max(ty.ofs[dy] + k, 0), sh - 1);
            if (!done[sy[k]]) {
                const uchar* S = src + (size_t)sy[k] * sstep;
                int* D = &H[(size_t)sy[k] * dw];
                for (int dx = 0; dx < dw; dx++) {
                    int s0 = tx.ofs[dx], s1 = std::min(s0 + 1, sw - 1);
                    D[dx] = S[s0] * tx.c0[dx] + S[s1] * tx.c1[dx];
                }
                done[sy[k]] = 1;
            }
        }
        const int *S0 = &H[(size_t)sy[0] * dw], *S1 = &H[(size_t)sy[1] * dw];
        int b0 = ty.c0[dy], b1 = ty.c1[dy];
        uchar* D = dst + (size_t)dy * dstep;
        for (int x = 0; x < dw; x++)
            D[x] = (uchar)((((b0 * (S0[x] >> 4)) >> 16) + ((b1 * (S1[x] >> 4)) >> 16) + 2) >> 2);
    }
}

// ---- border helpers ---------------------------------------------------------------------------
static inline int reflect101(int p, int len)
{
    if (len == 1) return 0;
    while (p < 0 || p >= len) { if (p < 0) p = -p; else p = 2 * len - 2 - p; }
    return p;
}

static inline void copy_make_border_reflect101(const uchar* src, int w, int h, size_t sstep,
                                               uchar* dst, size_t dstep, int top, int bottom, int left, int right)
{
    for (int y = -top; y < h + bottom; y++) {
        const uchar* S = src + (size_t)reflect101(y, h) * sstep;
        uchar* D = dst + (size_t)(y + top) * dstep;
        for (int x = -left; x < w + right; x++) D[x + left] = S[reflect101(x, w)];
    }
}

// ---- GaussianBlur 7x7, sigma 2, BORDER_REFLECT_101, u8 ------------------------------------------
// OpenCV's bit-exact fixed-point path: Q8 kernel {18,34,48,56,48,34,18}; no rounding between the
// passes; dst = (V + 32768) >> 16.
static inline void gaussian7_s2_u8(const uchar* src, int w, int h, size_t sstep, uchar* dst, size_t dstep)
{
    static const int k[7] = { 18, 34, 48, 56, 48, 34, 18 };
    // horizontal pass over a reflect-padded copy of each row (values <= 65280: exact in 16 bits)
    std::vector<unsigned short> H((size_t)w * h);
    std::vector<uchar> row((size_t)w + 6);
    for (int y = 0; y < h; y++) {
        const uchar* S = src + (size_t)y * sstep;
        for (int x = -3; x < w + 3; x++) row[x + 3] = S[reflect101(x, w)];
        unsigned short* D = &H[(size_t)y * w];
        const uchar* R = row.data();
        for (int x = 0; x < w; x++)
            D[x] = (unsigned short)(k[0] * (R[x] + R[x + 6]) + k[1] * (R[x + 1] + R[x + 5]) + k[2] * (R[x + 2] + R[x + 4]) + k[3] * R[x + 3]);
    }
    std::vector<uchar> out((size_t)w * h);   // allows in-place use
    for (int y = 0; y < h; y++) {
        const unsigned short* r[7];
        for (int j = 0; j < 7; j++) r[j] = &H[(size_t)reflect101(y + j - 3, h) * w];
        uchar* O = &out[(size_t)y * w];
        for (int x = 0; x < w; x++) {
            int v = k[0] * (r[0][x] + r[6][x]) + k[1] * (r[1][x] + r[5][x]) + k[2] * (r[2][x] + r[4][x]) + k[3] * r[3][x];
            O[x] = (uchar)((v + 32768) >> 16);
        }
    }
    for (int y = 0; y < h; y++) memcpy(dst + (size_t)y * dstep, &out[(size_t)y * w], w);
}

// ---- FAST 9_16 with non-max suppression --------------------------------------------------------
struct FastPt { int x, y, score; };

static const int fast_dx[16] = { 0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1 };
static const int fast_dy[16] = { 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3 };

// max over the 16 arcs of 9 contiguous ring pixels of min(c - ring) and of min(ring - c).
// A pixel is a corner at threshold t  <=>  fast_arc_strength > t ; cv's response = strength - 1.
static inline int fast_arc_strength(const uchar* p, size_t step)
{
    int d[25];
    int c = p[0];
    for (int k = 0; k < 16; k++) d[k] = c - p[(ptrdiff_t)fast_dy[k] * (ptrdiff_t)step + fast_dx[k]];
    for (int k = 16; k < 25; k++) d[k] = d[k - 16];
    int best = 0;
    for (int k = 0; k < 16; k++) {
        int mn = d[k], mx = d[k];
        for (int j = 1; j < 9; j++) { mn = std::min(mn, d[k + j]); mx = std::max(mx, d[k + j]); }
        best = std::max(best, mn);     // centre brighter than the whole arc by >= mn
        best = std::max(best, -mx);    // centre darker  than the whole arc by >= -mx
    }
    return best;
}

// Necessary condition: an arc of 9 contiguous ring pixels holds one pixel of each opposite pair (k, k+8),
// so every pair needs a member brighter than c+t, or every pair needs a member darker than c-t.
static inline bool fast_may_be_corner(const uchar* p, size_t step, int t)
{
    int c = p[0], hi = c + t, lo = c - t;
    bool br = true, dk = true;
    for (int k = 0; k < 8 && (br || dk); k++) {
        int a = p[(ptrdiff_t)fast_dy[k] * (ptrdiff_t)step + fast_dx[k]];
        int b = p[(ptrdiff_t)fast_dy[k + 8] * (ptrdiff_t)step + fast_dx[k + 8]];
        br = br && (a > hi || b > hi);
        dk = dk && (a < lo || b < lo);
    }
    return br || dk;
}

// keypoints of the w x h image in row-major order, coordinates relative to (0,0) of the image
static inline void fast9_16(const uchar* img, int w, int h, size_t step, int threshold, bool nonmax,
                            std::vector<FastPt>& out)
{
    out.clear();
    if (w < 7 || h < 7) return;
    std::vector<uchar> sc((size_t)w * h, 0), corner((size_t)w * h, 0);
    for (int y = 3; y < h - 3; y++)
        for (int x = 3; x < w - 3; x++) {
            const uchar* p = img + (size_t)y * step + x;
            if (!fast_may_be_corner(p, step, threshold)) continue;
            int s = fast_arc_strength(p, step);
            if (s > threshold) { sc[(size_t)y * w + x] = (uchar)(s - 1); corner[(size_t)y * w + x] = 1; }
        }
    for (int y = 3; y < h - 3; y++)
        for (int x = 3; x < w - 3; x++) {
            if (!corner[(size_t)y * w + x]) continue;
            int s = sc[(size_t)y * w + x];
            if (nonmax) {
                const uchar* r = &sc[(size_t)y * w + x];
                if (!(s > r[-1] && s > r[1] && s > r[-w - 1] && s > r[-w] && s > r[-w + 1] &&
                      s > r[w - 1] && s > r[w] && s > r[w + 1])) continue;
            }
            out.push_back(FastPt{ x, y, s });
        }
}

// ---- undistortPoints(src, dst, K, dist, R = Mat(), P = K) ------------------------------------------------------
// OpenCV's published iterative inverse of the Brown model: five fixed-point iterations, everything in double, then the
// projection through P = K.  K4 = {fx, fy, cx, cy} (float entries widened), dist = k1,k2,p1,p2[,k3[,k4,k5,k6[,s1..s4]]].
// Pinned against cv2.undistortPoints 4.13 for 4, 5 and 8 coefficients (tests/test_oracle_primitives.py).
// Call sites: Frame.cc:307 (UndistortKeyPoints), :338 (ComputeImageBounds).
static inline void undistort_points(const float* xy, int n, const float* K4, const float* dist, int ndist, float* out)
{
    const double fx = K4[0], fy = K4[1], cx = K4[2], cy = K4[3];
    double k[14] = { 0 };
    for (int i = 0; i < ndist && i < 14; i++) k[i] = dist[i];
    const double ifx = 1. / fx, ify = 1. / fy;
    for (int i = 0; i < n; i++) {
        double x = xy[2 * i], y = xy[2 * i + 1];
        x = (x - cx) * ifx; y = (y - cy) * ify;
        const double x0 = x, y0 = y;
        for (int j = 0; j < 5; j++) {
            const double r2 = x * x + y * y;
            const double icdist = (1 + ((k[7] * r2 + k[6]) * r2 + k[5]) * r2) / (1 + ((k[4] * r2 + k[1]) * r2 + k[0]) * r2);
            if (icdist < 0) { x = x0; y = y0; break; }
            const double deltaX = 2 * k[2] * x * y + k[3] * (r2 + 2 * x * x) + k[8] * r2 + k[9] * r2 * r2;
            const double deltaY = k[2] * (r2 + 2 * y * y) + 2 * k[3] * x * y + k[10] * r2 + k[11] * r2 * r2;
            x = (x0 - deltaX) * icdist;
            y = (y0 - deltaY) * icdist;
        }
        const double xx = fx * x + 0.0 * y + cx, yy = 0.0 * x + fy * y + cy, ww = 1. / (0.0 * x + 0.0 * y + 1.0);
        out[2 * i] = (float)(xx * ww); out[2 * i + 1] = (float)(yy * ww);
    }
}

} // namespace cvx
