// TEST INFRASTRUCTURE — NOT PRODUCT CODE (see oracle/cv_exact.h header).
//
// C wrapper around the UNMODIFIED bodies of the reference's matcher-side hot-path functions:
//   ORBmatcher::DescriptorDistance / ComputeThreeMaxima / SearchForInitialization / SearchByProjection (x2)
//                                                  /root/reference/src/ORBmatcher.cc:1605-1621, :1554-1595, :400-515, :45-124, :1292-1423
//   Frame::AssignFeaturesToGrid / GetFeaturesInArea / PosInGrid / UndistortKeyPoints / ComputeImageBounds /
//   antidistorsionarProyeccionEquidistante         /root/reference/src/Frame.cc:114-129, :209-274, :286-384
//   MapPoint::ComputeDistinctiveDescriptors        /root/reference/src/MapPoint.cc:227-292
//   Osmap::serialize / deserialize of a descriptor and of a keypoint    /root/reference/src/Osmap.cpp:775-787, :819-831
// The bodies are cut out of the reference tree by path and line range when this file is built (oracle/slice_reference.py
// -> oracle/_ref/gen_*.inc, git-ignored) and compiled between the stub declarations of stubs.hpp.  Built into
// oracle/_ref/libos1_refmatch.so; it validates the restated oracle (oracle/orb_oracle.cpp) and generates the committed
// golden vectors tests/golden/matcher_*.npz.
#include <climits>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <algorithm>
#include "stubs.hpp"

using namespace std;       // as the reference's translation units do (ORBmatcher.cc:33, Frame.cc, MapPoint.cc)

namespace ORB_SLAM2 {

float Frame::fx, Frame::fy, Frame::cx, Frame::cy;
float Frame::mnMinX, Frame::mnMaxX, Frame::mnMinY, Frame::mnMaxY;
float Frame::mfGridElementWidthInv, Frame::mfGridElementHeightInv;

#include "gen_orbmatcher.inc"
#include "gen_frame.inc"
#include "gen_mappoint.inc"
using namespace cv;        // Osmap.cpp:37
#include "gen_osmap.inc"

} // namespace ORB_SLAM2

using namespace ORB_SLAM2;

namespace {

// kp rows: x, y, angle, octave
void fill_keys(std::vector<cv::KeyPoint>& v, const float* kp, int n)
{
    v.resize(n);
    for (int i = 0; i < n; i++) {
        v[i].pt.x = kp[4 * i]; v[i].pt.y = kp[4 * i + 1]; v[i].angle = kp[4 * i + 2]; v[i].octave = (int)kp[4 * i + 3];
    }
}

// the part of Frame::Frame that follows ComputeImageBounds (Frame.cc:95-98): the grid pitch from the bounds
void set_bounds(const float* b4)
{
    Frame::mnMinX = b4[0]; Frame::mnMaxX = b4[1]; Frame::mnMinY = b4[2]; Frame::mnMaxY = b4[3];
    Frame::mfGridElementWidthInv = static_cast<float>(FRAME_GRID_COLS) / static_cast<float>(Frame::mnMaxX - Frame::mnMinX);
    Frame::mfGridElementHeightInv = static_cast<float>(FRAME_GRID_ROWS) / static_cast<float>(Frame::mnMaxY - Frame::mnMinY);
}

void make_frame(Frame& F, const float* kp, const uint8_t* D, int n)
{
    F.N = n;
    fill_keys(F.mvKeysUn, kp, n);
    F.mvKeys = F.mvKeysUn;
    F.mDescriptors = cv::Mat(n, 32, CV_8U);
    if (n) memcpy(F.mDescriptors.data, D, (size_t)n * 32);
    F.mvpMapPoints.assign(n, static_cast<MapPoint*>(NULL));
    F.mvbOutlier.assign(n, false);
    F.AssignFeaturesToGrid();
}

cv::Mat mat_f32(int r, int c, const float* p)
{
    cv::Mat m(r, c, CV_32F);
    memcpy(m.data, p, (size_t)r * c * 4);
    return m;
}

void set_camera(Frame& F, const float* K4, const float* dist, int ndist, int mode)
{
    const float K[9] = { K4[0], 0.f, K4[2], 0.f, K4[1], K4[3], 0.f, 0.f, 1.f };
    F.mK = mat_f32(3, 3, K);
    // Tracking stores the coefficients as a column (Tracking.cc:1226-1238); mode 1 has none (Frame.cc:88: camaraModo = rows ? 0 : 1)
    F.mDistCoef = (mode == 0 && ndist > 0) ? mat_f32(ndist, 1, dist) : cv::Mat();
    if (mode == 0 && ndist == 0) { const float z[4] = { 0, 0, 0, 0 }; F.mDistCoef = mat_f32(4, 1, z); }
    F.camaraModo = F.mDistCoef.rows ? 0 : 1;
}

} // namespace

extern "C" {

int rm_descriptor_distance(const uint8_t* a, const uint8_t* b)
{
    cv::Mat A(1, 32, CV_8U, (void*)a), B(1, 32, CV_8U, (void*)b);
    return ORBmatcher::DescriptorDistance(A, B);
}

// the two matrix expressions of ORBmatcher.cc:1304-1312,:1325 through this build's cv::Mat stand-in (pinned against cv2.gemm)
void rm_gemm3_abc(const float* T44, const float* x3, float* d3)
{
    const cv::Mat T = mat_f32(4, 4, T44), x = mat_f32(3, 1, x3);
    const cv::Mat R = T.rowRange(0, 3).colRange(0, 3), t = T.rowRange(0, 3).col(3);
    cv::Mat d = R * x + t;
    for (int i = 0; i < 3; i++) d3[i] = d.at<float>(i);
}
void rm_gemm3_atb(const float* T44, float* d3)
{
    const cv::Mat T = mat_f32(4, 4, T44);
    const cv::Mat R = T.rowRange(0, 3).colRange(0, 3), t = T.rowRange(0, 3).col(3);
    const cv::Mat d = -R.t() * t;
    for (int i = 0; i < 3; i++) d3[i] = d.at<float>(i);
}

// ComputeThreeMaxima on a histogram given by its bin sizes
void rm_three_maxima(const int* sizes, int L, int* ind3)
{
    std::vector<std::vector<int> > h(L);
    for (int i = 0; i < L; i++) h[i].assign(sizes[i], 0);
    int i1 = -1, i2 = -1, i3 = -1;
    ORBmatcher m(0.9f, true);
    m.ComputeThreeMaxima(h.data(), L, i1, i2, i3);
    ind3[0] = i1; ind3[1] = i2; ind3[2] = i3;
}

// Frame::AssignFeaturesToGrid + GetFeaturesInArea for a list of queries; CSR output in the reference's own order
int rm_features_in_area(const float* kp, int n, const float* bounds4, const float* qx, const float* qy, const float* qr, const int* qmin,
                        const int* qmax, int nq, int* ptr, int* cand, int cap)
{
    set_bounds(bounds4);
    Frame F;
    std::vector<uint8_t> D((size_t)std::max(n, 1) * 32, 0);
    make_frame(F, kp, D.data(), n);
    int total = 0;
    for (int q = 0; q < nq; q++) {
        ptr[q] = total;
        const std::vector<size_t> v = F.GetFeaturesInArea(qx[q], qy[q], qr[q], qmin ? qmin[q] : -1, qmax ? qmax[q] : -1);
        for (size_t k = 0; k < v.size(); k++) { if (total < cap) cand[total] = (int)v[k]; total++; }
    }
    ptr[nq] = total;
    return total;
}

// ORBmatcher::SearchForInitialization, whole function (candidate generation through Frame::GetFeaturesInArea included).
// prev: n1 x 2 (vbPrevMatched), updated in place as :507-510 does.  Returns nmatches.
int rm_search_for_initialization(const float* kp1, const uint8_t* D1, int n1, const float* kp2, const uint8_t* D2, int n2, const float* bounds4,
                                 float* prev, int windowSize, float nnratio, int checkOri, int* matches12)
{
    set_bounds(bounds4);
    Frame F1, F2;
    make_frame(F1, kp1, D1, n1);
    make_frame(F2, kp2, D2, n2);
    std::vector<cv::Point2f> vb(n1);
    for (int i = 0; i < n1; i++) vb[i] = cv::Point2f(prev[2 * i], prev[2 * i + 1]);
    std::vector<int> m12;
    ORBmatcher matcher(nnratio, checkOri != 0);
    const int nm = matcher.SearchForInitialization(F1, F2, vb, m12, windowSize);
    for (int i = 0; i < n1; i++) { matches12[i] = m12[i]; prev[2 * i] = vb[i].x; prev[2 * i + 1] = vb[i].y; }
    return nm;
}

// ORBmatcher::SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, th), whole function.
//   cur:  kpc (n x {x,y,angle,octave} undistorted), Dc, TcwC (4x4), taken[i2] = 0 none / 1 point with observations / 2 point without
//   last: kpl (angle from mvKeysUn, octave from mvKeys: the same here), has[i] (0 none, 1 point with observations, 2 point without) /
//         outlier[i], world positions (3 floats) and descriptors
//   K4, bounds4, scale factors.  out_match[i2] = index i of the last-frame keypoint whose point was attached, -1 none,
//   -2 = still holds the point it came in with.
int rm_search_by_projection_frame(const float* kpc, const uint8_t* Dc, int nc, const float* TcwC, const uint8_t* taken, const float* kpl, int nl,
                                  const float* TcwL, const uint8_t* has, const uint8_t* outlier, const float* wpos, const uint8_t* Dmp,
                                  const float* K4, const float* bounds4, const float* scaleFactors, int nlevels, float th, int checkOri,
                                  int* out_match)
{
    set_bounds(bounds4);
    Frame::fx = K4[0]; Frame::fy = K4[1]; Frame::cx = K4[2]; Frame::cy = K4[3];
    Frame C, L;
    make_frame(C, kpc, Dc, nc);
    std::vector<uint8_t> dl((size_t)std::max(nl, 1) * 32, 0);
    make_frame(L, kpl, dl.data(), nl);
    C.mTcw = mat_f32(4, 4, TcwC); L.mTcw = mat_f32(4, 4, TcwL);
    C.mvScaleFactors.assign(scaleFactors, scaleFactors + nlevels);
    L.mvScaleFactors = C.mvScaleFactors;
    std::vector<MapPoint> pre(nc), pts(nl);
    for (int i = 0; i < nc; i++)
        if (taken[i]) { pre[i].nObs = taken[i] == 1 ? 1 : 0; C.mvpMapPoints[i] = &pre[i]; }
    for (int i = 0; i < nl; i++) {
        if (!has[i]) continue;
        pts[i].nObs = has[i] == 2 ? 0 : 1;
        pts[i].worldPos = mat_f32(3, 1, wpos + 3 * i);
        pts[i].mDescriptor = cv::Mat(1, 32, CV_8U);
        memcpy(pts[i].mDescriptor.data, Dmp + (size_t)i * 32, 32);
        L.mvpMapPoints[i] = &pts[i];
        L.mvbOutlier[i] = outlier[i] != 0;
    }
    ORBmatcher matcher(0.9f, checkOri != 0);
    const int nm = matcher.SearchByProjection(C, L, th);
    for (int i = 0; i < nc; i++) {
        MapPoint* p = C.mvpMapPoints[i];
        out_match[i] = !p ? -1 : (p == &pre[i] ? -2 : (int)(p - pts.data()));
    }
    return nm;
}

// ORBmatcher::SearchByProjection(Frame& F, const vector<MapPoint*>&, th), whole function.
//   mp rows (6 floats): mTrackProjX, mTrackProjY, mTrackViewCos, mnTrackScaleLevel, flags (1 = mbTrackInView, 2 = isBad, 4 = plCandidato), unused
int rm_search_by_projection_points(const float* kpf, const uint8_t* Df, int nf, const uint8_t* taken, const float* mp, const uint8_t* Dmp, int nmp,
                                   const float* bounds4, const float* scaleFactors, int nlevels, float th, float nnratio, int* out_match)
{
    set_bounds(bounds4);
    Frame F;
    make_frame(F, kpf, Df, nf);
    F.mvScaleFactors.assign(scaleFactors, scaleFactors + nlevels);
    std::vector<MapPoint> pre(nf), pts(nmp);
    for (int i = 0; i < nf; i++)
        if (taken[i]) { pre[i].nObs = taken[i] == 1 ? 1 : 0; F.mvpMapPoints[i] = &pre[i]; }
    std::vector<MapPoint*> vp(nmp);
    for (int i = 0; i < nmp; i++) {
        MapPoint& P = pts[i];
        const int flags = (int)mp[6 * i + 4];
        P.mTrackProjX = mp[6 * i]; P.mTrackProjY = mp[6 * i + 1]; P.mTrackViewCos = mp[6 * i + 2]; P.mnTrackScaleLevel = (int)mp[6 * i + 3];
        P.mbTrackInView = (flags & 1) != 0; P.bad = (flags & 2) != 0; P.plCandidato = (flags & 4) != 0;
        P.mDescriptor = cv::Mat(1, 32, CV_8U);
        memcpy(P.mDescriptor.data, Dmp + (size_t)i * 32, 32);
        vp[i] = &P;
    }
    ORBmatcher matcher(nnratio, true);
    const int nm = matcher.SearchByProjection(F, vp, th);
    for (int i = 0; i < nf; i++) {
        MapPoint* p = F.mvpMapPoints[i];
        out_match[i] = !p ? -1 : (p == &pre[i] ? -2 : (int)(p - pts.data()));
    }
    return nm;
}

// Frame::UndistortKeyPoints (mode 0: pinhole with `ndist` coefficients, mode 1: os1's equidistant fisheye)
void rm_undistort(const float* xy, int n, const float* K4, const float* dist, int ndist, int mode, float* out)
{
    Frame F;
    set_camera(F, K4, dist, ndist, mode);
    F.N = n;
    F.mvKeys.resize(n);
    for (int i = 0; i < n; i++) { F.mvKeys[i].pt.x = xy[2 * i]; F.mvKeys[i].pt.y = xy[2 * i + 1]; }
    F.UndistortKeyPoints();
    for (int i = 0; i < n; i++) { out[2 * i] = F.mvKeysUn[i].pt.x; out[2 * i + 1] = F.mvKeysUn[i].pt.y; }
}

// Frame::ComputeImageBounds: bounds4 = mnMinX, mnMaxX, mnMinY, mnMaxY
void rm_image_bounds(int cols, int rows, const float* K4, const float* dist, int ndist, int mode, float* bounds4)
{
    Frame F;
    set_camera(F, K4, dist, ndist, mode);
    cv::Mat im; im.rows = rows; im.cols = cols;
    F.ComputeImageBounds(im);
    bounds4[0] = Frame::mnMinX; bounds4[1] = Frame::mnMaxX; bounds4[2] = Frame::mnMinY; bounds4[3] = Frame::mnMaxY;
}

// MapPoint::ComputeDistinctiveDescriptors for npoints points; observations of point m = rows ptr[m]..ptr[m+1] of desc, in map
// iteration order (the KeyFrame stubs live in one array, so std::map<KeyFrame*, size_t> iterates them by index); bad[k] marks
// observations whose keyframe isBad().  out_desc: 32 bytes per point (the chosen mDescriptor), valid[m] = 0 when the function
// returned without choosing.
void rm_distinctive_descriptors(const uint8_t* desc, const int* ptr, const uint8_t* bad, int npoints, uint8_t* out_desc, uint8_t* valid)
{
    for (int m = 0; m < npoints; m++) {
        const int n = ptr[m + 1] - ptr[m];
        std::vector<KeyFrame> kfs(n);
        MapPoint P;
        for (int k = 0; k < n; k++) {
            kfs[k].mDescriptors = cv::Mat(1, 32, CV_8U);
            memcpy(kfs[k].mDescriptors.data, desc + (size_t)(ptr[m] + k) * 32, 32);
            kfs[k].bad = bad && bad[ptr[m] + k];
            P.mObservations[&kfs[k]] = 0;
        }
        P.ComputeDistinctiveDescriptors();
        valid[m] = P.mDescriptor.empty() ? 0 : 1;
        if (valid[m]) memcpy(out_desc + (size_t)m * 32, P.mDescriptor.data, 32);
    }
}

// Osmap's descriptor record: the eight 32-bit blocks of each 32-byte row (Osmap.cpp:775-779) and the row rebuilt from them (:781-786)
void rm_osmap_descriptor_blocks(const uint8_t* desc, int n, unsigned int* blocks, uint8_t* back)
{
    Osmap om;
    for (int i = 0; i < n; i++) {
        cv::Mat row(1, 32, CV_8U, (void*)(desc + (size_t)i * 32));
        SerializedDescriptor sd;
        om.serialize(row, &sd);
        for (int k = 0; k < 8; k++) blocks[8 * i + k] = sd.block(k);
        cv::Mat m;
        om.deserialize(sd, m);
        memcpy(back + (size_t)i * 32, m.data, 32);
    }
}
// Osmap's keypoint record (Osmap.cpp:819-831): rows in = x, y, angle, octave; rec = ptx, pty, octave, angle as stored
void rm_osmap_keypoints(const float* kp, int n, float* rec_xy_angle, int* rec_octave)
{
    Osmap om;
    std::vector<cv::KeyPoint> v;
    fill_keys(v, kp, n);
    for (int i = 0; i < n; i++) {
        SerializedKeypoint sk;
        om.serialize(v[i], &sk);
        cv::KeyPoint k2;
        om.deserialize(sk, k2);
        rec_xy_angle[3 * i] = k2.pt.x; rec_xy_angle[3 * i + 1] = k2.pt.y; rec_xy_angle[3 * i + 2] = k2.angle; rec_octave[i] = k2.octave;
    }
}

} // extern "C"
