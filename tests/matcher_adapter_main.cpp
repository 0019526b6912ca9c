// Test program for include/ORBmatcher_b200.h: the three ORBmatcher member functions get the one-line bodies INTEGRATION.md
// prescribes, and are then called exactly as Tracking.cc calls the reference's (Tracking.cc:383-384, :608, :823) on Frame /
// MapPoint objects.  The class declarations come from the test stand-ins oracle/refmatch/stubs.hpp (same member names, types and
// static-ness as os1's include/Frame.h, MapPoint.h, ORBmatcher.h); a scenario is read from a file written by the Python test.
// usage: matcher_adapter_main <kind: init|proj_frame|proj_points> <case file> <out file>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#define OS1_STUBS_STANDALONE 1
#include "stubs.hpp"
#include "ORBmatcher_b200.h"

namespace ORB_SLAM2 {

float Frame::fx, Frame::fy, Frame::cx, Frame::cy;
float Frame::mnMinX, Frame::mnMaxX, Frame::mnMinY, Frame::mnMaxY;
float Frame::mfGridElementWidthInv, Frame::mfGridElementHeightInv;

const int ORBmatcher::TH_HIGH = 100;
const int ORBmatcher::TH_LOW = 50;
const int ORBmatcher::HISTO_LENGTH = 30;
ORBmatcher::ORBmatcher(float nnratio, bool checkOri) : mfNNratio(nnratio), mbCheckOrientation(checkOri) {}

// ---- the bodies a maintainer puts into src/ORBmatcher.cc ----------------------------------------
int ORBmatcher::SearchForInitialization(Frame& F1, Frame& F2, std::vector<cv::Point2f>& vbPrevMatched, std::vector<int>& vnMatches12, int windowSize)
{
    return os1_b200::SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize, mfNNratio, mbCheckOrientation);
}
int ORBmatcher::SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, const float th)
{
    return os1_b200::SearchByProjection(CurrentFrame, LastFrame, th, mbCheckOrientation);
}
int ORBmatcher::SearchByProjection(Frame& F, const std::vector<MapPoint*>& vpMapPoints, const float th)
{
    return os1_b200::SearchByProjection(F, vpMapPoints, th, mfNNratio);
}

} // namespace ORB_SLAM2

using namespace ORB_SLAM2;

static std::vector<char> g_buf;
static size_t g_off = 0;
template <typename T> static const T* take(size_t count)
{
    const T* p = (const T*)(g_buf.data() + g_off);
    g_off += (count * sizeof(T) + 7) & ~(size_t)7;
    return p;
}

static void make_frame(Frame& F, const float* kp4, const uint8_t* D, int n)
{
    F.N = n;
    F.mvKeysUn.resize(n);
    for (int i = 0; i < n; i++) {
        F.mvKeysUn[i].pt.x = kp4[4 * i]; F.mvKeysUn[i].pt.y = kp4[4 * i + 1]; F.mvKeysUn[i].angle = kp4[4 * i + 2]; F.mvKeysUn[i].octave = (int)kp4[4 * i + 3];
    }
    F.mvKeys = F.mvKeysUn;
    F.mDescriptors = cv::Mat(n, 32, CV_8U);
    if (n) memcpy(F.mDescriptors.data, D, (size_t)n * 32);
    F.mvpMapPoints.assign(n, static_cast<MapPoint*>(NULL));
    F.mvbOutlier.assign(n, false);
}

static void set_bounds(const float* b)
{
    Frame::mnMinX = b[0]; Frame::mnMaxX = b[1]; Frame::mnMinY = b[2]; Frame::mnMaxY = b[3];
}

int main(int argc, char** argv)
{
    if (argc < 4) return 2;
    const std::string kind = argv[1];
    FILE* f = fopen(argv[2], "rb");
    if (!f) return 3;
    fseek(f, 0, SEEK_END); const long sz = ftell(f); fseek(f, 0, SEEK_SET);
    g_buf.resize(sz);
    if (fread(g_buf.data(), 1, sz, f) != (size_t)sz) return 3;
    fclose(f);
    FILE* o = fopen(argv[3], "wb");
    if (!o) return 4;
    const int* hdr = take<int>(8);
    const float* fl = take<float>(8);
    try {
        if (kind == "init") {
            const int n1 = hdr[0], n2 = hdr[1], window = hdr[2], checkOri = hdr[3];
            const float nnratio = fl[0];
            const float* b = take<float>(4);
            const float* kp1 = take<float>((size_t)n1 * 4); const uint8_t* D1 = take<uint8_t>((size_t)n1 * 32);
            const float* kp2 = take<float>((size_t)n2 * 4); const uint8_t* D2 = take<uint8_t>((size_t)n2 * 32);
            const float* prev = take<float>((size_t)n1 * 2);
            set_bounds(b);
            Frame F1, F2;
            make_frame(F1, kp1, D1, n1); make_frame(F2, kp2, D2, n2);
            std::vector<cv::Point2f> vb(n1);
            for (int i = 0; i < n1; i++) vb[i] = cv::Point2f(prev[2 * i], prev[2 * i + 1]);
            std::vector<int> m12;
            ORBmatcher matcher(nnratio, checkOri != 0);                                         // Tracking.cc:383
            const int nm = matcher.SearchForInitialization(F1, F2, vb, m12, window);            // Tracking.cc:384
            fwrite(&nm, 4, 1, o);
            fwrite(m12.data(), 4, n1, o);
            for (int i = 0; i < n1; i++) { fwrite(&vb[i].x, 4, 1, o); fwrite(&vb[i].y, 4, 1, o); }
        } else if (kind == "proj_frame") {
            const int nc = hdr[0], nl = hdr[1], checkOri = hdr[2], nlev = hdr[3];
            const float th = fl[0];
            const float* b = take<float>(4); const float* K4 = take<float>(4); const float* T = take<float>(16); const float* sf = take<float>(nlev);
            const float* kpc = take<float>((size_t)nc * 4); const uint8_t* Dc = take<uint8_t>((size_t)nc * 32); const uint8_t* taken = take<uint8_t>(nc);
            const float* kpl = take<float>((size_t)nl * 4); const uint8_t* has = take<uint8_t>(nl); const uint8_t* outl = take<uint8_t>(nl);
            const float* wpos = take<float>((size_t)nl * 3); const uint8_t* Dmp = take<uint8_t>((size_t)nl * 32);
            set_bounds(b);
            Frame::fx = K4[0]; Frame::fy = K4[1]; Frame::cx = K4[2]; Frame::cy = K4[3];
            Frame C, L;
            std::vector<uint8_t> zeros((size_t)(nl > 0 ? nl : 1) * 32, 0);
            make_frame(C, kpc, Dc, nc); make_frame(L, kpl, zeros.data(), nl);
            C.mTcw = cv::Mat(4, 4, CV_32F); memcpy(C.mTcw.data, T, 64);
            C.mvScaleFactors.assign(sf, sf + nlev); L.mvScaleFactors = C.mvScaleFactors;
            std::vector<MapPoint> pre(nc), pts(nl);
            for (int i = 0; i < nc; i++)
                if (taken[i]) { pre[i].nObs = taken[i] == 1 ? 1 : 0; C.mvpMapPoints[i] = &pre[i]; }
            for (int i = 0; i < nl; i++) {
                if (!has[i]) continue;
                pts[i].nObs = has[i] == 2 ? 0 : 1;
                pts[i].worldPos = cv::Mat(3, 1, CV_32F); memcpy(pts[i].worldPos.data, wpos + 3 * i, 12);
                pts[i].mDescriptor = cv::Mat(1, 32, CV_8U); memcpy(pts[i].mDescriptor.data, Dmp + (size_t)i * 32, 32);
                L.mvpMapPoints[i] = &pts[i];
                L.mvbOutlier[i] = outl[i] != 0;
            }
            ORBmatcher matcher(0.9f, checkOri != 0);                                            // Tracking.cc:596
            const int nm = matcher.SearchByProjection(C, L, th);                                // Tracking.cc:608
            fwrite(&nm, 4, 1, o);
            for (int i = 0; i < nc; i++) {
                MapPoint* p = C.mvpMapPoints[i];
                const int v = !p ? -1 : (p == &pre[i] ? -2 : (int)(p - pts.data()));
                fwrite(&v, 4, 1, o);
            }
        } else if (kind == "proj_points") {
            const int nf = hdr[0], np = hdr[1], nlev = hdr[2];
            const float th = fl[0], nnratio = fl[1];
            const float* b = take<float>(4); const float* sf = take<float>(nlev);
            const float* kpf = take<float>((size_t)nf * 4); const uint8_t* Df = take<uint8_t>((size_t)nf * 32); const uint8_t* taken = take<uint8_t>(nf);
            const float* mp = take<float>((size_t)np * 6); const uint8_t* Dmp = take<uint8_t>((size_t)np * 32);
            set_bounds(b);
            Frame F;
            make_frame(F, kpf, Df, nf);
            F.mvScaleFactors.assign(sf, sf + nlev);
            std::vector<MapPoint> pre(nf), pts(np);
            for (int i = 0; i < nf; i++)
                if (taken[i]) { pre[i].nObs = taken[i] == 1 ? 1 : 0; F.mvpMapPoints[i] = &pre[i]; }
            std::vector<MapPoint*> vp(np);
            for (int i = 0; i < np; i++) {
                MapPoint& P = pts[i];
                const int flags = (int)mp[6 * i + 4];
                P.mTrackProjX = mp[6 * i]; P.mTrackProjY = mp[6 * i + 1]; P.mTrackViewCos = mp[6 * i + 2]; P.mnTrackScaleLevel = (int)mp[6 * i + 3];
                P.mbTrackInView = (flags & 1) != 0; P.bad = (flags & 2) != 0; P.plCandidato = (flags & 4) != 0;
                P.mDescriptor = cv::Mat(1, 32, CV_8U); memcpy(P.mDescriptor.data, Dmp + (size_t)i * 32, 32);
                vp[i] = &P;
            }
            ORBmatcher matcher(nnratio, true);                                                  // Tracking.cc:818
            const int nm = matcher.SearchByProjection(F, vp, th);                               // Tracking.cc:823
            fwrite(&nm, 4, 1, o);
            for (int i = 0; i < nf; i++) {
                MapPoint* p = F.mvpMapPoints[i];
                const int v = !p ? -1 : (p == &pre[i] ? -2 : (int)(p - pts.data()));
                fwrite(&v, 4, 1, o);
            }
        } else return 2;
    } catch (const std::exception& e) {
        fprintf(stderr, "%s\n", e.what());
        return 5;
    }
    fclose(o);
    return 0;
}
