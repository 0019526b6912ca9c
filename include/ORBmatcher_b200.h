/* ORBmatcher_b200.h — the three ORBmatcher functions of os1's hot path, behind their own signatures, on the GPU.
 *
 *   int ORBmatcher::SearchForInitialization(Frame&, Frame&, vector<Point2f>&, vector<int>&, int)   include/ORBmatcher.h:193, src/ORBmatcher.cc:400-515
 *   int ORBmatcher::SearchByProjection(Frame&, const Frame&, const float)                          include/ORBmatcher.h:118, src/ORBmatcher.cc:1292-1423
 *   int ORBmatcher::SearchByProjection(Frame&, const vector<MapPoint*>&, const float)              include/ORBmatcher.h:93,  src/ORBmatcher.cc:45-124
 *
 * The rest of os1's ORBmatcher (BoW, epipolar, Sim3 and fuse matchers) stays the reference's code, so this is not a replacement
 * class: the maintainer keeps include/ORBmatcher.h and src/ORBmatcher.cc and replaces the BODIES of these three member functions
 * by the one-line forwards shown in INTEGRATION.md (and exercised by tests/matcher_adapter_main.cpp).  Each forward gathers what the
 * function reads from its Frame / MapPoint arguments into plain arrays, calls the C ABI of include/os1_b200.h (where candidate
 * generation, distances, the ordered claim pass and the rotation filter run as CUDA kernels), and writes the results back into the
 * same members the reference writes: vnMatches12 / vbPrevMatched, CurrentFrame.mvpMapPoints, F.mvpMapPoints.
 *
 * Templates on the frame and map-point types, so this header includes none of os1's headers; it needs cv::KeyPoint, cv::Point2f
 * and cv::Mat from <opencv2/core/core.hpp> (the repo's tests compile it against a minimal stand-in).
 */
#ifndef ORBMATCHER_B200_H
#define ORBMATCHER_B200_H

#include <cstdlib>
#include <cstring>
#include <mutex>
#include <stdexcept>
#include <string>
#include <vector>

#include "os1_b200.h"

namespace os1_b200
{

static_assert(sizeof(cv::KeyPoint) == sizeof(os1_keypoint), "cv::KeyPoint must have the 28-byte layout of os1_keypoint");

namespace detail
{
struct Shared {
    std::mutex mu;
    os1_matcher* m = nullptr;
    os1_matcher* get()
    {
        if (!m) {
            int device = 0;
            if (const char* e = std::getenv("OS1_B200_DEVICE")) device = std::atoi(e);
            if (os1_matcher_create(&m, 16384, 16384, 4 << 20, device) != OS1_OK)
                throw std::runtime_error(std::string("os1_b200 matcher: ") + os1_last_error(nullptr));
        }
        return m;
    }
};
inline Shared& shared() { static Shared s; return s; }

inline void fail(os1_matcher* m) { throw std::runtime_error(std::string("os1_b200 matcher: ") + os1_matcher_last_error(m)); }

// mDescriptors (N x 32, CV_8U) as one run of bytes
inline const uint8_t* descriptor_rows(const cv::Mat& D, int n, std::vector<uint8_t>& tmp)
{
    if (n == 0) return nullptr;
    if ((size_t)D.step == 32) return D.ptr(0);
    tmp.resize((size_t)n * 32);
    for (int i = 0; i < n; i++) std::memcpy(&tmp[(size_t)i * 32], D.ptr(i), 32);
    return tmp.data();
}

template <class FrameT> inline void bounds_of(float (&b)[4])
{
    b[0] = FrameT::mnMinX; b[1] = FrameT::mnMaxX; b[2] = FrameT::mnMinY; b[3] = FrameT::mnMaxY;
}
} // namespace detail

/* ORBmatcher::SearchForInitialization — call site src/Tracking.cc:383-384 */
template <class FrameT>
int SearchForInitialization(FrameT& F1, FrameT& F2, std::vector<cv::Point2f>& vbPrevMatched, std::vector<int>& vnMatches12, int windowSize,
                            float mfNNratio, bool mbCheckOrientation)
{
    detail::Shared& S = detail::shared();
    std::lock_guard<std::mutex> lk(S.mu);
    os1_matcher* m = S.get();
    const int n1 = (int)F1.mvKeysUn.size(), n2 = (int)F2.mvKeysUn.size();
    vnMatches12 = std::vector<int>(n1, -1);                                                        // :403
    float b[4];
    detail::bounds_of<FrameT>(b);
    std::vector<uint8_t> t1, t2;
    std::vector<float> prev((size_t)n1 * 2);
    for (int i = 0; i < n1; i++) { prev[2 * i] = vbPrevMatched[i].x; prev[2 * i + 1] = vbPrevMatched[i].y; }
    int nmatches = 0;
    if (os1_matcher_search_for_initialization_host(m, (const os1_keypoint*)F1.mvKeysUn.data(), detail::descriptor_rows(F1.mDescriptors, n1, t1), n1,
                                                   (const os1_keypoint*)F2.mvKeysUn.data(), detail::descriptor_rows(F2.mDescriptors, n2, t2), n2, b,
                                                   prev.data(), windowSize, mfNNratio, mbCheckOrientation ? 1 : 0, vnMatches12.data(), &nmatches) != OS1_OK)
        detail::fail(m);
    for (int i = 0; i < n1; i++)
        if (vnMatches12[i] >= 0) { vbPrevMatched[i].x = prev[2 * i]; vbPrevMatched[i].y = prev[2 * i + 1]; }      // :507-510
    return nmatches;
}

/* ORBmatcher::SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, const float th) — call sites src/Tracking.cc:608,:614 */
template <class FrameT>
int SearchByProjection(FrameT& CurrentFrame, const FrameT& LastFrame, const float th, bool mbCheckOrientation)
{
    detail::Shared& S = detail::shared();
    std::lock_guard<std::mutex> lk(S.mu);
    os1_matcher* m = S.get();
    const int nc = (int)CurrentFrame.mvKeysUn.size(), nl = LastFrame.N;
    float b[4];
    detail::bounds_of<FrameT>(b);
    float T[16];
    for (int r = 0; r < 4; r++)
        for (int c = 0; c < 4; c++) T[4 * r + c] = CurrentFrame.mTcw.template at<float>(r, c);
    const float K4[4] = { CurrentFrame.fx, CurrentFrame.fy, CurrentFrame.cx, CurrentFrame.cy };
    std::vector<uint8_t> taken(nc), has(nl), outl(nl), dmp((size_t)nl * 32, 0), tc;
    std::vector<float> wpos((size_t)nl * 3, 0.f);
    std::vector<os1_keypoint> kl(nl);
    for (int i = 0; i < nc; i++) {
        auto* p = CurrentFrame.mvpMapPoints[i];
        taken[i] = p ? (p->Observations() > 0 ? 1 : 2) : 0;                                        // :1363-1365
    }
    for (int i = 0; i < nl; i++) {
        std::memcpy(&kl[i], &LastFrame.mvKeysUn[i], sizeof(os1_keypoint));                         // angle: mvKeysUn (:1386)
        kl[i].octave = LastFrame.mvKeys[i].octave;                                                 // octave: mvKeys  (:1344)
        auto* p = LastFrame.mvpMapPoints[i];
        has[i] = p ? (p->Observations() > 0 ? 1 : 2) : 0;
        outl[i] = LastFrame.mvbOutlier[i] ? 1 : 0;
        if (p && !outl[i]) {
            const cv::Mat x = p->GetWorldPos();                                                    // :1324
            for (int k = 0; k < 3; k++) wpos[3 * i + k] = x.template at<float>(k);
            const cv::Mat d = p->GetDescriptor();                                                  // :1354
            std::memcpy(&dmp[(size_t)i * 32], d.ptr(0), 32);
        }
    }
    std::vector<int> match(nc, OS1_MATCH_NONE);
    int nmatches = 0;
    if (os1_matcher_search_by_projection_frame_host(m, (const os1_keypoint*)CurrentFrame.mvKeysUn.data(), detail::descriptor_rows(CurrentFrame.mDescriptors, nc, tc),
                                                    nc, T, taken.data(), kl.data(), nl, has.data(), outl.data(), wpos.data(), dmp.data(), K4, b,
                                                    CurrentFrame.mvScaleFactors.data(), (int)CurrentFrame.mvScaleFactors.size(), th,
                                                    mbCheckOrientation ? 1 : 0, match.data(), &nmatches) != OS1_OK)
        detail::fail(m);
    for (int i = 0; i < nc; i++) {
        if (match[i] >= 0) CurrentFrame.mvpMapPoints[i] = LastFrame.mvpMapPoints[match[i]];        // :1381
        else if (match[i] == OS1_MATCH_REMOVED) CurrentFrame.mvpMapPoints[i] = NULL;               // :1413
    }
    return nmatches;
}

/* ORBmatcher::SearchByProjection(Frame& F, const vector<MapPoint*>& vpMapPoints, const float th) — call site src/Tracking.cc:823 */
template <class FrameT, class MapPointT>
int SearchByProjection(FrameT& F, const std::vector<MapPointT*>& vpMapPoints, const float th, float mfNNratio)
{
    detail::Shared& S = detail::shared();
    std::lock_guard<std::mutex> lk(S.mu);
    os1_matcher* m = S.get();
    const int nf = (int)F.mvKeysUn.size(), np = (int)vpMapPoints.size();
    float b[4];
    detail::bounds_of<FrameT>(b);
    std::vector<uint8_t> taken(nf), dmp((size_t)np * 32, 0), tf;
    std::vector<float> mp((size_t)np * 6, 0.f);
    for (int i = 0; i < nf; i++) {
        auto* p = F.mvpMapPoints[i];
        taken[i] = p ? (p->Observations() > 0 ? 1 : 2) : 0;                                        // :83-85
    }
    for (int i = 0; i < np; i++) {
        MapPointT* p = vpMapPoints[i];
        const bool view = p->mbTrackInView, bad = view ? p->isBad() : false;                       // :54-58
        mp[6 * i] = p->mTrackProjX; mp[6 * i + 1] = p->mTrackProjY; mp[6 * i + 2] = p->mTrackViewCos; mp[6 * i + 3] = (float)p->mnTrackScaleLevel;
        mp[6 * i + 4] = (float)((view ? 1 : 0) | (bad ? 2 : 0) | (p->plCandidato ? 4 : 0));
        if (view && !bad) {
            const cv::Mat d = p->GetDescriptor();                                                  // :76
            std::memcpy(&dmp[(size_t)i * 32], d.ptr(0), 32);
        }
    }
    std::vector<int> match(nf, OS1_MATCH_NONE);
    int nmatches = 0;
    if (os1_matcher_search_by_projection_points_host(m, (const os1_keypoint*)F.mvKeysUn.data(), detail::descriptor_rows(F.mDescriptors, nf, tf), nf, taken.data(),
                                                     mp.data(), dmp.data(), np, b, F.mvScaleFactors.data(), (int)F.mvScaleFactors.size(), th, mfNNratio,
                                                     match.data(), &nmatches) != OS1_OK)
        detail::fail(m);
    for (int i = 0; i < nf; i++)
        if (match[i] >= 0) F.mvpMapPoints[i] = vpMapPoints[match[i]];                              // :116
    return nmatches;
}

} // namespace os1_b200

#endif /* ORBMATCHER_B200_H */
