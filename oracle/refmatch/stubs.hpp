// TEST INFRASTRUCTURE — NOT PRODUCT CODE (see oracle/cv_exact.h header).
//
// Declarations of ORB_SLAM2::Frame / MapPoint / KeyFrame / ORBmatcher that hold exactly the members the sliced reference
// function bodies touch (names, types and static-ness as in /root/reference/include/Frame.h:242-260,:289-354,:388,
// :416,:427-433; MapPoint.h:308-319,:404; ORBmatcher.h:74-270).  The bodies themselves come, unmodified, from the
// reference tree (oracle/slice_reference.py).
#pragma once
#include <map>
#include <mutex>
#include <set>
#include <vector>
#include "cvmini.hpp"

#ifdef OS1_STUBS_STANDALONE          // used as plain declarations (tests/matcher_adapter_main.cpp): no sliced reference files around
#define FRAME_GRID_ROWS 48
#define FRAME_GRID_COLS 64
#else
#include "gen_frame_defs.inc"      // FRAME_GRID_ROWS / FRAME_GRID_COLS  (include/Frame.h:36-37)
#endif

namespace ORB_SLAM2 {

class KeyFrame {
public:
    bool bad = false;
    cv::Mat mDescriptors;
    bool isBad() { return bad; }
};

class MapPoint {
public:
    // what Tracking::SearchLocalPoints leaves in a map point before SearchByProjection (MapPoint.h:308-319, :404)
    float mTrackProjX = 0, mTrackProjY = 0;
    bool mbTrackInView = false;
    int mnTrackScaleLevel = 0;
    float mTrackViewCos = 0;
    bool plCandidato = false;
    // state behind the accessors
    bool bad = false;
    int nObs = 1;
    cv::Mat worldPos;          // 3x1 CV_32F
    cv::Mat mDescriptor;       // 1x32 CV_8U
    bool isBad() { return bad; }
    int Observations() { return nObs; }
    cv::Mat GetWorldPos() { return worldPos.clone(); }
    cv::Mat GetDescriptor() { return mDescriptor.clone(); }
    // ComputeDistinctiveDescriptors
    std::mutex mMutexFeatures;
    bool mbBad = false;
    std::map<KeyFrame*, size_t> mObservations;
    void ComputeDistinctiveDescriptors();
};

class Frame {
public:
    int N = 0;
    std::vector<cv::KeyPoint> mvKeys, mvKeysUn;
    cv::Mat mDescriptors;
    std::vector<MapPoint*> mvpMapPoints;
    std::vector<bool> mvbOutlier;
    cv::Mat mTcw;
    std::vector<float> mvScaleFactors;
    cv::Mat mK, mDistCoef;
    int camaraModo = 0;
    static float fx, fy, cx, cy;
    static float mnMinX, mnMaxX, mnMinY, mnMaxY;
    static float mfGridElementWidthInv, mfGridElementHeightInv;
    std::vector<std::size_t> mGrid[FRAME_GRID_COLS][FRAME_GRID_ROWS];

    void AssignFeaturesToGrid();
    std::vector<size_t> GetFeaturesInArea(const float& x, const float& y, const float& r, const int minLevel = -1, const int maxLevel = -1) const;
    bool PosInGrid(const cv::KeyPoint& kp, int& posX, int& posY);
    void UndistortKeyPoints();
    void ComputeImageBounds(const cv::Mat& imLeft);
    void antidistorsionarProyeccionEquidistante(cv::Mat& puntos);
};

class ORBmatcher {
public:
    ORBmatcher(float nnratio = 0.6, bool checkOri = true);
    static int DescriptorDistance(const cv::Mat& a, const cv::Mat& b);
    int SearchByProjection(Frame& F, const std::vector<MapPoint*>& vpMapPoints, const float th = 3);
    int SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, const float th);
    int SearchForInitialization(Frame& F1, Frame& F2, std::vector<cv::Point2f>& vbPrevMatched, std::vector<int>& vnMatches12, int windowSize = 10);
    static const int TH_LOW;
    static const int TH_HIGH;
    static const int HISTO_LENGTH;
    float RadiusByViewingCos(const float& viewCos);
    void ComputeThreeMaxima(std::vector<int>* histo, const int L, int& ind1, int& ind2, int& ind3);
    float mfNNratio;
    bool mbCheckOrientation;
};

// Osmap's wire records (protobuf messages in the reference, osmap.proto): only the accessors Osmap.cpp:775-787, :819-831 call
class SerializedDescriptor {
public:
    std::vector<unsigned int> blocks;
    void add_block(unsigned int v) { blocks.push_back(v); }
    unsigned int block(int i) const { return blocks[i]; }
    int block_size() const { return (int)blocks.size(); }
};
class SerializedKeypoint {
public:
    float x = 0, y = 0, a = 0; int o = 0;
    void set_ptx(float v) { x = v; }
    void set_pty(float v) { y = v; }
    void set_octave(int v) { o = v; }
    void set_angle(float v) { a = v; }
    float ptx() const { return x; }
    float pty() const { return y; }
    int octave() const { return o; }
    float angle() const { return a; }
};
class Osmap {
public:
    void serialize(const cv::Mat& m, SerializedDescriptor* serializedDescriptor);
    void deserialize(const SerializedDescriptor& serializedDescriptor, cv::Mat& m);
    void serialize(const cv::KeyPoint& kp, SerializedKeypoint* serializedKeypoint);
    void deserialize(const SerializedKeypoint& serializedKeypoint, cv::KeyPoint& kp);
};

} // namespace ORB_SLAM2
