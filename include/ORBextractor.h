/* ORBextractor.h — drop-in replacement for os1's include/ORBextractor.h.
 *
 * Keeps the reference's class name, namespace, constructor, operator() and getters
 * (reference include/ORBextractor.h:150-321; call sites src/Tracking.cc:65,67 and src/Frame.cc:69-75,133), so
 * Frame.cc and Tracking.cc compile and call it unchanged, and forwards to the C ABI of include/os1_b200.h, behind
 * which the whole extraction runs as CUDA kernels on a B200.  Header-only: add this directory to the include path
 * in front of the reference's own include/, drop src/ORBextractor.cc from the build, link libos1_b200.so.
 *
 * Needs <opencv2/core/core.hpp> for cv::Mat / cv::KeyPoint / cv::InputArray / cv::OutputArray — the real OpenCV in
 * os1's build; the repo's tests compile it against a minimal stand-in for those headers (test infrastructure).
 */
#ifndef ORBEXTRACTOR_H
#define ORBEXTRACTOR_H

#include <cassert>
#include <cstdlib>
#include <cstring>
#include <list>
#include <stdexcept>
#include <string>
#include <vector>
#include <opencv2/core/core.hpp>

#include "os1_b200.h"

namespace ORB_SLAM2
{

class ORBextractor
{
public:
    enum { HARRIS_SCORE = 0, FAST_SCORE = 1 };      // reference ORBextractor.h:159

    // reference src/ORBextractor.cc:442-502.  The CUDA device comes from OS1_B200_DEVICE (default 0).
    ORBextractor(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST) : h_(nullptr)
    {
        int device = 0;
        if (const char* e = std::getenv("OS1_B200_DEVICE")) device = std::atoi(e);
        if (os1_orb_create(&h_, nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, device) != OS1_OK)
            throw std::runtime_error(std::string("ORBextractor: ") + os1_last_error(nullptr));
        nlevels_ = nlevels;
        mvImagePyramid.resize(nlevels);              // :465
    }
    ~ORBextractor() { os1_orb_destroy(h_); }
    ORBextractor(const ORBextractor&) = delete;
    ORBextractor& operator=(const ORBextractor&) = delete;

    // reference src/ORBextractor.cc:907-969.  Mask is ignored, as in the reference (ORBextractor.h:184).
    void operator()(cv::InputArray _image, cv::InputArray /*_mask*/, std::vector<cv::KeyPoint>& _keypoints, cv::OutputArray _descriptors)
    {
        if (_image.empty()) return;                  // :910-911 — outputs untouched
        cv::Mat image = _image.getMat();
        assert(image.type() == CV_8UC1);             // :914
        const int cap = os1_orb_max_keypoints(h_);
        kp_.resize(cap);
        desc_.resize((size_t)cap * 32);
        int n = 0;
        if (os1_orb_extract(h_, image.data, image.rows, image.cols, (size_t)image.step, kp_.data(), cap, desc_.data(), &n) != OS1_OK)
            throw std::runtime_error(std::string("ORBextractor: ") + os1_last_error(h_));
        if (n == 0) _descriptors.release();          // :928-929
        else {
            _descriptors.create(n, 32, CV_8U);       // :932
            cv::Mat d = _descriptors.getMat();
            for (int i = 0; i < n; i++) std::memcpy(d.ptr(i), &desc_[(size_t)i * 32], 32);
        }
        _keypoints.clear();                          // :936-937
        _keypoints.reserve(n);
        for (int i = 0; i < n; i++) {
            const os1_keypoint& k = kp_[i];
            _keypoints.push_back(cv::KeyPoint(k.x, k.y, k.size, k.angle, k.response, k.octave, k.class_id));
        }
#ifdef OS1_B200_FILL_IMAGE_PYRAMID
        FillImagePyramid();
#endif
    }

    int inline GetLevels() { return os1_orb_get_levels(h_); }                     // :194
    float inline GetScaleFactor() { return os1_orb_get_scale_factor(h_); }        // :200
    std::vector<float> inline GetScaleFactors() { return table(0); }              // :210
    std::vector<float> inline GetInverseScaleFactors() { return table(1); }       // :219
    std::vector<float> inline GetScaleSigmaSquares() { return table(2); }         // :228
    std::vector<float> inline GetInverseScaleSigmaSquares() { return table(3); }  // :237

    // reference ORBextractor.h:245.  Nothing else in os1 reads it, so it is only filled on request (a device->host
    // copy of every level); the 19-px border the reference adds around each level is not reproduced.
    std::vector<cv::Mat> mvImagePyramid;
    void FillImagePyramid()
    {
        for (int l = 0; l < nlevels_; l++) {
            int w = 0, hh = 0;
            if (os1_orb_level_size(h_, l, &w, &hh) != OS1_OK) return;
            mvImagePyramid[l].create(hh, w, CV_8U);
            os1_orb_get_pyramid_level(h_, 0, l, mvImagePyramid[l].data);
        }
    }

    os1_orb* handle() { return h_; }

protected:
    std::vector<float> table(int which)
    {
        std::vector<float> t[4];
        for (int i = 0; i < 4; i++) t[i].resize(nlevels_);
        os1_orb_get_tables(h_, t[0].data(), t[1].data(), t[2].data(), t[3].data());
        return t[which];
    }
    os1_orb* h_;
    int nlevels_;
    std::vector<os1_keypoint> kp_;
    std::vector<unsigned char> desc_;
};

} // namespace ORB_SLAM2

#endif
