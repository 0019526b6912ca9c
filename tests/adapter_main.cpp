// Test program for include/ORBextractor.h: uses the adapter exactly as Frame.cc / Tracking.cc use the reference class
// (new ORBextractor(...), getters by value, (*ext)(im, cv::Mat(), keys, descriptors)) and dumps the result.
// usage: adapter_main <raw gray file> <rows> <cols> <nfeatures> <out file>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "ORBextractor.h"

int main(int argc, char** argv)
{
    if (argc < 6) return 2;
    const int rows = atoi(argv[2]), cols = atoi(argv[3]), nf = atoi(argv[4]);
    std::vector<unsigned char> buf((size_t)rows * cols);
    FILE* f = fopen(argv[1], "rb");
    if (!f || fread(buf.data(), 1, buf.size(), f) != buf.size()) return 3;
    fclose(f);
    ORB_SLAM2::ORBextractor* ext = new ORB_SLAM2::ORBextractor(nf, 1.2f, 8, 20, 7);     // Tracking.cc:65
    int nlev = ext->GetLevels();                                                          // Frame.cc:69-75
    std::vector<float> sf = ext->GetScaleFactors(), isf = ext->GetInverseScaleFactors();
    std::vector<float> s2 = ext->GetScaleSigmaSquares(), is2 = ext->GetInverseScaleSigmaSquares();
    cv::Mat im(rows, cols, CV_8UC1, buf.data(), (size_t)cols);
    std::vector<cv::KeyPoint> keys;
    cv::Mat desc;
    (*ext)(im, cv::Mat(), keys, desc);                                                    // Frame.cc:133
    cv::Mat empty;
    std::vector<cv::KeyPoint> keys2(3);
    (*ext)(empty, cv::Mat(), keys2, desc);                                                // empty image: outputs untouched
    FILE* o = fopen(argv[5], "wb");
    int n = (int)keys.size();
    fwrite(&n, 4, 1, o); fwrite(&nlev, 4, 1, o);
    int untouched = (int)keys2.size(); fwrite(&untouched, 4, 1, o);
    float scale1 = sf[1] * isf[1] * s2[0] * is2[0]; fwrite(&scale1, 4, 1, o);
    for (int i = 0; i < n; i++) {
        float rec[6] = { keys[i].pt.x, keys[i].pt.y, keys[i].size, keys[i].angle, keys[i].response, (float)keys[i].octave };
        fwrite(rec, 4, 6, o);
    }
    for (int i = 0; i < n; i++) fwrite(desc.ptr(i), 1, 32, o);
    fclose(o);
    delete ext;
    return 0;
}
