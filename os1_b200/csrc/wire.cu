// Descriptor / keypoint wire records of os1's map files (include/os1_b200.h "wire formats").
// Host code only: the records are re-interpretations of the arrays the extractor already returns.
//   Osmap::serialize(const Mat&, SerializedDescriptor*)      src/Osmap.cpp:775-779   eight native-endian 32-bit blocks per row
//   Osmap::deserialize(const SerializedDescriptor&, Mat&)    src/Osmap.cpp:781-786
//   Osmap::serialize / deserialize of a KeyPoint             src/Osmap.cpp:819-831   (pt.x, pt.y, octave, angle)
//   Converter::toDescriptorVector                            src/Converter.cc:27-36  one 1 x 32 row per descriptor: row i of the
//                                                            matrix as it is, nothing to convert
#include "../../include/os1_b200.h"
#include "os1_internal.cuh"
#include <cstring>

extern "C" {

int os1_osmap_descriptor_blocks(const uint8_t* desc, int n, uint32_t* blocks)
{
    if (n < 0 || (n > 0 && (!desc || !blocks))) { os1_internal_set_error("os1_osmap_descriptor_blocks: bad argument"); return OS1_E_INVALID; }
    // ((unsigned int*)m.data)[i]: the host's own byte order, so this is a copy of the rows
    for (int i = 0; i < n; i++)
        for (int k = 0; k < 8; k++) {
            uint32_t v;
            memcpy(&v, desc + (size_t)i * 32 + 4 * k, 4);
            blocks[(size_t)i * 8 + k] = v;
        }
    return OS1_OK;
}

int os1_osmap_descriptors_from_blocks(const uint32_t* blocks, int n, uint8_t* desc)
{
    if (n < 0 || (n > 0 && (!desc || !blocks))) { os1_internal_set_error("os1_osmap_descriptors_from_blocks: bad argument"); return OS1_E_INVALID; }
    for (int i = 0; i < n; i++)
        for (int k = 0; k < 8; k++) {
            const uint32_t v = blocks[(size_t)i * 8 + k];
            memcpy(desc + (size_t)i * 32 + 4 * k, &v, 4);
        }
    return OS1_OK;
}

int os1_osmap_keypoint_records(const os1_keypoint* kp, int n, float* ptx, float* pty, int32_t* octave, float* angle)
{
    if (n < 0 || (n > 0 && (!kp || !ptx || !pty || !octave || !angle))) { os1_internal_set_error("os1_osmap_keypoint_records: bad argument"); return OS1_E_INVALID; }
    for (int i = 0; i < n; i++) { ptx[i] = kp[i].x; pty[i] = kp[i].y; octave[i] = kp[i].octave; angle[i] = kp[i].angle; }
    return OS1_OK;
}

int os1_osmap_keypoints_from_records(const float* ptx, const float* pty, const int32_t* octave, const float* angle, int n, os1_keypoint* kp)
{
    if (n < 0 || (n > 0 && (!kp || !ptx || !pty || !octave || !angle))) { os1_internal_set_error("os1_osmap_keypoints_from_records: bad argument"); return OS1_E_INVALID; }
    for (int i = 0; i < n; i++) {
        // a default-constructed cv::KeyPoint with the four stored fields set (Osmap.cpp:826-831): size 0, response 0, class_id -1
        kp[i].x = ptx[i]; kp[i].y = pty[i]; kp[i].size = 0.f; kp[i].angle = angle[i]; kp[i].response = 0.f; kp[i].octave = octave[i]; kp[i].class_id = -1;
    }
    return OS1_OK;
}

} // extern "C"
