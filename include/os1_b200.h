/* os1_b200.h — C ABI of the B200-native ORB front end and Hamming consumer for os1 (ORB-SLAM2 monocular).
 *
 * Plain pointers and sizes only; no C++ or torch types cross this boundary and no exception escapes it.
 * Every function returns 0 on success or a negative OS1_E_* code; os1_last_error() gives the text.
 * Citations are into the reference tree (AlejandroSilvestri/os1).  The header-only C++ adapter
 * include/ORBextractor.h keeps the reference's ORB_SLAM2::ORBextractor class on top of these calls;
 * INTEGRATION.md shows the binding a maintainer adds to the reference.
 *
 * Threading: a handle owns one CUDA stream and its device buffers; calls on one handle must be serialised
 * by the caller (the reference calls its extractor from the tracker thread only, src/System.cc:64-66, and
 * the object is not re-entrant, src/ORBextractor.cc:979-986).  Distinct handles are independent and may
 * live on different GPUs.  There is no CPU fallback: without a CUDA device every call fails with OS1_E_CUDA.
 *
 * Current device: an entry point makes the GPU it works on current for the duration of the call and restores the
 * calling thread's previous current device before it returns, on success and on failure alike.
 */
#ifndef OS1_B200_H
#define OS1_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define OS1_OK 0
#define OS1_E_INVALID (-1)     /* bad argument */
#define OS1_E_CUDA (-2)        /* CUDA runtime / driver failure (text in os1_last_error) */
#define OS1_E_CAPACITY (-3)    /* a caller-supplied or internal capacity was exceeded */
#define OS1_E_UNSUPPORTED (-4) /* configuration outside the supported range (see DESIGN.md "limits") */

typedef struct os1_orb os1_orb;

/* Binary layout of cv::KeyPoint (28 bytes): what the reference stores in Frame::mvKeys. */
typedef struct os1_keypoint {
    float x, y;        /* pt, in level-0 pixel units (src/ORBextractor.cc:959-965) */
    float size;        /* (float)(int)(31 * mvScaleFactor[octave])  (:877,:888) */
    float angle;       /* degrees [0,360), cv::fastAtan2 of the intensity centroid (:86-113) */
    float response;    /* FAST score (cv::FAST response) */
    int32_t octave;    /* pyramid level (:887) */
    int32_t class_id;  /* always -1 */
} os1_keypoint;

/* ---- extractor: replaces ORB_SLAM2::ORBextractor (include/ORBextractor.h:150-321) ------------------- */

/* ORBextractor::ORBextractor(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST)  src/ORBextractor.cc:442-502.
 * `device` is the CUDA ordinal the handle lives on. */
int os1_orb_create(os1_orb** out, int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST, int device);
/* ~ORBextractor (include/ORBextractor.h:170).  Frees every device and pinned buffer of the handle. */
int os1_orb_destroy(os1_orb* h);
/* Text of the last failure on this handle (h == NULL: last failure of a call that had no handle). */
const char* os1_last_error(const os1_orb* h);

/* GetLevels / GetScaleFactor  include/ORBextractor.h:194-203 */
int os1_orb_get_levels(const os1_orb* h);
float os1_orb_get_scale_factor(const os1_orb* h);
/* GetScaleFactors / GetInverseScaleFactors / GetScaleSigmaSquares / GetInverseScaleSigmaSquares
 * include/ORBextractor.h:210-239 — each array receives nlevels floats (NULL pointers are skipped). */
int os1_orb_get_tables(const os1_orb* h, float* scale, float* inv_scale, float* sigma2, float* inv_sigma2);
/* mnFeaturesPerLevel (src/ORBextractor.cc:467-478): nlevels ints. */
int os1_orb_get_features_per_level(const os1_orb* h, int* quota);
/* Upper bound of keypoints one frame can return (nfeatures plus the quadtree's overshoot). */
int os1_orb_max_keypoints(const os1_orb* h);

/* Pre-size the device buffers for `max_frames` frames of rows x cols (otherwise done lazily by the first
 * extract call; a steady-state call allocates nothing). */
int os1_orb_reserve(os1_orb* h, int rows, int cols, int max_frames);

/* ORBextractor::operator()(image, mask, keypoints, descriptors)   src/ORBextractor.cc:907-969  (mask ignored,
 * include/ORBextractor.h:184).  HOST buffers; synchronous.  image: rows x cols CV_8UC1 with `step` bytes per
 * row.  kp: kp_cap records; desc: kp_cap x 32 bytes (row i <-> keypoint i); *n receives the count.
 * Empty image (rows or cols == 0, or image == NULL) returns OS1_OK and leaves the outputs untouched (:910-911). */
int os1_orb_extract(os1_orb* h, const uint8_t* image, int rows, int cols, size_t step, os1_keypoint* kp, int kp_cap,
                    uint8_t* desc, int* n);

/* The same on a batch of equally sized frames (frame f starts at frames + f*frame_stride).
 * frames_on_device / out_on_device say whether those pointers are device memory of the handle's GPU.
 * Outputs are slabs of cap_per_frame records per frame: kp[f*cap + i], desc[(f*cap + i)*32], counts[f].
 * With every buffer on the device the call only enqueues work on the handle's stream (see os1_orb_synchronize);
 * device-side failures are then reported by os1_orb_synchronize.
 * Host buffers are copied in chunks that overlap with the kernels: give it page-locked memory (cudaHostAlloc /
 * cudaHostRegister) — with pageable memory the copies are staged by the driver and nothing overlaps.  Slots beyond
 * counts[f] of a frame are not meaningful. */
int os1_orb_extract_batch(os1_orb* h, const uint8_t* frames, int nframes, int rows, int cols, size_t step, size_t frame_stride,
                          int frames_on_device, os1_keypoint* kp, uint8_t* desc, int cap_per_frame, int* counts, int out_on_device);
int os1_orb_synchronize(os1_orb* h);
/* Make the handle enqueue on a caller-owned cudaStream_t (NULL restores the handle's own stream). */
int os1_orb_set_stream(os1_orb* h, void* cuda_stream);
/* Number of kernels the handle's last extract call launched (for launch accounting): 4 — pyramid, FAST, quadtree, describe — per
 * chunk, or nlevels - 1 + 3 where a large device-resident batch builds its pyramid with one launch per level. */
int os1_orb_launches_per_call(const os1_orb* h);
/* Measurement aid (no reference counterpart): with timing enabled every extract call brackets its four kernels
 * (pyramid, FAST, quadtree, describe) with CUDA events on the handle's stream; os1_orb_get_stage_times waits for
 * the last call and returns the four durations in milliseconds. */
int os1_orb_set_timing(os1_orb* h, int enable);
int os1_orb_get_stage_times(os1_orb* h, float* ms4);

/* ---- stage accessors (parity tests; frame index refers to the last extract call) --------------------- */
/* public mvImagePyramid (include/ORBextractor.h:245), without the never-read 19-px border: w*h bytes, packed. */
int os1_orb_level_size(const os1_orb* h, int level, int* w, int* hh);
int os1_orb_get_pyramid_level(os1_orb* h, int frame, int level, uint8_t* out_host);
/* vToDistributeKeys of ComputeKeyPointsOctTree (src/ORBextractor.cc:816-870): n x {x, y, response}, coordinates
 * relative to (minBorderX, minBorderY), in the reference's order. */
int os1_orb_get_candidates(os1_orb* h, int frame, int level, int32_t* xys, int cap, int* n);
/* allKeypoints[level] after DistributeOctTree (:873-889): n x {x, y, response} in level pixels, list order. */
int os1_orb_get_level_keypoints(os1_orb* h, int frame, int level, int32_t* xys, int cap, int* n);

/* ---- Hamming consumer: replaces ORBmatcher::DescriptorDistance inside the matchers' scan loops ------------ */
/* static int ORBmatcher::DescriptorDistance(a, b)   src/ORBmatcher.cc:1605-1621 — scalar, host. */
int os1_descriptor_distance(const uint8_t* a, const uint8_t* b);

/* All-pairs best / second best (the scan of src/ORBmatcher.cc:431-452 with the whole of B as candidate list,
 * ascending index order, no claim bookkeeping): for every row i of A (nA x 32 bytes)
 *   best_idx[i] = first index of the minimum distance, best_dist[i] = that distance,
 *   second_dist[i] = second smallest distance VALUE (== best when the minimum occurs twice);
 * -1 / INT_MAX / INT_MAX when nB == 0.  on_device: all five pointers are device memory of `device`;
 * otherwise they are host memory and the call is synchronous.  cuda_stream may be NULL. */
int os1_hamming_knn2(const uint8_t* A, int nA, const uint8_t* B, int nB, int32_t* best_idx, int32_t* best_dist, int32_t* second_dist,
                     int on_device, int device, void* cuda_stream);
/* Windowed variant: row i is compared with B[cand[ptr[i] .. ptr[i+1])] in list order — the candidate lists
 * Frame::GetFeaturesInArea returns (src/Frame.cc:209-262) as used at src/ORBmatcher.cc:420, :1349, :73.
 * ptr has nA+1 entries.  best_idx is an index into B. */
int os1_hamming_window(const uint8_t* A, int nA, const uint8_t* B, int nB, const int32_t* ptr, const int32_t* cand, int32_t* best_idx,
                       int32_t* best_dist, int32_t* second_dist, int on_device, int device, void* cuda_stream);

/* MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:226-292) for npoints map points in one launch.
 * desc: the observation descriptors of all points back to back (32 bytes each); point m owns rows ptr[m] .. ptr[m+1]
 * in the order the reference iterates its observation map (:246-252).  For every point the N x N matrix of
 * DescriptorDistance values, the median of each row = element int(0.5*(N-1)) of the sorted row (:277-279), and
 *   best_idx[m]    = position (0 .. N-1, within the point's rows) of the FIRST row with the least median (:281-285),
 *                    -1 for a point without observations (the reference returns before touching mDescriptor, :240, :253);
 *   best_median[m] = that median (INT_MAX for an empty point); may be NULL.
 * ptr has npoints+1 entries.  A point with more than 256 observations is rejected (OS1_E_UNSUPPORTED).
 * on_device: all pointers are device memory of `device`; otherwise host memory.  Synchronous either way (the
 * size check of the device variant reads one status word back). */
int os1_distinctive_descriptors(const uint8_t* desc, const int32_t* ptr, int npoints, int32_t* best_idx, int32_t* best_median,
                                int on_device, int device, void* cuda_stream);

/* Many all-pairs scans in one launch — the cross-keyframe sweep of BASELINE.json configs[4].  DEVICE pointers.
 * desc: a slab of F descriptor sets, set f at desc + f*cap*32 with counts[f] valid rows; pairs: npairs x {i, j};
 * for every pair the rows of set i are scanned against the whole of set j exactly as os1_hamming_knn2 does.
 * best_idx / best_dist / second_dist: npairs x cap each (any of them may be NULL); accepted (may be NULL): per pair, the
 * number of rows with best <= th_low && best < nnratio * second (the acceptance test of src/ORBmatcher.cc:454-456).
 * Enqueues on cuda_stream and returns. */
int os1_hamming_knn2_pairs(const uint8_t* desc, const int32_t* counts, int cap, const int32_t* pairs, int npairs, int32_t* best_idx,
                           int32_t* best_dist, int32_t* second_dist, int32_t* accepted, float nnratio, int th_low, int device, void* cuda_stream);

/* Distance of every entry of the candidate lists: dist[k] = DescriptorDistance(A[i], B[cand[k]]) for ptr[i] <= k < ptr[i+1]
 * (src/ORBmatcher.cc:1605-1621 evaluated for the whole of a matcher's scan in one launch). */
int os1_hamming_csr(const uint8_t* A, int nA, const uint8_t* B, int nB, const int32_t* ptr, const int32_t* cand, uint16_t* dist,
                    int on_device, int device, void* cuda_stream);

/* The matching loops of the three consumers named in BASELINE.json, with their in-order bookkeeping.  HOST buffers.
 * Candidate lists are what Frame::GetFeaturesInArea returns per query (src/Frame.cc:209-262), concatenated CSR-style
 * in the reference's visiting order; a query the reference skips (`continue`) simply has an empty list.  All
 * distances are computed on the GPU in one launch; the ordered accept / claim pass runs on the host over them.
 *
 * ORBmatcher::SearchForInitialization, src/ORBmatcher.cc:400-515: ratio test with TH_LOW = 50, partners already
 * claimed with a smaller-or-equal distance are skipped (:439), a better later match steals the partner (:458-466),
 * rotation-histogram filter when both angle arrays are given (:468-505).  matches12[i1] = index in D2 or -1. */
int os1_search_for_initialization(const uint8_t* D1, int n1, const uint8_t* D2, int n2, const int32_t* ptr, const int32_t* cand,
                                  float nnratio, const float* angles1, const float* angles2, int32_t* matches12, int* nmatches, int device);
/* ORBmatcher::SearchByProjection(Frame&, const Frame&, th), src/ORBmatcher.cc:1292-1423 from the candidate lists on
 * (:1349): best only, accept <= TH_HIGH = 100 (:1379); taken[i2] != 0 <=> CurrentFrame.mvpMapPoints[i2] already holds
 * an observed point (:1363-1365) and is updated as matches are made (:1381); rotation filter when angles are given.
 * match_of_cand[i2] = query index or -1. */
int os1_search_by_projection_frame(const uint8_t* Dq, int nq, const uint8_t* Dc, int nc, const int32_t* ptr, const int32_t* cand,
                                   uint8_t* taken, const float* angles_q, const float* angles_c, int32_t* match_of_cand, int* nmatches, int device);
/* ORBmatcher::SearchByProjection(Frame&, const vector<MapPoint*>&, th), src/ORBmatcher.cc:45-124 from the candidate
 * lists on (:73): best and second best with the ratio applied only when both lie in the same pyramid level
 * (:113-114; octave_c = octave of every frame keypoint), accept <= TH_HIGH. */
int os1_search_by_projection_points(const uint8_t* Dq, int nq, const uint8_t* Dc, int nc, const int32_t* ptr, const int32_t* cand,
                                    uint8_t* taken, const int32_t* octave_c, float nnratio, int32_t* match_of_cand, int* nmatches, int device);

/* Frame::AssignFeaturesToGrid + Frame::GetFeaturesInArea for a batch of queries (src/Frame.cc:114-129, :209-274):
 * the 64 x 48 grid over [min_x, max_x) x [min_y, max_y) (Frame::ComputeImageBounds' mnMinX..mnMaxY), a keypoint's cell by
 * PosInGrid's round(), and for query q the keypoints with |dx| < r, |dy| < r whose octave passes the reference's
 * minLevel/maxLevel test (NULL arrays = -1 = no test), in the reference's visiting order (grid column, grid row, keypoint
 * index).  kp: the frame's (undistorted) keypoints.  Output CSR: ptr[nq+1], cand[ptr[nq]]; if cand_cap is too small the call
 * returns OS1_E_CAPACITY with the required size in ptr[nq] (host mode; device mode leaves ptr[nq] for the caller to check).
 * on_device: every pointer is device memory and the work is only enqueued.  At most 16384 keypoints. */
int os1_features_in_area(const os1_keypoint* kp, int n, float min_x, float max_x, float min_y, float max_y, const float* qx, const float* qy,
                         const float* qr, const int32_t* q_min_level, const int32_t* q_max_level, int nq, int32_t* ptr, int32_t* cand,
                         int cand_cap, int on_device, int device, void* cuda_stream);

/* Frame::UndistortKeyPoints (src/Frame.cc:286-320).  K4 = {fx, fy, cx, cy}.  mode 0 = pinhole: cv::undistortPoints(pts, pts, K,
 * dist, Mat(), K) with dist = k1,k2,p1,p2[,k3[,k4,k5,k6[,s1..s4]]] (ndist values), the identity when dist[0] == 0 (:288-292);
 * mode 1 = os1's equidistant fisheye without distortion (antidistorsionarProyeccionEquidistante, :355-384; dist ignored).
 * kp_un receives copies of the keypoints with pt replaced (kp_un may alias kp).
 * Parity of mode 1: the function body itself is the reference's (compiled from its source for the tests), but one product in it,
 * `Matx33f * Vec3d` (:376), is evaluated here in double with the float matrix entries widened — which overload real OpenCV picks
 * there cannot be checked without its headers, so that single step is an assumption (PARITY UNPINNED for it), and tan() is the
 * device's double-precision one (0 of 960 000 test keypoints differ from glibc's after the final rounding to float). */
int os1_undistort_keypoints(const os1_keypoint* kp, int n, const float* K4, const float* dist, int ndist, int mode, os1_keypoint* kp_un,
                            int on_device, int device, void* cuda_stream);
/* Frame::ComputeImageBounds (src/Frame.cc:322-353): bounds4 = {mnMinX, mnMaxX, mnMinY, mnMaxY}. */
int os1_image_bounds(int cols, int rows, const float* K4, const float* dist, int ndist, int mode, float* bounds4, int device);

/* ---- the three matching functions as WHOLE functions on the GPU ------------------------------------------------------------
 * An os1_matcher owns every scratch buffer the calls below need (grid, candidate lists, distances, claim state); it is sized
 * once, at creation, so a steady-state call allocates nothing.  The device entry points take DEVICE pointers — the keypoints and
 * descriptors os1_orb_extract_batch left on the GPU, undistorted by os1_undistort_keypoints — enqueue their kernels on
 * cuda_stream and return; results (matches and the match count) stay on the device.  Calls on one handle must be serialised
 * (one stream, or synchronised).  If the candidate lists of a call exceed max_candidates nothing is matched, the count is 0
 * and os1_matcher_synchronize reports OS1_E_CAPACITY.
 * Host-side quantities that the reference keeps in static members of Frame are plain arguments here: bounds4 = {mnMinX,
 * mnMaxX, mnMinY, mnMaxY} (Frame::ComputeImageBounds, os1_image_bounds), K4 = {fx, fy, cx, cy}. */
typedef struct os1_matcher os1_matcher;
#define OS1_MATCH_NONE (-1)     /* nothing attached / matched */
#define OS1_MATCH_REMOVED (-2)  /* attached during the call, detached again by the rotation filter (the reference leaves NULL) */
int os1_matcher_create(os1_matcher** out, int max_keypoints, int max_queries, int max_candidates, int device);
int os1_matcher_destroy(os1_matcher* m);
const char* os1_matcher_last_error(const os1_matcher* m);
/* Waits for cuda_stream and reports a device-side capacity failure of the calls queued before it. */
int os1_matcher_synchronize(os1_matcher* m, void* cuda_stream);

/* int ORBmatcher::SearchForInitialization(Frame& F1, Frame& F2, vector<Point2f>& vbPrevMatched, vector<int>& vnMatches12,
 * int windowSize)   src/ORBmatcher.cc:400-515, include/ORBmatcher.h:193 — candidate windows (GetFeaturesInArea around
 * vbPrevMatched, level 0 only), distances, claims and steals, rotation filter when check_orientation (mbCheckOrientation),
 * and the vbPrevMatched update (:507-510).  kp1/kp2 = mvKeysUn of the two frames, D1/D2 = mDescriptors, prev_xy = n1 x 2
 * floats in and out, matches12 = n1 ints (index into frame 2 or -1), nmatches = one int.  nnratio = mfNNratio. */
int os1_matcher_search_for_initialization(os1_matcher* m, const os1_keypoint* kp1, const uint8_t* D1, int n1, const os1_keypoint* kp2,
                                          const uint8_t* D2, int n2, const float* bounds4, float* prev_xy, int window_size, float nnratio,
                                          int check_orientation, int32_t* matches12, int32_t* nmatches, void* cuda_stream);
/* int ORBmatcher::SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, const float th)   src/ORBmatcher.cc:1292-1423,
 * include/ORBmatcher.h:118.  Tcw_cur = CurrentFrame.mTcw (16 floats, row major, host); per last-frame keypoint i: has_point[i]
 * (0 = mvpMapPoints[i] is NULL, 1 = a point with observations, 2 = a point whose Observations() is 0), outlier[i] (mvbOutlier[i]), world_pos = GetWorldPos() (3 floats), D_points = GetDescriptor()
 * (32 bytes), kp_last[i].octave / .angle.  taken[i2] (in/out, one byte per current keypoint): 0 = no map point, 1 = holds a point
 * with observations (:1363-1365 skips it), 2 = holds a point without; a keypoint that gets a point takes that point's 1 or 2, one
 * that loses it to the rotation filter becomes 0.  match_of_cur[i2] = index of the last-frame keypoint whose point was attached, OS1_MATCH_NONE
 * (untouched: it keeps what it had), or OS1_MATCH_REMOVED. */
int os1_matcher_search_by_projection_frame(os1_matcher* m, const os1_keypoint* kp_cur, const uint8_t* D_cur, int n_cur, const float* Tcw_cur,
                                           uint8_t* taken, const os1_keypoint* kp_last, int n_last, const uint8_t* has_point,
                                           const uint8_t* outlier, const float* world_pos, const uint8_t* D_points, const float* K4,
                                           const float* bounds4, const float* scale_factors, int nlevels, float th, int check_orientation,
                                           int32_t* match_of_cur, int32_t* nmatches, void* cuda_stream);
/* int ORBmatcher::SearchByProjection(Frame& F, const vector<MapPoint*>& vpMapPoints, const float th)   src/ORBmatcher.cc:45-124,
 * include/ORBmatcher.h:93.  map_points: n_points x 6 floats = mTrackProjX, mTrackProjY, mTrackViewCos, mnTrackScaleLevel, flags
 * (1 = mbTrackInView, 2 = isBad(), 4 = plCandidato), unused; D_points = GetDescriptor().  taken / match_of_frame as above
 * (no rotation filter in this function). */
int os1_matcher_search_by_projection_points(os1_matcher* m, const os1_keypoint* kp_frame, const uint8_t* D_frame, int n_frame, uint8_t* taken,
                                            const float* map_points, const uint8_t* D_points, int n_points, const float* bounds4,
                                            const float* scale_factors, int nlevels, float th, float nnratio, int32_t* match_of_frame,
                                            int32_t* nmatches, void* cuda_stream);
/* The same three calls with HOST buffers (used by the C++ adapter include/ORBmatcher.h): inputs are staged on the device in a
 * buffer the handle keeps, the pipeline runs, the results are copied back and the call returns when they have arrived. */
int os1_matcher_search_for_initialization_host(os1_matcher* m, const os1_keypoint* kp1, const uint8_t* D1, int n1, const os1_keypoint* kp2,
                                               const uint8_t* D2, int n2, const float* bounds4, float* prev_xy, int window_size, float nnratio,
                                               int check_orientation, int32_t* matches12, int32_t* nmatches);
int os1_matcher_search_by_projection_frame_host(os1_matcher* m, const os1_keypoint* kp_cur, const uint8_t* D_cur, int n_cur, const float* Tcw_cur,
                                                uint8_t* taken, const os1_keypoint* kp_last, int n_last, const uint8_t* has_point,
                                                const uint8_t* outlier, const float* world_pos, const uint8_t* D_points, const float* K4,
                                                const float* bounds4, const float* scale_factors, int nlevels, float th, int check_orientation,
                                                int32_t* match_of_cur, int32_t* nmatches);
int os1_matcher_search_by_projection_points_host(os1_matcher* m, const os1_keypoint* kp_frame, const uint8_t* D_frame, int n_frame, uint8_t* taken,
                                                 const float* map_points, const uint8_t* D_points, int n_points, const float* bounds4,
                                                 const float* scale_factors, int nlevels, float th, float nnratio, int32_t* match_of_frame,
                                                 int32_t* nmatches);

/* ---- wire records of os1's map files (host memory; no GPU work) ------------------------------------------------------
 * Osmap::serialize / deserialize of a descriptor (src/Osmap.cpp:775-786): every 1 x 32 CV_8U row is stored as eight 32-bit
 * blocks, `((unsigned int*)m.data)[i]` — the host's byte order.  desc: n x 32 bytes (the matrix os1_orb_extract returns),
 * blocks: n x 8.  DBoW2's per-descriptor rows (Converter::toDescriptorVector, src/Converter.cc:27-36) are the matrix rows
 * themselves and need no call. */
int os1_osmap_descriptor_blocks(const uint8_t* desc, int n, uint32_t* blocks);
int os1_osmap_descriptors_from_blocks(const uint32_t* blocks, int n, uint8_t* desc);
/* Osmap::serialize / deserialize of a keypoint (src/Osmap.cpp:819-831): the record keeps pt.x, pt.y, octave and angle; a keypoint
 * read back has size 0, response 0 and class_id -1 (cv::KeyPoint's defaults; angle and octave as stored). */
int os1_osmap_keypoint_records(const os1_keypoint* kp, int n, float* ptx, float* pty, int32_t* octave, float* angle);
int os1_osmap_keypoints_from_records(const float* ptx, const float* pty, const int32_t* octave, const float* angle, int n, os1_keypoint* kp);

/* os1_hamming_knn2 for callers that must not allocate.  This entry point has no counterpart in the reference
 * (os1 has no GPU code and so no allocation policy to mirror); it exists for the hosts this library gets embedded in.
 *
 * Who needs it: a tracking thread with a per-frame deadline, where a cudaMalloc / cudaFree (or a pool growing under
 * cudaMallocAsync) can take the driver's allocator lock or synchronise the device at an unpredictable moment; a
 * process whose device memory was divided up front between components, where a late allocation may simply fail;
 * a plug-in running inside a host application that forbids allocation on its worker threads.  os1_hamming_knn2
 * takes its scratch from a pool that grows with cudaMalloc whenever a problem is larger than any before it, or when a
 * call overlaps one on another stream, and is therefore not suitable for them.
 *
 * What it does differently: the scratch (one 8-byte partial result per row of A and per partition of B, plus 64 merge
 * tickets) is a fixed 1 MiB array inside the library's device image.  It is present on every GPU that loads the library, from load until
 * the process exits.  The call performs no cudaMalloc, cudaFree, cudaMallocAsync or any other allocating call.
 * A, B and the three outputs are DEVICE pointers on `device`, supplied by the caller; results are identical to
 * os1_hamming_knn2 with on_device = 1.  The work is enqueued on cuda_stream (may be NULL) and not waited for.
 *
 * Limits: nA <= 4096 and nB <= 65536; anything larger returns OS1_E_CAPACITY and launches nothing.
 * First use: the first call on a given device looks up the device address of the built-in array (a one-time
 * runtime query, made with that device current).  Make that first call during start-up, before any stream capture.
 * Sharing: all calls on one GPU use the same array, and it keeps whatever the previous call left in it.  The partial
 * results need no clearing between calls (each call overwrites every entry it later reads); the merge tickets must be
 * zero when a call starts, and every call leaves them at zero when its kernel ends.  Two calls whose kernels can overlap on
 * the same GPU — different streams, or different host threads — race on it and return garbage: serialise them. */
int os1_hamming_knn2_noalloc(const uint8_t* A, int nA, const uint8_t* B, int nB, int32_t* best_idx, int32_t* best_dist,
                             int32_t* second_dist, int device, void* cuda_stream);

#ifdef __cplusplus
}
#endif
#endif /* OS1_B200_H */
