"""Host-side mirror of the Hamming consumers of os1's ORBmatcher (src/ORBmatcher.cc) on top of the C ABI."""
import ctypes as C
import numpy as np
from . import _lib

TH_LOW = 50      # ORBmatcher.cc:35
TH_HIGH = 100    # ORBmatcher.cc:36
HISTO_LENGTH = 30


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def DescriptorDistance(a, b):
    """static int ORBmatcher::DescriptorDistance(const cv::Mat&, const cv::Mat&)  (ORBmatcher.cc:1605-1621)."""
    a = np.ascontiguousarray(a, np.uint8).reshape(32)
    b = np.ascontiguousarray(b, np.uint8).reshape(32)
    return _lib.load().os1_descriptor_distance(_ptr(a), _ptr(b))


def hamming_knn2(A, B, device=0):
    """All-pairs best index / best distance / second-best distance (scan semantics of ORBmatcher.cc:431-452)."""
    A = np.ascontiguousarray(A, np.uint8).reshape(-1, 32)
    B = np.ascontiguousarray(B, np.uint8).reshape(-1, 32)
    bi = np.zeros(len(A), np.int32); bd = np.zeros(len(A), np.int32); sd = np.zeros(len(A), np.int32)
    _lib.check(_lib.load().os1_hamming_knn2(_ptr(A), len(A), _ptr(B), len(B), _ptr(bi), _ptr(bd), _ptr(sd), 0, device, None))
    return bi, bd, sd


def hamming_knn2_raw(a_ptr, nA, b_ptr, nB, bi_ptr, bd_ptr, sd_ptr, device=0, stream=None):
    """Device-pointer pass-through (all five buffers on `device`); enqueues on `stream`."""
    _lib.check(_lib.load().os1_hamming_knn2(a_ptr, nA, b_ptr, nB, bi_ptr, bd_ptr, sd_ptr, 1, device, stream))


def hamming_knn2_noalloc_raw(a_ptr, nA, b_ptr, nB, bi_ptr, bd_ptr, sd_ptr, device=0, stream=None):
    """os1_hamming_knn2_noalloc: the all-pairs matcher for callers that must not allocate device memory (a frame loop
    with a deadline, a fully pre-partitioned GPU, a host that forbids allocation on worker threads).  Device pointers
    only; scratch is a fixed 1 MiB array inside the library, so nA <= 4096 and nB <= 65536, and calls that may overlap
    on one GPU must be serialised by the caller.  The first call on a device should happen at start-up."""
    _lib.check(_lib.load().os1_hamming_knn2_noalloc(a_ptr, nA, b_ptr, nB, bi_ptr, bd_ptr, sd_ptr, device, stream))


def hamming_knn2_pairs_raw(desc_ptr, counts_ptr, cap, pairs_ptr, npairs, bi_ptr, bd_ptr, sd_ptr, accepted_ptr, nnratio=0.9, th_low=TH_LOW,
                           device=0, stream=None):
    """os1_hamming_knn2_pairs: many frame-pair scans in one launch (device pointers; NULL outputs are skipped)."""
    _lib.check(_lib.load().os1_hamming_knn2_pairs(desc_ptr, counts_ptr, cap, pairs_ptr, npairs, bi_ptr, bd_ptr, sd_ptr, accepted_ptr,
                                                  nnratio, th_low, device, stream))


def hamming_window(A, B, ptr, cand, device=0):
    """Best / second best over per-row candidate lists (CSR), e.g. the lists Frame::GetFeaturesInArea returns."""
    A = np.ascontiguousarray(A, np.uint8).reshape(-1, 32)
    B = np.ascontiguousarray(B, np.uint8).reshape(-1, 32)
    ptr = np.ascontiguousarray(ptr, np.int32); cand = np.ascontiguousarray(cand, np.int32)
    if len(ptr) != len(A) + 1:
        raise ValueError("ptr must have len(A)+1 entries")
    bi = np.zeros(len(A), np.int32); bd = np.zeros(len(A), np.int32); sd = np.zeros(len(A), np.int32)
    _lib.check(_lib.load().os1_hamming_window(_ptr(A), len(A), _ptr(B), len(B), _ptr(ptr), _ptr(cand), _ptr(bi), _ptr(bd),
                                              _ptr(sd), 0, device, None))
    return bi, bd, sd


def distinctive_descriptors(desc, ptr, device=0):
    """MapPoint::ComputeDistinctiveDescriptors (MapPoint.cc:226-292) for len(ptr)-1 map points whose observation
    descriptors are rows ptr[m]..ptr[m+1] of desc.  Returns (best_idx, best_median); best_idx is -1 for an empty point."""
    desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
    ptr = np.ascontiguousarray(ptr, np.int32)
    m = len(ptr) - 1
    if m < 0:
        raise ValueError("ptr must have at least one entry")
    bi = np.zeros(max(m, 1), np.int32); bm = np.zeros(max(m, 1), np.int32)
    _lib.check(_lib.load().os1_distinctive_descriptors(_ptr(desc) if len(desc) else None, _ptr(ptr), m, _ptr(bi), _ptr(bm), 0, device, None))
    return bi[:m], bm[:m]


def _csr(ptr, cand):
    return np.ascontiguousarray(ptr, np.int32), np.ascontiguousarray(cand, np.int32)


def hamming_csr(A, B, ptr, cand, device=0):
    """Distance of every candidate-list entry (uint16)."""
    A = np.ascontiguousarray(A, np.uint8).reshape(-1, 32); B = np.ascontiguousarray(B, np.uint8).reshape(-1, 32)
    ptr, cand = _csr(ptr, cand)
    dist = np.zeros(max(len(cand), 1), np.uint16)
    _lib.check(_lib.load().os1_hamming_csr(_ptr(A), len(A), _ptr(B), len(B), _ptr(ptr), _ptr(cand), _ptr(dist), 0, device, None))
    return dist[:len(cand)]


def SearchForInitialization(D1, D2, ptr, cand, nnratio=0.9, angles1=None, angles2=None, device=0):
    """Matching loop of ORBmatcher::SearchForInitialization (ORBmatcher.cc:400-515) over CSR candidate lists.
    Returns (nmatches, vnMatches12)."""
    D1 = np.ascontiguousarray(D1, np.uint8).reshape(-1, 32); D2 = np.ascontiguousarray(D2, np.uint8).reshape(-1, 32)
    ptr, cand = _csr(ptr, cand)
    a1 = None if angles1 is None else np.ascontiguousarray(angles1, np.float32)
    a2 = None if angles2 is None else np.ascontiguousarray(angles2, np.float32)
    m = np.zeros(max(len(D1), 1), np.int32); n = C.c_int()
    _lib.check(_lib.load().os1_search_for_initialization(_ptr(D1), len(D1), _ptr(D2), len(D2), _ptr(ptr), _ptr(cand), nnratio,
                                                          None if a1 is None else _ptr(a1), None if a2 is None else _ptr(a2), _ptr(m),
                                                          C.byref(n), device))
    return n.value, m[:len(D1)]


def SearchByProjectionFrame(Dq, Dc, ptr, cand, taken, angles_q=None, angles_c=None, device=0):
    """Matching loop of ORBmatcher::SearchByProjection(Frame&, const Frame&, th) (ORBmatcher.cc:1349-1419).
    `taken` (uint8 per current-frame keypoint) is updated in place.  Returns (nmatches, match_of_cand)."""
    Dq = np.ascontiguousarray(Dq, np.uint8).reshape(-1, 32); Dc = np.ascontiguousarray(Dc, np.uint8).reshape(-1, 32)
    ptr, cand = _csr(ptr, cand)
    assert taken.dtype == np.uint8 and taken.flags.c_contiguous and len(taken) == len(Dc)
    aq = None if angles_q is None else np.ascontiguousarray(angles_q, np.float32)
    ac = None if angles_c is None else np.ascontiguousarray(angles_c, np.float32)
    m = np.zeros(max(len(Dc), 1), np.int32); n = C.c_int()
    _lib.check(_lib.load().os1_search_by_projection_frame(_ptr(Dq), len(Dq), _ptr(Dc), len(Dc), _ptr(ptr), _ptr(cand), _ptr(taken),
                                                           None if aq is None else _ptr(aq), None if ac is None else _ptr(ac), _ptr(m),
                                                           C.byref(n), device))
    return n.value, m[:len(Dc)]


def SearchByProjectionPoints(Dq, Dc, ptr, cand, taken, octave_c, nnratio=0.8, device=0):
    """Matching loop of ORBmatcher::SearchByProjection(Frame&, const vector<MapPoint*>&, th) (ORBmatcher.cc:73-121)."""
    Dq = np.ascontiguousarray(Dq, np.uint8).reshape(-1, 32); Dc = np.ascontiguousarray(Dc, np.uint8).reshape(-1, 32)
    ptr, cand = _csr(ptr, cand)
    octave_c = np.ascontiguousarray(octave_c, np.int32)
    assert taken.dtype == np.uint8 and taken.flags.c_contiguous and len(taken) == len(Dc)
    m = np.zeros(max(len(Dc), 1), np.int32); n = C.c_int()
    _lib.check(_lib.load().os1_search_by_projection_points(_ptr(Dq), len(Dq), _ptr(Dc), len(Dc), _ptr(ptr), _ptr(cand), _ptr(taken),
                                                            _ptr(octave_c), nnratio, _ptr(m), C.byref(n), device))
    return n.value, m[:len(Dc)]


def GetFeaturesInArea(kp, bounds, qx, qy, qr, min_level=None, max_level=None, device=0):
    """Frame::GetFeaturesInArea (Frame.cc:209-262) for a batch of queries over the frame's 64x48 grid
    (AssignFeaturesToGrid, Frame.cc:114-129).  kp: structured keypoint array (KEYPOINT_DTYPE); bounds = (mnMinX, mnMaxX,
    mnMinY, mnMaxY).  Returns CSR (ptr, cand) in the reference's visiting order."""
    kp = np.ascontiguousarray(kp)
    qx = np.ascontiguousarray(qx, np.float32); qy = np.ascontiguousarray(qy, np.float32); qr = np.ascontiguousarray(qr, np.float32)
    nq = len(qx)
    lo = None if min_level is None else np.ascontiguousarray(min_level, np.int32)
    hi = None if max_level is None else np.ascontiguousarray(max_level, np.int32)
    ptr = np.zeros(nq + 1, np.int32)
    cap = max(64 * nq, 1024)
    L = _lib.load()
    while True:
        cand = np.zeros(cap, np.int32)
        rc = L.os1_features_in_area(_ptr(kp), len(kp), bounds[0], bounds[1], bounds[2], bounds[3], _ptr(qx), _ptr(qy), _ptr(qr),
                                    None if lo is None else _ptr(lo), None if hi is None else _ptr(hi), nq, _ptr(ptr), _ptr(cand), cap, 0,
                                    device, None)
        if rc == _lib.OS1_E_CAPACITY and ptr[nq] > cap:
            cap = int(ptr[nq])
            continue
        _lib.check(rc)
        return ptr, cand[:ptr[nq]].copy()


def UndistortKeyPoints(kp, K4, dist=None, mode=0, device=0):
    """Frame::UndistortKeyPoints (Frame.cc:286-320): mode 0 pinhole (cv::undistortPoints), mode 1 os1's equidistant fisheye."""
    kp = np.ascontiguousarray(kp)
    out = np.zeros_like(kp)
    K4 = np.ascontiguousarray(K4, np.float32)
    d = np.ascontiguousarray(dist if dist is not None else [], np.float32)
    _lib.check(_lib.load().os1_undistort_keypoints(_ptr(kp), len(kp), _ptr(K4), _ptr(d) if len(d) else None, len(d), mode, _ptr(out), 0,
                                                    device, None))
    return out


def ComputeImageBounds(cols, rows, K4, dist=None, mode=0, device=0):
    """Frame::ComputeImageBounds (Frame.cc:322-353) -> (mnMinX, mnMaxX, mnMinY, mnMaxY)."""
    K4 = np.ascontiguousarray(K4, np.float32)
    d = np.ascontiguousarray(dist if dist is not None else [], np.float32)
    b = np.zeros(4, np.float32)
    _lib.check(_lib.load().os1_image_bounds(cols, rows, _ptr(K4), _ptr(d) if len(d) else None, len(d), mode, _ptr(b), device))
    return tuple(float(v) for v in b)


MATCH_NONE, MATCH_REMOVED = -1, -2


class ORBmatcher:
    """Host-side mirror of os1's ORBmatcher for the three functions the north star names (include/ORBmatcher.h:93,:118,:193),
    run as whole functions on the GPU through an os1_matcher handle.  Constructor arguments as in the reference
    (`ORBmatcher(float nnratio=0.6, bool checkOri=true)`, ORBmatcher.h:71).  Frames are passed as plain arrays:
    keypoints = structured KEYPOINT_DTYPE arrays of the UNDISTORTED keypoints, descriptors (n, 32) uint8, bounds =
    (mnMinX, mnMaxX, mnMinY, mnMaxY)."""

    TH_LOW, TH_HIGH, HISTO_LENGTH = TH_LOW, TH_HIGH, HISTO_LENGTH

    def __init__(self, nnratio=0.6, checkOri=True, device=0, max_keypoints=16384, max_queries=16384, max_candidates=4 << 20):
        self.mfNNratio = float(nnratio); self.mbCheckOrientation = bool(checkOri); self.device = device
        self._L = _lib.load()
        self._h = C.c_void_p()
        _lib.check(self._L.os1_matcher_create(C.byref(self._h), max_keypoints, max_queries, max_candidates, device))

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            self._L.os1_matcher_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc != _lib.OS1_OK:
            msg = self._L.os1_matcher_last_error(self._h)
            raise _lib.Os1Error(rc, msg.decode() if msg else "")

    @property
    def handle(self):
        return self._h

    DescriptorDistance = staticmethod(DescriptorDistance)

    def SearchForInitialization(self, kp1, D1, kp2, D2, bounds, vbPrevMatched, windowSize=10):
        """ORBmatcher.cc:400-515.  Returns (nmatches, vnMatches12, vbPrevMatched updated)."""
        kp1 = np.ascontiguousarray(kp1); kp2 = np.ascontiguousarray(kp2)
        D1 = np.ascontiguousarray(D1, np.uint8).reshape(-1, 32); D2 = np.ascontiguousarray(D2, np.uint8).reshape(-1, 32)
        b = np.ascontiguousarray(bounds, np.float32)
        prev = np.array(vbPrevMatched, np.float32).reshape(-1, 2).copy()
        m = np.full(max(len(kp1), 1), -1, np.int32); n = C.c_int(0)
        self._check(self._L.os1_matcher_search_for_initialization_host(self._h, _ptr(kp1), _ptr(D1), len(kp1), _ptr(kp2), _ptr(D2), len(kp2), _ptr(b),
                                                                       _ptr(prev), int(windowSize), self.mfNNratio, int(self.mbCheckOrientation), _ptr(m),
                                                                       C.byref(n)))
        return n.value, m[:len(kp1)], prev

    def SearchByProjectionFrame(self, kp_cur, D_cur, Tcw_cur, taken, kp_last, has_point, outlier, world_pos, D_points, K4, bounds, scale_factors, th):
        """SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, th), ORBmatcher.cc:1292-1423.  `taken` (uint8 per current
        keypoint: 0 none, 1 point with observations, 2 point without) is updated in place.  Returns (nmatches, match_of_cur)."""
        kp_cur = np.ascontiguousarray(kp_cur); kp_last = np.ascontiguousarray(kp_last)
        D_cur = np.ascontiguousarray(D_cur, np.uint8).reshape(-1, 32); D_points = np.ascontiguousarray(D_points, np.uint8).reshape(-1, 32)
        T = np.ascontiguousarray(Tcw_cur, np.float32).reshape(16); K4 = np.ascontiguousarray(K4, np.float32); b = np.ascontiguousarray(bounds, np.float32)
        sf = np.ascontiguousarray(scale_factors, np.float32)
        has_point = np.ascontiguousarray(has_point, np.uint8); outlier = np.ascontiguousarray(outlier, np.uint8)
        world_pos = np.ascontiguousarray(world_pos, np.float32).reshape(-1, 3)
        assert taken.dtype == np.uint8 and taken.flags.c_contiguous and len(taken) == len(kp_cur)
        m = np.full(max(len(kp_cur), 1), -1, np.int32); n = C.c_int(0)
        self._check(self._L.os1_matcher_search_by_projection_frame_host(self._h, _ptr(kp_cur), _ptr(D_cur), len(kp_cur), _ptr(T), _ptr(taken), _ptr(kp_last),
                                                                        len(kp_last), _ptr(has_point), _ptr(outlier), _ptr(world_pos), _ptr(D_points), _ptr(K4),
                                                                        _ptr(b), _ptr(sf), len(sf), float(th), int(self.mbCheckOrientation), _ptr(m),
                                                                        C.byref(n)))
        return n.value, m[:len(kp_cur)]

    def SearchByProjectionPoints(self, kp_frame, D_frame, taken, map_points, D_points, bounds, scale_factors, th=3):
        """SearchByProjection(Frame& F, const vector<MapPoint*>&, th), ORBmatcher.cc:45-124.  map_points rows: mTrackProjX, mTrackProjY,
        mTrackViewCos, mnTrackScaleLevel, flags (1 mbTrackInView, 2 isBad, 4 plCandidato), unused."""
        kp_frame = np.ascontiguousarray(kp_frame)
        D_frame = np.ascontiguousarray(D_frame, np.uint8).reshape(-1, 32); D_points = np.ascontiguousarray(D_points, np.uint8).reshape(-1, 32)
        mp = np.ascontiguousarray(map_points, np.float32).reshape(-1, 6); b = np.ascontiguousarray(bounds, np.float32)
        sf = np.ascontiguousarray(scale_factors, np.float32)
        assert taken.dtype == np.uint8 and taken.flags.c_contiguous and len(taken) == len(kp_frame)
        m = np.full(max(len(kp_frame), 1), -1, np.int32); n = C.c_int(0)
        self._check(self._L.os1_matcher_search_by_projection_points_host(self._h, _ptr(kp_frame), _ptr(D_frame), len(kp_frame), _ptr(taken), _ptr(mp),
                                                                         _ptr(D_points), len(mp), _ptr(b), _ptr(sf), len(sf), float(th), self.mfNNratio,
                                                                         _ptr(m), C.byref(n)))
        return n.value, m[:len(kp_frame)]
