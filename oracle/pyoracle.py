"""TEST INFRASTRUCTURE — ctypes loaders for the CPU oracles.  NOT PRODUCT CODE.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
  Oracle      -> oracle/liboracle.so            our restatement (oracle/orb_oracle.cpp)
  Reference   -> oracle/_ref/libos1_ref*.so      the unmodified reference ORBextractor.cc (oracle/ref_wrapper.cpp)
"""
import ctypes as C
import os
import subprocess
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
_vp, _i, _f = C.c_void_p, C.c_int, C.c_float


def build(ref: bool = True) -> None:
    """Compile the checkers (g++ only).  _ref is built only where /root/reference exists."""
    subprocess.check_call(["make", "-s", "-C", HERE, "liboracle.so"])
    if ref and os.path.exists("/root/reference/src/ORBextractor.cc"):
        subprocess.check_call(["make", "-s", "-C", HERE, "ref"])


def _load(path):
    if not os.path.exists(path):
        raise FileNotFoundError(path)
    return C.CDLL(path)


def _p(a):
    return a.ctypes.data_as(_vp)


class Oracle:
    """Our CPU restatement.  extract() mirrors ORBextractor::operator() (ORBextractor.cc:907-969)."""

    _lib = None

    @classmethod
    def lib(cls):
        if cls._lib is None:
            path = os.path.join(HERE, "liboracle.so")
            if not os.path.exists(path):
                build(ref=False)
            L = _load(path)
            L.orc_create.restype = _vp
            L.orc_create.argtypes = [_i, _f, _i, _i, _i]
            L.orc_destroy.argtypes = [_vp]
            L.orc_tables.argtypes = [_vp] * 7
            L.orc_extract.argtypes = [_vp, _vp, _i, _i, _i, _vp, _i, _vp]
            L.orc_extract_batch.restype = C.c_long
            L.orc_extract_batch.argtypes = [_i, _f, _i, _i, _i, _vp, _i, _i, _i, _i, _vp]
            L.orc_level_size.argtypes = [_vp, _i, _vp, _vp]
            L.orc_level_pixels.argtypes = [_vp, _i, _vp, _i]
            L.orc_candidates.argtypes = [_vp, _i, _vp, _i]
            L.orc_level_keypoints.argtypes = [_vp, _i, _vp, _vp, _i]
            L.orc_distribute.argtypes = [_vp, _i, _i, _i, _i, _i, _i, _vp, _i]
            L.orc_resize.argtypes = [_vp, _i, _i, _i, _vp, _i, _i, _i]
            L.orc_blur.argtypes = [_vp, _i, _i, _i, _vp, _i]
            L.orc_fast.argtypes = [_vp, _i, _i, _i, _i, _i, _vp, _i]
            L.orc_fast_atan2.restype = _f
            L.orc_fast_atan2.argtypes = [_f, _f]
            L.orc_fast_atan2_v.argtypes = [_vp, _vp, _vp, _i]
            L.orc_ic_angle.restype = _f
            L.orc_ic_angle.argtypes = [_vp, _i, _i, _i]
            L.orc_descriptor.argtypes = [_vp, _i, _i, _i, _f, _vp]
            L.orc_descriptor_distance.argtypes = [_vp, _vp]
            L.orc_hamming_knn2.argtypes = [_vp, _i, _vp, _i, _vp, _vp, _vp]
            L.orc_hamming_knn2_mt.argtypes = [_vp, _i, _vp, _i, _vp, _vp, _vp, _i]
            L.orc_hamming_window.argtypes = [_vp, _i, _vp, _vp, _vp, _vp, _vp, _vp]
            L.orc_search_for_initialization.argtypes = [_vp, _i, _vp, _i, _vp, _vp, _f, _vp, _vp, _vp]
            L.orc_undistort.argtypes = [_vp, _i, _vp, _vp, _i, _i, _vp]
            L.orc_features_in_area.argtypes = [_vp, _vp, _i, _f, _f, _f, _f, _vp, _vp, _vp, _vp, _vp, _i, _vp, _vp, _i]
            L.orc_search_by_projection_frame.argtypes = [_vp, _i, _vp, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]
            L.orc_search_by_projection_points.argtypes = [_vp, _i, _vp, _i, _vp, _vp, _vp, _vp, _f, _vp]
            L.orc_distinctive_descriptors.argtypes = [_vp, _vp, _i, _vp, _vp]
            L.orc_project_last_frame.argtypes = [_vp, _vp, _vp, _vp, _vp, _i, _vp, _vp, _vp, _f, _vp, _vp, _vp, _vp, _vp, _vp]
            L.orc_point_queries.argtypes = [_vp, _i, _vp, _f, _vp, _vp, _vp, _vp, _vp, _vp]
            L.orc_three_maxima.argtypes = [_vp, _i, _vp]
            L.orc_gemm3_abc.argtypes = [_vp, _vp, _vp]
            L.orc_gemm3_atb.argtypes = [_vp, _vp]
            L.orc_image_bounds.argtypes = [_i, _i, _vp, _vp, _i, _i, _vp]
            cls._lib = L
        return cls._lib

    def __init__(self, nfeatures=1000, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7):
        self.L = self.lib()
        self.params = (nfeatures, scale_factor, nlevels, ini_th, min_th)
        self.nlevels = nlevels
        self.h = self.L.orc_create(nfeatures, scale_factor, nlevels, ini_th, min_th)

    def __del__(self):
        if getattr(self, "h", None):
            self.L.orc_destroy(self.h)
            self.h = None

    def tables(self):
        n = self.nlevels
        sc, inv, s2, is2 = (np.zeros(n, np.float32) for _ in range(4))
        quota = np.zeros(n, np.int32)
        umax = np.zeros(16, np.int32)
        self.L.orc_tables(self.h, _p(sc), _p(inv), _p(s2), _p(is2), _p(quota), _p(umax))
        return dict(scale=sc, inv=inv, sigma2=s2, invsigma2=is2, quota=quota, umax=umax)

    def extract(self, img, cap=None):
        img = np.ascontiguousarray(img, np.uint8)
        cap = cap or self.params[0] * 2 + 256
        kp = np.zeros((cap, 6), np.float32)
        desc = np.zeros((cap, 32), np.uint8)
        n = self.L.orc_extract(self.h, _p(img), img.shape[0], img.shape[1], img.strides[0], _p(kp), cap, _p(desc))
        assert n <= cap
        return kp[:n].copy(), desc[:n].copy()

    def level_size(self, level):
        w, h = C.c_int(), C.c_int()
        self.L.orc_level_size(self.h, level, C.byref(w), C.byref(h))
        return w.value, h.value

    def level_pixels(self, level, blurred=False):
        w, h = self.level_size(level)
        out = np.zeros((h, w), np.uint8)
        rc = self.L.orc_level_pixels(self.h, level, _p(out), int(blurred))
        return out if rc == 0 else None

    def candidates(self, level):
        n = self.L.orc_candidates(self.h, level, None, 0)
        out = np.zeros((max(n, 1), 3), np.int32)
        self.L.orc_candidates(self.h, level, _p(out), n)
        return out[:n]

    def level_keypoints(self, level):
        n = self.L.orc_level_keypoints(self.h, level, None, None, 0)
        xys = np.zeros((max(n, 1), 3), np.int32)
        ang = np.zeros(max(n, 1), np.float32)
        self.L.orc_level_keypoints(self.h, level, _p(xys), _p(ang), n)
        return xys[:n], ang[:n]

    # ---- stand-alone pieces
    @classmethod
    def distribute(cls, xys, min_x, max_x, min_y, max_y, N):
        xys = np.ascontiguousarray(xys, np.int32).reshape(-1, 3)
        out = np.zeros(max(len(xys), 1), np.int32)
        n = cls.lib().orc_distribute(_p(xys), len(xys), min_x, max_x, min_y, max_y, N, _p(out), len(out))
        return out[:n].copy()

    @classmethod
    def resize(cls, src, dw, dh):
        src = np.ascontiguousarray(src, np.uint8)
        dst = np.zeros((dh, dw), np.uint8)
        cls.lib().orc_resize(_p(src), src.shape[1], src.shape[0], src.strides[0], _p(dst), dw, dh, dw)
        return dst

    @classmethod
    def blur(cls, src):
        src = np.ascontiguousarray(src, np.uint8)
        dst = np.zeros_like(src)
        cls.lib().orc_blur(_p(src), src.shape[1], src.shape[0], src.strides[0], _p(dst), dst.strides[0])
        return dst

    @classmethod
    def fast(cls, img, threshold, nonmax=True):
        img = np.ascontiguousarray(img, np.uint8)
        cap = img.size // 2 + 16
        out = np.zeros((cap, 3), np.int32)
        n = cls.lib().orc_fast(_p(img), img.shape[1], img.shape[0], img.strides[0], threshold, int(nonmax), _p(out), cap)
        return out[:n].copy()

    @classmethod
    def fast_atan2(cls, y, x):
        y = np.ascontiguousarray(y, np.float32); x = np.ascontiguousarray(x, np.float32)
        out = np.zeros_like(y)
        cls.lib().orc_fast_atan2_v(_p(y), _p(x), _p(out), y.size)
        return out

    @classmethod
    def hamming_knn2(cls, A, B, threads=1):
        A = np.ascontiguousarray(A, np.uint8); B = np.ascontiguousarray(B, np.uint8)
        bi = np.zeros(len(A), np.int32); bd = np.zeros(len(A), np.int32); sd = np.zeros(len(A), np.int32)
        if threads > 1:
            cls.lib().orc_hamming_knn2_mt(_p(A), len(A), _p(B), len(B), _p(bi), _p(bd), _p(sd), threads)
        else:
            cls.lib().orc_hamming_knn2(_p(A), len(A), _p(B), len(B), _p(bi), _p(bd), _p(sd))
        return bi, bd, sd

    @classmethod
    def hamming_window(cls, A, B, ptr, cand):
        A = np.ascontiguousarray(A, np.uint8); B = np.ascontiguousarray(B, np.uint8)
        ptr = np.ascontiguousarray(ptr, np.int32); cand = np.ascontiguousarray(cand, np.int32)
        bi = np.zeros(len(A), np.int32); bd = np.zeros(len(A), np.int32); sd = np.zeros(len(A), np.int32)
        cls.lib().orc_hamming_window(_p(A), len(A), _p(B), _p(ptr), _p(cand), _p(bi), _p(bd), _p(sd))
        return bi, bd, sd

    @classmethod
    def search_for_initialization(cls, D1, D2, ptr, cand, nnratio=0.9, angles1=None, angles2=None):
        D1 = np.ascontiguousarray(D1, np.uint8); D2 = np.ascontiguousarray(D2, np.uint8)
        ptr = np.ascontiguousarray(ptr, np.int32); cand = np.ascontiguousarray(cand, np.int32)
        m = np.zeros(len(D1), np.int32)
        a1 = None if angles1 is None else np.ascontiguousarray(angles1, np.float32)
        a2 = None if angles2 is None else np.ascontiguousarray(angles2, np.float32)
        n = cls.lib().orc_search_for_initialization(_p(D1), len(D1), _p(D2), len(D2), _p(ptr), _p(cand), nnratio,
                                                     None if a1 is None else _p(a1), None if a2 is None else _p(a2), _p(m))
        return n, m

    @classmethod
    def undistort(cls, xy, K4, dist, mode):
        xy = np.ascontiguousarray(xy, np.float32).reshape(-1, 2)
        K4 = np.ascontiguousarray(K4, np.float32); dist = np.ascontiguousarray(dist if dist is not None else [], np.float32)
        out = np.zeros_like(xy)
        cls.lib().orc_undistort(_p(xy), len(xy), _p(K4), _p(dist) if len(dist) else None, len(dist), mode, _p(out))
        return out

    @classmethod
    def features_in_area(cls, xy, octave, bounds, qx, qy, qr, min_level=None, max_level=None):
        xy = np.ascontiguousarray(xy, np.float32).reshape(-1, 2); octave = np.ascontiguousarray(octave, np.int32)
        qx = np.ascontiguousarray(qx, np.float32); qy = np.ascontiguousarray(qy, np.float32); qr = np.ascontiguousarray(qr, np.float32)
        lo = None if min_level is None else np.ascontiguousarray(min_level, np.int32)
        hi = None if max_level is None else np.ascontiguousarray(max_level, np.int32)
        nq = len(qx)
        ptr = np.zeros(nq + 1, np.int32)
        cap = len(xy) * max(nq, 1) + 1
        cand = np.zeros(cap, np.int32)
        tot = cls.lib().orc_features_in_area(_p(xy), _p(octave), len(xy), bounds[0], bounds[1], bounds[2], bounds[3], _p(qx), _p(qy), _p(qr),
                                             None if lo is None else _p(lo), None if hi is None else _p(hi), nq, _p(ptr), _p(cand), cap)
        return ptr, cand[:tot].copy()

    @classmethod
    def distinctive_descriptors(cls, desc, ptr):
        """MapPoint::ComputeDistinctiveDescriptors (MapPoint.cc:226-292) for len(ptr)-1 map points: (best_idx, best_median)."""
        desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32); ptr = np.ascontiguousarray(ptr, np.int32)
        m = len(ptr) - 1
        bi = np.zeros(m, np.int32); bm = np.zeros(m, np.int32)
        cls.lib().orc_distinctive_descriptors(_p(desc), _p(ptr), m, _p(bi), _p(bm))
        return bi, bm

    @classmethod
    def search_by_projection_frame(cls, Dq, Dc, ptr, cand, taken, angles_q=None, angles_c=None, touched=None, q_noobs=None):
        Dq = np.ascontiguousarray(Dq, np.uint8); Dc = np.ascontiguousarray(Dc, np.uint8)
        ptr = np.ascontiguousarray(ptr, np.int32); cand = np.ascontiguousarray(cand, np.int32)
        aq = None if angles_q is None else np.ascontiguousarray(angles_q, np.float32)
        ac = None if angles_c is None else np.ascontiguousarray(angles_c, np.float32)
        m = np.zeros(max(len(Dc), 1), np.int32)
        n = cls.lib().orc_search_by_projection_frame(_p(Dq), len(Dq), _p(Dc), len(Dc), _p(ptr), _p(cand), _p(taken),
                                                     None if aq is None else _p(aq), None if ac is None else _p(ac), _p(m),
                                                     None if touched is None else _p(touched), None if q_noobs is None else _p(q_noobs))
        return n, m[:len(Dc)]

    @classmethod
    def search_by_projection_points(cls, Dq, Dc, ptr, cand, taken, octave_c, nnratio=0.8):
        Dq = np.ascontiguousarray(Dq, np.uint8); Dc = np.ascontiguousarray(Dc, np.uint8)
        ptr = np.ascontiguousarray(ptr, np.int32); cand = np.ascontiguousarray(cand, np.int32)
        octave_c = np.ascontiguousarray(octave_c, np.int32)
        m = np.zeros(max(len(Dc), 1), np.int32)
        n = cls.lib().orc_search_by_projection_points(_p(Dq), len(Dq), _p(Dc), len(Dc), _p(ptr), _p(cand), _p(taken), _p(octave_c),
                                                      nnratio, _p(m))
        return n, m[:len(Dc)]

    @classmethod
    def gemm3(cls, T44, x3=None):
        """R*x+t (x3 given) or -R^T*t of a 4x4 float pose, as cv::gemm evaluates them."""
        T44 = np.ascontiguousarray(T44, np.float32); out = np.zeros(3, np.float32)
        if x3 is None:
            cls.lib().orc_gemm3_atb(_p(T44), _p(out))
        else:
            x3 = np.ascontiguousarray(x3, np.float32)
            cls.lib().orc_gemm3_abc(_p(T44), _p(x3), _p(out))
        return out

    @classmethod
    def three_maxima(cls, sizes):
        sizes = np.ascontiguousarray(sizes, np.int32); out = np.zeros(3, np.int32)
        cls.lib().orc_three_maxima(_p(sizes), len(sizes), _p(out))
        return tuple(int(v) for v in out)

    @classmethod
    def image_bounds(cls, cols, rows, K4, dist, mode):
        K4 = np.ascontiguousarray(K4, np.float32); dist = np.ascontiguousarray(dist if dist is not None else [], np.float32)
        out = np.zeros(4, np.float32)
        cls.lib().orc_image_bounds(cols, rows, _p(K4), _p(dist) if len(dist) else None, len(dist), mode, _p(out))
        return out

    @classmethod
    def project_last_frame(cls, TcwC, wpos, has, outlier, octave, K4, bounds, scale_factors, th):
        """ORBmatcher.cc:1304-1349: the GetFeaturesInArea query of every last-frame keypoint -> qx, qy, qr, qmin, qmax, valid."""
        TcwC = np.ascontiguousarray(TcwC, np.float32); wpos = np.ascontiguousarray(wpos, np.float32).reshape(-1, 3)
        has = np.ascontiguousarray(has, np.uint8); outlier = np.ascontiguousarray(outlier, np.uint8); octave = np.ascontiguousarray(octave, np.int32)
        K4 = np.ascontiguousarray(K4, np.float32); b = np.ascontiguousarray(bounds, np.float32); sf = np.ascontiguousarray(scale_factors, np.float32)
        n = len(has)
        qx, qy, qr = (np.zeros(n, np.float32) for _ in range(3)); qmin = np.zeros(n, np.int32); qmax = np.zeros(n, np.int32)
        valid = np.zeros(n, np.uint8)
        cls.lib().orc_project_last_frame(_p(TcwC), _p(wpos), _p(has), _p(outlier), _p(octave), n, _p(K4), _p(b), _p(sf), th, _p(qx), _p(qy), _p(qr),
                                         _p(qmin), _p(qmax), _p(valid))
        return qx, qy, qr, qmin, qmax, valid.astype(bool)

    @classmethod
    def point_queries(cls, mp, scale_factors, th):
        """ORBmatcher.cc:51-71: the GetFeaturesInArea query of every map point (mp rows as in RefMatch.search_by_projection_points)."""
        mp = np.ascontiguousarray(mp, np.float32).reshape(-1, 6); sf = np.ascontiguousarray(scale_factors, np.float32)
        n = len(mp)
        qx, qy, qr = (np.zeros(n, np.float32) for _ in range(3)); qmin = np.zeros(n, np.int32); qmax = np.zeros(n, np.int32)
        valid = np.zeros(n, np.uint8)
        cls.lib().orc_point_queries(_p(mp), n, _p(sf), th, _p(qx), _p(qy), _p(qr), _p(qmin), _p(qmax), _p(valid))
        return qx, qy, qr, qmin, qmax, valid.astype(bool)

    @classmethod
    def csr_for_queries(cls, xy, octave, bounds, qx, qy, qr, qmin, qmax, valid):
        """Candidate lists of a query batch; the queries the reference skips get an empty list."""
        idx = np.flatnonzero(valid)
        p, c = cls.features_in_area(xy, octave, bounds, qx[idx], qy[idx], qr[idx], qmin[idx], qmax[idx])
        cnt = np.zeros(len(valid), np.int64); cnt[idx] = np.diff(p)
        ptr = np.concatenate([[0], np.cumsum(cnt)]).astype(np.int32)
        return ptr, c

    # ---- the three matcher functions as WHOLE functions, composed from the pieces above exactly as the product composes them
    @classmethod
    def search_for_initialization_full(cls, kp1, D1, kp2, D2, bounds, prev, window=100, nnratio=0.9, check_ori=True):
        """ORBmatcher::SearchForInitialization (ORBmatcher.cc:400-515) incl. candidate generation and the vbPrevMatched update."""
        kp1 = np.ascontiguousarray(kp1, np.float32).reshape(-1, 4); kp2 = np.ascontiguousarray(kp2, np.float32).reshape(-1, 4)
        prev = np.array(prev, np.float32).reshape(-1, 2).copy()
        n1 = len(kp1)
        valid = kp1[:, 3] <= 0                                              # :416-418
        z = np.zeros(n1, np.int32)
        ptr, cand = cls.csr_for_queries(kp2[:, :2], kp2[:, 3].astype(np.int32), bounds, prev[:, 0].copy(), prev[:, 1].copy(),
                                        np.full(n1, float(window), np.float32), z, z, valid)
        n, m = cls.search_for_initialization(D1, D2, ptr, cand, nnratio, kp1[:, 2] if check_ori else None, kp2[:, 2] if check_ori else None)
        ok = m >= 0
        prev[ok] = kp2[m[ok], :2]                                           # :507-510
        return n, m, prev

    @classmethod
    def search_by_projection_frame_full(cls, kpc, Dc, TcwC, taken, kpl, has, outlier, wpos, Dmp, K4, bounds, scale_factors, th, check_ori=True):
        """ORBmatcher::SearchByProjection(Frame&, const Frame&, th) (ORBmatcher.cc:1292-1423) as a whole.  taken: 0 / 1 / 2 as in RefMatch.
        Returns nmatches and, per current keypoint, the last-frame index attached, -1, or -2 (kept what it had)."""
        kpc = np.ascontiguousarray(kpc, np.float32).reshape(-1, 4); kpl = np.ascontiguousarray(kpl, np.float32).reshape(-1, 4)
        q = cls.project_last_frame(TcwC, wpos, has, outlier, kpl[:, 3].astype(np.int32), K4, bounds, scale_factors, th)
        ptr, cand = cls.csr_for_queries(kpc[:, :2], kpc[:, 3].astype(np.int32), bounds, *q)
        tk = (np.asarray(taken) == 1).astype(np.uint8)                      # a point without observations does not block (:1363-1365)
        touched = np.zeros(max(len(kpc), 1), np.uint8)
        noobs = (np.asarray(has) == 2).astype(np.uint8)
        n, m = cls.search_by_projection_frame(Dmp, Dc, ptr, cand, tk, kpl[:, 2] if check_ori else None, kpc[:, 2] if check_ori else None, touched, noobs)
        out = m.copy()
        keep = (np.asarray(taken) != 0) & (touched[:len(kpc)] == 0)
        out[keep] = -2
        return n, out

    @classmethod
    def search_by_projection_points_full(cls, kpf, Df, taken, mp, Dmp, bounds, scale_factors, th, nnratio=0.8):
        """ORBmatcher::SearchByProjection(Frame&, const vector<MapPoint*>&, th) (ORBmatcher.cc:45-124) as a whole."""
        kpf = np.ascontiguousarray(kpf, np.float32).reshape(-1, 4)
        q = cls.point_queries(mp, scale_factors, th)
        ptr, cand = cls.csr_for_queries(kpf[:, :2], kpf[:, 3].astype(np.int32), bounds, *q)
        tk = (np.asarray(taken) == 1).astype(np.uint8)
        n, m = cls.search_by_projection_points(Dmp, Df, ptr, cand, tk, kpf[:, 3].astype(np.int32), nnratio)
        out = m.copy()
        keep = (np.asarray(taken) != 0) & (m < 0)
        out[keep] = -2
        return n, out

    @classmethod
    def extract_batch(cls, frames, params, threads):
        frames = np.ascontiguousarray(frames, np.uint8)
        counts = np.zeros(len(frames), np.int32)
        tot = cls.lib().orc_extract_batch(params[0], params[1], params[2], params[3], params[4], _p(frames), len(frames),
                                          frames.shape[1], frames.shape[2], threads, _p(counts))
        return tot, counts


class Reference:
    """The unmodified reference extractor (compiled by path).  canon=True selects the bump-allocator build
    whose equal-size tie-break is the declared canonical rule; canon=False is stock glibc behaviour."""

    _libs = {}

    @classmethod
    def available(cls, canon=True):
        return os.path.exists(os.path.join(HERE, "_ref", "libos1_ref_canon.so" if canon else "libos1_ref.so"))

    @classmethod
    def lib(cls, canon):
        if canon not in cls._libs:
            L = _load(os.path.join(HERE, "_ref", "libos1_ref_canon.so" if canon else "libos1_ref.so"))
            L.ref_create.restype = _vp
            L.ref_create.argtypes = [_i, _f, _i, _i, _i]
            L.ref_destroy.argtypes = [_vp]
            L.ref_tables.argtypes = [_vp] * 7
            L.ref_extract.argtypes = [_vp, _vp, _i, _i, _i, _vp, _i, _vp, _i]
            L.ref_pyramid_size.argtypes = [_vp, _i, _vp, _vp]
            L.ref_pyramid_level.argtypes = [_vp, _i, _vp, _i]
            L.ref_candidate_count.argtypes = [_i]
            L.ref_candidates.argtypes = [_i, _vp]
            L.ref_distribute.argtypes = [_vp, _vp, _i, _i, _i, _i, _i, _i, _i, _vp, _i]
            L.ref_fast_area.argtypes = [_vp]
            cls._libs[canon] = L
        return cls._libs[canon]

    def __init__(self, nfeatures=1000, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7, canon=True):
        self.L = self.lib(canon)
        self.params = (nfeatures, scale_factor, nlevels, ini_th, min_th)
        self.nlevels = nlevels
        self.h = self.L.ref_create(nfeatures, scale_factor, nlevels, ini_th, min_th)

    def __del__(self):
        if getattr(self, "h", None):
            self.L.ref_destroy(self.h)
            self.h = None

    def tables(self):
        n = self.nlevels
        sc, inv, s2, is2 = (np.zeros(n, np.float32) for _ in range(4))
        quota = np.zeros(n, np.int32)
        umax = np.zeros(16, np.int32)
        self.L.ref_tables(self.h, _p(sc), _p(inv), _p(s2), _p(is2), _p(quota), _p(umax))
        return dict(scale=sc, inv=inv, sigma2=s2, invsigma2=is2, quota=quota, umax=umax)

    def extract(self, img, cap=None, record=False):
        img = np.ascontiguousarray(img, np.uint8)
        cap = cap or self.params[0] * 2 + 256
        kp = np.zeros((cap, 6), np.float32)
        desc = np.zeros((cap, 32), np.uint8)
        n = self.L.ref_extract(self.h, _p(img), img.shape[0], img.shape[1], img.strides[0], _p(kp), cap, _p(desc), int(record))
        assert n <= cap
        return kp[:n].copy(), desc[:n].copy()

    def fast_area(self):
        """Pixels the last extract(record=True) handed to cv::FAST at iniThFAST and at the minThFAST retry."""
        px = np.zeros(2, np.int64)
        self.L.ref_fast_area(_p(px))
        return int(px[0]), int(px[1])

    def level_pixels(self, level, border=False):
        w, h = C.c_int(), C.c_int()
        self.L.ref_pyramid_size(self.h, level, C.byref(w), C.byref(h))
        b = 19 if border else 0
        out = np.zeros((h.value + 2 * b, w.value + 2 * b), np.uint8)
        self.L.ref_pyramid_level(self.h, level, _p(out), int(border))
        return out

    def candidates(self, level):
        n = self.L.ref_candidate_count(level)
        out = np.zeros((max(n, 1), 3), np.int32)
        self.L.ref_candidates(level, _p(out))
        return out[:n]

    def distribute(self, xys, min_x, max_x, min_y, max_y, N, level=0):
        xys = np.ascontiguousarray(xys, np.int32).reshape(-1, 3)
        cap = len(xys) + 8
        out = np.zeros((cap, 3), np.int32)
        n = self.L.ref_distribute(self.h, _p(xys), len(xys), min_x, max_x, min_y, max_y, N, level, _p(out), cap)
        return out[:n].copy()


class RefMatch:
    """The reference's own matcher-side function bodies (ORBmatcher.cc / Frame.cc / MapPoint.cc), cut out of the reference tree
    by path and line range at build time and compiled unmodified between stub declarations (oracle/refmatch/).
    Keypoints are passed as n x 4 float rows: x, y, angle, octave."""

    _lib = None

    @classmethod
    def available(cls):
        return os.path.exists(os.path.join(HERE, "_ref", "libos1_refmatch.so"))

    @classmethod
    def lib(cls):
        if cls._lib is None:
            L = _load(os.path.join(HERE, "_ref", "libos1_refmatch.so"))
            L.rm_descriptor_distance.argtypes = [_vp, _vp]
            L.rm_three_maxima.argtypes = [_vp, _i, _vp]
            L.rm_gemm3_abc.argtypes = [_vp, _vp, _vp]
            L.rm_gemm3_atb.argtypes = [_vp, _vp]
            L.rm_features_in_area.argtypes = [_vp, _i, _vp, _vp, _vp, _vp, _vp, _vp, _i, _vp, _vp, _i]
            L.rm_search_for_initialization.argtypes = [_vp, _vp, _i, _vp, _vp, _i, _vp, _vp, _i, _f, _i, _vp]
            L.rm_search_by_projection_frame.argtypes = [_vp, _vp, _i, _vp, _vp, _vp, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _f, _i, _vp]
            L.rm_search_by_projection_points.argtypes = [_vp, _vp, _i, _vp, _vp, _vp, _i, _vp, _vp, _i, _f, _f, _vp]
            L.rm_undistort.argtypes = [_vp, _i, _vp, _vp, _i, _i, _vp]
            L.rm_image_bounds.argtypes = [_i, _i, _vp, _vp, _i, _i, _vp]
            L.rm_distinctive_descriptors.argtypes = [_vp, _vp, _vp, _i, _vp, _vp]
            L.rm_osmap_descriptor_blocks.argtypes = [_vp, _i, _vp, _vp]
            L.rm_osmap_keypoints.argtypes = [_vp, _i, _vp, _vp]
            cls._lib = L
        return cls._lib

    @staticmethod
    def _kp(kp):
        return np.ascontiguousarray(kp, np.float32).reshape(-1, 4)

    @classmethod
    def descriptor_distance(cls, a, b):
        a = np.ascontiguousarray(a, np.uint8); b = np.ascontiguousarray(b, np.uint8)
        return cls.lib().rm_descriptor_distance(_p(a), _p(b))

    @classmethod
    def gemm3(cls, T44, x3=None):
        T44 = np.ascontiguousarray(T44, np.float32); out = np.zeros(3, np.float32)
        if x3 is None:
            cls.lib().rm_gemm3_atb(_p(T44), _p(out))
        else:
            x3 = np.ascontiguousarray(x3, np.float32)
            cls.lib().rm_gemm3_abc(_p(T44), _p(x3), _p(out))
        return out

    @classmethod
    def three_maxima(cls, sizes):
        sizes = np.ascontiguousarray(sizes, np.int32); out = np.zeros(3, np.int32)
        cls.lib().rm_three_maxima(_p(sizes), len(sizes), _p(out))
        return tuple(int(v) for v in out)

    @classmethod
    def features_in_area(cls, kp, bounds, qx, qy, qr, min_level=None, max_level=None):
        kp = cls._kp(kp); b = np.ascontiguousarray(bounds, np.float32)
        qx = np.ascontiguousarray(qx, np.float32); qy = np.ascontiguousarray(qy, np.float32); qr = np.ascontiguousarray(qr, np.float32)
        lo = None if min_level is None else np.ascontiguousarray(min_level, np.int32)
        hi = None if max_level is None else np.ascontiguousarray(max_level, np.int32)
        nq = len(qx); cap = len(kp) * max(nq, 1) + 1
        ptr = np.zeros(nq + 1, np.int32); cand = np.zeros(cap, np.int32)
        tot = cls.lib().rm_features_in_area(_p(kp), len(kp), _p(b), _p(qx), _p(qy), _p(qr), None if lo is None else _p(lo),
                                            None if hi is None else _p(hi), nq, _p(ptr), _p(cand), cap)
        return ptr, cand[:tot].copy()

    @classmethod
    def search_for_initialization(cls, kp1, D1, kp2, D2, bounds, prev, window=100, nnratio=0.9, check_ori=True):
        kp1 = cls._kp(kp1); kp2 = cls._kp(kp2)
        D1 = np.ascontiguousarray(D1, np.uint8); D2 = np.ascontiguousarray(D2, np.uint8)
        b = np.ascontiguousarray(bounds, np.float32)
        prev = np.array(prev, np.float32).reshape(-1, 2).copy()
        m = np.zeros(max(len(kp1), 1), np.int32)
        n = cls.lib().rm_search_for_initialization(_p(kp1), _p(D1), len(kp1), _p(kp2), _p(D2), len(kp2), _p(b), _p(prev), int(window), nnratio,
                                                   int(check_ori), _p(m))
        return n, m[:len(kp1)], prev

    @classmethod
    def search_by_projection_frame(cls, kpc, Dc, TcwC, taken, kpl, TcwL, has, outlier, wpos, Dmp, K4, bounds, scale_factors, th, check_ori=True):
        kpc = cls._kp(kpc); kpl = cls._kp(kpl)
        Dc = np.ascontiguousarray(Dc, np.uint8); Dmp = np.ascontiguousarray(Dmp, np.uint8)
        TcwC = np.ascontiguousarray(TcwC, np.float32); TcwL = np.ascontiguousarray(TcwL, np.float32)
        taken = np.ascontiguousarray(taken, np.uint8); has = np.ascontiguousarray(has, np.uint8); outlier = np.ascontiguousarray(outlier, np.uint8)
        wpos = np.ascontiguousarray(wpos, np.float32); K4 = np.ascontiguousarray(K4, np.float32); b = np.ascontiguousarray(bounds, np.float32)
        sf = np.ascontiguousarray(scale_factors, np.float32)
        m = np.zeros(max(len(kpc), 1), np.int32)
        n = cls.lib().rm_search_by_projection_frame(_p(kpc), _p(Dc), len(kpc), _p(TcwC), _p(taken), _p(kpl), len(kpl), _p(TcwL), _p(has), _p(outlier),
                                                    _p(wpos), _p(Dmp), _p(K4), _p(b), _p(sf), len(sf), th, int(check_ori), _p(m))
        return n, m[:len(kpc)]

    @classmethod
    def search_by_projection_points(cls, kpf, Df, taken, mp, Dmp, bounds, scale_factors, th, nnratio=0.8):
        kpf = cls._kp(kpf); Df = np.ascontiguousarray(Df, np.uint8); taken = np.ascontiguousarray(taken, np.uint8)
        mp = np.ascontiguousarray(mp, np.float32).reshape(-1, 6); Dmp = np.ascontiguousarray(Dmp, np.uint8)
        b = np.ascontiguousarray(bounds, np.float32); sf = np.ascontiguousarray(scale_factors, np.float32)
        m = np.zeros(max(len(kpf), 1), np.int32)
        n = cls.lib().rm_search_by_projection_points(_p(kpf), _p(Df), len(kpf), _p(taken), _p(mp), _p(Dmp), len(mp), _p(b), _p(sf), len(sf), th, nnratio,
                                                     _p(m))
        return n, m[:len(kpf)]

    @classmethod
    def undistort(cls, xy, K4, dist, mode):
        xy = np.ascontiguousarray(xy, np.float32).reshape(-1, 2); K4 = np.ascontiguousarray(K4, np.float32)
        dist = np.ascontiguousarray(dist if dist is not None else [], np.float32)
        out = np.zeros_like(xy)
        cls.lib().rm_undistort(_p(xy), len(xy), _p(K4), _p(dist) if len(dist) else None, len(dist), mode, _p(out))
        return out

    @classmethod
    def image_bounds(cls, cols, rows, K4, dist, mode):
        K4 = np.ascontiguousarray(K4, np.float32); dist = np.ascontiguousarray(dist if dist is not None else [], np.float32)
        out = np.zeros(4, np.float32)
        cls.lib().rm_image_bounds(cols, rows, _p(K4), _p(dist) if len(dist) else None, len(dist), mode, _p(out))
        return out

    @classmethod
    def distinctive_descriptors(cls, desc, ptr, bad=None):
        desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32); ptr = np.ascontiguousarray(ptr, np.int32)
        bad = None if bad is None else np.ascontiguousarray(bad, np.uint8)
        m = len(ptr) - 1
        out = np.zeros((max(m, 1), 32), np.uint8); valid = np.zeros(max(m, 1), np.uint8)
        cls.lib().rm_distinctive_descriptors(_p(desc), _p(ptr), None if bad is None else _p(bad), m, _p(out), _p(valid))
        return out[:m], valid[:m].astype(bool)

    @classmethod
    def osmap_descriptor(cls, desc):
        """Osmap::serialize + deserialize of every row (Osmap.cpp:775-786): (blocks n x 8 uint32, rows read back)."""
        desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        blocks = np.zeros((len(desc), 8), np.uint32); back = np.zeros_like(desc)
        cls.lib().rm_osmap_descriptor_blocks(_p(desc), len(desc), _p(blocks), _p(back))
        return blocks, back

    @classmethod
    def osmap_keypoints(cls, kp):
        """Osmap::serialize + deserialize of every keypoint (Osmap.cpp:819-831): (n x {x, y, angle}, octave)."""
        kp = cls._kp(kp)
        xya = np.zeros((len(kp), 3), np.float32); octv = np.zeros(len(kp), np.int32)
        cls.lib().rm_osmap_keypoints(_p(kp), len(kp), _p(xya), _p(octv))
        return xya, octv
