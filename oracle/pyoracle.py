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
            L.orc_search_by_projection_frame.argtypes = [_vp, _i, _vp, _i, _vp, _vp, _vp, _vp, _vp, _vp]
            L.orc_search_by_projection_points.argtypes = [_vp, _i, _vp, _i, _vp, _vp, _vp, _vp, _f, _vp]
            L.orc_distinctive_descriptors.argtypes = [_vp, _vp, _i, _vp, _vp]
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
    def search_by_projection_frame(cls, Dq, Dc, ptr, cand, taken, angles_q=None, angles_c=None):
        Dq = np.ascontiguousarray(Dq, np.uint8); Dc = np.ascontiguousarray(Dc, np.uint8)
        ptr = np.ascontiguousarray(ptr, np.int32); cand = np.ascontiguousarray(cand, np.int32)
        aq = None if angles_q is None else np.ascontiguousarray(angles_q, np.float32)
        ac = None if angles_c is None else np.ascontiguousarray(angles_c, np.float32)
        m = np.zeros(max(len(Dc), 1), np.int32)
        n = cls.lib().orc_search_by_projection_frame(_p(Dq), len(Dq), _p(Dc), len(Dc), _p(ptr), _p(cand), _p(taken),
                                                     None if aq is None else _p(aq), None if ac is None else _p(ac), _p(m))
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
