"""Host-side mirror of the reference's ORBextractor interface (include/ORBextractor.h:150-321 of os1) on top of the
C ABI.  Same constructor arguments, same getters, operator() -> __call__.  Keypoints come back as a structured
array with the binary layout of cv::KeyPoint; descriptors as an (n, 32) uint8 matrix (cv::Mat CV_8U), row i <->
keypoint i, levels concatenated in order as src/ORBextractor.cc:936-967 does."""
import ctypes as C
import numpy as np
from . import _lib

KEYPOINT_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
                           ("octave", "<i4"), ("class_id", "<i4")])
assert KEYPOINT_DTYPE.itemsize == 28


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


class ORBextractor:
    HARRIS_SCORE = 0
    FAST_SCORE = 1

    def __init__(self, nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, device=0):
        self._L = _lib.load()
        self._h = C.c_void_p()
        _lib.check(self._L.os1_orb_create(C.byref(self._h), int(nfeatures), float(scaleFactor), int(nlevels),
                                          int(iniThFAST), int(minThFAST), int(device)))
        self.device = device
        self.nfeatures = nfeatures
        self.nlevels = nlevels

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            self._L.os1_orb_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- getters (ORBextractor.h:194-239) — by value, as the reference returns them
    def GetLevels(self):
        return self._L.os1_orb_get_levels(self._h)

    def GetScaleFactor(self):
        return self._L.os1_orb_get_scale_factor(self._h)

    def _tables(self):
        n = self.nlevels
        t = [np.zeros(n, np.float32) for _ in range(4)]
        _lib.check(self._L.os1_orb_get_tables(self._h, *[_ptr(a) for a in t]), self._h)
        return t

    def GetScaleFactors(self):
        return self._tables()[0]

    def GetInverseScaleFactors(self):
        return self._tables()[1]

    def GetScaleSigmaSquares(self):
        return self._tables()[2]

    def GetInverseScaleSigmaSquares(self):
        return self._tables()[3]

    def features_per_level(self):
        q = np.zeros(self.nlevels, np.int32)
        _lib.check(self._L.os1_orb_get_features_per_level(self._h, _ptr(q)), self._h)
        return q

    def max_keypoints(self):
        return self._L.os1_orb_max_keypoints(self._h)

    # ---- operator() (ORBextractor.cc:907-969)
    def __call__(self, image, mask=None):
        """image: HxW uint8 (any row stride).  mask is ignored, as in the reference (ORBextractor.h:184).
        Returns (keypoints, descriptors); an empty image returns two empty arrays."""
        image = np.asarray(image)
        if image.size == 0:
            return np.zeros(0, KEYPOINT_DTYPE), np.zeros((0, 32), np.uint8)
        if image.dtype != np.uint8 or image.ndim != 2:
            raise TypeError("ORBextractor expects a CV_8UC1 image")     # assert(image.type() == CV_8UC1), :914
        if image.strides[1] != 1:
            image = np.ascontiguousarray(image)
        cap = self.max_keypoints()
        kp = np.zeros(cap, KEYPOINT_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        n = C.c_int(0)
        _lib.check(self._L.os1_orb_extract(self._h, _ptr(image), image.shape[0], image.shape[1], image.strides[0],
                                           _ptr(kp), cap, _ptr(desc), C.byref(n)), self._h)
        return kp[:n.value].copy(), desc[:n.value].copy()

    def extract_batch(self, frames):
        """frames: (B, H, W) uint8 host array.  Returns (kp [B, cap], desc [B, cap, 32], counts [B])."""
        frames = np.ascontiguousarray(frames, np.uint8)
        B, H, W = frames.shape
        cap = self.max_keypoints()
        kp = np.zeros((B, cap), KEYPOINT_DTYPE)
        desc = np.zeros((B, cap, 32), np.uint8)
        counts = np.zeros(B, np.int32)
        _lib.check(self._L.os1_orb_extract_batch(self._h, _ptr(frames), B, H, W, W, H * W, 0, _ptr(kp), _ptr(desc), cap,
                                                 _ptr(counts), 0), self._h)
        return kp, desc, counts

    def extract_batch_raw(self, frames_ptr, nframes, rows, cols, step, frame_stride, frames_on_device,
                          kp_ptr, desc_ptr, cap, counts_ptr, out_on_device):
        """Thin pass-through for callers that own the buffers (pinned host memory, torch tensors via data_ptr())."""
        _lib.check(self._L.os1_orb_extract_batch(self._h, frames_ptr, nframes, rows, cols, step, frame_stride,
                                                 int(frames_on_device), kp_ptr, desc_ptr, cap, counts_ptr,
                                                 int(out_on_device)), self._h)

    def reserve(self, rows, cols, max_frames):
        _lib.check(self._L.os1_orb_reserve(self._h, rows, cols, max_frames), self._h)

    def set_stream(self, cuda_stream):
        _lib.check(self._L.os1_orb_set_stream(self._h, cuda_stream), self._h)

    def synchronize(self):
        _lib.check(self._L.os1_orb_synchronize(self._h), self._h)

    def launches_per_call(self):
        return self._L.os1_orb_launches_per_call(self._h)

    def set_timing(self, enable=True):
        _lib.check(self._L.os1_orb_set_timing(self._h, int(enable)), self._h)

    def stage_times_ms(self):
        """Durations (ms) of the pyramid, FAST, quadtree and describe kernels of the last call."""
        t = np.zeros(4, np.float32)
        _lib.check(self._L.os1_orb_get_stage_times(self._h, _ptr(t)), self._h)
        return t

    # ---- stage accessors (parity tests)
    def level_size(self, level):
        w, h = C.c_int(), C.c_int()
        _lib.check(self._L.os1_orb_level_size(self._h, level, C.byref(w), C.byref(h)), self._h)
        return w.value, h.value

    def pyramid_level(self, level, frame=0):
        w, h = self.level_size(level)
        out = np.zeros((h, w), np.uint8)
        _lib.check(self._L.os1_orb_get_pyramid_level(self._h, frame, level, _ptr(out)), self._h)
        return out

    def candidates(self, level, frame=0):
        n = C.c_int()
        _lib.check(self._L.os1_orb_get_candidates(self._h, frame, level, None, 0, C.byref(n)), self._h)
        out = np.zeros((max(n.value, 1), 3), np.int32)
        _lib.check(self._L.os1_orb_get_candidates(self._h, frame, level, _ptr(out), n.value, C.byref(n)), self._h)
        return out[:n.value]

    def level_keypoints(self, level, frame=0):
        n = C.c_int()
        _lib.check(self._L.os1_orb_get_level_keypoints(self._h, frame, level, None, 0, C.byref(n)), self._h)
        out = np.zeros((max(n.value, 1), 3), np.int32)
        _lib.check(self._L.os1_orb_get_level_keypoints(self._h, frame, level, _ptr(out), n.value, C.byref(n)), self._h)
        return out[:n.value]
