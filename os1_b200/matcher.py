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
