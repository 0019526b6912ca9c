"""Host-side mirror of os1's map-file records for descriptors and keypoints (src/Osmap.cpp:775-786, :819-831) on top of the C ABI."""
import ctypes as C
import numpy as np
from . import _lib
from .orb import KEYPOINT_DTYPE


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def descriptor_blocks(desc):
    """(n, 32) uint8 descriptors -> (n, 8) uint32 blocks, as Osmap::serialize stores them."""
    desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
    out = np.zeros((len(desc), 8), np.uint32)
    _lib.check(_lib.load().os1_osmap_descriptor_blocks(_ptr(desc), len(desc), _ptr(out)))
    return out


def descriptors_from_blocks(blocks):
    blocks = np.ascontiguousarray(blocks, np.uint32).reshape(-1, 8)
    out = np.zeros((len(blocks), 32), np.uint8)
    _lib.check(_lib.load().os1_osmap_descriptors_from_blocks(_ptr(blocks), len(blocks), _ptr(out)))
    return out


def keypoint_records(kp):
    """Structured keypoints -> (ptx, pty, octave, angle), the fields Osmap keeps."""
    kp = np.ascontiguousarray(kp, KEYPOINT_DTYPE)
    n = len(kp)
    x = np.zeros(n, np.float32); y = np.zeros(n, np.float32); o = np.zeros(n, np.int32); a = np.zeros(n, np.float32)
    _lib.check(_lib.load().os1_osmap_keypoint_records(_ptr(kp), n, _ptr(x), _ptr(y), _ptr(o), _ptr(a)))
    return x, y, o, a


def keypoints_from_records(x, y, o, a):
    x = np.ascontiguousarray(x, np.float32); y = np.ascontiguousarray(y, np.float32)
    o = np.ascontiguousarray(o, np.int32); a = np.ascontiguousarray(a, np.float32)
    kp = np.zeros(len(x), KEYPOINT_DTYPE)
    _lib.check(_lib.load().os1_osmap_keypoints_from_records(_ptr(x), _ptr(y), _ptr(o), _ptr(a), len(x), _ptr(kp)))
    return kp
