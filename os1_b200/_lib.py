"""ctypes binding of libos1_b200.so (include/os1_b200.h).  No fallback: if the library is missing or will not load
the import of os1_b200 fails loudly."""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libos1_b200.so")

OS1_OK, OS1_E_INVALID, OS1_E_CUDA, OS1_E_CAPACITY, OS1_E_UNSUPPORTED = 0, -1, -2, -3, -4

_vp, _i, _f, _sz = C.c_void_p, C.c_int, C.c_float, C.c_size_t

# every symbol include/os1_b200.h declares: (name, restype, argtypes)
SYMBOLS = [
    ("os1_orb_create", _i, [C.POINTER(_vp), _i, _f, _i, _i, _i, _i]),
    ("os1_orb_destroy", _i, [_vp]),
    ("os1_last_error", C.c_char_p, [_vp]),
    ("os1_orb_get_levels", _i, [_vp]),
    ("os1_orb_get_scale_factor", _f, [_vp]),
    ("os1_orb_get_tables", _i, [_vp, _vp, _vp, _vp, _vp]),
    ("os1_orb_get_features_per_level", _i, [_vp, _vp]),
    ("os1_orb_max_keypoints", _i, [_vp]),
    ("os1_orb_reserve", _i, [_vp, _i, _i, _i]),
    ("os1_orb_extract", _i, [_vp, _vp, _i, _i, _sz, _vp, _i, _vp, C.POINTER(_i)]),
    ("os1_orb_extract_batch", _i, [_vp, _vp, _i, _i, _i, _sz, _sz, _i, _vp, _vp, _i, _vp, _i]),
    ("os1_orb_synchronize", _i, [_vp]),
    ("os1_orb_set_stream", _i, [_vp, _vp]),
    ("os1_orb_launches_per_call", _i, [_vp]),
    ("os1_orb_set_timing", _i, [_vp, _i]),
    ("os1_orb_get_stage_times", _i, [_vp, _vp]),
    ("os1_orb_level_size", _i, [_vp, _i, C.POINTER(_i), C.POINTER(_i)]),
    ("os1_orb_get_pyramid_level", _i, [_vp, _i, _i, _vp]),
    ("os1_orb_get_candidates", _i, [_vp, _i, _i, _vp, _i, C.POINTER(_i)]),
    ("os1_orb_get_level_keypoints", _i, [_vp, _i, _i, _vp, _i, C.POINTER(_i)]),
    ("os1_descriptor_distance", _i, [_vp, _vp]),
    ("os1_hamming_knn2", _i, [_vp, _i, _vp, _i, _vp, _vp, _vp, _i, _i, _vp]),
    ("os1_hamming_knn2_noalloc", _i, [_vp, _i, _vp, _i, _vp, _vp, _vp, _i, _vp]),
    ("os1_hamming_knn2_pairs", _i, [_vp, _vp, _i, _vp, _i, _vp, _vp, _vp, _vp, _f, _i, _i, _vp]),
    ("os1_hamming_csr", _i, [_vp, _i, _vp, _i, _vp, _vp, _vp, _i, _i, _vp]),
    ("os1_search_for_initialization", _i, [_vp, _i, _vp, _i, _vp, _vp, _f, _vp, _vp, _vp, C.POINTER(_i), _i]),
    ("os1_search_by_projection_frame", _i, [_vp, _i, _vp, _i, _vp, _vp, _vp, _vp, _vp, _vp, C.POINTER(_i), _i]),
    ("os1_search_by_projection_points", _i, [_vp, _i, _vp, _i, _vp, _vp, _vp, _vp, _f, _vp, C.POINTER(_i), _i]),
    ("os1_features_in_area", _i, [_vp, _i, _f, _f, _f, _f, _vp, _vp, _vp, _vp, _vp, _i, _vp, _vp, _i, _i, _i, _vp]),
    ("os1_undistort_keypoints", _i, [_vp, _i, _vp, _vp, _i, _i, _vp, _i, _i, _vp]),
    ("os1_image_bounds", _i, [_i, _i, _vp, _vp, _i, _i, _vp, _i]),
    ("os1_hamming_window", _i, [_vp, _i, _vp, _i, _vp, _vp, _vp, _vp, _vp, _i, _i, _vp]),
    ("os1_distinctive_descriptors", _i, [_vp, _vp, _i, _vp, _vp, _i, _i, _vp]),
    ("os1_matcher_create", _i, [C.POINTER(_vp), _i, _i, _i, _i]),
    ("os1_matcher_destroy", _i, [_vp]),
    ("os1_matcher_last_error", C.c_char_p, [_vp]),
    ("os1_matcher_synchronize", _i, [_vp, _vp]),
    ("os1_matcher_search_for_initialization", _i, [_vp, _vp, _vp, _i, _vp, _vp, _i, _vp, _vp, _i, _f, _i, _vp, _vp, _vp]),
    ("os1_matcher_search_by_projection_frame", _i, [_vp, _vp, _vp, _i, _vp, _vp, _vp, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _f, _i, _vp, _vp, _vp]),
    ("os1_matcher_search_by_projection_points", _i, [_vp, _vp, _vp, _i, _vp, _vp, _vp, _i, _vp, _vp, _i, _f, _f, _vp, _vp, _vp]),
    ("os1_matcher_search_for_initialization_host", _i, [_vp, _vp, _vp, _i, _vp, _vp, _i, _vp, _vp, _i, _f, _i, _vp, _vp]),
    ("os1_matcher_search_by_projection_frame_host", _i, [_vp, _vp, _vp, _i, _vp, _vp, _vp, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _f, _i, _vp, _vp]),
    ("os1_matcher_search_by_projection_points_host", _i, [_vp, _vp, _vp, _i, _vp, _vp, _vp, _i, _vp, _vp, _i, _f, _f, _vp, _vp]),
    ("os1_osmap_descriptor_blocks", _i, [_vp, _i, _vp]),
    ("os1_osmap_descriptors_from_blocks", _i, [_vp, _i, _vp]),
    ("os1_osmap_keypoint_records", _i, [_vp, _i, _vp, _vp, _vp, _vp]),
    ("os1_osmap_keypoints_from_records", _i, [_vp, _vp, _vp, _vp, _i, _vp]),
]

_lib = None


def load():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                "os1_b200: %s is missing — build it with `python -m os1_b200.build` (nvcc, sm_100a). "
                "There is no CPU fallback." % LIB_PATH)
        lib = C.CDLL(LIB_PATH)
        for name, res, args in SYMBOLS:
            fn = getattr(lib, name)      # AttributeError if the library does not export a declared symbol
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


class Os1Error(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("os1_b200 error %d: %s" % (code, msg))
        self.code = code


def check(rc, handle=None):
    if rc != OS1_OK:
        msg = load().os1_last_error(handle)
        raise Os1Error(rc, msg.decode() if msg else "")
