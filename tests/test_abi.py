"""CPU: the C-ABI library loads and exports every symbol include/os1_b200.h declares (no compute without a GPU)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "os1_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(os1_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from os1_b200 import build, _lib
    build.build()
    lib = ctypes.CDLL(_lib.LIB_PATH)
    names = declared_symbols()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), "missing export: " + n
    # and the Python binding covers exactly the declared set
    assert sorted(s[0] for s in _lib.SYMBOLS) == names


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    import os1_b200
    with pytest.raises(os1_b200.Os1Error) as e:
        os1_b200.ORBextractor(1000, 1.2, 8, 20, 7)
    assert e.value.code == -2 and "no CPU fallback" in str(e.value)


def test_argument_validation_needs_no_gpu():
    import os1_b200
    L = os1_b200.load()
    h = ctypes.c_void_p()
    assert L.os1_orb_create(ctypes.byref(h), 0, 1.2, 8, 20, 7, 0) == -1          # nfeatures < 1
    assert L.os1_orb_create(ctypes.byref(h), 1000, 1.0, 8, 20, 7, 0) == -1        # scaleFactor must exceed 1
    assert L.os1_orb_create(ctypes.byref(h), 1000, 1.2, 40, 20, 7, 0) == -4       # too many levels
    assert L.os1_orb_create(ctypes.byref(h), 1000, 1.2, 8, 0, 7, 0) == -4         # threshold 0 unsupported
    assert b"threshold" in L.os1_last_error(None)


def test_descriptor_distance_scalar():
    import numpy as np
    from os1_b200 import matcher
    rng = np.random.default_rng(0)
    a = rng.integers(0, 256, 32, dtype=np.uint8); b = rng.integers(0, 256, 32, dtype=np.uint8)
    assert matcher.DescriptorDistance(a, b) == int(np.unpackbits(a ^ b).sum())
    assert matcher.DescriptorDistance(a, a) == 0


def test_product_never_touches_the_oracle():
    """The oracle is test infrastructure: nothing under os1_b200/ or include/ may import, link or call it."""
    bad = []
    for d in ("os1_b200", "include"):
        for dirpath, _, files in os.walk(os.path.join(ROOT, d)):
            if "build" in dirpath.split(os.sep) or "__pycache__" in dirpath:
                continue
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp")):
                    txt = open(os.path.join(dirpath, f), errors="ignore").read()
                    if re.search(r"pyoracle|liboracle|orb_oracle|cv_exact|from oracle|import oracle|oracle/", txt):
                        bad.append(os.path.join(dirpath, f))
    assert not bad, bad


def test_argument_validation_happens_before_any_cuda_call():
    """Bad arguments are refused with OS1_E_INVALID / OS1_E_UNSUPPORTED by the entry point itself — no device needed."""
    import ctypes as C
    import numpy as np
    from os1_b200 import _lib
    L = _lib.load()
    K4 = np.array([500, 500, 320, 240], np.float32); b = np.zeros(4, np.float32)
    # dist == NULL with ndist > 0 (used to be dereferenced), ndist out of range, unknown mode
    assert L.os1_image_bounds(640, 480, K4.ctypes.data, None, 4, 0, b.ctypes.data, 0) == _lib.OS1_E_INVALID
    assert L.os1_image_bounds(640, 480, K4.ctypes.data, None, 99, 0, b.ctypes.data, 0) == _lib.OS1_E_INVALID
    assert L.os1_image_bounds(640, 480, K4.ctypes.data, None, 0, 7, b.ctypes.data, 0) == _lib.OS1_E_INVALID
    # the identity shortcut needs no GPU either (Frame.cc:288-292 / :345-350)
    assert L.os1_image_bounds(640, 480, K4.ctypes.data, None, 0, 0, b.ctypes.data, 0) == _lib.OS1_OK and list(b) == [0, 640, 0, 480]
    h = C.c_void_p()
    assert L.os1_matcher_create(C.byref(h), 0, 10, 10, 0) == _lib.OS1_E_INVALID
    assert L.os1_matcher_create(C.byref(h), 20000, 10, 10, 0) == _lib.OS1_E_UNSUPPORTED
    assert L.os1_matcher_create(None, 100, 10, 10, 0) == _lib.OS1_E_INVALID
    assert L.os1_osmap_descriptor_blocks(None, 3, None) == _lib.OS1_E_INVALID
    assert L.os1_hamming_knn2_noalloc(None, 5000, None, 10, None, None, None, 0, None) == _lib.OS1_E_INVALID
    assert L.os1_orb_create(C.byref(h), 1000, C.c_float(1.2), 40, 20, 7, 0) == _lib.OS1_E_UNSUPPORTED       # more than 12 levels
