"""The header-only C++ adapter include/ORBextractor.h, used the way Frame.cc / Tracking.cc use the reference class.
CPU: it compiles and links against the C ABI (with the oracle's cv:: shim standing in for OpenCV's headers).
GPU: the compiled program's keypoints/descriptors equal the Python path's."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "tests", "adapter_main")


def build_adapter():
    from os1_b200 import build, _lib
    build.build()
    cmd = ["g++", "-O2", "-std=c++11", "-w", "-I" + os.path.join(ROOT, "include"), "-I" + os.path.join(ROOT, "oracle", "cvshim"),
           os.path.join(ROOT, "tests", "adapter_main.cpp"), "-o", EXE, _lib.LIB_PATH, "-Wl,-rpath," + os.path.dirname(_lib.LIB_PATH)]
    subprocess.check_call(cmd)


def test_adapter_compiles_and_links():
    build_adapter()
    assert os.path.exists(EXE)


@pytest.mark.gpu
def test_adapter_output_equals_python_path(tmp_path):
    import os1_b200
    from os1_b200 import synth
    build_adapter()
    img = synth.frame(480, 640, 17)
    raw = tmp_path / "img.raw"; out = tmp_path / "out.bin"
    img.tofile(raw)
    subprocess.check_call([EXE, str(raw), "480", "640", "1000", str(out)])
    b = out.read_bytes()
    n, nlev, untouched = np.frombuffer(b[:12], np.int32)
    assert nlev == 8 and untouched == 3
    kp = np.frombuffer(b[16:16 + n * 24], np.float32).reshape(n, 6)
    desc = np.frombuffer(b[16 + n * 24:], np.uint8).reshape(n, 32)
    k2, d2 = os1_b200.ORBextractor(1000, 1.2, 8, 20, 7)(img)
    assert n == len(k2)
    assert np.array_equal(kp[:, 0], k2["x"]) and np.array_equal(kp[:, 3], k2["angle"]) and np.array_equal(kp[:, 5], k2["octave"].astype(np.float32))
    assert np.array_equal(desc, d2)
