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


# ---- include/ORBmatcher_b200.h: the three matcher functions behind the reference's member-function signatures ------------------
MEXE = os.path.join(ROOT, "tests", "matcher_adapter_main")


def build_matcher_adapter():
    from os1_b200 import build, _lib
    build.build()
    cmd = ["g++", "-O2", "-std=c++11", "-w", "-I" + os.path.join(ROOT, "include"), "-I" + os.path.join(ROOT, "oracle", "refmatch"),
           os.path.join(ROOT, "tests", "matcher_adapter_main.cpp"), "-o", MEXE, _lib.LIB_PATH, "-Wl,-rpath," + os.path.dirname(_lib.LIB_PATH), "-pthread"]
    subprocess.check_call(cmd)


def _pad8(b):
    return b + b"\0" * ((-len(b)) % 8)


def _case_bytes(kind, c):
    f32 = lambda a: _pad8(np.ascontiguousarray(a, np.float32).tobytes())
    u8 = lambda a: _pad8(np.ascontiguousarray(a, np.uint8).tobytes())
    hdr = np.zeros(8, np.int32); fl = np.zeros(8, np.float32)
    if kind == "init":
        hdr[:4] = [len(c["kp1"]), len(c["kp2"]), int(c["window"]), int(c["check_ori"])]; fl[0] = c["nnratio"]
        body = [f32(c["bounds"]), f32(c["kp1"]), u8(c["D1"]), f32(c["kp2"]), u8(c["D2"]), f32(c["prev"])]
    elif kind == "proj_frame":
        hdr[:4] = [len(c["kpc"]), len(c["kpl"]), int(c["check_ori"]), len(c["scale"])]; fl[0] = c["th"]
        body = [f32(c["bounds"]), f32(c["K4"]), f32(c["TcwC"]), f32(c["scale"]), f32(c["kpc"]), u8(c["Dc"]), u8(c["taken"]), f32(c["kpl"]), u8(c["has"]),
                u8(c["outlier"]), f32(c["wpos"]), u8(c["Dmp"])]
    else:
        hdr[:3] = [len(c["kpf"]), len(c["mp"]), len(c["scale"])]; fl[0] = c["th"]; fl[1] = c["nnratio"]
        body = [f32(c["bounds"]), f32(c["scale"]), f32(c["kpf"]), u8(c["Df"]), u8(c["taken"]), f32(c["mp"]), u8(c["Dmp"])]
    return _pad8(hdr.tobytes()) + _pad8(fl.tobytes()) + b"".join(body)


def test_matcher_adapter_compiles_and_links():
    build_matcher_adapter()
    assert os.path.exists(MEXE)


@pytest.mark.gpu
@pytest.mark.parametrize("kind", ["init", "proj_frame", "proj_points"])
def test_matcher_adapter_equals_reference_goldens(tmp_path, golden_dir, kind):
    """The member functions with the forwarding bodies, called on Frame / MapPoint objects as Tracking.cc does, leave in
    vnMatches12 / vbPrevMatched / mvpMapPoints exactly what the reference's own bodies leave (tests/golden/matcher_*.npz)."""
    from oracle import matcher_cases as mc
    build_matcher_adapter()
    g = np.load(os.path.join(golden_dir, "matcher_%s.npz" % kind))
    for s in g["seeds"]:
        c = mc.make_case(kind, int(s))
        inp = tmp_path / ("%s_%d.bin" % (kind, s)); out = tmp_path / ("%s_%d.out" % (kind, s))
        inp.write_bytes(_case_bytes(kind, c))
        subprocess.check_call([MEXE, kind, str(inp), str(out)])
        b = out.read_bytes()
        n = int(np.frombuffer(b[:4], np.int32)[0])
        assert n == int(g["n_%d" % s])
        nm = len(g["m_%d" % s])
        assert np.array_equal(np.frombuffer(b[4:4 + 4 * nm], np.int32), g["m_%d" % s])
        if kind == "init":
            prev = np.frombuffer(b[4 + 4 * nm:], np.float32).reshape(-1, 2)
            assert np.array_equal(prev, g["prev_%d" % s])
