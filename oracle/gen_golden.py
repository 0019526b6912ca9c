"""Generates tests/golden/*.npz — run in the build container, where cv2 4.13.0 and /root/reference exist.

  primitives.npz : small seeded inputs with the outputs of the REAL OpenCV primitives (cv2.resize INTER_LINEAR,
                   cv2.GaussianBlur 7x7 s=2 REFLECT_101, cv2 FAST 9_16 nonmax at 20 and 7, cv2.fastAtan2) — pins
                   oracle/cv_exact.h to third-party arithmetic that is not vendored by the reference.
  extract_*.npz  : keypoints + descriptors + per-level candidate counts produced by the UNMODIFIED reference
                   ORBextractor.cc (oracle/_ref/libos1_ref_canon.so, canonical tie rule) on seeded synthetic frames,
                   plus the count/difference seen with the stock glibc allocator build (libos1_ref.so).
  matcher_<kind>.npz : outputs of the reference's own matcher / Frame / MapPoint function bodies (oracle/_ref/libos1_refmatch.so:
                   ORBmatcher.cc, Frame.cc, MapPoint.cc sliced by path and compiled unmodified, oracle/refmatch/) on the seeded
                   scenarios of oracle/matcher_cases.py, with a digest of the scenario inputs.
  gemm in primitives.npz : cv2.gemm on 3x3 * 3x1 float matrices, the arithmetic behind `Rcw*x3Dw+tcw` (ORBmatcher.cc:1325).
The frames themselves are regenerated from the seed by os1_b200/synth.py, so the fixtures stay small.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from os1_b200 import synth  # noqa: E402
from oracle import pyoracle  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def primitives():
    import cv2
    cv2.setNumThreads(1)
    d = {"cv2_version": np.array(cv2.__version__)}
    img = synth.frame(240, 320, 11)
    d["img"] = img
    for k, (dw, dh) in enumerate([(267, 200), (222, 167), (149, 112)]):
        d["resize_%d" % k] = cv2.resize(img, (dw, dh), interpolation=cv2.INTER_LINEAR)
    d["blur"] = cv2.GaussianBlur(img, (7, 7), 2, 2, borderType=cv2.BORDER_REFLECT_101)
    for th in (20, 7):
        f = cv2.FastFeatureDetector_create(th, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
        d["fast_%d" % th] = np.array([(p.pt[0], p.pt[1], p.response) for p in f.detect(img)], np.int32).reshape(-1, 3)
    rng = np.random.default_rng(5)
    y = rng.integers(-3000000, 3000000, 4000).astype(np.float32)
    x = rng.integers(-3000000, 3000000, 4000).astype(np.float32)
    y[:4] = [0, -5, 0, 5]; x[:4] = [0, 0, -3, 5]
    d["atan_y"] = y; d["atan_x"] = x
    d["atan"] = np.array([cv2.fastAtan2(float(a), float(b)) for a, b in zip(y, x)], np.float32)
    # cv::undistortPoints(pts, pts, K, dist, Mat(), K) as Frame::UndistortKeyPoints calls it (Frame.cc:307)
    pts = np.stack([rng.uniform(0, 1280, 600), rng.uniform(0, 720, 600)], 1).astype(np.float32)
    d["und_pts"] = pts
    for name, K4, dist in [("a", (863.94, 865.22, 642.61, 356.83), (-0.3716, 0.2156, 0.0011, -0.0015)),
                           ("b", (500., 510., 320., 240.), (0.12, -0.25, 0.001, 0.0005, 0.05))]:
        K = np.array([[K4[0], 0, K4[2]], [0, K4[1], K4[3]], [0, 0, 1]], np.float32)
        d["und_K_" + name] = np.array(K4, np.float32); d["und_D_" + name] = np.array(dist, np.float32)
        d["und_out_" + name] = cv2.undistortPoints(pts.reshape(-1, 1, 2), K, np.array(dist, np.float32), None, K).reshape(-1, 2)
    # cv::gemm as the matrix expressions of ORBmatcher.cc:1304-1312,:1325 use it: A*B+C (flags 0) and -A.t()*B (GEMM_1_T)
    T = (rng.normal(size=(400, 4, 4)) * rng.choice([0.01, 1, 10, 100], (400, 1, 1))).astype(np.float32)
    X = (rng.normal(size=(400, 3, 1)) * rng.choice([0.1, 1, 50], (400, 1, 1))).astype(np.float32)
    d["gemm_T"] = T; d["gemm_x"] = X
    d["gemm_abc"] = np.stack([cv2.gemm(np.ascontiguousarray(t[:3, :3]), x, 1.0, np.ascontiguousarray(t[:3, 3:4]), 1.0) for t, x in zip(T, X)])
    d["gemm_at_b"] = np.stack([cv2.gemm(np.ascontiguousarray(t[:3, :3]), np.ascontiguousarray(t[:3, 3:4]), -1.0, None, 0.0, flags=cv2.GEMM_1_T) for t in T])
    np.savez_compressed(os.path.join(OUT, "primitives.npz"), **d)


def extraction():
    for name, (h, w, nf, seed) in {"vga": (480, 640, 1000, 0), "720p": (720, 1280, 1500, 1), "small": (200, 260, 300, 4)}.items():
        img = synth.frame(h, w, seed)
        rc = pyoracle.Reference(nf, canon=True)
        kp, desc = rc.extract(img, record=True)
        ncand = np.array([len(rc.candidates(l)) for l in range(8)], np.int32)
        rg = pyoracle.Reference(nf, canon=False)
        kg, _ = rg.extract(img)
        sc = set(map(tuple, kp[:, [0, 1, 5]].tolist())); sg = set(map(tuple, kg[:, [0, 1, 5]].tolist()))
        np.savez_compressed(os.path.join(OUT, "extract_%s.npz" % name), h=h, w=w, nfeatures=nf, seed=seed, kp=kp, desc=desc,
                            ncand=ncand, glibc_count=len(kg), glibc_only=len(sg - sc), canon_only=len(sc - sg),
                            tables_scale=rc.tables()["scale"], tables_quota=rc.tables()["quota"], tables_umax=rc.tables()["umax"])
        print(name, len(kp), "glibc build:", len(kg), "differs by", len(sg - sc), "/", len(sc - sg))


def case_digest(c):
    import hashlib
    h = hashlib.sha256()
    for k in sorted(c):
        if isinstance(c[k], np.ndarray):
            h.update(k.encode()); h.update(str(c[k].dtype).encode()); h.update(np.ascontiguousarray(c[k]).tobytes())
        else:
            h.update(("%s=%r" % (k, c[k])).encode())
    return h.hexdigest()


GOLDEN_SEEDS = {"init": [0, 1, 2, 5, 7, 9], "proj_frame": [0, 1, 2, 3, 5, 7], "proj_points": [0, 1, 2, 3, 4, 5], "area": [0, 1, 2, 3],
                "distinct": [0, 1, 5], "undistort": [0, 1, 2, 3, 4], "bounds": [0, 1, 2, 3, 4], "maxima": list(range(12)), "distance": [0], "wire": [0, 1]}


def matcher():
    from oracle import matcher_cases as mc
    for kind, seeds in GOLDEN_SEEDS.items():
        d = {"seeds": np.array(seeds, np.int32)}
        for s in seeds:
            c = mc.make_case(kind, s)
            d["digest_%d" % s] = np.array(case_digest(c))
            for k, v in mc.run_reference(kind, c).items():
                d["%s_%d" % (k, s)] = v
        np.savez_compressed(os.path.join(OUT, "matcher_%s.npz" % kind), **d)
        print("matcher_%s.npz: %d cases" % (kind, len(seeds)))


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    if len(sys.argv) > 1 and sys.argv[1] == "matcher":
        matcher()
        sys.exit(0)
    primitives()
    extraction()
    matcher()
