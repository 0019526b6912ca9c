"""Developer check: os1_undistort_keypoints mode 1 (equidistant fisheye, Frame.cc:355-384) against the oracle (glibc tan) on 960 000 points."""
import numpy as np, sys, os
sys.path.insert(0, os.getcwd())
from os1_b200 import matcher
from os1_b200.orb import KEYPOINT_DTYPE
from oracle import pyoracle
rng = np.random.default_rng(0)
n = 16000
bad = 0; tot = 0; maxd = 0
for rep in range(60):
    kp = np.zeros(n, KEYPOINT_DTYPE)
    kp["x"] = rng.uniform(0, 1280, n).astype(np.float32); kp["y"] = rng.uniform(0, 720, n).astype(np.float32)
    K4 = (520.0 + rep, 522.0, 640.0, 360.0)
    un = matcher.UndistortKeyPoints(kp, K4, None, 1)
    ref = pyoracle.Oracle.undistort(np.stack([kp["x"], kp["y"]], 1), K4, None, 1)
    ne = (un["x"] != ref[:, 0]) | (un["y"] != ref[:, 1])
    bad += int(ne.sum()); tot += n
    if ne.any(): maxd = max(maxd, float(np.abs(un["x"] - ref[:, 0]).max()), float(np.abs(un["y"] - ref[:, 1]).max()))
print("fisheye: %d of %d keypoints differ from glibc path, max diff %g" % (bad, tot, maxd))
