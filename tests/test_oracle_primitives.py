"""CPU: oracle/cv_exact.h against the REAL OpenCV primitives (golden vectors from cv2 4.13.0; live cv2 too when importable)."""
import os

import numpy as np
import pytest


@pytest.fixture(scope="module")
def gold(golden_dir):
    return np.load(os.path.join(golden_dir, "primitives.npz"))


def test_resize_linear_bit_exact(oracle_lib, gold):
    img = gold["img"]
    for k in range(3):
        ref = gold["resize_%d" % k]
        out = oracle_lib.Oracle.resize(img, ref.shape[1], ref.shape[0])
        assert np.array_equal(out, ref)


def test_gaussian_blur_bit_exact(oracle_lib, gold):
    assert np.array_equal(oracle_lib.Oracle.blur(gold["img"]), gold["blur"])


@pytest.mark.parametrize("th", [20, 7])
def test_fast_bit_exact_including_order(oracle_lib, gold, th):
    assert np.array_equal(oracle_lib.Oracle.fast(gold["img"], th), gold["fast_%d" % th])


def test_fast_atan2_bit_exact(oracle_lib, gold):
    out = oracle_lib.Oracle.fast_atan2(gold["atan_y"], gold["atan_x"])
    assert np.array_equal(out, gold["atan"])
    # the survey's known answers
    ka = oracle_lib.Oracle.fast_atan2(np.array([0, -5, 0, 5], np.float32), np.array([0, 0, -3, 5], np.float32))
    assert ka[0] == 0 and ka[1] == 270 and ka[2] == 180 and abs(ka[3] - 44.990455627) < 1e-5


def test_primitives_against_live_cv2(oracle_lib):
    cv2 = pytest.importorskip("cv2")
    from os1_b200 import synth
    cv2.setNumThreads(1)
    for (h, w, seed) in [(480, 640, 0), (333, 444, 2), (134, 179, 3)]:
        img = synth.frame(h, w, seed)
        dw, dh = int(round(w / 1.2)), int(round(h / 1.2))
        assert np.array_equal(cv2.resize(img, (dw, dh), interpolation=cv2.INTER_LINEAR), oracle_lib.Oracle.resize(img, dw, dh))
        assert np.array_equal(cv2.GaussianBlur(img, (7, 7), 2, 2, borderType=cv2.BORDER_REFLECT_101), oracle_lib.Oracle.blur(img))
        sub = img[10:90, 20:140]
        for th in (20, 7):
            f = cv2.FastFeatureDetector_create(th, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
            a = np.array([(p.pt[0], p.pt[1], p.response) for p in f.detect(sub)], np.int32).reshape(-1, 3)
            assert np.array_equal(a, oracle_lib.Oracle.fast(sub, th))


def test_fast_small_and_flat_images(oracle_lib):
    assert len(oracle_lib.Oracle.fast(np.zeros((6, 40), np.uint8), 7)) == 0          # fewer than 7 rows: nothing is tested
    assert len(oracle_lib.Oracle.fast(np.full((40, 40), 128, np.uint8), 7)) == 0


def test_undistort_points_bit_exact(oracle_lib, gold):
    """Frame::UndistortKeyPoints' pinhole branch = cv::undistortPoints with P = K (Frame.cc:307)."""
    for name in ("a", "b"):
        out = oracle_lib.Oracle.undistort(gold["und_pts"], gold["und_K_" + name], gold["und_D_" + name], 0)
        assert np.array_equal(out, gold["und_out_" + name])
    same = oracle_lib.Oracle.undistort(gold["und_pts"], gold["und_K_a"], [0.0, 0.3, 0, 0], 0)
    assert np.array_equal(same, gold["und_pts"])          # dist[0] == 0 -> identity (Frame.cc:288-292)
