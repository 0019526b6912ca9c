"""GPU: the three matcher loops named by the north star (SearchForInitialization, both SearchByProjection variants)
through the C ABI — distances on the GPU, ordered bookkeeping on the host — against the oracle's restatement of the
reference loops.  Candidate lists come from a restatement of Frame::GetFeaturesInArea's window test
(src/Frame.cc:248-255) on real extracted keypoints of two frames of a synthetic video."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from os1_b200 import matcher, synth  # noqa: E402


@pytest.fixture(scope="module")
def two_frames():
    import os1_b200
    fr = synth.video(2, 720, 1280, seed=5, max_shift=12)
    e = os1_b200.ORBextractor(2000, 1.2, 8, 20, 7)
    k0, d0 = e(fr[0]); k1, d1 = e(fr[1])
    return k0, d0, k1, d1


def windows(kq, kc, r, level_of_query=None, only_level0=False):
    """CSR lists: for query i, frame keypoints within |dx| < r and |dy| < r (Frame.cc:251-254), index order."""
    ptr = [0]; cand = []
    for i in range(len(kq)):
        if only_level0 and kq["octave"][i] > 0:
            ptr.append(len(cand)); continue
        m = (np.abs(kc["x"] - kq["x"][i]) < r) & (np.abs(kc["y"] - kq["y"][i]) < r)
        if only_level0:
            m &= kc["octave"] == 0
        elif level_of_query is not None:
            lv = level_of_query[i]
            m &= (kc["octave"] >= lv - 1) & (kc["octave"] <= lv + 1)
        cand.extend(np.nonzero(m)[0].tolist())
        ptr.append(len(cand))
    return np.array(ptr, np.int32), np.array(cand, np.int32)


def test_csr_distances(oracle_lib, two_frames):
    k0, d0, k1, d1 = two_frames
    ptr, cand = windows(k0[:300], k1, 60)
    dist = matcher.hamming_csr(d0[:300], d1, ptr, cand)
    rows = np.repeat(np.arange(300), np.diff(ptr))
    ref = np.unpackbits(d0[rows] ^ d1[cand], axis=1).sum(1)
    assert np.array_equal(dist, ref.astype(np.uint16))


@pytest.mark.parametrize("check_ori", [False, True])
def test_search_for_initialization(oracle_lib, two_frames, check_ori):
    k0, d0, k1, d1 = two_frames
    ptr, cand = windows(k0, k1, 100, only_level0=True)              # windowSize 100, level 0 only (ORBmatcher.cc:415-420)
    a = (k0["angle"], k1["angle"]) if check_ori else (None, None)
    n, m = matcher.SearchForInitialization(d0, d1, ptr, cand, 0.9, *a)
    rn, rm = oracle_lib.Oracle.search_for_initialization(d0, d1, ptr, cand, 0.9, *a)
    assert n == rn and np.array_equal(m, rm[:len(m)])
    assert n > 100 and n == (m >= 0).sum()
    got = m[m >= 0]
    assert len(np.unique(got)) == len(got)                           # one-to-one after the claim / steal bookkeeping


def test_search_for_initialization_claims_are_sequential(oracle_lib):
    """Hand-made case where the in-order rule matters: query 1 steals query 0's partner, then query 2 is refused it."""
    rng = np.random.default_rng(2)
    D2 = rng.integers(0, 256, (4, 32), dtype=np.uint8)
    D1 = np.zeros((3, 32), np.uint8)
    D1[0] = D2[1]; D1[0, 0] ^= 0b111
    D1[1] = D2[1]
    D1[2] = D2[1]; D1[2, 1] ^= 0b1
    ptr = np.array([0, 4, 8, 12]); cand = np.tile(np.arange(4), 3)
    n, m = matcher.SearchForInitialization(D1, D2, ptr, cand, 0.9)
    rn, rm = oracle_lib.Oracle.search_for_initialization(D1, D2, ptr, cand, 0.9)
    assert n == rn and np.array_equal(m, rm) and list(m) == [-1, 1, -1]


@pytest.mark.parametrize("check_ori", [False, True])
def test_search_by_projection_frame(oracle_lib, two_frames, check_ori):
    k0, d0, k1, d1 = two_frames
    q = np.arange(0, len(k0), 2)                                      # "map points" of the last frame
    ptr, cand = windows(k0[q], k1, 15 * 1.2, level_of_query=k0["octave"][q])
    rng = np.random.default_rng(1)
    taken0 = (rng.random(len(k1)) < 0.1).astype(np.uint8)
    a = (k0["angle"][q], k1["angle"]) if check_ori else (None, None)
    t1 = taken0.copy(); t2 = taken0.copy()
    n, m = matcher.SearchByProjectionFrame(d0[q], d1, ptr, cand, t1, *a)
    rn, rm = oracle_lib.Oracle.search_by_projection_frame(d0[q], d1, ptr, cand, t2, *a)
    assert n == rn and np.array_equal(m, rm) and np.array_equal(t1, t2)
    assert n > 50


def test_search_by_projection_points(oracle_lib, two_frames):
    k0, d0, k1, d1 = two_frames
    q = np.arange(1, len(k0), 3)
    lv = k0["octave"][q]
    ptr = [0]; cand = []
    for i, qi in enumerate(q):                                        # levels nPredictedLevel-1 .. nPredictedLevel (ORBmatcher.cc:73)
        m = (np.abs(k1["x"] - k0["x"][qi]) < 8) & (np.abs(k1["y"] - k0["y"][qi]) < 8) & (k1["octave"] >= lv[i] - 1) & (k1["octave"] <= lv[i])
        cand.extend(np.nonzero(m)[0].tolist()); ptr.append(len(cand))
    ptr = np.array(ptr, np.int32); cand = np.array(cand, np.int32)
    t1 = np.zeros(len(k1), np.uint8); t2 = t1.copy()
    n, m = matcher.SearchByProjectionPoints(d0[q], d1, ptr, cand, t1, k1["octave"], 0.8)
    rn, rm = oracle_lib.Oracle.search_by_projection_points(d0[q], d1, ptr, cand, t2, k1["octave"], 0.8)
    assert n == rn and np.array_equal(m, rm) and np.array_equal(t1, t2)
    assert n > 5
