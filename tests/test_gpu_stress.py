"""GPU: repeatability under changing co-residency — the stand-in for compute-sanitizer's racecheck, which is closed on this pool.

The kernels with hand-synchronised shared-memory reuse (describe: the horizontal blur pass overwrites the raw window in place under
__syncwarp; quadtree: the bitonic sort ping-pongs halves of one buffer; FAST: survivor lists reuse the image tile) would, if a
barrier were missing, depend on warp timing.  Timing is perturbed here the only way a test can: the same frames are extracted alone,
inside batches of different sizes and positions, on different streams and while another handle keeps the GPU busy, many times over —
every run must give byte-identical keypoints and descriptors, and they must equal the oracle's."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

import os1_b200  # noqa: E402
from os1_b200 import synth  # noqa: E402


def test_results_do_not_depend_on_timing(oracle_lib):
    import torch
    dev = torch.device("cuda", 0)
    H, W, NF = 480, 640, 1000
    vid = synth.video(24, H, W, seed=5)
    ext = os1_b200.ORBextractor(NF, 1.2, 8, 20, 7)
    other = os1_b200.ORBextractor(1500, 1.2, 8, 20, 7)
    cap = ext.max_keypoints()
    ko, do = oracle_lib.Oracle(NF).extract(vid[3])
    ref_kp, ref_desc = ext(vid[3])
    assert len(ref_kp) == len(ko) and np.array_equal(ref_desc, do) and np.array_equal(ref_kp["angle"], ko[:, 3])
    # a busy neighbour: a 720p batch running on another handle / stream while the small calls go through
    big = torch.from_numpy(synth.video(32, 720, 1280, seed=6)).to(dev)
    capo = other.max_keypoints(); other.reserve(720, 1280, 32)
    kb = torch.empty((32, capo, 28), dtype=torch.uint8, device=dev); db = torch.empty((32, capo, 32), dtype=torch.uint8, device=dev)
    cb = torch.zeros(32, dtype=torch.int32, device=dev)
    rng = np.random.default_rng(0)
    for rep in range(40):
        if rep % 2:
            other.extract_batch_raw(big.data_ptr(), 32, 720, 1280, 1280, 720 * 1280, True, kb.data_ptr(), db.data_ptr(), capo, cb.data_ptr(), True)
        n = int(rng.integers(1, 24)); pos = int(rng.integers(0, n))
        idx = rng.permutation(24)[:n]; idx[pos] = 3
        kp, desc, cnt = ext.extract_batch(vid[idx])
        assert cnt[pos] == len(ref_kp)
        assert np.array_equal(kp[pos, :cnt[pos]], ref_kp), "keypoints changed with batch composition (rep %d)" % rep
        assert np.array_equal(desc[pos, :cnt[pos]], ref_desc), "descriptors changed with batch composition (rep %d)" % rep
        k1, d1 = ext(vid[3])
        assert np.array_equal(k1, ref_kp) and np.array_equal(d1, ref_desc)
    other.synchronize()


def test_matchers_do_not_depend_on_timing(oracle_lib):
    from os1_b200 import matcher
    rng = np.random.default_rng(1)
    A = rng.integers(0, 256, (1999, 32), dtype=np.uint8); B = rng.integers(0, 256, (2003, 32), dtype=np.uint8)
    B[100:140] = A[50]                                # forty-fold tie: the first index must win every time
    ref = oracle_lib.Oracle.hamming_knn2(A, B)
    for rep in range(30):
        got = matcher.hamming_knn2(A, B)
        assert all(np.array_equal(g, r) for g, r in zip(got, ref)), "rep %d" % rep
