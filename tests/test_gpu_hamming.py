"""GPU: batched Hamming consumer against the oracle's scan loops (bit-exact distances and indices)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from os1_b200 import matcher, synth  # noqa: E402


@pytest.mark.parametrize("nA,nB,seed", [(2000, 2000, 0), (1, 1, 1), (130, 257, 2), (777, 5000, 3), (4096, 300, 4), (50, 0, 5)])
def test_knn2_matches_oracle(oracle_lib, nA, nB, seed):
    rng = np.random.default_rng(seed)
    A = rng.integers(0, 256, (nA, 32), dtype=np.uint8); B = rng.integers(0, 256, (nB, 32), dtype=np.uint8)
    if nB > 12:
        B[7] = A[0]; B[11] = A[0]              # duplicated minimum: first index wins, second == best
        B[nB - 1] = A[nA - 1]
    got = matcher.hamming_knn2(A, B); ref = oracle_lib.Oracle.hamming_knn2(A, B)
    for g, r in zip(got, ref):
        assert np.array_equal(g, r)


def test_knn2_on_real_descriptors_and_properties(oracle_lib):
    """config 4: two frames of a synthetic video, 2000-feature extractor (the init extractor, Tracking.cc:67)."""
    import os1_b200
    fr = synth.video(2, 720, 1280, seed=9)
    e = os1_b200.ORBextractor(2000, 1.2, 8, 20, 7)
    _, d0 = e(fr[0]); _, d1 = e(fr[1])
    bi, bd, sd = matcher.hamming_knn2(d0, d1)
    rbi, rbd, rsd = oracle_lib.Oracle.hamming_knn2(d0, d1)
    assert np.array_equal(bi, rbi) and np.array_equal(bd, rbd) and np.array_equal(sd, rsd)
    assert (bd <= sd).all() and (bd >= 0).all() and (sd <= 256).all()
    # identity: a set against itself finds itself at distance 0
    si, sdist, _ = matcher.hamming_knn2(d0, d0)
    assert (sdist == 0).all()
    # symmetry of the distance: d(a,b) == d(b,a) through the kernel
    j = bi[:50]
    back = np.array([matcher.hamming_knn2(d1[jj:jj + 1], d0[i:i + 1])[1][0] for i, jj in enumerate(j)])
    assert np.array_equal(back, bd[:50])
    # the shifted scene really matches: most best distances are under TH_LOW with the 0.9 ratio (ORBmatcher.cc:454-456)
    good = (bd <= matcher.TH_LOW) & (bd < 0.9 * sd)
    assert good.mean() > 0.3


def test_window_matches_oracle(oracle_lib):
    rng = np.random.default_rng(7)
    A = rng.integers(0, 256, (500, 32), dtype=np.uint8); B = rng.integers(0, 256, (800, 32), dtype=np.uint8)
    lens = rng.integers(0, 90, 500); lens[3] = 0; lens[10] = 1
    ptr = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    cand = rng.integers(0, 800, ptr[-1]).astype(np.int32)
    B[cand[ptr[5]]] = A[5]
    got = matcher.hamming_window(A, B, ptr, cand); ref = oracle_lib.Oracle.hamming_window(A, B, ptr, cand)
    for g, r in zip(got, ref):
        assert np.array_equal(g, r)
    assert got[0][3] == -1


def test_pairs_batch_matches_per_pair_oracle(oracle_lib):
    """os1_hamming_knn2_pairs: several frame pairs in one launch == the single-pair scan on each, plus the accepted count."""
    import torch
    import os1_b200
    fr = synth.video(4, 480, 640, seed=3)
    e = os1_b200.ORBextractor(1000, 1.2, 8, 20, 7)
    kp, desc, cnt = e.extract_batch(fr)
    cap = desc.shape[1]
    pairs = np.array([[0, 1], [0, 2], [1, 3], [2, 3], [3, 0]], np.int32)
    d = torch.from_numpy(desc).cuda(); c = torch.from_numpy(cnt).cuda(); p = torch.from_numpy(pairs).cuda()
    bi = torch.full((len(pairs), cap), -9, dtype=torch.int32, device="cuda"); bd = torch.zeros_like(bi); sd = torch.zeros_like(bi)
    acc = torch.zeros(len(pairs), dtype=torch.int32, device="cuda")
    matcher.hamming_knn2_pairs_raw(d.data_ptr(), c.data_ptr(), cap, p.data_ptr(), len(pairs), bi.data_ptr(), bd.data_ptr(), sd.data_ptr(),
                                   acc.data_ptr(), 0.9, 50)
    torch.cuda.synchronize()
    for k, (i, j) in enumerate(pairs):
        ri, rd, rs = oracle_lib.Oracle.hamming_knn2(desc[i, :cnt[i]], desc[j, :cnt[j]])
        n = cnt[i]
        assert np.array_equal(bi[k, :n].cpu().numpy(), ri) and np.array_equal(bd[k, :n].cpu().numpy(), rd)
        assert np.array_equal(sd[k, :n].cpu().numpy(), rs)
        ok = (rd <= 50) & (rd.astype(np.float32) < rs.astype(np.float32) * np.float32(0.9))
        assert int(acc[k]) == int(ok.sum())
