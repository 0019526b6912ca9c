"""CPU: the matcher oracle's own properties (what the GPU Hamming kernels are later held to)."""
import numpy as np


def brute(A, B):
    d = np.unpackbits(A[:, None, :] ^ B[None, :, :], axis=2).sum(2).astype(np.int64)
    bi = d.argmin(1); bd = d.min(1)
    d2 = d.copy(); d2[np.arange(len(A)), bi] = 1 << 30
    return bi, bd, d2.min(1)


def test_knn2_oracle_against_numpy(oracle_lib):
    rng = np.random.default_rng(0)
    A = rng.integers(0, 256, (200, 32), dtype=np.uint8); B = rng.integers(0, 256, (300, 32), dtype=np.uint8)
    B[7] = A[3]; B[11] = A[3]                      # duplicated minimum: second == best, first index wins
    bi, bd, sd = oracle_lib.Oracle.hamming_knn2(A, B)
    nbi, nbd, nsd = brute(A, B)
    assert np.array_equal(bi, nbi) and np.array_equal(bd, nbd) and np.array_equal(sd, nsd)
    assert bi[3] == 7 and bd[3] == 0 and sd[3] == 0


def test_knn2_oracle_threads_agree(oracle_lib):
    rng = np.random.default_rng(1)
    A = rng.integers(0, 256, (333, 32), dtype=np.uint8); B = rng.integers(0, 256, (257, 32), dtype=np.uint8)
    a = oracle_lib.Oracle.hamming_knn2(A, B); b = oracle_lib.Oracle.hamming_knn2(A, B, threads=4)
    assert all(np.array_equal(x, y) for x, y in zip(a, b))


def test_search_for_initialization_claims(oracle_lib):
    """Sequential claim bookkeeping of ORBmatcher.cc:439,:458-466: a later, better match steals the partner."""
    rng = np.random.default_rng(2)
    D2 = rng.integers(0, 256, (4, 32), dtype=np.uint8)
    D1 = np.zeros((2, 32), np.uint8)
    D1[0] = D2[1]; D1[0, 0] ^= 0b111          # distance 3 to D2[1]
    D1[1] = D2[1]                             # distance 0 to D2[1]: steals it
    ptr = np.array([0, 4, 8]); cand = np.array([0, 1, 2, 3, 0, 1, 2, 3])
    n, m = oracle_lib.Oracle.search_for_initialization(D1, D2, ptr, cand, 0.9)
    assert n == 1 and list(m) == [-1, 1]
