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


def distinctive_numpy(desc, ptr):
    """MapPoint.cc:226-292 restated independently: sorted row, element int(0.5*(N-1)), first least median."""
    bi = np.full(len(ptr) - 1, -1, np.int32); bm = np.full(len(ptr) - 1, np.iinfo(np.int32).max, np.int32)
    for m in range(len(ptr) - 1):
        D = desc[ptr[m]:ptr[m + 1]]
        if len(D) == 0:
            continue
        d = np.unpackbits(D[:, None, :] ^ D[None, :, :], axis=2).sum(2)
        med = np.sort(d, axis=1)[:, int(0.5 * (len(D) - 1))]
        bi[m] = int(np.argmin(med)); bm[m] = int(med.min())
    return bi, bm


def distinctive_case(seed, npoints=120, max_obs=40):
    rng = np.random.default_rng(seed)
    cnt = rng.integers(0, max_obs + 1, npoints)
    k = min(4, npoints, max_obs + 1); cnt[:k] = np.arange(k)
    ptr = np.concatenate([[0], np.cumsum(cnt)]).astype(np.int32)
    desc = rng.integers(0, 256, (int(ptr[-1]), 32), dtype=np.uint8)
    # observations of one map point look alike: noisy copies of a prototype, some exact duplicates (ties in the medians)
    for m in range(npoints):
        if cnt[m] >= 2 and m % 3:
            proto = desc[ptr[m]].copy()
            for i in range(ptr[m], ptr[m + 1]):
                flips = rng.integers(0, 256, rng.integers(0, 30))
                row = proto.copy()
                for f in flips:
                    row[f >> 3] ^= 1 << (f & 7)
                desc[i] = row
            if cnt[m] >= 4:
                desc[ptr[m] + 2] = desc[ptr[m] + 1]
    return desc, ptr


def test_distinctive_descriptors_oracle_against_numpy(oracle_lib):
    for seed in (0, 1):
        desc, ptr = distinctive_case(seed)
        bi, bm = oracle_lib.Oracle.distinctive_descriptors(desc, ptr)
        nbi, nbm = distinctive_numpy(desc, ptr)
        assert np.array_equal(bi, nbi) and np.array_equal(bm, nbm)
    # the median index: N = 2 -> element 0 of the sorted row = the self distance 0, so the first row wins
    D = np.zeros((2, 32), np.uint8); D[1, 0] = 0xFF
    bi, bm = oracle_lib.Oracle.distinctive_descriptors(D, np.array([0, 2], np.int32))
    assert bi[0] == 0 and bm[0] == 0
    # N = 3: element 1 = the smaller of the two distances to the others
    D = np.zeros((3, 32), np.uint8); D[1, 0] = 0x0F; D[2, 0] = 0xFF          # d01 = 4, d02 = 8, d12 = 4
    bi, bm = oracle_lib.Oracle.distinctive_descriptors(D, np.array([0, 3], np.int32))
    assert bi[0] == 0 and bm[0] == 4                                          # rows 0 and 1 tie at 4: the first wins
