"""Random inputs: the Hamming kernels, the three matcher loops, the frame grid and the distinctive-descriptor kernel
against the oracle (everything bit for bit)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from os1_b200 import matcher
from oracle.pyoracle import Oracle

n = int(sys.argv[1]) if len(sys.argv) > 1 else 40
seed0 = int(sys.argv[2]) if len(sys.argv) > 2 else 0
bad = 0


def descs(rng, n, protos):
    """n descriptors: noisy copies of a few prototypes (so that ties and near-ties are common) or pure noise."""
    if n == 0:
        return np.zeros((0, 32), np.uint8)
    d = rng.integers(0, 256, (n, 32), dtype=np.uint8)
    if protos is not None:
        for i in range(n):
            if rng.random() < 0.7:
                row = protos[rng.integers(0, len(protos))].copy()
                for f in rng.integers(0, 256, rng.integers(0, 40)):
                    row[f >> 3] ^= 1 << (f & 7)
                d[i] = row
    return d


def csr(rng, nq, nc, maxlen):
    ptr = [0]; cand = []
    for i in range(nq):
        k = int(rng.integers(0, maxlen + 1)) if nc > 0 and rng.random() < 0.85 else 0
        cand.extend(rng.integers(0, nc, k).tolist() if k else [])
        ptr.append(len(cand))
    return np.array(ptr, np.int32), np.array(cand, np.int32)


def check(name, case, *pairs):
    global bad
    for a, b in pairs:
        if not np.array_equal(np.asarray(a), np.asarray(b)):
            bad += 1
            print("MISMATCH", name, "case", case)
            return


for k in range(n):
    case = seed0 + k
    rng = np.random.default_rng(case)
    protos = rng.integers(0, 256, (int(rng.integers(1, 30)), 32), dtype=np.uint8) if k % 4 else None
    nA = int(rng.integers(0, 2500)); nB = int(rng.integers(0, 2500))
    if k % 7 == 0: nA = int(rng.integers(0, 5))
    if k % 11 == 0: nB = int(rng.integers(0, 3))
    A = descs(rng, nA, protos); B = descs(rng, nB, protos)
    if nA and nB:
        check("knn2", case, *zip(matcher.hamming_knn2(A, B), Oracle.hamming_knn2(A, B)))
    ptr, cand = csr(rng, nA, nB, int(rng.integers(1, 60)))
    if nA:
        check("window", case, *zip(matcher.hamming_window(A, B, ptr, cand), Oracle.hamming_window(A, B, ptr, cand)))
        a1 = (rng.random(nA) * 360).astype(np.float32); a2 = (rng.random(max(nB, 1)) * 360).astype(np.float32)[:nB]
        for ang in (False, True):
            g = matcher.SearchForInitialization(A, B, ptr, cand, 0.9, a1 if ang else None, a2 if ang else None)
            o = Oracle.search_for_initialization(A, B, ptr, cand, 0.9, a1 if ang else None, a2 if ang else None)
            check("SearchForInitialization", case, (g[0], o[0]), (g[1], o[1]))
            tg = (rng.random(nB) < 0.2).astype(np.uint8); to = tg.copy()
            g = matcher.SearchByProjectionFrame(A, B, ptr, cand, tg, a1 if ang else None, a2 if ang else None)
            o = Oracle.search_by_projection_frame(A, B, ptr, cand, to, a1 if ang else None, a2 if ang else None)
            check("SearchByProjectionFrame", case, (g[0], o[0]), (g[1], o[1]), (tg, to))
        tg = (rng.random(nB) < 0.2).astype(np.uint8); to = tg.copy()
        oc = rng.integers(0, 8, nB).astype(np.int32)
        g = matcher.SearchByProjectionPoints(A, B, ptr, cand, tg, oc, 0.8)
        o = Oracle.search_by_projection_points(A, B, ptr, cand, to, oc, 0.8)
        check("SearchByProjectionPoints", case, (g[0], o[0]), (g[1], o[1]), (tg, to))
    # frame grid: keypoints in an image, random query windows and level ranges
    nk = int(rng.integers(0, 3000)); W, H = float(rng.integers(100, 2000)), float(rng.integers(100, 1200))
    kp = np.zeros(nk, [('x', '<f4'), ('y', '<f4'), ('size', '<f4'), ('angle', '<f4'), ('response', '<f4'), ('octave', '<i4'), ('class_id', '<i4')])
    kp["x"] = np.floor(rng.random(nk) * W); kp["y"] = np.floor(rng.random(nk) * H); kp["octave"] = rng.integers(0, 8, nk)
    nq = int(rng.integers(0, 400))
    qx = (rng.random(nq) * (W + 200) - 100).astype(np.float32); qy = (rng.random(nq) * (H + 200) - 100).astype(np.float32)
    qr = (rng.random(nq) * 120).astype(np.float32)
    lo = rng.integers(-1, 6, nq).astype(np.int32); hi = rng.integers(-1, 8, nq).astype(np.int32)
    bounds = (0.0, W, 0.0, H)
    use_lv = k % 2 == 0
    g = matcher.GetFeaturesInArea(kp, bounds, qx, qy, qr, lo if use_lv else None, hi if use_lv else None)
    o = Oracle.features_in_area(np.stack([kp["x"], kp["y"]], 1), kp["octave"], bounds, qx, qy, qr, lo if use_lv else None, hi if use_lv else None)
    check("GetFeaturesInArea", case, (g[0], o[0]), (g[1], o[1]))
    # distinctive descriptors
    M = int(rng.integers(1, 300)); cnt = rng.integers(0, int(rng.choice([4, 20, 80, 257])), M)
    p2 = np.concatenate([[0], np.cumsum(cnt)]).astype(np.int32)
    D = descs(rng, int(p2[-1]), protos)
    check("distinctive", case, *zip(matcher.distinctive_descriptors(D, p2), Oracle.distinctive_descriptors(D, p2)))
print("cases %d, mismatches %d" % (n, bad))
sys.exit(1 if bad else 0)
