"""TEST INFRASTRUCTURE — seeded scenarios for the matcher / Frame / MapPoint functions, and the two ways of running them:

  run_reference(kind, case) -> the reference's own function bodies (oracle/_ref/libos1_refmatch.so, see oracle/refmatch/)
  run_oracle(kind, case)    -> the restated oracle (oracle/liboracle.so)

Used by tests/test_matcher_oracle.py (live comparison + committed goldens tests/golden/matcher_*.npz), by
oracle/gen_golden.py (which writes those goldens from the reference build) and by scripts/fuzz_matcher_vs_reference.py.
Every output is an integer / float array that must agree bit for bit.
"""
import numpy as np

from oracle import pyoracle as po

KINDS = ("init", "proj_frame", "proj_points", "area", "distinct", "undistort", "bounds", "maxima", "distance", "wire")
SCALE = np.array([1.0, 1.2000000477, 1.4400000572, 1.7280001640, 2.0736002922, 2.4883203506, 2.9859845638, 3.5831816196], np.float32)


def _keypoints(rng, n, w, h, lo=0.0, octave_p0=0.6, integer=False):
    x = rng.uniform(lo, w - lo, n); y = rng.uniform(lo, h - lo, n)
    octv = np.where(rng.random(n) < octave_p0, 0, rng.integers(0, 8, n))
    if integer:                                  # what the extractor really emits: integer level coordinates times the level scale
        x = (np.floor(x / SCALE[octv]).astype(np.float32) * SCALE[octv]); y = (np.floor(y / SCALE[octv]).astype(np.float32) * SCALE[octv])
    ang = rng.uniform(0, 360, n)
    return np.stack([x, y, ang, octv], 1).astype(np.float32)


def _noisy_copies(rng, D, idx, max_flips):
    out = D[idx].copy()
    for r in range(len(out)):
        k = int(rng.integers(0, max_flips + 1))
        bits = rng.integers(0, 256, k)
        for b in bits:
            out[r, b >> 3] ^= np.uint8(1 << (b & 7))
    return out


def _bounds(rng, w, h):
    if rng.random() < 0.5:
        return np.array([0, w, 0, h], np.float32)
    return np.array([-rng.uniform(0, 40), w + rng.uniform(0, 60), -rng.uniform(0, 30), h + rng.uniform(0, 45)], np.float32)


def make_case(kind, seed):
    rng = np.random.default_rng([seed, KINDS.index(kind)])
    w, h = [(640, 480), (1280, 720), (1920, 1080)][seed % 3]
    c = {"kind": kind, "seed": seed}
    if kind == "init":
        n2 = int(rng.integers(1, 2200)); n1 = int(rng.integers(1, 2200))
        if seed % 7 == 0:
            n1, n2 = int(rng.integers(1, 6)), int(rng.integers(1, 6))
        kp2 = _keypoints(rng, n2, w, h, integer=seed % 2 == 0)
        D2 = rng.integers(0, 256, (n2, 32), dtype=np.uint8)
        src = rng.integers(0, n2, n1)
        kp1 = kp2[src].copy()
        kp1[:, :2] += rng.normal(0, 25, (n1, 2)).astype(np.float32)
        kp1[:, 2] = np.mod(kp1[:, 2] + rng.choice([0, 0, 0, 7, 90], n1) + rng.normal(0, 3, n1), 360).astype(np.float32)
        kp1[:, 3] = np.where(rng.random(n1) < 0.7, 0, rng.integers(0, 8, n1))
        D1 = _noisy_copies(rng, D2, src, 70)
        rnd = rng.random(n1) < 0.3
        D1[rnd] = rng.integers(0, 256, (int(rnd.sum()), 32), dtype=np.uint8)
        prev = kp1[:, :2].copy()
        if seed % 3 == 1:
            prev += rng.normal(0, 40, prev.shape).astype(np.float32)
        c.update(kp1=kp1, D1=D1, kp2=kp2, D2=D2, prev=prev, bounds=_bounds(rng, w, h), window=int(rng.choice([100, 100, 30, 250])),
                 nnratio=float(rng.choice([0.9, 0.9, 0.7])), check_ori=int(seed % 4 != 3))
    elif kind == "proj_frame":
        nc = int(rng.integers(1, 2200)); nl = int(rng.integers(1, 2200))
        K4 = np.array([[520., 525., 320., 240.], [863.9, 865.2, 642.6, 356.8], [1379.2, 1378.6, 958.1, 546.9]][seed % 3], np.float32)
        kpc = _keypoints(rng, nc, w, h, octave_p0=0.3, integer=seed % 2 == 0)
        Dc = rng.integers(0, 256, (nc, 32), dtype=np.uint8)
        # current pose: small rotation + translation (rows of a 4x4, float)
        a = rng.normal(0, 0.03, 3)
        Rx = np.array([[1, 0, 0], [0, np.cos(a[0]), -np.sin(a[0])], [0, np.sin(a[0]), np.cos(a[0])]])
        Ry = np.array([[np.cos(a[1]), 0, np.sin(a[1])], [0, 1, 0], [-np.sin(a[1]), 0, np.cos(a[1])]])
        Rz = np.array([[np.cos(a[2]), -np.sin(a[2]), 0], [np.sin(a[2]), np.cos(a[2]), 0], [0, 0, 1]])
        T = np.eye(4); T[:3, :3] = Rx @ Ry @ Rz; T[:3, 3] = rng.normal(0, 0.1, 3)
        TcwC = T.astype(np.float32)
        TcwL = np.eye(4, dtype=np.float32)
        # world points: back-projections of (noisy) current keypoints, so that most projections land near one
        src = rng.integers(0, nc, nl)
        z = rng.uniform(0.5, 20, nl)
        z[rng.random(nl) < 0.05] *= -1                                    # behind the camera (:1332)
        uv = kpc[src, :2].astype(np.float64) + rng.normal(0, 6, (nl, 2))
        far = rng.random(nl) < 0.05
        uv[far] += rng.choice([-1, 1], (int(far.sum()), 2)) * rng.uniform(0.6, 1.5, (int(far.sum()), 2)) * [w, h]      # outside the bounds
        pc = np.stack([(uv[:, 0] - K4[2]) / K4[0] * z, (uv[:, 1] - K4[3]) / K4[1] * z, z], 1)
        Tinv = np.linalg.inv(TcwC.astype(np.float64))
        wpos = (pc @ Tinv[:3, :3].T + Tinv[:3, 3]).astype(np.float32)
        kpl = kpc[src].copy()
        kpl[:, 2] = np.mod(kpl[:, 2] + rng.choice([0, 0, 0, 12, 180], nl) + rng.normal(0, 3, nl), 360).astype(np.float32)
        kpl[:, 3] = np.clip(kpl[:, 3] + rng.integers(-1, 2, nl), 0, 7)
        Dmp = _noisy_copies(rng, Dc, src, 140)
        c.update(kpc=kpc, Dc=Dc, TcwC=TcwC, taken=rng.choice([0, 0, 0, 0, 1, 2], nc).astype(np.uint8), kpl=kpl, TcwL=TcwL,
                 has=((rng.random(nl) < 0.8) * np.where(rng.random(nl) < 0.1, 2, 1)).astype(np.uint8), outlier=(rng.random(nl) < 0.1).astype(np.uint8), wpos=wpos, Dmp=Dmp, K4=K4,
                 bounds=_bounds(rng, w, h), scale=SCALE, th=float(rng.choice([15, 30, 7])), check_ori=int(seed % 4 != 3))
    elif kind == "proj_points":
        nf = int(rng.integers(1, 2200)); nmp = int(rng.integers(1, 3000))
        kpf = _keypoints(rng, nf, w, h, octave_p0=0.3, integer=seed % 2 == 0)
        Df = rng.integers(0, 256, (nf, 32), dtype=np.uint8)
        src = rng.integers(0, nf, nmp)
        mp = np.zeros((nmp, 6), np.float32)
        mp[:, :2] = kpf[src, :2] + rng.normal(0, 3, (nmp, 2)).astype(np.float32)
        mp[:, 2] = rng.choice([0.5, 0.9, 0.998, 0.9981, 0.9999], nmp)
        mp[:, 3] = np.clip(kpf[src, 3] + rng.integers(0, 2, nmp), 0, 7)
        mp[:, 4] = (rng.random(nmp) < 0.9) * 1 + (rng.random(nmp) < 0.05) * 2 + (rng.random(nmp) < 0.1) * 4
        Dmp = _noisy_copies(rng, Df, src, 140)
        c.update(kpf=kpf, Df=Df, taken=rng.choice([0, 0, 0, 0, 1, 2], nf).astype(np.uint8), mp=mp, Dmp=Dmp, bounds=_bounds(rng, w, h), scale=SCALE,
                 th=float(rng.choice([1, 3, 5])), nnratio=float(rng.choice([0.8, 0.8, 0.6])))
    elif kind == "area":
        n = int(rng.integers(0, 3000)); nq = int(rng.integers(1, 400))
        kp = _keypoints(rng, n, w, h, lo=-30.0, octave_p0=0.3)
        qx = rng.uniform(-150, w + 150, nq).astype(np.float32); qy = rng.uniform(-150, h + 150, nq).astype(np.float32)
        qr = rng.choice([4, 15, 30, 100, 300], nq).astype(np.float32) * rng.uniform(0.5, 1.5, nq).astype(np.float32)
        qmin = rng.integers(-1, 7, nq).astype(np.int32); qmax = (qmin + rng.integers(-1, 3, nq)).astype(np.int32)
        c.update(kp=kp, bounds=_bounds(rng, w, h), qx=qx, qy=qy, qr=qr, qmin=qmin, qmax=qmax)
    elif kind == "distinct":
        npts = 150
        cnt = rng.integers(0, 41, npts); cnt[:4] = np.arange(4)
        if seed % 5 == 0:
            cnt[5] = 300
        ptr = np.concatenate([[0], np.cumsum(cnt)]).astype(np.int32)
        desc = np.zeros((int(ptr[-1]), 32), np.uint8)
        for m in range(npts):
            proto = rng.integers(0, 256, (1, 32), dtype=np.uint8)
            blk = _noisy_copies(rng, np.repeat(proto, cnt[m], 0), np.arange(cnt[m]), 60)
            if cnt[m] > 3 and rng.random() < 0.5:
                blk[rng.integers(0, cnt[m])] = blk[0]                      # exact duplicates: ties between medians
            desc[ptr[m]:ptr[m + 1]] = blk
        c.update(desc=desc, ptr=ptr, bad=(rng.random(len(desc)) < 0.1).astype(np.uint8))
    elif kind in ("undistort", "bounds"):
        n = 500
        xy = np.stack([rng.uniform(0, w, n), rng.uniform(0, h, n)], 1).astype(np.float32)
        K4 = np.array([w * rng.uniform(0.6, 0.8), w * rng.uniform(0.6, 0.8), w / 2 + rng.normal(0, 8), h / 2 + rng.normal(0, 8)], np.float32)
        mode = int(seed % 4 == 3)
        nd = [4, 5, 8, 0][seed % 4]
        dist = np.array([-0.37, 0.21, 0.0011, -0.0015, 0.05, 0.01, -0.02, 0.003][:nd], np.float32) * np.float32(rng.uniform(0.3, 1.0))
        if seed % 8 == 4:
            dist[0] = 0                                                   # the identity shortcut of Frame.cc:288-292
        c.update(xy=xy, K4=K4, dist=dist, mode=mode, cols=w, rows=h)
    elif kind == "maxima":
        sizes = rng.integers(0, [3, 30, 400][seed % 3], 30).astype(np.int32)
        if seed % 4 == 0:
            sizes[rng.integers(0, 30, 3)] = sizes.max()                   # ties
        c.update(sizes=sizes)
    elif kind == "distance":
        A = rng.integers(0, 256, (400, 32), dtype=np.uint8); B = rng.integers(0, 256, (400, 32), dtype=np.uint8)
        B[:50] = A[:50]; A[:8] = [[0] * 32, [255] * 32] * 4; B[:4] = 0; B[4:8] = 255
        c.update(A=A, B=B)
    elif kind == "wire":
        n = 300
        desc = rng.integers(0, 256, (n, 32), dtype=np.uint8)
        desc[0] = 0; desc[1] = 255; desc[2] = np.arange(32)
        c.update(desc=desc, kp=_keypoints(rng, n, w, h))
    else:
        raise ValueError(kind)
    return c


def run_reference(kind, c):
    R = po.RefMatch
    if kind == "init":
        n, m, prev = R.search_for_initialization(c["kp1"], c["D1"], c["kp2"], c["D2"], c["bounds"], c["prev"], int(c["window"]), float(c["nnratio"]),
                                                 bool(c["check_ori"]))
        return {"n": np.int32(n), "m": m, "prev": prev}
    if kind == "proj_frame":
        n, m = R.search_by_projection_frame(c["kpc"], c["Dc"], c["TcwC"], c["taken"], c["kpl"], c["TcwL"], c["has"], c["outlier"], c["wpos"], c["Dmp"],
                                            c["K4"], c["bounds"], c["scale"], float(c["th"]), bool(c["check_ori"]))
        return {"n": np.int32(n), "m": m}
    if kind == "proj_points":
        n, m = R.search_by_projection_points(c["kpf"], c["Df"], c["taken"], c["mp"], c["Dmp"], c["bounds"], c["scale"], float(c["th"]), float(c["nnratio"]))
        return {"n": np.int32(n), "m": m}
    if kind == "area":
        p, cd = R.features_in_area(c["kp"], c["bounds"], c["qx"], c["qy"], c["qr"], c["qmin"], c["qmax"])
        return {"ptr": p, "cand": cd}
    if kind == "distinct":
        keep = c["bad"] == 0
        d, valid = R.distinctive_descriptors(c["desc"], c["ptr"], c["bad"])
        return {"desc": d, "valid": valid}
    if kind == "undistort":
        return {"out": R.undistort(c["xy"], c["K4"], c["dist"], int(c["mode"]))}
    if kind == "bounds":
        return {"b": R.image_bounds(int(c["cols"]), int(c["rows"]), c["K4"], c["dist"], int(c["mode"]))}
    if kind == "maxima":
        return {"ind": np.array(R.three_maxima(c["sizes"]), np.int32)}
    if kind == "distance":
        return {"d": np.array([R.descriptor_distance(a, b) for a, b in zip(c["A"], c["B"])], np.int32)}
    if kind == "wire":
        blocks, back = R.osmap_descriptor(c["desc"])
        xya, octv = R.osmap_keypoints(c["kp"])
        return {"blocks": blocks, "back": back, "xya": xya, "octave": octv}
    raise ValueError(kind)


def run_oracle(kind, c):
    O = po.Oracle
    if kind == "init":
        n, m, prev = O.search_for_initialization_full(c["kp1"], c["D1"], c["kp2"], c["D2"], c["bounds"], c["prev"], int(c["window"]), float(c["nnratio"]),
                                                      bool(c["check_ori"]))
        return {"n": np.int32(n), "m": m, "prev": prev}
    if kind == "proj_frame":
        n, m = O.search_by_projection_frame_full(c["kpc"], c["Dc"], c["TcwC"], c["taken"], c["kpl"], c["has"], c["outlier"], c["wpos"], c["Dmp"], c["K4"],
                                                 c["bounds"], c["scale"], float(c["th"]), bool(c["check_ori"]))
        return {"n": np.int32(n), "m": m}
    if kind == "proj_points":
        n, m = O.search_by_projection_points_full(c["kpf"], c["Df"], c["taken"], c["mp"], c["Dmp"], c["bounds"], c["scale"], float(c["th"]),
                                                  float(c["nnratio"]))
        return {"n": np.int32(n), "m": m}
    if kind == "area":
        p, cd = O.features_in_area(c["kp"][:, :2], c["kp"][:, 3].astype(np.int32), c["bounds"], c["qx"], c["qy"], c["qr"], c["qmin"], c["qmax"])
        return {"ptr": p, "cand": cd}
    if kind == "distinct":
        # the reference drops observations of bad keyframes before anything else (MapPoint.cc:246-252): same list, filtered
        keep = c["bad"] == 0
        cnt = np.array([int(keep[c["ptr"][m]:c["ptr"][m + 1]].sum()) for m in range(len(c["ptr"]) - 1)])
        ptr = np.concatenate([[0], np.cumsum(cnt)]).astype(np.int32)
        desc = c["desc"][keep]
        bi, _ = O.distinctive_descriptors(desc, ptr)
        out = np.zeros((len(cnt), 32), np.uint8)
        valid = bi >= 0
        out[valid] = desc[ptr[:-1][valid] + bi[valid]]
        return {"desc": out, "valid": valid}
    if kind == "undistort":
        return {"out": O.undistort(c["xy"], c["K4"], c["dist"], int(c["mode"]))}
    if kind == "bounds":
        return {"b": O.image_bounds(int(c["cols"]), int(c["rows"]), c["K4"], c["dist"], int(c["mode"]))}
    if kind == "maxima":
        return {"ind": np.array(O.three_maxima(c["sizes"]), np.int32)}
    if kind == "distance":
        L = O.lib()
        return {"d": np.array([L.orc_descriptor_distance(po._p(np.ascontiguousarray(a)), po._p(np.ascontiguousarray(b))) for a, b in zip(c["A"], c["B"])], np.int32)}
    if kind == "wire":
        # Osmap.cpp:775-786, :819-831 restated: rows re-read as eight native 32-bit words; the record keeps x, y, octave, angle
        d = np.ascontiguousarray(c["desc"])
        return {"blocks": d.view(np.uint32).reshape(-1, 8).copy(), "back": d.copy(), "xya": np.ascontiguousarray(c["kp"][:, :3]),
                "octave": c["kp"][:, 3].astype(np.int32)}
    raise ValueError(kind)


def same(a, b):
    """Bit-for-bit equality of two result dicts (floats compared through their bytes)."""
    if a.keys() != b.keys():
        return False
    for k in a:
        x, y = np.asarray(a[k]), np.asarray(b[k])
        if x.shape != y.shape or x.dtype != y.dtype or x.tobytes() != y.tobytes():
            return False
    return True
