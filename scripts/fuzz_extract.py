"""Random geometries / parameters: GPU extraction against the oracle (final keypoints and descriptors, bit for bit)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import os1_b200
from os1_b200 import synth
from oracle.pyoracle import Oracle

n = int(sys.argv[1]) if len(sys.argv) > 1 else 40
seed0 = int(sys.argv[2]) if len(sys.argv) > 2 else 0
bad = 0; skipped = 0
for k in range(n):
    rng = np.random.default_rng(seed0 + k)
    h = int(rng.integers(70, 1300)); w = int(rng.integers(max(70, h // 2 + 30), min(2000, 4 * h)))
    nl = int(rng.integers(1, 9)); sf = float(rng.choice([1.1, 1.2, 1.25, 1.3, 1.41, 1.5, 1.6]))
    ini = int(rng.integers(8, 60)); mn = int(rng.integers(2, 30)); nf = int(rng.integers(50, 3000))
    img = synth.frame(h, w, seed0 + k)
    if k % 3 == 1:
        img = (img // 4 + 90).astype(np.uint8)              # low contrast: the minTh retry everywhere
    if k % 5 == 2:
        img = np.random.default_rng(k).integers(0, 256, (h, w), dtype=np.uint8)       # pure noise: dense corners
    try:
        e = os1_b200.ORBextractor(nf, sf, nl, ini, mn)
        kg, dg = e(img)
    except os1_b200.Os1Error as ex:
        if ex.code == -4:
            skipped += 1; continue
        raise
    o = Oracle(nf, sf, nl, ini, mn)
    r = o.extract(img); ko, do = r[0], r[1]
    ok = len(kg) == len(ko) and np.array_equal(dg, do) and np.array_equal(kg["x"], ko[:, 0]) and np.array_equal(kg["y"], ko[:, 1]) \
        and np.array_equal(kg["angle"], ko[:, 3]) and np.array_equal(kg["response"], ko[:, 4])
    if ok and k % 2 == 0:                 # the batch path (other pyramid plan, 512-thread quadtree, chunked upload) on the same image
        kb, db, cb = e.extract_batch(np.stack([img] * 5))
        for f in range(5):
            ok = ok and cb[f] == len(kg) and np.array_equal(db[f, :cb[f]], dg) and np.array_equal(kb[f, :cb[f]], kg)
    if not ok:
        bad += 1
        print("MISMATCH case", seed0 + k, (h, w, nl, sf, ini, mn, nf), "n gpu/oracle", len(kg), len(ko))
print("cases %d, skipped (unsupported geometry) %d, mismatches %d" % (n, skipped, bad))
sys.exit(1 if bad else 0)
