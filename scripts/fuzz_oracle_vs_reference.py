"""CPU only: the oracle restatement against the unmodified reference ORBextractor.cc (canonical-allocator build) on random
geometries and parameters — keypoints, angles and descriptors bit for bit."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from os1_b200 import synth
from oracle.pyoracle import Oracle, Reference

n = int(sys.argv[1]) if len(sys.argv) > 1 else 40
seed0 = int(sys.argv[2]) if len(sys.argv) > 2 else 0
if not Reference.available(canon=True):
    print("reference library not built (oracle/_ref): nothing to compare"); sys.exit(0)
bad = 0; skipped = 0
for k in range(n):
    case = seed0 + k
    rng = np.random.default_rng(case)
    h = int(rng.integers(70, 900)); w = int(rng.integers(max(70, h // 2 + 30), min(1600, 4 * h)))
    nl = int(rng.integers(1, 9)); sf = float(rng.choice([1.1, 1.2, 1.25, 1.3, 1.41, 1.5, 1.6]))
    ini = int(rng.integers(8, 60)); mn = int(rng.integers(2, 30)); nf = int(rng.integers(50, 2500))
    img = synth.frame(h, w, case)
    if k % 3 == 1: img = (img // 4 + 90).astype(np.uint8)
    if k % 5 == 2: img = np.random.default_rng(k).integers(0, 256, (h, w), dtype=np.uint8)
    # the reference divides by zero / reads out of bounds on geometries it was never meant for: keep to the supported ones
    W = w / sf ** (nl - 1); H = h / sf ** (nl - 1)
    if W < 70 or H < 70 or not (0.55 < (w - 26) / (h - 26) < 4.4):
        skipped += 1; continue
    try:
        ko, do = Oracle(nf, sf, nl, ini, mn).extract(img)[:2]
        kr, dr = Reference(nf, sf, nl, ini, mn, canon=True).extract(img)[:2]
    except Exception as ex:
        skipped += 1; continue
    ok = len(ko) == len(kr) and np.array_equal(do, dr) and np.array_equal(ko, kr)
    if not ok:
        bad += 1; print("MISMATCH case", case, (h, w, nl, sf, ini, mn, nf), len(ko), len(kr))
print("cases %d, skipped %d, mismatches %d" % (n, skipped, bad))
sys.exit(1 if bad else 0)
