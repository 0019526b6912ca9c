"""Fuzz: the restated matcher / Frame / MapPoint oracle (oracle/liboracle.so) against the reference's own function bodies
(oracle/_ref/libos1_refmatch.so: ORBmatcher.cc, Frame.cc, MapPoint.cc sliced by path and compiled unmodified).
CPU only; needs /root/reference (the build container).   python scripts/fuzz_matcher_vs_reference.py [cases-per-kind] [first-seed]
"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import matcher_cases as mc  # noqa: E402
from oracle import pyoracle as po  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 50
    s0 = int(sys.argv[2]) if len(sys.argv) > 2 else 1000
    po.build(ref=True)
    if not po.RefMatch.available():
        sys.exit("oracle/_ref/libos1_refmatch.so is missing (no /root/reference here?)")
    bad = 0
    t0 = time.time()
    for kind in mc.KINDS:
        nbad = 0; info = []
        for seed in range(s0, s0 + n):
            c = mc.make_case(kind, seed)
            r = mc.run_reference(kind, c); o = mc.run_oracle(kind, c)
            if not mc.same(r, o):
                nbad += 1
                diff = {k: int((np.asarray(r[k]).ravel() != np.asarray(o[k]).ravel()).sum()) if np.asarray(r[k]).shape == np.asarray(o[k]).shape else "shape"
                        for k in r}
                info.append((seed, diff))
        print("%-12s %4d cases  %d mismatching %s" % (kind, n, nbad, info[:3]), flush=True)
        bad += nbad
    print("total mismatching cases: %d  (%.1f s)" % (bad, time.time() - t0))
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
