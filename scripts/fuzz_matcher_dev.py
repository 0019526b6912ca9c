"""GPU fuzz: the three whole-function matchers on the device (os1_matcher) against the oracle's whole-function composition, on the
seeded scenarios of oracle/matcher_cases.py.   python scripts/fuzz_matcher_dev.py [cases-per-kind] [first-seed]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle import matcher_cases as mc  # noqa: E402
from oracle import pyoracle  # noqa: E402
import test_gpu_matcher_dev as t  # noqa: E402
from os1_b200 import matcher  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 100
s0 = int(sys.argv[2]) if len(sys.argv) > 2 else 5000
pyoracle.build(ref=False)
bad = 0
for kind in ("init", "proj_frame", "proj_points"):
    M = matcher.ORBmatcher(0.9, True)
    nb = 0
    for seed in range(s0, s0 + n):
        c = mc.make_case(kind, seed)
        if not mc.same(t.strip(t.run_gpu(kind, c, M)), mc.run_oracle(kind, c)):
            nb += 1; print("MISMATCH", kind, seed, flush=True)
    print("%-12s %d cases, %d mismatching" % (kind, n, nb), flush=True)
    bad += nb
sys.exit(1 if bad else 0)
