"""GPU: a short run of the two fuzzers (random geometries / extractor parameters, random matcher inputs, random batch layouts) against the
oracle.  The long runs (hundreds of cases, `scripts/fuzz_extract.py N SEED`, `scripts/fuzz_matcher.py N SEED`) are how
the cosf/sinf rounding difference of DESIGN.md section 4 was found."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("script,n,seed", [("fuzz_extract.py", 24, 7000), ("fuzz_matcher.py", 16, 7100), ("fuzz_batch.py", 16, 7200)])
def test_fuzz(script, n, seed):
    r = subprocess.run([sys.executable, os.path.join(ROOT, "scripts", script), str(n), str(seed)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "mismatches 0" in r.stdout
