"""os1_b200 — B200-native (sm_100a) ORB front end and Hamming consumer behind os1's ORBextractor / ORBmatcher API.

The compute lives in libos1_b200.so (hand-written CUDA, C ABI in include/os1_b200.h); this package is the thin
host-side mirror of the reference interface.  There is no CPU fallback."""
from . import synth  # noqa: F401  (numpy only)
from ._lib import Os1Error, load  # noqa: F401
from .orb import ORBextractor, KEYPOINT_DTYPE  # noqa: F401
from . import matcher  # noqa: F401
