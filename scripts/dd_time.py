"""Time os1_distinctive_descriptors (device pointers) for a map's worth of points."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from os1_b200 import _lib
rng = np.random.default_rng(4)
for M, lo, hi in ((50000, 2, 25), (5000, 100, 257), (200000, 2, 8)):
    cnt = rng.integers(lo, hi, M); ptr = np.concatenate([[0], np.cumsum(cnt)]).astype(np.int32)
    dd = torch.from_numpy(rng.integers(0, 256, (int(ptr[-1]), 32), dtype=np.uint8)).cuda(); dp = torch.from_numpy(ptr).cuda()
    bi = torch.empty(M, dtype=torch.int32, device="cuda"); bm = torch.empty_like(bi)
    s = torch.cuda.current_stream().cuda_stream
    f = lambda: _lib.check(_lib.load().os1_distinctive_descriptors(dd.data_ptr(), dp.data_ptr(), M, bi.data_ptr(), bm.data_ptr(), 1, 0, s))
    for _ in range(3): f()
    t = time.perf_counter()
    for _ in range(20): f()
    dt = (time.perf_counter() - t) / 20
    print("M=%d obs %d..%d: %.3f ms per call, %.1f M points/s, %.1f G pairs/s" % (M, lo, hi - 1, dt * 1e3, M / dt / 1e6, float((cnt.astype(float) ** 2).sum()) / dt / 1e9))
