"""Developer helper: per-kernel milliseconds of the extraction step for the three BASELINE resolutions (device-resident batches),
plus a quick bit-exactness check of one frame per resolution against the CPU oracle.   python scripts/kt.py [1080p 720p vga] [--nocheck]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import os1_b200  # noqa: E402
from os1_b200 import synth  # noqa: E402

W = {"vga": (480, 640, 1000, 512), "720p": (720, 1280, 1500, 256), "1080p": (1080, 1920, 2000, 128)}
keys = [a for a in sys.argv[1:] if a in W] or ["1080p", "720p"]
dev = torch.device("cuda", 0)
for k in keys:
    R, C, nf, B = W[k]
    if "--nocheck" not in sys.argv:
        from oracle.pyoracle import Oracle
        img = synth.frame(R, C, 3)
        e = os1_b200.ORBextractor(nf, 1.2, 8, 20, 7)
        kp, d = e(img); ko, do = Oracle(nf).extract(img)
        ok = len(kp) == len(ko) and np.array_equal(kp["x"], ko[:, 0]) and np.array_equal(kp["y"], ko[:, 1]) and np.array_equal(kp["angle"], ko[:, 3]) \
            and np.array_equal(kp["response"], ko[:, 4]) and np.array_equal(d, do)
        print("%s parity vs oracle: %s (%d keypoints)" % (k, "OK" if ok else "MISMATCH", len(kp)), flush=True)
        e.close()
    fr = torch.from_numpy(synth.video(B, R, C, seed=1000)).to(dev)
    e = os1_b200.ORBextractor(nf, 1.2, 8, 20, 7)
    cap = e.max_keypoints(); e.reserve(R, C, B)
    st = torch.cuda.Stream(dev); e.set_stream(st.cuda_stream)
    kpd = torch.empty((B, cap, 28), dtype=torch.uint8, device=dev); dd = torch.empty((B, cap, 32), dtype=torch.uint8, device=dev)
    cn = torch.zeros(B, dtype=torch.int32, device=dev)
    run = lambda: e.extract_batch_raw(fr.data_ptr(), B, R, C, C, R * C, True, kpd.data_ptr(), dd.data_ptr(), cap, cn.data_ptr(), True)
    for _ in range(3):
        run()
    e.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for _ in range(20):
        run()
    e1.record(st); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    e.set_timing(True)
    acc = np.zeros(4)
    for _ in range(5):
        run(); acc += e.stage_times_ms()
    acc /= 5
    print("%s B=%d: step %.3f ms = %.0f frames/s | pyramid %.3f  fast %.3f  quadtree %.3f  describe %.3f" % (k, B, ms, B / ms * 1e3, *acc), flush=True)
    e.close()
