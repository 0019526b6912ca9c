"""Pyramid time per frame against the batch size for the two paths (OS1_PYR_LEVELS_MIN=1: per-level launches; =100000: fused chain).
   RES=1080p|720p|vga python scripts/pyr_crossover.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import os1_b200
from os1_b200 import synth
RES = {"vga": (480, 640, 1000), "720p": (720, 1280, 1500), "1080p": (1080, 1920, 2000)}[os.environ.get("RES", "1080p")]
R, C, nf = RES
dev = torch.device("cuda", 0)
Bmax = 128
fr = torch.from_numpy(synth.video(Bmax, R, C, seed=5)).to(dev)
e = os1_b200.ORBextractor(nf, 1.2, 8, 20, 7)
cap = e.max_keypoints(); e.reserve(R, C, Bmax)
kd = torch.empty((Bmax, cap, 28), dtype=torch.uint8, device=dev); dd = torch.empty((Bmax, cap, 32), dtype=torch.uint8, device=dev)
cd = torch.zeros(Bmax, dtype=torch.int32, device=dev)
e.set_timing(True)
for B in (2, 4, 8, 16, 32, 64, 128):
    acc = np.zeros(4)
    for it in range(8):
        e.extract_batch_raw(fr.data_ptr(), B, R, C, C, R * C, True, kd.data_ptr(), dd.data_ptr(), cap, cd.data_ptr(), True)
        t = e.stage_times_ms()
        if it >= 3: acc += t
    print("B=%3d  pyramid %.4f ms = %.2f us/frame   (launches %d)" % (B, acc[0] / 5, acc[0] / 5 / B * 1e3, e.launches_per_call()), flush=True)
