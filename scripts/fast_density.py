"""FAST kernel time against corner density: the bench's synthetic frames (19 % of the pixels are corners at threshold 7) and the
same frames smoothed by a k x k box filter (real footage is far sparser).   python scripts/fast_density.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import os1_b200
from os1_b200 import synth
R, C, nf, B = 1080, 1920, 2000, 64
dev = torch.device("cuda", 0)
base = synth.video(B, R, C, seed=7).astype(np.float32)
def box(a, k):
    if k == 1: return a
    c = np.cumsum(np.cumsum(np.pad(a, ((0, 0), (k, 0), (k, 0)), mode="edge"), 1), 2)
    out = (c[:, k:, k:] - c[:, :-k, k:] - c[:, k:, :-k] + c[:, :-k, :-k]) / (k * k)
    return out
e = os1_b200.ORBextractor(nf, 1.2, 8, 20, 7)
cap = e.max_keypoints(); e.reserve(R, C, B)
kd = torch.empty((B, cap, 28), dtype=torch.uint8, device=dev); dd = torch.empty((B, cap, 32), dtype=torch.uint8, device=dev)
cd = torch.zeros(B, dtype=torch.int32, device=dev)
e.set_timing(True)
for k in (1, 2, 3, 5):
    fr = torch.from_numpy(np.clip(box(base, k), 0, 255).astype(np.uint8)[:, :R, :C].copy()).to(dev)
    acc = np.zeros(4)
    for it in range(8):
        e.extract_batch_raw(fr.data_ptr(), B, R, C, C, R * C, True, kd.data_ptr(), dd.data_ptr(), cap, cd.data_ptr(), True)
        t = e.stage_times_ms()
        if it >= 3: acc += t
    print("box %d: FAST %.3f ms per %d frames, keypoints/frame %.0f, digest %d" % (k, acc[1] / 5, B, cd.float().mean().item(), int(dd.sum().item())), flush=True)
