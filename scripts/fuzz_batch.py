"""Random batch layouts (row step, frame stride, host / device inputs and outputs, odd alignments, batch sizes) against
the single-frame path of the same handle."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import os1_b200
from os1_b200 import synth, KEYPOINT_DTYPE

n = int(sys.argv[1]) if len(sys.argv) > 1 else 30
seed0 = int(sys.argv[2]) if len(sys.argv) > 2 else 0
bad = 0
for k in range(n):
    case = seed0 + k
    rng = np.random.default_rng(case)
    h = int(rng.integers(90, 500)); w = int(rng.integers(100, 700)); B = int(rng.integers(1, 45))
    nf = int(rng.integers(100, 1200)); nl = int(rng.integers(2, 9))
    try:
        e = os1_b200.ORBextractor(nf, 1.2, nl, 20, 7)
        frames = synth.video(B, h, w, seed=case, max_shift=6)
        ref = [e(f) for f in frames]
    except os1_b200.Os1Error as ex:
        if ex.code == -4: continue
        raise
    step = w + int(rng.choice([0, 0, 1, 3, 16, 37])); fstride = step * h + int(rng.choice([0, 0, 5, 64, 1000]))
    off = int(rng.choice([0, 0, 1, 4, 16, 100]))
    buf = np.zeros(off + fstride * B + 64, np.uint8)
    for f in range(B):
        v = buf[off + f * fstride: off + f * fstride + step * h].reshape(h, step); v[:, :w] = frames[f]
    cap = e.max_keypoints() + int(rng.choice([0, 7, 100]))
    in_dev = bool(rng.integers(0, 2)); out_dev = bool(rng.integers(0, 2))
    if in_dev:
        tb = torch.from_numpy(buf).cuda(); src_ptr = tb.data_ptr() + off
    else:
        tb = torch.from_numpy(buf).pin_memory() if rng.random() < 0.5 else torch.from_numpy(buf); src_ptr = tb.data_ptr() + off
    if out_dev:
        kp = torch.zeros((B, cap, 28), dtype=torch.uint8, device="cuda"); de = torch.zeros((B, cap, 32), dtype=torch.uint8, device="cuda")
        cn = torch.zeros(B, dtype=torch.int32, device="cuda")
    else:
        kp = torch.zeros((B, cap, 28), dtype=torch.uint8); de = torch.zeros((B, cap, 32), dtype=torch.uint8); cn = torch.zeros(B, dtype=torch.int32)
    e.extract_batch_raw(src_ptr, B, h, w, step, fstride, in_dev, kp.data_ptr(), de.data_ptr(), cap, cn.data_ptr(), out_dev)
    e.synchronize()
    kpn = kp.cpu().numpy().view(KEYPOINT_DTYPE).reshape(B, cap); den = de.cpu().numpy(); cnn = cn.cpu().numpy()
    ok = True
    for f in range(B):
        rk, rd = ref[f]
        ok = ok and cnn[f] == len(rk) and np.array_equal(den[f, :cnn[f]], rd) and np.array_equal(kpn[f, :cnn[f]], rk)
    if not ok:
        bad += 1
        print("MISMATCH case", case, dict(h=h, w=w, B=B, nf=nf, nl=nl, step=step, fstride=fstride, off=off, cap=cap, in_dev=in_dev, out_dev=out_dev))
print("cases %d, mismatches %d" % (n, bad))
sys.exit(1 if bad else 0)
