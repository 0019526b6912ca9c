"""Pass rate of the compass pretest per 4-pixel group on the bench's frames (all pyramid levels via the oracle)."""
import numpy as np, sys
sys.path.insert(0, '.')
from os1_b200 import synth
from oracle.pyoracle import Oracle
fr = synth.video(2, 720, 1280, seed=1000)[1]
o = Oracle(1500, 1.2, 8, 20, 7); o.extract(fr)
ring = [(0,3),(1,3),(2,2),(3,1),(3,0),(3,-1),(2,-2),(1,-3),(0,-3),(-1,-3),(-2,-2),(-3,-1),(-3,0),(-3,1),(-2,2),(-1,3)]
tot = np.zeros(3); n = 0
for l in range(8):
    img = o.level_pixels(l).astype(np.int32); H, W = img.shape
    c = img[3:H-3, 3:W-3]
    R = np.stack([img[3+dy:H-3+dy, 3+dx:W-3+dx] for dx, dy in ring])
    def score(idx, arc):
        k = len(idx); r = R[idx]; best = np.zeros_like(c)
        for i in range(k):
            sel = [(i + j) % k for j in range(arc)]
            best = np.maximum(best, np.maximum(r[sel].min(0) - c, c - r[sel].max(0)))
        return best
    full = score(list(range(16)), 9) > 7; b4 = score([0,4,8,12], 2) > 7
    def grp(m):
        w = (m.shape[1] // 4) * 4
        return m[:, :w].reshape(m.shape[0], -1, 4).any(2)
    gf, g4 = grp(full), grp(b4)
    print("level", l, "pixel corner %.3f  group corner %.3f  group pretest pass %.3f" % (full.mean(), gf.mean(), g4.mean()))
