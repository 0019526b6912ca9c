"""Developer diagnostic: stage-by-stage GPU vs oracle comparison on seeded frames (prints mismatch counts)."""
import sys, os, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import os1_b200
from os1_b200 import synth
from oracle.pyoracle import Oracle

def run(h, w, nf, seed):
    img = synth.frame(h, w, seed)
    o = Oracle(nf)
    ko, do = o.extract(img)
    e = os1_b200.ORBextractor(nf, 1.2, 8, 20, 7)
    t = time.time(); kg, dg = e(img); t1 = time.time() - t
    t = time.time(); kg, dg = e(img); t2 = time.time() - t
    print(f"{w}x{h} nf={nf} seed={seed}: oracle n={len(ko)} gpu n={len(kg)}  first {t1*1e3:.1f} ms, second {t2*1e3:.2f} ms")
    for l in range(8):
        po = o.level_pixels(l); pg = e.pyramid_level(l)
        co = o.candidates(l); cg = e.candidates(l)
        xo, ao = o.level_keypoints(l); xg = e.level_keypoints(l)
        print(f"  L{l}: pyr diff {(po != pg).sum() if po.shape == pg.shape else 'shape'}  cands {len(co)}/{len(cg)} eq={np.array_equal(co, cg)}"
              f"  kept {len(xo)}/{len(xg)} eq={np.array_equal(xo, xg)}")
        if not np.array_equal(co, cg) and len(co) and len(cg):
            so = set(map(tuple, co.tolist())); sg = set(map(tuple, cg.tolist()))
            print("     cand only-oracle", len(so - sg), "only-gpu", len(sg - so), list(so - sg)[:3], list(sg - so)[:3])
        if not np.array_equal(xo, xg) and len(xo) and len(xg):
            so = set(map(tuple, xo.tolist())); sg = set(map(tuple, xg.tolist()))
            print("     kept only-oracle", len(so - sg), "only-gpu", len(sg - so))
    n = min(len(ko), len(kg))
    g = np.stack([kg[f] for f in ("x", "y", "size", "angle", "response")], 1)[:n]
    same_xy = np.array_equal(g[:, :2], ko[:n, :2])
    print("  final: xy eq", same_xy, " size eq", np.array_equal(g[:, 2], ko[:n, 2]), " angle maxdiff",
          np.abs(g[:, 3] - ko[:n, 3]).max() if n else 0, " angle bit-eq", np.array_equal(g[:, 3], ko[:n, 3]),
          " resp eq", np.array_equal(g[:, 4], ko[:n, 4]), " octave eq", np.array_equal(kg["octave"][:n], ko[:n, 5].astype(np.int32)),
          " desc rows differing", int((dg[:n] != do[:n]).any(1).sum()))

if __name__ == "__main__":
    run(480, 640, 1000, 0)
    run(720, 1280, 1500, 1)
    run(1080, 1920, 2000, 2)
    A = np.random.default_rng(0).integers(0, 256, (2000, 32), dtype=np.uint8)
    B = np.random.default_rng(1).integers(0, 256, (2000, 32), dtype=np.uint8)
    B[5] = A[7]; B[9] = A[7]
    r = os1_b200.matcher.hamming_knn2(A, B); q = Oracle.hamming_knn2(A, B)
    print("hamming knn2 eq:", [np.array_equal(a, b) for a, b in zip(r, q)])
