import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import os1_b200
from os1_b200 import synth
from oracle.pyoracle import Oracle
img = synth.frame(480, 640, 0)
o = Oracle(1000); o.extract(img)
e = os1_b200.ORBextractor(1000, 1.2, 8, 20, 7); e(img)
co = o.candidates(0); cg = e.candidates(0)
print(len(co), len(cg))
print(co[:12].tolist()); print(cg[:12].tolist())
so = {(x, y): s for x, y, s in co.tolist()}; sg = {(x, y): s for x, y, s in cg.tolist()}
for dx in range(-4, 5):
    for dy in range(-4, 5):
        m = sum(1 for (x, y) in sg if (x + dx, y + dy) in so)
        if m > len(sg) // 3: print("shift", dx, dy, m)
both = [k for k in sg if k in so]
print("same pos", len(both), "same score", sum(1 for k in both if sg[k] == so[k]))
