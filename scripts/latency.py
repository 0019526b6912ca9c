"""Single-frame latency through the C ABI with host buffers (what Frame.cc:133 would see)."""
import sys, os, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import os1_b200
from os1_b200 import synth
for (h, w, nf) in [(480, 640, 1000), (720, 1280, 1500), (1080, 1920, 2000)]:
    img = synth.frame(h, w, 0)
    e = os1_b200.ORBextractor(nf, 1.2, 8, 20, 7)
    for _ in range(20): e(img)
    ts = []
    for _ in range(200):
        t = time.perf_counter(); k, d = e(img); ts.append(time.perf_counter() - t)
    ts = np.array(ts) * 1e3
    e.set_timing(True); e(img); st = e.stage_times_ms()
    print("%dx%d nf=%d: n=%d  median %.3f ms  p10 %.3f  p90 %.3f   kernels(ms) pyr %.3f fast %.3f quad %.3f desc %.3f sum %.3f" % (
        w, h, nf, len(k), np.median(ts), np.percentile(ts, 10), np.percentile(ts, 90), st[0], st[1], st[2], st[3], st.sum()))
