"""One 1920x1080 frame through the extractor, a few times (profiling target for the latency path)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import os1_b200
from os1_b200 import synth
R, C, nf = {"vga": (480, 640, 1000), "720p": (720, 1280, 1500), "1080p": (1080, 1920, 2000)}[os.environ.get("RES", "1080p")]
img = synth.frame(R, C, 0)
e = os1_b200.ORBextractor(nf, 1.2, 8, 20, 7)
for _ in range(6): k, d = e(img)
print(len(k))
