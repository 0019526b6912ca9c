"""Quick A/B of the batch paths: device-resident step and host-to-host step, in ms.  RES=vga|720p|1080p B=<frames> STEPS=<n>;
the library's knobs (OS1_CHUNK, OS1_RAMP0, OS1_RAMPPCT, OS1_ALT, OS1_TRACE) are read from the environment."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import os1_b200
from os1_b200 import synth
RES = {"vga": (480, 640, 1000), "720p": (720, 1280, 1500), "1080p": (1080, 1920, 2000)}[os.environ.get("RES", "720p")]
B, ROWS, COLS = int(os.environ.get("B", 256)), RES[0], RES[1]
steps = int(os.environ.get("STEPS", 30))
dev = torch.device("cuda", 0)
fh = torch.from_numpy(synth.video(B, ROWS, COLS, seed=1000)).pin_memory(); fd = fh.to(dev)
ext = os1_b200.ORBextractor(RES[2], 1.2, 8, 20, 7, device=0)
cap = ext.max_keypoints(); ext.reserve(ROWS, COLS, B)
st = torch.cuda.current_stream(dev); ext.set_stream(st.cuda_stream)
kd = torch.empty((B, cap, 28), dtype=torch.uint8, device=dev); dd = torch.empty((B, cap, 32), dtype=torch.uint8, device=dev)
cd = torch.zeros(B, dtype=torch.int32, device=dev)
kh = torch.empty((B, cap, 28), dtype=torch.uint8).pin_memory(); dh = torch.empty((B, cap, 32), dtype=torch.uint8).pin_memory()
ch = torch.zeros(B, dtype=torch.int32).pin_memory()
def sd(): ext.extract_batch_raw(fd.data_ptr(), B, ROWS, COLS, COLS, ROWS * COLS, True, kd.data_ptr(), dd.data_ptr(), cap, cd.data_ptr(), True)
def se(): ext.extract_batch_raw(fh.data_ptr(), B, ROWS, COLS, COLS, ROWS * COLS, False, kh.data_ptr(), dh.data_ptr(), cap, ch.data_ptr(), False)
def sm(): ext.extract_batch_raw(fd.data_ptr(), B, ROWS, COLS, COLS, ROWS * COLS, True, kh.data_ptr(), dh.data_ptr(), cap, ch.data_ptr(), False)
for f, name in ((sd, "device"), (se, "e2e"), (sm, "device frames -> host outputs (chunked)")):
    for _ in range(5): f()
    torch.cuda.synchronize(); t = time.perf_counter()
    for _ in range(steps): f()
    torch.cuda.synchronize(); dt = (time.perf_counter() - t) / steps
    print("%s: %.3f ms/step  %.1f k frames/s" % (name, dt * 1e3, B / dt / 1e3), flush=True)
assert np.array_equal(ch.numpy(), cd.cpu().numpy())
if os.environ.get("STAGES"):
    ext.set_timing(True); acc = np.zeros(4)
    for _ in range(steps): sd(); acc += ext.stage_times_ms()
    print("stages ms:", np.round(acc / steps, 3))
