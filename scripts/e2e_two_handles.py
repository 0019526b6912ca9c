"""Host-to-host batches through TWO handles driven by two host threads (each call is the synchronous C-ABI call; ctypes drops the GIL
inside it): while one handle drains its last chunks the other is already uploading, so the host->device link never idles between
batches.  RES=vga|720p|1080p B=<frames> STEPS=<n> NH=<handles>"""
import os, sys, threading, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import os1_b200
from os1_b200 import synth
RES = {"vga": (480, 640, 1000), "720p": (720, 1280, 1500), "1080p": (1080, 1920, 2000)}[os.environ.get("RES", "1080p")]
B, ROWS, COLS = int(os.environ.get("B", 256)), RES[0], RES[1]
steps = int(os.environ.get("STEPS", 20)); NH = int(os.environ.get("NH", 2))
dev = torch.device("cuda", 0)
fh = torch.from_numpy(synth.video(B, ROWS, COLS, seed=1000)).pin_memory()
H = []
for i in range(NH):
    ext = os1_b200.ORBextractor(RES[2], 1.2, 8, 20, 7, device=0)
    cap = ext.max_keypoints(); ext.reserve(ROWS, COLS, B)
    kh = torch.empty((B, cap, 28), dtype=torch.uint8).pin_memory(); dh = torch.empty((B, cap, 32), dtype=torch.uint8).pin_memory()
    ch = torch.zeros(B, dtype=torch.int32).pin_memory()
    H.append((ext, kh, dh, ch, cap))
def step(i):
    ext, kh, dh, ch, cap = H[i]
    ext.extract_batch_raw(fh.data_ptr(), B, ROWS, COLS, COLS, ROWS * COLS, False, kh.data_ptr(), dh.data_ptr(), cap, ch.data_ptr(), False)
for i in range(NH):
    for _ in range(3): step(i)
for nh in range(1, NH + 1):
    def work(i, n):
        for _ in range(n): step(i)
    torch.cuda.synchronize(); t = time.perf_counter()
    th = [threading.Thread(target=work, args=(i, steps // nh)) for i in range(nh)]
    for x in th: x.start()
    for x in th: x.join()
    torch.cuda.synchronize(); dt = (time.perf_counter() - t) / (steps // nh * nh)
    print("%d handle(s): %.3f ms/step  %.2f k frames/s" % (nh, dt * 1e3, B / dt / 1e3), flush=True)
for i in range(1, NH):
    assert np.array_equal(H[0][3].numpy(), H[i][3].numpy()) and np.array_equal(H[0][2].numpy(), H[i][2].numpy())
print("outputs of all handles identical")
