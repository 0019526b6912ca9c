import torch, time
x = torch.empty(256*720*1280, dtype=torch.uint8).pin_memory(); d = torch.empty_like(x, device="cuda")
o = torch.empty(24*1024*1024, dtype=torch.uint8, device="cuda"); oh = torch.empty_like(o, device="cpu").pin_memory()
for _ in range(3): d.copy_(x, non_blocking=True)
torch.cuda.synchronize(); t=time.perf_counter()
for _ in range(10): d.copy_(x, non_blocking=True)
torch.cuda.synchronize(); dt=(time.perf_counter()-t)/10
print("H2D 236MB: %.2f ms  %.1f GB/s" % (dt*1e3, x.numel()/dt/1e9))
s2 = torch.cuda.Stream()
torch.cuda.synchronize(); t=time.perf_counter()
for _ in range(10):
    d.copy_(x, non_blocking=True)
    with torch.cuda.stream(s2): oh.copy_(o, non_blocking=True)
torch.cuda.synchronize(); dt=(time.perf_counter()-t)/10
print("H2D 236MB + concurrent D2H 24MB: %.2f ms" % (dt*1e3))
