"""Developer helper: os1_hamming_knn2 (device pointers) timed back to back for a few problem sizes."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, numpy as np, time
from os1_b200 import matcher
dev=torch.device("cuda",0)
rng=np.random.default_rng(0)
for (na,nb) in ((32,32),(2000,2000),(1000,1000),(2000,4000),(500,2000),(2400,2000)):
    a=torch.from_numpy(rng.integers(0,256,(na,32),dtype=np.uint8)).to(dev); b=torch.from_numpy(rng.integers(0,256,(nb,32),dtype=np.uint8)).to(dev)
    o=[torch.zeros(na,dtype=torch.int32,device=dev) for _ in range(3)]
    s=torch.cuda.current_stream(dev)
    call=lambda: matcher.hamming_knn2_raw(a.data_ptr(),na,b.data_ptr(),nb,o[0].data_ptr(),o[1].data_ptr(),o[2].data_ptr(),device=0,stream=s.cuda_stream)
    for _ in range(20): call()
    e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record(s)
    for _ in range(200): call()
    e1.record(s); torch.cuda.synchronize()
    us=e0.elapsed_time(e1)*1e3/200
    print("knn2 %dx%d: %.2f us  %.0f Gpairs/s"%(na,nb,us,na*nb/us/1e3))
# the same 2000 x 2000 call replayed from a CUDA graph: no host work between launches, so this is the GPU's own time per call
na, nb = 2000, 2000
a=torch.from_numpy(rng.integers(0,256,(na,32),dtype=np.uint8)).to(dev); b=torch.from_numpy(rng.integers(0,256,(nb,32),dtype=np.uint8)).to(dev)
o=[torch.zeros(na,dtype=torch.int32,device=dev) for _ in range(3)]
st = torch.cuda.Stream(dev)
call=lambda: matcher.hamming_knn2_noalloc_raw(a.data_ptr(),na,b.data_ptr(),nb,o[0].data_ptr(),o[1].data_ptr(),o[2].data_ptr(),device=0,stream=st.cuda_stream)
call(); torch.cuda.synchronize()
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g, stream=st):
    for _ in range(50): call()
g.replay(); torch.cuda.synchronize()
e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10): g.replay()
e1.record(); torch.cuda.synchronize()
print("knn2 2000x2000 replayed from a CUDA graph (50 launches per graph): %.2f us per call" % (e0.elapsed_time(e1)*1e3/500))
