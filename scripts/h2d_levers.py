"""Host->device feed experiment for N ranks on one box (run under torchrun): what limits the end-to-end path when every GPU is
fed from pinned host memory at once?  Each rank copies the bench's step (256 x 1080p frames = 531 MB) to its GPU repeatedly with
plain asynchronous copies, all ranks barrier-aligned, and the aggregate GB/s is printed for a few variants:
  default     cudaHostAlloc(default), whatever CPU affinity the launcher gave
  wc          cudaHostAllocWriteCombined (the host never reads the frames back)
  affinity    rank r pinned to its own slice of the host cores before allocating and touching its buffer (first-touch NUMA placement)
  chunked     the same bytes as 16 copies (what the library's pipeline issues) instead of one
Prints one JSON line on rank 0.   torchrun --nproc-per-node N scripts/h2d_levers.py"""
import ctypes as C
import json
import os
import subprocess
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); lr = int(os.environ.get("LOCAL_RANK", "0"))
dev = torch.device("cuda", lr); torch.cuda.set_device(dev)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
cudart = C.CDLL("libcudart.so.12") if os.path.exists("/usr/local/cuda/lib64/libcudart.so.12") else C.CDLL("libcudart.so")
BYTES = 256 * 1080 * 1920
dst = torch.empty(BYTES, dtype=torch.uint8, device=dev)
st = torch.cuda.Stream(dev)


def host_alloc(flags):
    p = C.c_void_p()
    rc = cudart.cudaHostAlloc(C.byref(p), C.c_size_t(BYTES), C.c_uint(flags))
    assert rc == 0, rc
    return p


def touch(p):
    C.memset(p, 7, BYTES)


def barrier():
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)


def measure(p, chunks=1, reps=8):
    sz = BYTES // chunks
    def once():
        for c in range(chunks):
            rc = cudart.cudaMemcpyAsync(C.c_void_p(dst.data_ptr() + c * sz), C.c_void_p(p.value + c * sz), C.c_size_t(sz), C.c_int(1), C.c_void_p(st.cuda_stream))
            assert rc == 0, rc
    once(); barrier()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for _ in range(reps):
        once()
    e1.record(st)
    barrier()
    t = torch.tensor([e0.elapsed_time(e1) / reps], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return world * BYTES / (float(t.item()) * 1e-3) / 1e9


out = {"ranks": world, "bytes_per_rank_per_step": BYTES, "host_cpus": os.cpu_count(), "affinity_at_start": len(os.sched_getaffinity(0))}
p0 = host_alloc(0); touch(p0)
out["default_gbs"] = measure(p0)
out["default_16chunks_gbs"] = measure(p0, 16)
pw = host_alloc(4); touch(pw)          # cudaHostAllocWriteCombined
out["write_combined_gbs"] = measure(pw)
# affinity: an equal slice of the allowed cores per rank, then allocate + first-touch from there
cores = sorted(os.sched_getaffinity(0))
k = max(1, len(cores) // world)
mine = cores[lr * k:(lr + 1) * k] or cores
os.sched_setaffinity(0, mine)
pa = host_alloc(0); touch(pa)
out["affinity_slice_gbs"] = measure(pa)
out["affinity_slice_16chunks_gbs"] = measure(pa, 16)
# one rank at a time (what a single GPU gets alone), measured on rank 0 only while the others wait
if world > 1:
    dist.barrier()
alone = None
if rank == 0:
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for _ in range(8):
        cudart.cudaMemcpyAsync(C.c_void_p(dst.data_ptr()), p0, C.c_size_t(BYTES), C.c_int(1), C.c_void_p(st.cuda_stream))
    e1.record(st); torch.cuda.synchronize(dev)
    alone = BYTES / (e0.elapsed_time(e1) / 8 * 1e-3) / 1e9
if world > 1:
    dist.barrier()
if rank == 0:
    out["one_rank_alone_gbs"] = alone
    try:
        out["topo"] = subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True, timeout=20).stdout[-3000:]
        out["numa"] = subprocess.run(["bash", "-c", "lscpu | grep -i -E 'numa|socket|model name|^cpu\\(s\\)'"], capture_output=True, text=True, timeout=20).stdout
    except Exception as ex:
        out["topo"] = str(ex)
    print(json.dumps(out), flush=True)
if world > 1:
    dist.destroy_process_group()
