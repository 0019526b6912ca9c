"""GPU: os1_hamming_knn2_noalloc — the all-pairs matcher for callers that cannot allocate device memory.

Some hosts of this library may not call cudaMalloc at all while they run: a tracker with a per-frame deadline (an
allocation can stall on the driver's lock or synchronise the device), a process that handed out all device memory at
start-up, a plug-in whose host forbids allocation on worker threads.  For them the library carries its own fixed
scratch.  This file is named so that pytest collects it last: if it fails, every other test has already run.

Checked here, through the C ABI, with torch only as the owner of the input/output buffers: the new entry point gives
exactly the outputs of the existing os1_hamming_knn2 (bit-exact — integer distances and indices, the tolerance every
Hamming test in this suite uses), on two calls in a row with different data (the second call must not see anything the
first one left in the shared scratch), and sizes beyond the built-in scratch are refused with a clear error.  The same
outputs are also held against the CPU oracle's scan, and the device's free memory is read before and after a call to
show that the call itself takes none."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _run_both(A, B):
    import torch
    from os1_b200 import matcher
    dev = torch.device("cuda", 0)
    a = torch.from_numpy(A).to(dev); b = torch.from_numpy(B).to(dev)
    outs = [[torch.full((len(A),), -7, dtype=torch.int32, device=dev) for _ in range(3)] for _ in range(2)]
    torch.cuda.synchronize()
    s = torch.cuda.current_stream(dev).cuda_stream
    matcher.hamming_knn2_raw(a.data_ptr(), len(A), b.data_ptr(), len(B), *[t.data_ptr() for t in outs[0]], device=0, stream=s)
    matcher.hamming_knn2_noalloc_raw(a.data_ptr(), len(A), b.data_ptr(), len(B), *[t.data_ptr() for t in outs[1]], device=0, stream=s)
    torch.cuda.synchronize()
    return [[t.cpu().numpy() for t in o] for o in outs]


def test_noalloc_equals_existing_entry_point_twice_in_a_row():
    rng = np.random.default_rng(123)
    # first call: the headline 2000 x 2000 problem
    A = rng.integers(0, 256, (2000, 32), dtype=np.uint8); B = rng.integers(0, 256, (2000, 32), dtype=np.uint8)
    B[5] = A[9]; B[17] = A[9]
    ref, got = _run_both(A, B)
    for r, g in zip(ref, got):
        assert np.array_equal(r, g)
    assert got[0][9] == 5 and got[1][9] == 0 and got[2][9] == 0
    # second call, straight after, smaller and with other data: stale partials of the first call must not leak in
    A2 = rng.integers(0, 256, (300, 32), dtype=np.uint8); B2 = rng.integers(0, 256, (4097, 32), dtype=np.uint8)
    ref2, got2 = _run_both(A2, B2)
    for r, g in zip(ref2, got2):
        assert np.array_equal(r, g)
    # and the limits of the built-in scratch are usable right up to the edge
    A3 = rng.integers(0, 256, (4096, 32), dtype=np.uint8); B3 = rng.integers(0, 256, (70, 32), dtype=np.uint8)
    ref3, got3 = _run_both(A3, B3)
    for r, g in zip(ref3, got3):
        assert np.array_equal(r, g)


def test_noalloc_agrees_with_the_oracle_and_takes_no_device_memory(oracle_lib):
    import torch
    from os1_b200 import matcher
    rng = np.random.default_rng(321)
    A = rng.integers(0, 256, (1500, 32), dtype=np.uint8); B = rng.integers(0, 256, (2600, 32), dtype=np.uint8)
    B[40] = A[3]; B[41] = A[3]
    dev = torch.device("cuda", 0)
    a = torch.from_numpy(A).to(dev); b = torch.from_numpy(B).to(dev)
    o = [torch.full((len(A),), -7, dtype=torch.int32, device=dev) for _ in range(3)]
    s = torch.cuda.current_stream(dev).cuda_stream
    call = lambda: matcher.hamming_knn2_noalloc_raw(a.data_ptr(), len(A), b.data_ptr(), len(B), *[t.data_ptr() for t in o], device=0, stream=s)
    call(); torch.cuda.synchronize()                       # first use on this device: the one-time address look-up
    free0, _ = torch.cuda.mem_get_info(dev)
    for _ in range(3):
        call()
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info(dev)
    assert free1 == free0, "the allocation-free entry point changed the device's free memory (%d -> %d bytes)" % (free0, free1)
    ref = oracle_lib.Oracle.hamming_knn2(A, B)
    for r, g in zip(ref, o):
        assert np.array_equal(r, g.cpu().numpy())


def test_noalloc_rejects_sizes_beyond_its_scratch():
    import torch
    import os1_b200
    from os1_b200 import matcher
    dev = torch.device("cuda", 0)
    a = torch.zeros((4097, 32), dtype=torch.uint8, device=dev); b = torch.zeros((16, 32), dtype=torch.uint8, device=dev)
    o = [torch.zeros(4097, dtype=torch.int32, device=dev) for _ in range(3)]
    with pytest.raises(os1_b200.Os1Error) as e:
        matcher.hamming_knn2_noalloc_raw(a.data_ptr(), 4097, b.data_ptr(), 16, *[t.data_ptr() for t in o])
    assert e.value.code == -3 and "4096" in str(e.value)
    with pytest.raises(os1_b200.Os1Error) as e:
        matcher.hamming_knn2_noalloc_raw(a.data_ptr(), 16, b.data_ptr(), 65537, *[t.data_ptr() for t in o])
    assert e.value.code == -3
