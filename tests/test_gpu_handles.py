"""GPU: handle lifecycle — what the reference does with its two long-lived extractors (Tracking.cc:65,67): coexisting
handles, calls from two host threads on distinct handles, a handle re-planning when the resolution changes, stream
hand-over, reserve(), and loud capacity errors."""
import threading

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from os1_b200 import synth  # noqa: E402


def test_two_extractors_coexist_and_interleave(oracle_lib):
    import os1_b200
    a = os1_b200.ORBextractor(1000, 1.2, 8, 20, 7)          # mpORBextractorLeft
    b = os1_b200.ORBextractor(2000, 1.2, 8, 20, 7)          # mpIniORBextractor (2 * nFeatures)
    img = synth.frame(480, 640, 30)
    ra = oracle_lib.Oracle(1000).extract(img); rb = oracle_lib.Oracle(2000).extract(img)
    for _ in range(3):
        ka, da = a(img); kb, db = b(img)
        assert np.array_equal(da, ra[1]) and np.array_equal(db, rb[1])
        assert np.array_equal(ka["x"], ra[0][:, 0]) and np.array_equal(kb["x"], rb[0][:, 0])


def test_distinct_handles_from_two_threads(oracle_lib):
    import os1_b200
    imgs = [synth.frame(480, 640, 31), synth.frame(720, 1280, 32)]
    nf = [1000, 1500]
    refs = [oracle_lib.Oracle(n).extract(i) for n, i in zip(nf, imgs)]
    errs = []

    def work(t):
        try:
            e = os1_b200.ORBextractor(nf[t], 1.2, 8, 20, 7)
            for _ in range(10):
                k, d = e(imgs[t])
                assert np.array_equal(d, refs[t][1]) and np.array_equal(k["angle"], refs[t][0][:, 3])
        except Exception as ex:        # surfaced in the main thread below
            errs.append(ex)
    ts = [threading.Thread(target=work, args=(t,)) for t in range(2)]
    [t.start() for t in ts]; [t.join() for t in ts]
    assert not errs, errs


def test_resolution_change_replans(oracle_lib):
    import os1_b200
    e = os1_b200.ORBextractor(1000, 1.2, 8, 20, 7)
    o = oracle_lib.Oracle(1000)
    for shape, seed in [((480, 640), 1), ((720, 1280), 2), ((480, 640), 3), ((300, 400), 4)]:
        img = synth.frame(shape[0], shape[1], seed)
        k, d = e(img); ko, do = o.extract(img)
        assert len(k) == len(ko) and np.array_equal(d, do)
    # batch after single, larger batch after smaller (buffers grow, results unchanged)
    fr = synth.video(6, 480, 640, seed=9)
    kp2, d2, c2 = e.extract_batch(fr[:2]); kp6, d6, c6 = e.extract_batch(fr)
    assert np.array_equal(c2, c6[:2]) and np.array_equal(d2[0, :c2[0]], d6[0, :c6[0]])


def test_caller_capacity_too_small_is_an_error():
    import ctypes as C
    import os1_b200
    from os1_b200 import _lib
    e = os1_b200.ORBextractor(1000, 1.2, 8, 20, 7)
    img = synth.frame(480, 640, 0)
    kp = np.zeros(100, os1_b200.KEYPOINT_DTYPE); desc = np.zeros((100, 32), np.uint8); n = C.c_int(-5)
    rc = e._L.os1_orb_extract(e._h, img.ctypes.data_as(C.c_void_p), 480, 640, 640, kp.ctypes.data_as(C.c_void_p), 100,
                              desc.ctypes.data_as(C.c_void_p), C.byref(n))
    assert rc == _lib.OS1_E_CAPACITY and b"capacity" in e._L.os1_last_error(e._h)
    k, d = e(img)                                               # and the handle is still usable
    assert len(k) > 900


def test_reserve_and_foreign_stream(oracle_lib):
    import torch
    import os1_b200
    e = os1_b200.ORBextractor(1000, 1.2, 8, 20, 7)
    e.reserve(480, 640, 4)
    s = torch.cuda.Stream()
    e.set_stream(s.cuda_stream)
    fr = synth.video(4, 480, 640, seed=12)
    cap = e.max_keypoints()
    with torch.cuda.stream(s):
        fd = torch.from_numpy(fr).cuda(non_blocking=False)
        kp = torch.zeros((4, cap, 28), dtype=torch.uint8, device="cuda"); desc = torch.zeros((4, cap, 32), dtype=torch.uint8, device="cuda")
        cnt = torch.zeros(4, dtype=torch.int32, device="cuda")
        e.extract_batch_raw(fd.data_ptr(), 4, 480, 640, 640, 480 * 640, True, kp.data_ptr(), desc.data_ptr(), cap, cnt.data_ptr(), True)
    s.synchronize()
    e.set_stream(None)
    o = oracle_lib.Oracle(1000)
    for f in range(4):
        ko, do = o.extract(fr[f])
        assert int(cnt[f]) == len(ko) and np.array_equal(desc[f, :len(ko)].cpu().numpy(), do)


def test_latency_graph_replays_follow_the_input(oracle_lib):
    """From the third call of a geometry on, a single host frame runs as one CUDA graph: every replay must see the new
    pixels, and the graph must be rebuilt when a batch call reallocates the buffers it points to."""
    import os1_b200
    from os1_b200 import synth
    e = os1_b200.ORBextractor(800, 1.2, 8, 20, 7)
    o = oracle_lib.Oracle(800, 1.2, 8, 20, 7)
    frames = synth.video(8, 360, 480, seed=77, max_shift=10)

    def same(i):
        kg, dg = e(frames[i])
        r = o.extract(frames[i])
        return len(kg) == len(r[0]) and np.array_equal(dg, r[1]) and np.array_equal(kg["angle"], r[0][:, 3])

    assert all(same(i) for i in range(5))                      # plain, plain, capture, replay, replay
    kb, db, cb = e.extract_batch(frames)                       # grows the per-frame buffers: the graph is dropped
    for i in range(8):
        r = o.extract(frames[i])
        assert cb[i] == len(r[0]) and np.array_equal(db[i, :cb[i]], r[1])
    assert all(same(i) for i in (7, 6, 5, 4, 3))               # and rebuilt
    e.set_timing(True)                                         # timed calls bypass the graph
    assert same(0) and len(e.stage_times_ms()) == 4
    e.set_timing(False)
    assert same(1)


def test_feature_limit_is_reported_as_unsupported():
    """A level's quadtree state must fit one CTA's shared memory: too many features in one level is OS1_E_UNSUPPORTED at
    create time (or at the first frame, when the image's cell table tips it over), never a CUDA launch failure."""
    import os1_b200
    with pytest.raises(os1_b200.Os1Error) as ei:
        os1_b200.ORBextractor(6000, 1.2, 1, 20, 7)
    assert ei.value.code == -4
    e = os1_b200.ORBextractor(2200, 1.2, 1, 20, 7)            # fits
    kp, d = e(synth.frame(480, 640, 3))
    assert len(kp) > 1500
