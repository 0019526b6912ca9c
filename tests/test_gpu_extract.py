"""GPU parity tests proper: the CUDA path, called through the C ABI, against the CPU oracle on the same seeded frames.

Bars (BASELINE.json north_star): keypoint coordinates / octaves / counts after distribution bit-exact (including the
reference's output ORDER under the canonical tie rule), FAST responses bit-exact, angles within 1e-3 degrees (we
assert bit-equality and report), descriptors bit-exact, pyramid bytes bit-exact."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from os1_b200 import synth  # noqa: E402

ANGLE_TOL_DEG = 1e-3


def gpu_ext(*a, **k):
    import os1_b200
    return os1_b200.ORBextractor(*a, **k)


def compare_full(oracle_lib, img, nf, sf=1.2, nl=8, ini=20, mn=7, stages=True):
    o = oracle_lib.Oracle(nf, sf, nl, ini, mn)
    ko, do = o.extract(img)
    e = gpu_ext(nf, sf, nl, ini, mn)
    kg, dg = e(img)
    if stages:
        for l in range(nl):
            assert np.array_equal(o.level_pixels(l), e.pyramid_level(l)), "pyramid level %d" % l
            assert np.array_equal(o.candidates(l), e.candidates(l)), "FAST candidates level %d" % l
            xo, _ = o.level_keypoints(l)
            assert np.array_equal(xo, e.level_keypoints(l)), "kept keypoints level %d" % l
    assert len(kg) == len(ko)
    assert np.array_equal(kg["x"], ko[:, 0]) and np.array_equal(kg["y"], ko[:, 1])
    assert np.array_equal(kg["size"], ko[:, 2])
    assert np.array_equal(kg["response"], ko[:, 4])
    assert np.array_equal(kg["octave"], ko[:, 5].astype(np.int32))
    assert (kg["class_id"] == -1).all()
    dang = np.abs(kg["angle"] - ko[:, 3]); dang = np.minimum(dang, 360 - dang)
    assert dang.max(initial=0) <= ANGLE_TOL_DEG
    moved = int((kg["angle"] != ko[:, 3]).sum())          # angle tolerance cases are counted and reported
    bad_desc = int((dg != do).any(1).sum())
    assert bad_desc <= moved, "descriptor rows differ where angles are bit-equal"
    assert moved == 0 and bad_desc == 0, "angles/descriptors expected bit-exact (moved=%d, desc=%d)" % (moved, bad_desc)
    e.close()
    return len(kg)


@pytest.mark.parametrize("shape,nf,seed", [((480, 640), 1000, 0), ((720, 1280), 1500, 1), ((1080, 1920), 2000, 2)])
def test_baseline_configs_bit_exact(oracle_lib, shape, nf, seed):
    n = compare_full(oracle_lib, synth.frame(shape[0], shape[1], seed), nf)
    assert n >= nf        # the quadtree overshoots nFeatures slightly


@pytest.mark.parametrize("name", ["vga", "720p", "small"])
def test_against_committed_reference_golden(golden_dir, name):
    """Directly against outputs of the unmodified reference (no oracle in the loop)."""
    g = np.load(os.path.join(golden_dir, "extract_%s.npz" % name))
    img = synth.frame(int(g["h"]), int(g["w"]), int(g["seed"]))
    e = gpu_ext(int(g["nfeatures"]), 1.2, 8, 20, 7)
    kp, desc = e(img)
    ref = g["kp"]
    assert len(kp) == len(ref)
    for i, f in enumerate(["x", "y", "size", "angle", "response"]):
        assert np.array_equal(kp[f], ref[:, i]), f
    assert np.array_equal(kp["octave"], ref[:, 5].astype(np.int32))
    assert np.array_equal(desc, g["desc"])
    assert [len(e.candidates(l)) for l in range(8)] == list(g["ncand"])


@pytest.mark.parametrize("shape,nf,sf,nl,ini,mn,seed", [
    ((479, 641), 700, 1.2, 8, 20, 7, 3),        # odd sizes, pitch not a multiple of 16
    ((300, 500), 500, 1.2, 6, 20, 7, 4),
    ((200, 260), 300, 1.2, 4, 20, 7, 5),
    ((480, 640), 1000, 1.5, 5, 20, 7, 6),       # another scale factor
    ((480, 640), 250, 1.2, 8, 40, 12, 7),       # other thresholds, few features
    ((360, 1000), 800, 1.2, 5, 20, 7, 8),       # 3 quadtree roots
    ((480, 640), 3000, 1.2, 8, 20, 7, 9),       # more features than good candidates in the small levels
    ((85, 85), 60, 1.2, 1, 20, 7, 11),          # one FAST cell of 59 x 59: the widest and tallest a cell can be
    ((80, 100), 100, 1.2, 2, 20, 7, 12),        # cells wider than 32 columns, one cell row taller than 32
    ((220, 800), 400, 1.2, 3, 20, 7, 13),       # four quadtree roots (the widest aspect supported)
    ((2160, 3840), 4000, 1.2, 8, 20, 7, 21),    # 4K, 4000 features
    ((1080, 1920), 2000, 1.2, 12, 20, 7, 22),   # twelve levels (the maximum)
    ((600, 800), 1000, 1.1, 10, 20, 7, 23),     # fine scale steps
    ((1000, 1000), 1500, 1.3, 6, 20, 7, 24),    # square image, coarse scale steps
    ((1231, 1319), 873, 1.25, 2, 25, 28, 1194),  # a keypoint whose cosf is not the correctly rounded cosine (found by scripts/fuzz_extract.py)
    ((762, 1004), 2302, 1.41, 1, 31, 5, 1198),   # the same for sinf; one level with 2302 features
])
def test_other_geometries(oracle_lib, shape, nf, sf, nl, ini, mn, seed):
    compare_full(oracle_lib, synth.frame(shape[0], shape[1], seed), nf, sf, nl, ini, mn)


def test_getters_match_reference_tables(oracle_lib):
    e = gpu_ext(1000, 1.2, 8, 20, 7)
    t = oracle_lib.Oracle(1000).tables()
    assert e.GetLevels() == 8 and e.GetScaleFactor() == np.float32(1.2)
    assert np.array_equal(e.GetScaleFactors(), t["scale"]) and np.array_equal(e.GetInverseScaleFactors(), t["inv"])
    assert np.array_equal(e.GetScaleSigmaSquares(), t["sigma2"]) and np.array_equal(e.GetInverseScaleSigmaSquares(), t["invsigma2"])
    assert np.array_equal(e.features_per_level(), t["quota"])


def test_empty_and_textureless_images():
    e = gpu_ext(1000, 1.2, 8, 20, 7)
    kp, d = e(np.zeros((0, 0), np.uint8))
    assert len(kp) == 0 and d.shape == (0, 32)                 # if(_image.empty()) return;
    kp, d = e(np.full((480, 640), 77, np.uint8))               # no corner anywhere: descriptors.release()
    assert len(kp) == 0 and d.shape == (0, 32)
    kp, d = e(synth.frame(480, 640, 0))                         # and the handle still works afterwards
    assert len(kp) > 900


def test_levels_too_small_for_a_cell_row_yield_nothing(oracle_lib):
    """100x120 with 8 levels: levels below 62 px have no FAST cell row/column; the reference's loops simply do not
    run there (ORBextractor.cc:826,:837), so those levels are built but contribute no keypoints."""
    compare_full(oracle_lib, synth.frame(100, 120, 0), 500)
    compare_full(oracle_lib, synth.frame(70, 200, 1), 200, nl=3)


def test_unsupported_configurations_fail_loudly():
    import os1_b200
    e = gpu_ext(1000, 1.2, 8, 20, 7)
    with pytest.raises(os1_b200.Os1Error) as ex:
        e(synth.frame(400, 150, 0))                              # taller than 2:1 -> the reference's nIni is 0 (:575)
    assert ex.value.code == -4
    with pytest.raises(TypeError):
        e(np.zeros((480, 640), np.float32))                      # assert(image.type() == CV_8UC1)  (:914)


def test_non_contiguous_rows(oracle_lib):
    big = synth.frame(500, 700, 12)
    view = big[10:490, 30:670]                                   # step 700, cols 640
    ko, do = oracle_lib.Oracle(1000).extract(np.ascontiguousarray(view))
    kg, dg = gpu_ext(1000, 1.2, 8, 20, 7)(view)
    assert np.array_equal(kg["x"], ko[:, 0]) and np.array_equal(dg, do)


def test_batch_equals_single_and_is_deterministic(oracle_lib):
    frames = synth.video(5, 480, 640, seed=21)
    e = gpu_ext(1000, 1.2, 8, 20, 7)
    kp, desc, counts = e.extract_batch(frames)
    kp2, desc2, counts2 = e.extract_batch(frames)
    assert np.array_equal(counts, counts2) and np.array_equal(desc, desc2) and np.array_equal(kp, kp2)
    o = oracle_lib.Oracle(1000)
    for f in range(len(frames)):
        ko, do = o.extract(frames[f])
        n = counts[f]
        assert n == len(ko)
        assert np.array_equal(kp[f, :n]["x"], ko[:, 0]) and np.array_equal(kp[f, :n]["y"], ko[:, 1])
        assert np.array_equal(kp[f, :n]["angle"], ko[:, 3]) and np.array_equal(desc[f, :n], do)


def test_device_resident_batch_path(oracle_lib):
    """The bench's path: frames and outputs live in device memory (torch is only the allocator here)."""
    import torch
    frames = synth.video(3, 720, 1280, seed=33)
    e = gpu_ext(1500, 1.2, 8, 20, 7)
    cap = e.max_keypoints()
    fd = torch.from_numpy(frames).cuda()
    kp = torch.zeros((3, cap, 28), dtype=torch.uint8, device="cuda"); desc = torch.zeros((3, cap, 32), dtype=torch.uint8, device="cuda")
    cnt = torch.zeros(3, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    e.extract_batch_raw(fd.data_ptr(), 3, 720, 1280, 1280, 720 * 1280, True, kp.data_ptr(), desc.data_ptr(), cap, cnt.data_ptr(), True)
    e.synchronize()
    from os1_b200 import KEYPOINT_DTYPE
    o = oracle_lib.Oracle(1500)
    for f in range(3):
        ko, do = o.extract(frames[f])
        n = int(cnt[f])
        assert n == len(ko)
        k = kp[f, :n].cpu().numpy().view(KEYPOINT_DTYPE).reshape(-1)
        assert np.array_equal(k["x"], ko[:, 0]) and np.array_equal(k["angle"], ko[:, 3])
        assert np.array_equal(desc[f, :n].cpu().numpy(), do)


def test_full_size_properties_1080p_batch():
    """BASELINE full size, size-independent properties: determinism across batch positions (the same frame at
    different batch slots gives identical bytes), level-major ordering, border margins, quota overshoot bounds."""
    fr = synth.frame(1080, 1920, 40)
    frames = np.stack([fr, synth.frame(1080, 1920, 41), fr, fr])
    e = gpu_ext(2000, 1.2, 8, 20, 7)
    kp, desc, counts = e.extract_batch(frames)
    assert counts[0] == counts[2] == counts[3]
    n = counts[0]
    assert np.array_equal(kp[0, :n], kp[2, :n]) and np.array_equal(desc[0, :n], desc[3, :n])
    k = kp[0, :n]
    assert (np.diff(k["octave"]) >= 0).all()                     # levels concatenated in order (:936-967)
    quota = e.features_per_level()
    sc = e.GetScaleFactors()
    for l in range(8):
        m = k["octave"] == l
        assert quota[l] <= m.sum() <= quota[l] + 2               # Appendix C: N, N+1 or N+2
        x = k["x"][m] / sc[l]; y = k["y"][m] / sc[l]
        w, h = e.level_size(l)
        assert x.min() >= 19 - 1e-3 and y.min() >= 19 - 1e-3 and x.max() <= w - 20 + 1e-3 and y.max() <= h - 20 + 1e-3
    assert (k["response"] >= 7).all() and (k["angle"] >= 0).all() and (k["angle"] < 360).all()


@pytest.mark.parametrize("shape,nf,sf,nl,seed", [((240, 320), 500, 1.2, 8, 70), ((333, 517), 700, 1.2, 8, 71), ((300, 420), 400, 1.5, 5, 72),
                                                 ((131, 257), 300, 1.2, 6, 73)])
def test_per_level_pyramid_launches_for_batches(oracle_lib, monkeypatch, shape, nf, sf, nl, seed):
    """Large batches (from about 32 M pixels per call; here from 64 frames) build the pyramid with one launch per level (csrc/pyramid.cu: pyramid_level_kernel) instead of
    the fused chain: same bytes at every level (tiles cut by the image edge, images narrower than a tile, a tile row of one
    pixel row), and the same keypoints and descriptors as the frames extracted one by one."""
    import torch
    from os1_b200 import KEYPOINT_DTYPE
    monkeypatch.setenv("OS1_PYR_LEVELS_MIN", "64")          # (by default the switch is at about 32 M pixels per call)
    B = 66
    frames = synth.video(B, shape[0], shape[1], seed=seed)
    e = gpu_ext(nf, sf, nl, 20, 7)
    cap = e.max_keypoints()
    fd = torch.from_numpy(frames).cuda()
    kp = torch.zeros((B, cap, 28), dtype=torch.uint8, device="cuda"); desc = torch.zeros((B, cap, 32), dtype=torch.uint8, device="cuda")
    cnt = torch.zeros(B, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    e.extract_batch_raw(fd.data_ptr(), B, shape[0], shape[1], shape[1], shape[0] * shape[1], True, kp.data_ptr(), desc.data_ptr(), cap, cnt.data_ptr(), True)
    e.synchronize()
    assert e.launches_per_call() == (nl - 1) + 3, "the batch did not take the per-level launches"
    o = oracle_lib.Oracle(nf, sf, nl, 20, 7)
    for f in (0, 1, B // 2, B - 1):
        ko, do = o.extract(frames[f])
        for l in range(nl):
            assert np.array_equal(o.level_pixels(l), e.pyramid_level(l, f)), "frame %d pyramid level %d" % (f, l)
        n = int(cnt[f])
        assert n == len(ko)
        k = kp[f, :n].cpu().numpy().view(KEYPOINT_DTYPE).reshape(-1)
        assert np.array_equal(k["x"], ko[:, 0]) and np.array_equal(k["y"], ko[:, 1]) and np.array_equal(k["angle"], ko[:, 3])
        assert np.array_equal(desc[f, :n].cpu().numpy(), do)
    # the fused chain on the same frames (a batch below the threshold)
    kq, dq, cq = e.extract_batch(frames[:5])
    assert e.launches_per_call() == 4
    for f in range(5):
        n = int(cnt[f])
        assert cq[f] == n and np.array_equal(dq[f, :n], desc[f, :n].cpu().numpy())
    # host frames: the chunked pipeline (chunks of up to 64 frames, so both paths inside one call)
    kh, dh, ch = e.extract_batch(frames)
    assert np.array_equal(ch, cnt.cpu().numpy())
    dd = desc.cpu().numpy()
    for f in range(B):
        assert np.array_equal(dh[f, :ch[f]], dd[f, :ch[f]]), "frame %d" % f
    e.close()


@pytest.mark.parametrize("shape,nf", [((1080, 1920), 2000), ((720, 1280), 1500), ((481, 643), 1000)])
def test_per_level_and_fused_pyramids_agree_at_full_size(oracle_lib, monkeypatch, shape, nf):
    """BASELINE sizes, 64 frames on the device: the per-level launches (whole batch) and the fused chain (the same frames in
    batches of 16) are two independent kernels — every keypoint and descriptor byte of every frame must agree, and the pyramid bytes
    of a frame equal the oracle's."""
    import torch
    monkeypatch.setenv("OS1_PYR_LEVELS_MIN", "64")
    B = 64
    frames = synth.video(B, shape[0], shape[1], seed=90)
    e = gpu_ext(nf, 1.2, 8, 20, 7)
    cap = e.max_keypoints()
    fd = torch.from_numpy(frames).cuda()
    out = []
    for step in (B, 16):
        kp = torch.zeros((B, cap, 28), dtype=torch.uint8, device="cuda"); desc = torch.zeros((B, cap, 32), dtype=torch.uint8, device="cuda")
        cnt = torch.zeros(B, dtype=torch.int32, device="cuda")
        torch.cuda.synchronize()
        for f0 in range(0, B, step):
            e.extract_batch_raw(fd[f0:].data_ptr(), step, shape[0], shape[1], shape[1], shape[0] * shape[1], True,
                                kp[f0:].data_ptr(), desc[f0:].data_ptr(), cap, cnt[f0:].data_ptr(), True)
            e.synchronize()
            assert e.launches_per_call() == (10 if step == B else 4)
        if step == B:
            o = oracle_lib.Oracle(nf, 1.2, 8, 20, 7)
            o.extract(frames[B - 1])
            for l in range(8):
                assert np.array_equal(o.level_pixels(l), e.pyramid_level(l, B - 1)), "pyramid level %d" % l
        out.append((kp.cpu().numpy(), desc.cpu().numpy(), cnt.cpu().numpy()))
    (k0, d0, c0), (k1, d1, c1) = out
    assert np.array_equal(c0, c1) and c0.min() > 0
    for f in range(B):
        n = int(c0[f])
        assert np.array_equal(k0[f, :n], k1[f, :n]) and np.array_equal(d0[f, :n], d1[f, :n]), "frame %d" % f
    e.close()


def test_pyramid_path_switch_by_pixels_per_call():
    """Without the override the per-level launches start at about 32 M pixels per call: 16 frames of 1920x1080."""
    import torch
    assert "OS1_PYR_LEVELS_MIN" not in os.environ
    frames = torch.from_numpy(synth.video(16, 1080, 1920, seed=91)).cuda()
    e = gpu_ext(2000, 1.2, 8, 20, 7)
    cap = e.max_keypoints()
    kp = torch.zeros((16, cap, 28), dtype=torch.uint8, device="cuda"); desc = torch.zeros((16, cap, 32), dtype=torch.uint8, device="cuda")
    cnt = torch.zeros(16, dtype=torch.int32, device="cuda")
    res = []
    for n, launches in ((16, 10), (15, 4)):
        e.extract_batch_raw(frames.data_ptr(), n, 1080, 1920, 1920, 1080 * 1920, True, kp.data_ptr(), desc.data_ptr(), cap, cnt.data_ptr(), True)
        e.synchronize()
        assert e.launches_per_call() == launches
        res.append((cnt[:15].cpu().numpy().copy(), desc[:15].cpu().numpy().copy()))
    assert np.array_equal(res[0][0], res[1][0])
    for f in range(15):
        assert np.array_equal(res[0][1][f, :res[0][0][f]], res[1][1][f, :res[0][0][f]])
    e.close()
