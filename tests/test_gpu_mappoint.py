"""GPU: MapPoint::ComputeDistinctiveDescriptors for many map points in one launch through the C ABI,
against the oracle's restatement of src/MapPoint.cc:226-292."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from os1_b200 import matcher, Os1Error  # noqa: E402
from test_matcher_oracle import distinctive_case  # noqa: E402


@pytest.mark.parametrize("seed,npoints,max_obs", [(0, 200, 12), (1, 150, 40), (2, 40, 256), (3, 2, 1)])
def test_distinctive_descriptors_bit_exact(oracle_lib, seed, npoints, max_obs):
    desc, ptr = distinctive_case(seed, npoints, max_obs)
    bi, bm = matcher.distinctive_descriptors(desc, ptr)
    obi, obm = oracle_lib.Oracle.distinctive_descriptors(desc, ptr)
    assert np.array_equal(bi, obi) and np.array_equal(bm, obm)


def test_distinctive_descriptors_real_descriptors(oracle_lib):
    """Map points observed in several frames of a video: the same physical corner seen through integer shifts."""
    import os1_b200
    from os1_b200 import synth
    fr = synth.video(6, 480, 640, seed=9, max_shift=8)
    e = os1_b200.ORBextractor(500, 1.2, 8, 20, 7)
    kd = [e(f) for f in fr]
    k0, d0 = kd[0]
    lv0 = np.flatnonzero(k0["octave"] == 0)[:200]
    obs = []; ptr = [0]
    for i in lv0:                                   # observations = nearest level-0 descriptor in every other frame
        rows = [d0[i]]
        for k, d in kd[1:]:
            dist = np.unpackbits(d ^ d0[i], axis=1).sum(1)
            rows.append(d[int(dist.argmin())])
        obs.extend(rows); ptr.append(len(obs))
    desc = np.array(obs, np.uint8); ptr = np.array(ptr, np.int32)
    bi, bm = matcher.distinctive_descriptors(desc, ptr)
    obi, obm = oracle_lib.Oracle.distinctive_descriptors(desc, ptr)
    assert np.array_equal(bi, obi) and np.array_equal(bm, obm)


def test_distinctive_descriptors_limits_and_empty():
    bi, bm = matcher.distinctive_descriptors(np.zeros((0, 32), np.uint8), np.array([0, 0, 0], np.int32))
    assert list(bi) == [-1, -1] and list(bm) == [np.iinfo(np.int32).max] * 2
    with pytest.raises(Os1Error) as ei:
        matcher.distinctive_descriptors(np.zeros((257, 32), np.uint8), np.array([0, 257], np.int32))
    assert ei.value.code == -4
    with pytest.raises(Os1Error):
        matcher.distinctive_descriptors(np.zeros((4, 32), np.uint8), np.array([0, 3, 2], np.int32))      # not monotone


def test_distinctive_descriptors_device_pointers(oracle_lib):
    torch = pytest.importorskip("torch")
    from os1_b200 import _lib
    desc, ptr = distinctive_case(5, 300, 20)
    dd = torch.from_numpy(desc).cuda(); dp = torch.from_numpy(ptr).cuda()
    bi = torch.empty(len(ptr) - 1, dtype=torch.int32, device="cuda"); bm = torch.empty_like(bi)
    _lib.check(_lib.load().os1_distinctive_descriptors(dd.data_ptr(), dp.data_ptr(), len(ptr) - 1, bi.data_ptr(), bm.data_ptr(), 1, 0,
                                                       torch.cuda.current_stream().cuda_stream))
    obi, obm = oracle_lib.Oracle.distinctive_descriptors(desc, ptr)
    assert np.array_equal(bi.cpu().numpy(), obi) and np.array_equal(bm.cpu().numpy(), obm)
