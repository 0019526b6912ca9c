"""CPU (host-only entry points): os1's map-file records for descriptors and keypoints through the C ABI, against the outputs of
the reference's own Osmap::serialize / deserialize (src/Osmap.cpp:775-786, :819-831; sliced and compiled by oracle/refmatch/,
committed as tests/golden/matcher_wire.npz) — and the round trip at the sizes the extractor produces."""
import os

import numpy as np

from oracle import matcher_cases as mc
from os1_b200 import wire
from os1_b200.orb import KEYPOINT_DTYPE


def _kp_struct(kp4):
    kp = np.zeros(len(kp4), KEYPOINT_DTYPE)
    kp["x"] = kp4[:, 0]; kp["y"] = kp4[:, 1]; kp["angle"] = kp4[:, 2]; kp["octave"] = kp4[:, 3].astype(np.int32)
    kp["size"] = 31.0; kp["response"] = 40.0; kp["class_id"] = -1
    return kp


def test_records_equal_the_reference_serialiser(golden_dir):
    g = np.load(os.path.join(golden_dir, "matcher_wire.npz"))
    for s in g["seeds"]:
        c = mc.make_case("wire", int(s))
        blocks = wire.descriptor_blocks(c["desc"])
        assert blocks.dtype == np.uint32 and np.array_equal(blocks, g["blocks_%d" % s])
        assert np.array_equal(wire.descriptors_from_blocks(g["blocks_%d" % s]), g["back_%d" % s])
        assert np.array_equal(g["back_%d" % s], c["desc"])                      # the reference's own round trip is the identity
        x, y, o, a = wire.keypoint_records(_kp_struct(c["kp"]))
        ref = g["xya_%d" % s]
        assert np.array_equal(x, ref[:, 0]) and np.array_equal(y, ref[:, 1]) and np.array_equal(a, ref[:, 2])
        assert np.array_equal(o, g["octave_%d" % s])
        back = wire.keypoints_from_records(x, y, o, a)
        assert np.array_equal(back["x"], ref[:, 0]) and np.array_equal(back["octave"], g["octave_%d" % s])
        assert (back["size"] == 0).all() and (back["response"] == 0).all() and (back["class_id"] == -1).all()


def test_row_as_eight_native_words_round_trip():
    rng = np.random.default_rng(11)
    for n in (0, 1, 2018, 100000):
        d = rng.integers(0, 256, (n, 32), dtype=np.uint8)
        b = wire.descriptor_blocks(d)
        assert np.array_equal(b, d.view(np.uint32).reshape(-1, 8))               # `((unsigned int*)m.data)[i]`
        assert np.array_equal(wire.descriptors_from_blocks(b), d)
    one = np.zeros((1, 32), np.uint8); one[0, 0] = 1; one[0, 7] = 0x80
    assert wire.descriptor_blocks(one)[0, 0] == 1 and wire.descriptor_blocks(one)[0, 1] == 0x80000000   # little-endian hosts
