"""CPU tests of the host-side logic: weight-file synthesis / parsing, segment planning."""
import os
import struct

import numpy as np
import pytest

import util
from demucs_cpp_b200 import weights
from oracle import demucs_oracle as O

# file sizes the reference's own converter produces for these architectures
EXPECT = {"htdemucs_4s": (533, 83994361), "htdemucs_6s": (525, 54855129), "hdemucs_mmi": (395, 167294211)}


@pytest.mark.parametrize("arch", list(EXPECT))
def test_manifest_matches_reference_format(arch, tmp_path):
    man = weights.load_manifest(arch)
    assert man["n_tensors"] == EXPECT[arch][0]
    size = 4
    for t in man["tensors"]:
        size += 8 + 4 * len(t["shape"]) + len(t["name"].encode()) + 2 * int(np.prod(t["shape"]))
    assert size == EXPECT[arch][1]


def test_synth_write_read_roundtrip(tmp_path):
    p = str(tmp_path / "w.bin")
    weights.make_weights("htdemucs_6s", p, seed=3)
    assert os.path.getsize(p) == EXPECT["htdemucs_6s"][1]
    with open(p, "rb") as f:
        assert struct.unpack("<I", f.read(4))[0] == 0x646D6336
    magic, tens = weights.read_ggml(p)
    assert magic == weights.MAGIC["htdemucs_6s"] and len(tens) == 525
    again = dict(weights.synth_tensors("htdemucs_6s", seed=3))
    for k in ("encoder.0.conv.weight", "crosstransformer.layers.1.cross_attn.in_proj_weight"):
        np.testing.assert_array_equal(tens[k], again[k].astype(np.float32))
    assert tens["decoder.3.conv_tr.weight"].shape == (48, 24, 8)           # 6 sources * 4 CaC channels
    assert tens["tdecoder.3.conv_tr.weight"].shape == (48, 12, 8)
    # stress variant only touches LayerScale tensors
    st = dict(weights.synth_tensors("htdemucs_6s", seed=3, stress=True))
    assert np.all(st["encoder.0.dconv.layers.0.6.scale"] == np.float16(0.3))
    np.testing.assert_array_equal(st["encoder.0.conv.weight"], again["encoder.0.conv.weight"])


def test_segment_plan_matches_reference_arithmetic():
    # 3-minute track, shift 4033: 31 chunks, the last one 216467 samples
    L, shift = 7938000, 4033
    plan = O.segment_plan(L + O.MAX_SHIFT - shift)
    assert len(plan) == 31 and plan[-1][1] == 216467
    plan10 = O.segment_plan(441000 + O.MAX_SHIFT - shift)
    assert [n for _, n in plan10] == [343980, 201032]


def test_num_segments_cabi_host_function():
    # dmx_num_segments is pure host arithmetic and must agree with the oracle's plan for ragged lengths
    from demucs_cpp_b200 import api
    for L, shift in [(7938000, 4033), (441000, 4033), (343980, 0), (257985, 22049), (2, 0), (1031941, 12436)]:
        assert api.num_segments(L, shift) == len(O.segment_plan(L + O.MAX_SHIFT - shift))


def test_triangle_weight():
    w = O.triangle_weight()
    assert w.shape == (util.SEG,) and w.max() == 1.0 and w[0] == w[-1] == np.float32(1.0 / 171990)
    assert np.all(w[:171990] == w[171990:][::-1])
