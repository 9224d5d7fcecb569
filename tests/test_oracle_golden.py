"""CPU tests pinning the ORACLE (oracle/demucs_oracle.py, the numpy/torch restatement of the
reference's hot path) against golden vectors produced by the reference's own compiled code
(tests/golden/make_golden.py via oracle/_ref).  Tolerances: the fp32 bar of BASELINE.md §4
(max relative error 1e-4 per stage, stems >= 80 dB SNR)."""
import os

import numpy as np
import pytest
import torch

import util
from demucs_cpp_b200 import weights
from oracle import demucs_oracle as O

TOL = 1e-4  # fp32-mode per-layer tolerance stated by BASELINE.json north_star


def sub_idx(size, nsub=2048):
    return np.linspace(0, size - 1, nsub).astype(np.int64)


def check_summary(name, full, gsub, gstat, tol=TOL):
    a = np.ascontiguousarray(full, dtype=np.float32).reshape(-1)
    sub = a[sub_idx(a.size)]
    scale = max(np.abs(gsub).max(), 1e-30)
    err = np.abs(sub.astype(np.float64) - gsub).max() / scale
    assert err <= tol, "%s: max rel err %.3e on golden samples" % (name, err)
    ss = (a.astype(np.float64) ** 2).sum()
    assert abs(ss - gstat[1]) <= 1e-3 * max(gstat[1], 1e-30), "%s: energy %.6e vs golden %.6e" % (name, ss, gstat[1])


@pytest.fixture(scope="module")
def torch_threads():
    torch.set_num_threads(os.cpu_count() or 4)


def test_stft_istft_against_reference(golden_dir, torch_threads):
    g = np.load(os.path.join(golden_dir, "dsp.npz"))
    spec = O.stft(torch.from_numpy(g["wave"])).numpy()
    assert spec.shape == g["spec"].shape
    assert util.max_rel_err(spec.real, g["spec"].real) <= TOL and util.max_rel_err(spec.imag, g["spec"].imag) <= TOL
    wave = O.istft(torch.from_numpy(g["spec_in"]), g["wave"].shape[1]).numpy()
    assert util.max_rel_err(wave, g["wave_out"]) <= TOL
    # the reference's own asserting test (test/test_dsp.cpp:106-196): STFT -> iSTFT round trip within 1e-4
    rt = O.istft(O.stft(torch.from_numpy(g["wave"])), g["wave"].shape[1]).numpy()
    assert np.abs(rt - g["wave"]).max() <= 1e-4


@pytest.mark.parametrize("arch,stress,tag", [("htdemucs_4s", False, "htdemucs_4s"),
                                             ("htdemucs_4s", True, "htdemucs_4s_stress"),
                                             ("htdemucs_6s", False, "htdemucs_6s"),
                                             ("hdemucs_mmi", False, "hdemucs_mmi"),
                                             ("hdemucs_mmi", True, "hdemucs_mmi_stress")])
def test_segment_forward_against_reference(golden_dir, torch_threads, arch, stress, tag):
    g = np.load(os.path.join(golden_dir, "segment_%s.npz" % tag))
    _, tens = weights.read_ggml(weights.cached_weights(arch, 0, stress))
    W = O.Weights(tens, torch.float32)
    taps = {}
    fwd = O.segment_forward_v3 if W.is_v3 else O.segment_forward
    with torch.no_grad():
        out = fwd(W, torch.from_numpy(util.synth_audio(util.SEG)), taps).numpy()
    for n in (util.STAGE_TAPS_V3 if W.is_v3 else util.STAGE_TAPS):
        if n + ".sub" not in g.files:
            continue
        check_summary(n, util.oracle_tap_to_cfirst(n, taps[n].numpy()), g[n + ".sub"], g[n + ".stat"])
    check_summary("targets", out, g["targets.sub"], g["targets.stat"])
    sub = out.reshape(-1)[sub_idx(out.size)]
    assert util.snr_db(sub, g["targets.sub"]) >= 80.0


def test_track_separation_against_reference(golden_dir, torch_threads):
    """BASELINE.json configs[0]: htdemucs 4s on the 10 s synthetic clip through demucs_inference."""
    g = np.load(os.path.join(golden_dir, "separate_htdemucs_4s_10s.npz"))
    assert int(g["shift"][0]) == 4033            # what an unseeded glibc rand() % 22050 gives
    _, tens = weights.read_ggml(weights.cached_weights("htdemucs_4s", 0))
    W = O.Weights(tens, torch.float32)
    with torch.no_grad():
        out = O.separate(W, util.synth_audio(441000), 4033).numpy()
    assert tuple(g["shape"]) == out.shape
    check_summary("stems", out, g["sub"], g["stat"])
    assert util.snr_db(out.reshape(-1)[sub_idx(out.size)], g["sub"]) >= 80.0
