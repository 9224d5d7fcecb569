"""Generates tests/golden/*.npz by running the REFERENCE ITSELF (oracle/_ref/libdemucs_ref.so, the
unmodified sevagh/demucs.cpp library compiled by oracle/Makefile) on seeded inputs and seeded
random-init weights (demucs_cpp_b200/weights.py).  Run in the build container only:

    make -C oracle && python tests/golden/make_golden.py

The reference's own tests hold no golden numbers for the layers (they print, never
assert), so these vectors -- outputs of the reference's compiled code -- are what pins both the CPU
oracle (tests/test_oracle_golden.py) and the CUDA path (tests/test_gpu_parity.py).

Each stage tensor is stored as NSUB evenly spaced samples of its (C, positions) flattening plus its
float64 sum and sum of squares.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import util  # noqa: E402
from demucs_cpp_b200 import weights  # noqa: E402
from oracle import ref  # noqa: E402

NSUB = 2048


def sub_idx(size):
    return np.linspace(0, size - 1, NSUB).astype(np.int64)


def summarize(a):
    a = np.ascontiguousarray(a, dtype=np.float32).reshape(-1)
    return a[sub_idx(a.size)].copy(), np.array([a.astype(np.float64).sum(), (a.astype(np.float64) ** 2).sum()])


def golden_segment(arch, stress, tag):
    path = weights.cached_weights(arch, 0, stress)
    m = ref.RefModel(path)
    mix = util.synth_audio(util.SEG)
    out = m.segment(mix)
    d = {}
    names = util.REF_TAPS_V3 if arch == "hdemucs_mmi" else \
        [n for n in util.STAGE_TAPS if not (arch != "htdemucs_4s" and n.endswith("_up"))]
    for n in names:
        v = util.ref_view_to_cfirst(n, m.view(n))
        d[n + ".sub"], d[n + ".stat"] = summarize(v)
        d[n + ".shape"] = np.array(v.shape)
    d["targets.sub"], d["targets.stat"] = summarize(out)
    d["targets.shape"] = np.array(out.shape)
    np.savez_compressed(os.path.join(HERE, "segment_%s.npz" % tag), **d)
    m.close()
    print("wrote segment_%s.npz" % tag)


def golden_separate():
    path = weights.cached_weights("htdemucs_4s", 0, False)
    m = ref.RefModel(path)
    audio = util.synth_audio(441000)           # the 10 s clip of BASELINE.json configs[0]
    out, shift = m.separate(audio, seed=1)     # srand(1) == fresh-process rand(): offset 4033
    sub, stat = summarize(out)
    np.savez_compressed(os.path.join(HERE, "separate_htdemucs_4s_10s.npz"), sub=sub, stat=stat,
                        shape=np.array(out.shape), shift=np.array([shift]))
    m.close()
    print("wrote separate_htdemucs_4s_10s.npz shift", shift)


def golden_dsp():
    rng = np.random.default_rng(7)
    n = 8192
    wave = rng.standard_normal((2, n)).astype(np.float32)
    spec = ref.stft(wave)                                   # (2, 2049, 9)
    spec_in = (rng.standard_normal((2, 2049, 9)) + 1j * rng.standard_normal((2, 2049, 9))).astype(np.complex64)
    wave_out = ref.istft(spec_in, n)
    # full-size segment STFT (the shape the model uses), subsampled
    seg = util.synth_audio(util.SEG)
    pad = np.concatenate([seg[:, :1536][:, ::-1], seg, seg[:, util.SEG - 1620:][:, ::-1]], axis=1)
    full = ref.stft(pad)[:, :2048, 2:338]                   # what src/model_inference.cpp:75-99 keeps
    fsub, fstat = summarize(np.stack([full[0].real, full[0].imag, full[1].real, full[1].imag]))
    np.savez_compressed(os.path.join(HERE, "dsp.npz"), wave=wave, spec=spec, spec_in=spec_in, wave_out=wave_out,
                        seg_spec_sub=fsub, seg_spec_stat=fstat)
    print("wrote dsp.npz")


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "v3":
        golden_segment("hdemucs_mmi", False, "hdemucs_mmi")
        golden_segment("hdemucs_mmi", True, "hdemucs_mmi_stress")
        sys.exit(0)
    golden_dsp()
    golden_segment("htdemucs_4s", False, "htdemucs_4s")
    golden_segment("htdemucs_4s", True, "htdemucs_4s_stress")
    golden_segment("htdemucs_6s", False, "htdemucs_6s")
    golden_separate()
    golden_segment("hdemucs_mmi", False, "hdemucs_mmi")
    golden_segment("hdemucs_mmi", True, "hdemucs_mmi_stress")
