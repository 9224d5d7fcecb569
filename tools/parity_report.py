"""Prints the measured stem SNRs of the CUDA path against the reference's golden vectors (tests/golden) for both
precision modes and writes them to gpurun_out/parity_report.json -- the numbers quoted in DESIGN.md / BASELINE.md.
Test infrastructure (uses the golden fixtures and, for the stress case, the CPU oracle); run on the GPU box."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import util  # noqa: E402
from demucs_cpp_b200 import api, weights  # noqa: E402
from oracle import demucs_oracle as O  # noqa: E402

NSUB = 2048


def sub_idx(size):
    return np.linspace(0, size - 1, NSUB).astype(np.int64)


rep = {}
gd = os.path.join(ROOT, "tests", "golden")
for arch, stress, tag in [("htdemucs_4s", False, "htdemucs_4s"), ("htdemucs_4s", True, "htdemucs_4s_stress"),
                          ("htdemucs_6s", False, "htdemucs_6s"), ("hdemucs_mmi", False, "hdemucs_mmi"),
                          ("hdemucs_mmi", True, "hdemucs_mmi_stress")]:
    g = np.load(os.path.join(gd, "segment_%s.npz" % tag))
    for prec, pname in ((api.PRECISION_FP32, "fp32"), (api.PRECISION_BF16, "bf16")):
        m = api.DemucsModel(weights.cached_weights(arch, 0, stress), max_batch=1, precision=prec)
        out = m.model_inference(util.synth_audio(util.SEG)[None])[0]
        snr = float(util.snr_db(out.reshape(-1)[sub_idx(out.size)], g["targets.sub"]))
        rep["%s/%s" % (tag, pname)] = round(snr, 1)
        print(tag, pname, "stems vs reference golden: %.1f dB" % snr)
        del m
# the stress case of the test-suite (seed 31 segment) against the full-tensor oracle
seg = util.synth_audio(util.SEG, seed=31)
_, tens = weights.read_ggml(weights.cached_weights("htdemucs_4s", 0, True))
with torch.no_grad():
    ref = O.segment_forward(O.Weights(tens, torch.float32), torch.from_numpy(seg)).numpy()
for prec, pname in ((api.PRECISION_FP32, "fp32"), (api.PRECISION_BF16, "bf16")):
    m = api.DemucsModel(weights.cached_weights("htdemucs_4s", 0, True), max_batch=1, precision=prec)
    out = m.model_inference(seg[None])[0]
    rep["htdemucs_4s_stress_seed31_full/%s" % pname] = round(float(util.snr_db(out, ref)), 1)
    print("stress seed31 full tensor", pname, rep["htdemucs_4s_stress_seed31_full/%s" % pname])
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(rep, open(os.path.join(ROOT, "gpurun_out", "parity_report.json"), "w"), indent=1)
