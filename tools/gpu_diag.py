"""GPU-box diagnostic: runs the CUDA forward and the CPU oracle on the same seeded segments and prints
the per-stage parity table (max relative error, SNR).  Writes gpurun_out/diag_<arch>.json.
Usage: python tools/gpu_diag.py [--arch htdemucs_4s] [--stress] [--batch 2]"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import torch  # noqa: E402

import util  # noqa: E402
from demucs_cpp_b200 import api, weights  # noqa: E402
from oracle import demucs_oracle as O  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--arch", default="htdemucs_4s")
    ap.add_argument("--stress", action="store_true")
    ap.add_argument("--batch", type=int, default=2)
    ap.add_argument("--precision", type=int, default=0)
    args = ap.parse_args()
    torch.set_num_threads(os.cpu_count() or 8)
    path = weights.cached_weights(args.arch, 0, args.stress)
    _, tens = weights.read_ggml(path)
    W = O.Weights(tens, torch.float32)
    segs = []
    base = util.synth_audio(util.SEG + 1000 * args.batch)
    for b in range(args.batch):
        segs.append(base[:, 1000 * b:1000 * b + util.SEG] * (1.0 if b == 0 else 0.5 + 0.25 * b))
    segs = np.stack(segs)
    t0 = time.time()
    m = api.DemucsModel(path, max_batch=args.batch, precision=args.precision)
    print("model load %.2fs" % (time.time() - t0))
    t0 = time.time()
    out = m.model_inference(segs)
    print("gpu forward (first call) %.3fs" % (time.time() - t0))
    t0 = time.time()
    out = m.model_inference(segs)
    print("gpu forward (second call, incl. H2D/D2H) %.3fs, launches so far %d" % (time.time() - t0, m.launch_count))
    v3 = args.arch == "hdemucs_mmi"
    names = util.STAGE_TAPS_V3 if v3 else [n for n in util.STAGE_TAPS if not (args.arch != "htdemucs_4s" and n.endswith("_up"))]
    taps = {n: m.tap(n) for n in names}
    rows = []
    for b in range(args.batch):
        otaps = {}
        t0 = time.time()
        with torch.no_grad():
            oref = (O.segment_forward_v3 if v3 else O.segment_forward)(W, torch.from_numpy(segs[b]), otaps).numpy()
        print("oracle forward b=%d %.1fs" % (b, time.time() - t0))
        for n in taps:
            a = util.tap_to_cfirst(taps[n])[b]
            r = util.oracle_tap_to_cfirst(n, otaps[n].numpy())
            rows.append({"b": b, "tap": n, "max_rel": util.max_rel_err(a, r), "snr_db": util.snr_db(a, r),
                         "nan": bool(np.isnan(a).any())})
        rows.append({"b": b, "tap": "targets", "max_rel": util.max_rel_err(out[b], oref), "snr_db": util.snr_db(out[b], oref),
                     "nan": bool(np.isnan(out[b]).any())})
    for r in rows:
        print("b=%d %-10s max_rel %.3e  snr %7.1f dB %s" % (r["b"], r["tap"], r["max_rel"], r["snr_db"], "NaN!" if r["nan"] else ""))
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    tag = args.arch + ("_stress" if args.stress else "") + ("_p%d" % args.precision)
    with open(os.path.join(ROOT, "gpurun_out", "diag_%s.json" % tag), "w") as f:
        json.dump(rows, f, indent=1)


if __name__ == "__main__":
    main()
