"""One batch of --batch segments (default 16, what bench.py launches) of the bench workload, bracketed by cudaProfilerStart/Stop, for ncu:
  ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file X python tools/profile_step.py
Same model, precision and segments-per-launch as bench.py; the track is cut to exactly 8 segments."""
import argparse
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from demucs_cpp_b200 import api, weights  # noqa: E402
from demucs_cpp_b200.synth import synth_audio  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--precision", default="bf16")
ap.add_argument("--batch", type=int, default=16)
ap.add_argument("--arch", default="htdemucs_4s")
args = ap.parse_args()
L = args.batch * 257985 - 22050 + 4033 - 1000          # -> exactly --batch segments at shift 4033
assert api.num_segments(L, 4033) == args.batch
m = api.DemucsModel(weights.cached_weights(args.arch, 0), max_batch=args.batch,
                    precision=api.PRECISION_BF16 if args.precision == "bf16" else api.PRECISION_FP32)
d_audio = torch.from_numpy(np.ascontiguousarray(synth_audio(L).T)).cuda()
d_out = torch.empty((L, 2, m.nb_sources), dtype=torch.float32, device="cuda")
for _ in range(2):
    m.demucs_inference_device(d_audio.data_ptr(), L, d_out.data_ptr(), 4033)
torch.cuda.synchronize()
m.profile_begin()
torch.cuda.profiler.start()
m.demucs_inference_device(d_audio.data_ptr(), L, d_out.data_ptr(), 4033)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
seq = m.profile_sequence()
prof = m.profile_end()
import json
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
with open(os.path.join(ROOT, "gpurun_out", "profile_step_classes.json"), "w") as f:
    json.dump({"sequence": seq, "classes": prof}, f)
for name, c in sorted(prof.items(), key=lambda kv: -kv[1]["ms"]):
    if c["launches"]:
        print("  %-18s %8.3f ms  %5d launches" % (name, c["ms"], c["launches"]))
print("device ms for %d segments: %.2f" % (args.batch, m.last_device_ms))
