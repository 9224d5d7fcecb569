"""Model load time: file -> resident, repacked device weights, per architecture and precision."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from demucs_cpp_b200 import api, weights  # noqa: E402

import torch  # noqa: E402  (creates the CUDA context up front so it is not billed to the first load)

torch.zeros(1, device="cuda")
for arch in ("htdemucs_4s", "htdemucs_6s", "hdemucs_mmi"):
    path = weights.cached_weights(arch, 0)
    for prec, name in ((api.PRECISION_FP32, "fp32"), (api.PRECISION_BF16, "bf16")):
        ts = []
        for _ in range(3):
            t0 = time.perf_counter()
            m = api.DemucsModel(path, max_batch=1, precision=prec)
            m.sync()
            ts.append(time.perf_counter() - t0)
            m.close()
        print("%-12s %s  file %.0f MB  load %.3f s (first) %.3f s (best of 3)" % (arch, name, os.path.getsize(path) / 1e6, ts[0], min(ts)))
