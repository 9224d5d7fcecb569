"""Scratch driver (GPU box): the TMA-fed convolution kernel at the decoder-rewrite shapes of htdemucs (8 segments)."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from demucs_cpp_b200 import api

T3x3 = [(a, b) for a in (-1, 0, 1) for b in (-1, 0, 1)]
K3 = [(0, -1), (0, 0), (0, 1)]
B = int(os.environ.get("NB", "8"))
SHAPES = [("dec0 rw", 8, 336, 384, 768, T3x3), ("dec1 rw", 32, 336, 192, 384, T3x3), ("dec2 rw", 128, 336, 96, 192, T3x3),
          ("dec3 rw", 512, 336, 48, 96, T3x3), ("tdec0 rw", 1, 1344, 384, 768, K3), ("tdec1 rw", 1, 5375, 192, 384, K3),
          ("tdec2 rw", 1, 21499, 96, 192, K3), ("tdec3 rw", 1, 85995, 48, 96, K3)]
only = sys.argv[1:] 
rng = np.random.default_rng(0)
for name, U, V, C, N, taps in SHAPES:
    if only and name.split()[0] not in only:
        continue
    img = rng.standard_normal((B, U, V, C)).astype(np.float32)
    W = (rng.standard_normal((N, len(taps) * C)) / np.sqrt(len(taps) * C)).astype(np.float32)
    bias = rng.standard_normal(N).astype(np.float32)
    out, ms = api.debug_conv(img, W, taps, bias, 0, 1, 1, iters=10)
    flops = 2.0 * B * U * V * N * len(taps) * C
    byt = B * U * V * (C * 2 + N // 2 * 4)
    print("%-9s B%d: %.1f us  %.0f TFLOP/s  %.0f GB/s" % (name, B, ms * 1e3, flops / ms / 1e9, byt / ms / 1e6), flush=True)
