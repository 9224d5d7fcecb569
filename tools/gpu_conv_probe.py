"""Scratch probes of the TMA conv kernel (GPU box)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from demucs_cpp_b200 import api
T3x3 = [(a, b) for a in (-1, 0, 1) for b in (-1, 0, 1)]
ONE = [(0, 0)]
P3 = [(-1, 0), (0, 0), (1, 0)]
K3 = [(0, -1), (0, 0), (0, 1)]
B = 8
rng = np.random.default_rng(0)
cases = [("C48 3x3", 512, 336, 48, 96, T3x3, 1, 1), ("C48 1x1", 512, 336, 48, 96, ONE, 1, 1), ("C48 3planes", 512, 336, 48, 96, P3, 1, 1),
         ("C48 k3", 512, 336, 48, 96, K3, 1, 1), ("C64 3x3", 512, 336, 64, 96, T3x3, 1, 1), ("C64 1x1", 512, 336, 64, 96, ONE, 1, 1),
         ("C64 k3", 512, 336, 64, 96, K3, 1, 1), ("C64 3planes", 512, 336, 64, 96, P3, 1, 1)]
only = sys.argv[1:]
for name, U, V, C, N, taps, share, glu in cases:
    if only and name not in only:
        continue
    img = rng.standard_normal((B, U, V, C)).astype(np.float32)
    W = (rng.standard_normal((N, len(taps) * C)) / np.sqrt(len(taps) * C)).astype(np.float32)
    out, ms = api.debug_conv(img, W, taps, None, 0, glu, share, iters=10)
    tiles = B * U * 3
    print("%-16s %.1f us  %.2f us/tile/SM" % (name, ms * 1e3, ms * 1e3 / (tiles / 148.0)), flush=True)
