"""Scratch driver (GPU box): the TMA-fed convolution kernel (dmx_debug_conv) against numpy on bf16-rounded operands, one
case per subprocess so a trapped kernel does not take the other cases down.  `python tools/gpu_conv_check.py` runs all
cases; `python tools/gpu_conv_check.py <case index>` runs one in this process."""
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

T3x3 = [(a, b) for a in (-1, 0, 1) for b in (-1, 0, 1)]
K3 = [(0, -1), (0, 0), (0, 1)]
K3D2 = [(0, -2), (0, 0), (0, 2)]
CT = [(-1, 0), (0, 0)]
ONE = [(0, 0)]
#        B  U   V    Cin  N    taps  act glu share
CASES = [(1, 2, 128, 64, 64, ONE, 0, 0, 1),
         (1, 2, 336, 64, 64, ONE, 0, 0, 1),
         (2, 3, 336, 64, 96, T3x3, 0, 0, 0),
         (2, 3, 336, 64, 96, T3x3, 0, 0, 1),
         (2, 4, 336, 48, 96, T3x3, 0, 1, 0),
         (2, 4, 336, 48, 96, T3x3, 0, 1, 1),
         (1, 4, 336, 96, 192, T3x3, 0, 1, 1),
         (1, 4, 336, 192, 384, T3x3, 0, 1, 1),
         (1, 8, 336, 384, 768, T3x3, 1, 1, 1),
         (2, 1, 1344, 384, 768, K3, 0, 1, 1),
         (1, 1, 5375, 96, 24, K3D2, 0, 0, 1),
         (1, 5, 336, 96, 192, CT, 1, 0, 1),
         (1, 1, 85995, 48, 96, K3, 0, 1, 1)]


def bf16_round(x):
    u = np.ascontiguousarray(x, dtype=np.float32).view(np.uint32).astype(np.uint64)
    u = (u + 0x7FFF + ((u >> 16) & 1)) >> 16 << 16
    return u.astype(np.uint32).view(np.float32)


def reference(img, W, taps, bias, act, glu):
    B, U, V, C = img.shape
    N = W.shape[0]
    a = bf16_round(img).astype(np.float64)
    w = bf16_round(W).astype(np.float64).reshape(N, len(taps), C)
    out = np.zeros((B, U, V, N))
    for j, (du, dv) in enumerate(taps):
        sh = np.zeros_like(a)
        u0, u1 = max(0, -du), min(U, U - du)
        v0, v1 = max(0, -dv), min(V, V - dv)
        if u1 > u0 and v1 > v0:
            sh[:, u0:u1, v0:v1] = a[:, u0 + du:u1 + du, v0 + dv:v1 + dv]
        out += sh @ w[:, j, :].T
    out += bias
    if act:
        from scipy.special import erf
        out = 0.5 * out * (1.0 + erf(out / np.sqrt(2.0)))
    if glu:
        out = out[..., :N // 2] / (1.0 + np.exp(-out[..., N // 2:]))
    return out


def run_case(i):
    from demucs_cpp_b200 import api
    B, U, V, C, N, taps, act, glu, share = CASES[i]
    rng = np.random.default_rng(100 + i)
    img = rng.standard_normal((B, U, V, C)).astype(np.float32)
    W = (rng.standard_normal((N, len(taps) * C)) / np.sqrt(len(taps) * C)).astype(np.float32)
    bias = rng.standard_normal(N).astype(np.float32)
    out, ms = api.debug_conv(img, W, taps, bias, act, glu, share, iters=5)
    ref = reference(img, W, taps, bias, act, glu)
    err = np.abs(out - ref).max() / np.abs(ref).max()
    flops = 2.0 * B * U * V * N * len(taps) * C
    print("case %2d B%d U%d V%d C%d N%d taps%d act%d glu%d share%d: max rel err %.3e  %.3f ms  %.1f TFLOP/s %s"
          % (i, B, U, V, C, N, len(taps), act, glu, share, err, ms, flops / ms / 1e9, "OK" if err < 2e-4 else "FAIL"), flush=True)
    return err < 2e-4


if __name__ == "__main__":
    if len(sys.argv) > 1:
        sys.exit(0 if run_case(int(sys.argv[1])) else 1)
    bad = 0
    for i in range(len(CASES)):
        try:
            r = subprocess.run([sys.executable, os.path.abspath(__file__), str(i)], timeout=180, capture_output=True, text=True)
            sys.stdout.write(r.stdout)
            if r.returncode != 0:
                bad += 1
                sys.stdout.write("case %d rc=%d %s\n" % (i, r.returncode, r.stderr[-400:]))
        except subprocess.TimeoutExpired:
            bad += 1
            print("case %d TIMEOUT" % i)
    print("failed cases:", bad)
