"""GPU-box check of the GEMM kernels in isolation through dmx_debug_linear (fp32 and bf16/tcgen05)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from demucs_cpp_b200 import api  # noqa: E402


def bf16_round(x):
    u = np.ascontiguousarray(x, dtype=np.float32).view(np.uint32).astype(np.uint64)
    u = (u + 0x7FFF + ((u >> 16) & 1)) >> 16 << 16
    return u.astype(np.uint32).view(np.float32)


def gelu(x):
    from math import erf
    return 0.5 * x * (1 + np.vectorize(erf)(x / np.sqrt(2)))


def main():
    rng = np.random.default_rng(0)
    shapes = [(128, 16, 64), (128, 256, 64), (300, 48, 100), (257, 6, 6), (1000, 384, 512), (512, 2048, 512),
              (129, 96, 1536), (4000, 192, 192)]
    ok = True
    for (M, N, K) in shapes:
        A = rng.standard_normal((M, K)).astype(np.float32)
        W = (rng.standard_normal((N, K)) / np.sqrt(K)).astype(np.float32)
        b = rng.standard_normal(N).astype(np.float32)
        for prec in (0, 1):
            for act in (0, 1):
                out = api.debug_linear(A, W, b, act=act, precision=prec)
                if prec == 1:
                    ref = bf16_round(A).astype(np.float64) @ bf16_round(W).astype(np.float64).T + b
                else:
                    ref = A.astype(np.float64) @ W.astype(np.float64).T + b
                if act:
                    ref = gelu(ref)
                err = np.abs(out - ref).max() / np.abs(ref).max()
                good = err < 2e-5
                ok &= good
                print("M=%5d N=%5d K=%5d prec=%d act=%d  max_rel %.3e %s" % (M, N, K, prec, act, err, "" if good else "FAIL"))
    print("ALL OK" if ok else "FAILURES")
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
