"""GPU-box micro-benchmark of the GEMM kernels on the model's representative shapes (batch of 8 segments)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from demucs_cpp_b200 import api  # noqa: E402

SHAPES = [  # (name, M, N, K)
    ("ffn1      T=2688x8  512->2048", 21504, 2048, 512),
    ("ffn2      T=2688x8 2048-> 512", 21504, 512, 2048),
    ("qkv       T=2688x8  512->1536", 21504, 1536, 512),
    ("outproj   T=2688x8  512-> 512", 21504, 512, 512),
    ("fdec0 rewrite-like K=3456 N=768", 21504, 768, 3456),
    ("fenc1 conv-like K=384 N=96 ", 344064, 96, 384),
    ("fdec3 rewrite-like K=432 N=96", 1376256, 96, 432),
    ("dconv c2 K=48 N=768       ", 21504, 768, 48),
    ("dconv c2 K=12 N=192       ", 344064, 192, 12),
    ("dconv c2 K=6  N=96        ", 1376256, 96, 6),
    ("dconv c1-like K=144 N=6   ", 1376256, 6, 144),
]


def main():
    precs = [int(a) for a in sys.argv[1:]] or [1]
    rng = np.random.default_rng(0)
    for name, M, N, K in SHAPES:
        A = rng.standard_normal((M, K), dtype=np.float32)
        W = (rng.standard_normal((N, K), dtype=np.float32) / np.sqrt(K)).astype(np.float32)
        for prec in precs:
            _, ms = api.debug_linear(A, W, None, act=0, precision=prec, iters=10)
            fl = 2.0 * M * N * K
            by = 4.0 * M * (K + N)
            print("%-36s prec=%d  %8.3f ms  %8.1f TFLOP/s  %8.1f GB/s (A+C fp32)" % (name, prec, ms, fl / ms / 1e9, by / ms / 1e6))


if __name__ == "__main__":
    main()
