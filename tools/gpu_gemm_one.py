"""One GEMM shape through dmx_debug_linear (for ncu captures).  Usage: gpu_gemm_one.py M N K prec"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from demucs_cpp_b200 import api  # noqa: E402

M, N, K, prec = (int(a) for a in sys.argv[1:5])
rng = np.random.default_rng(0)
A = rng.standard_normal((M, K), dtype=np.float32)
W = (rng.standard_normal((N, K), dtype=np.float32) / np.sqrt(K)).astype(np.float32)
_, ms = api.debug_linear(A, W, None, act=0, precision=prec, iters=3)
print("M=%d N=%d K=%d prec=%d: %.3f ms, %.1f TFLOP/s" % (M, N, K, prec, ms, 2.0 * M * N * K / ms / 1e9))
