"""Shared helpers for the parity tests (test infrastructure)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

SEG = 343980
STAGE_TAPS = ["x", "xt", "saved_0", "savedt_0", "saved_1", "savedt_1", "saved_2", "savedt_2", "saved_3", "savedt_3",
              "x_3_up", "xt_3_up", "x_3", "xt_3", "x_2", "xt_2", "x_1", "xt_1", "x_0", "xt_0", "x_out", "xt_out"]


from demucs_cpp_b200.synth import synth_audio  # noqa: E402,F401


STAGE_TAPS_V3 = ["x", "xt", "saved_0", "savedt_0", "saved_1", "savedt_1", "saved_2", "savedt_2", "saved_3", "savedt_3",
                 "saved_4", "x_shared_5", "x_4", "x_3", "xt_3", "x_2", "xt_2", "x_1", "xt_1", "x_0", "xt_0", "x_out", "xt_out"]
# the subset of the v3 taps that the reference's own scratch struct keeps after a forward (oracle/ref_harness.cpp)
REF_TAPS_V3 = ["x", "saved_0", "savedt_0", "saved_1", "savedt_1", "saved_2", "savedt_2", "saved_3", "savedt_3", "saved_4",
               "x_shared_5", "x_out", "xt_out"]


def max_rel_err(a, b):
    """max |a-b| / max |b|  -- the per-layer metric of BASELINE.md §4."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


def snr_db(a, b):
    """10 log10(|b|^2 / |a-b|^2) with b the reference."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return float(10 * np.log10((b ** 2).sum() / max(((a - b) ** 2).sum(), 1e-300)))


def tap_to_cfirst(a):
    """Device tap (B, d1, d2, C) channels-last -> (B, C, d1*d2)."""
    B, d1, d2, C = a.shape
    return a.transpose(0, 3, 1, 2).reshape(B, C, d1 * d2)


def oracle_tap_to_cfirst(name, t):
    """Oracle / reference tap in the reference's logical order -> (C, positions).
    Freq taps are (C, F, T); time taps are (1, C, L)."""
    a = np.asarray(t)
    if name.startswith("xt") or name.startswith("savedt") or name in ("x_shared_5", "x_4"):
        a = a[0]                      # (1, C, L) time-like tensors
    return a.reshape(a.shape[0], -1)


def ref_view_to_cfirst(name, a):
    """Reference harness views: like the oracle's, except the post-transformer freq tokens
    `x_3_up`, which the reference leaves as (1, C, T1*Fr) with token index t*Fr+f."""
    a = np.asarray(a)
    if name == "x_3_up":
        C = a.shape[1]
        return a[0].reshape(C, 336, 8).transpose(0, 2, 1).reshape(C, -1)
    return oracle_tap_to_cfirst(name, a)
