"""Synthetic stereo 44.1 kHz audio used by the benchmarks and tests: 220/330 Hz tones
plus 0.1*N(0,1) noise, numpy seed 1234.  There is no network for datasets."""
import numpy as np


def synth_audio(n, seed=1234):
    rng = np.random.default_rng(seed)
    t = np.arange(n) / 44100.0
    left = 0.3 * np.sin(2 * np.pi * 220 * t) + 0.1 * rng.standard_normal(n)
    right = 0.3 * np.sin(2 * np.pi * 330 * t) + 0.1 * rng.standard_normal(n)
    return np.stack([left, right]).astype(np.float32)
