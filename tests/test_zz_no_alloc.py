"""GPU test of dmx_stft_segments_noalloc, the entry point for callers that may not allocate device
memory at call time (real-time audio threads, fixed-pool hosts): its statistics workspace is a static
array inside the library rather than cudaMalloc'd memory.

The file name sorts last on purpose: pytest collects it after every other test, so with `-x` a failure
here cannot hide the rest of the suite.  It builds its own model, so it gives the same answer run alone
in a fresh process or after the other tests.

What is checked, on two calls in a row with different inputs and batch sizes (the second call would see
the first call's leftover sums if the shared workspace were not cleared):
  * the spectrogram equals the existing entry point's (dmx_stft_segment, same kernel) within the fp32
    tolerance the other GPU tests use (1e-4 max relative error);
  * the per-segment (mean, unbiased std) derived from the workspace match the statistics of the existing
    entry point's output;
  * a batch above DMX_NOALLOC_MAX_SEGMENTS is refused with a clear message.
Device buffers for the call come from torch (plumbing only); the call itself goes through the C ABI."""
import ctypes

import numpy as np
import pytest
import torch

import util
from demucs_cpp_b200 import api, weights

pytestmark = pytest.mark.gpu

TOL_FP32 = 1e-4
MAX_SEGMENTS = 64  # DMX_NOALLOC_MAX_SEGMENTS in include/demucs_b200.h


def test_noalloc_stft_matches_existing_entry_point_on_two_calls_in_a_row():
    m = api.DemucsModel(weights.cached_weights("htdemucs_6s", 0), max_batch=1)
    L = api.lib()
    try:
        base = util.synth_audio(util.SEG + 4 * 7000, seed=77)
        for call, (n, gain) in enumerate([(3, 1.0), (2, 3.0)]):
            segs = np.stack([gain * (b + 1) * base[:, (call * 2 + b) * 7000:(call * 2 + b) * 7000 + util.SEG]
                             for b in range(n)]).astype(np.float32)                 # (n, 2, SEG)
            d_segs = torch.from_numpy(np.ascontiguousarray(segs.transpose(0, 2, 1))).cuda()
            d_spec = torch.full((n, 2048, 336, 4), float("nan"), dtype=torch.float32, device="cuda")
            d_ms = torch.full((n, 2), float("nan"), dtype=torch.float32, device="cuda")
            torch.cuda.synchronize()
            rc = L.dmx_stft_segments_noalloc(m._h, ctypes.c_void_p(d_segs.data_ptr()), n,
                                             ctypes.c_void_p(d_spec.data_ptr()), ctypes.c_void_p(d_ms.data_ptr()), None)
            assert rc == 0, api.last_error()
            torch.cuda.synchronize()
            spec = d_spec.cpu().numpy()
            ms = d_ms.cpu().numpy()
            for b in range(n):
                ref = m.stft(segs[b])                                               # existing entry point
                cac = np.stack([ref[0].real, ref[0].imag, ref[1].real, ref[1].imag], axis=-1)  # (2048, 336, 4)
                assert util.max_rel_err(spec[b], cac) <= TOL_FP32, "call %d segment %d" % (call, b)
                c64 = cac.astype(np.float64)
                assert abs(ms[b, 0] - c64.mean()) <= TOL_FP32 * c64.std(ddof=1) + 1e-7
                assert abs(ms[b, 1] - c64.std(ddof=1)) <= TOL_FP32 * c64.std(ddof=1)
        # sizes beyond the fixed workspace are rejected, loudly
        rc = L.dmx_stft_segments_noalloc(m._h, ctypes.c_void_p(d_segs.data_ptr()), MAX_SEGMENTS + 1,
                                         ctypes.c_void_p(d_spec.data_ptr()), None, None)
        assert rc != 0 and "between 1 and 64" in api.last_error()
        rc = L.dmx_stft_segments_noalloc(m._h, ctypes.c_void_p(d_segs.data_ptr()), 0,
                                         ctypes.c_void_p(d_spec.data_ptr()), None, None)
        assert rc != 0
    finally:
        m.close()


def test_noalloc_stft_on_a_caller_stream_without_statistics():
    """The asynchronous form: the caller passes its own stream (the call returns without waiting) and no
    statistics buffer.  Two calls back to back on that stream, then one synchronisation; the caller's current
    device must be what it was."""
    m = api.DemucsModel(weights.cached_weights("htdemucs_6s", 0), max_batch=1)
    L = api.lib()
    try:
        st = torch.cuda.Stream()
        base = util.synth_audio(util.SEG + 3 * 5000, seed=78)
        segs = np.stack([(b + 1) * base[:, b * 5000:b * 5000 + util.SEG] for b in range(3)]).astype(np.float32)
        d_segs = torch.from_numpy(np.ascontiguousarray(segs.transpose(0, 2, 1))).cuda()
        outs = [torch.full((n, 2048, 336, 4), float("nan"), dtype=torch.float32, device="cuda") for n in (3, 1)]
        torch.cuda.synchronize()
        dev_before = torch.cuda.current_device()
        for d_spec in outs:
            rc = L.dmx_stft_segments_noalloc(m._h, ctypes.c_void_p(d_segs.data_ptr()), d_spec.shape[0],
                                             ctypes.c_void_p(d_spec.data_ptr()), None, ctypes.c_void_p(st.cuda_stream))
            assert rc == 0, api.last_error()
        st.synchronize()
        assert torch.cuda.current_device() == dev_before
        for d_spec in outs:
            spec = d_spec.cpu().numpy()
            for b in range(spec.shape[0]):
                ref = m.stft(segs[b])
                cac = np.stack([ref[0].real, ref[0].imag, ref[1].real, ref[1].imag], axis=-1)
                assert util.max_rel_err(spec[b], cac) <= TOL_FP32
    finally:
        m.close()
