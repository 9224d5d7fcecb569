"""Host-side Python mirror of the reference's model API (src/model.hpp) over the C ABI in
include/demucs_b200.h.  It only marshals numpy buffers through ctypes; all compute is in
lib/libdemucs_b200.so (CUDA, sm_100a).  There is no CPU fallback: if the library or a CUDA device
is missing, loading a model raises.

    reference (C++)                                   here
    ----------------------------------------------    ---------------------------------------------
    load_demucs_model(path, &model)                   m = DemucsModel(path)
    model.is_4sources                                 m.is_4sources
    demucs_inference(model, audio(2,L), cb)           m.demucs_inference(audio(2,L), cb, shift_offset)
    model_inference(model, buffers, ...)              m.model_inference(mix(B,2,343980))
    stft / istft (dsp.hpp)                            m.stft(seg) / m.istft(spec)
"""
import ctypes
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "lib", "libdemucs_b200.so")
SEG = 343980
MAX_SHIFT = 22050

PRECISION_FP32, PRECISION_BF16 = 0, 1
PROGRESS_FN = ctypes.CFUNCTYPE(None, ctypes.c_float, ctypes.c_char_p, ctypes.c_void_p)

EXPORTS = [
    "dmx_device_count", "dmx_last_error", "dmx_model_load_file", "dmx_model_create", "dmx_model_set_tensor",
    "dmx_model_finalize", "dmx_model_free", "dmx_model_arch", "dmx_model_sources", "dmx_model_max_batch",
    "dmx_model_set_precision", "dmx_separate", "dmx_separate_device", "dmx_num_segments", "dmx_forward_segments",
    "dmx_debug_tap", "dmx_launch_count", "dmx_stft_segment", "dmx_istft_segment", "dmx_profile_classes",
    "dmx_profile_class_name", "dmx_profile_begin", "dmx_profile_end", "dmx_forward_shard_device",
    "dmx_overlap_add_device", "dmx_last_device_ms", "dmx_stft_segments_noalloc",
    "dmx_debug_linear", "dmx_profile_sequence", "dmx_plan_block", "dmx_track_stats_device", "dmx_forward_block_device",
    "dmx_overlap_add_block_device", "dmx_model_sync", "dmx_separate_multi", "dmx_separate_bag", "dmx_debug_conv",
    "dmx_separate_shifts", "dmx_model_set_layer_taps", "dmx_segment_progress",
]


class BlockPlan(ctypes.Structure):
    """dmx_block_plan (include/demucs_b200.h): one worker's share of a track under the contiguous-block split."""
    _fields_ = [("seg_begin", ctypes.c_int32), ("seg_end", ctypes.c_int32), ("n_segments", ctypes.c_int32),
                ("tail_frames", ctypes.c_int32), ("in_begin", ctypes.c_int64), ("in_end", ctypes.c_int64),
                ("out_begin", ctypes.c_int64), ("out_end", ctypes.c_int64)]


TAIL_FRAMES = 85995

_lib = None


def lib():
    """Load the CUDA library; raises if it has not been built (no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError("CUDA extension missing: %s (run `python -c 'import __graft_entry__ as g; g.build()'`)"
                               % LIB_PATH)
        L = ctypes.CDLL(LIB_PATH)
        vp, ci, i64 = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64
        L.dmx_device_count.restype = ci
        L.dmx_last_error.restype = ctypes.c_char_p
        L.dmx_model_load_file.restype = vp
        L.dmx_model_load_file.argtypes = [ctypes.c_char_p, ci, ci]
        L.dmx_model_create.restype = vp
        L.dmx_model_create.argtypes = [ci, ci, ci]
        L.dmx_model_set_tensor.argtypes = [vp, ctypes.c_char_p, vp, ctypes.POINTER(ctypes.c_int32), ci]
        L.dmx_model_finalize.argtypes = [vp]
        L.dmx_model_free.argtypes = [vp]
        L.dmx_model_free.restype = None
        for f in ("dmx_model_arch", "dmx_model_sources", "dmx_model_max_batch"):
            getattr(L, f).argtypes = [vp]
        L.dmx_model_set_precision.argtypes = [vp, ci]
        L.dmx_separate.argtypes = [vp, vp, i64, ci, vp, PROGRESS_FN, vp]
        L.dmx_separate_device.argtypes = [vp, vp, i64, ci, vp]
        L.dmx_num_segments.argtypes = [i64, ci]
        L.dmx_forward_segments.argtypes = [vp, vp, ci, vp]
        L.dmx_debug_tap.restype = i64
        L.dmx_debug_tap.argtypes = [vp, ctypes.c_char_p, vp, i64, ctypes.POINTER(i64)]
        L.dmx_launch_count.restype = i64
        L.dmx_launch_count.argtypes = [vp]
        L.dmx_forward_shard_device.argtypes = [vp, vp, i64, ci, ci, ci, vp]
        L.dmx_overlap_add_device.argtypes = [vp, vp, i64, ci, vp, ci, vp]
        L.dmx_last_device_ms.restype = ctypes.c_double
        L.dmx_last_device_ms.argtypes = [vp]
        L.dmx_stft_segments_noalloc.argtypes = [vp, vp, ci, vp, vp, vp]
        L.dmx_debug_linear.argtypes = [ci, ci, ci, ci, ci, vp, vp, vp, ci, vp, ci, vp]
        L.dmx_profile_class_name.restype = ctypes.c_char_p
        L.dmx_profile_class_name.argtypes = [ci]
        L.dmx_profile_begin.argtypes = [vp]
        L.dmx_profile_end.argtypes = [vp, vp, vp, vp, vp]
        L.dmx_profile_sequence.argtypes = [vp, vp, ci]
        L.dmx_stft_segment.argtypes = [vp, vp, vp]
        L.dmx_istft_segment.argtypes = [vp, vp, vp]
        L.dmx_plan_block.argtypes = [i64, ci, ci, ci, ctypes.POINTER(BlockPlan)]
        L.dmx_track_stats_device.argtypes = [vp, vp, i64, vp]
        L.dmx_forward_block_device.argtypes = [vp, vp, i64, ci, ctypes.POINTER(BlockPlan), vp, vp]
        L.dmx_overlap_add_block_device.argtypes = [vp, i64, ci, ctypes.POINTER(BlockPlan), vp, vp, vp, ci, ci, vp]
        L.dmx_model_sync.argtypes = [vp]
        L.dmx_separate_multi.argtypes = [vp, ci, vp, i64, ci, vp, PROGRESS_FN, vp]
        L.dmx_separate_bag.argtypes = [vp, ci, vp, i64, vp, vp, PROGRESS_FN, vp]
        L.dmx_debug_conv.argtypes = [ci] * 7 + [vp] * 5 + [ci] * 3 + [vp, ci, vp]
        L.dmx_separate_shifts.argtypes = [vp, vp, i64, vp, ci, vp, PROGRESS_FN, vp]
        L.dmx_model_set_layer_taps.argtypes = [vp, ci]
        _lib = L
    return _lib


def last_error():
    return lib().dmx_last_error().decode()


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


class DemucsError(RuntimeError):
    pass


def _check(rc, what):
    if rc != 0:
        raise DemucsError("%s failed: %s" % (what, last_error()))


class DemucsModel:
    """Device-resident model = demucscpp::demucs_model after load_demucs_model
    (reference: src/model.hpp:285-554,649-650)."""

    def __init__(self, path, device=0, max_batch=4, precision=PRECISION_FP32):
        self._h = lib().dmx_model_load_file(os.fsencode(path), device, max_batch)
        if not self._h:
            raise DemucsError("load_demucs_model(%s) failed: %s" % (path, last_error()))
        self.arch = lib().dmx_model_arch(self._h)
        self.nb_sources = lib().dmx_model_sources(self._h)
        self.is_4sources = self.nb_sources == 4
        self.max_batch = max_batch
        self.device = device
        if precision != PRECISION_FP32:
            self.set_precision(precision)

    def close(self):
        if getattr(self, "_h", None):
            lib().dmx_model_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_layer_taps(self, on=True):
        """Keep the transformer's token streams after norm_in and after every layer (taps x_ct_* / xt_ct_*)."""
        _check(lib().dmx_model_set_layer_taps(self._h, 1 if on else 0), "dmx_model_set_layer_taps")

    def set_precision(self, precision):
        _check(lib().dmx_model_set_precision(self._h, precision), "dmx_model_set_precision")

    @property
    def launch_count(self):
        return int(lib().dmx_launch_count(self._h))

    def profile_begin(self):
        _check(lib().dmx_profile_begin(self._h), "dmx_profile_begin")

    def profile_sequence(self):
        """Class name of every launch scope since profile_begin(), in order (call before profile_end())."""
        n = lib().dmx_profile_sequence(self._h, None, 0)
        buf = np.zeros(max(n, 1), dtype=np.int32)
        lib().dmx_profile_sequence(self._h, _p(buf), n)
        return [lib().dmx_profile_class_name(int(c)).decode() for c in buf[:n]]

    def profile_end(self):
        """-> {class name: {"ms", "launches", "flops", "bytes"}} accumulated since profile_begin()."""
        n = lib().dmx_profile_classes()
        ms, fl, by = (np.zeros(n, dtype=np.float64) for _ in range(3))
        la = np.zeros(n, dtype=np.int64)
        _check(lib().dmx_profile_end(self._h, _p(ms), _p(la), _p(fl), _p(by)), "dmx_profile_end")
        return {lib().dmx_profile_class_name(i).decode(): {"ms": float(ms[i]), "launches": int(la[i]),
                                                           "flops": float(fl[i]), "bytes": float(by[i])}
                for i in range(n)}

    # -- demucscpp::demucs_inference (src/model_apply.cpp:60-91)
    def demucs_inference(self, audio, cb=None, shift_offset=4033):
        """audio: float32 (2, L).  Returns stems float32 (S, 2, L).  `shift_offset` stands in for the
        reference's `rand() % 22050` (4033 is what a fresh glibc process draws)."""
        audio = np.asarray(audio, dtype=np.float32)
        if audio.ndim != 2 or audio.shape[0] != 2:
            raise ValueError("audio must be (2, L)")
        L = audio.shape[1]
        inter = np.ascontiguousarray(audio.T)            # (L, 2) == MatrixXf(2, L) column-major
        out = np.empty((L, 2, self.nb_sources), dtype=np.float32)  # Tensor3dXf(S, 2, L) column-major
        if cb is not None:
            fn = PROGRESS_FN(lambda p, msg, _u: cb(p, msg.decode()))
        else:
            fn = ctypes.cast(None, PROGRESS_FN)
        _check(lib().dmx_separate(self._h, _p(inter), L, int(shift_offset), _p(out), fn, None), "dmx_separate")
        return np.ascontiguousarray(out.transpose(2, 1, 0))

    def demucs_inference_shifts(self, audio, shift_offsets, cb=None):
        """Upstream's `shifts > 1` mode (vendor/demucs/demucs/apply.py:187-199): the mean of demucs_inference over the given
        shift offsets.  (2, L) -> (S, 2, L)."""
        L, inter, out = _track_io(audio, self.nb_sources)
        sh = (ctypes.c_int * len(shift_offsets))(*[int(x) for x in shift_offsets])
        fn = PROGRESS_FN(lambda p, msg, _u: cb(p, msg.decode())) if cb is not None else ctypes.cast(None, PROGRESS_FN)
        _check(lib().dmx_separate_shifts(self._h, _p(inter), L, sh, len(shift_offsets), _p(out), fn, None),
               "dmx_separate_shifts")
        return np.ascontiguousarray(out.transpose(2, 1, 0))

    def demucs_inference_device(self, d_audio_ptr, n_frames, d_out_ptr, shift_offset=4033):
        """Device-resident variant: raw device pointers (e.g. torch tensors' data_ptr())."""
        _check(lib().dmx_separate_device(self._h, ctypes.c_void_p(d_audio_ptr), n_frames, int(shift_offset),
                                         ctypes.c_void_p(d_out_ptr)), "dmx_separate_device")

    def forward_shard_device(self, d_audio_ptr, n_frames, shift_offset, rank, world, d_y_ptr):
        """This rank's segments (rank, rank+world, ...) of the track -> d_y [ceil(nseg/world)][S][SEG][2]."""
        _check(lib().dmx_forward_shard_device(self._h, ctypes.c_void_p(d_audio_ptr), n_frames, int(shift_offset),
                                              rank, world, ctypes.c_void_p(d_y_ptr)), "dmx_forward_shard_device")

    def overlap_add_device(self, d_audio_ptr, n_frames, shift_offset, d_y_all_ptr, world, d_out_ptr):
        _check(lib().dmx_overlap_add_device(self._h, ctypes.c_void_p(d_audio_ptr), n_frames, int(shift_offset),
                                            ctypes.c_void_p(d_y_all_ptr), world, ctypes.c_void_p(d_out_ptr)),
               "dmx_overlap_add_device")

    # -- contiguous-block split (multi-GPU data path); raw device pointers, asynchronous on the model's stream
    def track_stats_device(self, d_audio_ptr, n_frames, d_sums_ptr):
        _check(lib().dmx_track_stats_device(self._h, ctypes.c_void_p(d_audio_ptr), n_frames, ctypes.c_void_p(d_sums_ptr)),
               "dmx_track_stats_device")

    def forward_block_device(self, d_win_ptr, n_frames, shift_offset, plan, d_sums_ptr, d_y_ptr):
        _check(lib().dmx_forward_block_device(self._h, ctypes.c_void_p(d_win_ptr), n_frames, int(shift_offset),
                                              ctypes.byref(plan), ctypes.c_void_p(d_sums_ptr), ctypes.c_void_p(d_y_ptr)),
               "dmx_forward_block_device")

    def overlap_add_block_device(self, n_frames, shift_offset, plan, d_sums_ptr, d_y_ptr, d_tail_ptr, d_out_ptr,
                                 src_begin=0, src_end=None):
        _check(lib().dmx_overlap_add_block_device(self._h, n_frames, int(shift_offset), ctypes.byref(plan),
                                                  ctypes.c_void_p(d_sums_ptr), ctypes.c_void_p(d_y_ptr),
                                                  ctypes.c_void_p(d_tail_ptr) if d_tail_ptr else None, src_begin,
                                                  self.nb_sources if src_end is None else src_end,
                                                  ctypes.c_void_p(d_out_ptr)), "dmx_overlap_add_block_device")

    def sync(self):
        _check(lib().dmx_model_sync(self._h), "dmx_model_sync")

    @property
    def last_device_ms(self):
        return float(lib().dmx_last_device_ms(self._h))

    # -- demucscpp::model_inference (src/model_inference.cpp:48-475) on B full segments
    def model_inference(self, mix):
        """mix: float32 (B, 2, 343980) -> targets (B, S, 2, 343980)."""
        mix = np.asarray(mix, dtype=np.float32)
        if mix.ndim == 2:
            mix = mix[None]
        B = mix.shape[0]
        assert mix.shape[1:] == (2, SEG), mix.shape
        inter = np.ascontiguousarray(mix.transpose(0, 2, 1))
        out = np.empty((B, self.nb_sources, SEG, 2), dtype=np.float32)
        _check(lib().dmx_forward_segments(self._h, _p(inter), B, _p(out)), "dmx_forward_segments")
        return np.ascontiguousarray(out.transpose(0, 1, 3, 2))

    def tap(self, name):
        """Named activation of the last model_inference as (B, d1, d2, C) channels-last."""
        dims = (ctypes.c_int64 * 4)()
        n = lib().dmx_debug_tap(self._h, name.encode(), None, 0, dims)
        if n < 0:
            raise DemucsError(last_error())
        buf = np.empty([int(d) for d in dims], dtype=np.float32)
        if lib().dmx_debug_tap(self._h, name.encode(), _p(buf), buf.size, dims) < 0:
            raise DemucsError(last_error())
        return buf

    def stft(self, seg):
        """seg (2, 343980) -> complex64 (2, 2048, 336): frames 2..337, Nyquist dropped
        (demucscpp::stft + the slice of src/model_inference.cpp:75-99)."""
        seg = np.asarray(seg, dtype=np.float32)
        inter = np.ascontiguousarray(seg.T)
        spec = np.empty((2048, 336, 4), dtype=np.float32)
        _check(lib().dmx_stft_segment(self._h, _p(inter), _p(spec)), "dmx_stft_segment")
        return np.stack([spec[..., 0] + 1j * spec[..., 1], spec[..., 2] + 1j * spec[..., 3]]).astype(np.complex64)

    def istft(self, spec):
        """complex (2, 2048, 336) -> (2, 343980) (demucscpp::istft + crop of src/model_inference.cpp:439-455)."""
        s = np.empty((2048, 336, 4), dtype=np.float32)
        s[..., 0], s[..., 1], s[..., 2], s[..., 3] = spec[0].real, spec[0].imag, spec[1].real, spec[1].imag
        out = np.empty((SEG, 2), dtype=np.float32)
        _check(lib().dmx_istft_segment(self._h, _p(s), _p(out)), "dmx_istft_segment")
        return np.ascontiguousarray(out.T)


def debug_linear(A, W, bias=None, act=0, precision=PRECISION_FP32, device=0, iters=0):
    """act(A (M,K) @ W (N,K)^T + bias) through the production GEMM kernels (test hook).
    With iters > 0 returns (out, average kernel ms)."""
    A = np.ascontiguousarray(A, dtype=np.float32)
    W = np.ascontiguousarray(W, dtype=np.float32)
    M, K = A.shape
    N = W.shape[0]
    out = np.empty((M, N), dtype=np.float32)
    b = None if bias is None else np.ascontiguousarray(bias, dtype=np.float32)
    ms = ctypes.c_float(0.0)
    _check(lib().dmx_debug_linear(device, precision, M, N, K, _p(A), _p(W), None if b is None else _p(b), act,
                                  _p(out), iters, ctypes.byref(ms)), "dmx_debug_linear")
    return (out, ms.value) if iters > 0 else out


def plan_block(n_frames, shift_offset, rank, world):
    """This worker's run of segments and its input / output stretches (dmx_plan_block)."""
    pl = BlockPlan()
    _check(lib().dmx_plan_block(n_frames, int(shift_offset), rank, world, ctypes.byref(pl)), "dmx_plan_block")
    return pl


def _track_io(audio, S):
    audio = np.asarray(audio, dtype=np.float32)
    if audio.ndim != 2 or audio.shape[0] != 2:
        raise ValueError("audio must be (2, L)")
    L = audio.shape[1]
    return L, np.ascontiguousarray(audio.T), np.empty((L, 2, S), dtype=np.float32)


def demucs_inference_multi(models, audio, cb=None, shift_offset=4033):
    """demucs_inference over several GPUs of one process: `models` = one replica of the same model per GPU
    (replaces the thread fan-out of cli-apps/threaded_inference.hpp:50-190).  (2, L) -> (S, 2, L)."""
    L, inter, out = _track_io(audio, models[0].nb_sources)
    hs = (ctypes.c_void_p * len(models))(*[m._h for m in models])
    fn = PROGRESS_FN(lambda p, msg, _u: cb(p, msg.decode())) if cb is not None else ctypes.cast(None, PROGRESS_FN)
    _check(lib().dmx_separate_multi(hs, len(models), _p(inter), L, int(shift_offset), _p(out), fn, None),
           "dmx_separate_multi")
    return np.ascontiguousarray(out.transpose(2, 1, 0))


def demucs_ft_inference(models, audio, shift_offsets=(4033, 12436, 5427, 6865), cb=None):
    """htdemucs_ft: four fine-tuned members, stem i from member i (cli-apps/demucs_ft.cpp:221-241); the default
    shifts are the successive rand() % 22050 draws of a fresh process.  (2, L) -> (4, 2, L)."""
    L, inter, out = _track_io(audio, models[0].nb_sources)
    hs = (ctypes.c_void_p * len(models))(*[m._h for m in models])
    sh = (ctypes.c_int * len(models))(*[int(x) for x in shift_offsets[:len(models)]])
    fn = PROGRESS_FN(lambda p, msg, _u: cb(p, msg.decode())) if cb is not None else ctypes.cast(None, PROGRESS_FN)
    _check(lib().dmx_separate_bag(hs, len(models), _p(inter), L, sh, _p(out), fn, None), "dmx_separate_bag")
    return np.ascontiguousarray(out.transpose(2, 1, 0))


def debug_conv(img, W, taps, bias=None, act=0, glu=0, share_dv=1, device=0, iters=0):
    """The TMA-fed convolution kernel alone (test hook): img (B, U, V, Cin), W (N, ntaps * Cin) with k = tap * Cin + c,
    taps = [(du, dv), ...] -> (B, U, V, N or N / 2).  With iters > 0 returns (out, average kernel ms)."""
    img = np.ascontiguousarray(img, dtype=np.float32)
    W = np.ascontiguousarray(W, dtype=np.float32)
    B, U, V, Cin = img.shape
    N = W.shape[0]
    du = (ctypes.c_int * len(taps))(*[int(t[0]) for t in taps])
    dv = (ctypes.c_int * len(taps))(*[int(t[1]) for t in taps])
    out = np.empty((B, U, V, N // 2 if glu else N), dtype=np.float32)
    b = None if bias is None else np.ascontiguousarray(bias, dtype=np.float32)
    ms = ctypes.c_float(0.0)
    _check(lib().dmx_debug_conv(device, B, U, V, Cin, N, len(taps), du, dv, _p(img), _p(W), None if b is None else _p(b),
                                act, glu, share_dv, _p(out), iters, ctypes.byref(ms)), "dmx_debug_conv")
    return (out, ms.value) if iters > 0 else out


def load_demucs_model(path, **kw):
    """Name-compatible helper (reference: demucscpp::load_demucs_model, src/model.hpp:649)."""
    return DemucsModel(path, **kw)


def num_segments(n_frames, shift_offset):
    return lib().dmx_num_segments(n_frames, shift_offset)
