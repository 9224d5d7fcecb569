"""TEST INFRASTRUCTURE ONLY (see oracle/README in DESIGN.md §oracle): ctypes binding of
oracle/_ref/libdemucs_ref.so = the UNMODIFIED reference library compiled by
oracle/Makefile + oracle/ref_harness.cpp.  Only tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs may import this module.

All arrays are numpy float32.  The reference's Eigen tensors are column-major
(src/tensor.hpp:24-28); helpers below convert to ordinary C-order numpy shapes.
"""
import ctypes
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "_ref", "libdemucs_ref.so")
SEG = 343980

_lib = None


def available():
    return os.path.exists(LIB_PATH)


def lib():
    global _lib
    if _lib is None:
        L = ctypes.CDLL(LIB_PATH)
        L.ref_load.restype = ctypes.c_void_p
        L.ref_load.argtypes = [ctypes.c_char_p]
        L.ref_free.argtypes = [ctypes.c_void_p]
        L.ref_arch.argtypes = [ctypes.c_void_p]
        L.ref_nb_sources.argtypes = [ctypes.c_void_p]
        L.ref_separate.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p,
                                   ctypes.c_uint, ctypes.POINTER(ctypes.c_int)]
        L.ref_segment.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
        L.ref_view_dims.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.POINTER(ctypes.c_int64)]
        L.ref_view_copy.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_void_p]
        L.ref_stft.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
        L.ref_istft.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


class RefModel:
    def __init__(self, path):
        self.h = lib().ref_load(path.encode())
        if not self.h:
            raise RuntimeError("reference loader rejected " + path)
        self.arch = lib().ref_arch(self.h)
        self.S = lib().ref_nb_sources(self.h)

    def close(self):
        if self.h:
            lib().ref_free(self.h)
            self.h = None

    def separate(self, audio, seed=1):
        """audio: (2, L) float32 C-order.  Returns (stems (S,2,L), shift_offset)."""
        audio = np.asarray(audio, dtype=np.float32)
        L = audio.shape[1]
        inter = np.ascontiguousarray(audio.T)  # (L,2) C-order == (2,L) column-major
        out = np.empty((L, 2, self.S), dtype=np.float32)  # (S,2,L) column-major
        shift = ctypes.c_int(0)
        lib().ref_separate(self.h, _p(inter), L, _p(out), seed, ctypes.byref(shift))
        return np.ascontiguousarray(out.transpose(2, 1, 0)), shift.value

    def segment(self, mix):
        """mix: (2, 343980).  Returns targets (S,2,343980); views via .view(name)."""
        mix = np.asarray(mix, dtype=np.float32)
        assert mix.shape == (2, SEG)
        inter = np.ascontiguousarray(mix.T)
        out = np.empty((SEG, 2, self.S), dtype=np.float32)
        lib().ref_segment(self.h, _p(inter), _p(out))
        return np.ascontiguousarray(out.transpose(2, 1, 0))

    def view(self, name):
        """Named scratch tensor left by the last segment(), as C-order (d0,d1,d2)."""
        dims = (ctypes.c_int64 * 3)()
        if lib().ref_view_dims(self.h, name.encode(), dims) != 0:
            raise KeyError(name)
        d = [int(x) for x in dims]
        buf = np.empty(d[::-1], dtype=np.float32)
        lib().ref_view_copy(self.h, name.encode(), _p(buf))
        return np.ascontiguousarray(buf.transpose(2, 1, 0))


def stft(wave):
    """wave (2,n) -> complex64 (2, 2049, n//1024+1)  (demucscpp::stft, src/dsp.cpp:51-86)."""
    wave = np.asarray(wave, dtype=np.float32)
    n = wave.shape[1]
    frames = n // 1024 + 1
    inter = np.ascontiguousarray(wave.T)
    out = np.empty((frames, 2049, 2), dtype=np.complex64)
    lib().ref_stft(_p(inter), n, _p(out))
    return np.ascontiguousarray(out.transpose(2, 1, 0))


def istft(spec, n):
    """spec complex64 (2,2049,frames) -> (2,n)  (demucscpp::istft, src/dsp.cpp:88-117)."""
    s = np.ascontiguousarray(np.asarray(spec, dtype=np.complex64).transpose(2, 1, 0))
    out = np.empty((n, 2), dtype=np.float32)
    lib().ref_istft(_p(s), n, _p(out))
    return np.ascontiguousarray(out.T)
