"""CPU tests of the C-ABI library: it loads, exports every symbol include/*.h declares, and fails
loudly (no CPU fallback) when there is no CUDA device.  No compute calls here."""
import ctypes
import glob
import os
import re

import pytest

import util
from demucs_cpp_b200 import api


def declared_symbols():
    syms = set()
    for h in glob.glob(os.path.join(util.ROOT, "include", "*.h")):
        src = open(h).read()
        src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
        syms |= set(re.findall(r"\b(dmx_[a-z0-9_]+)\s*\(", src))
    return syms


def test_library_exports_every_declared_symbol():
    L = ctypes.CDLL(api.LIB_PATH)
    decl = declared_symbols()
    assert decl, "no declarations found"
    for s in sorted(decl):
        assert hasattr(L, s), "libdemucs_b200.so does not export " + s
    assert set(api.EXPORTS) <= decl


def test_fails_loudly_without_gpu_or_with_bad_input(tmp_path):
    n = api.lib().dmx_device_count()
    bad = tmp_path / "bad.bin"
    bad.write_bytes(b"\x00\x01\x02\x03" + b"\x00" * 64)
    with pytest.raises(api.DemucsError, match="bad magic"):
        api.DemucsModel(str(bad))
    with pytest.raises(api.DemucsError, match="cannot open"):
        api.DemucsModel(str(tmp_path / "missing.bin"))
    if n == 0:
        from demucs_cpp_b200 import weights
        p = weights.cached_weights("htdemucs_6s", 0)
        with pytest.raises(api.DemucsError, match="no CUDA device"):
            api.DemucsModel(p)


def test_product_does_not_import_oracle():
    # the shipped package must never reach into oracle/ (that would be a hidden CPU path)
    pat = re.compile(r"^\s*(from\s+oracle|import\s+oracle|#\s*include\s*[<\"].*oracle)", re.M)
    for f in glob.glob(os.path.join(util.ROOT, "demucs_cpp_b200", "**", "*"), recursive=True):
        if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
            assert not pat.search(open(f).read()), f
