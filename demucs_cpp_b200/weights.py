"""ggml-f16 weight files: synthesis of random-init models and a plain reader.

On-disk format (reference: scripts/convert-pth-to-ggml.py:111-140, read back by
src/model_load.cpp:71-147):

    u32 magic                       'dmc4' 0x646d6334 | 'dmc6' 0x646d6336 | 'dmc3' 0x646d6333
    repeat until EOF:
        i32 n_dims, i32 name_len, i32 ne[n_dims], char name[name_len],
        f16 data[prod(ne)]          row-major (C order), tensor squeezed first

There is no network for checkpoints, so benchmarks and parity tests run on
random-init weights of each architecture.  The tensor names/shapes and the scale
of upstream's own initialisation come from `manifests/<arch>.json` (generated once
from the vendored upstream package by oracle/gen_manifest.py); values are drawn
here with numpy only, so the same file can be produced on any box from a seed.

"stress" variant: every LayerScale tensor (`*.scale`) is set to 0.3
instead of 1e-3/1e-4 so that DConv / transformer residual branches actually matter
in end-to-end parity.
"""
import json
import os
import struct

import numpy as np

MANIFEST_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "manifests")
ARCHS = ("htdemucs_4s", "htdemucs_6s", "hdemucs_mmi")
MAGIC = {"htdemucs_4s": 0x646D6334, "htdemucs_6s": 0x646D6336, "hdemucs_mmi": 0x646D6333}


def load_manifest(arch):
    with open(os.path.join(MANIFEST_DIR, arch + ".json")) as f:
        return json.load(f)


def synth_tensors(arch, seed=0, stress=False):
    """Yield (name, float16 ndarray) in state_dict order."""
    man = load_manifest(arch)
    rng = np.random.default_rng(seed)
    for t in man["tensors"]:
        shape = tuple(t["shape"])
        init = t["init"]
        if init["kind"] == "const":
            v = np.full(shape, init["value"], dtype=np.float32)
        else:
            a = init["std"] * np.sqrt(3.0)
            v = rng.uniform(-a, a, size=shape).astype(np.float32)
        if stress and t["name"].endswith(".scale"):
            v = np.full(shape, 0.3, dtype=np.float32)
        yield t["name"], v.astype(np.float16)


def write_ggml(path, magic, named_tensors):
    with open(path, "wb") as f:
        f.write(struct.pack("<I", magic))
        for name, d in named_tensors:
            d = np.ascontiguousarray(d, dtype=np.float16)
            s = name.encode()
            f.write(struct.pack("<ii", d.ndim, len(s)))
            for k in d.shape:
                f.write(struct.pack("<i", k))
            f.write(s)
            f.write(d.tobytes())


def make_weights(arch, path, seed=0, stress=False):
    """Write a random-init ggml-f16 file for `arch`; returns path."""
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    tmp = path + ".tmp%d" % os.getpid()
    write_ggml(tmp, MAGIC[arch], synth_tensors(arch, seed, stress))
    os.replace(tmp, path)
    return path


def cached_weights(arch, seed=0, stress=False, cache_dir=None):
    """Synthesise once per (arch, seed, stress) under a cache dir and return the path."""
    cache_dir = cache_dir or os.environ.get("DEMUCS_B200_CACHE", "/tmp/demucs_b200_weights")
    name = "ggml-model-%s-seed%d%s-f16.bin" % (arch, seed, "-stress" if stress else "")
    path = os.path.join(cache_dir, name)
    if not os.path.exists(path):
        make_weights(arch, path, seed, stress)
    return path


def read_ggml(path):
    """Return (magic, {name: float32 ndarray}) -- f16 widened exactly like
    src/model_load.cpp:1092-1300 does (values are exactly representable)."""
    out = {}
    with open(path, "rb") as f:
        magic = struct.unpack("<I", f.read(4))[0]
        while True:
            hdr = f.read(8)
            if len(hdr) < 8:
                break
            n_dims, name_len = struct.unpack("<ii", hdr)
            ne = struct.unpack("<%di" % n_dims, f.read(4 * n_dims)) if n_dims else ()
            name = f.read(name_len).decode()
            n = int(np.prod(ne)) if n_dims else 1
            d = np.frombuffer(f.read(2 * n), dtype=np.float16).astype(np.float32).reshape(ne)
            out[name] = d
    return magic, out
