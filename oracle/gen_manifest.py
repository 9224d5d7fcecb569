"""TEST/TOOLING ONLY.  Generates demucs_cpp_b200/manifests/*.json from the vendored
upstream PyTorch package (reference: /root/reference/vendor/demucs).  Runs only in
the build container (the GPU box has no /root/reference); its OUTPUT is committed.

A manifest lists, in `state_dict()` order, every tensor the reference loader
demands (src/model_load.cpp:120-147 record format; names = PyTorch state_dict
keys, shapes after `.squeeze()`, scripts/convert-pth-to-ggml.py:123-140) plus the
statistics of upstream's own random initialisation (after its `rescale_module`),
so that demucs_cpp_b200/weights.py can synthesise random-init weight files of the
same architecture and scale anywhere, without torch or the reference tree.

Architectures:
  htdemucs_4s : HTDemucs(4 sources, dconv_mode=3, bottom_channels=512, segment=39/5)  -> 'dmc4'
  htdemucs_6s : HTDemucs(6 sources, dconv_mode=3, bottom_channels=0,   segment=39/5)  -> 'dmc6'
  hdemucs_mmi : HDemucs(4 sources) defaults                                            -> 'dmc3'
"""
import json
import os
import sys
import types
from fractions import Fraction

import torch

REF = os.environ.get("DEMUCS_REF", "/root/reference")
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "demucs_cpp_b200", "manifests")


def _stub_imports():
    # module-scope-only third-party imports of the upstream package; unused for construction
    for n in ["julius", "openunmix", "openunmix.filtering", "omegaconf", "diffq", "dora", "dora.log"]:
        sys.modules[n] = types.ModuleType(n)
    sys.modules["openunmix.filtering"].wiener = None
    sys.modules["omegaconf"].OmegaConf = object
    for a in ["DiffQuantizer", "UniformQuantizer", "restore_quantized_state"]:
        setattr(sys.modules["diffq"], a, None)
    sys.modules["dora.log"].fatal = sys.modules["dora.log"].bold = None
    sys.path.insert(0, os.path.join(REF, "vendor", "demucs"))


def build(arch):
    from demucs.hdemucs import HDemucs
    from demucs.htdemucs import HTDemucs

    s4 = ["drums", "bass", "other", "vocals"]
    torch.manual_seed(0)
    if arch == "htdemucs_4s":
        return HTDemucs(sources=s4, dconv_mode=3, bottom_channels=512, segment=Fraction(39, 5)), 0x646D6334
    if arch == "htdemucs_6s":
        return HTDemucs(sources=s4 + ["guitar", "piano"], dconv_mode=3, bottom_channels=0,
                        segment=Fraction(39, 5)), 0x646D6336
    if arch == "hdemucs_mmi":
        return HDemucs(sources=s4), 0x646D6333
    raise ValueError(arch)


def main():
    _stub_imports()
    os.makedirs(OUT, exist_ok=True)
    for arch in ["htdemucs_4s", "htdemucs_6s", "hdemucs_mmi"]:
        model, magic = build(arch)
        tensors = []
        nparam = 0
        for name, t in model.state_dict().items():
            d = t.squeeze().float()
            nparam += d.numel()
            std = float(d.std(unbiased=False)) if d.numel() > 1 else 0.0
            mean = float(d.mean())
            if std < 1e-12:
                init = {"kind": "const", "value": mean}
            else:
                init = {"kind": "uniform", "mean": mean, "std": std}
            tensors.append({"name": name, "shape": list(d.shape), "init": init})
        with open(os.path.join(OUT, arch + ".json"), "w") as f:
            json.dump({"arch": arch, "magic": magic, "n_tensors": len(tensors), "n_params": nparam,
                       "tensors": tensors}, f, indent=0)
        print(arch, len(tensors), "tensors", nparam, "params")


if __name__ == "__main__":
    main()
