"""CPU test of the C++17 drop-in tier: the reference's UNMODIFIED command-line drivers must compile against
compat/model.hpp (our header with the reference's public names) and every demucscpp symbol they reference
must be provided by compat/model_b200.cpp + libdemucs_b200.so.  Needs the reference tree for the drivers,
Eigen and libnyquist headers (present in the build container only); skipped elsewhere."""
import os
import subprocess

import pytest

import util

REF = os.environ.get("DEMUCS_REF", "/root/reference")
CLIS = ["demucs.cpp", "demucs_ft.cpp", "demucs_v3.cpp", "demucs_mt.cpp", "demucs_ft_mt.cpp", "demucs_v3_mt.cpp"]

pytestmark = pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "cli-apps")), reason="reference tree not present")


def cxx(args, **kw):
    return subprocess.run(["/usr/bin/g++", "-std=c++17", "-O0", "-w"] + args, capture_output=True, text=True, **kw)


def test_reference_clis_compile_and_resolve_against_the_shim(tmp_path):
    inc = ["-I", os.path.join(util.ROOT, "compat"), "-I", os.path.join(util.ROOT, "include"),
           "-I", os.path.join(REF, "vendor", "eigen"), "-I", os.path.join(REF, "vendor", "libnyquist", "include"),
           "-I", os.path.join(REF, "cli-apps")]
    shim = str(tmp_path / "model_b200.o")
    r = cxx(inc + ["-fPIC", "-c", os.path.join(util.ROOT, "compat", "model_b200.cpp"), "-o", shim])
    assert r.returncode == 0, r.stderr[-2000:]
    provided = subprocess.run(["nm", "-C", "--defined-only", shim], capture_output=True, text=True).stdout
    for cli in CLIS:
        obj = str(tmp_path / (cli + ".o"))
        r = cxx(inc + ["-c", os.path.join(REF, "cli-apps", cli), "-o", obj])
        assert r.returncode == 0, "%s does not compile against compat/model.hpp:\n%s" % (cli, r.stderr[-2000:])
        und = subprocess.run(["nm", "-C", "--undefined-only", obj], capture_output=True, text=True).stdout
        wanted = [ln.split(" U ", 1)[1].strip() for ln in und.splitlines() if " U demucscpp" in ln]
        assert wanted, cli + " references no demucscpp symbol?"
        for sym in wanted:
            assert sym in provided, "%s needs %s, which the shim does not define" % (cli, sym)
    # the shim itself only needs the C ABI
    und = subprocess.run(["nm", "-C", "--undefined-only", shim], capture_output=True, text=True).stdout
    dmx = sorted({ln.split(" U ", 1)[1].strip() for ln in und.splitlines() if " U dmx_" in ln})
    lib = subprocess.run(["nm", "-D", "--defined-only", os.path.join(util.ROOT, "demucs_cpp_b200", "lib", "libdemucs_b200.so")],
                         capture_output=True, text=True).stdout
    for s in dmx:
        assert (" T " + s) in lib, "libdemucs_b200.so does not export " + s


def test_segment_tier_of_the_reference_api_compiles_and_resolves(tmp_path):
    """demucs_segment_buffers / model_inference and the v3 twins (src/model.hpp:569-647,662-666,1213-1415): the driver that
    uses them like the reference's tests do compiles against compat/model.hpp, and the shim defines what it needs."""
    inc = ["-I", os.path.join(util.ROOT, "compat"), "-I", os.path.join(util.ROOT, "include"),
           "-I", os.path.join(REF, "vendor", "eigen")]
    shim = str(tmp_path / "model_b200.o")
    assert cxx(inc + ["-fPIC", "-c", os.path.join(util.ROOT, "compat", "model_b200.cpp"), "-o", shim]).returncode == 0
    obj = str(tmp_path / "segment_tier.o")
    r = cxx(inc + ["-c", os.path.join(util.ROOT, "compat", "segment_tier_main.cpp"), "-o", obj])
    assert r.returncode == 0, r.stderr[-2000:]
    provided = subprocess.run(["nm", "-C", "--defined-only", shim], capture_output=True, text=True).stdout
    und = subprocess.run(["nm", "-C", "--undefined-only", obj], capture_output=True, text=True).stdout
    wanted = [ln.split(" U ", 1)[1].strip() for ln in und.splitlines() if " U demucscpp" in ln]
    assert any("model_inference" in w for w in wanted) and any("demucs_segment_buffers" in w for w in wanted)
    assert any("model_v3_inference" in w for w in wanted)
    for sym in wanted:
        assert sym in provided, "segment_tier_main.cpp needs %s, which the shim does not define" % sym
