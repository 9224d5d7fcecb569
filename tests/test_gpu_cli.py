"""GPU tests (`-m gpu`) of the drop-in boundary at the level a user of the reference sees it: the reference's UNMODIFIED
command-line drivers (cli-apps/demucs.cpp, demucs_mt.cpp, demucs_ft.cpp, demucs_v3.cpp), compiled against compat/model.hpp,
linked with compat/model_b200.cpp + libdemucs_b200.so + the reference's own libnyquist by compat/Makefile
(compat/_build/*.cpp.main, built in the container where /root/reference lives; the binaries travel to the GPU box), are
RUN on a WAV file and the stems they write are compared with the reference's golden stems / this library's API."""
import os
import subprocess

import numpy as np
import pytest
from scipy.io import wavfile

import util
from demucs_cpp_b200 import api, weights

pytestmark = pytest.mark.gpu

BUILD = os.path.join(util.ROOT, "compat", "_build")
NAMES = ["drums", "bass", "other", "vocals", "guitar", "piano"]


def cli(name):
    p = os.path.join(BUILD, name + ".cpp.main")
    if not os.path.exists(p):
        pytest.skip("compat/_build/%s.cpp.main not built (needs the reference tree; run __graft_entry__.build())" % name)
    return p


def run(args, env=None, timeout=600, with_stderr=False):
    e = dict(os.environ)
    e.update(env or {})
    r = subprocess.run(args, capture_output=True, text=True, env=e, timeout=timeout)
    assert r.returncode == 0, "%s failed:\n%s\n%s" % (args[0], r.stdout[-1500:], r.stderr[-1500:])
    return (r.stdout, r.stderr) if with_stderr else r.stdout


def write_wav(path, audio):
    wavfile.write(path, 44100, np.ascontiguousarray(audio.T.astype(np.float32)))


def read_stems(outdir, S):
    out = []
    for s in range(S):
        sr, d = wavfile.read(os.path.join(outdir, "target_%d_%s.wav" % (s, NAMES[s])))
        assert sr == 44100 and d.dtype == np.float32
        out.append(d.T)
    return np.stack(out)


def test_demucs_cpp_main_writes_the_reference_stems(tmp_path, golden_dir):
    """BASELINE.json configs[0] through the reference's own driver: `demucs.cpp.main <model> <wav> <out dir>` on the 10 s
    clip.  A fresh process draws rand() % 22050 = 4033, the offset the golden stems were produced with."""
    g = np.load(os.path.join(golden_dir, "separate_htdemucs_4s_10s.npz"))
    wav = str(tmp_path / "in.wav")
    write_wav(wav, util.synth_audio(441000))
    sub = np.linspace(0, 4 * 2 * 441000 - 1, 2048).astype(np.int64)
    for prec, bar in (("fp32", 80.0), ("bf16", 40.0)):
        out = str(tmp_path / ("out_" + prec))
        log = run([cli("demucs"), weights.cached_weights("htdemucs_4s", 0), wav, out], {"DEMUCS_PRECISION": prec})
        assert "offset: %d" % int(g["shift"][0]) in log and "Writing wav file" in log
        stems = read_stems(out, 4)
        assert stems.shape == tuple(g["shape"])
        assert util.snr_db(stems.reshape(-1)[sub], g["sub"]) >= bar, prec


def test_demucs_mt_and_v3_and_ft_mains_run(tmp_path):
    """The threaded driver (chunks of the track on std::threads sharing one model, cli-apps/threaded_inference.hpp), the v3
    driver and the fine-tuned bag driver, against this library's API on the same input."""
    audio = util.synth_audio(30 * 44100, seed=41)
    wav = str(tmp_path / "in.wav")
    write_wav(wav, audio)
    env = {"DEMUCS_PRECISION": "fp32", "DEMUCS_MAX_BATCH": "4"}
    # demucs_mt: 2 threads.  threaded_inference.hpp cuts the track into one overlapping chunk per thread, each normalised by
    # its own statistics and shifted by whatever rand() hands the racing threads, so its stems are not comparable sample by
    # sample with a whole-track separation (the reference says the same of its own *_mt drivers): check that the two
    # threads went through one shared model and produced sane stems of the same level
    out = str(tmp_path / "mt")
    run([cli("demucs_mt"), weights.cached_weights("htdemucs_4s", 0), wav, out, "2"], env)
    m = api.DemucsModel(weights.cached_weights("htdemucs_4s", 0), max_batch=4)
    ref = m.demucs_inference(audio, shift_offset=4033)
    m.close()
    mt = read_stems(out, 4)
    assert mt.shape == ref.shape and np.isfinite(mt).all()
    for s in range(4):
        r = np.sqrt((mt[s].astype(np.float64) ** 2).mean()) / np.sqrt((ref[s].astype(np.float64) ** 2).mean())
        assert 0.5 <= r <= 2.0, "stem %d RMS ratio %.2f" % (s, r)
    # demucs_v3
    out = str(tmp_path / "v3")
    run([cli("demucs_v3"), weights.cached_weights("hdemucs_mmi", 0), wav, out], env)
    m = api.DemucsModel(weights.cached_weights("hdemucs_mmi", 0), max_batch=4)
    ref = m.demucs_inference(audio, shift_offset=4033)
    m.close()
    assert util.snr_db(read_stems(out, 4), ref) >= 80.0
    # demucs_ft: a directory with the four members, found by name (cli-apps/demucs_ft.cpp:136-168)
    mdir = tmp_path / "ft_models"
    mdir.mkdir()
    for i, name in enumerate(["drums", "bass", "other", "vocals"]):
        os.symlink(weights.cached_weights("htdemucs_4s", i), str(mdir / ("ggml-model-htdemucs_ft_%s-4s-f16.bin" % name)))
    out = str(tmp_path / "ft")
    run([cli("demucs_ft"), str(mdir), wav, out], env)
    members = [api.DemucsModel(weights.cached_weights("htdemucs_4s", i), max_batch=4) for i in range(4)]
    ref = api.demucs_ft_inference(members, audio)
    for mm in members:
        mm.close()
    assert util.snr_db(read_stems(out, 4), ref) >= 80.0


def test_replicas_split_a_lone_track_and_spread_the_chunk_threads(tmp_path):
    """DEMUCS_NGPU=2 (needs two GPUs): the single-threaded driver splits its track over both replicas and writes the same
    stems as one GPU; the chunk threads of demucs_mt (cli-apps/threaded_inference.hpp:110-123) each get a replica of
    their own instead of queueing on one context."""
    if api.lib().dmx_device_count() < 2:
        pytest.skip("needs two GPUs")
    audio = util.synth_audio(40 * 44100, seed=43)
    wav = str(tmp_path / "in.wav")
    write_wav(wav, audio)
    env = {"DEMUCS_PRECISION": "fp32", "DEMUCS_MAX_BATCH": "4", "DEMUCS_VERBOSE_DEVICES": "1"}
    w = weights.cached_weights("htdemucs_4s", 0)
    run([cli("demucs"), w, wav, str(tmp_path / "one")], env)
    _, err = run([cli("demucs"), w, wav, str(tmp_path / "two")], dict(env, DEMUCS_NGPU="2"), with_stderr=True)
    assert "replica(s) 0,1" in err
    assert util.snr_db(read_stems(str(tmp_path / "two"), 4), read_stems(str(tmp_path / "one"), 4)) >= 80.0
    _, err = run([cli("demucs_mt"), w, wav, str(tmp_path / "mt"), "2"], dict(env, DEMUCS_NGPU="2"), with_stderr=True)
    used = sorted(l.split("replica(s) ")[1].strip() for l in err.splitlines() if "replica(s) " in l)
    assert used == ["0", "1"], err[-800:]
    assert np.isfinite(read_stems(str(tmp_path / "mt"), 4)).all()


@pytest.mark.parametrize("arch,v3", [("htdemucs_4s", 0), ("htdemucs_6s", 0), ("hdemucs_mmi", 1)])
def test_segment_tier_buffers_and_model_inference(tmp_path, arch, v3):
    """The reference's public per-segment tier: `demucs_segment_buffers buffers(2, 343980, S);
    stft_buffers stft_buf(...); model_inference(model, buffers, stft_buf, cb, current, segment)` as its own tests use it,
    provided by compat/model.hpp over the C ABI.  compat/segment_tier_main.cpp fills `mix`, calls it and dumps every
    public tensor; here each one must equal the library's stage activation of the same segment, in the REFERENCE's
    layout (frequency branch Tensor3dXf(C, F, T) column-major, time branch (1, C, T), targets_out (S, 2, L)), and the
    progress callback must deliver the numbered milestones scaled into [current, current + segment]."""
    exe = os.path.join(BUILD, "segment_tier.main")
    if not os.path.exists(exe):
        pytest.skip("compat/_build/segment_tier.main not built (needs the reference tree; run __graft_entry__.build())")
    seg = util.synth_audio(api.SEG, seed=61)
    (tmp_path / "seg.f32").write_bytes(np.ascontiguousarray(seg.T.astype(np.float32)).tobytes())
    out = tmp_path / "out"
    out.mkdir()
    run([exe, weights.cached_weights(arch, 0), str(v3), str(tmp_path / "seg.f32"), str(out)], {"DEMUCS_PRECISION": "fp32"})
    lines = (out / "shapes.txt").read_text().splitlines()
    shapes = {l.split()[0]: tuple(int(v) for v in l.split()[1:4]) for l in lines if not l.startswith("#")}
    dims = [l for l in lines if l.startswith("# dims")][0].split()[2:]
    assert [int(v) for v in dims] == [336, 1536, 1620, 347136, 336, 2049]          # src/model.hpp:622-628 on 343980 samples
    prog = [l for l in lines if l.startswith("# progress")][0]
    nmsg = 22 if v3 else 26
    assert prog.split()[2] == str(nmsg) and abs(float(prog.split()[3]) - 0.75) < 1e-6
    assert "Time encoder 0" in prog or v3

    def load(name):
        d0, d1, d2 = shapes[name]
        return np.fromfile(str(out / (name + ".f32")), dtype=np.float32).reshape(d2, d1, d0)      # column-major (d0, d1, d2)

    m = api.DemucsModel(weights.cached_weights(arch, 0), max_batch=1)
    ref_out = m.model_inference(seg)[0]                                                            # (S, 2, L)
    S = m.nb_sources
    assert shapes["targets_out"] == (S, 2, api.SEG)
    assert util.max_rel_err(load("targets_out").transpose(2, 1, 0), ref_out) <= 1e-5
    names = [n for n in shapes if n != "targets_out"]
    assert len(names) >= 22
    upsampled = {"x_3_channel_upsampled": "x_3_up", "xt_3_channel_upsampled": "xt_3_up"}
    checked = 0
    for n in names:
        tapname = upsampled.get(n, n)
        if arch == "htdemucs_6s" and n in upsampled:
            continue                                                                               # 4-source models only
        if arch == "htdemucs_6s" and n in ("x_3", "xt_3"):
            tapname = tapname + "_up"                                                              # no down-sampler: the transformer's output
        t = m.tap(tapname)[0]                                                                      # (d1, d2, C) channels-last
        got = load(n)                                                                              # (d2, d1, d0)
        if n.startswith("xt") or n.startswith("savedt") or n == "x_shared_5":
            assert shapes[n] == (1, t.shape[2], t.shape[1]), n                                     # (1, C, T)
            got = got[:, :, 0][None]                                                               # (1, T, C)
        else:
            assert shapes[n] == (t.shape[2], t.shape[0], t.shape[1]), n                            # (C, F, T)
            got = got.transpose(1, 0, 2)                                                           # (F, T, C)
        assert util.max_rel_err(got, t) <= 1e-5, n
        checked += 1
    assert checked >= 20
    m.close()
