"""GPU parity tests (`-m gpu`) of the CONFIGURATIONS that are benchmarked and shipped, on top of the per-kernel /
per-stage tests of test_gpu_parity.py:

  * the launch shape bench.py times: bf16 mode, 16 segments per launch (and hdemucs_mmi at 16, which crosses the
    LSTM kernel's chunk-of-8 boundary) -- every item against the same segment run alone and against the oracle;
  * BASELINE.json configs[1], [2], [4] at full size: the 3-minute, 31-segment track in bf16 at max_batch = 16 for
    htdemucs 4s / 6s / hdemucs_mmi, against the oracle's track pipeline with the ORACLE forward on the first, a
    middle and the ragged last segment;
  * configs[3]: the htdemucs_ft bag (four members, stem i from member i) through dmx_separate_bag;
  * the multi-GPU data path: the contiguous-block split (dmx_plan_block / dmx_forward_block_device /
    dmx_overlap_add_block_device) run worker by worker, and dmx_separate_multi, against dmx_separate.

Tolerances are BASELINE.json's: fp32 <= 1e-4 / >= 80 dB, bf16 <= 1e-2 / >= 40 dB."""
import os

import numpy as np
import pytest
import torch

import util
from demucs_cpp_b200 import api, weights
from oracle import demucs_oracle as O

pytestmark = pytest.mark.gpu

TOL_FP32, SNR_FP32 = 1e-4, 80.0
TOL_BF16, SNR_BF16 = 1e-2, 40.0
SR = 44100

_models = {}


def model(arch="htdemucs_4s", stress=False, max_batch=4, precision=api.PRECISION_FP32, seed=0, device=0, tag=0):
    key = (arch, stress, max_batch, precision, seed, device, tag)
    if key not in _models:
        _models[key] = api.DemucsModel(weights.cached_weights(arch, seed, stress), device=device, max_batch=max_batch,
                                       precision=precision)
    return _models[key]


def drop_models():
    for m in _models.values():
        m.close()
    _models.clear()


@pytest.fixture(scope="module", autouse=True)
def _cleanup():
    torch.set_num_threads(os.cpu_count() or 4)
    yield
    drop_models()


def oracle_weights(arch, seed=0, stress=False):
    _, tens = weights.read_ggml(weights.cached_weights(arch, seed, stress))
    return O.Weights(tens, torch.float32)


def oracle_forward(W, seg):
    fwd = O.segment_forward_v3 if getattr(W, "is_v3", False) else O.segment_forward
    with torch.no_grad():
        return fwd(W, torch.from_numpy(np.ascontiguousarray(seg))).numpy()


# ------------------------------------------------------------------------------ the benchmarked launch shape
@pytest.mark.parametrize("arch,precision", [("htdemucs_4s", api.PRECISION_BF16), ("hdemucs_mmi", api.PRECISION_BF16),
                                            ("hdemucs_mmi", api.PRECISION_FP32), ("htdemucs_6s", api.PRECISION_BF16)])
def test_batch16_matches_single_segment_and_oracle(arch, precision):
    """16 distinct segments in ONE launch (stress weights: the residual branches count).  Every item must equal the
    same segment run alone (the batch only changes grid sizes, index ranges and the order of the statistics'
    atomic adds), and items 0, 9 (second LSTM chunk for hdemucs_mmi) and the zero-padded last one must meet the
    mode's bar against the oracle."""
    B = 16
    base = util.synth_audio(util.SEG + B * 3000, seed=77)
    segs = np.stack([(0.4 + 0.1 * b) * base[:, b * 3000:b * 3000 + util.SEG] for b in range(B)]).astype(np.float32)
    segs[B - 1] = 0.0
    segs[B - 1, :, 60000:260000] = base[:, :200000]          # a centre-padded short chunk
    bf16 = precision == api.PRECISION_BF16
    out16 = model(arch, True, 16, precision).model_inference(segs)
    assert np.isfinite(out16).all()
    m1 = model(arch, True, 1, precision)
    for b in range(B):
        o1 = m1.model_inference(segs[b:b + 1])[0]
        if bf16:
            assert util.snr_db(out16[b], o1) >= SNR_BF16 + 10.0, "item %d differs from the single-segment run" % b
        else:
            assert util.max_rel_err(out16[b], o1) <= 1e-5, "item %d differs from the single-segment run" % b
    W = oracle_weights(arch, 0, True)
    for b in (0, 9, B - 1):
        ref = oracle_forward(W, segs[b])
        snr, err = util.snr_db(out16[b], ref), util.max_rel_err(out16[b], ref)
        assert snr >= (SNR_BF16 if bf16 else SNR_FP32) and err <= (TOL_BF16 if bf16 else TOL_FP32), \
            "item %d vs oracle: %.1f dB, max rel err %.2e" % (b, snr, err)
    drop_models()


# ------------------------------------------------------------------------------ full-size tracks, bf16, 16 per launch
@pytest.mark.parametrize("arch", ["htdemucs_4s", "htdemucs_6s", "hdemucs_mmi"])
def test_bf16_three_minute_track_against_oracle_segments(arch):
    """BASELINE.json configs[1] / [2] / [4] as benchmarked: 31 segments, bf16, max_batch 16.  Checker: the oracle's
    demucs_inference pipeline (normalise, shift, split, triangular overlap-add) with the ORACLE forward on segments
    0, 15 and the ragged last one (30) and this library's fp32 forward (itself pinned to the oracle) on the others."""
    L, shift = 180 * SR, 4033
    audio = util.synth_audio(L, seed=9)
    assert api.num_segments(L, shift) == 31
    out = model(arch, False, 16, api.PRECISION_BF16).demucs_inference(audio, shift_offset=shift)
    assert out.shape[1:] == (2, L) and np.isfinite(out).all()
    W = oracle_weights(arch)
    m32 = model(arch, False, 1, api.PRECISION_FP32)
    count = [0]

    def forward(seg):
        k = count[0]
        count[0] += 1
        if k in (0, 15, 30):
            return torch.from_numpy(oracle_forward(W, seg.numpy()))
        return torch.from_numpy(m32.model_inference(seg.numpy()[None])[0])

    ref = O.separate(W, audio, shift, forward=forward).numpy()
    assert count[0] == 31
    assert util.snr_db(out, ref) >= SNR_BF16 and util.max_rel_err(out, ref) <= TOL_BF16
    # the stretches that ONLY the oracle-run segments cover (between the overlaps with their neighbours)
    lead = 22050 - shift
    for k in (0, 15, 30):
        a = max(0, k * 257985 + (85995 if k else 0) - lead)
        b = min(L, (k + 1) * 257985 - lead)
        assert util.snr_db(out[:, :, a:b], ref[:, :, a:b]) >= SNR_BF16, "segment %d" % k
    drop_models()


# ------------------------------------------------------------------------------ htdemucs_ft
def test_ft_bag_stem_i_from_member_i():
    """BASELINE.json configs[3] on the 10 s clip: four members (seeds 0-3), each with its own shift (the successive
    rand() % 22050 draws of a fresh process), output stem i = stem i of member i, no averaging
    (cli-apps/demucs_ft.cpp:221-241).  Member 1 is checked against the oracle end to end, the others against this
    library's own demucs_inference of that member (pinned to the reference's golden stems elsewhere)."""
    shifts = (4033, 12436, 5427, 6865)
    audio = util.synth_audio(441000)
    members = [model("htdemucs_4s", False, 4, seed=s) for s in range(4)]
    calls = []
    out = api.demucs_ft_inference(members, audio, shifts, cb=lambda p, msg: calls.append(p))
    assert out.shape == (4, 2, 441000) and np.isfinite(out).all()
    assert len(calls) == 4 * 2 * 26
    for i in (0, 2, 3):
        own = members[i].demucs_inference(audio, shift_offset=shifts[i])
        assert util.max_rel_err(out[i], own[i]) <= 1e-5
    with torch.no_grad():
        ref = O.separate(oracle_weights("htdemucs_4s", 1), audio, shifts[1]).numpy()
    assert util.snr_db(out[1], ref[1]) >= SNR_FP32 and util.max_rel_err(out[1], ref[1]) <= TOL_FP32
    with pytest.raises(api.DemucsError, match="one member per source"):
        api.demucs_ft_inference(members[:3], audio, shifts)


# ------------------------------------------------------------------------------ multi-GPU data path
def run_block_split(m, audio, shift, world):
    """The contiguous-block protocol of include/demucs_b200.h, worker by worker on one device (what bench.py does
    with one process per GPU): windows in, partial statistics summed, forward, tail hand-over, overlap-add."""
    L = audio.shape[1]
    S = m.nb_sources
    inter = np.ascontiguousarray(audio.T)
    plans = [api.plan_block(L, shift, r, world) for r in range(world)]
    assert plans[0].out_begin == 0 and plans[-1].out_end == L
    assert all(plans[r].out_end == plans[r + 1].out_begin for r in range(world - 1))
    assert sum(p.seg_end - p.seg_begin for p in plans) == plans[0].n_segments
    d_sums = torch.zeros(2, dtype=torch.float64, device="cuda")
    wins = []
    for p in plans:
        w = torch.from_numpy(inter[p.in_begin:p.in_end]).cuda() if p.in_end > p.in_begin else None
        wins.append(w)
        if w is not None:
            m.track_stats_device(w[p.out_begin - p.in_begin:].data_ptr(), p.out_end - p.out_begin, d_sums.data_ptr())
    m.sync()
    out = torch.full((L, 2, S), float("nan"), dtype=torch.float32, device="cuda")
    tail = None
    for p, w in zip(plans, wins):
        n = p.seg_end - p.seg_begin
        if n == 0:
            continue
        d_y = torch.empty((n, S, api.SEG, 2), dtype=torch.float32, device="cuda")
        m.forward_block_device(w.data_ptr(), L, shift, p, d_sums.data_ptr(), d_y.data_ptr())
        assert (p.tail_frames > 0) == (tail is not None)
        m.overlap_add_block_device(L, shift, p, d_sums.data_ptr(), d_y.data_ptr(), tail.data_ptr() if tail is not None else 0,
                                   out[p.out_begin:].data_ptr())
        m.sync()
        tail = d_y[n - 1, :, api.SEG - api.TAIL_FRAMES:, :].contiguous()
    return np.ascontiguousarray(out.cpu().numpy().transpose(2, 1, 0))


@pytest.mark.parametrize("seconds,shift,world", [(180, 4033, 8), (60, 12436, 2), (20, 0, 7)])
def test_block_split_equals_single_worker(seconds, shift, world):
    """8 workers on the 3-minute track (4,4,4,4,4,4,4,3 segments), 2 workers, and more workers than segments."""
    m = model(max_batch=4)
    audio = util.synth_audio(seconds * SR, seed=13)
    ref = m.demucs_inference(audio, shift_offset=shift)
    out = run_block_split(m, audio, shift, world)
    assert np.isfinite(out).all()
    assert util.max_rel_err(out, ref) <= 1e-5 and util.snr_db(out, ref) >= 100.0


def test_separate_multi_one_process():
    """dmx_separate_multi: one host thread per replica.  With fewer GPUs than replicas the replicas share GPU 0
    (the thread / tail / statistics protocol is the same); on a multi-GPU box each replica gets its own GPU."""
    ngpu = api.lib().dmx_device_count()
    audio = util.synth_audio(60 * SR, seed=17)
    ref = model(max_batch=4).demucs_inference(audio, shift_offset=4033)
    for n in (2, 3):
        reps = [model(max_batch=4, device=(r if ngpu >= n else 0), tag=100 + r) for r in range(n)]
        calls = []
        out = api.demucs_inference_multi(reps, audio, cb=lambda p, msg: calls.append(p), shift_offset=4033)
        assert util.max_rel_err(out, ref) <= 1e-5 and util.snr_db(out, ref) >= 100.0
        assert len(calls) == api.num_segments(60 * SR, 4033) * 26 and abs(calls[-1] - 1.0) < 1e-6


# ------------------------------------------------------------------------------ per-layer tier (transformer)
@pytest.mark.parametrize("arch,precision,stress", [("htdemucs_4s", api.PRECISION_FP32, True),
                                                   ("htdemucs_6s", api.PRECISION_FP32, True),
                                                   ("htdemucs_4s", api.PRECISION_BF16, False)])
def test_transformer_layer_by_layer(arch, precision, stress):
    """The reference tests one transformer layer at a time (common_encoder_layer, src/layers.cpp:377-531, in
    test/test_layers.cpp:1973-2157; norms :2159-2234).  Here: the token streams after norm_in + positional embedding and
    after EACH of the five self / cross layers of both branches, for 4s (d = 512, with up / down-samplers) and 6s (d = 384),
    against the oracle's per-layer values: fp32 mode with stress weights (LayerScale 0.3, so the attention and FFN branches
    count) at 1e-4, bf16 mode with the default initialisation at the bf16 per-stage bar of 1e-2 (as the other bf16 stage
    tests; with stress weights the encoders' accumulated bf16 error alone reaches 1.06e-2 at the transformer's input)."""
    m = model(arch, stress, 1, precision)
    m.set_layer_taps(True)
    seg = util.synth_audio(util.SEG, seed=23)
    m.model_inference(seg[None])
    ot = {}
    with torch.no_grad():
        O.segment_forward(oracle_weights(arch, 0, stress), torch.from_numpy(seg), ot)
    tol = TOL_FP32 if precision == api.PRECISION_FP32 else TOL_BF16
    for tag in ["in"] + [str(i) for i in range(5)]:
        for name in ("x_ct_" + tag, "xt_ct_" + tag):
            got = util.tap_to_cfirst(m.tap(name))[0]
            ref = util.oracle_tap_to_cfirst(name, ot[name].numpy())
            err = util.max_rel_err(got, ref)
            assert err <= tol, "%s: max rel err %.3e" % (name, err)
    m.set_layer_taps(False)
    drop_models()


# ------------------------------------------------------------------------------ shifts > 1, host buffer kinds
def test_shift_averaging_matches_oracle():
    """dmx_separate_shifts = upstream's `shifts` option (vendor/demucs/demucs/apply.py:187-199): the mean of the
    single-shift separations.  Against the mean of the oracle's demucs_inference over the same offsets."""
    shifts = (4033, 12436, 7)
    audio = util.synth_audio(300000, seed=29)
    m = model(max_batch=4)
    out = m.demucs_inference_shifts(audio, shifts)
    own = np.mean([m.demucs_inference(audio, shift_offset=s) for s in shifts], axis=0)
    assert util.max_rel_err(out, own) <= 1e-6
    W = oracle_weights("htdemucs_4s")
    with torch.no_grad():
        ref = np.mean([O.separate(W, audio, s).numpy() for s in shifts], axis=0)
    assert util.snr_db(out, ref) >= SNR_FP32 and util.max_rel_err(out, ref) <= TOL_FP32


def test_pageable_and_pinned_host_buffers_agree():
    """dmx_separate copies pinned caller buffers directly and stages ordinary (pageable) ones -- what the reference's CLIs
    hold, cli-apps/demucs.cpp:162-230 -- through its pinned ring and worker thread: same bytes either way."""
    import ctypes
    L = 45 * SR
    audio = util.synth_audio(L, seed=31)
    m = model(max_batch=4)
    ref = m.demucs_inference(audio, shift_offset=4033)                  # numpy buffers: the staged path
    h_in = torch.from_numpy(np.ascontiguousarray(audio.T)).pin_memory()
    h_out = torch.empty((L, 2, 4), dtype=torch.float32).pin_memory()
    rc = api.lib().dmx_separate(m._h, ctypes.c_void_p(h_in.data_ptr()), L, 4033, ctypes.c_void_p(h_out.data_ptr()),
                                ctypes.cast(None, api.PROGRESS_FN), None)
    assert rc == 0, api.last_error()
    got = h_out.numpy().transpose(2, 1, 0)
    assert util.max_rel_err(got, ref) <= 1e-6 and np.isfinite(got).all()


@pytest.mark.parametrize("arch,precision", [("htdemucs_4s", api.PRECISION_BF16), ("hdemucs_mmi", api.PRECISION_FP32)])
def test_replayed_forward_graph_equals_the_eager_forward(arch, precision):
    """The track-level entry points run their first forward of a (buffers, batch) key eagerly, capture the second into a
    CUDA graph and replay it afterwards (csrc/track.cu track_forward).  Same model, same device buffers: call 1 (eager)
    on track A, call 2 (captured) on track B, calls 3 / 4 (replays) on A and B again -- replays must reproduce the eager
    results and follow the CONTENT of the buffers, and DMX_NO_GRAPH-style eager runs of another model object
    are the cross-check for B."""
    L = 5 * 257985 + 1000                                             # 6 segments: launches of 4 and 2 at max_batch 4
    a = torch.from_numpy(np.ascontiguousarray(util.synth_audio(L, seed=51).T)).cuda()
    b = torch.from_numpy(np.ascontiguousarray(util.synth_audio(L, seed=52).T)).cuda()
    d_in = torch.empty_like(a)
    m = model(arch, max_batch=4, precision=precision, tag=77)
    d_out = torch.empty((L, 2, m.nb_sources), dtype=torch.float32, device="cuda")
    outs = []
    for src in (a, b, a, b, a):
        d_in.copy_(src)
        torch.cuda.synchronize()
        m.demucs_inference_device(d_in.data_ptr(), L, d_out.data_ptr(), 4033)
        torch.cuda.synchronize()
        outs.append(d_out.cpu().numpy().copy())
    assert np.isfinite(outs[0]).all() and np.abs(outs[0] - outs[1]).max() > 1e-3
    same = lambda x, y: util.max_rel_err(x, y) <= 1e-5    # statistics are summed with atomics: last-bit differences only
    assert same(outs[0], outs[2]) and same(outs[0], outs[4])                           # eager A == replayed A
    assert same(outs[1], outs[3])                                                      # captured-call B == replayed B
    m2 = model(arch, max_batch=4, precision=precision, tag=78)                         # fresh model: its first call is eager
    d_in.copy_(b)
    torch.cuda.synchronize()
    m2.demucs_inference_device(d_in.data_ptr(), L, d_out.data_ptr(), 4033)
    torch.cuda.synchronize()
    assert same(outs[1], d_out.cpu().numpy())
    if precision == api.PRECISION_BF16:
        # a captured forward holds the kernels of the mode it was captured in: switching the mode must drop it
        m.set_precision(api.PRECISION_FP32)
        d_in.copy_(a)
        torch.cuda.synchronize()
        m.demucs_inference_device(d_in.data_ptr(), L, d_out.data_ptr(), 4033)
        torch.cuda.synchronize()
        got = d_out.cpu().numpy().copy()
        m3 = model(arch, max_batch=4, precision=api.PRECISION_FP32, tag=79)
        m3.demucs_inference_device(d_in.data_ptr(), L, d_out.data_ptr(), 4033)
        torch.cuda.synchronize()
        assert same(got, d_out.cpu().numpy()) and util.snr_db(got, outs[0]) < 75.0      # fp32 result, not the bf16 replay
        m.set_precision(api.PRECISION_BF16)
