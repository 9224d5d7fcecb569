"""CPU tests (gloo, world_size 2) of the multi-rank host logic bench.py uses on N GPUs.

  * the contiguous-block split: the PRODUCT's planner (dmx_plan_block, pure host code in the C-ABI library) gives each
    rank a run of segments, an input window and an output stretch; ranks exchange the two statistics sums
    (all_reduce) and the 85995-frame tail of their last segment (send/recv) exactly as bench.py does over NCCL, and
    the concatenated stretches must equal the oracle's demucs_inference pipeline.  The per-segment forward is a
    deterministic stand-in (no GPU here); the device kernels of the same protocol are covered by
    tests/test_gpu_configs.py::test_block_split_equals_single_worker;
  * the older round-robin shard + gather layout (dmx_forward_shard_device / dmx_overlap_add_device), with its slot
    mapping slot(k) = (k % world) * ceil(nseg/world) + k // world (csrc/track.cu)."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import util
from oracle import demucs_oracle as O

SEGL = 64          # shrunken segment length for the stand-in pipeline
STRIDE = 48


def plan(length):
    return [(o, min(SEGL, length - o)) for o in range(0, length, STRIDE)]


def fake_forward(k, n):
    """Stand-in for segment k's stems: depends on (k, position) only."""
    return torch.arange(SEGL, dtype=torch.float32)[None, :] * 0.01 + float(k) + torch.tensor([[0.0], [0.5]])


def worker(rank, world, port, length, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    segs = plan(length)
    nseg = len(segs)
    cnt = (nseg + world - 1) // world
    mine = torch.zeros(cnt, 2, SEGL)
    for j, k in enumerate(range(rank, nseg, world)):       # shard: k = rank, rank + world, ...
        mine[j] = fake_forward(k, segs[k][1])
    gathered = [torch.zeros_like(mine) for _ in range(world)] if rank == 0 else None
    dist.gather(mine, gathered, dst=0)
    if rank == 0:
        y_all = torch.stack(gathered)                       # [world][cnt][...]: what d_y_all holds on rank 0
        flat = y_all.reshape(world * cnt, 2, SEGL)
        out = torch.zeros(2, length)
        den = torch.zeros(length)
        w = torch.minimum(torch.arange(1, SEGL + 1), torch.arange(SEGL, 0, -1)).float() / (SEGL // 2)
        for k, (off, n) in enumerate(segs):
            slot = (k % world) * cnt + k // world           # the device kernel's mapping
            out[:, off:off + n] += w[:n] * flat[slot][:, :n]
            den[off:off + n] += w[:n]
        ret.put((out / den).numpy())
    dist.barrier()
    dist.destroy_process_group()


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_round_robin_shard_gather_overlap_add_world2():
    length = 1000                                            # 21 ragged segments over 2 ranks (11 + 10)
    ctx = mp.get_context("spawn")
    ret = ctx.Queue()
    port = free_port()
    procs = [ctx.Process(target=worker, args=(r, 2, port, length, ret)) for r in range(2)]
    for p in procs:
        p.start()
    got = ret.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    # single-process reference of the same pipeline
    segs = plan(length)
    out = torch.zeros(2, length)
    den = torch.zeros(length)
    w = torch.minimum(torch.arange(1, SEGL + 1), torch.arange(SEGL, 0, -1)).float() / (SEGL // 2)
    for k, (off, n) in enumerate(segs):
        out[:, off:off + n] += w[:n] * fake_forward(k, n)[:, :n]
        den[off:off + n] += w[:n]
    np.testing.assert_allclose(got, (out / den).numpy(), rtol=0, atol=1e-6)


def test_shard_counts_cover_every_segment_once():
    for L, shift in [(7938000, 4033), (441000, 4033), (1000, 0)]:
        nseg = len(O.segment_plan(L + O.MAX_SHIFT - shift))
        for world in (1, 2, 4, 8):
            owned = sorted(k for r in range(world) for k in range(r, nseg, world))
            assert owned == list(range(nseg))
            cnt = (nseg + world - 1) // world
            slots = {(k % world) * cnt + k // world for k in range(nseg)}
            assert len(slots) == nseg and max(slots) < world * cnt


# ------------------------------------------------------------------------------------------------ contiguous blocks
def standin_forward(seg):
    """(2, SEG) -> (3, 2, SEG): three "stems" that depend on the segment's content and on the position."""
    ramp = torch.linspace(0.5, 1.5, seg.shape[1], dtype=seg.dtype)
    return torch.stack([seg * ramp, -0.5 * seg.flip(0), seg * seg])


def block_worker(rank, world, port, L, shift, ret):
    from demucs_cpp_b200 import api
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    audio = torch.from_numpy(util.synth_audio(L, seed=5)).double()
    pl = api.plan_block(L, shift, rank, world)
    nxt = api.plan_block(L, shift, rank + 1, world) if rank + 1 < world else None
    SEG, STRIDE, TAIL, lead = api.SEG, 257985, api.TAIL_FRAMES, 22050 - shift
    Ls = L + lead
    # (1) partial statistics over the OWN stretch, summed over ranks (src/model_apply.cpp:72-80)
    mono = audio[:, pl.out_begin:pl.out_end].mean(0)
    sums = torch.tensor([mono.sum(), (mono * mono).sum()], dtype=torch.float64)
    dist.all_reduce(sums)
    mean = sums[0] / L
    std = torch.sqrt((sums[1] - sums[0] * mean) / (L - 1))
    # (2) forward of the own run of segments, reading only the own window
    win = audio[:, pl.in_begin:pl.in_end]
    ys = []
    for k in range(pl.seg_begin, pl.seg_end):
        off = k * STRIDE
        n = min(SEG, Ls - off)
        left = (SEG - n) // 2
        seg = torch.zeros(2, SEG, dtype=torch.float64)
        src = torch.arange(off, off + n) - lead                 # track frames of the chunk (may fall outside [0, L))
        ok = (src >= 0) & (src < L)
        assert ((src[ok] >= pl.in_begin) & (src[ok] < pl.in_end)).all(), "the window does not cover the block"
        seg[:, left:left + n][:, ok] = (win[:, src[ok] - pl.in_begin] - mean) / std
        ys.append((standin_forward(seg), left, n))
    # (3) the one exchange: my last segment's tail to the next rank
    tail = torch.zeros(3, 2, TAIL, dtype=torch.float64)
    reqs = []
    if nxt is not None and nxt.tail_frames > 0:
        reqs.append(dist.isend(ys[-1][0][:, :, SEG - TAIL:].contiguous(), rank + 1))
    if pl.tail_frames > 0:
        reqs.append(dist.irecv(tail, rank - 1))
    for r in reqs:
        r.wait()
    # (4) overlap-add of the own stretch
    w = torch.from_numpy(O.triangle_weight()).double()
    out = torch.zeros(3, 2, pl.out_end - pl.out_begin, dtype=torch.float64)
    den = torch.zeros(pl.out_end - pl.out_begin, dtype=torch.float64)

    def add(y, w_seg, t_first):
        """y (3,2,n) starts at shifted-track position t_first"""
        a = max(t_first - lead, pl.out_begin)
        b = min(t_first - lead + y.shape[2], pl.out_end)
        if b > a:
            sl = slice(a - (t_first - lead), b - (t_first - lead))
            out[:, :, a - pl.out_begin:b - pl.out_begin] += w_seg[sl] * y[:, :, sl]
            den[a - pl.out_begin:b - pl.out_begin] += w_seg[sl]

    if pl.tail_frames > 0:
        add(tail, w[SEG - TAIL:], (pl.seg_begin - 1) * STRIDE + SEG - TAIL)
    for j, (y, left, n) in enumerate(ys):
        add(y[:, :, left:left + n], w[:n], (pl.seg_begin + j) * STRIDE)
    res = (out / den * std + mean) if out.shape[2] else out
    ret.put((rank, pl.out_begin, pl.out_end, res.numpy()))
    dist.barrier()
    dist.destroy_process_group()


def test_contiguous_block_split_world2_matches_oracle_pipeline():
    L, shift = 900000, 12436                                  # 4 segments, the last one short: 2 + 2
    ctx = mp.get_context("spawn")
    ret = ctx.Queue()
    port = free_port()
    procs = [ctx.Process(target=block_worker, args=(r, 2, port, L, shift, ret)) for r in range(2)]
    for p in procs:
        p.start()
    parts = sorted([ret.get(timeout=300) for _ in range(2)])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert parts[0][1] == 0 and parts[0][2] == parts[1][1] and parts[1][2] == L
    got = np.concatenate([parts[0][3], parts[1][3]], axis=2)
    W = type("W", (), {"dtype": torch.float64, "S": 3})()
    ref = O.separate(W, util.synth_audio(L, seed=5).astype(np.float64), shift, forward=standin_forward).numpy()
    np.testing.assert_allclose(got, ref, rtol=0, atol=1e-9)


def test_block_plans_partition_the_track():
    from demucs_cpp_b200 import api
    for L, shift in [(7938000, 4033), (441000, 4033), (1000, 0), (600001, 22049)]:
        nseg = api.num_segments(L, shift)
        for world in (1, 2, 3, 4, 8, 40):
            plans = [api.plan_block(L, shift, r, world) for r in range(world)]
            segs = [k for p in plans for k in range(p.seg_begin, p.seg_end)]
            assert segs == list(range(nseg)) and all(p.n_segments == nseg for p in plans)
            sizes = [p.seg_end - p.seg_begin for p in plans]
            assert max(sizes) - min(sizes) <= 1 and sizes == sorted(sizes, reverse=True)
            edges = [plans[0].out_begin] + [p.out_end for p in plans]
            assert edges[0] == 0 and edges[-1] == L and edges == sorted(edges)
            for r, p in enumerate(plans):
                assert (p.tail_frames == api.TAIL_FRAMES) == (r > 0 and p.seg_end > p.seg_begin)
                assert p.in_begin <= p.out_begin and p.in_end >= p.out_end


# ------------------------------------------------------------------------------------------------ htdemucs_ft bag
BAG_SHIFTS = [4033, 12436, 5427, 6865]          # the four successive rand() % 22050 draws of a fresh process


def bag_value(i, lo, hi):
    """What stem i holds on frames [lo, hi): depends on (stem, frame, channel) only, whoever computed it."""
    fr = torch.arange(lo, hi, dtype=torch.float64)[:, None]
    return (fr * 1e-3 + 0.25 * i + torch.tensor([[0.0, 0.125]], dtype=torch.float64)).float()


def bag_worker(rank, world, port, L, ret):
    import sys
    sys.path.insert(0, util.ROOT)
    import bench
    from demucs_cpp_b200 import api
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    plans = [api.plan_block(L, sh, rank, world) for sh in BAG_SHIFTS]
    olo, ohi = min(p.out_begin for p in plans), max(p.out_end for p in plans)
    d_out = torch.full((max(ohi - olo, 1), 2, 4), float("nan"))
    for i, p in enumerate(plans):                               # member i's forward + overlap-add: its own stretch of stem i
        d_out[p.out_begin - olo:p.out_end - olo, :, i] = bag_value(i, p.out_begin, p.out_end)
    bench.bag_fixup(torch, dist, rank, world, plans, d_out, olo, torch.device("cpu"))
    p0 = plans[0]
    own = d_out[p0.out_begin - olo:p0.out_end - olo]            # final ownership follows member 0
    ok = all(torch.equal(own[:, :, i], bag_value(i, p0.out_begin, p0.out_end)) for i in range(4))
    moved = sum(abs(p.out_begin - p0.out_begin) + abs(p.out_end - p0.out_end) for p in plans[1:])
    ret.put((rank, p0.out_begin, p0.out_end, bool(ok), moved))
    dist.barrier()
    dist.destroy_process_group()


def test_bag_slivers_cross_to_the_owning_rank_world3():
    """bench.py's htdemucs_ft path on N ranks (BASELINE.json configs[3]): every rank runs its block of each of the four
    members; the members' shifts differ, so their block boundaries differ by up to 0.5 s and those slivers of stem i are
    handed to the rank that owns the frames (bench.bag_fixup, the code the NCCL run uses, here over gloo)."""
    L, world = 40 * 44100, 3
    ctx = mp.get_context("spawn")
    ret = ctx.Queue()
    port = free_port()
    procs = [ctx.Process(target=bag_worker, args=(r, world, port, L, ret)) for r in range(world)]
    for p in procs:
        p.start()
    parts = sorted(ret.get(timeout=300) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert parts[0][1] == 0 and parts[-1][2] == L
    assert all(parts[r][2] == parts[r + 1][1] for r in range(world - 1))
    assert all(ok for _, _, _, ok, _ in parts)
    assert sum(m for *_, m in parts) > 0, "the members' boundaries coincide: the test exercised nothing"
