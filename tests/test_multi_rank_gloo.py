"""CPU test (gloo, world_size 2) of the multi-rank host logic bench.py uses on N GPUs: round-robin segment
sharding, the rank-major gather layout and the slot mapping the device overlap-add assumes
(slot(k) = (k % world) * ceil(nseg/world) + k // world, csrc/track.cu).  The per-segment forward is replaced
by a deterministic stand-in so no GPU is needed; the real kernels are covered by the -m gpu tests."""
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
