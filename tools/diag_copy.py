"""Scratch (GPU box): pageable vs registered vs pinned host copies of the stems of a 3-minute track (254 MB)."""
import time, numpy as np, torch
torch.cuda.set_device(0)
rt = torch.cuda.cudart()
n = 254016000 // 4
d = torch.zeros(n, device="cuda")
for trial in range(2):
    h = np.empty(n, dtype=np.float32); h[:] = 1.0          # pageable, touched
    t = torch.from_numpy(h)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    t.copy_(d); torch.cuda.synchronize()
    print("pageable D2H 254MB: %.1f ms" % ((time.perf_counter() - t0) * 1e3))
    t0 = time.perf_counter()
    rc = rt.cudaHostRegister(t.data_ptr(), n * 4, 0)
    t1 = time.perf_counter()
    t.copy_(d, non_blocking=True); torch.cuda.synchronize()
    t2 = time.perf_counter()
    rt.cudaHostUnregister(t.data_ptr())
    t3 = time.perf_counter()
    print("register %.1f ms (rc %d), copy %.1f ms, unregister %.1f ms" % ((t1 - t0) * 1e3, int(rc), (t2 - t1) * 1e3, (t3 - t2) * 1e3))
    # staged: pinned ring + host memcpy
    ring = [torch.empty(8 << 20, dtype=torch.uint8).pin_memory() for _ in range(4)]
    hb = t.view(torch.uint8)
    db = d.view(torch.uint8)
    t0 = time.perf_counter()
    evs = [None] * 4
    off = 0; k = 0; pending = []
    while off < n * 4 or pending:
        if off < n * 4 and len(pending) < 4:
            sz = min(8 << 20, n * 4 - off)
            r = ring[k % 4]
            r[:sz].copy_(db[off:off + sz], non_blocking=True)
            ev = torch.cuda.Event(); ev.record()
            pending.append((ev, r, off, sz)); off += sz; k += 1
        else:
            ev, r, o, sz = pending.pop(0)
            ev.synchronize()
            hb[o:o + sz].copy_(r[:sz])
    print("staged (4 x 8 MB pinned ring, 1 host thread) D2H: %.1f ms" % ((time.perf_counter() - t0) * 1e3))
