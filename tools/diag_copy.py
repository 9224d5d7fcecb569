import os, sys, time, numpy as np, torch, tempfile
sys.path.insert(0, "/root/repo")
torch.cuda.set_device(0)
n = 32 * 1024 * 1024 // 4
d = torch.zeros(n, device="cuda")
fd, path = tempfile.mkstemp(prefix="dmx_diag_", dir="/dev/shm"); os.ftruncate(fd, n * 4); os.close(fd)
mm = np.memmap(path, dtype=np.float32, mode="r+", shape=(n,))
t = torch.from_numpy(mm)
rc = torch.cuda.cudart().cudaHostRegister(t.data_ptr(), n * 4, 0)
print("register rc", int(rc))
pin = torch.empty(n).pin_memory()
for name, h in (("shm-registered", t), ("pin_memory", pin)):
    for _ in range(3):
        h.copy_(d, non_blocking=True); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(10):
        h.copy_(d, non_blocking=True); torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / 10
    print(name, "D2H 32MB: %.2f ms  %.1f GB/s" % (dt * 1e3, n * 4 / dt / 1e9))
    t0 = time.perf_counter()
    for _ in range(10):
        d.copy_(h, non_blocking=True); torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / 10
    print(name, "H2D 32MB: %.2f ms  %.1f GB/s" % (dt * 1e3, n * 4 / dt / 1e9))
os.unlink(path)
