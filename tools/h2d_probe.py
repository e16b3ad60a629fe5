"""Pinned H2D bandwidth on this box: one stream vs two streams vs 4 MB chunks (the e2e leg of bench.py is PCIe-bound).
python tools/h2d_probe.py"""
import time, torch
dev = torch.device("cuda", 0)
n = 512 * 16000
hx, hy = torch.randn(n).pin_memory(), torch.randn(n).pin_memory()
dx, dy = torch.empty(n, device=dev), torch.empty(n, device=dev)
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def run(fn, reps=20):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps): fn()
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / reps
    return 2 * n * 4 / dt / 1e9, dt * 1e3
def one():
    with torch.cuda.stream(s1):
        dx.copy_(hx, non_blocking=True); dy.copy_(hy, non_blocking=True)
def two():
    with torch.cuda.stream(s1): dx.copy_(hx, non_blocking=True)
    with torch.cuda.stream(s2): dy.copy_(hy, non_blocking=True)
def chunks(k=8):
    m = n // k
    for i in range(k):
        with torch.cuda.stream(s1 if i % 2 == 0 else s2):
            dx[i * m:(i + 1) * m].copy_(hx[i * m:(i + 1) * m], non_blocking=True)
            dy[i * m:(i + 1) * m].copy_(hy[i * m:(i + 1) * m], non_blocking=True)
for name, fn in (("one stream", one), ("two streams", two), ("8 chunks, two streams", chunks)):
    g, ms = run(fn)
    print(f"{name:26s} {g:6.1f} GB/s  {ms:.3f} ms per 65.5 MB")
import os
print("cpus", len(os.sched_getaffinity(0)), "numa nodes", [d for d in os.listdir("/sys/devices/system/node") if d.startswith("node")] if os.path.exists("/sys/devices/system/node") else None)
