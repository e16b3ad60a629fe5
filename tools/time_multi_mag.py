"""cfg 4 (5 scales, 512 x 10 s): re / im planes vs ONE magnitude / power / log-power plane per scale from the same launch family.
    gpurun -- 'python tools/time_multi_mag.py'"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "task-aware-audio-coding-perceptual_b200"))
import tac_spectral as ts  # noqa: E402

n4, h4 = [2048, 1024, 512, 256, 128], [512, 256, 128, 64, 32]
B, T = 512, 160000
x = 0.1 * torch.randn(B, 1, T, device="cuda")
ms = ts.MultiScaleSTFT(n4, h4, n4)
fk = sum((1 + (T - n) // h) * (n // 2 + 1) for n, h in zip(n4, h4))


def timeit(fn, iters=5):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters


for name, fn, per in (("planes re/im", lambda: ms(x), 8), ("magnitude", lambda: ms.magnitudes(x), 4),
                      ("power", lambda: ms.magnitudes(x, power=2.0), 4),
                      ("10 log10(power + 1e-10)", lambda: ms.magnitudes(x, power=2.0, log_multiplier=10.0), 4)):
    t = timeit(fn)
    nb = 4 * B * T + per * B * fk
    print(f"{name:28s} {t:7.3f} ms  {nb / 1e6:8.1f} MB  {nb / t / 1e6:7.1f} GB/s", flush=True)
    torch.cuda.empty_cache()
