"""Write-only and copy bandwidth of this GPU (torch fill_ / copy_ on 4 GiB): the ceiling a write-bound kernel can see."""
import torch
n = 1 << 30
a = torch.empty(n, dtype=torch.float32, device="cuda"); b = torch.empty(n, dtype=torch.float32, device="cuda")
def t(fn, reps=5):
    for _ in range(2): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps
ms = t(lambda: a.fill_(1.0)); print("fill  4 GiB  %.3f ms  %.0f GB/s written" % (ms, 4 * n / ms / 1e6))
ms = t(lambda: a.zero_()); print("zero  4 GiB  %.3f ms  %.0f GB/s written" % (ms, 4 * n / ms / 1e6))
ms = t(lambda: b.copy_(a)); print("copy  4 GiB  %.3f ms  %.0f GB/s read+written" % (ms, 8 * n / ms / 1e6))
