"""Per-scale cost of the fused loss (cfg2 shapes): python tools/time_scales.py  -> ms and ns per frame-point for each n_fft alone."""
import os, sys, math
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "task-aware-audio-coding-perceptual_b200")); sys.path.insert(0, ROOT)
import torch
from bench import synth_host, WORKLOADS
from tac_spectral.step import SpectralStep
w = WORKLOADS["cfg2"]
dev = torch.device("cuda", 0)
sets = []
for i in range(6):
    x, y, zs = synth_host(w["batch"], w["views"], w["T"], w["d"], seed=i % 2)
    sets.append((x.to(dev) * (1 - 0.001 * i), y.to(dev) * (1 - 0.001 * i)))
def timeit(fn, n=20):
    for i in range(3): fn(i)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(n): fn(i)
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n
print("lib", os.environ.get("TAC_SPECTRAL_LIB", "default"))
for nf in ((2048,), (1024,), (512,), (256,), (128,), (1024, 2048, 512)):
    st = SpectralStep(w["batch"] * w["views"], w["T"], w["views"], w["batch"], w["d"], w["dim_p"], nf, [n // 4 for n in nf], device=dev, overlap=False, frontend=False)
    ms = timeit(lambda i: st.leg_loss_fused(*sets[i % 6]))
    work = sum((1 + (w["T"] - n) // (n // 4)) * n * math.log2(n) for n in nf) * w["batch"] * w["views"]
    print(f"n_fft {nf}: fused {ms:.4f} ms   {ms * 1e6 / work * 1e3:.3f} ps per (frame point x log2 n)")
