import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "task-aware-audio-coding-perceptual_b200")]
import torch
import tac_spectral as ts
from oracle import spectral_oracle as O
for T in (8192, 16000):
    for n_ffts in [(2048,), (2048, 1024), (2048, 512), (1024, 512), (512, 1024), (1024, 2048, 512), (512, 1024, 2048)]:
        (x,), (y,) = O.synth_audio(1, T, 1, seed=77)
        y64 = y.double().requires_grad_(True); O.mrstft_loss_ref(x.double(), y64, n_ffts, eps=1e-2).backward()
        yg = y.cuda().requires_grad_(True); ts.mrstft_loss(x.cuda(), yg, n_ffts, eps=1e-2).backward()
        r = y64.grad[0]; e = (yg.grad.double().cpu()[0] - r)
        blocks = [f"{float(e[i:i+1024].norm()/ (r[i:i+1024].norm()+1e-30)):.1e}" for i in range(0, T, 1024)]
        print(f"T={T} {n_ffts}: L2rel {float(e.norm()/r.norm()):.2e}  per-1024 blocks: {' '.join(blocks)}")
