"""Diagnostic (not product): locate the loss-gradient error vs a float64 oracle per n_fft / T."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "task-aware-audio-coding-perceptual_b200")]
import torch
import tac_spectral as ts
from oracle import spectral_oracle as O
for T in (6002, 8192 + 2048, 16000, 20000):
    for n in (64, 128, 256, 512, 1024, 2048):
        (x,), (y,) = O.synth_audio(1, T, 1, seed=77)
        y64 = y.double().requires_grad_(True); O.mrstft_loss_ref(x.double(), y64, (n,)).backward()
        y32 = y.clone().requires_grad_(True); O.mrstft_loss_ref(x, y32, (n,)).backward()
        yg = y.cuda().requires_grad_(True); ts.mrstft_loss(x.cuda(), yg, (n,)).backward()
        r = y64.grad[0]
        e = (yg.grad.double().cpu()[0] - r).abs()
        e32 = (y32.grad.double()[0] - r).abs()
        top = torch.topk(e, 5).indices.tolist()
        print(f"T={T} n={n}: L2rel ours {float(e.norm()/r.norm()):.2e} torch32 {float(e32.norm()/r.norm()):.2e}  max err {float(e.max()):.2e} (|g|max {float(r.abs().max()):.2e}) at {sorted(top)}")
