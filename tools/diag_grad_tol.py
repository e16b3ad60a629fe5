"""Measured gradient-tolerance statistics of tests/test_gpu_loss.py::_check_grad for every case (bad-element fraction at
rtol 1e-4 + atol 1e-4 max, median error, relative L2) -- the thresholds in the test are 10x these."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "task-aware-audio-coding-perceptual_b200"), os.path.join(ROOT, "tests")]
import torch
import tac_spectral as ts
from oracle import spectral_oracle as O
CASES = [(2, 16000, (1024, 2048, 512)), (3, 6002, (512, 128)), (1, 20000, (2048, 64, 256)), (2, 16000, (2048, 1024, 512, 256, 128))]
for eps in (1e-2, O.LOG_FLOOR):
    for B, T, n_ffts in CASES:
        (x,), (y,) = O.synth_audio(B, T, 1, seed=77)
        y64 = y.double().requires_grad_(True); O.mrstft_loss_ref(x.double(), y64, n_ffts, eps=eps).backward()
        y32 = y.clone().requires_grad_(True); O.mrstft_loss_ref(x, y32, n_ffts, eps=eps).backward()
        yg = y.cuda().requires_grad_(True); ts.mrstft_loss(x.cuda(), yg, n_ffts, eps=eps).backward()
        ref = y64.grad
        for name, got in (("kernel", yg.grad.double().cpu()), ("torch32", y32.grad.double())):
            scale = float(ref.abs().max()); err = (got - ref).abs()
            bad = err > 1e-4 * scale + 1e-4 * ref.abs()
            print(f"eps={eps:g} B={B} T={T} n={n_ffts} {name}: bad_frac {float(bad.float().mean()):.3e} median/scale {float(err.median())/scale:.2e} L2rel {float(err.norm()/ref.norm()):.2e}")
# smoke()'s case
views, decs = O.synth_audio(4, 16000, 2)
x, y = torch.stack(views, 1), torch.stack(decs, 1)
yr = y.clone().requires_grad_(True); O.mrstft_loss_ref(x.reshape(-1, 16000), yr.reshape(-1, 16000)).backward()
yg = y.cuda().requires_grad_(True); ts.mrstft_loss(x.cuda(), yg).backward()
gerr = (yg.grad.cpu() - yr.grad).abs(); gscale = float(yr.grad.abs().max())
print("smoke case (vs torch fp32): bad_frac %.3e median/scale %.2e" % (float((gerr > 1e-4 * gscale + 1e-4 * yr.grad.abs()).float().mean()), float(gerr.median()) / gscale))
