"""Diagnostic (not product): fp32 accuracy of the STFT kernel and the loss gradient vs a float64 oracle,
next to torch's own fp32 error on the same inputs."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "task-aware-audio-coding-perceptual_b200")]
import torch
import tac_spectral as ts
from oracle import spectral_oracle as O

(x,), (y,) = O.synth_audio(4, 16000, 1, seed=77)
for n in (64, 128, 256, 512, 1024, 2048):
    ref = O.disc_inputs_ref(x[:, None].double(), n, n // 4)
    cpu32 = O.disc_inputs_ref(x[:, None], n, n // 4).double()
    gpu32 = cpu32
    ours = ts.stft_multi(x[:, None].cuda(), [n], [n // 4])[0].double().cpu()
    m = ref.abs().max()
    f = lambda a: (float((a - ref).abs().max() / m), float((a - ref).norm() / ref.norm()))
    print(f"n={n}: max/L2 err vs f64  ours {f(ours)}  torch-cpu {f(cpu32)}  torch-cuda {f(gpu32)}")

for B, T, n_ffts in [(2, 16000, (1024, 2048, 512)), (3, 6002, (512, 128)), (1, 20000, (2048, 64, 256)), (1, 20000, (64,)), (1, 20000, (2048,))]:
    (x,), (y,) = O.synth_audio(B, T, 1, seed=77)
    y64 = y.double().requires_grad_(True); O.mrstft_loss_ref(x.double(), y64, n_ffts).backward()
    y32 = y.clone().requires_grad_(True); O.mrstft_loss_ref(x, y32, n_ffts).backward()
    yc = y32
    yg = y.cuda().requires_grad_(True); ts.mrstft_loss(x.cuda(), yg, n_ffts).backward()
    r = y64.grad
    e = lambda g: float((g.double().cpu() - r).norm() / r.norm())
    print(f"grad {B}x{T} {n_ffts}: L2 rel vs f64  ours {e(yg.grad):.3e}  torch-cpu32 {e(y32.grad):.3e}  torch-cuda32 {e(yc.grad):.3e}")
