import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "task-aware-audio-coding-perceptual_b200")); sys.path.insert(0, ROOT)
import torch
import tac_spectral as ts
from bench import synth_host
_, _, zs = synth_host(256, 2, 64, 128, seed=5)
zs = [z.cuda() for z in zs]
stats = ts.dpca_stats(zs, "tf32x3")
for _ in range(3):
    out = ts.dpca_eig(stats, 128)
torch.cuda.synchronize()
print("ok")
