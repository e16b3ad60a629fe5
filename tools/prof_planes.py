import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "task-aware-audio-coding-perceptual_b200")); sys.path.insert(0, ROOT)
import torch
import tac_spectral as ts
x = 0.1 * torch.randn(256, 1, 160000, device="cuda")
m = ts.MultiScaleSTFT((2048, 1024, 512, 256, 128), (512, 256, 128, 64, 32), (2048, 1024, 512, 256, 128))
for _ in range(3):
    out = m(x)
torch.cuda.synchronize()
print("ok")
