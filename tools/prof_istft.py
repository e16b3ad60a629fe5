import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "task-aware-audio-coding-perceptual_b200")); sys.path.insert(0, ROOT)
import torch
import tac_spectral as ts
dec = torch.randn(128, 2, 1025, 600, device="cuda"); dec[:, 0].abs_()
for _ in range(2): ts.batch_reconstruct_waveform(dec[:, 0], dec[:, 1], 16000)
torch.cuda.synchronize(); print("ok")
