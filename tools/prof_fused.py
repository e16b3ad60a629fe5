import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "task-aware-audio-coding-perceptual_b200")); sys.path.insert(0, ROOT)
import torch
from bench import synth_host, WORKLOADS, N_FFTS, HOPS
from tac_spectral.step import SpectralStep
w = WORKLOADS["cfg2"]
dev = torch.device("cuda", 0)
x, y, zs = synth_host(w["batch"], w["views"], w["T"], w["d"], seed=1)
x, y, zs = x.to(dev), y.to(dev), [z.to(dev) for z in zs]
st = SpectralStep(w["batch"] * w["views"], w["T"], w["views"], w["batch"], w["d"], w["dim_p"], N_FFTS, HOPS, device=dev, overlap=False)
which = sys.argv[1] if len(sys.argv) > 1 else "fused"
for _ in range(3):
    if which == "fused": st.leg_loss_fused(x, y)
    elif which == "frontend": st.leg_frontend(x)
torch.cuda.synchronize()
print("ok")
