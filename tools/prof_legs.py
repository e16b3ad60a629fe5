"""Launch each leg of the cfg2 step a few times (for ncu / launch lists).  python tools/prof_legs.py [reps] [legs]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "task-aware-audio-coding-perceptual_b200")); sys.path.insert(0, ROOT)
import torch
from bench import synth_host, WORKLOADS, N_FFTS, HOPS
from tac_spectral.step import SpectralStep
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 2
legs = sys.argv[2].split(",") if len(sys.argv) > 2 else ["frontend", "fwd", "bwd", "dpca"]
w = WORKLOADS[os.environ.get("WORKLOAD", "cfg2")]
dev = torch.device("cuda", 0)
x, y, zs = synth_host(w["batch"], w["views"], w["T"], w["d"], seed=1)
x, y, zs = x.to(dev), y.to(dev), [z.to(dev) for z in zs]
st = SpectralStep(w["batch"] * w["views"], w["T"], w["views"], w["batch"], w["d"], w["dim_p"], N_FFTS, HOPS, device=dev, overlap=False)
for _ in range(reps):
    if "frontend" in legs: st.leg_frontend(x)
    if "fwd" in legs: st.leg_loss_fwd(x, y)
    if "bwd" in legs: st.leg_loss_bwd(x, y)
    if "dpca" in legs:
        st.leg_dpca_stats(zs); st.leg_dpca_solve(); st.leg_dpca_project(zs)
torch.cuda.synchronize()
print("ok", float(st.sums.sum()))
