"""Time the fused loss and front-end legs (CUDA events, cfg2).  python tools/time_legs.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "task-aware-audio-coding-perceptual_b200")); sys.path.insert(0, ROOT)
import torch
from bench import synth_host, WORKLOADS, N_FFTS, HOPS
from tac_spectral.step import SpectralStep
w = WORKLOADS[os.environ.get("WORKLOAD", "cfg2")]
dev = torch.device("cuda", 0)
sets = []
for i in range(6):
    x, y, zs = synth_host(w["batch"], w["views"], w["T"], w["d"], seed=i % 2)
    sets.append((x.to(dev) * (1 - 0.001 * i), y.to(dev) * (1 - 0.001 * i), [z.to(dev) for z in zs]))
st = SpectralStep(w["batch"] * w["views"], w["T"], w["views"], w["batch"], w["d"], w["dim_p"], N_FFTS, HOPS, device=dev, overlap=False)
def timeit(fn, n=30):
    for i in range(3): fn(i)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(n): fn(i)
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n
print("lib", os.environ.get("TAC_SPECTRAL_LIB", "default"))
print("fused    ms %.4f" % timeit(lambda i: st.leg_loss_fused(*sets[i % 6][:2])))
print("fwd      ms %.4f" % timeit(lambda i: st.leg_loss_fwd(*sets[i % 6][:2])))
print("frontend ms %.4f" % timeit(lambda i: st.leg_frontend(sets[i % 6][0])))
print("solve    ms %.4f" % timeit(lambda i: (st.leg_dpca_stats(sets[i % 6][2]), st.leg_dpca_solve())))
print("sums", st.sums.tolist())
print("stats    ms %.4f" % timeit(lambda i: st.leg_dpca_stats(sets[i % 6][2])))
