"""Accuracy + timing of tac_dpca_eig against numpy.linalg.eigh on float64 statistics (GPU box)."""
import os, sys, math
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "task-aware-audio-coding-perceptual_b200")); sys.path.insert(0, ROOT)
import numpy as np, torch
import tac_spectral as ts
from oracle import spectral_oracle as O

def case(z, name):
    z = z.float(); B, d = z.shape
    stats = ts.dpca_stats([z.cuda()], "fp64")
    s, v, mu, rank = ts.dpca_eig(stats, d)
    torch.cuda.synchronize()
    zc = (z.double() - z.double().mean(0)).numpy(); C = zc.T @ zc
    lr, Vr = np.linalg.eigh(C); lr = lr[::-1]; Vr = Vr[:, ::-1].copy()
    s = s[0].cpu().double().numpy(); V = v[0].cpu().double().numpy()
    fro = np.sqrt((C * C).sum()); r = int((lr > 1e-13 * fro).sum()); r = min(r, int(rank[0]))
    sref = np.sqrt(np.maximum(lr, 0))
    serr = np.abs(s[:r] - sref[:r]).max() / max(sref[0], 1e-300) if r else 0.0
    ks = [k for k in (1, 4, 8, 16, 32, r // 2, r) if 1 <= k <= r and (k == d or lr[k - 1] - lr[k] > 1e-6 * lr[0])]
    ang = max([O.subspace_angle(torch.from_numpy(V[:, :k].copy()), torch.from_numpy(Vr[:, :k].copy())) for k in ks] or [0])
    orth = np.abs(V[:, :r].T @ V[:, :r] - np.eye(r)).max() if r else 0.0
    print(f"{name:14s} r={r:3d} s err/s1 {serr:.1e} angle {ang:.1e} orth {orth:.1e}")

case(O.synth_latents(256, 128, 2)[0], "cfg2")
case(O.synth_latents(1024, 128, 4)[3], "cfg3")
case(O.synth_latents(2, 64, 2)[0], "B=2")
case(O.synth_latents(77, 50, 2)[1], "77x50")
case(O.synth_latents(40, 128, 1)[0], "wide40x128")
case(torch.randn(300, 96), "iso300x96")
case(O.synth_latents(256, 128, 1, decay=0.9)[0], "decay0.9")
zz = torch.randn(200, 64); zz[:, 10] = zz[:, 3]; zz[:, 20] = zz[:, 5] * 2
case(zz, "dupcols")
zz = torch.randn(500, 32); case(torch.cat([zz, zz.flip(1)], 1), "mirror")
case(torch.eye(64) * 3, "identity")
g, _ = torch.linalg.qr(torch.randn(400, 48)); q, _ = torch.linalg.qr(torch.randn(48, 48)); sig = torch.ones(48); sig[:10] = 5; sig[10:30] = 2.0
case((g * sig) @ q.T, "exactclusters")
case(torch.randn(50, 1), "d=1"); case(torch.randn(50, 2), "d=2"); case(torch.randn(50, 3), "d=3"); case(torch.zeros(10, 16), "zeros")
# timing
zs = [z.cuda() for z in O.synth_latents(256, 128, 2)]
stats = ts.dpca_stats(zs, "tf32x3")
for _ in range(3): ts.dpca_eig(stats, 128)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20): ts.dpca_eig(stats, 128)
e1.record(); torch.cuda.synchronize()
print("eig ms (incl. allocs)", e0.elapsed_time(e1) / 20)
import ctypes
from tac_spectral import _lib as L
buf = (ctypes.c_longlong * 8)()
L.lib().tac_debug_eig_profile(buf)
t = list(buf)
names = ["form C", "tridiag", "split", "multisection", "twisted", "back-transform", "outputs"]
for i, nm in enumerate(names):
    print(f"  {nm:16s} {t[i+1]-t[i]:9d} cycles")
