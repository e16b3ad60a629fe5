"""Rows a9 and a5-i: the bottleneck that is live in the SpectralRes models (pkl_file_models.py:1108-1127, :692-718,
:335-354) with the DPCA hook (:302-332, :636-679), and the loss tail of task_aware() (train_pkl_file_task_aware.py:312-364),
against the reference's own outputs (tests/golden/bottleneck.npz, task_aware_tail.npz) and the oracle."""
import os

import numpy as np
import pytest
import torch

from oracle import spectral_oracle as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ts():
    import tac_spectral
    return tac_spectral


def _g(golden_dir, name):
    return np.load(os.path.join(golden_dir, name))


def test_bottleneck_mix_golden(ts, golden_dir):
    g = _g(golden_dir, "bottleneck.npz")
    z1, z2 = torch.from_numpy(g["e2d2_z1"]).cuda(), torch.from_numpy(g["e2d2_z2"]).cuda()
    tfa, d = int(g["e2d2_tfa"]), z1.shape[1]
    alpha = torch.from_numpy(g["e2d2_alpha"]).cuda()
    out = ts.bottleneck_mix([z1, z2], d // 2, tfa // 4, [alpha])
    w = out.shape[1] // 2
    assert [w, w] == list(g["e2d2_after"])
    # one FMA-free expression per element, same operation order as the reference: bit-equal
    assert np.array_equal(out[:, :w].cpu().numpy(), g["e2d2_mix_s1"]) and np.array_equal(out[:, w:].cpu().numpy(), g["e2d2_mix_s2"])
    out = ts.bottleneck_mix([z1, z2], d // 2, tfa // 4)
    assert np.array_equal(out[:, :w].cpu().numpy(), g["e2d2_nomix_s1"]) and np.array_equal(out[:, w:].cpu().numpy(), g["e2d2_nomix_s2"])
    zs = [torch.from_numpy(g[f"e4d1_z{i + 1}"]).cuda() for i in range(4)]
    samples, info = ts.spectral_bottleneck(zs, tfa, True, model="E4D1")
    assert np.array_equal(torch.cat(samples, 1).cpu().numpy(), g["e4d1_mix"])
    assert info["before_z3"] == d and info["after_z4"] == tfa // 4


def test_bottleneck_mix_alpha_stream_and_backward(ts):
    torch.manual_seed(5)
    B, d, tfa = 37, 20, 12
    zs = [torch.randn(B, d) for _ in range(4)]
    gen = torch.Generator(device="cuda").manual_seed(123)
    samples, _ = ts.spectral_bottleneck([z.cuda() for z in zs], tfa, True, model="E2D2", generator=gen)
    gen = torch.Generator(device="cuda").manual_seed(123)
    alphas = [torch.rand(B, 1, device="cuda", generator=gen).cpu() for _ in range(2)]      # the same draws (:1116)
    ref = O.bottleneck_mix_ref(zs, d // 2, tfa // 4, alphas)
    for s, r in zip(samples, ref):
        assert torch.allclose(s.cpu(), r, rtol=0, atol=1e-6)
    # adjoint against torch autograd of the oracle
    zg = [z.clone().cuda().requires_grad_(True) for z in zs]
    wgt = torch.randn(B, 4 * (tfa // 4) * 2)
    (ts.bottleneck_mix(zg, d // 2, tfa // 4, [a.cuda() for a in alphas]) * wgt.cuda()).sum().backward()
    zr = [z.clone().requires_grad_(True) for z in zs]
    (torch.cat(O.bottleneck_mix_ref(zr, d // 2, tfa // 4, alphas), 1) * wgt).sum().backward()
    for a, b in zip(zg, zr):
        assert torch.allclose(a.grad.cpu(), b.grad, rtol=1e-6, atol=1e-6)


def test_dpca_hook_golden(ts, golden_dir):
    # the DPCA projection swapped in at the bottleneck (pkl_file_models.py:302-332 and :636-679, run as written upstream)
    g = _g(golden_dir, "bottleneck.npz")
    z1, z2 = torch.from_numpy(g["e2d2_z1"]).cuda(), torch.from_numpy(g["e2d2_z2"]).cuda()
    outs, info = ts.spectral_bottleneck([z1, z2], 0, True, mode="dpca")
    assert [info["after_z1"], info["after_z2"]] == list(g["dpca2_after"])
    ref = g["dpca2_out"]
    assert np.abs(torch.cat(outs, 1).cpu().numpy() - ref).max() <= 1e-4 * np.abs(ref).max()
    zs = [torch.from_numpy(g[f"e4d1_z{i + 1}"]).cuda() for i in range(4)]
    outs, info = ts.spectral_bottleneck(zs, 0, True, mode="dpca", dim_p=int(g["dpca4_dim_p"]))
    assert [info[f"after_z{i + 1}"] for i in range(4)] == list(g["dpca4_after"])
    ref = g["dpca4_out"]
    assert np.abs(torch.cat(outs, 1).cpu().numpy() - ref).max() <= 1e-4 * np.abs(ref).max()
    same, info = ts.spectral_bottleneck(zs, 0, False, mode="dpca")
    assert all(a is b for a, b in zip(same, zs)) and info["after_z1"] == zs[0].shape[1]


def test_task_aware_tail_golden(ts, golden_dir):
    g = _g(golden_dir, "task_aware_tail.npz")

    def pack(tag):
        lg = [torch.from_numpy(g[f"{tag}_logit_{i}"]).cuda() for i in range(3)]
        fm = [[torch.from_numpy(g[f"{tag}_fmap_{i}_{j}"]).cuda() for j in range(5)] for i in range(3)]
        return lg, fm
    mse, dl = ts.task_aware_tail(torch.from_numpy(g["enhanced"]).cuda(), torch.from_numpy(g["clean"]).cuda(), pack("enh"), pack("gt"))
    assert abs(float(mse) - float(g["mse"])) <= 1e-4 * float(g["mse"])             # north star: losses at fp32 rtol 1e-4
    assert abs(float(dl) - float(g["disc_loss"])) <= 1e-4 * float(g["disc_loss"])


@pytest.mark.parametrize("La,Lb", [(958, 306688), (306688, 958), (5000, 5000), (7, 3), (1, 9)])
def test_interp_mse_vs_oracle(ts, La, Lb):
    torch.manual_seed(La + Lb)
    a, b = torch.randn(3, La), torch.randn(3, Lb)
    got = float(ts.interp_mse(a.cuda(), b.cuda()))
    ref = float(O.interp_mse_ref(a.double(), b.double()))
    assert abs(got - ref) <= 1e-4 * abs(ref)


def test_multi_reduce_many_segments(ts):
    torch.manual_seed(9)
    a = [torch.randn(n) for n in (1, 31, 8192, 8193, 100000)] * 12          # 60 segments: more than one call
    b = [torch.randn_like(t) for t in a]
    out = ts.multi_reduce([t.cuda() for t in a[:40]], [t.cuda() for t in b[:40]], [0, 1] * 20)
    ref = [(x - y).abs().mean().double() if i % 2 == 0 else ((x - y) ** 2).mean().double() for i, (x, y) in enumerate(zip(a[:40], b[:40]))]
    assert torch.allclose(out[1:].cpu(), torch.stack(ref), rtol=1e-5)
    assert abs(float(out[0]) - float(sum(ref))) <= 1e-5 * float(sum(ref))
    # deterministic: a second run is bit-identical
    out2 = ts.multi_reduce([t.cuda() for t in a[:40]], [t.cuda() for t in b[:40]], [0, 1] * 20)
    assert torch.equal(out, out2)
    with pytest.raises(RuntimeError):
        ts.multi_reduce([t.cuda() for t in a[:49]], [t.cuda() for t in b[:49]], [0] * 49)
