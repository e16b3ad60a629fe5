"""K3 parity: DPCA statistics / eigen-solve / selection / projection vs the oracle and the reference's
golden outputs.  Tolerances (north star / SURVEY.md 8c): largest principal angle < 1e-3 rad (float64 sine
formulation), singular values rtol 1e-4, ||z_out - z_ref||_inf <= 1e-4 * ||z_ref||_inf, per-view selected
component counts exactly equal."""
import os

import numpy as np
import pytest
import torch

from oracle import spectral_oracle as O

pytestmark = pytest.mark.gpu

ANGLE = 1e-3


@pytest.fixture(scope="module")
def ts():
    import tac_spectral
    return tac_spectral


def _zclose(got, ref, what=""):
    got, ref = got.detach().cpu(), ref.detach().cpu()
    err = float((got - ref).abs().max())
    assert err <= 1e-4 * float(ref.abs().max()), f"{what}: max err {err:.3e} vs scale {float(ref.abs().max()):.3e}"


@pytest.mark.parametrize("algo", ["tf32x3", "fp64"])
@pytest.mark.parametrize("tag", ["tall", "wide"])
def test_data_pca_golden(ts, golden_dir, algo, tag):
    """data_pca of the reference (pkl_file_model_updated.py:37-42) on tall (B > d) and wide (B < d) batches."""
    g = np.load(os.path.join(golden_dir, "dpca.npz"))
    z = torch.from_numpy(g[f"pca_{tag}_z"])
    s, v, mu = ts.data_pca(z.cuda(), algo=algo)
    r = min(z.shape)
    assert s.shape == (r,) and v.shape == (z.shape[1], r) and mu.shape == (z.shape[1],)
    s_ref = g[f"pca_{tag}_s"]
    live = s_ref > 1e-3 * s_ref.max()          # B <= d: the centred batch has rank B - 1, the last value is ~0
    np.testing.assert_allclose(s.cpu().numpy()[live], s_ref[live], rtol=1e-4, atol=1e-4 * s_ref.max())
    assert np.all(np.abs(s.cpu().numpy()[~live]) <= 2e-3 * s_ref.max())     # sqrt() of a ~1e-7-relative eigenvalue
    np.testing.assert_allclose(mu.cpu().numpy(), g[f"pca_{tag}_mu"], rtol=1e-5, atol=1e-6)
    vr = torch.from_numpy(g[f"pca_{tag}_v"])
    for k in (1, 4, 8):
        assert O.subspace_angle(v[:, :k].cpu(), vr[:, :k]) < ANGLE
    _zclose(ts.apply_pca(z.cuda(), 8, algo=algo), torch.from_numpy(g[f"pca_{tag}_apply8"]), "apply_pca")


@pytest.mark.parametrize("algo", ["tf32x3", "fp64"])
def test_dpca_block_golden(ts, golden_dir, algo):
    """ResE2D1 / ResE4D1 DPCA blocks run by the reference itself (tests/golden/dpca.npz)."""
    g = np.load(os.path.join(golden_dir, "dpca.npz"))
    zs = [torch.from_numpy(g["e2_z1"]), torch.from_numpy(g["e2_z2"])]
    outs = ts.dpca_project([z.cuda() for z in zs], int(g["e2_dim_p"]), algo=algo)
    _zclose(torch.cat(outs, 1), torch.from_numpy(g["e2_out"]), "E2 block")
    zs = [torch.from_numpy(g[f"e4_z{i}"]) for i in range(1, 5)]
    outs = ts.dpca_project([z.cuda() for z in zs], int(g["e4_dim_p"]), algo=algo)
    _zclose(torch.cat(outs, 1), torch.from_numpy(g["e4_out"]), "E4 block")


@pytest.mark.parametrize("B,d,nv,dim_p,algo", [
    (256, 128, 2, 32, "tf32x3"),       # cfg 2
    (1024, 128, 4, 128, "tf32x3"),     # cfg 3
    (1024, 128, 4, 8, "fp64"),
    (1024, 128, 4, 256, "tf32x3"),     # cfg 5 sweep end
    (300, 96, 3, 40, "tf32x3"),        # ragged: B not a multiple of the 64-row chunk, d < 128
    (512, 256, 2, 64, "fp64"),         # d = 256 stress variant (global-memory Jacobi)
    (77, 50, 2, 17, "fp64"),
])
def test_dpca_vs_oracle(ts, B, d, nv, dim_p, algo):
    zs = O.synth_latents(B, d, nv)
    ref_out, ref_idx = O.dpca_project_ref(zs, dim_p)
    outs, info = ts.dpca_project([z.cuda() for z in zs], dim_p, algo=algo, return_info=True)
    cnt = info["cnt"].cpu().tolist()
    assert cnt == [len(i) for i in ref_idx], (cnt, [len(i) for i in ref_idx])
    assert sum(cnt) == min(dim_p, nv * min(B, d))
    for v in range(nv):
        s_ref, v_ref, mu_ref = O.data_pca_ref(zs[v])
        r = min(B, d)
        np.testing.assert_allclose(info["s"][v, :r].cpu().numpy(), s_ref.numpy(), rtol=1e-4, atol=1e-4 * float(s_ref.max()))
        got_idx = info["idx"][v, :cnt[v]].cpu().long()
        assert sorted(got_idx.tolist()) == sorted(ref_idx[v].tolist())
        if cnt[v]:
            ang = O.subspace_angle(info["v"][v].cpu()[:, got_idx], v_ref[:, ref_idx[v]])
            assert ang < ANGLE, f"view {v}: subspace angle {ang:.3e}"
        _zclose(outs[v], ref_out[v], f"view {v}")


def test_rank_deficient_batch(ts):
    """Default training batch of the reference is 2 (train_pkl_file_task_aware.py:709): B < d."""
    zs = O.synth_latents(2, 64, 2)
    ref_out, ref_idx = O.dpca_project_ref(zs, 8)
    outs, info = ts.dpca_project([z.cuda() for z in zs], 8, algo="fp64", return_info=True)
    assert info["rank"].cpu().tolist() == [2, 2]
    assert info["cnt"].cpu().tolist() == [len(i) for i in ref_idx]
    for o, r in zip(outs, ref_out):
        _zclose(o, r, "B=2")


def test_var_topk(ts, golden_dir):
    g = np.load(os.path.join(golden_dir, "var_topk.npz"))
    z = torch.from_numpy(g["z"])
    sel, idx = ts.var_topk(z.cuda(), int(g["k"]))
    assert np.array_equal(sel.cpu().numpy(), g["z_sel"])
    z2 = O.synth_latents(333, 256, 1)[0]
    s2, i2 = ts.var_topk(z2.cuda(), 8)
    r2, ri2 = O.var_topk_ref(z2, 8)
    assert i2.cpu().tolist() == ri2.tolist() and torch.equal(s2.cpu(), r2)


def test_sharded_stats_equal_full_batch(ts):
    """Multi-GPU equivalence on one GPU: summing the statistics of batch shards
    and solving once reproduces the full-batch DPCA."""
    zs = O.synth_latents(512, 128, 4)
    full = ts.dpca_stats([z.cuda() for z in zs], "tf32x3")
    parts = sum(ts.dpca_stats([z[i * 128:(i + 1) * 128].cuda() for z in zs], "tf32x3") for i in range(4))
    scale = full.abs().max()
    assert float((full - parts).abs().max()) <= 1e-6 * float(scale)
    s, v, mu, rank = ts.dpca_eig(parts, 128)
    idx, cnt = ts.dpca_select(s, rank, 128)
    _, ref_idx = O.dpca_project_ref(zs, 128)
    assert cnt.cpu().tolist() == [len(i) for i in ref_idx]


@pytest.mark.parametrize("shape", [(256, 128, 2, 32), (64, 32, 2, 16), (24, 32, 2, 10)])
def test_project_backward_matches_reference_autograd(ts, shape):
    """The reference back-propagates through torch.svd inside data_pca (pkl_file_model_updated.py:37-42, 489-511):
    d loss / d z of dpca_project (basis_grad=True, the default) against torch autograd of the oracle's restatement,
    float64 on the CPU.  Includes a rank-deficient batch (B < d).  Tolerance: 1e-3 of the largest gradient entry
    (float32 eigenvectors and singular values enter 1 / (s_i^2 - s_j^2))."""
    B, d, nv, dim_p = shape
    zs = O.synth_latents(B, d, nv)
    torch.manual_seed(2)
    gs = [torch.randn(B, d, dtype=torch.float64) for _ in zs]
    zr = [z.double().requires_grad_(True) for z in zs]
    ref_out, ref_idx = O.dpca_project_ref(zr, dim_p)
    torch.autograd.backward(ref_out, gs)
    zg = [z.cuda().requires_grad_(True) for z in zs]
    outs, info = ts.dpca_project(zg, dim_p, algo="fp64", return_info=True)
    assert info["cnt"].cpu().tolist() == [len(i) for i in ref_idx]
    torch.autograd.backward(outs, [g.float().cuda() for g in gs])
    for v in range(nv):
        ref = zr[v].grad
        err = float((zg[v].grad.cpu().double() - ref).abs().max())
        assert err <= 1e-3 * float(ref.abs().max()), f"view {v}: {err:.3e} vs scale {float(ref.abs().max()):.3e}"


def test_data_pca_warns_when_graph_is_cut(ts):
    z = torch.randn(32, 16, device="cuda", requires_grad=True)
    with pytest.warns(UserWarning, match="detached"):
        ts.data_pca(z)


def test_project_backward(ts):
    """basis_grad=False: grad_z = (g - mean g) P + mean g with the projector held constant (stop-gradient form)."""
    zs = O.synth_latents(64, 32, 2)
    zg = [z.cuda().requires_grad_(True) for z in zs]
    outs, info = ts.dpca_project(zg, 16, algo="fp64", return_info=True, basis_grad=False)
    torch.manual_seed(1)
    gs = [torch.randn(64, 32) for _ in zs]
    torch.autograd.backward(outs, [g.cuda() for g in gs])
    for v in range(2):
        c = int(info["cnt"][v])
        Vs = info["v"][v].cpu()[:, info["idx"][v, :c].cpu().long()].double()
        P = Vs @ Vs.T
        g = gs[v].double()
        ref = (g - g.mean(0)) @ P + g.mean(0)
        assert torch.allclose(zg[v].grad.cpu().double(), ref, rtol=1e-4, atol=1e-5)


def test_dpca_errors(ts):
    z = torch.randn(8, 16, device="cuda")
    with pytest.raises(ValueError):
        ts.dpca_project([z, z[:, :8]], 4)
    with pytest.raises(RuntimeError):
        ts.dpca_project([z.cpu()], 4)
    with pytest.raises(ValueError):
        ts.dpca_project([torch.randn(8, 300, device="cuda")], 4)
    with pytest.raises(RuntimeError):
        ts.dpca_project([torch.randn(8, 200, device="cuda")], 4, algo="tf32x3")     # tensor-core Gram needs d <= 128


@pytest.mark.gpu
@pytest.mark.parametrize("B,d", [(256, 128), (64, 32), (300, 256), (16, 64)])
def test_nuclear_loss_forward_backward(ts, B, d):
    """Next row 8f-1: |zc/|zc|_F|_* / B (pkl_file_model_updated.py:526-532) from the Gram eigenvalues; fp32 rtol 1e-4
    on the value, gradient against torch autograd of the float64 oracle (rtol 1e-3 of the largest entry: 1/s amplifies
    the fp32 rounding of the small singular values)."""
    z = O.synth_latents(B, d, 1, seed=91)[0]
    ref_in = z.double().requires_grad_(True)
    ref = O.nuclear_loss_ref(ref_in)
    ref.backward()
    zg = z.cuda().requires_grad_(True)
    got = ts.nuclear_loss(zg)
    assert abs(float(got.detach()) - float(ref.detach())) <= 1e-4 * abs(float(ref.detach())), (float(got.detach()), float(ref.detach()))
    got.backward()
    err = (zg.grad.cpu().double() - ref_in.grad).abs().max()
    assert float(err) <= 1e-3 * float(ref_in.grad.abs().max()), float(err / ref_in.grad.abs().max())


def test_non_finite_latents_do_not_fault(ts):
    # NaN / Inf latents (an exploded encoder, or an uninitialised buffer in a warm-up pass) must give garbage values, never
    # a wild index: the rank-by-counting of the eigenvalues is made on sanitised values (dpca_eig_tri.cuh)
    torch.manual_seed(0)
    for bad in (float("nan"), float("inf")):
        zs = [torch.randn(64, 128, device="cuda") for _ in range(2)]
        zs[0][3, 5] = bad
        zs[1][:] = bad
        outs = ts.dpca_project(zs, 16)
        torch.cuda.synchronize()
        assert all(o.shape == (64, 128) for o in outs)
    z = torch.randn(64, 128, device="cuda")
    ref = ts.dpca_project([z], 16)[0]
    torch.cuda.synchronize()
    assert bool(torch.isfinite(ref).all())                           # the context is still healthy
