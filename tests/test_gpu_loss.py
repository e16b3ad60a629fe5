"""K2 / K2' parity: fused MR-STFT loss forward + backward and the reference's spectral reductions.

Tolerances (north star): loss values fp32 rtol 1e-4; gradients rtol 1e-4 with atol = 1e-4 * max|grad_ref|
plus a small allowance for the sign() kinks of |.| (elements where X == Y or a magnitude sits on the log floor).
"""
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


def _rel(a, b):
    return abs(float(a) - float(b)) / max(abs(float(b)), 1e-30)


@pytest.mark.parametrize("B,T", [(3, 16000), (2, 4099), (1, 2048)])
def test_mrstft_forward_default_scales(ts, B, T):
    (x,), (y,) = O.synth_audio(B, T, 1)
    ref, terms = O.mrstft_loss_ref(x, y, return_terms=True)
    got, gterms = ts.mrstft_loss(x.cuda(), y.cuda(), return_terms=True)
    assert _rel(got, ref) < 1e-4, (float(got), float(ref))
    for (a, b), (c, d) in zip(gterms, terms):
        assert _rel(a, c) < 1e-4 and _rel(b, d) < 1e-4


def test_mrstft_forward_five_scales_and_views(ts):
    views, decs = O.synth_audio(4, 16000, 2)
    n_ffts = [2048, 1024, 512, 256, 128]
    x = torch.stack(views, 1)                       # [B, V, T]: views ride on the channel axis
    y = torch.stack(decs, 1)
    ref = O.mrstft_loss_ref(x.reshape(-1, 16000), y.reshape(-1, 16000), n_ffts)
    got = ts.mrstft_loss(x.cuda(), y.cuda(), n_ffts)
    assert _rel(got, ref) < 1e-4


def test_mrstft_identical_inputs_and_floor(ts):
    (x,), _ = O.synth_audio(2, 8000, 1)
    assert float(ts.mrstft_loss(x.cuda(), x.cuda())) == 0.0
    z = torch.zeros(2, 8000)
    ref = O.mrstft_loss_ref(x, z)                   # every |Y| sits on the log floor
    got = ts.mrstft_loss(x.cuda(), z.cuda())
    assert _rel(got, ref) < 1e-4


def test_mrstft_deterministic(ts):
    (x,), (y,) = O.synth_audio(8, 16000, 1)
    a = ts.mrstft_sums(x.cuda(), y.cuda(), [1024, 2048, 512], [256, 512, 128])
    b = ts.mrstft_sums(x.cuda(), y.cuda(), [1024, 2048, 512], [256, 512, 128])
    assert torch.equal(a, b)


def _grads(ts, x, y, n_ffts, eps):
    """(kernel, torch-fp32 oracle, float64 oracle) gradients of the loss w.r.t. y."""
    y64 = y.double().requires_grad_(True)
    O.mrstft_loss_ref(x.double(), y64, n_ffts, eps=eps).backward()
    y32 = y.clone().requires_grad_(True)
    O.mrstft_loss_ref(x, y32, n_ffts, eps=eps).backward()
    yg = y.cuda().requires_grad_(True)
    ts.mrstft_loss(x.cuda(), yg, n_ffts, eps=eps).backward()
    return yg.grad.double().cpu(), y32.grad.double(), y64.grad


CASES = [(2, 16000, (1024, 2048, 512)), (3, 6002, (512, 128)), (1, 20000, (2048, 64, 256)), (2, 16000, (2048, 1024, 512, 256, 128))]


def _check_grad(got, ref, what):
    """fp32 rtol 1e-4 (atol = 1e-4 * max|grad|) element-wise, robust to kink flips.

    The loss is piecewise smooth: sign(X - Y) of one bin flips when |X - Y| is within fp32 rounding of 0
    (|X - Y| < ~1e-7 happens for ~1e-5 of the bins of the seeded clips, i.e. a handful per test case), and one
    flipped bin perturbs the n_fft samples of its frame by ~1/sqrt(4K) of their value.  torch's own fp32 path
    shows the same isolated frames against a float64 run (measured: 1e-4..3e-2 relative L2).  So: all but a
    few frames must meet the element-wise tolerance, the typical element must be at fp32 rounding level, and
    the flips must stay isolated."""
    assert got.shape == ref.shape
    scale = float(ref.abs().max())
    err = (got - ref).abs()
    bad = err > 1e-4 * scale + 1e-4 * ref.abs()
    bad_frac, med, l2 = float(bad.float().mean()), float(err.median()) / scale, float(err.norm() / ref.norm())
    print(f"{what}: bad fraction {bad_frac:.3e}, median error / scale {med:.2e}, relative L2 {l2:.2e}")
    # Measured on B200 over the four CASES and both floors (tools/diag_grad_tol.py, gpurun_out/grad_tol_r02.log): bad fraction
    # 0 .. 1.3e-2 (ONE flipped bin of a 1024-sample frame marks ~400 of 32 000 elements), median 3e-8 .. 6e-7, relative L2
    # 4e-7 .. 8.3e-4; torch's own fp32 path against float64 shows flips of the same size on these inputs (4.3e-3 / 1.8e-4 on
    # the five-scale case).  Bounds = 2.5x the worst measured case; every element is checked strictly where no bin can flip
    # (test_mrstft_backward_no_kinks, all CASES, both signs).
    assert bad_frac < 0.03, f"{what}: {int(bad.sum())} of {bad.numel()} elements off, max {float(err.max()):.2e} vs scale {scale:.2e}"
    assert med < 2e-6, f"{what}: median error {float(err.median()):.2e} vs scale {scale:.2e}"
    assert l2 < 2e-3, what


@pytest.mark.parametrize("B,T,n_ffts", CASES)
def test_mrstft_backward_well_conditioned(ts, B, T, n_ffts):
    """Log floor at 1e-2: the 1/|Y| factor of the log-magnitude gradient is bounded."""
    (x,), (y,) = O.synth_audio(B, T, 1, seed=77)
    got, _, ref = _grads(ts, x, y, n_ffts, eps=1e-2)
    _check_grad(got, ref, "eps=1e-2")


@pytest.mark.parametrize("B,T,n_ffts", CASES)
def test_mrstft_backward_default_floor(ts, B, T, n_ffts):
    """Definition-of-record floor (1e-5): bins with tiny |Y| additionally carry gradients ~ 1/|Y| that amplify
    the fp32 rounding of the transform (torch-fp32 vs float64: ~1e-4 relative L2 on these inputs)."""
    (x,), (y,) = O.synth_audio(B, T, 1, seed=77)
    got, t32, ref = _grads(ts, x, y, n_ffts, eps=O.LOG_FLOOR)
    _check_grad(got, ref, "eps=1e-5")
    e_ours = float((got - ref).norm() / ref.norm())
    e_t32 = float((t32 - ref).norm() / ref.norm())
    print(f"relative L2 vs float64: kernel {e_ours:.2e}, torch-fp32 {e_t32:.2e}")


@pytest.mark.parametrize("eps", [1e-2, O.LOG_FLOOR])
@pytest.mark.parametrize("B,T,n_ffts", CASES)
def test_mrstft_backward_strict_off_kinks(ts, B, T, n_ffts, eps):
    """Realistic, NON-proportional (x, y) -- the inputs of the statistical check above -- with the frames that hold a bin near a
    kink masked out by the float64 oracle: EVERY remaining sample must meet rtol 1e-4 / atol 1e-4 * max|grad|.  A systematic
    error confined to a few samples (an edge frame per row, a run boundary of the streaming kernel's hand-over) cannot hide
    behind the kink allowance here; the masked samples keep the statistical bounds of _check_grad."""
    (x,), (y,) = O.synth_audio(B, T, 1, seed=77)
    got, _, ref = _grads(ts, x, y, n_ffts, eps=eps)
    mask = O.kink_sample_mask(x, y, n_ffts, eps)
    keep = ~mask
    frac = float(mask.float().mean())
    assert frac < 0.65, f"mask covers {frac:.2f} of the samples: the strict check would be vacuous"
    scale = float(ref.abs().max())
    err = (got - ref).abs()
    bad = (err > 1e-4 * scale + 1e-4 * ref.abs()) & keep
    worst = float((err * keep).max()) / scale
    print(f"eps={eps:g} {n_ffts}: masked {frac:.3f}, strict samples {int(keep.sum())}, worst error / scale {worst:.2e}, off {int(bad.sum())}")
    assert not bool(bad.any()), (f"{int(bad.sum())} of {int(keep.sum())} kink-free samples off; worst {worst:.2e} of the scale at "
                                 f"(row, sample) {torch.nonzero(bad)[:4].tolist()}")
    # edge coverage: the first and last n_fft samples of some row are among the strictly checked ones
    nmax = max(n_ffts)
    assert bool(keep[:, :nmax].any()) and bool(keep[:, -nmax:].any())


@pytest.mark.parametrize("gain", [0.5, 1.5])
@pytest.mark.parametrize("B,T,n_ffts", CASES)
def test_mrstft_backward_no_kinks(ts, B, T, n_ffts, gain):
    """Strict check where no bin sits near a kink: y = gain * x makes Y = gain * X in every bin, so sign(X - Y) is the
    same everywhere (+1 for gain 0.5, -1 for 1.5) and the float64 oracle must be matched to rtol 1e-4 on EVERY element --
    all plans, ragged lengths, run boundaries and edge frames of CASES included."""
    (x,), _ = O.synth_audio(B, T, 1, seed=21)
    y = gain * x
    got, _, ref = _grads(ts, x, y, n_ffts, eps=1e-2)
    scale = float(ref.abs().max())
    err = (got - ref).abs()
    assert bool((err <= 1e-4 * scale + 1e-4 * ref.abs()).all()), float(err.max() / scale)
    assert float(err.norm() / ref.norm()) < 1e-5


def test_mrstft_backward_scaled_upstream(ts):
    (x,), (y,) = O.synth_audio(2, 9000, 1, seed=3)
    yg = y.cuda().requires_grad_(True)
    (3.0 * ts.mrstft_loss(x.cuda(), yg)).backward()
    g3 = yg.grad.clone()
    yg.grad = None
    ts.mrstft_loss(x.cuda(), yg).backward()
    assert torch.allclose(g3, 3.0 * yg.grad, rtol=1e-4, atol=1e-5 * float(g3.abs().max()))
    yg.grad = None
    ts.mrstft_loss(x.cuda(), yg).backward()
    g1 = yg.grad.clone()
    yg.grad = None
    ts.mrstft_loss(x.cuda(), yg).backward()
    assert torch.equal(g1, yg.grad), "backward must be bit-stable run to run"


def test_mrstft_gradcheck_direction(ts):
    """Directional central difference in float64 on the oracle vs the kernel's gradient (h small enough not
    to cross sign(X - Y) kinks: the perturbation is 1e-6 against |X - Y| ~ 1e-2)."""
    (x,), (y,) = O.synth_audio(1, 5000, 1, seed=8)
    torch.manual_seed(0)
    v = torch.randn_like(y)
    yg = y.cuda().requires_grad_(True)
    ts.mrstft_loss(x.cuda(), yg, (512, 256)).backward()
    d_kernel = float((yg.grad.cpu().double() * v.double()).sum())
    h = 1e-6
    f = lambda t: float(O.mrstft_loss_ref(x.double(), t, (512, 256)))
    d_fd = (f(y.double() + h * v.double()) - f(y.double() - h * v.double())) / (2 * h)
    assert abs(d_kernel - d_fd) <= 2e-3 * abs(d_fd), (d_kernel, d_fd)


def test_spec_losses_golden(ts, golden_dir):
    """Reference outputs of ResE2D1.forward's loss block (pkl_file_model_updated.py:520-540)."""
    g = np.load(os.path.join(golden_dir, "spec_losses.npz"))
    r = ts.spectral_losses(torch.from_numpy(g["obs"]).cuda(), torch.from_numpy(g["dec"]).cuda())
    for k in ("magnitude_loss", "phase_loss", "total_loss", "spec_snr"):
        assert _rel(r[k], g[k]) < 1e-5, k
    assert _rel(r["mse_mean"], g["mse_mean"]) < 1e-5
    assert _rel(r["mse"].mean(), g["mse_mean"]) < 1e-5


@pytest.mark.parametrize("B,K,F", [(4, 1025, 600), (3, 65, 21), (1, 7, 3)])
def test_spec_losses_vs_oracle_and_backward(ts, B, K, F):
    torch.manual_seed(B * K)
    obs = torch.rand(B, 2, K, F) * 2.0
    dec = obs + 0.1 * torch.randn(B, 2, K, F)
    ref = O.spectral_losses_ref(obs, dec)
    got = ts.spectral_losses(obs.cuda(), dec.cuda())
    for k in ("magnitude_loss", "phase_loss", "total_loss", "spec_snr"):
        assert _rel(got[k], ref[k]) < 1e-5, k
    assert torch.allclose(got["mse"].cpu(), ref["mse_per_sample"], rtol=1e-5)
    # backward: a generic combination of all outputs (train loss adds mse.mean(), spec total, snr is logged)
    wv = torch.linspace(0.5, 1.5, B)

    def combo(r, mse):
        return 0.7 * r["magnitude_loss"] + 1.3 * r["phase_loss"] + 0.4 * r["total_loss"] - 0.01 * r["spec_snr"] + (mse * wv.to(mse.device)).sum()

    dr = dec.clone().requires_grad_(True)
    rr = O.spectral_losses_ref(obs, dr)
    combo(rr, rr["mse_per_sample"]).backward()
    dg = dec.cuda().requires_grad_(True)
    rg = ts.spectral_losses(obs.cuda(), dg)
    (combo(rg, rg["mse"]) + 0.2 * rg["mse_mean"]).backward()
    dr2 = dec.clone().requires_grad_(True)
    rr2 = O.spectral_losses_ref(obs, dr2)
    (combo(rr2, rr2["mse_per_sample"]) + 0.2 * rr2["mse_per_sample"].mean()).backward()
    assert torch.allclose(dg.grad.cpu(), dr2.grad, rtol=1e-4, atol=1e-4 * float(dr2.grad.abs().max()))


def test_loss_errors(ts):
    x = torch.randn(2, 4000, device="cuda")
    with pytest.raises(ValueError):
        ts.mrstft_loss(x, x[:, :3000])
    with pytest.raises(RuntimeError):
        ts.mrstft_loss(x.cpu(), x.cpu())
    with pytest.raises(RuntimeError):
        ts.mrstft_loss(x, x, (1000,))
    with pytest.raises(ValueError):
        ts.spectral_losses(torch.zeros(2, 3, 4, 5, device="cuda"), torch.zeros(2, 3, 4, 5, device="cuda"))
