"""K1 parity: tac_stft_multi (through the C ABI) vs the CPU oracle and the reference's golden outputs.

Tolerance of record (BASELINE.json north_star / SURVEY.md 8c): fp32 rtol 1e-4 with
atol = 1e-4 * max|ref| for near-zero bins; phase compared modulo 2*pi where |X| > 1e-3 * max|X|.
"""
import os

import numpy as np
import pytest
import torch

from oracle import spectral_oracle as O

pytestmark = pytest.mark.gpu

RTOL = 1e-4


def _close(got: torch.Tensor, ref: torch.Tensor, what=""):
    got, ref = got.detach().cpu(), ref.detach().cpu()
    assert got.shape == ref.shape, f"{what}: shape {tuple(got.shape)} vs {tuple(ref.shape)}"
    atol = RTOL * float(ref.abs().max())
    err = (got - ref).abs()
    bad = err > atol + RTOL * ref.abs()
    assert not bool(bad.any()), f"{what}: {int(bad.sum())} of {bad.numel()} off, max err {float(err.max()):.3e} (atol {atol:.3e})"


@pytest.fixture(scope="module")
def ts():
    import tac_spectral
    return tac_spectral


def test_golden_disc_inputs(ts, golden_dir):
    """The reference's own spec_transform + cat + rearrange outputs (tests/golden/stft_disc.npz)."""
    g = np.load(os.path.join(golden_dir, "stft_disc.npz"))
    x = torch.from_numpy(g["x"]).cuda()
    cfgs = [tuple(int(v) for v in g[f"cfg_{i}"]) for i in range(3)]
    outs = ts.MultiScaleSTFT([c[0] for c in cfgs], [c[1] for c in cfgs], [c[2] for c in cfgs])(x)
    for i, o in enumerate(outs):
        _close(o, torch.from_numpy(g[f"planes_{i}"]), f"planes scale {i}")
    for i, (n, h, w) in enumerate(cfgs):
        sp = ts.Spectrogram(n_fft=n, hop_length=h, win_length=w, window_fn=torch.hann_window, normalized=True,
                            center=False, pad_mode=None, power=None)          # msstftd.py:125-127 verbatim
        z = sp(x)
        ref = torch.from_numpy(g[f"complex_{i}"])
        assert z.dtype == torch.complex64 and z.shape == ref.shape
        _close(torch.view_as_real(z), torch.view_as_real(ref), f"complex scale {i}")
        _close(sp.planes(x), torch.from_numpy(g[f"planes_{i}"]), f"Spectrogram.planes scale {i}")


def test_kat_shapes(ts):
    """testing_scripts.ipynb:2021-2023: (2,1,24000) -> F x K = 90x513, 43x1025, 184x257."""
    x = torch.randn(2, 1, 24000, device="cuda")
    outs = ts.MultiScaleSTFT()(x)
    assert [tuple(o.shape) for o in outs] == [(2, 2, 90, 513), (2, 2, 43, 1025), (2, 2, 184, 257)]


@pytest.mark.parametrize("n", [64, 128, 256, 512, 1024, 2048])
def test_single_scale_all_sizes(ts, n):
    torch.manual_seed(n)
    T = 5 * n + 37                                   # ragged: not a multiple of hop, unaligned rows
    x = torch.randn(3, 2, T)
    got = ts.stft_multi(x.cuda(), [n], [n // 4])[0]
    _close(got, O.disc_inputs_ref(x, n, n // 4), f"n={n}")


def test_five_scales_one_launch(ts):
    """cfg 4 scale list (n = 2048..128, hop = n/4), small batch."""
    (x,), _ = O.synth_audio(3, 16000, 1)
    n_ffts = [2048, 1024, 512, 256, 128]
    hops = [n // 4 for n in n_ffts]
    outs = ts.stft_multi(x[:, None, :].cuda(), n_ffts, hops)
    refs = O.multiscale_disc_inputs_ref(x[:, None, :], n_ffts, hops)
    for n, o, r in zip(n_ffts, outs, refs):
        _close(o, r, f"5-scale n={n}")


def test_multichannel_and_hop_variants(ts):
    torch.manual_seed(7)
    x = torch.randn(2, 3, 9001)
    for n, h, w in [(512, 64, 512), (1024, 512, 1024), (256, 256, 256), (1024, 256, 800), (512, 128, 400)]:
        got = ts.stft_multi(x.cuda(), [n], [h], [w])[0]
        _close(got, O.disc_inputs_ref(x, n, h, w), f"n={n} hop={h} win={w}")


def test_minimum_length_and_single_frame(ts):
    x = torch.randn(2, 1, 2048)
    got = ts.stft_multi(x.cuda(), [2048, 512], [512, 128])
    refs = O.multiscale_disc_inputs_ref(x, [2048, 512], [512, 128])
    assert got[0].shape[2] == 1
    for o, r in zip(got, refs):
        _close(o, r, "minimum length")


@pytest.mark.parametrize("pad_mode", ["reflect", "constant"])
def test_centered_complex(ts, pad_mode):
    torch.manual_seed(11)
    x = torch.randn(3, 7000)
    for n, h in [(2048, 512), (512, 128), (128, 32)]:
        sp = ts.Spectrogram(n_fft=n, hop_length=h, win_length=n, power=None, normalized=False, center=True, pad_mode=pad_mode)
        z = sp(x.cuda())
        ref = O.stft_ref(x, n, h, n, center=True, pad_mode=pad_mode, normalized=False)
        _close(torch.view_as_real(z), torch.view_as_real(ref), f"center {pad_mode} n={n}")


def test_frontend_mag_phase(ts):
    """Notebook front end: |D|, angle(D) of the centred 2048/512 STFT, [B, 2, 1025, F] and padded to 600."""
    (y,), _ = O.synth_audio(4, 16000, 1, seed=99)
    mag, ph = O.frontend_ref(y)
    out = ts.complex_spectrogram(y.cuda())
    assert out.shape == (4, 2, 1025, 32)
    _close(out[:, 0], mag, "magnitude")
    big = mag > 1e-3 * mag.max()
    dphi = torch.angle(torch.exp(1j * (out[:, 1].cpu() - ph)))
    assert float(dphi[big].abs().max()) < 1e-3, float(dphi[big].abs().max())
    out600 = ts.complex_spectrogram(y.cuda(), f_pad=600)
    assert out600.shape == (4, 2, 1025, 600)
    assert torch.equal(out600[..., :32], out) and float(out600[..., 32:].abs().sum()) == 0.0
    crop = ts.complex_spectrogram(y.cuda(), f_pad=20)
    assert torch.equal(crop, out[..., :20])
    d = ts.create_complex_spectrogram(y[0].cuda(), 16000)
    assert set(d) == {"magnitude", "phase", "params"} and d["magnitude"].shape == (1025, 32)
    assert torch.equal(d["magnitude"], out[0, 0])


def test_long_clip_property(ts):
    """Full-size clip (10 s): Parseval-type identity instead of an element-wise oracle.
    With hop = n/4 and a periodic Hann window sum_f w^2(t - f*hop) = 1.5 for interior samples, so
    sum_{f,k} c_k |X_f[k]|^2 * (sum w^2) / n = 1.5 * sum_interior x^2 + edge terms; checked against the
    float64 evaluation of the same quantity by the oracle on one row."""
    (x,), _ = O.synth_audio(2, 160000, 1, seed=5)
    for n in (2048, 128):
        got = ts.stft_multi(x[:, None, :].cuda(), [n], [n // 4])[0].double().cpu()      # [B, 2, F, K]
        ref = O.disc_inputs_ref(x[:, None, :].double(), n, n // 4)
        e_got = (got ** 2).sum(dim=(1, 3))
        e_ref = (ref ** 2).sum(dim=(1, 3))
        assert torch.allclose(e_got, e_ref, rtol=1e-4), f"frame energies differ (n={n})"


def test_errors(ts):
    x = torch.randn(1, 1, 4000, device="cuda")
    with pytest.raises(RuntimeError):
        ts.stft_multi(x, [1000], [250])                      # unsupported n_fft surfaces tac_last_error()
    with pytest.raises(ValueError):
        ts.stft_multi(x, [8192 // 2], [1024])                # signal shorter than n_fft
    with pytest.raises(RuntimeError):
        ts.stft_multi(torch.randn(1, 1, 4000), [512], [128])  # CPU tensor: no fallback
    with pytest.raises(ValueError):
        ts.Spectrogram(n_fft=512, window_fn=torch.hamming_window, power=None)


@pytest.mark.gpu
def test_amplitude_to_db(ts):
    """dB channel of the notebook front end (cell 9): librosa.amplitude_to_db(|D|, ref=np.max), per utterance; absolute
    tolerance 1e-3 dB (20*log10 amplifies the 1e-7 relative error of |D| by 8.7)."""
    (y,), _ = O.synth_audio(3, 16000, 1, seed=17)
    y[1] *= 1e-3                                              # a quiet utterance: its own reference level
    mag, _ = O.frontend_ref(y)
    ref = O.amplitude_to_db_ref(mag)
    spec = ts.complex_spectrogram(y.cuda(), f_pad=40)         # padded layout, 32 real frames
    got = ts.amplitude_to_db(spec, n_frames=32)
    assert got.shape == (3, 1025, 32)
    assert float((got.cpu() - ref).abs().max()) < 1e-3
    assert float(got.max()) == 0.0 and float(got.min()) >= -80.0
    plain = ts.amplitude_to_db(mag.cuda().contiguous())
    assert float((plain.cpu() - ref).abs().max()) < 1e-3
    with pytest.raises(ValueError):
        ts.amplitude_to_db(spec, n_frames=41)


# ------------------------------------------------------------------------------------------------ backward (autograd)
def _torch_stft(x, n, h, win, center, pad_mode, normalized):
    w = torch.hann_window(win, dtype=x.dtype)
    X = torch.stft(x, n, h, win, w, center=center, pad_mode=pad_mode if center else "reflect", return_complex=True)
    return X / w.pow(2).sum().sqrt() if normalized else X


@pytest.mark.parametrize("cfg", [(1024, 256, 1024, False, "reflect", True), (2048, 512, 2048, True, "constant", False),
                                 (512, 128, 512, True, "reflect", True), (256, 64, 200, False, "reflect", True)])
def test_stft_backward_matches_torch(ts, cfg):
    """torchaudio's Spectrogram (msstftd.py:125-127) is differentiable: d loss / d waveform through the complex
    layout [B, C, K, F] and through the discriminator planes [B, 2C, F, K] against torch.stft autograd (float64, CPU).
    Tolerance: 1e-4 of the largest gradient entry."""
    n, h, win, center, pad_mode, normalized = cfg
    torch.manual_seed(3)
    x = 0.3 * torch.randn(3, 2, 6000)
    xr = x.double().requires_grad_(True)
    Xr = _torch_stft(xr.reshape(6, -1), n, h, win, center, pad_mode, normalized).reshape(3, 2, n // 2 + 1, -1)
    G = torch.randn(Xr.shape, dtype=torch.float64) + 1j * torch.randn(Xr.shape, dtype=torch.float64)
    (Xr.real * G.real + Xr.imag * G.imag).sum().backward()
    ref = xr.grad
    # complex layout
    from tac_spectral import _lib as L
    xg = x.cuda().requires_grad_(True)
    Xc = ts.stft_multi(xg, [n], [h], [win], center=center, pad_mode=pad_mode, normalized=normalized, out_mode=L.OUT_COMPLEX)[0]
    Gc = G.to(torch.complex64).cuda()
    (Xc.real * Gc.real + Xc.imag * Gc.imag).sum().backward()
    err = float((xg.grad.cpu().double() - ref).abs().max())
    assert err <= 1e-4 * float(ref.abs().max()), f"complex: {err:.3e} vs {float(ref.abs().max()):.3e}"
    # planes layout (cat(real, imag) + rearrange b c w t -> b c t w)
    xg2 = x.cuda().requires_grad_(True)
    Pl = ts.stft_multi(xg2, [n], [h], [win], center=center, pad_mode=pad_mode, normalized=normalized, out_mode=L.OUT_PLANES)[0]
    Gp = torch.cat([G.real, G.imag], 1).permute(0, 1, 3, 2).float().cuda()
    (Pl * Gp).sum().backward()
    err = float((xg2.grad.cpu().double() - ref).abs().max())
    assert err <= 1e-4 * float(ref.abs().max()), f"planes: {err:.3e} vs {float(ref.abs().max()):.3e}"


def test_frontend_backward_matches_torch(ts):
    """Magnitude / phase front end (spectrogram.ipynb cell 10): polar chain rule on top of the adjoint transform."""
    torch.manual_seed(4)
    x = 0.3 * torch.randn(2, 8000)
    xr = x.double().requires_grad_(True)
    D = torch.stft(xr, 2048, 512, 2048, torch.hann_window(2048, dtype=torch.float64), center=True, pad_mode="constant",
                   return_complex=True)
    gm, gp = torch.randn(D.shape, dtype=torch.float64), 0.01 * torch.randn(D.shape, dtype=torch.float64)
    (D.abs() * gm + D.angle() * gp).sum().backward()
    xg = x.cuda().requires_grad_(True)
    out = ts.complex_spectrogram(xg)
    (out[:, 0] * gm.float().cuda() + out[:, 1] * gp.float().cuda()).sum().backward()
    ref = xr.grad
    err = float((xg.grad.cpu().double() - ref).abs().max())
    assert err <= 2e-4 * float(ref.abs().max()), f"{err:.3e} vs {float(ref.abs().max()):.3e}"


def test_multiscale_module_is_differentiable(ts):
    x = (0.1 * torch.randn(2, 1, 16000)).cuda().requires_grad_(True)
    outs = ts.MultiScaleSTFT()(x)
    sum(o.square().sum() for o in outs).backward()
    assert x.grad is not None and float(x.grad.abs().max()) > 0


# ---- persistent double-buffered planes kernel (single-scale calls with >= 4 * 148 (row, chunk) units) -------------------------

@pytest.mark.parametrize("n,T,center", [(2048, 16000, False), (1024, 16000, False), (512, 12288, False), (256, 16000, True),
                                        (128, 9000, False), (64, 8192, True), (2048, 70000, True)])
def test_persistent_single_scale(ts, n, T, center):
    """Enough rows that the single-scale call takes the persistent kernel (bulk-copied, double-buffered tiles; full / empty
    mbarrier pipeline): spot rows against the oracle (rtol 1e-4), and the whole output bit-equal to the one-shot kernel, which
    the same scale takes when a second scale rides along.  Ragged row ends (T not a multiple of the chunk), zero centre padding
    (tiles that hang over both row ends) and every plan are covered."""
    B = 640 if T <= 16000 else 160
    gen = torch.Generator().manual_seed(n + T)
    x = (0.3 * torch.randn(B, 1, T, generator=gen)).cuda()
    other = 512 if n != 512 else 256
    one = ts.stft_multi(x, [n], [n // 4], center=center, pad_mode="constant")[0]
    two = ts.stft_multi(x, [n, other], [n // 4, other // 4], center=center, pad_mode="constant")[0]
    assert torch.equal(one, two), "persistent and one-shot planes kernels disagree"
    rows = [0, 1, B // 2, B - 1]
    xc = x[rows].cpu()
    if center:
        xc = torch.nn.functional.pad(xc, (n // 2, n // 2))
    _close(one[rows], O.disc_inputs_ref(xc, n, n // 4), f"persistent n={n} T={T} center={center}")


@pytest.mark.parametrize("n_ffts", [(1024, 2048, 512), (2048, 1024, 512, 256, 128), (64,)])
@pytest.mark.parametrize("power,logm", [(1.0, 0.0), (2.0, 0.0), (2.0, 10.0), (0.5, 0.0)])
def test_multi_scale_magnitude_epilogues_one_launch(ts, n_ffts, power, logm):
    """tac_stft_multi_epi: |z|^p and p1 log10(|z|^p + eps) of several scales from ONE launch of the multi-scale kernel (north
    star kernel 1), against the oracle's per-scale loop (msstftd.py:185-192 convention, magnitude / power taken in float64
    from stft_ref; testing_scripts.ipynb:246-249 for the log form).  fp32 rtol 1e-4, atol 1e-4 * max|ref|."""
    L = ts._lib
    (x,), _ = O.synth_audio(3, 9001 + 2 * max(n_ffts), 1, seed=31)
    x = torch.stack([x, 0.5 * x.flip(-1)], 1)                        # [B, C = 2, T]: two channels, ragged T
    hops = [n // 4 for n in n_ffts]
    c0 = int(L.lib().tac_launch_count())
    outs = ts.stft_multi_power(x.cuda(), list(n_ffts), hops, power=power, log_multiplier=logm, log_eps=1e-10)
    assert int(L.lib().tac_launch_count()) - c0 == 1                 # one kernel for all scales
    for n, h, o in zip(n_ffts, hops, outs):
        B, Cc, T = x.shape
        ref = O.stft_ref(x.reshape(B * Cc, T).double(), n, h, n).abs() ** power       # [B*C, K, F]
        if logm:
            ref = logm * torch.log10(ref + 1e-10)
        ref = ref.reshape(B, Cc, *ref.shape[-2:]).transpose(-1, -2).float()           # -> [B, C, F, K]
        assert o.shape == ref.shape, (o.shape, ref.shape)
        _close(o, ref, f"n={n} power={power} log={logm}")
    # the module form and the planes form agree: sqrt(re^2 + im^2) of the same launch family
    if power == 1.0 and not logm:
        ms = ts.MultiScaleSTFT(list(n_ffts), hops, list(n_ffts))
        mags = ms.magnitudes(x.cuda())
        planes = ms(x.cuda())
        for m, p in zip(mags, planes):
            Cc = x.shape[1]
            want = torch.sqrt(p[:, :Cc] ** 2 + p[:, Cc:] ** 2)
            assert torch.allclose(m, want, rtol=1e-6, atol=1e-7 * float(want.max()))


def test_multi_scale_magnitude_errors(ts):
    L = ts._lib
    x = torch.randn(2, 1, 5000, device="cuda")
    with pytest.raises(RuntimeError, match="positive"):
        ts.stft_multi_power(x, [512], [128], power=0.0)
    with pytest.raises(NotImplementedError):
        ts.stft_multi_power(x.clone().requires_grad_(True), [512], [128])
    with pytest.raises(ValueError):
        ts.stft_multi_power(x, [8192 * 2], [128])                     # shorter than n_fft is caught first
