"""tac_stft_ex: any-n_fft STFT (warp core or chirp-z) with fused epilogues, against the oracle / the reference's outputs.

Call sites: spectrogram.ipynb cell 9 (n_fft 2046 + amplitude_to_db), sgmse/sgmse/data_module.py:162-175,212-214 (n_fft 510,
power-law compression, pad_spec), testing_scripts.ipynb:246-249 (torchaudio Spectrogram() + 10 log10), SURVEY.md 8c item 5 (mel).
Tolerance (north star): fp32 rtol 1e-4 with atol = 1e-4 * max|ref|."""
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


def _close(got, ref, what, rtol=1e-4):
    ref = ref.to(got.dtype) if not ref.is_complex() else ref
    err = (got.cpu() - ref).abs()
    tol = rtol * ref.abs() + rtol * float(ref.abs().max())
    assert got.shape == ref.shape, (what, got.shape, ref.shape)
    assert bool((err <= tol).all()), f"{what}: max err {float(err.max()):.3e} at scale {float(ref.abs().max()):.3e}"


@pytest.mark.parametrize("n,h,center,pad_mode", [(2046, 512, True, "constant"), (510, 128, True, "reflect"), (400, 200, True, "reflect"),
                                                   (400, 200, False, "reflect"), (1000, 250, True, "reflect"), (90, 45, True, "constant"),
                                                   (2730, 683, True, "reflect"), (2048, 512, True, "reflect"), (512, 128, False, "reflect"),
                                                   (256, 77, True, "reflect")])
def test_complex_any_nfft(ts, n, h, center, pad_mode):
    L = ts._lib
    x, _ = O.synth_audio(3, 3 * n + 1234, 1)
    x = x[0]
    ref = O.torch_stft_ref(x, n, h, n, center=center, pad_mode=pad_mode)
    got = ts.stft_ex(x.cuda(), n, h, n, center=center, pad_mode=pad_mode, kind=L.EPI_COMPLEX)
    _close(got, ref, f"complex n={n}")
    # torchaudio's 'window' normalisation and a window shorter than the frame
    w = n - n // 4
    refn = O.torch_stft_ref(x, n, h, w, center=center, pad_mode=pad_mode, normalized=True)
    gotn = ts.stft_ex(x.cuda(), n, h, w, center=center, pad_mode=pad_mode, normalized=True, kind=L.EPI_COMPLEX)
    _close(gotn, refn, f"complex normalized n={n} win={w}")


def test_chirpz_agrees_with_warp_core(ts):
    # the same frames through both transforms: sqrt-hann forces the chirp-z path for a power-of-two length
    L = ts._lib
    x, _ = O.synth_audio(2, 9000, 1)
    x = x[0].cuda()
    a = ts.stft_ex(x, 1024, 256, 1024, window="sqrthann", center=True, kind=L.EPI_COMPLEX)
    ref = O.torch_stft_ref(x.cpu(), 1024, 256, 1024, window="sqrthann", center=True)
    _close(a, ref, "sqrt-hann chirp-z")


def test_sgmse_front_end_golden(ts, golden_dir):
    g = np.load(os.path.join(golden_dir, "sgmse_spec.npz"))
    sig = torch.from_numpy(g["sig"])
    fe = ts.SpecsFrontEnd()
    got = fe(sig.cuda())
    ref = torch.from_numpy(g["Yp"])
    _close(got, ref, "sgmse spec_fwd + pad_spec")
    assert bool((got[..., int(g["F"]):] == 0).all())                  # pad_spec's zero frames
    z = fe.stft(sig.cuda())
    _close(z, O.torch_stft_ref(sig, 510, 128, 510), "sgmse stft")
    none = ts.SpecsFrontEnd(transform_type="none")(sig.cuda(), pad64=False)
    _close(none, O.torch_stft_ref(sig, 510, 128, 510), "sgmse transform none")
    # leading dimensions are kept ([B, 1, T] -> [B, 1, K, F]) as test_single.py:148-149 builds them
    got4 = fe(sig[:, None, :].cuda())
    assert got4.shape == (2, 1, 256, 64) and torch.equal(got4[:, 0], got)


def test_power_and_log_spectrogram(ts):
    x, _ = O.synth_audio(4, 16000, 1)
    x = x[0]
    sp = ts.Spectrogram()                                            # testing_scripts.ipynb:246
    S = sp(x.cuda())
    _close(S, O.power_spectrogram_ref(x), "Spectrogram()")
    logS = sp.log_power(x.cuda())
    ref = O.power_spectrogram_ref(x, log=True)
    err = (logS.cpu() - ref).abs()
    # 10 log10(S + 1e-10): compare where S is above the fp32 noise floor of the transform
    mask = O.power_spectrogram_ref(x) > 1e-6 * float(O.power_spectrogram_ref(x).max())
    assert float(err[mask].max()) <= 1e-3, float(err[mask].max())
    S1 = ts.Spectrogram(n_fft=1024, hop_length=256, power=1.0, normalized=True, center=False)(x.cuda())
    _close(S1, O.torch_stft_ref(x, 1024, 256, 1024, center=False, normalized=True).abs(), "power=1")
    # differentiable like torchaudio's (power-of-two core): d sum(S) / dx against torch autograd
    xg = x[:2].cuda().requires_grad_(True)
    ts.Spectrogram(n_fft=512, hop_length=128, power=2.0)(xg).sum().backward()
    xr = x[:2].clone().requires_grad_(True)
    O.power_spectrogram_ref(xr, n_fft=512, hop_length=128).sum().backward()
    _close(xg.grad, xr.grad, "d Spectrogram(power=2) / dx", rtol=2e-4)


def test_db_front_end_2046(ts):
    # spectrogram.ipynb cell 9: n_fft = win_length = 2046, hop 512, magnitude / phase / amplitude_to_db(ref=np.max)
    x, _ = O.synth_audio(3, 16000, 1)
    x = x[0]
    d = ts.create_complex_spectrogram(x.cuda(), 16000, n_fft=2046, hop_length=512, win_length=2046, db_scale=True)
    mag, ph = O.frontend_ref(x, 2046, 512, 2046)
    assert d["magnitude"].shape == (3, 1024, 32)
    _close(d["magnitude"], mag, "magnitude 2046")
    sel = mag > 1e-3 * float(mag.max())
    dphi = (d["phase"].cpu() - ph + np.pi) % (2 * np.pi) - np.pi
    assert float(dphi[sel].abs().max()) <= 1e-3
    db = O.amplitude_to_db_ref(mag)
    # a magnitude within rtol 1e-4 + atol 1e-4 max moves its dB value by at most 20 log10(1 + tol / mag)
    allow = 20.0 * torch.log10(1.0 + (1e-4 * mag + 1e-4 * float(mag.max())) / mag.clamp_min(1e-5)) + 1e-4
    assert bool(((d["db_scale"].cpu() - db).abs() <= allow).all())
    assert float((d["db_scale"].cpu() - db).abs()[mag > 0.05 * float(mag.max())].max()) <= 2e-3
    one = ts.create_complex_spectrogram(x[0].cuda(), 16000, n_fft=2046, hop_length=512, win_length=2046, db_scale=True)
    assert one["db_scale"].shape == (1024, 32) and torch.equal(one["db_scale"], d["db_scale"][0])
    # fixed-reference dB and the mag + dB pair in one pass
    L = ts._lib
    fixed = ts.stft_ex(x.cuda(), 2046, 512, 2046, pad_mode="constant", kind=L.EPI_DB, p0=1e-5, p1=0.0, p2=-80.0)
    assert bool(((fixed.cpu() - O.amplitude_to_db_fixed_ref(mag, 1.0)).abs() <= allow).all())
    pair = ts.stft_ex(x.cuda(), 2046, 512, 2046, pad_mode="constant", kind=L.EPI_MAGDB, p0=1e-5, p1=0.0, p2=-80.0, f_pad=40)
    assert pair.shape == (3, 2, 1024, 40) and bool((pair[..., 32:] == 0).all())
    _close(pair[:, 0, :, :32], mag, "magdb: magnitude")
    assert torch.equal(pair[:, 1, :, :32], fixed)
    crop = ts.stft_ex(x.cuda(), 2046, 512, 2046, pad_mode="constant", kind=L.EPI_MAGPHASE, f_pad=20)
    assert crop.shape == (3, 2, 1024, 20) and torch.equal(crop[:, 0], d["magnitude"][..., :20])


@pytest.mark.parametrize("n,h,center", [(1024, 256, False), (400, 160, True), (2048, 512, False)])
def test_mel_epilogue(ts, n, h, center):
    x, _ = O.synth_audio(3, 16000, 1)
    x = x[0]
    m = ts.MelSpectrogram(16000, n, n, h, 0.0, 8000.0, n_mels=80, power=1.0, normalized=True, center=center)
    got = m(x.cuda())
    ref = O.mel_spectrogram_ref(x, n, h, center=center)
    _close(got, ref, f"mel n={n}", rtol=2e-4)
    lg = m.log_mel(x.cuda(), 10.0, 1e-10)
    assert float((lg.cpu() - 10 * torch.log10(ref + 1e-10)).abs().max()) <= 2e-3


def test_prepare_makes_first_call_capturable(ts):
    """the chirp-z tables are built by tac_prepare_stft_ex (constructors call it), never inside a capture; a
    configuration nobody prepared is refused during capture with a clear error instead of a cudaMalloc."""
    L = ts._lib
    x = torch.randn(3, 6000, device="cuda")
    sp = ts.Spectrogram(n_fft=346, hop_length=100, power=2.0)        # a size no other test uses: tables come from __init__
    ref = O.power_spectrogram_ref(x.cpu(), n_fft=346, hop_length=100, power=2.0)
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        with torch.cuda.graph(g, stream=s):
            out = sp(x)                                             # FIRST call of this configuration, under capture
    torch.cuda.current_stream().wait_stream(s)
    g.replay()
    torch.cuda.synchronize()
    eager = sp(x)
    assert torch.equal(out, eager)
    assert torch.allclose(out.cpu(), ref, rtol=1e-4, atol=1e-4 * float(ref.abs().max()))
    # unprepared configuration under capture: TAC_EINVAL naming the prepare call, and no allocation happened
    g2 = torch.cuda.CUDAGraph()
    with pytest.raises(RuntimeError, match="tac_prepare_stft_ex"):
        with torch.cuda.stream(s):
            with torch.cuda.graph(g2, stream=s):
                ts.stft_ex(x, 350, 100, 350, center=True, pad_mode="reflect", kind=L.EPI_COMPLEX)
    torch.cuda.synchronize()
    assert float(sp(x).sum()) == float(eager.sum())                  # context and library still fine


def test_specs_front_end_odd_n_fft_frame_count(ts):
    """centred frames = (T + 2 (n // 2) - n) // hop + 1, which is (T - 1) // hop + 1 for odd n_fft (not T // hop + 1)."""
    T, n, hop = 128 * 40, 511, 128                                   # T a multiple of hop: the two formulas differ by one
    sig = torch.randn(2, T)
    fe = ts.SpecsFrontEnd(n_fft=n, hop_length=hop)
    z = fe.stft(sig.cuda())
    ref = torch.stft(sig, n, hop, n, torch.hann_window(n, periodic=True), center=True, pad_mode="reflect", return_complex=True)
    assert z.shape == ref.shape and z.shape[-1] == (T - 1) // hop + 1 == 40
    assert torch.allclose(z.cpu(), ref, rtol=1e-4, atol=1e-4 * float(ref.abs().max()))
    out = fe(sig.cuda())
    assert out.shape[-1] == 64                                       # 40 frames pad to ONE 64-block (T // hop + 1 = 41 would too; see below)
    T2 = 128 * 64
    assert fe(torch.randn(1, T2).cuda()).shape[-1] == 64             # 64 frames exactly: the old formula (65) would have padded to 128
