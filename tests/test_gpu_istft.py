"""K4 parity: tac_istft (through the C ABI) vs torch.istft on the CPU (oracle) and the reference's own
batch_reconstruct_waveform outputs (tests/golden/istft.npz).

Tolerance of record (north star): fp32 rtol 1e-4 with atol = 1e-4 * max|ref|.
"""
import os

import numpy as np
import pytest
import torch

from oracle import spectral_oracle as O

pytestmark = pytest.mark.gpu

RTOL = 1e-4


def _close(got, ref, what=""):
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


def test_golden_batch_reconstruct(ts, golden_dir):
    g = np.load(os.path.join(golden_dir, "istft.npz"))
    mag, ph = torch.from_numpy(g["mag"]).cuda(), torch.from_numpy(g["phase"]).cuda()
    wav = ts.batch_reconstruct_waveform(mag, ph, 16000)
    _close(wav, torch.from_numpy(g["wav"]), "batch_reconstruct_waveform")
    one = ts.reconstruct_waveform(mag[0], ph[0], 16000)
    _close(one, torch.from_numpy(g["one"]), "reconstruct_waveform")
    # bit-stable: fixed summation order
    assert torch.equal(wav, ts.batch_reconstruct_waveform(mag, ph, 16000))


def test_reference_errors(ts):
    m = torch.zeros(2, 1025, 8, device="cuda")
    with pytest.raises(ValueError, match="shapes of magnitude and phase must match"):
        ts.batch_reconstruct_waveform(m, m[:, :, :-1], 16000)                 # train_pkl_file_task_aware.py:179-180
    with pytest.raises(ValueError, match="not supported"):
        ts.batch_reconstruct_waveform(m, m, 16000, window="hamming")          # :190-191
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ts.batch_reconstruct_waveform(m.cpu(), m.cpu(), 16000)
    spec = torch.zeros(2, 513, 6, dtype=torch.complex64, device="cuda")
    with pytest.raises(RuntimeError, match="envelope"):                       # torch.istft: window overlap add min
        ts.istft(spec, 1024, 1024, center=True)


def test_decoder_output_views_in_place(ts):
    """The two channels of the [B, 2, 1025, F_pad] decoder output are consumed as strided views (no copy)."""
    gen = torch.Generator().manual_seed(3)
    B, K, F = 5, 1025, 37
    dec = torch.randn(B, 2, K, F, generator=gen)
    dec[:, 0].abs_()
    ref = O.batch_reconstruct_waveform_ref(dec[:, 0], dec[:, 1])
    d = dec.cuda()
    got = ts.batch_reconstruct_waveform(d[:, 0], d[:, 1], 16000)
    _close(got, ref, "strided channel views")
    got2 = ts.batch_reconstruct_waveform(d[:, 0, :, :20], d[:, 1, :, :20], 16000)     # cropped frames, ld > F
    _close(got2, O.batch_reconstruct_waveform_ref(dec[:, 0, :, :20], dec[:, 1, :, :20]), "cropped views")


@pytest.mark.parametrize("n_fft,hop,win,F,B", [(2048, 512, 2048, 70, 3), (2048, 1024, 2048, 9, 2), (2048, 256, 2048, 41, 2),
                                                 (1024, 256, 1024, 33, 4), (1024, 256, 800, 20, 2), (512, 128, 512, 150, 3),
                                                 (512, 192, 512, 19, 2), (256, 64, 256, 77, 5), (128, 32, 128, 301, 3),
                                                 (64, 16, 64, 45, 9), (2048, 512, 2048, 2, 1)])
def test_complex_istft_matches_torch(ts, n_fft, hop, win, F, B):
    gen = torch.Generator().manual_seed(n_fft + hop + F)
    spec = torch.randn(B, n_fft // 2 + 1, F, dtype=torch.complex64, generator=gen)
    ref = O.istft_ref(spec, n_fft, hop, win)
    got = ts.istft(spec.cuda(), n_fft, hop, win, window=torch.hann_window(win).cuda())
    _close(got, ref, f"istft n={n_fft} hop={hop} win={win}")
    # single item, default hop / win
    if hop * 4 == n_fft and win == n_fft:
        _close(ts.istft(spec[0].cuda(), n_fft), O.istft_ref(spec[:1], n_fft, hop, win)[0], "single item")


def test_length_and_normalized(ts):
    gen = torch.Generator().manual_seed(12)
    spec = torch.randn(3, 513, 40, dtype=torch.complex64, generator=gen)
    nat = 256 * 39
    for length in (nat - 1000, nat, nat + 300):
        _close(ts.istft(spec.cuda(), 1024, 256, length=length), O.istft_ref(spec, 1024, 256, length=length), f"length {length}")
    # torch.istft(length > natural) reads on into the centre-padded tail, where the window envelope falls to ~1e-10 and
    # the quotient is ill-conditioned (torch itself returns O(1e3) there): compare where the envelope is sane, and the
    # zero fill behind n_fft + hop * (F - 1) exactly
    got, ref = ts.istft(spec.cuda(), 1024, 256, length=nat + 700).cpu(), O.istft_ref(spec, 1024, 256, length=nat + 700)
    _close(got[:, :nat + 400], ref[:, :nat + 400], "length nat + 700 (well-conditioned part)")
    assert torch.isfinite(got).all() and float(got[:, nat + 512:].abs().max()) == 0.0 and float(ref[:, nat + 512:].abs().max()) == 0.0
    _close(ts.istft(spec.cuda(), 1024, 256, normalized=True), O.istft_ref(spec, 1024, 256, normalized=True), "normalized")


def test_round_trip_full_size(ts):
    """Size-independent property at the cfg-2 batch: front end (K1) -> inverse (K4) returns the signal."""
    views, _ = O.synth_audio(256, 16000, 2)
    y = torch.cat(views, 0).cuda()                                            # [512, 16000]
    fe = ts.complex_spectrogram(y)                                            # [512, 2, 1025, 32]
    back = ts.batch_reconstruct_waveform(fe[:, 0], fe[:, 1], 16000)
    L = back.shape[1]
    assert L == 512 * (fe.shape[3] - 1)
    err = (back - y[:, :L]).abs().max()
    assert float(err) <= 2e-5 * float(y.abs().max()), float(err)
    # padded model input [B, 2, 1025, 600] (pad_tensor + stack): the zero frames synthesise silence behind the clip
    fe600 = ts.complex_spectrogram(y[:8], f_pad=600)
    back600 = ts.batch_reconstruct_waveform(fe600[:, 0], fe600[:, 1], 16000)
    assert back600.shape == (8, 512 * 599)                                    # 306688, cf. sgmse/test_single.py:202
    assert float((back600[:, :15000] - y[:8, :15000]).abs().max()) <= 2e-5 * float(y.abs().max())
    assert float(back600[:, 18000:].abs().max()) == 0.0


def test_gradients(ts):
    gen = torch.Generator().manual_seed(5)
    B, K, F = 2, 513, 12
    mag = torch.rand(B, K, F, generator=gen) + 0.1
    ph = (torch.rand(B, K, F, generator=gen) - 0.5) * 6.0
    wgt = torch.randn(B, 256 * (F - 1), generator=gen)
    mr, pr = mag.clone().requires_grad_(True), ph.clone().requires_grad_(True)
    (O.batch_reconstruct_waveform_ref(mr, pr, 1024, 256, 1024) * wgt).sum().backward()
    mg, pg = mag.cuda().requires_grad_(True), ph.cuda().requires_grad_(True)
    (ts.batch_reconstruct_waveform(mg, pg, 16000, 1024, 256, 1024) * wgt.cuda()).sum().backward()
    _close(mg.grad, mr.grad, "d/d magnitude")
    _close(pg.grad, pr.grad, "d/d phase")
    spec = torch.randn(B, K, F, dtype=torch.complex64, generator=gen)
    sr = spec.clone().requires_grad_(True)
    (O.istft_ref(sr, 1024, 256) * wgt).sum().backward()
    sg = spec.cuda().requires_grad_(True)
    (ts.istft(sg, 1024, 256) * wgt.cuda()).sum().backward()
    # torch's autograd keeps (ignored) imaginary gradients at k = 0, n/2 at zero as well
    _close(torch.view_as_real(sg.grad), torch.view_as_real(sr.grad), "d/d spec")


# ---- chain: decoder spectrogram -> waveform -> discriminator STFT inputs (tac_istft_disc_inputs) -------------------------

def test_chain_golden_and_oracle(ts, golden_dir):
    """Waveform against the reference's own batch_reconstruct_waveform output (istft.npz); the per-scale discriminator inputs
    against the oracle (pinned by stft_disc.npz) applied to that waveform.  Tolerance: rtol 1e-4, atol 1e-4 max|ref|."""
    g = np.load(os.path.join(golden_dir, "istft.npz"))
    mag, ph = torch.from_numpy(g["mag"]), torch.from_numpy(g["phase"])
    wav, zs = ts.reconstruct_and_disc_inputs(mag.cuda(), ph.cuda(), 16000)
    _close(wav, torch.from_numpy(g["wav"]), "chain waveform")
    ref_wav, ref_zs = O.reconstruct_and_disc_inputs_ref(mag, ph)
    _close(wav, ref_wav, "chain waveform vs oracle")
    assert len(zs) == len(ref_zs) == 3
    for z, r, n in zip(zs, ref_zs, (1024, 2048, 512)):
        _close(z, r, f"chain disc input n_fft={n}")


@pytest.mark.parametrize("rows,group", [(5, 1), (5, 2), (5, 0), (5, 64), (3, 3)])
def test_chain_equals_separate_calls(ts, rows, group):
    """Row groups (ragged last group included) change nothing: bit-identical to batch_reconstruct_waveform + MultiScaleSTFT."""
    gen = torch.Generator().manual_seed(11)
    F = 40
    dec = torch.randn(rows, 2, 1025, F + 8, generator=gen).cuda()              # decoder layout, padded time axis
    mag, ph = dec[:, 0, :, :F].abs(), dec[:, 1, :, :F] * 3.0                   # strided views, used in place
    wav, zs = ts.reconstruct_and_disc_inputs(mag, ph, 16000, rows_per_group=group)
    wav2 = ts.batch_reconstruct_waveform(mag, ph, 16000)
    zs2 = ts.MultiScaleSTFT()(wav2.unsqueeze(1))
    assert torch.equal(wav, wav2)
    for a, b in zip(zs, zs2):
        assert a.shape == b.shape and torch.equal(a, b)


def test_chain_five_scales_and_errors(ts):
    gen = torch.Generator().manual_seed(12)
    mag, ph = torch.rand(2, 1025, 24, generator=gen), torch.randn(2, 1025, 24, generator=gen)
    n_ffts, hops = (2048, 1024, 512, 256, 128), (512, 256, 128, 64, 32)
    wav, zs = ts.reconstruct_and_disc_inputs(mag.cuda(), ph.cuda(), 16000, disc_n_ffts=n_ffts, disc_hop_lengths=hops,
                                             disc_win_lengths=n_ffts)
    ref_wav, ref_zs = O.reconstruct_and_disc_inputs_ref(mag, ph, n_ffts, hops)
    _close(wav, ref_wav, "chain waveform (5 scales)")
    for z, r, n in zip(zs, ref_zs, n_ffts):
        _close(z, r, f"chain disc input n_fft={n}")
    with pytest.raises(ValueError, match="shapes of magnitude and phase must match"):
        ts.reconstruct_and_disc_inputs(mag.cuda(), ph.cuda()[:, :, :-1])
    with pytest.raises(ValueError, match="shorter than n_fft"):
        ts.reconstruct_and_disc_inputs(mag.cuda()[:, :, :3], ph.cuda()[:, :, :3])   # 1024 samples < n_fft = 2048
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ts.reconstruct_and_disc_inputs(mag, ph)
