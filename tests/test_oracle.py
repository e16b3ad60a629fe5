"""The oracle against the reference's own outputs (tests/golden/*.npz, produced by
tests/golden/make_golden.py from /root/reference) and against an independent float64 DFT.
CPU only."""
import os

import numpy as np
import pytest
import torch

from oracle import spectral_oracle as O


def _load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name))


def test_shape_known_answers(golden_dir):
    # testing_scripts.ipynb:2021-2023 -- logits of MultiScaleSTFTDiscriminator(filters=32) on (2,1,24000)
    g = _load(golden_dir, "stft_disc.npz")
    assert g["kat_logit_shapes"].tolist() == [[2, 1, 90, 65], [2, 1, 43, 129], [2, 1, 184, 33]]
    assert g["kat_n_fmaps"].tolist() == [5, 5, 5]
    # frame counts behind those shapes: F = 1 + (T - n)//hop, K = n/2 + 1
    for (n, h), F in zip(((1024, 256), (2048, 512), (512, 128)), (90, 43, 184)):
        assert O.num_frames(24000, n, h, False) == F
    # spectrogram.ipynb:4210-4211 -- (1025, 123) for one utterance, centred STFT 2048/512
    T = 122 * 512 + 17
    assert O.num_frames(T, 2048, 512, True) == 123
    mag, ph = O.frontend_ref(torch.zeros(1, 5000))
    assert mag.shape == (1, 1025, 1 + 5000 // 512) == ph.shape


def test_stft_matches_reference_spectrogram(golden_dir):
    g = _load(golden_dir, "stft_disc.npz")
    x = torch.from_numpy(g["x"])
    for i in range(3):
        n, h, w = (int(v) for v in g[f"cfg_{i}"])
        z = O.stft_ref(x[:, 0], n, h, w, center=False, normalized=True)
        assert np.array_equal(z.numpy(), g[f"complex_{i}"][:, 0])          # bit-equal: same torch.stft call
        planes = O.disc_inputs_ref(x, n, h, w)
        assert np.array_equal(planes.numpy(), g[f"planes_{i}"])
    outs = O.multiscale_disc_inputs_ref(x)
    assert [tuple(o.shape) for o in outs] == [tuple(g[f"planes_{i}"].shape) for i in range(3)]


@pytest.mark.parametrize("n,h", [(128, 32), (256, 64), (512, 128)])
def test_stft_against_direct_dft(n, h):
    torch.manual_seed(3)
    x = torch.randn(2, 3 * n + 17)
    z = O.stft_ref(x, n, h, n).numpy()
    ref = O.dft_frames_f64(x.numpy(), n, h)
    scale = np.abs(ref).max()
    assert np.abs(z - ref).max() <= 2e-6 * scale


def test_frontend_conventions():
    torch.manual_seed(4)
    y = torch.randn(2, 4000)
    mag, ph = O.frontend_ref(y, pad_mode="constant")
    # centred frames of a zero-padded signal == uncentred frames of the explicitly padded one
    yp = torch.nn.functional.pad(y, (1024, 1024))
    ref = O.dft_frames_f64(yp.numpy(), 2048, 512, normalized=False)
    assert np.abs(mag.numpy() - np.abs(ref)).max() <= 2e-6 * np.abs(ref).max()
    big = np.abs(ref) > 1e-2 * np.abs(ref).max()
    dphi = np.angle(np.exp(1j * (ph.numpy() - np.angle(ref))))
    assert np.abs(dphi[big]).max() < 1e-4
    m600 = O.pad_frames(mag, 600)
    assert m600.shape[-1] == 600 and torch.equal(m600[..., :mag.shape[-1]], mag) and m600[..., mag.shape[-1]:].abs().sum() == 0
    assert O.pad_frames(mag, 5).shape[-1] == 5
    db = O.amplitude_to_db_ref(mag)
    assert db.max() <= 1e-6 and db.min() >= -80.0 - 1e-4


def test_data_pca_matches_reference(golden_dir):
    g = _load(golden_dir, "dpca.npz")
    for tag in ("tall", "wide"):
        z = torch.from_numpy(g[f"pca_{tag}_z"])
        s, v, mu = O.data_pca_ref(z)
        assert s.shape[0] == min(z.shape)
        np.testing.assert_allclose(s.numpy(), g[f"pca_{tag}_s"], rtol=1e-5, atol=1e-6)
        np.testing.assert_allclose(mu.numpy(), g[f"pca_{tag}_mu"], rtol=0, atol=0)
        # eigenvector signs are arbitrary -> compare projectors
        vr = torch.from_numpy(g[f"pca_{tag}_v"])
        r = 8
        assert O.subspace_angle(v[:, :r], vr[:, :r]) < 1e-5
        np.testing.assert_allclose(O.apply_pca_ref(z, 8).numpy(), g[f"pca_{tag}_apply8"], rtol=1e-4, atol=1e-5)


def test_dpca_block_matches_reference(golden_dir):
    g = _load(golden_dir, "dpca.npz")
    zs = [torch.from_numpy(g["e2_z1"]), torch.from_numpy(g["e2_z2"])]
    outs, idx = O.dpca_project_ref(zs, int(g["e2_dim_p"]))
    assert sum(len(i) for i in idx) == int(g["e2_dim_p"])
    np.testing.assert_allclose(torch.cat(outs, 1).numpy(), g["e2_out"], rtol=1e-4, atol=1e-5)
    zs = [torch.from_numpy(g[f"e4_z{i}"]) for i in range(1, 5)]
    outs, idx = O.dpca_project_ref(zs, int(g["e4_dim_p"]))
    assert sum(len(i) for i in idx) == int(g["e4_dim_p"])
    np.testing.assert_allclose(torch.cat(outs, 1).numpy(), g["e4_out"], rtol=1e-4, atol=1e-5)


def test_spec_losses_match_reference(golden_dir):
    g = _load(golden_dir, "spec_losses.npz")
    r = O.spectral_losses_ref(torch.from_numpy(g["obs"]), torch.from_numpy(g["dec"]))
    for k in ("magnitude_loss", "phase_loss", "total_loss", "spec_snr"):
        np.testing.assert_allclose(r[k].numpy(), g[k], rtol=1e-6)
    np.testing.assert_allclose(r["mse_per_sample"].mean().numpy(), g["mse_mean"], rtol=1e-6)
    np.testing.assert_allclose(O.nuclear_loss_ref(torch.from_numpy(g["z_sample"])).numpy(), g["nuc"], rtol=1e-5)


def test_var_topk_matches_reference(golden_dir):
    g = _load(golden_dir, "var_topk.npz")
    sel, idx = O.var_topk_ref(torch.from_numpy(g["z"]), int(g["k"]))
    assert np.array_equal(sel.numpy(), g["z_sel"])


def test_mrstft_loss_definition():
    (x,), (y,) = O.synth_audio(3, 4000, 1)
    total, terms = O.mrstft_loss_ref(x, y, return_terms=True)
    assert len(terms) == 3 and total > 0
    assert O.mrstft_loss_ref(x, x).item() == 0.0
    # hand evaluation of one scale from the float64 DFT
    X = np.abs(O.dft_frames_f64(x.numpy(), 512, 128)); Y = np.abs(O.dft_frames_f64(y.numpy(), 512, 128))
    l1 = np.abs(X - Y).mean(); ll = np.abs(np.log(np.maximum(X, 1e-5)) - np.log(np.maximum(Y, 1e-5))).mean()
    np.testing.assert_allclose(terms[2][0].item(), l1, rtol=1e-4)
    np.testing.assert_allclose(terms[2][1].item(), ll, rtol=1e-4)
    y2 = y.clone().requires_grad_(True)
    O.mrstft_loss_ref(x, y2).backward()
    assert y2.grad.shape == y.shape and torch.isfinite(y2.grad).all()


def test_subspace_angle_and_synth():
    zs = O.synth_latents(64, 16, 2)
    assert zs[0].shape == (64, 16)
    s, v, mu = O.data_pca_ref(zs[0])
    assert O.subspace_angle(v[:, :4], v[:, :4] @ torch.linalg.qr(torch.randn(4, 4))[0]) < 1e-6
    assert abs(O.subspace_angle(v[:, :1], v[:, 1:2]) - np.pi / 2) < 1e-6


def test_istft_matches_reference(golden_dir):
    """batch_reconstruct_waveform / reconstruct_waveform of train_pkl_file_task_aware.py:174-226 (tests/golden/istft.npz,
    produced by the reference's own functions) and an independent float64 inverse DFT + overlap-add."""
    g = np.load(os.path.join(golden_dir, "istft.npz"))
    mag, ph = torch.from_numpy(g["mag"]), torch.from_numpy(g["phase"])
    wav = O.batch_reconstruct_waveform_ref(mag, ph)
    assert wav.shape == (3, 512 * (14 - 1))                       # hop * (F - 1), cf. 512*(600-1) = 306688 in the reference
    assert torch.allclose(wav, torch.from_numpy(g["wav"]), rtol=1e-6, atol=1e-7)
    assert torch.allclose(wav[:1], torch.from_numpy(g["one"]), rtol=1e-6, atol=1e-7)
    spec = (mag * torch.exp(1j * ph)).to(torch.complex64)
    assert torch.allclose(O.istft_ref(spec, 2048, 512), wav, rtol=1e-6, atol=1e-7)
    # float64 direct synthesis of item 0
    X = spec[0].to(torch.complex128).numpy()
    n, hop, F = 2048, 512, X.shape[1]
    w = O.hann_periodic(n, torch.float64).numpy()
    ola, env = np.zeros(n + hop * (F - 1)), np.zeros(n + hop * (F - 1))
    for f in range(F):
        ola[f * hop:f * hop + n] += np.fft.irfft(X[:, f], n) * w
        env[f * hop:f * hop + n] += w * w
    ref = (ola / np.maximum(env, 1e-30))[n // 2:n // 2 + hop * (F - 1)]
    assert np.abs(wav[0].numpy() - ref).max() <= 2e-6 * np.abs(ref).max()
    with pytest.raises(ValueError):
        O.batch_reconstruct_waveform_ref(mag, ph[:, :-1])


def test_cosine_loss_matches_reference(golden_dir):
    """cos_loss of ResE4D1.forward / ResE2D1.forward (pkl_file_model_updated.py:429-430, 513-514)."""
    g = np.load(os.path.join(golden_dir, "cosine.npz"))
    zs = [torch.from_numpy(g[f"c4_z{i + 1}"]) for i in range(4)]
    assert torch.allclose(O.pairwise_cosine_loss_ref(zs), torch.from_numpy(g["c4_cos"]), rtol=1e-6, atol=1e-7)
    assert torch.allclose(O.pairwise_cosine_loss_ref(zs[:2]), torch.from_numpy(g["c2_cos"]), rtol=1e-6, atol=1e-7)


def test_sgmse_front_end_matches_reference(golden_dir):
    # sgmse/sgmse/data_module.py:162-175,212-214 + util/other.py:76-90, compiled from the reference by make_golden.gen_sgmse
    g = _load(golden_dir, "sgmse_spec.npz")
    got = O.sgmse_spec_ref(torch.from_numpy(g["sig"]))
    assert got.shape == g["Yp"].shape == (2, 256, 64) and int(g["F"]) == 6000 // 128 + 1
    assert np.array_equal(got.numpy(), g["Yp"])                       # same torch calls -> bit-equal
    assert np.all(g["Yp"][..., int(g["F"]):] == 0)


@pytest.mark.parametrize("n,h", [(2046, 512), (510, 128), (400, 200)])
def test_any_nfft_stft_against_direct_dft(n, h):
    # the non-power-of-two frame lengths of spectrogram.ipynb:2897-2901, data_module.py:109-110 and torchaudio's default
    torch.manual_seed(5)
    x = torch.randn(2, 2 * n + 123)
    z = O.torch_stft_ref(x, n, h, n, center=False).numpy()
    ref = O.dft_frames_f64(x.numpy(), n, h, normalized=False)
    assert np.abs(z - ref).max() <= 2e-6 * np.abs(ref).max()


def test_power_and_mel_refs_shapes():
    x = torch.randn(2, 4000)
    S = O.power_spectrogram_ref(x)
    assert S.shape == (2, 201, 21) and float(S.min()) >= 0.0          # torchaudio defaults: n_fft 400, hop 200, centred
    assert torch.allclose(O.power_spectrogram_ref(x, log=True), 10 * torch.log10(S + 1e-10))
    M = O.mel_spectrogram_ref(x, 1024, 256)
    assert M.shape == (2, 80, 1 + (4000 - 1024) // 256)


def _tail_golden(golden_dir):
    g = _load(golden_dir, "task_aware_tail.npz")
    def pack(tag):
        lg = [torch.from_numpy(g[f"{tag}_logit_{i}"]) for i in range(3)]
        fm = [[torch.from_numpy(g[f"{tag}_fmap_{i}_{j}"]) for j in range(5)] for i in range(3)]
        return lg, fm
    return g, pack("enh"), pack("gt")


def test_bottleneck_matches_reference(golden_dir):
    # pkl_file_models.py:1108-1127 / :692-718 through the reference's own forward (make_golden.gen_bottleneck)
    g = _load(golden_dir, "bottleneck.npz")
    z1, z2 = torch.from_numpy(g["e2d2_z1"]), torch.from_numpy(g["e2d2_z2"])
    tfa, d = int(g["e2d2_tfa"]), z1.shape[1]
    s1, s2 = O.bottleneck_mix_ref([z1, z2], d // 2, tfa // 4, [torch.from_numpy(g["e2d2_alpha"])])
    assert np.array_equal(s1.numpy(), g["e2d2_mix_s1"]) and np.array_equal(s2.numpy(), g["e2d2_mix_s2"])
    assert list(g["e2d2_after"]) == [s1.shape[1], s2.shape[1]]
    n1, n2 = O.bottleneck_mix_ref([z1, z2], d // 2, tfa // 4)
    assert np.array_equal(n1.numpy(), g["e2d2_nomix_s1"]) and np.array_equal(n2.numpy(), g["e2d2_nomix_s2"])
    zs = [torch.from_numpy(g[f"e4d1_z{i + 1}"]) for i in range(4)]
    torch.manual_seed(92)
    al = [torch.rand(48, 1), torch.rand(48, 1)]                      # :698-699 (the shared parts are empty: no effect)
    e4 = torch.cat(O.bottleneck_mix_ref(zs, 4 * d, tfa // 4, al), 1)
    assert np.array_equal(e4.numpy(), g["e4d1_mix"])
    # the commented-out DPCA blocks (:302-332, :636-679) == the DPCA oracle pinned by pkl_file_model_updated.py
    o2, idx2 = O.dpca_project_ref([z1, z2], (2 * d) // 2)
    assert [len(i) for i in idx2] == list(g["dpca2_after"])
    assert np.abs(torch.cat(o2, 1).numpy() - g["dpca2_out"]).max() <= 1e-5
    o4, idx4 = O.dpca_project_ref(zs, int(g["dpca4_dim_p"]))
    assert [len(i) for i in idx4] == list(g["dpca4_after"])
    assert np.abs(torch.cat(o4, 1).numpy() - g["dpca4_out"]).max() <= 1e-5


def test_task_aware_tail_matches_reference(golden_dir):
    # train_pkl_file_task_aware.py:312-364 executed as written (make_golden.gen_task_aware_tail)
    g, enh, gt = _tail_golden(golden_dir)
    mse, dl = O.task_aware_tail_ref(torch.from_numpy(g["enhanced"]), torch.from_numpy(g["clean"]), enh, gt)
    assert abs(float(mse) - float(g["mse"])) <= 1e-6 * float(g["mse"])
    assert abs(float(dl) - float(g["disc_loss"])) <= 1e-6 * float(g["disc_loss"])
