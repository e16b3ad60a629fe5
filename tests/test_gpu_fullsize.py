"""BASELINE.json's configurations at FULL size, through properties that do not need a full-size CPU oracle
(linearity, Parseval, shard/whole-batch equality, round trips) plus element-wise spot checks of a few rows against
the oracle.  Tolerances as everywhere: fp32 rtol 1e-4, atol = 1e-4 * max|ref|."""
import pytest
import torch

from oracle import spectral_oracle as O

pytestmark = pytest.mark.gpu

N5 = (2048, 1024, 512, 256, 128)          # cfg 4: msstftd front end with 5 window sizes, hop = n / 4
N3 = (1024, 2048, 512)


@pytest.fixture(scope="module")
def ts():
    import tac_spectral
    return tac_spectral


def _tile_rows(x, B):
    """B rows built from the seeded rows of x by per-row gains (keeps the host generation cheap)."""
    reps = (B + x.shape[0] - 1) // x.shape[0]
    gains = 1.0 - 0.5 * torch.arange(B, dtype=torch.float32) / B
    return (x.repeat(reps, 1)[:B] * gains[:, None]).contiguous()


def test_cfg4_disc_front_end_full_size(ts):
    """cfg 4: B = 512 clips of 10 s, 5 scales in ONE launch (6.57 GB of planes)."""
    (x64,), (n64,) = O.synth_audio(16, 160000, 1, seed=9)
    B = 512
    x = _tile_rows(x64, B).cuda()
    v = _tile_rows(n64, B).cuda()
    hops = [n // 4 for n in N5]
    ms = ts.MultiScaleSTFT(list(N5), hops, list(N5))
    ox = ms(x[:, None, :])
    # (1) spot rows against the oracle, element-wise
    rows = [0, 255, 511]
    ref = O.multiscale_disc_inputs_ref(x[rows].cpu()[:, None, :], N5, hops)
    for o, r in zip(ox, ref):
        assert o.shape[0] == B and tuple(o.shape[1:]) == tuple(r.shape[1:])
        atol = 1e-4 * float(r.abs().max())
        assert float((o[rows].cpu() - r).abs().max()) <= atol + 1e-4 * float(r.abs().max())
    # (2) linearity over the whole batch: STFT(0.5 x - 2 v) == 0.5 STFT(x) - 2 STFT(v)
    ov = ms(v[:, None, :])
    oc = ms((0.5 * x - 2.0 * v)[:, None, :])
    for a, b, c in zip(ox, ov, oc):
        lin = 0.5 * a - 2.0 * b
        assert float((c - lin).abs().max()) <= 2e-4 * float(lin.abs().max())
        del lin
    # (3) Parseval per row and scale: frame energies against a float64 evaluation of every 64th row
    sub = list(range(0, B, 64))
    for n, h, o in zip(N5, hops, ox):
        r = O.disc_inputs_ref(x[sub].cpu().double()[:, None, :], n, h)
        e_ref = (r ** 2).sum(dim=(1, 2, 3))
        e_got = (o[sub].double() ** 2).sum(dim=(1, 2, 3)).cpu()
        assert torch.allclose(e_got, e_ref, rtol=1e-4), n


def test_cfg5_long_clip_loss_and_grad(ts):
    """cfg 5 clips: 60 s (960 000 samples).  Loss value against the oracle on the whole (small) batch; the gradient
    through linearity in the upstream scale and a float64 spot check of one row."""
    views, decs = O.synth_audio(4, 960000, 1, seed=21)
    x, y = views[0], decs[0]
    ref = O.mrstft_loss_ref(x, y, N3)
    yg = y.cuda().requires_grad_(True)
    got = ts.mrstft_loss(x.cuda(), yg, N3)
    assert abs(float(got) - float(ref)) <= 1e-4 * abs(float(ref))
    got.backward()
    g1 = yg.grad.clone()
    yg.grad = None
    (3.0 * ts.mrstft_loss(x.cuda(), yg, N3)).backward()
    assert float((yg.grad - 3.0 * g1).abs().max()) <= 1e-5 * float(g1.abs().max())
    # row 0 alone has the same per-row gradient up to the batch normaliser (the loss is a mean over rows)
    y0 = y[:1].double().requires_grad_(True)
    O.mrstft_loss_ref(x[:1].double(), y0, N3, eps=1e-5).backward()
    r = y0.grad[0] / x.shape[0]
    err = (g1[0].double().cpu() - r).abs()
    scale = float(r.abs().max())
    bad = err > 1e-4 * scale + 1e-4 * r.abs()
    assert float(bad.float().mean()) < 0.03                          # isolated sign(X - Y) flips, see test_gpu_loss._check_grad
    keep = ~O.kink_sample_mask(x[:1], y[:1], N3, 1e-5)[0]             # strict on every sample no kink / tiny bin can reach
    assert float(keep.float().mean()) > 0.3 and not bool((bad & keep).any()), int((bad & keep).sum())
    assert float(err.median()) < 2e-6 * scale


def test_cfg3_sharded_step_equals_whole_batch(ts):
    """cfg 3: 4 views, batch 1024 split into 8 shards of 128 -- summed per-shard statistics (what the NCCL all-reduce
    produces) give the whole-batch DPCA; the loss sums of the shards add up to the whole-batch sums."""
    B, d, nv, dim_p, shards = 1024, 128, 4, 128, 8
    zs = [z.cuda() for z in O.synth_latents(B, d, nv)]
    # float64 Gram: the statistics are sum-reducible to rounding; the default 3xTF32 Gram (locally centred shards) to ~1e-6
    whole = ts.dpca_stats(zs, "fp64")
    parts = sum(ts.dpca_stats([z[i * 128:(i + 1) * 128] for z in zs], "fp64") for i in range(shards))
    assert torch.allclose(parts, whole, rtol=1e-10, atol=1e-11 * float(whole.abs().max()))
    tc = sum(ts.dpca_stats([z[i * 128:(i + 1) * 128] for z in zs]) for i in range(shards))
    assert torch.allclose(tc, whole, rtol=1e-5, atol=2e-6 * float(whole.abs().max()))
    s_w, v_w, mu_w, rank_w = ts.dpca_eig(whole, d)
    s_p, v_p, mu_p, rank_p = ts.dpca_eig(parts, d)
    assert torch.allclose(s_w, s_p, rtol=1e-5) and torch.equal(rank_w, rank_p)
    ref_out, ref_idx = O.dpca_project_ref([z.cpu() for z in zs], dim_p)
    got, info = ts.dpca_project(zs, dim_p, return_info=True)
    assert info["cnt"].cpu().tolist() == [len(i) for i in ref_idx]
    for g, r in zip(got, ref_out):
        assert float((g.cpu() - r).abs().max()) <= 1e-4 * float(r.abs().max())
    views, decs = O.synth_audio(64, 16000, 4, seed=3)
    x, y = torch.cat(views, 0).cuda(), torch.cat(decs, 0).cuda()                      # 256 clip-views
    hops = [n // 4 for n in N3]
    whole_sums = ts.mrstft_sums(x, y, list(N3), hops)
    part_sums = sum(ts.mrstft_sums(x[i::4].contiguous(), y[i::4].contiguous(), list(N3), hops) for i in range(4))
    assert torch.allclose(part_sums, whole_sums, rtol=1e-6)


def test_cfg2_front_end_and_inverse_full_size(ts):
    """cfg 2 front end at full size with the padded [512, 2, 1025, 600] model input (2.5 GB), spectral losses on it,
    and the inverse transform back to the waveform."""
    views, decs = O.synth_audio(256, 16000, 2)
    x, y = torch.cat(views, 0).cuda(), torch.cat(decs, 0).cuda()
    obs, dec = ts.complex_spectrogram(x, f_pad=600), ts.complex_spectrogram(y, f_pad=600)
    assert tuple(obs.shape) == (512, 2, 1025, 600)
    assert float(obs[:, :, :, 32:].abs().max()) == 0.0                               # pad_tensor's zeros
    rows = [0, 300, 511]
    mag, ph = O.frontend_ref(x[rows].cpu(), f_pad=600)
    assert float((obs[rows, 0].cpu() - mag).abs().max()) <= 1e-4 * float(mag.max())
    d = ts.spectral_losses(obs, dec)
    r = O.spectral_losses_ref(obs[:32].cpu(), dec[:32].cpu())
    d32 = ts.spectral_losses(obs[:32].contiguous(), dec[:32].contiguous())
    for k in ("magnitude_loss", "phase_loss", "total_loss"):
        assert abs(float(d32[k]) - float(r[k])) <= 1e-4 * abs(float(r[k])), k
        assert torch.isfinite(d[k])
    wav = ts.batch_reconstruct_waveform(obs[:, 0], obs[:, 1], 16000)
    assert tuple(wav.shape) == (512, 512 * 599)
    assert float((wav[:, :15000] - x[:, :15000]).abs().max()) <= 2e-5 * float(x.abs().max())


def test_cfg2_step_schedules_agree(ts, monkeypatch):
    """cfg 2 through SpectralStep: the three-stream schedule (DPCA chain | loss chain | front end filling the loss kernel's
    last wave), the two-stream schedule and the serial one give bit-identical results (every kernel is deterministic and
    the schedules only reorder independent work), and the loss equals the oracle's on the same inputs."""
    from tac_spectral.step import SpectralStep
    batch, nv, T, d, dim_p = 256, 2, 16000, 128, 32
    views, decs = O.synth_audio(32, T, nv, seed=11)
    x = _tile_rows(torch.cat(views, 0), batch * nv).cuda()
    y = _tile_rows(torch.cat(decs, 0), batch * nv).cuda()
    zs = [z.cuda() for z in O.synth_latents(batch, d, nv)]
    hops = [n // 4 for n in N3]
    outs = []
    for overlap, fill in ((False, "0"), (True, "0"), (True, "1")):
        monkeypatch.setenv("TAC_STEP_FILL", fill)
        st = SpectralStep(batch * nv, T, nv, batch, d, dim_p, list(N3), hops, device="cuda", overlap=overlap)
        for _ in range(2):
            loss = st.run(x, y, zs)
        torch.cuda.synchronize()
        outs.append((loss.clone(), st.grad_y.clone(), st.fe_out.clone(), [z.clone() for z in st.z_out]))
    for o in outs[1:]:
        assert torch.equal(o[0], outs[0][0]) and torch.equal(o[1], outs[0][1]) and torch.equal(o[2], outs[0][2])
        assert all(torch.equal(a, b) for a, b in zip(o[3], outs[0][3]))
    ref = O.mrstft_loss_ref(x[:16].cpu(), y[:16].cpu(), list(N3))                      # spot check on 16 rows
    got = ts.mrstft_loss(x[:16], y[:16], list(N3))
    assert abs(float(got) - float(ref)) <= 1e-4 * abs(float(ref))
