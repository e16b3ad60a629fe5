"""SpectralStep -- the object bench.py times -- against the ORACLE, output by output.

The legs' entry points are each checked against the oracle elsewhere; this file checks the step's own buffers after
``run``: ``loss`` (O.mrstft_loss_ref), ``grad_y`` (float64 autograd of the same, strict on every sample no kink can touch),
``fe_out`` magnitude AND phase (O.frontend_ref: spectrogram.ipynb cell 10), ``z_out`` / ``cnt`` / ``idx``
(O.dpca_project_ref: pkl_file_model_updated.py:489-511), eager and through CUDA-graph replay.
Tolerances (north star): fp32 rtol 1e-4 with atol = 1e-4 * max|ref|; phase within 1e-3 rad where |X| > 1e-3 max; DPCA
outputs within 1e-4 of their largest entry, per-view counts exactly equal.
"""
import pytest
import torch

from oracle import spectral_oracle as O

pytestmark = pytest.mark.gpu

N3 = (1024, 2048, 512)


@pytest.fixture(scope="module")
def ts():
    import tac_spectral
    return tac_spectral


def _check_step_outputs(st, x, y, zs, dim_p, what):
    nv = len(zs)
    rows, T = x.shape
    # ---- loss value
    ref_loss = O.mrstft_loss_ref(x.double(), y.double(), list(N3), eps=st.eps)
    assert abs(float(st.loss) - float(ref_loss)) <= 1e-4 * abs(float(ref_loss)), (what, float(st.loss), float(ref_loss))
    # ---- gradient w.r.t. the decoded signal: strict off the kinks, statistical on them
    y64 = y.double().requires_grad_(True)
    O.mrstft_loss_ref(x.double(), y64, list(N3), eps=st.eps).backward()
    ref_g = y64.grad
    got_g = st.grad_y.double().cpu()
    keep = ~O.kink_sample_mask(x, y, N3, st.eps)
    scale = float(ref_g.abs().max())
    err = (got_g - ref_g).abs()
    bad = err > 1e-4 * scale + 1e-4 * ref_g.abs()
    assert float(keep.float().mean()) > 0.3, float(keep.float().mean())
    assert not bool((bad & keep).any()), f"{what}: {int((bad & keep).sum())} kink-free gradient samples off"
    assert float(bad.float().mean()) < 0.03 and float(err.norm() / ref_g.norm()) < 2e-3, what
    # ---- front end: magnitude and phase planes
    mag, ph = O.frontend_ref(x)
    fe = st.fe_out.cpu()
    assert fe.shape == (rows, 2, 1025, mag.shape[-1])
    assert bool(((fe[:, 0] - mag).abs() <= 1e-4 * mag.abs() + 1e-4 * float(mag.max())).all()), what
    big = mag > 1e-3 * mag.max()
    dphi = torch.angle(torch.exp(1j * (fe[:, 1] - ph)))
    assert float(dphi[big].abs().max()) < 1e-3, (what, float(dphi[big].abs().max()))
    # ---- DPCA: per-view counts, selected indices, projected latents
    ref_out, ref_idx = O.dpca_project_ref(zs, dim_p)
    cnt = st.cnt.cpu().tolist()
    assert cnt == [len(i) for i in ref_idx], (what, cnt)
    for v in range(nv):
        assert sorted(st.idx[v, :cnt[v]].cpu().tolist()) == sorted(ref_idx[v].tolist()), (what, v)
        zo = st.z_out[v].cpu()
        assert float((zo - ref_out[v]).abs().max()) <= 1e-4 * float(ref_out[v].abs().max()), (what, v)


@pytest.mark.parametrize("overlap,eps", [(True, O.LOG_FLOOR), (False, O.LOG_FLOOR), (True, 1e-2)])
def test_step_outputs_match_oracle_eager_and_graph(ts, overlap, eps):
    """cfg 2's shape at a batch the oracle finishes in seconds: 2 views x 144 clips x 1 s (B > d: full-rank PCA), d = 128, dim_p = 32."""
    from tac_spectral.step import SpectralStep
    batch, nv, T, d, dim_p = 144, 2, 16000, 128, 32
    views, decs = O.synth_audio(batch, T, nv, seed=5)
    x, y = torch.cat(views, 0), torch.cat(decs, 0)                  # [batch * nv, T], view-major like bench.py
    zs = O.synth_latents(batch, d, nv)
    xg, yg, zg = x.cuda(), y.cuda(), [z.cuda() for z in zs]
    st = SpectralStep(batch * nv, T, nv, batch, d, dim_p, list(N3), [n // 4 for n in N3], eps=eps, device="cuda", overlap=overlap)
    st.run(xg, yg, zg)
    torch.cuda.synchronize()
    _check_step_outputs(st, x, y, zs, dim_p, f"eager overlap={overlap}")
    eager = (st.loss.clone(), st.grad_y.clone(), st.fe_out.clone(), [z.clone() for z in st.z_out], st.cnt.clone(), st.idx.clone())
    # ---- the same step captured and replayed: outputs poisoned first, so the replay has to produce them
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        st.run(xg, yg, zg)
    st.loss.fill_(float("nan")); st.grad_y.fill_(float("nan")); st.fe_out.fill_(float("nan"))
    for z in st.z_out:
        z.fill_(float("nan"))
    st.cnt.zero_()
    g.replay()
    torch.cuda.synchronize()
    assert torch.equal(st.loss, eager[0]) and torch.equal(st.grad_y, eager[1]) and torch.equal(st.fe_out, eager[2])
    assert all(torch.equal(a, b) for a, b in zip(st.z_out, eager[3])) and torch.equal(st.cnt, eager[4])
    _check_step_outputs(st, x, y, zs, dim_p, f"graph replay overlap={overlap}")
    st.check_exchange()                                             # no exchange at world 1: must be a no-op
    st.close()
