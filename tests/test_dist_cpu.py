"""N > 1 host logic on CPU: world-size-2 ``gloo`` runs of the step's exchange (plan.py).

What is checked is what makes the sharded step equal to the reference's whole-batch computation
(SURVEY.md 8e): contiguous batch shards, ONE packed float64 all-reduce of [loss sums | DPCA statistics],
global normalisers known from shapes.  The per-shard sums / statistics are produced here by the CPU oracle
(on the GPU they come from tac_mrstft_loss_fwd_bwd / tac_dpca_stats; tests/test_gpu_dpca.py and
tests/test_gpu_loss.py check those against the same oracle).
"""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "task-aware-audio-coding-perceptual_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

N_FFTS, HOPS = (256, 512, 128), (64, 128, 32)
BATCH, VIEWS, T, D, DIM_P = 6, 2, 2048, 16, 8


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _inputs():
    from oracle import spectral_oracle as O
    views, decs = O.synth_audio(BATCH, T, VIEWS, seed=7)
    x = torch.stack(views, 1).reshape(BATCH * VIEWS, T)          # item-major, view-minor rows
    y = torch.stack(decs, 1).reshape(BATCH * VIEWS, T)
    zs = O.synth_latents(BATCH * 8, D, VIEWS, seed=11)            # 48 latent rows per view
    return x, y, zs


def _raw_sums(x, y):
    """Per-scale raw sums [2S] float64 on the msstftd.py:125-127 convention (what the CUDA kernel emits)."""
    from oracle import spectral_oracle as O
    out = []
    for n, h in zip(N_FFTS, HOPS):
        X = O.stft_ref(x, n, h).abs().double()
        Y = O.stft_ref(y, n, h).abs().double()
        out += [(X - Y).abs().sum(), (X.clamp_min(O.LOG_FLOOR).log() - Y.clamp_min(O.LOG_FLOOR).log()).abs().sum()]
    return torch.stack(out)


def _raw_stats(z):
    z = z.double()
    return torch.cat([torch.tensor([float(z.shape[0])], dtype=torch.float64), z.sum(0), (z.T @ z).reshape(-1)])


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from tac_spectral.plan import StepPlan, shard_bounds
        x, y, zs = _inputs()
        lo, hi = shard_bounds(BATCH, world, rank)
        rows = slice(lo * VIEWS, hi * VIEWS)
        zlo, zhi = shard_bounds(zs[0].shape[0], world, rank)
        plan = StepPlan((hi - lo) * VIEWS, T, N_FFTS, HOPS, VIEWS, D, world)
        packed = torch.zeros(plan.packed_len, dtype=torch.float64)
        sums, stats = plan.views(packed)
        sums.copy_(_raw_sums(x[rows], y[rows]))
        for v in range(VIEWS):
            stats[v].copy_(_raw_stats(zs[v][zlo:zhi]))
        local_sums = sums.clone()
        plan.allreduce(packed, None, "stats")                     # two slices, as SpectralStep does
        assert torch.equal(sums, local_sums), "the 'stats' slice must not touch the loss sums"
        plan.allreduce(packed, None, "sums")
        q.put((rank, packed.clone(), float(plan.loss_from_sums(sums)), (lo, hi)))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(180)
def test_two_rank_exchange_equals_full_batch():
    from oracle import spectral_oracle as O
    from tac_spectral.plan import StepPlan
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=150) for _ in range(world)], key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    # shards are contiguous, disjoint and cover the batch
    assert [r[3] for r in res] == [(0, 3), (3, 6)]
    # every rank holds the same reduced buffer
    assert torch.equal(res[0][1], res[1][1])
    x, y, zs = _inputs()
    full = StepPlan(BATCH * VIEWS, T, N_FFTS, HOPS, VIEWS, D, 1)
    sums, stats = full.views(res[0][1].clone())
    # loss sums: sharded == whole batch (float64 sums of the same terms, order differs)
    ref_sums = _raw_sums(x, y)
    assert torch.allclose(sums, ref_sums, rtol=1e-12, atol=0)
    # loss value with the global normaliser == the oracle's whole-batch loss (fp32 rtol 1e-4 is the spec; this is tighter)
    ref_loss = float(O.mrstft_loss_ref(x, y, N_FFTS, HOPS))
    assert abs(res[0][2] - ref_loss) <= 1e-5 * abs(ref_loss)
    assert abs(res[0][2] - res[1][2]) == 0.0
    # DPCA statistics: n, sum z, sum z z^T of the whole batch; data_pca from them matches torch.svd (a6)
    for v in range(VIEWS):
        assert torch.allclose(stats[v], _raw_stats(zs[v]), rtol=1e-12, atol=1e-12)
        n, s1, S2 = stats[v][0], stats[v][1:1 + D], stats[v][1 + D:].view(D, D)
        C = S2 - torch.outer(s1, s1) / n
        lam, V = torch.linalg.eigh(C)
        s_ref, v_ref, mu_ref = O.data_pca_ref(zs[v])
        assert torch.allclose(lam.flip(0).clamp_min(0).sqrt().float(), s_ref, rtol=1e-4, atol=1e-5)
        assert torch.allclose((s1 / n).float(), mu_ref, rtol=1e-5, atol=1e-6)
        assert O.subspace_angle(V.flip(1)[:, :DIM_P].float(), v_ref[:, :DIM_P]) < 1e-3


def test_shard_bounds_and_plan_layout():
    from tac_spectral.plan import StepPlan, shard_bounds
    for batch, world in ((1024, 8), (10, 4), (3, 2), (7, 7)):
        b = [shard_bounds(batch, world, r) for r in range(world)]
        assert b[0][0] == 0 and b[-1][1] == batch
        assert all(b[i][1] == b[i + 1][0] for i in range(world - 1))
        assert max(h - l for l, h in b) - min(h - l for l, h in b) <= 1
    with pytest.raises(ValueError):
        shard_bounds(4, 2, 2)
    p = StepPlan(8, 16000, (1024, 2048, 512), (256, 512, 128), 4, 128, 8)
    assert p.packed_len == 6 + 4 * (1 + 128 + 128 * 128)          # 528 KB float64 at d = 128 x 4 views
    assert p.counts == [8 * 513 * 59, 8 * 1025 * 28, 8 * 257 * 122]
    assert abs(sum(1.0 / i for i in p.inv_counts()[::2]) - 3 * 8 * sum(p.counts)) < 1e-3
    with pytest.raises(ValueError):
        StepPlan(8, 1000, (2048,), (512,), 1, 8)
    with pytest.raises(ValueError):
        p.views(torch.zeros(3, dtype=torch.float64))
