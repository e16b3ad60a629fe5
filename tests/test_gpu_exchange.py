"""The step's one-shot peer exchange (include/tac_spectral.h "multi-GPU exchange", csrc/xchg.cuh).

The protocol (slots, flags, parities, sequence counters) only needs device pointers, so a WORLD OF TWO IS SIMULATED ON
ONE GPU: two buffers, two descriptors (rank 0 / rank 1) whose ``peer`` tables point at both.  Pushes of both "ranks" are
enqueued before the pops, so nothing spins.  The real thing -- two processes, two GPUs, NVLink, CUDA-graph replay -- is
``test_two_gpu_step_matches_whole_batch`` (skipped below two devices; run with ``gpurun --gpus 2``), which also checks
the whole-batch semantics of the reference (pkl_file_model_updated.py:489-511, train_pkl_file_task_aware.py:597-629).
"""
import ctypes as C
import os
import subprocess
import sys

import pytest
import torch

from oracle import spectral_oracle as O

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def ts():
    import tac_spectral
    return tac_spectral


def _fake_world(L, payload, world=2):
    lib = L.lib()
    n = int(lib.tac_xchg_bytes(world, payload))
    assert n > 0
    bufs = [torch.zeros(n, dtype=torch.uint8, device="cuda") for _ in range(world)]
    descs = []
    for r in range(world):
        d = L.XchgStruct()
        d.world, d.rank, d.payload = world, r, payload
        for p in range(world):
            d.peer[p] = bufs[p].data_ptr()
        descs.append(d)
    return bufs, descs


def test_push_pop_rank_ordered_sum(ts):
    L = ts._lib
    lib = L.lib()
    payload, world = 1000, 3
    bufs, descs = _fake_world(L, payload, world)
    st = torch.cuda.current_stream().cuda_stream
    for it in range(5):                                              # both parities, sequence counters advancing
        srcs = [torch.randn(payload, dtype=torch.float64, device="cuda") * (it + 1) for _ in range(world)]
        for r in range(world):
            L.check(lib.tac_xchg_push(C.byref(descs[r]), srcs[r].data_ptr(), st))
        outs = [torch.empty(payload, dtype=torch.float64, device="cuda") for _ in range(world)]
        for r in range(world):
            L.check(lib.tac_xchg_pop(C.byref(descs[r]), outs[r].data_ptr(), st))
        torch.cuda.synchronize()
        ref = (srcs[0] + srcs[1]) + srcs[2]                          # rank order
        for r in range(world):
            assert torch.equal(outs[r], ref), f"iteration {it}, rank {r}: not the rank-ordered sum"


def test_pop_gives_up_after_the_timeout_and_reports_it(ts):
    """A peer that never pushes must neither hang the GPU nor trap the context: the pop stops waiting after
    tac_xchg_set_timeout seconds, the host finds the failed pop in tac_xchg_status, and the context stays usable."""
    L = ts._lib
    lib = L.lib()
    bufs, descs = _fake_world(L, 8, 2)
    st = torch.cuda.current_stream().cuda_stream
    w = C.c_uint(7)
    L.check(lib.tac_xchg_status(C.byref(descs[0]), C.byref(w)))
    assert w.value == 0
    src = torch.ones(8, dtype=torch.float64, device="cuda")
    dst = torch.zeros(8, dtype=torch.float64, device="cuda")
    L.check(lib.tac_xchg_set_timeout(0.05))
    try:
        L.check(lib.tac_xchg_push(C.byref(descs[0]), src.data_ptr(), st))       # "rank 1" stays silent
        t0 = torch.cuda.Event(enable_timing=True)
        t1 = torch.cuda.Event(enable_timing=True)
        t0.record()
        L.check(lib.tac_xchg_pop(C.byref(descs[0]), dst.data_ptr(), st))
        t1.record()
        torch.cuda.synchronize()                                                # would raise after a trap
        assert 40.0 <= t0.elapsed_time(t1) <= 2000.0                             # waited ~50 ms, not ~0 and not for ever
    finally:
        L.check(lib.tac_xchg_set_timeout(-1.0))                                 # back to TAC_XCHG_TIMEOUT_S / 600 s
    L.check(lib.tac_xchg_status(C.byref(descs[0]), C.byref(w)))
    assert w.value == 1                                                         # pop number 0 gave up
    L.check(lib.tac_xchg_status(C.byref(descs[1]), C.byref(w)))
    assert w.value == 0                                                         # the other rank's buffer is untouched
    assert float((src * 2).sum()) == 16.0                                       # context alive
    with pytest.raises(RuntimeError):
        L.check(lib.tac_xchg_set_timeout(float("nan")))


def test_loss_finalize_single_rank(ts):
    L = ts._lib
    sums = torch.rand(6, dtype=torch.float64, device="cuda")
    inv = torch.rand(6, dtype=torch.float64, device="cuda")
    loss = torch.zeros((), dtype=torch.float64, device="cuda")
    L.check(L.lib().tac_loss_finalize(sums.data_ptr(), 6, inv.data_ptr(), None, loss.data_ptr(), torch.cuda.current_stream().cuda_stream))
    ref = 0.0
    for a, b in zip(sums.tolist(), inv.tolist()):
        ref += a * b
    assert float(loss) == ref


def test_loss_finalize_exchanges_sums(ts):
    L = ts._lib
    lib = L.lib()
    bufs, descs = _fake_world(L, 6, 2)
    st = torch.cuda.current_stream().cuda_stream
    inv = torch.rand(6, dtype=torch.float64, device="cuda")
    # a single-CTA kernel pushes AND pops: on one GPU the two "ranks" must run concurrently -> two streams
    s1 = torch.cuda.Stream()
    for it in range(3):
        a = torch.rand(6, dtype=torch.float64, device="cuda")
        b = torch.rand(6, dtype=torch.float64, device="cuda")
        want = a + b
        la = torch.zeros((), dtype=torch.float64, device="cuda")
        lb = torch.zeros((), dtype=torch.float64, device="cuda")
        torch.cuda.synchronize()
        L.check(lib.tac_loss_finalize(a.data_ptr(), 6, inv.data_ptr(), C.byref(descs[0]), la.data_ptr(), st))
        with torch.cuda.stream(s1):
            L.check(lib.tac_loss_finalize(b.data_ptr(), 6, inv.data_ptr(), C.byref(descs[1]), lb.data_ptr(), s1.cuda_stream))
        torch.cuda.synchronize()
        assert torch.equal(a, want) and torch.equal(b, want)         # global sums written back on both ranks
        assert float(la) == float(lb)
        assert abs(float(la) - float((want * inv).sum())) <= 1e-12 * abs(float(la))


@pytest.mark.parametrize("algo", ["tf32x3", "fp64"])
def test_fused_stats_push_and_eig_pop(ts, algo):
    """tac_dpca_stats_push (shards of two "ranks") -> tac_dpca_eig_pop == DPCA of the whole batch."""
    L = ts._lib
    lib = L.lib()
    nv, B, d, dim_p = 2, 256, 128, 32
    zs = [z.cuda() for z in O.synth_latents(B, d, nv)]
    shards = [[z[:96].contiguous() for z in zs], [z[96:].contiguous() for z in zs]]            # uneven shards
    stride = 1 + d + d * d
    bufs, descs = _fake_world(L, nv * stride, 2)
    st = torch.cuda.current_stream().cuda_stream
    code = {"tf32x3": L.GRAM_TF32X3, "fp64": L.GRAM_FP64}[algo]
    for it in range(3):
        stats = [torch.zeros((nv, stride), dtype=torch.float64, device="cuda") for _ in range(2)]
        for r in range(2):
            Br = shards[r][0].shape[0]
            nb = int(lib.tac_dpca_workspace_bytes(nv, Br, d))
            ws = torch.empty(max(nb, 256), dtype=torch.uint8, device="cuda")
            L.check(lib.tac_dpca_stats_push(L.ptr_array(shards[r]), nv, Br, d, code, stats[r].data_ptr(), ws.data_ptr(), nb,
                                            C.byref(descs[r]), st))
        res = []
        for r in range(2):
            s = torch.empty((nv, d), dtype=torch.float32, device="cuda")
            v = torch.empty((nv, d, d), dtype=torch.float32, device="cuda")
            mu = torch.empty((nv, d), dtype=torch.float32, device="cuda")
            rk = torch.empty(nv, dtype=torch.int32, device="cuda")
            nb = int(lib.tac_dpca_workspace_bytes(nv, B, d))
            ws = torch.empty(max(nb, 256), dtype=torch.uint8, device="cuda")
            L.check(lib.tac_dpca_eig_pop(stats[r].data_ptr(), nv, d, s.data_ptr(), v.data_ptr(), mu.data_ptr(), rk.data_ptr(),
                                         ws.data_ptr(), nb, C.byref(descs[r]), st))
            res.append((s, v, mu, rk))
        torch.cuda.synchronize()
        assert torch.equal(stats[0], stats[1])                       # both ranks hold the bit-identical global statistics
        assert float(stats[0][0, 0]) == B
        for a, b in zip(res[0], res[1]):
            assert torch.equal(a, b)
        s_ref = torch.stack([O.data_pca_ref(z.cpu())[0] for z in zs])
        assert torch.allclose(res[0][0].cpu(), s_ref, rtol=1e-4, atol=1e-4 * float(s_ref.max()))
        for vw in range(nv):
            v_ref = O.data_pca_ref(zs[vw].cpu())[1]
            assert O.subspace_angle(res[0][1][vw, :, :8].cpu(), v_ref[:, :8]) < 1e-3


def test_two_devices_one_process(ts):
    """cudaFuncSetAttribute is per device: every kernel family must also run on a second GPU of the same process."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    views, decs = O.synth_audio(2, 16000, 1)
    x, y = views[0], decs[0]
    zs = O.synth_latents(64, 128, 2)
    ref = None
    for dev in ("cuda:0", "cuda:1"):
        out = ts.MultiScaleSTFT()(x[:, None].to(dev))[1]
        fe = ts.complex_spectrogram(x.to(dev), f_pad=600)
        loss = ts.mrstft_loss(x.to(dev), y.to(dev).requires_grad_(True))
        zo = ts.dpca_project([z.to(dev) for z in zs], 16)
        wav = ts.batch_reconstruct_waveform(fe[:, 0, :, :32], fe[:, 1, :, :32], 16000)
        got = [out.cpu(), fe.cpu(), loss.detach().cpu(), zo[0].cpu(), wav.cpu()]
        if ref is None:
            ref = got
        else:
            for a, b in zip(ref, got):
                assert torch.allclose(a, b, rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("exchange", ["p2p", "nccl"])
def test_two_gpu_step_matches_whole_batch(exchange):
    """torchrun x 2: SpectralStep(group), eager and CUDA-graph replay, against the one-GPU whole-batch step."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (gpurun --gpus 2)")
    env = dict(os.environ, TAC_STEP_EXCHANGE=exchange)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29533", os.path.join(ROOT, "tests", "dist_gpu_worker.py")]
    r = subprocess.run(cmd, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert r.returncode == 0 and "PARITY_N_OK" in r.stdout, r.stdout[-4000:]
