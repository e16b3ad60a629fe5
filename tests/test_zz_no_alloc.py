"""tac_spec_losses_static: the spectral reductions (K2') for hosts that cannot allocate device memory.

Some users of libtac_spectral hold device pointers to their spectrograms but have no allocator at hand when they need
the losses: a plugin embedded in another framework's memory pool, a real-time audio / callback thread in which
cudaMalloc (an implicit device synchronisation, driver locks) is not allowed, or a job whose device-memory budget was
carved up at start-up.  For them the library offers one entry point whose scratch is a fixed 1 MiB array inside the
library image (B <= TAC_SPEC_LOSSES_STATIC_MAX_B = 4096), shared by all calls on a device.

The test drives both entry points through the C ABI (ctypes) on the same inputs, several calls in a row with different
data and a different B -- each later call finds the earlier call's sums still sitting in the shared array and must not
be affected by them -- and asks for identical results (same kernels, same summation order: compared exactly, which is
tighter than the 1e-5 the other spec-loss tests allow against the oracle).  It also checks the size limit's error.

The file sorts last on purpose (pytest -x stops at the first failure) and has to pass both alone in a fresh process,
where its first call performs the per-device address look-up, and after all the other tests.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_B = 4096     # TAC_SPEC_LOSSES_STATIC_MAX_B (include/tac_spectral.h)


@pytest.fixture(scope="module")
def L():
    import tac_spectral
    return tac_spectral._lib


def _both(L, obs, dec):
    """(out5, per_sample) of tac_spec_losses and of tac_spec_losses_static on the same device buffers."""
    lib = L.lib()
    B, _, K, F = obs.shape
    ws_bytes = int(lib.tac_spec_losses_workspace_bytes(B))
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=obs.device)
    res = []
    for static in (False, True):
        out5 = torch.full((5,), float("nan"), device=obs.device)
        per = torch.full((B,), float("nan"), device=obs.device)
        with L.device_of(obs):
            if static:
                L.check(lib.tac_spec_losses_static(obs.data_ptr(), dec.data_ptr(), B, K, F, out5.data_ptr(),
                                                   per.data_ptr(), L.stream_of(obs)))
            else:
                L.check(lib.tac_spec_losses(obs.data_ptr(), dec.data_ptr(), B, K, F, out5.data_ptr(), per.data_ptr(),
                                            ws.data_ptr(), ws_bytes, L.stream_of(obs)))
        torch.cuda.synchronize()
        res.append((out5.cpu(), per.cpu()))
    return res


def test_static_scratch_matches_workspace_entry_point_twice(L):
    g = torch.Generator().manual_seed(77)
    # calls in a row on the shared array; the second is SMALLER than the first, so stale sums of call 1 lie beyond
    # (and, had the kernels accumulated instead of overwriting, under) what call 2 uses; the third fills the array to
    # its last slot (B = the documented maximum)
    for B, K, F, noise in ((7, 1025, 40, 0.1), (3, 65, 21, 0.5), (MAX_B, 5, 3, 0.2)):
        obs = (torch.rand(B, 2, K, F, generator=g) * 2.0).cuda()
        dec = (obs.cpu() + noise * torch.randn(B, 2, K, F, generator=g)).cuda()
        (ref5, refper), (got5, gotper) = _both(L, obs, dec)
        assert torch.isfinite(ref5).all() and torch.isfinite(refper).all()
        assert torch.equal(got5, ref5), (got5, ref5)
        assert torch.equal(gotper, refper)
        # sanity against plain torch on the host (the oracle proper is exercised in test_gpu_loss.py), rtol 1e-5
        o, d = obs.cpu().double(), dec.cpu().double()
        tot = ((o - d) ** 2).mean()
        assert abs(float(got5[2]) - float(tot)) <= 1e-5 * float(tot)
        assert torch.allclose(gotper.double(), 0.5 * ((o - d) ** 2).mean(dim=(1, 2, 3)), rtol=1e-5)


def test_static_scratch_without_per_sample_and_size_limit(L):
    lib = L.lib()
    obs = torch.rand(2, 2, 9, 5, device="cuda")
    dec = obs * 0.9
    out_a = torch.empty(5, device="cuda")
    out_b = torch.empty(5, device="cuda")
    ws_bytes = int(lib.tac_spec_losses_workspace_bytes(2))
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device="cuda")
    with L.device_of(obs):
        L.check(lib.tac_spec_losses(obs.data_ptr(), dec.data_ptr(), 2, 9, 5, out_a.data_ptr(), None, ws.data_ptr(),
                                    ws_bytes, L.stream_of(obs)))
        L.check(lib.tac_spec_losses_static(obs.data_ptr(), dec.data_ptr(), 2, 9, 5, out_b.data_ptr(), None,
                                           L.stream_of(obs)))
        # one item too many for the built-in array: refused on the host, before anything is launched
        n0 = lib.tac_launch_count()
        rc = lib.tac_spec_losses_static(obs.data_ptr(), dec.data_ptr(), MAX_B + 1, 1, 1, out_b.data_ptr(), None,
                                        L.stream_of(obs))
        assert rc != 0 and lib.tac_launch_count() == n0
        assert b"at most 4096" in lib.tac_last_error()
        with pytest.raises(RuntimeError, match="at most 4096"):
            L.check(rc)
    torch.cuda.synchronize()
    assert torch.equal(out_a.cpu(), out_b.cpu())
