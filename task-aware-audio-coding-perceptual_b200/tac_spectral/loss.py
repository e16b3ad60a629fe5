"""Host-side mirror of the loss reductions on the spectral hot path (K2 / K2').

* :func:`mrstft_loss` -- multi-resolution STFT loss (L1 + log-magnitude per scale) on the STFT
  convention of ``msstftd.py:125-127``.  The reference holds no such function; the definition of
  record is SURVEY.md section 8c item 2 (restated in ``oracle/spectral_oracle.py:mrstft_loss_ref``).
  Forward and backward are one fused kernel each (``tac_mrstft_loss_fwd`` / ``_bwd``): no
  spectrogram is ever written to HBM and no host synchronisation happens.
* :func:`spectral_losses` -- the reference's own reductions over the stacked magnitude / phase
  tensors (``pkl_file_models.py:924,933-940``; ``pkl_file_model_updated.py:520-540``), one pass
  over ``(obs, obs_dec)`` instead of six, with an analytic backward (``tac_spec_losses_bwd``).

Multi-GPU: pass ``group`` (a ``torch.distributed`` process group) and the per-scale sums are
all-reduced, so that every rank returns the loss of the *global* batch exactly as the reference
computes it on one device; the input gradient needs no collective.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence

import torch

from . import _lib as L
from .stft import num_frames

LOG_FLOOR = 1e-5          # librosa `amin`, the only log floor the reference's dB channel uses


def _rows(t: torch.Tensor, name: str) -> torch.Tensor:
    L.require_cuda(t, name)
    if t.dim() not in (2, 3):
        raise ValueError(f"{name} must be [B, T] or [B, C, T], got {tuple(t.shape)}")
    return t.reshape(-1, t.shape[-1])


class _MRSTFTLoss(torch.autograd.Function):
    """sums[2S] (float64) of the global batch; differentiable w.r.t. ``y`` only (``x`` is the target)."""

    @staticmethod
    def forward(ctx, x, y, n_ffts, hops, wins, eps):
        xr = x.contiguous().float()
        yr = y.contiguous().float()
        B, T = xr.shape
        S = len(n_ffts)
        lib = L.lib()
        nf, hp, wn = L.int_array(n_ffts), L.int_array(hops), L.int_array(wins)
        ws_bytes = int(lib.tac_mrstft_workspace_bytes(B, T, S, nf, hp))
        ws = torch.empty(max(ws_bytes, 16), dtype=torch.uint8, device=xr.device)
        sums = torch.empty(2 * S, dtype=torch.float64, device=xr.device)
        with L.device_of(xr):
            L.check(lib.tac_mrstft_loss_fwd(xr.data_ptr(), yr.data_ptr(), B, T, S, nf, hp, wn, float(eps),
                                            sums.data_ptr(), ws.data_ptr(), ws_bytes, L.stream_of(xr)))
        ctx.save_for_backward(xr, yr)
        ctx.cfg = (list(n_ffts), list(hops), list(wins), float(eps), y.shape, y.dtype)
        return sums

    @staticmethod
    def backward(ctx, grad_sums):
        xr, yr = ctx.saved_tensors
        n_ffts, hops, wins, eps, yshape, ydtype = ctx.cfg
        B, T = xr.shape
        S = len(n_ffts)
        coef = grad_sums.to(torch.float32).contiguous()
        grad = torch.empty_like(yr)
        lib = L.lib()
        nf, hp = L.int_array(n_ffts), L.int_array(hops)
        ws_bytes = int(lib.tac_mrstft_workspace_bytes(B, T, S, nf, hp))
        ws = torch.empty(max(ws_bytes, 16), dtype=torch.uint8, device=xr.device)
        with L.device_of(xr):
            L.check(lib.tac_mrstft_loss_bwd(xr.data_ptr(), yr.data_ptr(), B, T, S, nf, hp, L.int_array(wins), eps,
                                            coef.data_ptr(), grad.data_ptr(), ws.data_ptr(), ws_bytes, L.stream_of(xr)))
        return None, grad.reshape(yshape).to(ydtype), None, None, None, None


class _MRSTFTLossFused(torch.autograd.Function):
    """The scalar loss with its gradient produced in the SAME pass (``tac_mrstft_loss_fwd_bwd``): since
    d loss / d sums is a table of shape-only constants, the kernel that forms the sums already knows the
    gradient direction; backward is one scaling by the upstream scalar."""

    @staticmethod
    def forward(ctx, x, y, n_ffts, hops, wins, eps, inv, group):
        xr = x.contiguous().float()
        yr = y.contiguous().float()
        B, T = xr.shape
        S = len(n_ffts)
        lib = L.lib()
        nf, hp, wn = L.int_array(n_ffts), L.int_array(hops), L.int_array(wins)
        ws_bytes = int(lib.tac_mrstft_workspace_bytes(B, T, S, nf, hp))
        ws = torch.empty(max(ws_bytes, 16), dtype=torch.uint8, device=xr.device)
        sums = torch.empty(2 * S, dtype=torch.float64, device=xr.device)
        grad = torch.empty_like(yr)
        coef = inv.to(torch.float32)
        with L.device_of(xr):
            L.check(lib.tac_mrstft_loss_fwd_bwd(xr.data_ptr(), yr.data_ptr(), B, T, S, nf, hp, wn, float(eps), coef.data_ptr(),
                                                sums.data_ptr(), grad.data_ptr(), ws.data_ptr(), ws_bytes, L.stream_of(xr)))
        if group is not None:
            import torch.distributed as dist
            dist.all_reduce(sums, group=group)
        ctx.save_for_backward(grad)
        ctx.meta = (y.shape, y.dtype)
        return torch.dot(sums, inv).float()

    @staticmethod
    def backward(ctx, g):
        (grad,) = ctx.saved_tensors
        shape, dtype = ctx.meta
        return None, (grad * g.to(grad.dtype)).reshape(shape).to(dtype), None, None, None, None, None, None


def mrstft_sums(x: torch.Tensor, y: torch.Tensor, n_ffts: Sequence[int], hops: Sequence[int],
                wins: Optional[Sequence[int]] = None, eps: float = LOG_FLOOR) -> torch.Tensor:
    """Raw per-scale sums ``[2S]`` float64: ``sum|X-Y|`` and ``sum|ln max(X,eps) - ln max(Y,eps)|``."""
    xr, yr = _rows(x, "x"), _rows(y, "y")
    if xr.shape != yr.shape:
        raise ValueError(f"Shape mismatch: x {tuple(x.shape)} vs y {tuple(y.shape)}")   # cf. train_pkl_file_task_aware.py:179-180
    wins = list(n_ffts) if wins is None else list(wins)
    assert len(n_ffts) == len(hops) == len(wins)                                        # msstftd.py:177
    return _MRSTFTLoss.apply(xr, yr, list(n_ffts), list(hops), wins, eps)


def mrstft_counts(B: int, T: int, n_ffts: Sequence[int], hops: Sequence[int]) -> List[int]:
    """Elements per scale: ``B * K_s * F_s`` (``center=False`` framing of ``msstftd.py:127``)."""
    return [B * (n // 2 + 1) * num_frames(T, n, h, False) for n, h in zip(n_ffts, hops)]


def mrstft_loss(x: torch.Tensor, y: torch.Tensor, n_ffts: Sequence[int] = (1024, 2048, 512),
                hops: Optional[Sequence[int]] = None, wins: Optional[Sequence[int]] = None,
                eps: float = LOG_FLOOR, return_terms: bool = False, group=None):
    """``L = (1/S) sum_s (mean|X_s - Y_s| + mean|ln max(X_s,eps) - ln max(Y_s,eps)|)``.

    ``x`` is the target, ``y`` the decoded signal gradients flow to.  Default scales are the
    discriminator's (``msstftd.py:174-175``), ``hop = n_fft/4``.  With ``group`` the means run over
    the batch of ALL ranks (one all-reduce of ``2S`` doubles)."""
    hops = [n // 4 for n in n_ffts] if hops is None else list(hops)
    rows = x.numel() // x.shape[-1]
    counts = mrstft_counts(rows, x.shape[-1], n_ffts, hops)
    if not return_terms and torch.is_grad_enabled() and isinstance(y, torch.Tensor) and y.requires_grad \
            and x.shape[-1] % 2 == 0:
        # training step: value and gradient in one fused pass
        xr, yr = _rows(x, "x"), _rows(y, "y")
        if xr.shape != yr.shape:
            raise ValueError(f"Shape mismatch: x {tuple(x.shape)} vs y {tuple(y.shape)}")
        wl = list(n_ffts) if wins is None else list(wins)
        assert len(n_ffts) == len(hops) == len(wl)                                      # msstftd.py:177
        w = 1
        if group is not None:
            import torch.distributed as dist
            w = dist.get_world_size(group)
        inv = torch.tensor([1.0 / (len(n_ffts) * c * w) for c in counts for _ in (0, 1)], dtype=torch.float64, device=x.device)
        return _MRSTFTLossFused.apply(xr, yr, list(n_ffts), list(hops), wl, eps, inv, group)
    sums = mrstft_sums(x, y, n_ffts, hops, wins, eps)
    world = 1
    if group is not None:
        import torch.distributed as dist
        world = dist.get_world_size(group)
        # value = global sums; gradient flows to the local shard only, with the global normaliser
        # (the input gradient needs no collective)
        local = sums.detach()
        sums = sums + (_allreduce_(local, group) - local)
    inv = torch.tensor([1.0 / (c * world) for c in counts for _ in (0, 1)], dtype=torch.float64, device=sums.device)
    terms = sums * inv
    total = terms.sum() / len(n_ffts)
    total = total.float()
    if return_terms:
        t = terms.float()
        return total, [(t[2 * s], t[2 * s + 1]) for s in range(len(n_ffts))]
    return total


def _allreduce_(t: torch.Tensor, group) -> torch.Tensor:
    import torch.distributed as dist
    t = t.clone()
    dist.all_reduce(t, group=group)
    return t


class _SpecLosses(torch.autograd.Function):
    """out5 = [magnitude_loss, phase_loss, total_loss, spec_snr, mean(mse)], per_sample [B]."""

    @staticmethod
    def forward(ctx, obs, dec):
        o = obs.contiguous().float()
        d = dec.contiguous().float()
        B, Cc, K, F = o.shape
        lib = L.lib()
        ws_bytes = int(lib.tac_spec_losses_workspace_bytes(B))
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=o.device)
        out5 = torch.empty(5, dtype=torch.float32, device=o.device)
        per = torch.empty(B, dtype=torch.float32, device=o.device)
        with L.device_of(o):
            L.check(lib.tac_spec_losses(o.data_ptr(), d.data_ptr(), B, K, F, out5.data_ptr(), per.data_ptr(),
                                        ws.data_ptr(), ws_bytes, L.stream_of(o)))
        ctx.save_for_backward(o, d, out5)
        ctx.meta = (dec.shape, dec.dtype)
        return out5, per

    @staticmethod
    def backward(ctx, g5, gper):
        o, d, out5 = ctx.saved_tensors
        B, Cc, K, F = o.shape
        n1 = float(B * K * F)
        # every output is  sum_{b,c} a[b,c] * SE[b,c]  (+ the log of it for spec_snr), so the gradient is
        # grad_dec[b,c,:,:] = w[b,c] * (dec - obs) with the [B, 2] weight table below (device math, no sync)
        g5 = g5.float()
        w_tot = g5[2] / (2.0 * n1) + g5[4] * 0.5 / (2.0 * n1) \
            - g5[3] * (10.0 / 2.302585092994046) / (out5[2] * 2.0 * n1)
        w = torch.stack([g5[0] / n1 + w_tot, g5[1] / n1 + w_tot]).reshape(1, 2).expand(B, 2)
        w = w + gper.float().reshape(B, 1) * (0.5 / (2.0 * K * F))
        w = (2.0 * w).contiguous()
        grad = torch.empty_like(d)
        with L.device_of(o):
            L.check(L.lib().tac_spec_losses_bwd(o.data_ptr(), d.data_ptr(), B, K, F, w.data_ptr(), grad.data_ptr(),
                                                L.stream_of(o)))
        shape, dtype = ctx.meta
        return None, grad.reshape(shape).to(dtype)


def spectral_losses(obs: torch.Tensor, obs_dec: torch.Tensor) -> Dict[str, torch.Tensor]:
    """The reference's spectral reductions for ``[B, 2, K, F]`` stacks (channel 0 magnitude, 1 phase).

    Keys follow ``pkl_file_models.py:933-937``: ``magnitude_loss``, ``phase_loss``, ``total_loss``; plus
    ``spec_snr`` (``:940``), ``mse`` (per-sample ``0.5 * mean sq``, ``:924``) and ``mse_mean``."""
    L.require_cuda(obs, "obs")
    L.require_cuda(obs_dec, "obs_dec")
    if obs.shape != obs_dec.shape:
        raise ValueError(f"Shape mismatch: obs {tuple(obs.shape)} vs obs_dec {tuple(obs_dec.shape)}")
    if obs.dim() != 4 or obs.shape[1] != 2:
        raise ValueError(f"expected [B, 2, K, F], got {tuple(obs.shape)}")
    out5, per = _SpecLosses.apply(obs, obs_dec)
    return {"magnitude_loss": out5[0], "phase_loss": out5[1], "total_loss": out5[2], "spec_snr": out5[3],
            "mse_mean": out5[4], "mse": per}
