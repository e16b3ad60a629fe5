"""Host-side mirror of the two small call sites either side of the spectral path (``csrc/tail.cu``).

* :func:`bottleneck_mix` / :func:`spectral_bottleneck` -- the bottleneck that is live in the SpectralRes models
  (SURVEY.md 8a row a9): ``pkl_file_models.py:1108-1127`` (SpectralResE2D2), ``:692-718`` (SpectralResE4D1), ``:335-354``
  (SpectralResE2D1).  ``spectral_bottleneck(..., mode="dpca")`` is the hook where the DPCA projection of
  ``:300-333`` / ``:636-679`` (commented out upstream) is swapped back in through :func:`tac_spectral.dpca_project`.
* :func:`task_aware_tail` -- the loss tail of ``task_aware()`` (row a5-i, ``train_pkl_file_task_aware.py:307-364``):
  waveform MSE after ``F.interpolate`` and the feature-matching / adversarial sum over the discriminator's outputs.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence, Tuple

import torch

from . import _lib as L


def _mix_raw(zs: Sequence[torch.Tensor], alphas, split: int, trunc: int) -> torch.Tensor:
    z0 = zs[0]
    B, d = z0.shape
    w = int(L.lib().tac_bottleneck_width(d, split, trunc))
    out = torch.empty((B, len(zs) * w), dtype=torch.float32, device=z0.device)
    if out.numel() == 0:
        return out
    al = (C.c_void_p * ((len(zs) + 1) // 2))(*[(a.data_ptr() if a is not None else None) for a in alphas])
    with L.device_of(z0):
        L.check(L.lib().tac_bottleneck_mix(L.ptr_array(zs), len(zs), B, d, split, trunc, al, out.data_ptr(), L.stream_of(z0)))
    return out


class _Mix(torch.autograd.Function):
    @staticmethod
    def forward(ctx, split, trunc, n_alpha, *tensors):
        alphas, zs = list(tensors[:n_alpha]), [z.contiguous().float() for z in tensors[n_alpha:]]
        ctx.cfg = (split, trunc, n_alpha, len(zs), tuple(zs[0].shape))
        ctx.save_for_backward(*[a for a in alphas if a is not None])
        ctx.have = [a is not None for a in alphas]
        return _mix_raw(zs, alphas, split, trunc)

    @staticmethod
    def backward(ctx, g):
        split, trunc, n_alpha, nv, (B, d) = ctx.cfg
        saved = list(ctx.saved_tensors)
        alphas = [saved.pop(0) if h else None for h in ctx.have]
        g = g.contiguous().float()
        gz = [torch.empty((B, d), dtype=torch.float32, device=g.device) for _ in range(nv)]
        al = (C.c_void_p * ((nv + 1) // 2))(*[(a.data_ptr() if a is not None else None) for a in alphas])
        with L.device_of(g):
            L.check(L.lib().tac_bottleneck_mix_bwd(g.data_ptr(), nv, B, d, split, trunc, al, L.ptr_array(gz), L.stream_of(g)))
        return (None, None, None) + (None,) * n_alpha + tuple(gz)


def bottleneck_mix(zs: Sequence[torch.Tensor], split: int, trunc: int,
                   alphas: Optional[Sequence[Optional[torch.Tensor]]] = None) -> torch.Tensor:
    """``zs``: the views' latents ``[B, d]``; returns ``[B, nv * w]`` with view ``i`` in columns ``[i w, (i+1) w)``:
    ``cat(z_i[:, :split][:, :trunc], share_i[:, :trunc])`` where ``share_i`` is ``z_i[:, split:]`` or, for the pair
    ``(2p, 2p+1)`` with ``alphas[p]`` given (``[B]`` or ``[B, 1]``), ``alpha * share_2p + (1 - alpha) * share_2p+1``."""
    if len(zs) < 1:
        raise ValueError("need at least one view")
    B, d = zs[0].shape
    for z in zs:
        L.require_cuda(z, "z")
        if tuple(z.shape) != (B, d):
            raise ValueError("all views must have the same [B, d] shape")
    n_pairs = (len(zs) + 1) // 2
    al: List[Optional[torch.Tensor]] = [None] * n_pairs
    if alphas is not None:
        if len(alphas) != n_pairs:
            raise ValueError(f"expected {n_pairs} alpha tensors (one per view pair)")
        for p, a in enumerate(alphas):
            if a is not None:
                L.require_cuda(a, "alpha")
                if a.numel() != B:
                    raise ValueError("alpha must hold one value per batch row")
                al[p] = a.detach().reshape(B).contiguous().float()
    return _Mix.apply(int(split), int(trunc), n_pairs, *al, *zs)


def spectral_bottleneck(zs: Sequence[torch.Tensor], total_features_after: int, random_bottle_neck: bool, *, model: str = "E2D2",
                        mode: str = "mix", dim_p: Optional[int] = None, generator: Optional[torch.Generator] = None
                        ) -> Tuple[List[torch.Tensor], dict]:
    """The bottleneck block of the SpectralRes models as one call; returns ``(z_samples, dimension_info)`` with one
    ``[B, w]`` sample per view (views of ONE buffer: ``torch.cat(z_samples, 1)`` is ``z_sample`` of ``:348`` / ``:711``).

    ``mode="mix"`` (what upstream runs today): split / alpha-mix / truncate -- ``model`` picks the split column:
    ``"E2D2"`` / ``"E2D1"``: ``d // 2`` (``:1110``, ``:337``); ``"E4D1"``: the SUM of the views' widths (``:633``), which leaves
    the shared part empty exactly as upstream does.  ``alpha = torch.rand(B, 1)`` per pair is drawn here from ``generator``
    (``:1116``, ``:698-699``, ``:345``) when ``random_bottle_neck``.
    ``mode="dpca"``: the DPCA projection upstream has commented out at ``:300-333`` (``dim_p = num_features / 2``) and
    ``:636-679`` (``dim_p = randint(8, num_features / 2)``) -- ``dpca_project(zs, dim_p)``; samples keep their width ``d``."""
    from .dpca import dpca_project
    nv = len(zs)
    B, d = zs[0].shape
    info = {f"before_z{i + 1}": d for i in range(nv)}
    if mode == "dpca":
        if random_bottle_neck:
            num_features = nv * d
            if dim_p is None:
                dim_p = int(num_features / 2) if nv == 2 else int(torch.randint(8, int(num_features / 2), (1,), generator=generator).item())
            outs, pca = dpca_project(list(zs), int(dim_p), return_info=True)
            info.update({f"after_z{i + 1}": int(c) for i, c in enumerate(pca["cnt"].tolist())})      # dimension_info of :323-328
        else:
            outs = list(zs)
            info.update({f"after_z{i + 1}": d for i in range(nv)})
        return outs, info
    if mode != "mix":
        raise ValueError(f"unknown bottleneck mode {mode!r} (mix | dpca)")
    if model in ("E2D2", "E2D1"):
        split = d // 2
    elif model == "E4D1":
        split = nv * d
    else:
        raise ValueError(f"unknown model {model!r} (E2D2 | E2D1 | E4D1)")
    trunc = int(total_features_after / 4)
    alphas = None
    if random_bottle_neck:
        alphas = [torch.rand(B, 1, device=zs[0].device, generator=generator) for _ in range((nv + 1) // 2)]
    out = bottleneck_mix(zs, split, trunc, alphas)
    w = out.shape[1] // nv
    samples = [out[:, i * w:(i + 1) * w] for i in range(nv)]
    info.update({f"after_z{i + 1}": w for i in range(nv)})
    return samples, info


# ------------------------------------------------------------------------------------------------ task_aware tail
def interp_mse(a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    """``F.mse_loss`` of two ``[B, L]`` waveforms after the shorter one is resampled to the longer length with
    ``F.interpolate(mode='linear', align_corners=False)`` (``train_pkl_file_task_aware.py:311-333``); float64 device scalar."""
    L.require_cuda(a, "a")
    L.require_cuda(b, "b")
    if a.dim() != 2 or b.dim() != 2 or a.shape[0] != b.shape[0]:
        raise ValueError(f"expected two [B, L] tensors with the same B, got {tuple(a.shape)}, {tuple(b.shape)}")
    a, b = a.contiguous().float(), b.contiguous().float()
    out = torch.empty(1, dtype=torch.float64, device=a.device)
    nb = int(L.lib().tac_interp_mse_workspace_bytes())
    ws = torch.empty(nb, dtype=torch.uint8, device=a.device)
    with L.device_of(a):
        L.check(L.lib().tac_interp_mse(a.data_ptr(), a.shape[1], b.data_ptr(), b.shape[1], a.shape[0], out.data_ptr(), ws.data_ptr(),
                                       nb, L.stream_of(a)))
    return out[0]


L1, MSE, MSE_CONST = 0, 1, 2


def multi_reduce(a: Sequence[torch.Tensor], b: Sequence[Optional[torch.Tensor]], modes: Sequence[int],
                 consts: Optional[Sequence[float]] = None) -> torch.Tensor:
    """``out[0] = sum_s mean_s`` and ``out[1 + s] = mean_s`` over tensor pairs of any shapes in two launches (float64)."""
    n = len(a)
    if not (n == len(b) == len(modes)) or n < 1:
        raise ValueError("a, b and modes must have the same non-zero length")
    ac, bc = [], []
    for s in range(n):
        L.require_cuda(a[s], "a")
        ac.append(a[s].contiguous().float())
        if modes[s] != MSE_CONST:
            L.require_cuda(b[s], "b")
            if b[s].shape != a[s].shape:
                raise ValueError(f"pair {s}: shapes differ {tuple(a[s].shape)} vs {tuple(b[s].shape)}")
            bc.append(b[s].contiguous().float())
        else:
            bc.append(None)
    dev = ac[0].device
    sizes = (C.c_int64 * n)(*[t.numel() for t in ac])
    md = L.int_array(modes)
    cs = (C.c_float * n)(*[float(c) for c in (consts if consts is not None else [0.0] * n)])
    bp = (C.c_void_p * n)(*[(t.data_ptr() if t is not None else None) for t in bc])
    out = torch.empty(1 + n, dtype=torch.float64, device=dev)
    nb = int(L.lib().tac_multi_reduce_workspace_bytes(n, sizes))
    ws = torch.empty(nb, dtype=torch.uint8, device=dev)
    with L.device_of(ac[0]):
        L.check(L.lib().tac_multi_reduce(n, L.ptr_array(ac), bp, sizes, md, cs, out.data_ptr(), ws.data_ptr(), nb, L.stream_of(ac[0])))
    return out


def task_aware_tail(enhanced_audio: torch.Tensor, reconstructed_clean_waveforms: torch.Tensor,
                    disc_enhanced, disc_ground_truth) -> Tuple[torch.Tensor, torch.Tensor]:
    """The reductions of ``task_aware()`` after the discriminator ran (``train_pkl_file_task_aware.py:307-364``):

    * ``mse_loss`` (``:311-333``): the two ``[B, L]`` waveforms, the shorter resampled linearly to the longer;
    * ``discriminator_loss`` (``:352-361``): ``sum F.l1_loss`` over every feature-map pair, plus
      ``F.mse_loss(y_enh, 1) + F.mse_loss(y_gt, 0)`` over every logit pair.

    ``disc_enhanced`` / ``disc_ground_truth`` are the discriminator's ``(logits, fmaps)`` results (``msstftd.py:185-192``) for
    the enhanced and the ground-truth waveform.  Returns two float64 device scalars (the reference returns ``.item()``s)."""
    mse = interp_mse(enhanced_audio.float(), reconstructed_clean_waveforms.float())
    (y_e, f_e), (y_g, f_g) = disc_enhanced, disc_ground_truth
    a, b, modes, consts = [], [], [], []
    for fe, fg in zip(f_e, f_g):
        for xe, xg in zip(fe, fg):
            a.append(xe); b.append(xg); modes.append(L1); consts.append(0.0)
    for ye, yg in zip(y_e, y_g):
        a.append(ye); b.append(None); modes.append(MSE_CONST); consts.append(1.0)
        a.append(yg); b.append(None); modes.append(MSE_CONST); consts.append(0.0)
    if len(a) <= 48:
        return mse, multi_reduce(a, b, modes, consts)[0]
    total = None
    for i in range(0, len(a), 48):
        part = multi_reduce(a[i:i + 48], b[i:i + 48], modes[i:i + 48], consts[i:i + 48])[0]
        total = part if total is None else total + part
    return mse, total
