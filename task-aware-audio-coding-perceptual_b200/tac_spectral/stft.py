"""Host-side mirror of the reference's STFT call sites, backed by ``tac_stft_multi`` (K1).

* :class:`Spectrogram` -- drop-in for ``torchaudio.transforms.Spectrogram`` as the reference
  constructs it at ``msstftd.py:125-127`` (same keyword arguments, same ``[..., K, F]`` complex64
  result for ``power=None``), plus :meth:`Spectrogram.planes` which returns the tensor the
  discriminator actually consumes after ``msstftd.py:152-153`` (``cat(real, imag)`` + rearrange).
* :class:`MultiScaleSTFT` -- the per-scale loop of ``MultiScaleSTFTDiscriminator.forward``
  (``msstftd.py:185-192``), front end only, all scales in ONE kernel launch.
* :func:`complex_spectrogram` / :func:`create_complex_spectrogram` -- the notebook front end
  (``spectrogram.ipynb`` cell 10 L26-52, raw ``:4120-4134``) batched on the GPU, optionally writing
  the padded ``[B, 2, 1025, 600]`` model input of ``SpectralDataset.pad_tensor`` + ``torch.stack``
  (``train_pkl_file_task_aware.py:64-68``, ``pkl_file_models.py:889-892``) directly.
"""
from __future__ import annotations

import ctypes as C
import math
from typing import Callable, Dict, List, Optional, Sequence

import torch
from torch import nn

from . import _lib as L

SUPPORTED_NFFT = (64, 128, 256, 512, 1024, 2048)


def num_frames(T: int, n_fft: int, hop: int, center: bool) -> int:
    return int(L.lib().tac_stft_num_frames(int(T), int(n_fft), int(hop), int(bool(center))))


def _pad_mode_code(pad_mode: Optional[str]) -> int:
    if pad_mode in (None, "reflect"):
        return L.PAD_REFLECT
    if pad_mode in ("constant", "zeros"):
        return L.PAD_ZERO
    raise ValueError(f"unsupported pad_mode {pad_mode!r} (reflect | constant)")


def _check_window(window_fn: Callable, wkwargs) -> None:
    # same error class the reference raises for an unsupported window (train_pkl_file_task_aware.py:190-191)
    if window_fn is not torch.hann_window or (wkwargs and wkwargs.get("periodic", True) is not True):
        raise ValueError("Unsupported window: only the periodic torch.hann_window is implemented")


def ex_supported(n_fft: int) -> bool:
    """Frame lengths ``tac_stft_ex`` accepts (warp FFT core or chirp-z)."""
    n = int(n_fft)
    return n >= 8 and n + n // 2 <= 4096


_EX_PLANES = {L.EPI_COMPLEX: None, L.EPI_COMPRESS: None, L.EPI_MAGPHASE: 2, L.EPI_MAGDB: 2, L.EPI_POWER: 1,
              L.EPI_LOGPOWER: 1, L.EPI_DB: 1}


def stft_ex(x: torch.Tensor, n_fft: int, hop: int, win: Optional[int] = None, *, window: str = "hann", center: bool = True,
            pad_mode: Optional[str] = "reflect", normalized: bool = False, kind: int = L.EPI_COMPLEX, p0: float = 0.0,
            p1: float = 0.0, p2: float = 0.0, f_pad: int = 0, mel: Optional["MelBank"] = None,
            rowmax: Optional[torch.Tensor] = None) -> torch.Tensor:
    """One scale of ``x [rows, T]`` through ``tac_stft_ex`` (any ``n_fft``; epilogue ``kind`` of include/tac_spectral.h).
    Returns complex64 ``[rows, K, f_pad]``, float32 ``[rows, 2, K, f_pad]`` / ``[rows, K, f_pad]`` or ``[rows, n_mels, f_pad]``."""
    L.require_cuda(x, "x")
    if x.dim() != 2:
        raise ValueError(f"expected x of shape [rows, T], got {tuple(x.shape)}")
    if window not in ("hann", "sqrthann"):
        raise NotImplementedError(f"Window type {window} not implemented!")           # data_module.py:19
    x = x.contiguous().float()
    rows, T = x.shape
    n_fft, hop = int(n_fft), int(hop)
    win = n_fft if win is None else int(win)
    F = (T + (2 * (n_fft // 2) if center else 0) - n_fft) // hop + 1 if T + (2 * (n_fft // 2) if center else 0) >= n_fft else 0
    if F < 1:
        raise ValueError(f"signal of {T} samples is shorter than n_fft={n_fft}")
    K, Fp = n_fft // 2 + 1, int(f_pad) if f_pad else F
    ep = L.EpilogueStruct()
    ep.kind, ep.p0, ep.p1, ep.p2 = int(kind), float(p0), float(p1), float(p2)
    if kind == L.EPI_MEL:
        if mel is None or mel.n_freqs != K:
            raise ValueError(f"the mel epilogue needs a MelBank with n_freqs={K}")
        fb, rng = mel.on(x.device)
        ep.n_mels, ep.mel_fb, ep.mel_range = mel.n_mels, fb.data_ptr(), rng.data_ptr()
        out = torch.empty((rows, mel.n_mels, Fp), dtype=torch.float32, device=x.device)
    elif _EX_PLANES[kind] is None:
        out = torch.empty((rows, K, Fp), dtype=torch.complex64, device=x.device)
    elif _EX_PLANES[kind] == 2:
        out = torch.empty((rows, 2, K, Fp), dtype=torch.float32, device=x.device)
    else:
        out = torch.empty((rows, K, Fp), dtype=torch.float32, device=x.device)
    if rowmax is not None:
        if rowmax.dtype != torch.int32 or rowmax.numel() != rows or not rowmax.is_cuda:
            raise ValueError("rowmax must be a zero-filled CUDA int32 tensor of `rows` elements")
        ep.rowmax = rowmax.data_ptr()
    with L.device_of(x):
        L.check(L.lib().tac_stft_ex(x.data_ptr(), rows, T, n_fft, hop, win, L.WINDOW_SQRT_HANN if window == "sqrthann" else L.WINDOW_HANN,
                                    int(bool(center)), _pad_mode_code(pad_mode), int(bool(normalized)), C.byref(ep), Fp,
                                    out.data_ptr(), L.stream_of(x)))
    return out


def _hz_to_mel(f: float) -> float:
    return 2595.0 * math.log10(1.0 + f / 700.0)


class MelBank:
    """Triangular HTK mel filter bank ``[n_freqs, n_mels]`` -- the matrix ``torchaudio.functional.melscale_fbanks(n_freqs,
    f_min, f_max, n_mels, sample_rate, norm=None, mel_scale='htk')`` defines -- generated on the host in
    float64, plus each filter's non-zero row range for the kernel."""

    def __init__(self, n_freqs: int, f_min: float = 0.0, f_max: float = 8000.0, n_mels: int = 80, sample_rate: int = 16000):
        self.n_freqs, self.n_mels = int(n_freqs), int(n_mels)
        freqs = [i * (sample_rate // 2) / (self.n_freqs - 1) for i in range(self.n_freqs)]
        m_lo, m_hi = _hz_to_mel(f_min), _hz_to_mel(f_max)
        pts = [700.0 * (10.0 ** ((m_lo + (m_hi - m_lo) * i / (self.n_mels + 1)) / 2595.0) - 1.0) for i in range(self.n_mels + 2)]
        fb = torch.zeros((self.n_freqs, self.n_mels), dtype=torch.float64)
        rng = torch.zeros((self.n_mels, 2), dtype=torch.int32)
        for m in range(self.n_mels):
            lo, ce, hi = pts[m], pts[m + 1], pts[m + 2]
            first, last = self.n_freqs, 0
            for k, f in enumerate(freqs):
                v = max(0.0, min((f - lo) / (ce - lo), (hi - f) / (hi - ce)))
                if v > 0.0:
                    fb[k, m] = v
                    first, last = min(first, k), max(last, k + 1)
            rng[m, 0], rng[m, 1] = (first, last) if last > first else (0, 0)
        self.fb, self.range = fb.to(torch.float32), rng
        self._dev = {}

    def on(self, device):
        key = str(device)
        if key not in self._dev:
            self._dev[key] = (self.fb.to(device).contiguous(), self.range.to(device).contiguous())
        return self._dev[key]


def stft_multi(x: torch.Tensor, n_ffts: Sequence[int], hops: Sequence[int], wins: Optional[Sequence[int]] = None, *,
               center: bool = False, pad_mode: Optional[str] = "reflect", normalized: bool = True,
               out_mode: int = L.OUT_PLANES, f_pad: int = 0) -> List[torch.Tensor]:
    """All scales of ``x [B, C, T]`` in one call of ``tac_stft_multi``; see ``include/tac_spectral.h``.

    Differentiable like the ``torchaudio.transforms.Spectrogram`` it replaces (``msstftd.py:125-127``): when ``x``
    requires grad the call goes through :class:`_STFTMulti`, whose backward is the adjoint transform
    (``tac_stft_adjoint``), so adversarial / feature-matching losses reach the generated waveform."""
    L.require_cuda(x, "x")
    if x.dim() != 3:
        raise ValueError(f"expected x of shape [B, C, T], got {tuple(x.shape)}")
    wins = list(n_ffts) if wins is None else list(wins)
    assert len(n_ffts) == len(hops) == len(wins)                       # msstftd.py:177
    if torch.is_grad_enabled() and x.requires_grad:
        cfg = (tuple(n_ffts), tuple(hops), tuple(wins), bool(center), pad_mode, bool(normalized), int(out_mode), int(f_pad))
        return list(_STFTMulti.apply(x, cfg))
    return _stft_multi_raw(x, n_ffts, hops, wins, center, pad_mode, normalized, out_mode, f_pad)


def _stft_multi_raw(x, n_ffts, hops, wins, center, pad_mode, normalized, out_mode, f_pad) -> List[torch.Tensor]:
    x = x.contiguous().float()
    B, Cc, T = x.shape
    outs = []
    for n, h in zip(n_ffts, hops):
        F = num_frames(T, n, h, center)
        if F < 1:
            raise ValueError(f"signal of {T} samples is shorter than n_fft={n}")
        K = n // 2 + 1
        if out_mode == L.OUT_PLANES:
            outs.append(torch.empty((B, 2 * Cc, F, K), device=x.device, dtype=torch.float32))
        elif out_mode == L.OUT_COMPLEX:
            outs.append(torch.empty((B, Cc, K, F), device=x.device, dtype=torch.complex64))
        else:
            outs.append(torch.empty((B * Cc, 2, K, f_pad if f_pad > 0 else F), device=x.device, dtype=torch.float32))
    with L.device_of(x):
        L.check(L.lib().tac_stft_multi(x.data_ptr(), B, Cc, T, len(n_ffts), L.int_array(n_ffts), L.int_array(hops),
                                       L.int_array(wins), int(bool(center)), _pad_mode_code(pad_mode),
                                       int(bool(normalized)), out_mode, int(f_pad), L.ptr_array(outs), L.stream_of(x)))
    return outs


def stft_multi_power(x: torch.Tensor, n_ffts: Sequence[int], hops: Sequence[int], wins: Optional[Sequence[int]] = None, *,
                     center: bool = False, pad_mode: Optional[str] = "reflect", normalized: bool = True, power: float = 1.0,
                     log_multiplier: float = 0.0, log_eps: float = 1e-10) -> List[torch.Tensor]:
    """Magnitude (``power=1``) / power (``power=2``) spectrograms of ALL scales of ``x [B, C, T]`` from one launch of the
    multi-scale kernel (``tac_stft_multi_epi``): one ``[B, C, F_s, K_s]`` float32 tensor per scale, frequency innermost
    (transpose the last two axes for torchaudio's ``[..., K, F]``).  ``log_multiplier != 0`` gives
    ``log_multiplier * log10(|z|^power + log_eps)`` -- ``testing_scripts.ipynb:246-249`` is ``power=2, log_multiplier=10``.
    The magnitude is formed where the transform ends, so half the bytes of the re / im planes reach HBM.  Forward only."""
    L.require_cuda(x, "x")
    if x.dim() != 3:
        raise ValueError(f"expected x of shape [B, C, T], got {tuple(x.shape)}")
    if torch.is_grad_enabled() and x.requires_grad:
        raise NotImplementedError("stft_multi_power is forward-only; use stft_multi (planes) for a differentiable path")
    wins = list(n_ffts) if wins is None else list(wins)
    assert len(n_ffts) == len(hops) == len(wins)                       # msstftd.py:177
    x = x.contiguous().float()
    B, Cc, T = x.shape
    outs = []
    for n, h in zip(n_ffts, hops):
        F = num_frames(T, n, h, center)
        if F < 1:
            raise ValueError(f"signal of {T} samples is shorter than n_fft={n}")
        outs.append(torch.empty((B, Cc, F, n // 2 + 1), device=x.device, dtype=torch.float32))
    epi = L.EpilogueStruct()
    epi.kind = L.EPI_LOGPOWER if log_multiplier != 0.0 else L.EPI_POWER
    epi.p0, epi.p1, epi.p2 = float(power), float(log_multiplier), float(log_eps)
    with L.device_of(x):
        L.check(L.lib().tac_stft_multi_epi(x.data_ptr(), B, Cc, T, len(n_ffts), L.int_array(n_ffts), L.int_array(hops),
                                           L.int_array(wins), int(bool(center)), _pad_mode_code(pad_mode),
                                           int(bool(normalized)), C.byref(epi), L.ptr_array(outs), L.stream_of(x)))
    return outs


def _window_norm(win: int) -> float:
    """``1 / sqrt(sum w^2)`` of the periodic Hann window (torchaudio's ``normalized=True`` / 'window')."""
    w = torch.hann_window(win, periodic=True, dtype=torch.float64)
    return float(1.0 / torch.sqrt((w * w).sum()))


def _stft_adjoint(a: torch.Tensor, b: Optional[torch.Tensor], in_mode: int, rows: int, K: int, F: int, n_fft: int, hop: int,
                  win: int, scale: float, out_len: int) -> torch.Tensor:
    """``[rows, out_len]`` = adjoint of the ``center=False`` analysis STFT applied to ``a + i b`` ([rows, K, F])."""
    if n_fft not in SUPPORTED_NFFT or hop * 8 < n_fft or hop > n_fft:
        raise NotImplementedError(f"STFT backward needs n_fft in {SUPPORTED_NFFT} and hop in [n_fft/8, n_fft] "
                                  f"(got n_fft={n_fft}, hop={hop})")
    out = torch.empty((rows, out_len), dtype=torch.float32, device=a.device)
    with L.device_of(a):
        L.check(L.lib().tac_stft_adjoint(a.data_ptr(), b.data_ptr() if b is not None else None, in_mode, rows, K * F, F, F,
                                         n_fft, hop, win, float(scale), out.data_ptr(), out_len, L.stream_of(a)))
    return out


class _STFTMulti(torch.autograd.Function):
    """``tac_stft_multi`` with the backward torchaudio's Spectrogram has: per scale the adjoint of framing + window +
    one-sided DFT (``tac_stft_adjoint``, the inverse-STFT kernel without edge weights and envelope), then the adjoint of
    the centre padding (crop for zeros, fold-back for reflect); magnitude / phase outputs add the polar chain rule."""

    @staticmethod
    def forward(ctx, x, cfg):
        n_ffts, hops, wins, center, pad_mode, normalized, out_mode, f_pad = cfg
        outs = _stft_multi_raw(x.detach(), n_ffts, hops, wins, center, pad_mode, normalized, out_mode, f_pad)
        ctx.cfg, ctx.xshape, ctx.xdtype = cfg, tuple(x.shape), x.dtype
        if out_mode == L.OUT_MAGPHASE:
            ctx.save_for_backward(*outs)
        return tuple(outs)

    @staticmethod
    def backward(ctx, *gs):
        n_ffts, hops, wins, center, pad_mode, normalized, out_mode, f_pad = ctx.cfg
        B, Cc, T = ctx.xshape
        rows = B * Cc
        total = None
        for i, (n, h, w) in enumerate(zip(n_ffts, hops, wins)):
            g = gs[i]
            if g is None:
                continue
            K, pad = n // 2 + 1, (n // 2 if center else 0)
            F = num_frames(T, n, h, center)
            if out_mode == L.OUT_COMPLEX:                                   # [B, C, K, F] complex64
                a, b, mode = torch.view_as_real(g.contiguous().to(torch.complex64)), None, L.IN_COMPLEX
            elif out_mode == L.OUT_PLANES:                                  # [B, 2C, F, K] -> two [B, C, K, F] planes
                g = g.float()
                a = g[:, :Cc].permute(0, 1, 3, 2).contiguous()
                b = g[:, Cc:].permute(0, 1, 3, 2).contiguous()
                mode = L.IN_PLANES
            else:                                                           # [rows, 2, K, f_pad]: polar chain rule
                out = ctx.saved_tensors[i]
                Fv = min(F, out.shape[-1])
                mag, ph = out[:, 0, :, :Fv], out[:, 1, :, :Fv]
                gm, gp = g[:, 0, :, :Fv].float(), g[:, 1, :, :Fv].float()
                c, sn = torch.cos(ph), torch.sin(ph)
                gpm = torch.where(mag > 0, gp / mag.clamp_min(1e-30), torch.zeros_like(gp))
                a = (gm * c - gpm * sn).contiguous()
                b = (gm * sn + gpm * c).contiguous()
                mode, F = L.IN_PLANES, Fv
            scale = _window_norm(w) if normalized else 1.0
            gp_sig = _stft_adjoint(a, b, mode, rows, K, F, n, h, w, scale, T + 2 * pad)   # grad w.r.t. the padded signal
            gx = gp_sig[:, pad:pad + T]
            if pad and _pad_mode_code(pad_mode) == L.PAD_REFLECT:
                gx = gx.clone()
                gx[:, 1:pad + 1] += gp_sig[:, :pad].flip(-1)                 # xp[i] = x[pad - i],          i < pad
                gx[:, T - 1 - pad:T - 1] += gp_sig[:, pad + T:].flip(-1)     # xp[pad + T + j] = x[T - 2 - j]
            total = gx if total is None else total + gx
        if total is None:
            return None, None
        return total.reshape(B, Cc, T).to(ctx.xdtype), None


class Spectrogram(nn.Module):
    """``torchaudio.transforms.Spectrogram`` replacement for the configuration of ``msstftd.py:125-127``.

    Accepts the same keyword arguments.  Implemented: periodic Hann window, ``power=None`` (complex
    result), ``normalized`` in ``{False, True, "window"}``, ``center`` with reflect / constant padding,
    ``onesided=True``, ``pad=0``.  Anything else raises (no silent fallback to torch.stft).
    """

    def __init__(self, n_fft: int = 400, win_length: Optional[int] = None, hop_length: Optional[int] = None,
                 pad: int = 0, window_fn: Callable = torch.hann_window, power: Optional[float] = 2.0,
                 normalized=False, wkwargs: Optional[dict] = None, center: bool = True,
                 pad_mode: Optional[str] = "reflect", onesided: bool = True, return_complex: Optional[bool] = None):
        super().__init__()
        self.n_fft = int(n_fft)
        self.win_length = int(win_length) if win_length is not None else self.n_fft
        self.hop_length = int(hop_length) if hop_length is not None else self.win_length // 2
        _check_window(window_fn, wkwargs)
        if not ex_supported(self.n_fft):
            raise ValueError(f"n_fft={n_fft} unsupported (8 <= n_fft, n_fft + n_fft // 2 <= 4096)")
        if power is not None and not power > 0:
            raise ValueError("power must be None (complex) or positive")
        if pad != 0 or not onesided:
            raise NotImplementedError("pad != 0 and onesided=False are not implemented")
        if normalized not in (False, True, "window"):
            raise NotImplementedError("normalized must be False, True or 'window'")
        self.normalized = bool(normalized)
        self.center = bool(center)
        self.pad_mode = pad_mode
        self.power = None if power is None else float(power)
        # torchaudio keeps the window as a buffer; kept so that state_dict()s stay interchangeable
        self.register_buffer("window", torch.hann_window(self.win_length), persistent=False)
        self._core = self.n_fft in SUPPORTED_NFFT and self.hop_length % 2 == 0      # the warp FFT core has a plan for it
        if torch.cuda.is_available():                                # tables now, so that forward() is capturable from call one
            if self._core:
                L.check(L.lib().tac_prepare_fft(self.n_fft, self.win_length, int(self.normalized)))
            else:
                L.check(L.lib().tac_prepare_stft_ex(self.n_fft, self.win_length, int(self.normalized), 0))

    def _run(self, x: torch.Tensor, out_mode: int) -> torch.Tensor:
        lead = x.shape[:-1]
        x3 = x.reshape(-1, 1, x.shape[-1]) if x.dim() != 3 else x
        out = stft_multi(x3, [self.n_fft], [self.hop_length], [self.win_length], center=self.center,
                         pad_mode=self.pad_mode, normalized=self.normalized, out_mode=out_mode)[0]
        return out, lead

    def _ex(self, waveform: torch.Tensor, kind: int, p0: float = 0.0, p1: float = 0.0, p2: float = 0.0) -> torch.Tensor:
        lead = waveform.shape[:-1]
        out = stft_ex(waveform.reshape(-1, waveform.shape[-1]), self.n_fft, self.hop_length, self.win_length,
                      center=self.center, pad_mode=self.pad_mode, normalized=self.normalized, kind=kind, p0=p0, p1=p1, p2=p2)
        return out.reshape(*lead, out.shape[-2], out.shape[-1])

    def forward(self, waveform: torch.Tensor) -> torch.Tensor:
        """``[..., T]`` -> complex64 ``[..., K, F]`` for ``power=None`` (what ``msstftd.py:151`` receives), else the
        real ``|X|^power`` (``torchaudio.transforms.Spectrogram()`` of ``testing_scripts.ipynb:246``: n_fft 400, power 2)."""
        needs_grad = torch.is_grad_enabled() and waveform.requires_grad
        if self.power is None:
            if self._core:
                out, lead = self._run(waveform, L.OUT_COMPLEX)
                return out.reshape(*lead, out.shape[-2], out.shape[-1])
            if needs_grad:
                raise NotImplementedError(f"the STFT backward is implemented for n_fft in {SUPPORTED_NFFT} only")
            return self._ex(waveform, L.EPI_COMPLEX)
        if needs_grad:                                   # differentiable like torchaudio's: |.|^p of the complex result
            if not self._core:
                raise NotImplementedError(f"the STFT backward is implemented for n_fft in {SUPPORTED_NFFT} only")
            out, lead = self._run(waveform, L.OUT_COMPLEX)
            return out.reshape(*lead, out.shape[-2], out.shape[-1]).abs().pow(self.power)
        return self._ex(waveform, L.EPI_POWER, self.power)

    def log_power(self, waveform: torch.Tensor, multiplier: float = 10.0, eps: float = 1e-10) -> torch.Tensor:
        """``multiplier * log10(Spectrogram(power)(x) + eps)`` in the same kernel (``testing_scripts.ipynb:246-249``)."""
        return self._ex(waveform, L.EPI_LOGPOWER, 2.0 if self.power is None else self.power, float(multiplier), float(eps))

    def planes(self, x: torch.Tensor) -> torch.Tensor:
        """``[B, C, T]`` -> float32 ``[B, 2C, F, K]``: ``msstftd.py:151-153`` fused (no complex
        intermediate, no ``cat`` copy, no non-contiguous rearrange)."""
        if x.dim() != 3:
            raise ValueError(f"expected [B, C, T], got {tuple(x.shape)}")
        return self._run(x, L.OUT_PLANES)[0]


class MultiScaleSTFT(nn.Module):
    """Front end of ``MultiScaleSTFTDiscriminator`` (``msstftd.py:162-192``): constructor lists as in
    ``:171-177`` (defaults ``:174-175``); ``forward`` returns one ``[B, 2C, F_s, K_s]`` tensor per
    scale, all produced by a single kernel launch that stages the signal once."""

    def __init__(self, n_ffts: Sequence[int] = (1024, 2048, 512), hop_lengths: Sequence[int] = (256, 512, 128),
                 win_lengths: Sequence[int] = (1024, 2048, 512), normalized: bool = True):
        super().__init__()
        assert len(n_ffts) == len(hop_lengths) == len(win_lengths)    # msstftd.py:177
        self.n_ffts, self.hop_lengths, self.win_lengths = list(n_ffts), list(hop_lengths), list(win_lengths)
        self.normalized = normalized
        for n in self.n_ffts:
            if n not in SUPPORTED_NFFT:
                raise ValueError(f"n_fft={n} unsupported; choose one of {SUPPORTED_NFFT}")
        if torch.cuda.is_available():
            for n, w in zip(self.n_ffts, self.win_lengths):
                L.check(L.lib().tac_prepare_fft(n, w, int(bool(normalized))))

    def forward(self, x: torch.Tensor) -> List[torch.Tensor]:
        return stft_multi(x, self.n_ffts, self.hop_lengths, self.win_lengths, center=False,
                          normalized=self.normalized, out_mode=L.OUT_PLANES)

    def magnitudes(self, x: torch.Tensor, power: float = 1.0, log_multiplier: float = 0.0, log_eps: float = 1e-10) -> List[torch.Tensor]:
        """``|STFT_s(x)|^power`` (optionally ``log_multiplier * log10(. + log_eps)``) of every scale, ``[B, C, F_s, K_s]``, from
        the same single launch (:func:`stft_multi_power`)."""
        return stft_multi_power(x, self.n_ffts, self.hop_lengths, self.win_lengths, center=False, normalized=self.normalized,
                                power=power, log_multiplier=log_multiplier, log_eps=log_eps)


def complex_spectrogram(y: torch.Tensor, n_fft: int = 2048, hop_length: int = 512, win_length: int = 2048,
                        pad_mode: str = "constant", f_pad: Optional[int] = None) -> torch.Tensor:
    """Batched notebook front end: ``y [B, T]`` -> ``[B, 2, n_fft/2+1, F]`` with channel 0 =
    ``np.abs(D)``, channel 1 = ``np.angle(D)``, ``D = librosa.stft(y, n_fft, hop, win, 'hann')``
    (centred, un-normalised; librosa 0.10 pads with zeros).  With ``f_pad`` the time axis is
    zero-padded / cropped to ``f_pad`` frames -- the reference's ``[B, 2, 1025, 600]`` model input."""
    L.require_cuda(y, "y")
    if y.dim() != 2:
        raise ValueError(f"expected y of shape [B, T], got {tuple(y.shape)}")
    if n_fft not in SUPPORTED_NFFT or hop_length % 2:
        # the dB-channel variant of the notebook frames with n_fft = win_length = 2046 (cell 9, raw :2897-2901): chirp-z path
        if torch.is_grad_enabled() and y.requires_grad:
            raise NotImplementedError(f"the STFT backward is implemented for n_fft in {SUPPORTED_NFFT} only")
        return stft_ex(y, n_fft, hop_length, win_length, center=True, pad_mode=pad_mode, normalized=False,
                       kind=L.EPI_MAGPHASE, f_pad=int(f_pad or 0))
    return stft_multi(y[:, None, :], [n_fft], [hop_length], [win_length], center=True, pad_mode=pad_mode,
                      normalized=False, out_mode=L.OUT_MAGPHASE, f_pad=int(f_pad or 0))[0]


def create_complex_spectrogram(y: torch.Tensor, sr: int, n_fft: int = 2048, hop_length: int = 512,
                               win_length: int = 2048, db_scale: bool = False) -> Dict[str, object]:
    """Same name, argument meaning and result keys as the notebook function (cell 10 L26-52):
    ``{'magnitude', 'phase', 'params'}`` for one utterance ``y [T]`` (or a batch ``[B, T]``).  ``db_scale=True`` is the
    cell 9 variant (raw ``:2894-2913``; it frames with ``n_fft = win_length = 2046``) and adds
    ``'db_scale' = librosa.amplitude_to_db(magnitude, ref=np.max)``: the per-utterance maximum leaves the STFT kernel with
    the magnitude plane, so the dB pass reads that plane once."""
    single = y.dim() == 1
    yb = y[None] if single else y
    out_d = {}
    if db_scale:
        L.require_cuda(yb, "y")
        rowmax = torch.zeros(yb.shape[0], dtype=torch.int32, device=yb.device)
        out = stft_ex(yb, n_fft, hop_length, win_length, center=True, pad_mode="constant", normalized=False,
                      kind=L.EPI_MAGPHASE, rowmax=rowmax)
        B, _, K, F = out.shape
        db = torch.empty((B, K, F), dtype=torch.float32, device=out.device)
        with L.device_of(out):
            L.check(L.lib().tac_db_from_rowmax(out.data_ptr(), B, 2 * K * F, K, F, F, rowmax.data_ptr(), 1e-5, 80.0,
                                               db.data_ptr(), L.stream_of(out)))
        out_d["db_scale"] = db[0] if single else db
    else:
        out = complex_spectrogram(yb, n_fft, hop_length, win_length)
    mag, ph = out[:, 0], out[:, 1]
    if single:
        mag, ph = mag[0], ph[0]
    return {"magnitude": mag, "phase": ph, **out_d,
            "params": {"sr": sr, "n_fft": n_fft, "hop_length": hop_length, "win_length": win_length, "window": "hann"}}


class MelSpectrogram(nn.Module):
    """``torchaudio.transforms.MelSpectrogram``-shaped module for the north star's mel alternative (SURVEY.md 8c item 5;
    the reference itself has no mel path): framing, window, FFT, ``|X|^power`` and the HTK filter bank in ONE kernel,
    ``[..., T]`` -> ``[..., n_mels, F]``.  ``log_mel`` adds ``multiplier * log10(mel + eps)`` in the same pass."""

    def __init__(self, sample_rate: int = 16000, n_fft: int = 400, win_length: Optional[int] = None,
                 hop_length: Optional[int] = None, f_min: float = 0.0, f_max: Optional[float] = None, pad: int = 0,
                 n_mels: int = 128, window_fn: Callable = torch.hann_window, power: float = 2.0, normalized: bool = False,
                 wkwargs: Optional[dict] = None, center: bool = True, pad_mode: str = "reflect", onesided: Optional[bool] = None,
                 norm: Optional[str] = None, mel_scale: str = "htk"):
        super().__init__()
        _check_window(window_fn, wkwargs)
        if norm is not None or mel_scale != "htk" or pad != 0:
            raise NotImplementedError("only norm=None, mel_scale='htk', pad=0 are implemented")
        if not ex_supported(n_fft):
            raise ValueError(f"n_fft={n_fft} unsupported (8 <= n_fft, n_fft + n_fft // 2 <= 4096)")
        self.n_fft = int(n_fft)
        self.win_length = int(win_length) if win_length is not None else self.n_fft
        self.hop_length = int(hop_length) if hop_length is not None else self.win_length // 2
        self.power, self.normalized, self.center, self.pad_mode = float(power), bool(normalized), bool(center), pad_mode
        self.bank = MelBank(self.n_fft // 2 + 1, f_min, float(f_max if f_max is not None else sample_rate // 2), n_mels, sample_rate)
        if torch.cuda.is_available():
            L.check(L.lib().tac_prepare_stft_ex(self.n_fft, self.win_length, int(self.normalized), 0))

    def _run(self, waveform, p1, p2):
        lead = waveform.shape[:-1]
        out = stft_ex(waveform.reshape(-1, waveform.shape[-1]), self.n_fft, self.hop_length, self.win_length, center=self.center,
                      pad_mode=self.pad_mode, normalized=self.normalized, kind=L.EPI_MEL, p0=self.power, p1=p1, p2=p2, mel=self.bank)
        return out.reshape(*lead, out.shape[-2], out.shape[-1])

    def forward(self, waveform: torch.Tensor) -> torch.Tensor:
        return self._run(waveform, 0.0, 0.0)

    def log_mel(self, waveform: torch.Tensor, multiplier: float = 10.0, eps: float = 1e-10) -> torch.Tensor:
        return self._run(waveform, float(multiplier), float(eps))


class SpecsFrontEnd(nn.Module):
    """The SGMSE+ enhancer's spectrogram front end: ``SpecsDataModule.stft`` + ``spec_fwd``
    (``sgmse/sgmse/data_module.py:162-175,212-214``; defaults ``:109-118``: n_fft 510, hop 128, periodic Hann,
    ``abs(z)**0.5 * exp(1j * angle(z)) * 0.15``) and ``pad_spec`` (``sgmse/sgmse/util/other.py:76-90``, frames zero-padded to a
    multiple of 64) as ONE kernel.  Same constructor keywords as the data module."""

    def __init__(self, n_fft: int = 510, hop_length: int = 128, window: str = "hann", spec_factor: float = 0.15,
                 spec_abs_exponent: float = 0.5, transform_type: str = "exponent", **ignored):
        super().__init__()
        if window not in ("hann", "sqrthann"):
            raise NotImplementedError(f"Window type {window} not implemented!")       # data_module.py:19
        if transform_type not in ("exponent", "none"):
            raise NotImplementedError(f"transform_type {transform_type!r}: only 'exponent' and 'none' are implemented")
        self.n_fft, self.hop_length, self.window = int(n_fft), int(hop_length), window
        self.spec_factor, self.spec_abs_exponent, self.transform_type = float(spec_factor), float(spec_abs_exponent), transform_type
        if torch.cuda.is_available() and ex_supported(self.n_fft):
            L.check(L.lib().tac_prepare_stft_ex(self.n_fft, self.n_fft, 0, 1 if window == "sqrthann" else 0))

    def stft(self, sig: torch.Tensor) -> torch.Tensor:
        """``torch.stft(sig, n_fft, hop, window, center=True, return_complex=True)`` (data_module.py:212-214)."""
        lead = sig.shape[:-1]
        out = stft_ex(sig.reshape(-1, sig.shape[-1]), self.n_fft, self.hop_length, self.n_fft, window=self.window, center=True,
                      pad_mode="reflect", kind=L.EPI_COMPLEX)
        return out.reshape(*lead, out.shape[-2], out.shape[-1])

    def forward(self, sig: torch.Tensor, pad64: bool = True) -> torch.Tensor:
        """``pad_spec(spec_fwd(stft(sig)))``: ``[..., T]`` -> complex64 ``[..., n_fft/2+1, ceil64(F)]``."""
        lead, T = sig.shape[:-1], sig.shape[-1]
        F = num_frames(T, self.n_fft, self.hop_length, True)        # (T + 2 (n_fft // 2) - n_fft) // hop + 1: odd n_fft too
        f_pad = (F + 63) // 64 * 64 if pad64 else F
        if self.transform_type == "none":
            kind, p0, p1 = L.EPI_COMPLEX, 0.0, 0.0
        else:
            kind, p0, p1 = L.EPI_COMPRESS, self.spec_abs_exponent, self.spec_factor
        out = stft_ex(sig.reshape(-1, T), self.n_fft, self.hop_length, self.n_fft, window=self.window, center=True,
                      pad_mode="reflect", kind=kind, p0=p0, p1=p1, f_pad=f_pad)
        return out.reshape(*lead, out.shape[-2], out.shape[-1])


def amplitude_to_db(spec: torch.Tensor, n_frames: Optional[int] = None, amin: float = 1e-5, top_db: float = 80.0) -> torch.Tensor:
    """``librosa.amplitude_to_db(magnitude, ref=np.max)`` per utterance (notebook cell 9 L31-63): ``spec`` is the
    ``[B, 2, K, f_pad]`` output of :func:`complex_spectrogram` (its magnitude plane is used in place) or a plain
    magnitude ``[B, K, F]``; ``n_frames`` = number of real (unpadded) frames.  Returns ``[B, K, n_frames]``."""
    L.require_cuda(spec, "spec")
    if spec.dim() == 4 and spec.shape[1] == 2:
        B, _, K, ld = spec.shape
        row_stride = 2 * K * ld
    elif spec.dim() == 3:
        B, K, ld = spec.shape
        row_stride = K * ld
    else:
        raise ValueError(f"expected [B, 2, K, F] or [B, K, F], got {tuple(spec.shape)}")
    if spec.dtype != torch.float32 or not spec.is_contiguous():
        raise ValueError("spec must be contiguous float32")
    F = int(n_frames) if n_frames is not None else ld
    if not 1 <= F <= ld:
        raise ValueError(f"n_frames must be in [1, {ld}]")
    out = torch.empty((B, K, F), dtype=torch.float32, device=spec.device)
    ws = torch.empty(max(4 * B, 16), dtype=torch.uint8, device=spec.device)
    with L.device_of(spec):
        L.check(L.lib().tac_amplitude_to_db(spec.data_ptr(), B, row_stride, K, F, ld, float(amin), float(top_db), out.data_ptr(),
                                            ws.data_ptr(), 4 * B, L.stream_of(spec)))
    return out


# ------------------------------------------------------------------------------------------------
# inverse STFT: batch_reconstruct_waveform, train_pkl_file_task_aware.py:174-208
# ------------------------------------------------------------------------------------------------
_ENV_CACHE: Dict[tuple, torch.Tensor] = {}


def _istft_envelope(n_fft: int, hop: int, win: int, F: int, center: bool, length: int, device) -> torch.Tensor:
    """Overlap-added squared synthesis window over the kept range (what torch.istft divides by); cached per shape."""
    key = (n_fft, hop, win, F, bool(center), int(length), str(device))
    env = _ENV_CACHE.get(key)
    if env is None:
        w = torch.zeros(n_fft, dtype=torch.float64)
        left = (n_fft - win) // 2
        w[left:left + win] = torch.hann_window(win, periodic=True, dtype=torch.float64)
        full = torch.zeros(n_fft + hop * (F - 1), dtype=torch.float64)
        for f in range(F):
            full[f * hop:f * hop + n_fft] += w * w
        start = n_fft // 2 if center else 0
        nat = full.numel() - start                      # torch.istft(length=) reads on into the padded tail
        env = torch.ones(length, dtype=torch.float64)
        m = min(nat, length)
        env[:m] = full[start:start + m]
        env = env.to(torch.float32).to(device)
        _ENV_CACHE[key] = env
    return env


def _istft_call(a: torch.Tensor, b: Optional[torch.Tensor], in_mode: int, rows: int, row_stride: int, ld: int, F: int,
                n_fft: int, hop: int, win: int, center: bool, normalized: bool, length: int) -> torch.Tensor:
    out = torch.empty((rows, length), dtype=torch.float32, device=a.device)
    with L.device_of(a):
        L.check(L.lib().tac_istft(a.data_ptr(), b.data_ptr() if b is not None else None, in_mode, rows, row_stride, ld, F,
                                  n_fft, hop, win, int(center), int(normalized), out.data_ptr(), length, L.stream_of(a)))
    return out


class _ISTFTPolar(torch.autograd.Function):
    """waveform = istft(mag * exp(1j * phase)).  Backward: the adjoint of the overlap-add synthesis is an analysis STFT
    of grad / envelope (K1 kernel, ``center=False`` on the re-padded signal), scaled by the one-sided weights c_k / n,
    followed by the polar chain rule."""

    @staticmethod
    def forward(ctx, mag, phase, n_fft, hop, win, center, normalized, length):
        B, K, F = mag.shape
        mag_c, ph_c = mag.contiguous().float(), phase.contiguous().float()
        ctx.save_for_backward(mag_c, ph_c)
        ctx.cfg = (n_fft, hop, win, center, normalized, length)
        return _istft_call(mag_c, ph_c, L.IN_POLAR, B, K * F, F, F, n_fft, hop, win, center, normalized, length)

    @staticmethod
    def backward(ctx, g):
        mag, ph = ctx.saved_tensors
        n_fft, hop, win, center, normalized, length = ctx.cfg
        G = _istft_adjoint(g.contiguous().float(), mag.shape[2], n_fft, hop, win, center, normalized, length)   # [B, K, F] complex
        c, s = torch.cos(ph), torch.sin(ph)
        gm = G.real * c + G.imag * s
        gp = mag * (G.imag * c - G.real * s)
        return gm, gp, None, None, None, None, None, None


class _ISTFTComplex(torch.autograd.Function):
    @staticmethod
    def forward(ctx, spec, n_fft, hop, win, center, normalized, length):
        B, K, F = spec.shape
        spec_c = spec.contiguous()
        ctx.cfg = (n_fft, hop, win, center, normalized, length, F)
        return _istft_call(torch.view_as_real(spec_c), None, L.IN_COMPLEX, B, K * F, F, F, n_fft, hop, win, center,
                           normalized, length)

    @staticmethod
    def backward(ctx, g):
        n_fft, hop, win, center, normalized, length, F = ctx.cfg
        return _istft_adjoint(g.contiguous().float(), F, n_fft, hop, win, center, normalized, length), None, None, None, None, None, None


def _istft_adjoint(g: torch.Tensor, F: int, n_fft: int, hop: int, win: int, center: bool, normalized: bool,
                   length: int) -> torch.Tensor:
    """dL/dX [B, K, F] (complex64) from dL/dwaveform [B, length]."""
    B = g.shape[0]
    env = _istft_envelope(n_fft, hop, win, F, center, length, g.device)
    start = n_fft // 2 if center else 0
    full = n_fft + hop * (F - 1)
    nat = min(full - start, length)
    gp = torch.zeros((B, 1, full), dtype=torch.float32, device=g.device)
    gp[:, 0, start:start + nat] = g[:, :nat] / env[:nat]
    G = stft_multi(gp, [n_fft], [hop], [win], center=False, normalized=False, out_mode=L.OUT_COMPLEX)[0][:, 0]   # [B, K, F]
    ck = torch.full((n_fft // 2 + 1, 1), 2.0 / n_fft, dtype=torch.float32, device=g.device)
    ck[0] = ck[-1] = 1.0 / n_fft
    if normalized:
        ck = ck * float(n_fft) ** 0.5
    G = G * ck
    # the imaginary parts of X[0] and X[n/2] do not reach the waveform
    G.imag[:, 0] = 0.0
    G.imag[:, -1] = 0.0
    return G


def istft(spec: torch.Tensor, n_fft: int, hop_length: Optional[int] = None, win_length: Optional[int] = None,
          window: Optional[torch.Tensor] = None, center: bool = True, normalized: bool = False, onesided: Optional[bool] = None,
          length: Optional[int] = None, return_complex: bool = False) -> torch.Tensor:
    """``torch.istft`` for a batch: ``spec`` complex64 ``[B, n_fft/2+1, F]`` (or ``[K, F]``) -> ``[B, length]`` in ONE
    launch of ``tac_istft``.  Same keyword arguments as ``torch.istft``; implemented: periodic Hann window (``window``
    must be ``None`` or equal ``torch.hann_window(win_length)``), ``onesided`` in (None, True), real output."""
    L.require_cuda(spec, "spec")
    if not spec.is_complex():
        raise RuntimeError("istft requires a complex-valued input tensor")          # torch.istft's own error class
    if onesided is False or return_complex:
        raise ValueError("only onesided=True, return_complex=False is implemented")
    single = spec.dim() == 2
    if single:
        spec = spec[None]
    if spec.dim() != 3 or spec.shape[1] != n_fft // 2 + 1:
        raise ValueError(f"expected spec of shape [B, {n_fft // 2 + 1}, F], got {tuple(spec.shape)}")
    hop = int(hop_length) if hop_length is not None else n_fft // 4
    win = int(win_length) if win_length is not None else n_fft
    if window is not None:
        ref_w = torch.hann_window(win, periodic=True, dtype=torch.float32, device=window.device)
        if window.numel() != win or not torch.allclose(window.float(), ref_w, atol=1e-6):
            raise ValueError("Unsupported window: only the periodic torch.hann_window is implemented")
    F = spec.shape[2]
    nat = int(L.lib().tac_istft_length(F, n_fft, hop, int(center)))
    out_len = int(length) if length is not None else nat
    y = _ISTFTComplex.apply(spec.to(torch.complex64), n_fft, hop, win, bool(center), bool(normalized), out_len)
    return y[0] if single else y


def batch_reconstruct_waveform(magnitude_batch: torch.Tensor, phase_batch: torch.Tensor, sample_rate: int, n_fft: int = 2048,
                               hop_length: int = 512, win_length: int = 2048, window: str = "hann") -> torch.Tensor:
    """Same name, arguments and errors as ``train_pkl_file_task_aware.py:174-208``: ``[B, K, F]`` magnitude and phase
    -> ``[B, hop*(F-1)]`` waveforms; the complex product, the per-item ``torch.istft`` loop and the stack are one kernel.
    Strided views (e.g. the two channels of the ``[B, 2, 1025, 600]`` decoder output) are used in place."""
    if magnitude_batch.shape != phase_batch.shape:
        raise ValueError("The shapes of magnitude and phase must match for batched processing.")
    if window != "hann":
        raise ValueError(f"Window type '{window}' is not supported. Use 'hann'.")
    L.require_cuda(magnitude_batch, "magnitude_batch")
    L.require_cuda(phase_batch, "phase_batch")
    if magnitude_batch.dim() != 3 or magnitude_batch.shape[1] != n_fft // 2 + 1:
        raise ValueError(f"expected [B, {n_fft // 2 + 1}, F], got {tuple(magnitude_batch.shape)}")
    B, K, F = magnitude_batch.shape
    nat = int(L.lib().tac_istft_length(F, n_fft, hop_length, 1))
    needs_grad = torch.is_grad_enabled() and (magnitude_batch.requires_grad or phase_batch.requires_grad)
    m, p = magnitude_batch, phase_batch
    in_place = (not needs_grad and m.dtype == torch.float32 and p.dtype == torch.float32 and m.stride() == p.stride()
                and m.stride(2) == 1 and m.stride(1) >= F and m.stride(0) >= K * m.stride(1))
    if in_place:
        return _istft_call(m, p, L.IN_POLAR, B, m.stride(0), m.stride(1), F, n_fft, hop_length, win_length, True, False, nat)
    return _ISTFTPolar.apply(m, p, n_fft, hop_length, win_length, True, False, nat)


def reconstruct_waveform(magnitude: torch.Tensor, phase: torch.Tensor, sample_rate: int, n_fft: int = 2048,
                         hop_length: int = 512, win_length: int = 2048, window: str = "hann") -> torch.Tensor:
    """Single-utterance form (``train_pkl_file_task_aware.py:210-226``): ``[K, F]`` -> ``[1, hop*(F-1)]`` (the reference
    returns ``y.unsqueeze(0)``)."""
    return batch_reconstruct_waveform(magnitude[None], phase[None], sample_rate, n_fft, hop_length, win_length, window)


def reconstruct_and_disc_inputs(magnitude_batch: torch.Tensor, phase_batch: torch.Tensor, sample_rate: int = 16000,
                                n_fft: int = 2048, hop_length: int = 512, win_length: int = 2048, window: str = "hann",
                                disc_n_ffts: Sequence[int] = (1024, 2048, 512), disc_hop_lengths: Sequence[int] = (256, 512, 128),
                                disc_win_lengths: Sequence[int] = (1024, 2048, 512), disc_normalized: bool = True,
                                rows_per_group: int = 0):
    """The ground-truth branch of ``task_aware()`` as one call: ``batch_reconstruct_waveform(clean_mag, clean_phase)``
    (``train_pkl_file_task_aware.py:290``), ``.unsqueeze(1)`` (``:347``) and the spectrogram front of every scale
    discriminator that ``disc(ground_truth_clean_waveforms)`` (``:353``) runs first (``msstftd.py:188-191`` ->
    ``:151-153``).  Returns ``(waveform [B, hop*(F-1)], [z_s [B, 2, F_s, K_s] per scale])``; the waveform is what ``:333``'s
    MSE consumes, ``z_s`` is what ``DiscriminatorSTFT.forward`` feeds its first conv.  One C-ABI call enqueues both kernels
    (``tac_istft_disc_inputs``); ``rows_per_group`` > 0 walks the batch in row groups (``tac_chain_rows_per_group`` sizes them
    so that a group's waveform stays in the L2 between synthesis and analysis; measured slower than one group on B200, see
    ``csrc/chain.cu``).  Results are bit-identical to ``batch_reconstruct_waveform`` + ``MultiScaleSTFT``.  No autograd:
    the reference detaches this branch (``torch.tensor(reconstructed_clean_waveforms)``, ``:303``)."""
    if magnitude_batch.shape != phase_batch.shape:
        raise ValueError("The shapes of magnitude and phase must match for batched processing.")
    if window != "hann":
        raise ValueError(f"Window type '{window}' is not supported. Use 'hann'.")
    L.require_cuda(magnitude_batch, "magnitude_batch")
    L.require_cuda(phase_batch, "phase_batch")
    if magnitude_batch.dim() != 3 or magnitude_batch.shape[1] != n_fft // 2 + 1:
        raise ValueError(f"expected [B, {n_fft // 2 + 1}, F], got {tuple(magnitude_batch.shape)}")
    if not (len(disc_n_ffts) == len(disc_hop_lengths) == len(disc_win_lengths)) or not 1 <= len(disc_n_ffts) <= L.MAX_SCALES:
        raise ValueError("disc_n_ffts, disc_hop_lengths, disc_win_lengths must have the same length in [1, 8]")
    m, p = magnitude_batch.detach(), phase_batch.detach()
    B, K, F = m.shape
    usable = (m.dtype == torch.float32 and p.dtype == torch.float32 and m.stride() == p.stride() and m.stride(2) == 1
              and m.stride(1) >= F and m.stride(0) >= K * m.stride(1))
    if not usable:
        m, p = m.float().contiguous(), p.float().contiguous()
    T = int(L.lib().tac_istft_length(F, n_fft, hop_length, 1))
    wave = torch.empty((B, T), dtype=torch.float32, device=m.device)
    outs = []
    for n, h in zip(disc_n_ffts, disc_hop_lengths):
        Fs = num_frames(T, n, h, False)
        if Fs < 1:
            raise ValueError(f"signal of {T} samples is shorter than n_fft={n}")
        outs.append(torch.empty((B, 2, Fs, n // 2 + 1), dtype=torch.float32, device=m.device))
    with L.device_of(m):
        L.check(L.lib().tac_istft_disc_inputs(m.data_ptr(), p.data_ptr(), L.IN_POLAR, B, m.stride(0), m.stride(1), F, n_fft,
                                              hop_length, win_length, 1, 0, wave.data_ptr(), T, len(disc_n_ffts),
                                              L.int_array(disc_n_ffts), L.int_array(disc_hop_lengths),
                                              L.int_array(disc_win_lengths), int(bool(disc_normalized)), L.ptr_array(outs),
                                              int(rows_per_group), L.stream_of(m)))
    return wave, outs
