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


def stft_multi(x: torch.Tensor, n_ffts: Sequence[int], hops: Sequence[int], wins: Optional[Sequence[int]] = None, *,
               center: bool = False, pad_mode: Optional[str] = "reflect", normalized: bool = True,
               out_mode: int = L.OUT_PLANES, f_pad: int = 0) -> List[torch.Tensor]:
    """All scales of ``x [B, C, T]`` in one call of ``tac_stft_multi``; see ``include/tac_spectral.h``."""
    L.require_cuda(x, "x")
    if x.dim() != 3:
        raise ValueError(f"expected x of shape [B, C, T], got {tuple(x.shape)}")
    wins = list(n_ffts) if wins is None else list(wins)
    assert len(n_ffts) == len(hops) == len(wins)                       # msstftd.py:177
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
        if self.n_fft not in SUPPORTED_NFFT:
            raise ValueError(f"n_fft={n_fft} unsupported; choose one of {SUPPORTED_NFFT}")
        if power is not None:
            raise NotImplementedError("only power=None (complex spectrogram, msstftd.py:127) is implemented")
        if pad != 0 or not onesided:
            raise NotImplementedError("pad != 0 and onesided=False are not implemented")
        if normalized not in (False, True, "window"):
            raise NotImplementedError("normalized must be False, True or 'window'")
        self.normalized = bool(normalized)
        self.center = bool(center)
        self.pad_mode = pad_mode
        self.power = None
        # torchaudio keeps the window as a buffer; kept so that state_dict()s stay interchangeable
        self.register_buffer("window", torch.hann_window(self.win_length), persistent=False)
        if torch.cuda.is_available():
            L.check(L.lib().tac_prepare_fft(self.n_fft, self.win_length, int(self.normalized)))

    def _run(self, x: torch.Tensor, out_mode: int) -> torch.Tensor:
        lead = x.shape[:-1]
        x3 = x.reshape(-1, 1, x.shape[-1]) if x.dim() != 3 else x
        out = stft_multi(x3, [self.n_fft], [self.hop_length], [self.win_length], center=self.center,
                         pad_mode=self.pad_mode, normalized=self.normalized, out_mode=out_mode)[0]
        return out, lead

    def forward(self, waveform: torch.Tensor) -> torch.Tensor:
        """``[..., T]`` -> complex64 ``[..., K, F]`` (what ``msstftd.py:151`` receives)."""
        out, lead = self._run(waveform, L.OUT_COMPLEX)
        return out.reshape(*lead, out.shape[-2], out.shape[-1])

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


def complex_spectrogram(y: torch.Tensor, n_fft: int = 2048, hop_length: int = 512, win_length: int = 2048,
                        pad_mode: str = "constant", f_pad: Optional[int] = None) -> torch.Tensor:
    """Batched notebook front end: ``y [B, T]`` -> ``[B, 2, n_fft/2+1, F]`` with channel 0 =
    ``np.abs(D)``, channel 1 = ``np.angle(D)``, ``D = librosa.stft(y, n_fft, hop, win, 'hann')``
    (centred, un-normalised; librosa 0.10 pads with zeros).  With ``f_pad`` the time axis is
    zero-padded / cropped to ``f_pad`` frames -- the reference's ``[B, 2, 1025, 600]`` model input."""
    L.require_cuda(y, "y")
    if y.dim() != 2:
        raise ValueError(f"expected y of shape [B, T], got {tuple(y.shape)}")
    return stft_multi(y[:, None, :], [n_fft], [hop_length], [win_length], center=True, pad_mode=pad_mode,
                      normalized=False, out_mode=L.OUT_MAGPHASE, f_pad=int(f_pad or 0))[0]


def create_complex_spectrogram(y: torch.Tensor, sr: int, n_fft: int = 2048, hop_length: int = 512,
                               win_length: int = 2048) -> Dict[str, object]:
    """Same name, argument meaning and result keys as the notebook function (cell 10 L26-52):
    ``{'magnitude', 'phase', 'params'}`` for one utterance ``y [T]`` (or a batch ``[B, T]``)."""
    single = y.dim() == 1
    out = complex_spectrogram(y[None] if single else y, n_fft, hop_length, win_length)
    mag, ph = out[:, 0], out[:, 1]
    if single:
        mag, ph = mag[0], ph[0]
    return {"magnitude": mag, "phase": ph,
            "params": {"sr": sr, "n_fft": n_fft, "hop_length": hop_length, "win_length": win_length, "window": "hann"}}


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
