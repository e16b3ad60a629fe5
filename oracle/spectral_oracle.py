"""CPU oracle (checker) for the spectral hot path.  TEST INFRASTRUCTURE, NOT PRODUCT.

A CPU restatement, in torch-CPU / numpy, of what the reference computes on the hot path that
``BASELINE.json:north_star`` names.  The reference does its arithmetic by calling third-party
library routines (``torchaudio.transforms.Spectrogram`` -> ``torch.stft``; ``librosa.stft``;
``torch.svd``; ``torch.matmul``), unpinned at top level (``requirements.txt:5,11``: ``torch>=1.12.1``,
``torchaudio``), so each function below restates the *call* the reference makes, with the
``file:line`` it follows.  All paths are relative to ``/root/reference``.

Pinning status
--------------
* The reference's own tests hold **shape** known-answers only (``testing_scripts.ipynb:2021-2023``,
  ``spectrogram.ipynb:4210-4211``); those are asserted in ``tests/test_oracle.py``.
* Numeric parity is pinned against **outputs of the reference itself run in the build container**:
  ``tests/golden/make_golden.py`` imports ``msstftd``, ``pkl_file_model_updated`` and
  ``pkl_file_models`` from ``/root/reference`` and stores their outputs on seeded inputs in
  ``tests/golden/*.npz``; ``tests/test_oracle.py`` checks every function here against them.
* Two things have **no reference implementation** and are therefore "parity unpinned" by the
  reference (definition of record = SURVEY.md section 8c): the L1 + log-magnitude multi-resolution
  STFT loss (``mrstft_loss``) and the ``librosa.stft`` front end (librosa is not installed here;
  restated through ``torch.stft`` with librosa 0.10.1's defaults and cross-checked against a
  float64 direct DFT).

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline legs may import this.
"""
from __future__ import annotations

import math
from typing import Dict, List, Sequence, Tuple

import numpy as np
import torch

LOG_FLOOR = 1e-5  # librosa ``amin`` -- the only log floor the reference's dB channel uses


# --------------------------------------------------------------------------------------------
# STFT conventions
# --------------------------------------------------------------------------------------------

def num_frames(T: int, n_fft: int, hop: int, center: bool) -> int:
    """Frame count of torch.stft: ``1 + (T_padded - n_fft) // hop``."""
    Tp = T + (2 * (n_fft // 2) if center else 0)
    if Tp < n_fft:
        return 0
    return 1 + (Tp - n_fft) // hop


def hann_periodic(n: int, dtype=torch.float32) -> torch.Tensor:
    """``torch.hann_window(n)`` (periodic=True), the ``window_fn`` of ``msstftd.py:126``."""
    return torch.hann_window(n, periodic=True, dtype=dtype)


def stft_ref(x: torch.Tensor, n_fft: int, hop: int, win: int | None = None, *, center: bool = False,
             pad_mode: str = "reflect", normalized: bool = True) -> torch.Tensor:
    """Complex STFT ``[B, K, F]`` as ``torchaudio.transforms.Spectrogram(power=None)`` computes it.

    Follows ``msstftd.py:125-127`` (constructor arguments) and ``msstftd.py:151`` (the call):
    one-sided ``torch.stft`` with a periodic Hann window of ``win`` samples, and, for
    ``normalized=True``, division by ``sqrt(sum(window**2))`` ("window" normalisation of
    ``torchaudio.functional.spectrogram``), *not* torch.stft's own ``normalized`` flag.
    """
    win = n_fft if win is None else win
    if x.dim() != 2:
        raise ValueError("stft_ref expects [B, T]")
    w = hann_periodic(win, x.dtype)
    if center:
        p = n_fft // 2
        mode = "constant" if pad_mode in ("constant", "zeros") else pad_mode
        x = torch.nn.functional.pad(x[:, None, :], (p, p), mode=mode)[:, 0, :]
    z = torch.stft(x, n_fft, hop_length=hop, win_length=win, window=w, center=False,
                   normalized=False, onesided=True, return_complex=True)
    if normalized:
        z = z / w.pow(2.0).sum().sqrt()
    return z


def disc_inputs_ref(x: torch.Tensor, n_fft: int, hop: int, win: int | None = None,
                    normalized: bool = True) -> torch.Tensor:
    """Discriminator input ``[B, 2C, F, K]`` of one scale: ``msstftd.py:151-153``.

    ``x`` is ``[B, C, T]``; the complex spectrogram ``[B, C, K, F]`` is split into real and
    imaginary parts concatenated on the channel axis (all real channels, then all imaginary
    ones) and transposed to time-major / frequency-innermost.
    """
    B, C, T = x.shape
    z = stft_ref(x.reshape(B * C, T), n_fft, hop, win, center=False, normalized=normalized)
    z = z.reshape(B, C, z.shape[-2], z.shape[-1])
    planes = torch.cat([z.real, z.imag], dim=1)          # [B, 2C, K, F]
    return planes.transpose(2, 3).contiguous()           # [B, 2C, F, K]


def multiscale_disc_inputs_ref(x: torch.Tensor, n_ffts: Sequence[int] = (1024, 2048, 512),
                               hops: Sequence[int] = (256, 512, 128),
                               wins: Sequence[int] | None = None) -> List[torch.Tensor]:
    """The per-scale loop of ``MultiScaleSTFTDiscriminator.forward`` (``msstftd.py:185-192``),
    front end only; default scale lists are ``msstftd.py:174-175``."""
    wins = n_ffts if wins is None else wins
    assert len(n_ffts) == len(hops) == len(wins)          # msstftd.py:177
    return [disc_inputs_ref(x, n, h, w) for n, h, w in zip(n_ffts, hops, wins)]


def dft_frames_f64(x: np.ndarray, n_fft: int, hop: int, normalized: bool = True) -> np.ndarray:
    """Direct O(n^2) float64 DFT of Hann-windowed frames, ``[B, K, F]`` complex128.

    Independent of torch.stft -- used to check ``stft_ref`` itself on small cases.
    """
    x = np.asarray(x, dtype=np.float64)
    B, T = x.shape
    F = num_frames(T, n_fft, hop, False)
    i = np.arange(n_fft)
    w = 0.5 - 0.5 * np.cos(2.0 * np.pi * i / n_fft)
    K = n_fft // 2 + 1
    basis = np.exp(-2j * np.pi * np.outer(np.arange(K), i) / n_fft)     # [K, n]
    out = np.empty((B, K, F), dtype=np.complex128)
    for f in range(F):
        seg = x[:, f * hop:f * hop + n_fft] * w
        out[:, :, f] = seg @ basis.T
    if normalized:
        out /= math.sqrt(float(np.sum(w * w)))
    return out


# --------------------------------------------------------------------------------------------
# Front end (offline spectrogram writer of the notebooks)
# --------------------------------------------------------------------------------------------

def frontend_ref(y: torch.Tensor, n_fft: int = 2048, hop: int = 512, win: int = 2048,
                 pad_mode: str = "constant", f_pad: int | None = None) -> Tuple[torch.Tensor, torch.Tensor]:
    """``create_complex_spectrogram`` of ``spectrogram.ipynb`` cell 10 L26-52 (raw ``:4120-4134``).

    ``librosa.stft(y, n_fft=2048, hop_length=512, win_length=2048, window='hann')`` is a centred,
    un-normalised STFT with a periodic Hann window; librosa 0.10.1 (the version the repo pins in
    ``sgmse/requirements_version.txt:3``) pads with zeros (``pad_mode='constant'``), older
    releases reflect.  Returns ``(magnitude, phase)`` = ``(np.abs(D), np.angle(D))``, each
    ``[B, 1025, F]``; with ``f_pad`` they are zero-padded / cropped along time to ``f_pad`` frames
    like ``SpectralDataset.pad_tensor`` (``train_pkl_file_task_aware.py:64-68``).
    """
    D = stft_ref(y, n_fft, hop, win, center=True, pad_mode=pad_mode, normalized=False)
    mag, ph = D.abs(), D.angle()
    if f_pad is not None:
        mag, ph = pad_frames(mag, f_pad), pad_frames(ph, f_pad)
    return mag, ph


def pad_frames(t: torch.Tensor, f_pad: int) -> torch.Tensor:
    """Zero-pad (or crop, when negative) the last axis to ``f_pad`` -- ``pad_tensor`` semantics."""
    F = t.shape[-1]
    if F >= f_pad:
        return t[..., :f_pad].contiguous()
    return torch.nn.functional.pad(t, (0, f_pad - F))


def stack_mag_phase_ref(mag: torch.Tensor, ph: torch.Tensor) -> torch.Tensor:
    """``torch.stack([magnitude, phase], dim=1)`` -> ``[B, 2, K, F]`` (``pkl_file_models.py:889-892``)."""
    return torch.stack([mag, ph], dim=1).float()


def amplitude_to_db_ref(mag: torch.Tensor, amin: float = 1e-5, top_db: float = 80.0) -> torch.Tensor:
    """``librosa.amplitude_to_db(S, ref=np.max)`` per utterance (notebook cell 9 L31-63, ``:2894-2913``):
    ``20*log10(max(S, amin)) - 20*log10(max(amin, max S))`` floored at ``max(out) - top_db``."""
    ref = mag.amax(dim=(-2, -1), keepdim=True)
    db = 20.0 * torch.log10(torch.clamp(mag, min=amin)) - 20.0 * torch.log10(torch.clamp(ref, min=amin))
    return torch.maximum(db, db.amax(dim=(-2, -1), keepdim=True) - top_db)


def torch_stft_ref(x: torch.Tensor, n_fft: int, hop: int, win: int | None = None, *, window: str = "hann", center: bool = True,
                   pad_mode: str = "reflect", normalized: bool = False) -> torch.Tensor:
    """``torch.stft`` for ANY ``n_fft`` -- the library call behind ``librosa.stft(n_fft=2046, ...)`` (``spectrogram.ipynb`` cell 9,
    raw ``:2897-2906``), ``SpecsDataModule.stft`` (``sgmse/sgmse/data_module.py:212-214``; window per ``:13-17``) and
    ``torchaudio.transforms.Spectrogram()`` (``testing_scripts.ipynb:246``).  ``normalized`` is torchaudio's 'window' form."""
    win = n_fft if win is None else win
    w = hann_periodic(win)
    if window == "sqrthann":
        w = torch.sqrt(w)
    lead = x.shape[:-1]
    out = torch.stft(x.reshape(-1, x.shape[-1]).float(), n_fft, hop, win, w, center=center, pad_mode=pad_mode, normalized=False,
                     onesided=True, return_complex=True)
    if normalized:
        out = out / w.pow(2).sum().sqrt()
    return out.reshape(*lead, out.shape[-2], out.shape[-1])


def sgmse_spec_ref(sig: torch.Tensor, n_fft: int = 510, hop: int = 128, window: str = "hann", spec_factor: float = 0.15,
                   spec_abs_exponent: float = 0.5, pad64: bool = True) -> torch.Tensor:
    """``pad_spec(spec_fwd(stft(sig)))`` of the SGMSE+ enhancer: ``sgmse/sgmse/data_module.py:212-214`` (centred reflect-padded
    ``torch.stft``), ``:162-169`` (``spec.abs()**e * exp(1j * spec.angle())`` then ``* spec_factor``; defaults ``:109-118``) and
    ``sgmse/sgmse/util/other.py:76-90`` (zero-pad the frame axis to a multiple of 64)."""
    spec = torch_stft_ref(sig, n_fft, hop, n_fft, window=window, center=True, pad_mode="reflect")
    if spec_abs_exponent != 1:
        spec = spec.abs() ** spec_abs_exponent * torch.exp(1j * spec.angle())
    spec = spec * spec_factor
    if pad64:
        F = spec.shape[-1]
        if F % 64 != 0:
            spec = torch.nn.functional.pad(spec, (0, 64 - F % 64))
    return spec


def power_spectrogram_ref(x: torch.Tensor, log: bool = False, **kwargs) -> torch.Tensor:
    """``torchaudio.transforms.Spectrogram(**kwargs)(x)`` (``testing_scripts.ipynb:246``: all defaults -- n_fft 400, hop 200,
    power 2, centred, reflect); ``log``: ``10 * torch.log10(S + 1e-10)`` (``:249``)."""
    import torchaudio
    S = torchaudio.transforms.Spectrogram(**kwargs)(x.float())
    return 10.0 * torch.log10(S + 1e-10) if log else S


def mel_spectrogram_ref(x: torch.Tensor, n_fft: int, hop: int, n_mels: int = 80, f_min: float = 0.0, f_max: float = 8000.0,
                        sample_rate: int = 16000, power: float = 1.0, center: bool = False, normalized: bool = True) -> torch.Tensor:
    """SURVEY.md 8c item 5 (no mel path exists in the reference; definition of record): the magnitude spectrogram of the
    ``msstftd.py:125-127`` convention times ``torchaudio.functional.melscale_fbanks(K, f_min, f_max, n_mels, sample_rate,
    norm=None, mel_scale='htk')``."""
    import torchaudio
    S = torch_stft_ref(x, n_fft, hop, n_fft, center=center, normalized=normalized).abs() ** power          # [..., K, F]
    fb = torchaudio.functional.melscale_fbanks(n_fft // 2 + 1, f_min, f_max, n_mels, sample_rate, norm=None, mel_scale="htk")
    return torch.matmul(S.transpose(-1, -2), fb).transpose(-1, -2)


def amplitude_to_db_fixed_ref(mag: torch.Tensor, ref: float = 1.0, amin: float = 1e-5, floor_db: float = -80.0) -> torch.Tensor:
    """``20 log10(max(S, amin)) - 20 log10(max(amin, ref))`` floored at ``floor_db`` (``librosa.amplitude_to_db`` with a
    fixed reference; with ``ref`` = the array maximum this is the ``ref=np.max`` form of the notebook)."""
    db = 20.0 * torch.log10(torch.clamp(mag, min=amin)) - 20.0 * math.log10(max(amin, ref))
    return torch.clamp(db, min=floor_db)


# --------------------------------------------------------------------------------------------
# Losses
# --------------------------------------------------------------------------------------------

def mrstft_loss_ref(x: torch.Tensor, y: torch.Tensor, n_ffts: Sequence[int] = (1024, 2048, 512),
                    hops: Sequence[int] | None = None, eps: float = LOG_FLOOR,
                    return_terms: bool = False):
    """Multi-resolution STFT loss -- DEFINITION OF RECORD.

    The reference holds no such function ("parity unpinned" by the reference); it is defined on
    the reference's own STFT convention (``msstftd.py:125-127``: periodic Hann, ``win = n_fft``,
    ``center=False``, ``normalized=True``):

        X = |STFT_s(x)|, Y = |STFT_s(y)|
        L1_s   = mean |X - Y|
        Llog_s = mean |ln max(X, eps) - ln max(Y, eps)|
        L      = (1/S) * sum_s (L1_s + Llog_s)

    means run over all ``B*K_s*F_s`` elements.  ``x`` is the target, ``y`` the decoded signal
    (the one gradients flow to).  Differentiable (torch autograd) -- the backward oracle.
    """
    hops = [n // 4 for n in n_ffts] if hops is None else list(hops)
    total = x.new_zeros(())
    terms = []
    for n, h in zip(n_ffts, hops):
        X = stft_ref(x, n, h, n).abs()
        Y = stft_ref(y, n, h, n).abs()
        l1 = (X - Y).abs().mean()
        ll = (torch.log(torch.clamp(X, min=eps)) - torch.log(torch.clamp(Y, min=eps))).abs().mean()
        terms.append((l1, ll))
        total = total + l1 + ll
    total = total / len(n_ffts)
    return (total, terms) if return_terms else total


def kink_sample_mask(x: torch.Tensor, y: torch.Tensor, n_ffts: Sequence[int] = (1024, 2048, 512), eps: float = LOG_FLOOR,
                     rel: float = 1e-6, cond: float = 5e-4) -> torch.Tensor:
    """Test helper for the gradient of :func:`mrstft_loss_ref` (definition of record, not in the reference): which samples of
    ``y`` [B, T] can an fp32 implementation NOT be held to rtol 1e-4 on?  Evaluated in float64:

    * kinks -- a bin with ``|X - Y| < rel * max|X|`` (the L1 and log terms switch sign there) or a magnitude within
      ``rel * max|X|`` of the log floor (the clamp switches the log term on / off).  fp32 transform rounding is
      ~1e-7 * max|X|, so ``rel = 1e-6`` leaves a 10x margin;
    * ill-conditioned bins -- unclamped ``Y < cond * max|X|``: the log term's 1 / Y turns that rounding into a relative
      error ``dY / Y >= 1e-7 / cond = 2e-4`` of the bin's gradient (torch's own fp32 path included).

    A marked bin taints the ``n_fft`` samples of its frame (hop = n_fft / 4).  True = excluded from a strict check."""
    shape = x.shape
    x2, y2 = x.reshape(-1, shape[-1]).double(), y.reshape(-1, shape[-1]).double()
    mask = torch.zeros(x2.shape, dtype=torch.bool)
    for n in n_ffts:
        h = n // 4
        X = stft_ref(x2, n, h, n).abs()
        Y = stft_ref(y2, n, h, n).abs()
        m = float(X.max())
        t = rel * m
        bins = ((X - Y).abs() < t) | ((X - eps).abs() < t) | ((Y - eps).abs() < t) | ((Y > eps) & (Y < cond * m))
        for b, f in torch.nonzero(bins.any(dim=-2)).tolist():         # [rows, F]
            mask[b, f * h:f * h + n] = True
    return mask.reshape(shape)


def spectral_losses_ref(obs: torch.Tensor, dec: torch.Tensor) -> Dict[str, torch.Tensor]:
    """The reference's own spectral reductions on ``[B, 2, K, F]`` stacks.

    ``pkl_file_models.py:924`` (per-sample ``0.5*mean sq``), ``:933-937`` (magnitude / phase / total
    MSE), ``:940`` (spectral SNR); identical block in ``pkl_file_model_updated.py:520-540``.
    """
    d = obs - dec
    sq = d * d
    return {
        "mse_per_sample": 0.5 * sq.mean(dim=(1, 2, 3)),
        "magnitude_loss": sq[:, 0].mean(),
        "phase_loss": sq[:, 1].mean(),
        "total_loss": sq.mean(),
        "spec_snr": -10.0 * torch.log10(sq.mean() / (obs * obs).mean()),
    }


# --------------------------------------------------------------------------------------------
# DPCA
# --------------------------------------------------------------------------------------------

def data_pca_ref(z: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """``data_pca`` (``pkl_file_model_updated.py:37-42``; same at ``pkl_file_models.py:51-56``):
    column mean, thin SVD of the centred batch; returns ``(s[r], v[d, r], mean[d])`` with
    ``r = min(B, d)`` and ``s`` descending.  ``s`` are singular values of ``Z - mean`` (no 1/(B-1))."""
    mu = z.mean(dim=0)
    _, s, vh = torch.linalg.svd(z - mu, full_matrices=False)
    return s, vh.transpose(0, 1), mu


def dpca_select_ref(s_list: Sequence[torch.Tensor], dim_p: int) -> List[torch.Tensor]:
    """Global top-``dim_p`` over the concatenated per-view singular values, split back per view
    (``pkl_file_model_updated.py:496-500`` for two views, ``:404-412`` for four).  Each returned
    index vector keeps the order in which its entries appear in the global descending sort."""
    cat = torch.cat(list(s_list), 0)
    top = torch.argsort(cat, descending=True)[:dim_p]
    out, lo = [], 0
    for s in s_list:
        hi = lo + s.shape[0]
        out.append(top[(top >= lo) & (top < hi)] - lo)
        lo = hi
    return out


def dpca_project_ref(zs: Sequence[torch.Tensor], dim_p: int) -> Tuple[List[torch.Tensor], List[torch.Tensor]]:
    """The DPCA block as a function: ``pkl_file_model_updated.py:489-511`` (E2D1) and ``:398-427``
    (E4D1) with the host draw of ``dim_p`` (``:489`` / ``:398``) left to the caller.
    Returns ``(z_out per view, selected indices per view)``."""
    pcs = [data_pca_ref(z) for z in zs]
    idx = dpca_select_ref([p[0] for p in pcs], dim_p)
    outs = []
    for z, (s, v, mu), ind in zip(zs, pcs, idx):
        basis = v[:, ind]
        outs.append(torch.matmul(torch.matmul(z - mu, basis), basis.transpose(0, 1)) + mu)
    return outs, idx


def apply_pca_ref(z: torch.Tensor, dim_p: int) -> torch.Tensor:
    """``apply_pca`` (``pkl_file_models.py:224-228``): single-view rank-``dim_p`` reconstruction."""
    s, v, mu = data_pca_ref(z)
    b = v[:, :dim_p]
    return torch.matmul(torch.matmul(z - mu, b), b.transpose(0, 1)) + mu


def var_topk_ref(z: torch.Tensor, k: int) -> Tuple[torch.Tensor, torch.Tensor]:
    """E1D1 bottleneck (``pkl_file_models.py:901-907``): unbiased per-feature variance over the
    batch, indices of the ``k`` largest, gathered columns."""
    idx = torch.topk(z.var(dim=0), k).indices
    return z[:, idx], idx


def nuclear_loss_ref(z_sample: torch.Tensor) -> torch.Tensor:
    """Latent regulariser (``pkl_file_model_updated.py:526-532``, ``pkl_file_models.py:943-947``):
    centre, divide by the Frobenius norm, nuclear norm / batch size."""
    zc = z_sample - z_sample.mean(dim=0)
    zc = zc / torch.norm(zc, p=2)
    return torch.linalg.svdvals(zc).sum() / z_sample.shape[0]


def pairwise_cosine_loss_ref(zs: Sequence[torch.Tensor]) -> torch.Tensor:
    """``cos_sim = torch.nn.CosineSimilarity(); cos_loss = torch.mean(cos_sim(z1, z2) + cos_sim(z1, z3) + ...)`` over all
    view pairs (``pkl_file_models.py:390-391, 761-763``; ``pkl_file_model_updated.py:429-430, 513-514``)."""
    cos_sim = torch.nn.CosineSimilarity()
    total = None
    for i in range(len(zs)):
        for j in range(i + 1, len(zs)):
            c = cos_sim(zs[i], zs[j])
            total = c if total is None else total + c
    return torch.mean(total)


# --------------------------------------------------------------------------------------------
# Inverse STFT
# --------------------------------------------------------------------------------------------

def istft_ref(spec: torch.Tensor, n_fft: int, hop: int, win: int | None = None, *, center: bool = True,
              normalized: bool = False, length: int | None = None) -> torch.Tensor:
    """``torch.istft`` with the periodic Hann window the reference builds (``torch.hann_window(win_length)``,
    ``train_pkl_file_task_aware.py:189``); ``spec`` complex ``[B, K, F]``."""
    win = n_fft if win is None else win
    return torch.istft(spec, n_fft=n_fft, hop_length=hop, win_length=win, window=hann_periodic(win), center=center,
                       normalized=normalized, length=length, return_complex=False)


def batch_reconstruct_waveform_ref(magnitude_batch: torch.Tensor, phase_batch: torch.Tensor, n_fft: int = 2048,
                                   hop_length: int = 512, win_length: int = 2048) -> torch.Tensor:
    """``batch_reconstruct_waveform`` (``train_pkl_file_task_aware.py:174-208``): ``mag * exp(1j * phase)``, then
    ``torch.istft`` item by item (list comprehension of ``:195-205``), then ``torch.stack`` (``:208``)."""
    if magnitude_batch.shape != phase_batch.shape:
        raise ValueError("The shapes of magnitude and phase must match for batched processing.")
    spec = magnitude_batch * torch.exp(1j * phase_batch)
    w = hann_periodic(win_length)
    items = [torch.istft(spec[i], n_fft=n_fft, hop_length=hop_length, win_length=win_length, window=w, return_complex=False)
             for i in range(magnitude_batch.shape[0])]
    return torch.stack(items, dim=0)


def reconstruct_and_disc_inputs_ref(magnitude_batch: torch.Tensor, phase_batch: torch.Tensor,
                                    n_ffts: Sequence[int] = (1024, 2048, 512), hops: Sequence[int] = (256, 512, 128),
                                    wins: Sequence[int] | None = None) -> Tuple[torch.Tensor, List[torch.Tensor]]:
    """The ground-truth branch of ``task_aware()``: ``batch_reconstruct_waveform(clean_mag, clean_phase)``
    (``train_pkl_file_task_aware.py:290``), ``.unsqueeze(1)`` (``:347``), then what ``disc(...)`` (``:353``) computes first
    for every scale (``msstftd.py:188-191`` -> ``:151-153``).  Returns (waveform ``[B, T]``, per-scale ``[B, 2, F, K]``)."""
    wav = batch_reconstruct_waveform_ref(magnitude_batch, phase_batch)
    return wav, multiscale_disc_inputs_ref(wav.unsqueeze(1), n_ffts, hops, wins)


# --------------------------------------------------------------------------------------------
# The bottleneck that is live in the SpectralRes models (row a9) and the task_aware() loss tail (row a5-i)
# --------------------------------------------------------------------------------------------

def bottleneck_mix_ref(zs: Sequence[torch.Tensor], split: int, trunc: int,
                       alphas: Sequence[torch.Tensor | None] | None = None) -> List[torch.Tensor]:
    """``pkl_file_models.py:1108-1123`` (E2D2; the same statements at ``:335-347`` E2D1 and ``:692-711`` E4D1): private /
    shared split at column ``split``, per-row alpha mix of the shared parts of the view pairs (1,2), (3,4), truncation of both
    parts to ``trunc`` columns, concatenation.  Returns one ``z_i_sample`` per view."""
    priv = [z[:, :split] for z in zs]
    share = [z[:, split:] for z in zs]
    if alphas is not None:
        for p, a in enumerate(alphas):
            if a is not None and 2 * p + 1 < len(zs):
                mixed = a.reshape(-1, 1) * share[2 * p] + (1 - a.reshape(-1, 1)) * share[2 * p + 1]
                share[2 * p] = share[2 * p + 1] = mixed
    return [torch.cat((pv[:, :trunc], sh[:, :trunc]), dim=1) for pv, sh in zip(priv, share)]


def interp_mse_ref(a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    """``train_pkl_file_task_aware.py:312-333``: the shorter of two ``[B, L]`` waveforms is resampled with
    ``F.interpolate(x.unsqueeze(1), size=longer, mode='linear', align_corners=False)``, then ``F.mse_loss``."""
    import torch.nn.functional as F
    a, b = a.float(), b.float()
    if a.shape[-1] < b.shape[-1]:
        a = F.interpolate(a.unsqueeze(1), size=b.shape[-1], mode="linear", align_corners=False).squeeze(1)
    elif b.shape[-1] < a.shape[-1]:
        b = F.interpolate(b.unsqueeze(1), size=a.shape[-1], mode="linear", align_corners=False).squeeze(1)
    return F.mse_loss(a, b, reduction="mean")


def task_aware_tail_ref(enhanced_audio: torch.Tensor, clean: torch.Tensor, disc_enhanced, disc_ground_truth):
    """``train_pkl_file_task_aware.py:333,352-361`` with the discriminator's results passed in: waveform MSE, then
    ``sum F.l1_loss(feat_enh, feat_gt)`` over all feature maps ``+ F.mse_loss(y_enh, 1) + F.mse_loss(y_gt, 0)`` per scale."""
    import torch.nn.functional as F
    mse = interp_mse_ref(enhanced_audio, clean)
    (y_e, f_e), (y_g, f_g) = disc_enhanced, disc_ground_truth
    loss = torch.zeros(())
    for fe, fg in zip(f_e, f_g):
        for xe, xg in zip(fe, fg):
            loss = loss + F.l1_loss(xe, xg)
    for ye, yg in zip(y_e, y_g):
        loss = loss + F.mse_loss(ye, torch.ones_like(ye)) + F.mse_loss(yg, torch.zeros_like(yg))
    return mse, loss


def subspace_angle(Va: torch.Tensor, Vb: torch.Tensor) -> float:
    """Largest principal angle (rad) between column spaces, float64 sine formulation
    ``asin ||(I - Qa Qa^T) Qb||_2`` after QR (SURVEY.md section 7 hard part 5)."""
    if Va.shape[1] == 0 and Vb.shape[1] == 0:
        return 0.0
    Qa, _ = torch.linalg.qr(Va.double())
    Qb, _ = torch.linalg.qr(Vb.double())
    R = Qb - Qa @ (Qa.transpose(0, 1) @ Qb)
    s = torch.linalg.matrix_norm(R, ord=2).item()
    return math.asin(min(1.0, s))


# --------------------------------------------------------------------------------------------
# Seeded synthetic inputs -- shared by tests and bench
# --------------------------------------------------------------------------------------------

def synth_audio(B: int, T: int, n_views: int = 1, seed: int = 1234, sr: int = 16000):
    """Correlated multi-view 16 kHz clips in [-1, 1]; returns ``(views [V][B,T], decoded [V][B,T])``."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    n = 0.1 * torch.randn(B, T + 1, generator=g)
    src = 0.5 * (n[:, 1:] + n[:, :-1])
    t = torch.arange(T, dtype=torch.float64) / sr
    for f0 in (220.0, 880.0, 3520.0):
        ph = 2 * math.pi * torch.rand(B, 1, generator=g, dtype=torch.float64)
        src = src + 0.05 * torch.sin(2 * math.pi * f0 * t[None, :] + ph).float()
    amps = (1.0, 0.9, 0.8, 0.7)
    views, decs = [], []
    for i in range(n_views):
        gi = torch.Generator(device="cpu").manual_seed(seed + 1 + i)
        v = amps[i % 4] * src + 0.05 * torch.randn(B, T, generator=gi)
        gd = torch.Generator(device="cpu").manual_seed(seed + 1001 + i)
        d = v + 0.02 * torch.randn(B, T, generator=gd)
        views.append(v.clamp(-1, 1).contiguous())
        decs.append(d.clamp(-1, 1).contiguous())
    return views, decs


def synth_latents(B: int, d: int, n_views: int, seed: int = 3234, decay: float | None = None):
    """Latents with a well-separated decaying spectrum: ``Z_i = (G_i * sigma) Q_i^T + 0.3``;
    correlated views share half of ``G``."""
    decay = (0.98 if d >= 256 else 0.97) if decay is None else decay
    sigma = decay ** torch.arange(d, dtype=torch.float32)
    g0 = torch.Generator(device="cpu").manual_seed(seed - 1)
    shared = torch.randn(B, d, generator=g0)
    out = []
    for i in range(n_views):
        g = torch.Generator(device="cpu").manual_seed(seed + i)
        own = torch.randn(B, d, generator=g)
        G = (shared + own) / math.sqrt(2.0)
        # distinct per-view gains keep the concatenated singular values free of exact ties
        Q, _ = torch.linalg.qr(torch.randn(d, d, generator=g))
        out.append((((G * sigma) @ Q.transpose(0, 1)) * (1.0 + 0.03 * i) + 0.3).contiguous())
    return out
