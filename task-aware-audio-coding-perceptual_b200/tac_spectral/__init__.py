"""tac_spectral -- B200-native spectral hot path of the task-aware audio coder.

Host-side mirror (PyTorch plumbing over the C ABI of ``include/tac_spectral.h``) of the reference
call sites on the path ``BASELINE.json:north_star`` names:

* :mod:`.stft`  -- ``Spectrogram`` / ``MultiScaleSTFT`` (``msstftd.py:125-127,151-153,185-192``) and the
  notebook front end ``create_complex_spectrogram`` (``spectrogram.ipynb`` cells 9 and 10); any-``n_fft`` STFT with fused
  epilogues (``stft_ex``: power / log / dB / mel, and the SGMSE+ front end ``SpecsFrontEnd``, ``sgmse/sgmse/data_module.py:162-214``);
* :mod:`.loss`  -- ``mrstft_loss`` and ``spectral_losses`` (``pkl_file_models.py:933-940``);
* :mod:`.dpca`  -- ``data_pca`` / ``dpca_project`` / ``apply_pca`` / ``var_topk``
  (``pkl_file_model_updated.py:37-42,489-511,398-427``; ``pkl_file_models.py:224-228,901-907``).
* :mod:`.tail`  -- the bottleneck that is live in the SpectralRes models (``pkl_file_models.py:1108-1127,692-718,335-354``) with
  the DPCA hook, and the loss tail of ``task_aware()`` (``train_pkl_file_task_aware.py:307-364``).

There is no CPU or eager-PyTorch fallback: every op raises ``RuntimeError`` when
``libtac_spectral.so`` is missing or its input is not on a CUDA device.
"""
from . import _lib
from .stft import (Spectrogram, MultiScaleSTFT, stft_multi, stft_multi_power, complex_spectrogram, create_complex_spectrogram, amplitude_to_db,
                   num_frames, SUPPORTED_NFFT, istft, batch_reconstruct_waveform, reconstruct_waveform, stft_ex, ex_supported,
                   MelBank, MelSpectrogram, SpecsFrontEnd, reconstruct_and_disc_inputs)
from .tail import bottleneck_mix, spectral_bottleneck, task_aware_tail, interp_mse, multi_reduce
from .loss import mrstft_loss, mrstft_sums, mrstft_counts, spectral_losses, LOG_FLOOR
from .dpca import data_pca, dpca_project, dpca_select, dpca_stats, dpca_eig, apply_pca, var_topk, nuclear_loss, pairwise_cosine_loss

__all__ = ["Spectrogram", "MultiScaleSTFT", "stft_multi", "stft_multi_power", "complex_spectrogram", "create_complex_spectrogram", "amplitude_to_db",
           "num_frames", "SUPPORTED_NFFT", "mrstft_loss", "mrstft_sums", "mrstft_counts", "spectral_losses",
           "LOG_FLOOR", "data_pca", "dpca_project", "dpca_select", "dpca_stats", "dpca_eig", "apply_pca", "var_topk", "nuclear_loss",
           "pairwise_cosine_loss", "istft", "batch_reconstruct_waveform", "reconstruct_waveform", "stft_ex", "ex_supported",
           "MelBank", "MelSpectrogram", "SpecsFrontEnd", "reconstruct_and_disc_inputs", "bottleneck_mix", "spectral_bottleneck", "task_aware_tail", "interp_mse",
           "multi_reduce"]
