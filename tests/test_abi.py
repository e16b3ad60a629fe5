"""CPU-side checks of the drop-in boundary: the C-ABI library builds, loads without a GPU and exports
every symbol include/tac_spectral.h declares; host-side argument validation raises before any launch."""
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def ts():
    import __graft_entry__ as g
    g.build()
    import tac_spectral
    return tac_spectral


def test_header_symbols_exported(ts):
    hdr = open(os.path.join(ROOT, "include", "tac_spectral.h")).read()
    declared = set(re.findall(r"TAC_API\s+[\w\s\*]+?\b(tac_\w+)\s*\(", hdr))
    assert len(declared) >= 18
    lib = ts._lib.lib()
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in the header but not exported"
    assert declared == set(ts._lib.SIGNATURES), declared ^ set(ts._lib.SIGNATURES)


def test_host_only_entry_points(ts):
    lib = ts._lib.lib()
    assert lib.tac_version() >= 100
    # frame counts follow torch.stft: 1 + (T + 2*(n/2)*center - n) // hop
    assert lib.tac_stft_num_frames(24000, 1024, 256, 0) == 90
    assert lib.tac_stft_num_frames(24000, 2048, 512, 0) == 43
    assert lib.tac_stft_num_frames(24000, 512, 128, 0) == 184
    assert lib.tac_stft_num_frames(16000, 2048, 512, 1) == 32
    assert lib.tac_stft_num_frames(100, 2048, 512, 0) == 0
    assert ts.mrstft_counts(2, 24000, [1024, 2048, 512], [256, 512, 128]) == [2 * 513 * 90, 2 * 1025 * 43, 2 * 257 * 184]
    assert lib.tac_dpca_workspace_bytes(4, 1024, 128) >= 4 * 128 * 128 * 8
    # the split eigen-solve parks T, the reflectors and the float64 eigenvectors at FIXED per-view offsets (dpca_eig_tri.cuh:
    # kTriWsDoubles = 9216 + 128 * 128), also for small d
    assert lib.tac_dpca_workspace_bytes(2, 256, 8) >= 2 * (9216 + 128 * 128) * 8
    assert lib.tac_dpca_workspace_bytes(1, 64, 128) >= (9216 + 128 * 128) * 8
    assert lib.tac_spec_losses_workspace_bytes(4) > 0
    # exchange wait bound: host-side setter validates; the status of "no exchange" is 0 without touching a device
    import ctypes as C
    assert lib.tac_xchg_set_timeout(C.c_double(30.0)) == 0 and lib.tac_xchg_set_timeout(C.c_double(-1.0)) == 0
    assert lib.tac_xchg_set_timeout(C.c_double(float("nan"))) != 0
    w = C.c_uint(5)
    assert lib.tac_xchg_status(None, C.byref(w)) == 0 and w.value == 0
    # chain row groups: the group's waveform stays inside the L2 budget, groups are even, never empty
    g = lib.tac_chain_rows_per_group
    assert g(128, 306688, 0) == 32                                   # 128 x 19 s decoder outputs: 4 groups of 32 rows (37.4 MiB each)
    assert g(5, 16000, 0) == 5 and g(1, 10 ** 9, 0) == 1 and g(0, 16000, 0) == 0
    assert g(100, 306688, 10 << 20) == 8                             # 13 groups under a 10 MiB budget


def test_no_cpu_fallback(ts):
    x = torch.randn(2, 1, 4000)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ts.stft_multi(x, [512], [128])
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ts.mrstft_loss(x, x)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ts.dpca_project([torch.randn(8, 16)], 4)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ts.spectral_losses(torch.zeros(1, 2, 3, 4), torch.zeros(1, 2, 3, 4))


def test_constructor_validation(ts):
    with pytest.raises(ValueError):
        ts.Spectrogram(n_fft=4000, power=None)                      # beyond the chirp-z path (n + n/2 <= 4096)
    with pytest.raises(ValueError):
        ts.Spectrogram(n_fft=512, window_fn=torch.hamming_window, power=None)
    with pytest.raises(ValueError):
        ts.Spectrogram(n_fft=512, power=-1.0)
    sp = ts.Spectrogram()                                            # testing_scripts.ipynb:246: torchaudio's defaults
    assert (sp.n_fft, sp.hop_length, sp.win_length, sp.power, sp.center) == (400, 200, 400, 2.0, True)
    assert ts.ex_supported(2046) and ts.ex_supported(510) and ts.ex_supported(400) and not ts.ex_supported(4096)
    with pytest.raises(NotImplementedError):
        ts.SpecsFrontEnd(window="hamming")                           # data_module.py:19
    fe = ts.SpecsFrontEnd()                                          # data_module.py:109-118 defaults
    assert (fe.n_fft, fe.hop_length, fe.spec_factor, fe.spec_abs_exponent) == (510, 128, 0.15, 0.5)
    with pytest.raises(AssertionError):
        ts.MultiScaleSTFT([512, 1024], [128], [512, 1024])          # msstftd.py:177
    sp = ts.Spectrogram(n_fft=1024, hop_length=256, win_length=1024, window_fn=torch.hann_window, normalized=True,
                        center=False, pad_mode=None, power=None)     # msstftd.py:125-127 verbatim
    assert (sp.n_fft, sp.hop_length, sp.win_length, sp.normalized, sp.center) == (1024, 256, 1024, True, False)


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "task-aware-audio-coding-perceptual_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in src and "from oracle" not in src, f
