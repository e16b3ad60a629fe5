"""ctypes binding of ``libtac_spectral.so`` (the C ABI declared in ``include/tac_spectral.h``).

There is no CPU fallback: if the shared library is missing or a call fails, a ``RuntimeError``
carrying ``tac_last_error()`` is raised -- mirroring how the reference's own native ops surface
``TORCH_CHECK`` failures as Python ``RuntimeError`` (``sgmse/.../op/fused_bias_act.cpp:7-14``).
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Sequence

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
# TAC_SPECTRAL_LIB: developer override used to A/B kernel variants built with build.py --out=...
LIB_PATH = os.environ.get("TAC_SPECTRAL_LIB") or os.path.join(_HERE, "libtac_spectral.so")

# name -> (restype, argtypes): every symbol of include/tac_spectral.h
_I, _I64, _SZ, _VP, _F = C.c_int, C.c_int64, C.c_size_t, C.c_void_p, C.c_float
_IP = C.POINTER(C.c_int)
_VPP = C.POINTER(C.c_void_p)
XCHG_MAX_WORLD = 8


class XchgStruct(C.Structure):
    """``tac_xchg`` of include/tac_spectral.h (host struct, passed by pointer; NULL = no exchange)."""
    _fields_ = [("world", C.c_int), ("rank", C.c_int), ("payload", C.c_int64), ("peer", C.c_void_p * XCHG_MAX_WORLD)]


class EpilogueStruct(C.Structure):
    """``tac_epilogue`` of include/tac_spectral.h (host struct, passed by pointer)."""
    _fields_ = [("kind", C.c_int), ("p0", C.c_float), ("p1", C.c_float), ("p2", C.c_float), ("n_mels", C.c_int),
                ("mel_fb", C.c_void_p), ("mel_range", C.c_void_p), ("rowmax", C.c_void_p)]


_XP = C.POINTER(XchgStruct)
_EP = C.POINTER(EpilogueStruct)
_UBP = C.POINTER(C.c_ubyte)
SIGNATURES = {
    "tac_last_error": (C.c_char_p, []),
    "tac_version": (_I, []),
    "tac_launch_count": (C.c_uint64, []),
    "tac_prepare_fft": (_I, [_I, _I, _I]),
    "tac_shutdown": (_I, []),
    "tac_stft_num_frames": (_I64, [_I64, _I, _I, _I]),
    "tac_stft_multi": (_I, [_VP, _I64, _I64, _I64, _I, _IP, _IP, _IP, _I, _I, _I, _I, _I64, _VPP, _VP]),
    "tac_stft_multi_epi": (_I, [_VP, _I64, _I64, _I64, _I, _IP, _IP, _IP, _I, _I, _I, _EP, _VPP, _VP]),
    "tac_stft_ex_supported": (_I, [_I]),
    "tac_prepare_stft_ex": (_I, [_I, _I, _I, _I]),
    "tac_stft_ex": (_I, [_VP, _I64, _I64, _I, _I, _I, _I, _I, _I, _I, _EP, _I64, _VP, _VP]),
    "tac_db_from_rowmax": (_I, [_VP, _I64, _I64, _I, _I, _I64, _VP, _F, _F, _VP, _VP]),
    "tac_mrstft_workspace_bytes": (_SZ, [_I64, _I64, _I, _IP, _IP]),
    "tac_mrstft_loss_fwd": (_I, [_VP, _VP, _I64, _I64, _I, _IP, _IP, _IP, _F, _VP, _VP, _SZ, _VP]),
    "tac_mrstft_loss_bwd": (_I, [_VP, _VP, _I64, _I64, _I, _IP, _IP, _IP, _F, _VP, _VP, _VP, _SZ, _VP]),
    "tac_mrstft_loss_fwd_bwd": (_I, [_VP, _VP, _I64, _I64, _I, _IP, _IP, _IP, _F, _VP, _VP, _VP, _VP, _SZ, _VP]),
    "tac_spec_losses_workspace_bytes": (_SZ, [_I64]),
    "tac_spec_losses": (_I, [_VP, _VP, _I64, _I64, _I64, _VP, _VP, _VP, _SZ, _VP]),
    "tac_spec_losses_static": (_I, [_VP, _VP, _I64, _I64, _I64, _VP, _VP, _VP]),
    "tac_spec_losses_bwd": (_I, [_VP, _VP, _I64, _I64, _I64, _VP, _VP, _VP]),
    "tac_dpca_workspace_bytes": (_SZ, [_I, _I64, _I]),
    "tac_dpca_stats": (_I, [_VPP, _I, _I64, _I, _I, _VP, _VP, _SZ, _VP]),
    "tac_dpca_eig": (_I, [_VP, _I, _I, _VP, _VP, _VP, _VP, _VP, _SZ, _VP]),
    "tac_dpca_select": (_I, [_VP, _VP, _I, _I, _I, _VP, _VP, _VP]),
    "tac_dpca_project": (_I, [_VPP, _I, _I64, _I, _VP, _VP, _VP, _VP, _VPP, _VP, _SZ, _VP]),
    "tac_var_topk": (_I, [_VP, _I, _I, _VP, _VP, _VP]),
    "tac_amplitude_to_db": (_I, [_VP, _I64, _I64, _I, _I, _I64, _F, _F, _VP, _VP, _SZ, _VP]),
    "tac_nuclear_from_s": (_I, [_VP, _VP, _VP, _I, _I, _VP, _VP, _VP]),
    "tac_nuclear_grad_matrix": (_I, [_VP, _VP, _VP, _VP, _I, _I, _VP, _VP]),
    "tac_dpca_apply": (_I, [_VPP, _I, _I64, _I, _VP, _VP, _I, _VPP, _VP]),
    "tac_cosine_pairs_workspace_bytes": (_SZ, [_I64]),
    "tac_cosine_pairs_fwd": (_I, [_VPP, _I, _I64, _I, _F, _VP, _VP, _SZ, _VP]),
    "tac_cosine_pairs_bwd": (_I, [_VPP, _I, _I64, _I, _F, _VP, _F, _VPP, _VP]),
    "tac_istft_length": (_I64, [_I64, _I, _I, _I]),
    "tac_istft": (_I, [_VP, _VP, _I, _I64, _I64, _I64, _I64, _I, _I, _I, _I, _I, _VP, _I64, _VP]),
    "tac_chain_rows_per_group": (_I64, [_I64, _I64, _I64]),
    "tac_istft_disc_inputs": (_I, [_VP, _VP, _I, _I64, _I64, _I64, _I64, _I, _I, _I, _I, _I, _VP, _I64, _I, _IP, _IP, _IP, _I, _VPP, _I64, _VP]),
    "tac_stft_adjoint": (_I, [_VP, _VP, _I, _I64, _I64, _I64, _I64, _I, _I, _I, _F, _VP, _I64, _VP]),
    "tac_bottleneck_width": (_I, [_I, _I, _I]),
    "tac_bottleneck_mix": (_I, [_VPP, _I, _I64, _I, _I, _I, _VPP, _VP, _VP]),
    "tac_bottleneck_mix_bwd": (_I, [_VP, _I, _I64, _I, _I, _I, _VPP, _VPP, _VP]),
    "tac_interp_mse_workspace_bytes": (_SZ, []),
    "tac_interp_mse": (_I, [_VP, _I64, _VP, _I64, _I64, _VP, _VP, _SZ, _VP]),
    "tac_multi_reduce_workspace_bytes": (_SZ, [_I, C.POINTER(C.c_int64)]),
    "tac_multi_reduce": (_I, [_I, _VPP, _VPP, C.POINTER(C.c_int64), _IP, C.POINTER(C.c_float), _VP, _VP, _SZ, _VP]),
    "tac_xchg_bytes": (_SZ, [_I, _I64]),
    "tac_xchg_set_timeout": (_I, [C.c_double]),
    "tac_xchg_status": (_I, [_XP, C.POINTER(C.c_uint)]),
    "tac_xchg_push": (_I, [_XP, _VP, _VP]),
    "tac_xchg_pop": (_I, [_XP, _VP, _VP]),
    "tac_dpca_stats_push": (_I, [_VPP, _I, _I64, _I, _I, _VP, _VP, _SZ, _XP, _VP]),
    "tac_dpca_eig_pop": (_I, [_VP, _I, _I, _VP, _VP, _VP, _VP, _VP, _SZ, _XP, _VP]),
    "tac_loss_finalize": (_I, [_VP, _I, _VP, _XP, _VP, _VP]),
    "tac_ipc_alloc": (_I, [_SZ, _VPP, _UBP]),
    "tac_ipc_open": (_I, [_UBP, _VPP]),
    "tac_ipc_close": (_I, [_VP]),
    "tac_ipc_free": (_I, [_VP]),
}

PAD_ZERO, PAD_REFLECT = 0, 1
OUT_COMPLEX, OUT_PLANES, OUT_MAGPHASE = 0, 1, 2
EPI_COMPLEX, EPI_COMPRESS, EPI_MAGPHASE, EPI_MAGDB, EPI_POWER, EPI_LOGPOWER, EPI_DB, EPI_MEL = range(8)
WINDOW_HANN, WINDOW_SQRT_HANN = 0, 1
GRAM_TF32X3, GRAM_FP64 = 0, 1
IN_POLAR, IN_PLANES, IN_COMPLEX = 0, 1, 2
MAX_SCALES = 8             # TAC_MAX_SCALES

_lib = None


def lib() -> C.CDLL:
    """Load (once) the shared library; fails loudly when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python task-aware-audio-coding-perceptual_b200/build.py` "
                "(there is no CPU or PyTorch fallback for the spectral hot path)")
        handle = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)          # AttributeError if the ABI and the header disagree
            fn.restype, fn.argtypes = res, args
        _lib = handle
    return _lib


def check(rc: int) -> None:
    if rc != 0:
        msg = lib().tac_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"tac_spectral error {rc}: {msg}")


def int_array(values: Sequence[int]):
    return (C.c_int * len(values))(*[int(v) for v in values])


def ptr_array(tensors: Sequence[torch.Tensor]):
    return (C.c_void_p * len(tensors))(*[t.data_ptr() for t in tensors])


def stream_of(t: torch.Tensor) -> int:
    """The caller's current CUDA stream on ``t``'s device, exactly as the reference's native
    launchers take it (``upfirdn2d_kernel.cu:213-215``)."""
    return torch.cuda.current_stream(t.device).cuda_stream


def require_cuda(t: torch.Tensor, name: str) -> None:
    if not isinstance(t, torch.Tensor):
        raise TypeError(f"{name} must be a torch.Tensor")
    if not t.is_cuda:
        raise RuntimeError(f"{name} must live on a CUDA device: the spectral hot path has no CPU fallback")


class device_of:
    """Context manager: make ``t``'s device current for the duration of a C-ABI call."""

    def __init__(self, t: torch.Tensor):
        self._g = torch.cuda.device(t.device)

    def __enter__(self):
        self._g.__enter__()
        return self

    def __exit__(self, *exc):
        return self._g.__exit__(*exc)
