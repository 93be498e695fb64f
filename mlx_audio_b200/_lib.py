"""ctypes binding of libkokoro_b200.so (the C ABI declared in include/kokoro_b200.h).

There is NO fallback: if the shared library is missing or a call fails, this module raises.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import torch

_PKG = os.path.dirname(os.path.abspath(__file__))
# KK_LIB selects another build of the same ABI (experiments: the -DKK_TUNE / -DKK_FUSED_DBG "lab" library)
LIB_PATH = os.environ.get("KK_LIB") or os.path.join(_PKG, "libkokoro_b200.so")

KK_ACT_NONE, KK_ACT_LEAKY, KK_ACT_SNAKE, KK_ACT_SNAKE_FAST = 0, 1, 2, 3
KK_ERR_INVALID_ARG, KK_ERR_UNSUPPORTED, KK_ERR_CUDA = -1, -2, -3


class KokoroB200Error(RuntimeError):
    pass


class ConvDesc(C.Structure):
    """Mirror of ``kk_conv_desc`` (include/kokoro_b200.h)."""
    _fields_ = [
        ("B", C.c_int), ("L_in", C.c_int), ("C_in", C.c_int), ("x_pitch", C.c_int),
        ("taps", C.c_int), ("dil", C.c_int), ("pad_left", C.c_int), ("stride", C.c_int),
        ("Cw", C.c_int),
        ("M", C.c_int), ("N", C.c_int),
        ("C_out", C.c_int), ("up", C.c_int), ("up_off", C.c_int), ("L_true", C.c_int), ("left_pad", C.c_int),
        ("L_out", C.c_int),
        ("bias", C.c_void_p),
        ("res", C.c_void_p), ("res_pitch", C.c_int), ("res_shift", C.c_int), ("L_res", C.c_int),
        ("res2", C.c_void_p), ("res2_pitch", C.c_int),
        ("scale", C.c_float),
        ("out", C.c_void_p), ("out_pitch", C.c_int),
        ("stats", C.c_void_p),
        ("out_act", C.c_void_p), ("act_pitch", C.c_int), ("act_is_bf16", C.c_int), ("act_slope", C.c_float),
        ("out_add", C.c_void_p), ("add_pitch", C.c_int), ("res_is_bf16", C.c_int),
        ("lens", C.c_void_p), ("len_mul", C.c_int), ("len_add", C.c_int),
    ]


class AdainIn(C.Structure):
    """Mirror of ``kk_adain_in`` (include/kokoro_b200.h)."""
    _fields_ = [
        ("x", C.c_void_p), ("x_is_bf16", C.c_int), ("x_pitch", C.c_int),
        ("stats", C.c_void_p),
        ("h", C.c_void_p), ("h_stride", C.c_int),
        ("alpha", C.c_void_p),
        ("act", C.c_int), ("slope", C.c_float),
    ]


_I, _F, _P, _LL, _ULL = C.c_int, C.c_float, C.c_void_p, C.c_longlong, C.c_ulonglong

# name -> argtypes; every function returns int except kk_last_error
PROTOTYPES = {
    "kk_version": [],
    "kk_device_is_sm100": [],
    "kk_fold_weight_norm": [_P, _P, _P, _I, _I, _P],
    "kk_pack_conv_weight": [_P, _P, _I, _I, _I, _I, _I, _I, _I, _P],
    "kk_style_affine": [_P, _P, _P, _P, _I, _I, _I, _P],
    "kk_instnorm_stats": [_P, _P, _I, _I, _I, _I, _P, _I, _I, _P],
    "kk_adain_act": [_P, _I, _P, _P, _I, _P, _I, _F, _P, _I, _I, _I, _I, _I, _I, _P, _P, _P, _I, _I, _P],
    "kk_front_assemble": [_P, _P, _P, _P, _P, _P, _I, _I, _I, _I, _I, _I, _I, _P, _P],
    "kk_copy_channels": [_P, _I, _I, _P, _I, _I, _I, _I, _P],
    "kk_convert_act": [_P, _I, _P, _I, _I, _LL, _I, _F, _P],
    "kk_conv1d_f32": [_P, _P, C.POINTER(ConvDesc), _P],
    "kk_conv1d_bf16_tc": [_P, _P, C.POINTER(ConvDesc), _I, _P],
    "kk_conv1d_adain_bf16_tc": [C.POINTER(AdainIn), _P, C.POINTER(ConvDesc), _P],
    "kk_split_tf32": [_P, _P, _P, _LL, _P],
    "kk_conv1d_tf32_tc": [_P, _P, _P, C.POINTER(ConvDesc), _I, _P],
    "kk_hnsf_source": [_P, _P, _ULL, _P, _P, _P, _P, _P, _I, _I, _I, _P, _P],
    "kk_hnsf_source_noalloc": [_P, _P, _ULL, _P, _P, _P, _P, _I, _I, _I, _P, _P],
    "kk_stft_n20": [_P, _P, _I, _I, _I, _P, _I, _P],
    "kk_stft_n20_bf16": [_P, _P, _I, _P, _I, _I, _I, _I, _P, _I, _P],
    "kk_istft_head_n20": [_P, _I, _P, _I, _I, _P, _I, _I, _P],
    "kk_clip_f32": [_P, _P, _F, _F, _LL, _P],
    "kk_axpby_f32": [_P, _P, _P, _F, _F, _LL, _P],
    "kk_aa_act_f32": [_P, _P, _P, _P, _P, _P, _I, _I, _I, _I, _F, _P],
    "kk_lstm_bidir_f32": [_P, _P, _P, _I, _I, _I, _P, _P],
    "kk_stft_f32": [_P, _P, _P, _I, _I, _I, _I, _I, _I, _P, _P],
    "kk_logmel_f32": [_P, _P, _P, _P, _I, _I, _I, _I, _I, _I, _I, _I, _I, _F, _P, _P],
    "kk_istft_f32": [_P, _P, _P, _I, _I, _I, _I, _I, _I, _P, _I, _P],
    "kk_polar_to_spec": [_P, _I, _P, _LL, _I, _F, _P],
    "kk_clamp_below_max_affine": [_P, _I, _LL, _F, _F, _F, _P],
    "kk_normalize_features": [_P, _I, _I, _I, _I, _F, _P],
    "kk_snake_f32": [_P, _P, _P, _P, _LL, _I, _LL, _I, _F, _P],
    "kk_dwconv1d_f32": [_P, _P, _P, _P, _P, _I, _I, _I, _I, _I, _I, _P],
    "kk_noise_mix_f32": [_P, _P, _P, _P, _I, _LL, _I, _P],
    "kk_tanh_f32": [_P, _P, _LL, _P],
    "kk_fold_weight_norm_ex": [_P, _P, _P, _I, _I, _F, _I, _P],
}

_lib: Optional[C.CDLL] = None


def load() -> C.CDLL:
    """Load the shared library (once).  Raises if it has not been built - there is no CPU path."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise KokoroB200Error(
            f"{LIB_PATH} not found: build it with `python -m mlx_audio_b200._build` "
            "(or __graft_entry__.build()).  mlx_audio_b200 has no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, argtypes in PROTOTYPES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is missing
        fn.argtypes = argtypes
        fn.restype = C.c_int
    lib.kk_last_error.argtypes = []
    lib.kk_last_error.restype = C.c_char_p
    _lib = lib
    return lib


def last_error() -> str:
    return load().kk_last_error().decode("utf-8", "replace")


def check(rc: int, what: str = "") -> None:
    if rc == 0:
        return
    msg = last_error()
    if rc == KK_ERR_INVALID_ARG:
        raise ValueError(f"{what}: {msg}")
    if rc == KK_ERR_UNSUPPORTED:
        raise KokoroB200Error(f"{what}: unsupported: {msg}")
    raise KokoroB200Error(f"{what}: {msg} (code {rc})")


def ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    """Device pointer of a CUDA tensor (None -> NULL)."""
    if t is None:
        return None
    if not t.is_cuda:
        raise KokoroB200Error("mlx_audio_b200 kernels need CUDA tensors; there is no CPU fallback")
    return t.data_ptr()


def stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def require_cuda() -> None:
    if not torch.cuda.is_available():
        raise KokoroB200Error("no CUDA device: mlx_audio_b200 runs only on a B200 (sm_100a); no CPU fallback")
