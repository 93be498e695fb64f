"""Oracle restatement of mlx_audio/tts/models/interpolate.py.

TEST INFRASTRUCTURE ONLY - see oracle/__init__.py.  float32 NumPy with the
reference's operation order (coordinates are computed in float32 as separate
mul / add / sub; a floor of -1 is used as a wrap-around gather index).
"""
from __future__ import annotations

from typing import Optional

import numpy as np


def interpolate(inp: np.ndarray, size=None, scale_factor=None, mode: str = "nearest",
                align_corners: Optional[bool] = None) -> np.ndarray:
    """interpolate.py:6-54 (1-D spatial only)."""
    inp = np.asarray(inp)
    ndim = inp.ndim
    if ndim < 3:
        raise ValueError(f"Expected at least 3D input (N, C, D1), got {ndim}D")
    spatial_dims = ndim - 2
    if size is not None and scale_factor is not None:
        raise ValueError("Only one of size or scale_factor should be defined")
    if size is None and scale_factor is None:
        raise ValueError("One of size or scale_factor must be defined")
    if size is not None and not isinstance(size, (list, tuple)):
        size = [size] * spatial_dims
    if scale_factor is not None and not isinstance(scale_factor, (list, tuple)):
        scale_factor = [scale_factor] * spatial_dims
    if size is None:
        size = []
        for i in range(spatial_dims):
            # interpolate.py:45: int(mx.ceil(shape * scale)) evaluated in float32
            sf = scale_factor[i]
            if isinstance(sf, np.generic) and sf.dtype == np.float32:
                prod = np.float32(inp.shape[i + 2]) * sf  # mx float32 scalar scale (e.g. 1/upsample_scale)
            else:
                prod = np.float32(inp.shape[i + 2] * sf)  # python scalar: double product, then float32
            size.append(max(1, int(np.ceil(prod))))
    if spatial_dims == 1:
        return interpolate1d(inp, size[0], mode, align_corners)
    raise ValueError(f"Only 1D interpolation currently supported, got {spatial_dims}D")


def interpolate1d(inp: np.ndarray, size: int, mode: str = "linear", align_corners: Optional[bool] = None) -> np.ndarray:
    """interpolate.py:57-108."""
    inp = np.asarray(inp)
    batch, channels, in_width = inp.shape
    dt = inp.dtype.type if inp.dtype.kind == "f" else np.float32
    if size < 1:
        size = 1
    if in_width < 1:
        in_width = 1
    if mode == "nearest":  # interpolate.py:71-78
        if size == 1:
            idx = np.array([0])
        else:
            scale = in_width / size
            idx = np.floor(np.arange(size).astype(np.float32) * np.float32(scale)).astype(np.int32)
            idx = np.clip(idx, 0, in_width - 1)
        return inp[:, :, idx]
    # linear (interpolate.py:81-108)
    ar = np.arange(size).astype(np.float32)
    if align_corners and size > 1:
        x = ar * np.float32((in_width - 1) / (size - 1))
    else:
        if size == 1:
            x = np.array([0.0], np.float32)
        else:
            x = ar * np.float32(in_width / size)
            if not align_corners:
                # (x + 0.5*(in/size)) - 0.5 : two separate float32 roundings (interpolate.py:90)
                x = (x + np.float32(0.5 * (in_width / size))).astype(np.float32) - np.float32(0.5)
    if in_width == 1:
        return np.broadcast_to(inp, (batch, channels, size)).copy()
    x_low = np.floor(x).astype(np.int32)
    x_high = np.minimum(x_low + 1, in_width - 1)
    x_frac = (x - x_low.astype(np.float32)).astype(np.float32)
    # x_low == -1 wraps to the last element (MLX / NumPy negative-index gather, interpolate.py:97-102)
    y_low = inp[:, :, x_low]
    y_high = inp[:, :, x_high]
    one_m = (np.float32(1) - x_frac).astype(np.float32)
    out = y_low * one_m.astype(dt)[None, None, :] + y_high * x_frac.astype(dt)[None, None, :]
    return out.astype(inp.dtype if inp.dtype.kind == "f" else np.float32)
