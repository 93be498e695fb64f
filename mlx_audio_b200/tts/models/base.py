"""Helpers mirrored from mlx_audio/tts/models/base.py (plain Python, no arithmetic on the hot path)."""
from __future__ import annotations

import inspect
from dataclasses import dataclass


@dataclass
class BaseModelArgs:
    @classmethod
    def from_dict(cls, params):
        """base.py:9-18: keep only the keys the dataclass declares."""
        return cls(**{k: v for k, v in params.items() if k in inspect.signature(cls).parameters})


def check_array_shape(arr) -> bool:
    """base.py:21-34: True when a 3-D conv weight already looks like the MLX layout (O, K, I) with K == I
    and O the largest dimension (the reference's heuristic, reproduced as is)."""
    shape = tuple(arr.shape)
    if len(shape) != 3:
        return False
    out_channels, kH, KW = shape
    return (out_channels >= kH) and (out_channels >= KW) and (kH == KW)


@dataclass
class GenerationResult:
    """base.py:72-84."""
    audio: object
    samples: int
    sample_rate: int
    segment_idx: int
    token_count: int
    audio_samples: dict
    audio_duration: str
    real_time_factor: float
    prompt: dict
    processing_time_seconds: float
    peak_memory_usage: float
