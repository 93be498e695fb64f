"""WAV output for generated audio.  The reference writes through ``soundfile``
(``sf.write(file_name, result.audio, result.sample_rate)``, tts/generate.py:313 - PCM_16 is soundfile's default WAV
subtype); soundfile is not in this image, so the same container is written with the stdlib ``wave`` module."""
from __future__ import annotations

import wave
from typing import Tuple

import numpy as np


def to_pcm16(audio) -> np.ndarray:
    """float waveform in [-1, 1] -> int16, libsndfile's float->PCM_16 rule: scale by 32767 + 1 = 0x8000, round to
    nearest, clip to [-32768, 32767]."""
    if hasattr(audio, "detach"):
        audio = audio.detach().float().cpu().numpy()
    a = np.asarray(audio, dtype=np.float32).reshape(-1)
    return np.clip(np.rint(a.astype(np.float64) * 32768.0), -32768, 32767).astype("<i2")


def save_wav(path: str, audio, sample_rate: int = 24000) -> int:
    """Mono PCM16 WAV; returns the number of samples written."""
    pcm = to_pcm16(audio)
    with wave.open(path, "wb") as f:
        f.setnchannels(1)
        f.setsampwidth(2)
        f.setframerate(int(sample_rate))
        f.writeframes(pcm.tobytes())
    return int(pcm.shape[0])


def load_wav(path: str) -> Tuple[np.ndarray, int]:
    """PCM16 mono WAV -> (float32 in [-1, 1), sample_rate)."""
    with wave.open(path, "rb") as f:
        if f.getsampwidth() != 2:
            raise ValueError(f"{path}: only PCM16 is supported")
        n, ch, sr = f.getnframes(), f.getnchannels(), f.getframerate()
        pcm = np.frombuffer(f.readframes(n), dtype="<i2").reshape(-1, ch)
    return (pcm[:, 0].astype(np.float32) / np.float32(32768.0)), sr
