"""Drop-in for mlx_audio/tts/models/indextts/mel.py ``log_mel_spectrogram`` on ``kk_logmel_f32``."""
from __future__ import annotations

import torch

from .... import utils as U


def log_mel_spectrogram(audio, sample_rate: int = 24_000, n_mels: int = 100, n_fft: int = 1024, hop_length: int = 256,
                        padding: int = 0) -> torch.Tensor:
    """indextts/mel.py:6-37: ``log(max(|stft(audio, "hann", n_fft, hop_length, n_fft)| @ mel_filters(htk).T, 1e-5))[None]``
    -> (1, frames, n_mels); a (B, N) batch returns (B, frames, n_mels)."""
    audio = U._as_cuda(audio)
    if padding > 0:
        audio = torch.nn.functional.pad(audio, (0, padding))
    mel = U.log_mel_spectrogram(audio, sample_rate=sample_rate, n_fft=n_fft, hop_length=hop_length, n_mels=n_mels,
                                window="hann", power=1, log="ln", floor=1e-5, norm=None, mel_scale="htk")
    return mel[None] if mel.dim() == 2 else mel
