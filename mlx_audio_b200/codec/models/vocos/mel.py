"""Drop-in for mlx_audio/codec/models/vocos/mel.py ``log_mel_spectrogram`` on the fused
STFT -> |X| -> mel -> log kernel (``kk_logmel_f32``; the spectrogram never reaches HBM)."""
from __future__ import annotations

import torch

from .... import utils as U


def log_mel_spectrogram(audio, sample_rate: int = 24_000, n_mels: int = 100, n_fft: int = 1024, hop_length: int = 256,
                        padding: int = 0) -> torch.Tensor:
    """vocos/mel.py:8-33: ``log(max(|stft(audio, window=hanning(n_fft))[:-1]| @ mel_filters(htk).T, 1e-5))[None]``.
    Quirks kept: the reference passes ``win_length=hop_length`` but an explicit window array, so the window is the full
    symmetric hanning(n_fft) (utils.py:72-76) and the hop is stft's default n_fft // 4 - ``hop_length`` only shapes the
    (ignored) win_length; the last frame is dropped (:22).  Returns (1, frames - 1, n_mels); a (B, N) batch returns
    (B, frames - 1, n_mels)."""
    audio = U._as_cuda(audio)
    if padding > 0:
        audio = torch.nn.functional.pad(audio, (0, padding))                  # mx.pad(audio, (0, padding)) :19-20
    mel = U.log_mel_spectrogram(audio, sample_rate=sample_rate, n_fft=n_fft, hop_length=n_fft // 4, n_mels=n_mels,
                                window=U.hanning(n_fft), power=1, log="ln", floor=1e-5, norm=None, mel_scale="htk")
    return mel[None, :-1] if mel.dim() == 2 else mel[:, :-1]
