"""Drop-in for ``log_mel_spectrogram`` of mlx_audio/codec/models/s3/utils.py:8-42 on the fused
log-mel kernel: periodic Hann(400), hop 160, power spectrum, slaney filterbank, log10, max - 8 clamp, (x + 4) / 4."""
from __future__ import annotations

import torch

from .... import _lib as L
from .... import utils as U


def log_mel_spectrogram(audio, sample_rate: int = 16_000, n_mels: int = 128, n_fft: int = 400, hop_length: int = 160,
                        padding: int = 0) -> torch.Tensor:
    """(N,) -> (n_mels, frames) like the reference (every frame kept, mel-major); (B, N) -> (B, n_mels, frames)."""
    audio = U._as_cuda(audio)
    if padding > 0:
        audio = torch.nn.functional.pad(audio, (0, padding))
    window = U.hanning(n_fft + 1)[:-1]                                   # s3/utils.py:22 (periodic)
    mel = U.log_mel_spectrogram(audio, sample_rate=sample_rate, n_fft=n_fft, hop_length=hop_length, n_mels=n_mels,
                                window=window, power=2, log="log10", floor=1e-10, norm="slaney", mel_scale="slaney")
    squeeze = mel.dim() == 2
    mel = (mel[None] if squeeze else mel).contiguous()
    B, frames, _ = mel.shape
    L.check(L.load().kk_clamp_below_max_affine(L.ptr(mel), B, frames * n_mels, 8.0, 4.0, 0.25, L.stream()),
            "kk_clamp_below_max_affine")                                 # :40-41
    mel = mel.transpose(1, 2)
    return mel[0] if squeeze else mel
