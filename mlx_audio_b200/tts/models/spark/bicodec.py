"""Drop-in for ``mel_spectrogram`` of mlx_audio/tts/models/spark/bicodec.py:22-52 on the fused
STFT -> |X| -> mel kernel (no logarithm): periodic Hann(640) zero-extended to n_fft 1024, hop 320, slaney filterbank
from f_min 10 Hz.  The BiCodec networks themselves are outside the rebuilt path."""
from __future__ import annotations

from typing import Optional

import torch

from .... import utils as U


def mel_spectrogram(audio, sample_rate: int = 16_000, n_mels: int = 128, n_fft: int = 1024, f_min: int = 10,
                    f_max: Optional[int] = None, hop_length: int = 320, win_length: int = 640,
                    padding: int = 0) -> torch.Tensor:
    """(N,) -> (1, frames, n_mels); (B, N) -> (B, frames, n_mels)."""
    audio = U._as_cuda(audio)
    if padding > 0:
        audio = torch.nn.functional.pad(audio, (0, padding))
    window = U.hanning(win_length + 1)[:-1]                              # bicodec.py:37; stft zero-extends it (utils.py:74-76)
    mel = U.log_mel_spectrogram(audio, sample_rate=sample_rate, n_fft=n_fft, hop_length=hop_length, n_mels=n_mels,
                                window=window, power=1, log=None, norm="slaney", mel_scale="slaney", f_min=f_min,
                                f_max=f_max)
    return mel[None] if mel.dim() == 2 else mel
