"""Drop-in for the log-mel front end of mlx_audio/stt/models/whisper/audio.py on ``kk_logmel_f32``
(power spectrum, slaney filterbank, log10 fused) + ``kk_clamp_below_max_affine`` (the max - 8 clamp and (x + 4) / 4)."""
from __future__ import annotations

import torch

from .... import _lib as L
from .... import utils as U

# hard-coded audio hyperparameters (audio.py:11-21)
SAMPLE_RATE = 16000
N_FFT = 400
HOP_LENGTH = 160
CHUNK_LENGTH = 30
N_SAMPLES = CHUNK_LENGTH * SAMPLE_RATE
N_FRAMES = N_SAMPLES // HOP_LENGTH
N_SAMPLES_PER_TOKEN = HOP_LENGTH * 2
FRAMES_PER_SECOND = SAMPLE_RATE // HOP_LENGTH
TOKENS_PER_SECOND = SAMPLE_RATE // N_SAMPLES_PER_TOKEN


def pad_or_trim(array, length: int = N_SAMPLES, *, axis: int = -1):
    """audio.py:24-39."""
    array = U._as_cuda(array)
    if array.shape[axis] > length:
        array = array.narrow(axis, 0, length)
    if array.shape[axis] < length:
        pad = [0, 0] * array.dim()
        pad[2 * (array.dim() - 1 - (axis % array.dim())) + 1] = length - array.shape[axis]
        array = torch.nn.functional.pad(array, pad)
    return array


def log_mel_spectrogram(audio, n_mels: int = 80, padding: int = 0) -> torch.Tensor:
    """audio.py:42-83 for array input (file loading is outside the path): (N,) -> (frames, n_mels) with
    frames = N // 160 (the last STFT frame is dropped, :73); (B, N) -> (B, frames, n_mels), normalised per signal."""
    if isinstance(audio, str):
        raise TypeError("audio file loading (stt/utils.load_audio) is outside the rebuilt path: pass the waveform")
    audio = U._as_cuda(audio)
    if padding > 0:
        audio = torch.nn.functional.pad(audio, (0, padding))
    mel = U.log_mel_spectrogram(audio, sample_rate=SAMPLE_RATE, n_fft=N_FFT, hop_length=HOP_LENGTH, n_mels=n_mels,
                                window=U.hanning(N_FFT), power=2, log="log10", floor=1e-10, norm="slaney",
                                mel_scale=None)
    squeeze = mel.dim() == 2
    mel = (mel[None] if squeeze else mel)[:, :-1].contiguous()
    B, frames, _ = mel.shape
    L.check(L.load().kk_clamp_below_max_affine(L.ptr(mel), B, frames * n_mels, 8.0, 4.0, 0.25, L.stream()),
            "kk_clamp_below_max_affine")
    return mel[0] if squeeze else mel
