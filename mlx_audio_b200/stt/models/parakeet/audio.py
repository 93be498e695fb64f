"""Drop-in for mlx_audio/stt/models/parakeet/audio.py: ``PreprocessArgs`` and
``log_mel_spectrogram`` on ``kk_logmel_f32`` (power spectrum, slaney-scale filterbank, ``log(x + 1e-5)`` fused) and
``kk_normalize_features`` (per-feature or per-signal mean / std)."""
from __future__ import annotations

from dataclasses import dataclass

import torch

from .... import _lib as L
from .... import utils as U


@dataclass
class PreprocessArgs:
    """parakeet/audio.py:16-35."""
    sample_rate: int
    normalize: str
    window_size: float
    window_stride: float
    window: str
    features: int
    n_fft: int
    dither: float
    pad_to: int = 0
    pad_value: float = 0

    @property
    def win_length(self) -> int:
        return int(self.window_size * self.sample_rate)

    @property
    def hop_length(self) -> int:
        return int(self.window_stride * self.sample_rate)


def log_mel_spectrogram(x, args: PreprocessArgs) -> torch.Tensor:
    """parakeet/audio.py:38-70: (N,) -> (1, frames, features); extension: (B, N) -> (B, frames, features).
    ``norm=args.normalize`` reaches ``mel_filters`` as is (:52-54): only the literal "slaney" area-normalises."""
    x = U._as_cuda(x)
    if args.pad_to > 0 and x.shape[-1] < args.pad_to:                                        # :41-44
        x = torch.nn.functional.pad(x, (0, args.pad_to - x.shape[-1]), value=args.pad_value)
    fn = U.STR_TO_WINDOW_FN.get(args.window, None)                                           # :46-47
    window = fn(args.win_length) if fn else U.hanning(args.win_length)
    mel = U.log_mel_spectrogram(x, sample_rate=args.sample_rate, n_fft=args.n_fft, hop_length=args.hop_length,
                                n_mels=args.features, window=window, power=2, log="ln_add", floor=1e-5,
                                norm=args.normalize, mel_scale=None)                         # :49-57
    squeeze = mel.dim() == 2
    mel = (mel[None] if squeeze else mel).contiguous()
    B, frames, feats = mel.shape
    L.check(L.load().kk_normalize_features(L.ptr(mel), B, frames, feats, int(args.normalize == "per_feature"), 1e-5,
                                           L.stream()), "kk_normalize_features")             # :59-66
    return mel
