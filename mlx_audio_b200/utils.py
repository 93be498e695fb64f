"""B200-native drop-in for mlx_audio/utils.py: window functions, ``stft``, ``istft``, ``mel_filters``.

Same names, argument order, defaults, output shapes and error behaviour as the reference
(mlx_audio/utils.py:11-237); tensors are torch CUDA tensors instead of mx.arrays.  The transforms run in
hand-written shared-memory FFT kernels (csrc/fft.cu) through the C ABI; there is no CPU fallback.

Extensions: ``stft`` / ``istft`` accept a leading batch dimension, and ``log_mel_spectrogram`` exposes the
fused STFT -> |X| / |X|^2 -> mel -> log kernel used by the reference's log-mel front ends
(indextts/mel.py:20-37, codec/models/vocos/mel.py:22-33, stt/models/whisper/audio.py:72-82).
"""
from __future__ import annotations

import math
from functools import lru_cache
from typing import Optional, Union

import numpy as np
import torch

from . import _lib as L


def _dev() -> torch.device:
    L.require_cuda()
    return torch.device("cuda", torch.cuda.current_device())


def _window_tensor(values) -> torch.Tensor:
    return torch.tensor(np.asarray(values, dtype=np.float32), device=_dev())


# ---- window functions (utils.py:10-47): evaluated in Python doubles, stored as float32
@lru_cache(maxsize=None)
def hanning(size):
    return _window_tensor([0.5 * (1 - math.cos(2 * math.pi * n / (size - 1))) for n in range(size)])


@lru_cache(maxsize=None)
def hamming(size):
    return _window_tensor([0.54 - 0.46 * math.cos(2 * math.pi * n / (size - 1)) for n in range(size)])


@lru_cache(maxsize=None)
def blackman(size):
    return _window_tensor([0.42 - 0.5 * math.cos(2 * math.pi * n / (size - 1))
                           + 0.08 * math.cos(4 * math.pi * n / (size - 1)) for n in range(size)])


@lru_cache(maxsize=None)
def bartlett(size):
    return _window_tensor([1 - 2 * abs(n - (size - 1) / 2) / (size - 1) for n in range(size)])


STR_TO_WINDOW_FN = {"hann": hanning, "hanning": hanning, "hamming": hamming, "blackman": blackman,
                    "bartlett": bartlett}

_PAD_MODES = {"constant": 2, "reflect": 1}


def _as_cuda(x, dtype=torch.float32) -> torch.Tensor:
    if isinstance(x, torch.Tensor):
        return x.to(device=_dev(), dtype=dtype)
    return torch.as_tensor(np.asarray(x)).to(device=_dev(), dtype=dtype)


def _resolve_window(window, length: int, periodic: bool) -> torch.Tensor:
    if isinstance(window, str):
        fn = STR_TO_WINDOW_FN.get(window.lower())
        if fn is None:
            raise ValueError(f"Unknown window function: {window}")
        return fn(length + 1)[:-1] if periodic else fn(length)
    return _as_cuda(window)


def _prep_stft(x, n_fft, hop_length, win_length, window, center, pad_mode):
    if hop_length is None:
        hop_length = n_fft // 4
    if win_length is None:
        win_length = n_fft
    w = _resolve_window(window, win_length, periodic=False)
    if w.shape[0] < n_fft:  # zero-extend on the right (utils.py:74-76)
        w = torch.cat([w, torch.zeros(n_fft - w.shape[0], device=w.device)])
    if w.shape[0] != n_fft:
        raise ValueError(f"window length {w.shape[0]} exceeds n_fft={n_fft}")
    if center:
        if pad_mode not in _PAD_MODES:
            raise ValueError(f"Invalid pad_mode {pad_mode}")
        mode = _PAD_MODES[pad_mode]
    else:
        mode = 0
    x = _as_cuda(x)
    squeeze = x.dim() == 1
    if squeeze:
        x = x[None]
    if x.dim() != 2:
        raise ValueError(f"stft expects a 1-D signal (or a (B, N) batch), got shape {tuple(x.shape)}")
    x = x.contiguous()
    n = x.shape[1]
    padded = n + (2 * (n_fft // 2) if center else 0)
    num_frames = 1 + (padded - n_fft) // hop_length
    if num_frames <= 0:
        raise ValueError(f"Input is too short (length={padded}) for n_fft={n_fft} with "
                         f"hop_length={hop_length} and center={center}.")
    if mode == 1 and n <= n_fft // 2:
        raise ValueError(f"Input is too short (length={n}) for reflect padding of {n_fft // 2}.")
    return x, w.contiguous(), hop_length, mode, num_frames, squeeze


def _ragged_lengths(lengths, B: int, limit: int, what: str) -> Optional[torch.Tensor]:
    """``lengths=`` of the batched transforms -> (B,) int32 on the device (None stays None).  Host-resident lengths are
    validated; device-resident ones are trusted (the kernels clamp them to [0, limit])."""
    if lengths is None:
        return None
    if isinstance(lengths, torch.Tensor) and lengths.is_cuda:
        if lengths.numel() != B:
            raise ValueError(f"lengths: expected {B} entries, got {lengths.numel()}")
        return lengths.to(torch.int32).contiguous()
    host = np.asarray(lengths.cpu() if isinstance(lengths, torch.Tensor) else lengths).astype(np.int64).reshape(-1)
    if host.size != B:
        raise ValueError(f"lengths: expected {B} entries, got {host.size}")
    if host.size and (host.min() < 0 or host.max() > limit):
        raise ValueError(f"lengths must lie in [0, {limit}] ({what}), got [{host.min()}, {host.max()}]")
    return torch.from_numpy(host.astype(np.int32)).to(_dev())


def stft(x, n_fft=800, hop_length=None, win_length=None, window: Union[torch.Tensor, str] = "hann", center=True,
         pad_mode="reflect", lengths=None):
    """utils.py:52-101.  (N,) -> (num_frames, n_fft//2+1) complex64 [batched: (B, N) -> (B, frames, bins)].
    ``lengths`` (batched calls; not in the reference, which transforms one signal per call): valid samples per signal
    of a zero-padded ragged batch.  Row b equals ``stft(x[b, :lengths[b]])`` - its own reflect padding and last frame -
    followed by zero frames (all zeros when the signal is too short to transform)."""
    x, w, hop, mode, frames, squeeze = _prep_stft(x, n_fft, hop_length, win_length, window, center, pad_mode)
    B, n = x.shape
    lens = _ragged_lengths(lengths, B, n, "samples")
    out = torch.empty(B, frames, n_fft // 2 + 1, 2, dtype=torch.float32, device=x.device)
    L.check(L.load().kk_stft_f32(L.ptr(x), L.ptr(w), L.ptr(out), B, n, n_fft, hop, mode, frames,
                                 None if lens is None else L.ptr(lens), L.stream()), "kk_stft_f32")
    spec = torch.view_as_complex(out)
    return spec[0] if squeeze else spec


def istft(x, hop_length=None, win_length=None, window="hann", center=True, length=None, lengths=None):
    """utils.py:104-158.  x (bins, frames) complex -> 1-D signal [batched: (B, bins, frames) -> (B, len)].
    Normalised by the window sum (not its square), exactly like the reference.
    ``lengths`` (batched calls; not in the reference): valid frames per signal of a ragged batch.  Row b equals
    ``istft(x[b, :, :lengths[b]], win_length=..., ...)`` followed by zeros."""
    if not isinstance(x, torch.Tensor):
        x = torch.as_tensor(np.asarray(x))
    x = x.to(device=_dev())
    if not x.is_complex():
        x = x.to(torch.complex64)
    x = x.to(torch.complex64)
    squeeze = x.dim() == 2
    if squeeze:
        x = x[None]
    if x.dim() != 3:
        raise ValueError(f"istft expects (bins, frames) or (B, bins, frames), got {tuple(x.shape)}")
    B, bins, frames = x.shape
    n_fft = (bins - 1) * 2
    if win_length is None:
        win_length = (frames - 1) * 2  # sic: the reference derives it from shape[1] (utils.py:112-113)
    if hop_length is None:
        hop_length = win_length // 4
    w = _resolve_window(window, win_length, periodic=True)
    if w.shape[0] < win_length:
        w = torch.cat([w, torch.zeros(win_length - w.shape[0], device=w.device)])
    if w.shape[0] != n_fft:
        raise ValueError(f"window length {w.shape[0]} does not match the frame length {n_fft}")
    # the kernel reads frames-major (B, frames, bins); a `.T` view of an stft result already is
    xt = x.transpose(1, 2)
    if not xt.is_contiguous():
        xt = xt.contiguous()
    total = (frames - 1) * hop_length + win_length
    trim, out_len = 0, total
    if center and length is None:  # utils.py:152-153
        trim = win_length // 2
        out_len = len(range(total)[win_length // 2: -win_length // 2])
    if length is not None:  # utils.py:155-156: no centre trim in this case
        out_len = min(length, total)
    if out_len <= 0:
        raise ValueError("istft: empty output")
    lens = _ragged_lengths(lengths, B, frames, "frames")
    tail = -(-win_length // 2) if (center and length is None) else 0      # range(total)[w//2 : -w//2]
    out = torch.empty(B, out_len, dtype=torch.float32, device=x.device)
    spec = torch.view_as_real(xt)
    L.check(L.load().kk_istft_f32(L.ptr(spec), L.ptr(w.contiguous()), L.ptr(out), B, frames, n_fft, hop_length, trim,
                                  out_len, None if lens is None else L.ptr(lens), tail, L.stream()), "kk_istft_f32")
    return out[0] if squeeze else out


@lru_cache(maxsize=None)
def _mel_filters_np(sample_rate, n_fft, n_mels, f_min, f_max, norm, mel_scale):
    f32 = np.float32

    def hz_to_mel(freq):
        if mel_scale == "htk":
            return 2595.0 * math.log10(1.0 + freq / 700.0)
        f_sp = 200.0 / 3
        mels = freq / f_sp
        if freq >= 1000.0:
            mels = 1000.0 / f_sp + math.log(freq / 1000.0) / (math.log(6.4) / 27.0)
        return mels

    def mel_to_hz(mels):
        if mel_scale == "htk":
            return (f32(700.0) * (np.power(f32(10.0), mels / f32(2595.0)) - f32(1.0))).astype(f32)
        f_sp = 200.0 / 3
        min_log_mel = 1000.0 / f_sp
        logstep = math.log(6.4) / 27.0
        lin = (f32(f_sp) * mels).astype(f32)
        return np.where(mels >= f32(min_log_mel), f32(1000.0) * np.exp(f32(logstep) * (mels - f32(min_log_mel))),
                        lin).astype(f32)

    def linspace(a, b, num):
        t = np.arange(num, dtype=f32) / f32(num - 1)
        return (f32(a) * (f32(1) - t) + f32(b) * t).astype(f32)

    f_max = f_max or sample_rate / 2
    n_freqs = n_fft // 2 + 1
    all_freqs = linspace(0, sample_rate // 2, n_freqs)
    f_pts = mel_to_hz(linspace(hz_to_mel(f_min), hz_to_mel(f_max), n_mels + 2))
    f_diff = f_pts[1:] - f_pts[:-1]
    slopes = f_pts[None, :] - all_freqs[:, None]
    down = (-slopes[:, :-2]) / f_diff[:-1]
    up = slopes[:, 2:] / f_diff[1:]
    fb = np.maximum(np.zeros_like(down), np.minimum(down, up))
    if norm == "slaney":
        fb = fb * (f32(2.0) / (f_pts[2:n_mels + 2] - f_pts[:n_mels]))[None, :]
    return np.ascontiguousarray(fb.T.astype(f32))


def mel_filters(sample_rate: int, n_fft: int, n_mels: int, f_min: float = 0, f_max: Optional[float] = None,
                norm: Optional[str] = None, mel_scale: str = "htk") -> torch.Tensor:
    """utils.py:164-237: triangular filterbank (n_mels, n_fft//2+1); any mel_scale other than "htk" is slaney."""
    fb = _mel_filters_np(sample_rate, n_fft, n_mels, f_min, f_max, norm, mel_scale)
    return torch.from_numpy(fb).to(_dev())


def log_mel_spectrogram(x, sample_rate=24000, n_fft=1024, hop_length=256, n_mels=80, window="hann", power=1,
                        log="ln", floor=1e-5, norm=None, mel_scale="htk", center=True, pad_mode="reflect",
                        f_min=0, f_max=None, lengths=None):
    """Fused STFT -> |X|^power -> mel -> log.  Defaults = indextts/mel.py:20-37 (``log(max(|X| @ fb.T, 1e-5))``).
    Returns (frames, n_mels) [batched: (B, frames, n_mels)]; the spectrogram is never written to HBM.
    ``lengths``: valid samples per signal of a ragged batch (see ``stft``); frames past a signal's end are zeros."""
    x, w, hop, mode, frames, squeeze = _prep_stft(x, n_fft, hop_length, n_fft, window, center, pad_mode)
    fb = mel_filters(sample_rate, n_fft, n_mels, f_min, f_max, norm, mel_scale).contiguous()
    B, n = x.shape
    lens = _ragged_lengths(lengths, B, n, "samples")
    out = torch.empty(B, frames, n_mels, dtype=torch.float32, device=x.device)
    log_mode = {None: 0, "none": 0, "ln": 1, "log": 1, "log10": 2, "ln_add": 3}[log]   # ln_add: log(x + floor)
    L.check(L.load().kk_logmel_f32(L.ptr(x), L.ptr(w), L.ptr(fb), L.ptr(out), B, n, n_fft, hop, mode, frames, n_mels,
                                   int(power), log_mode, float(floor), None if lens is None else L.ptr(lens),
                                   L.stream()), "kk_logmel_f32")
    return out[0] if squeeze else out
