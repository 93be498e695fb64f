"""Oracle restatement of mlx_audio/utils.py (windows, stft, istft, mel_filters).

TEST INFRASTRUCTURE ONLY - see oracle/__init__.py.  float32 NumPy, reference
operation order; ``dtype=np.float64`` gives the double-precision shadow.
"""
from __future__ import annotations

import math
from typing import Optional, Union

import numpy as np

ArrayOrStr = Union[np.ndarray, str]


# ---- window functions: utils.py:10-47 (evaluated in Python doubles, stored fp32)
def hanning(size: int, dtype=np.float32) -> np.ndarray:  # utils.py:11-14
    return np.array([0.5 * (1 - math.cos(2 * math.pi * n / (size - 1))) for n in range(size)], dtype=dtype)


def hamming(size: int, dtype=np.float32) -> np.ndarray:  # utils.py:18-21
    return np.array([0.54 - 0.46 * math.cos(2 * math.pi * n / (size - 1)) for n in range(size)], dtype=dtype)


def blackman(size: int, dtype=np.float32) -> np.ndarray:  # utils.py:25-34
    return np.array(
        [0.42 - 0.5 * math.cos(2 * math.pi * n / (size - 1)) + 0.08 * math.cos(4 * math.pi * n / (size - 1))
         for n in range(size)], dtype=dtype)


def bartlett(size: int, dtype=np.float32) -> np.ndarray:  # utils.py:38-39
    return np.array([1 - 2 * abs(n - (size - 1) / 2) / (size - 1) for n in range(size)], dtype=dtype)


STR_TO_WINDOW_FN = {  # utils.py:42-48
    "hann": hanning, "hanning": hanning, "hamming": hamming, "blackman": blackman, "bartlett": bartlett,
}


def _resolve_window(window: ArrayOrStr, length: int, periodic: bool, dtype) -> np.ndarray:
    if isinstance(window, str):
        fn = STR_TO_WINDOW_FN.get(window.lower())
        if fn is None:
            raise ValueError(f"Unknown window function: {window}")  # utils.py:69, 120
        # stft: symmetric fn(win_length) (utils.py:70); istft: periodic fn(win_length+1)[:-1] (utils.py:121)
        return fn(length + 1, dtype)[:-1] if periodic else fn(length, dtype)
    return np.asarray(window, dtype=dtype)


def stft(x, n_fft: int = 800, hop_length: Optional[int] = None, win_length: Optional[int] = None,
         window: ArrayOrStr = "hann", center: bool = True, pad_mode: str = "reflect", dtype=np.float32):
    """utils.py:52-101.  1-D x -> (num_frames, n_fft//2+1) complex."""
    x = np.asarray(x, dtype=dtype)
    if hop_length is None:
        hop_length = n_fft // 4
    if win_length is None:
        win_length = n_fft
    w = _resolve_window(window, win_length, periodic=False, dtype=dtype)
    if w.shape[0] < n_fft:  # zero-extend on the right (utils.py:74-76)
        w = np.concatenate([w, np.zeros(n_fft - w.shape[0], dtype)])
    if center:
        p = n_fft // 2
        if pad_mode == "constant":  # utils.py:79-80
            x = np.pad(x, (p, p))
        elif pad_mode == "reflect":  # utils.py:81-84
            prefix = x[1:p + 1][::-1]
            suffix = x[-(p + 1):-1][::-1]
            x = np.concatenate([prefix, x, suffix])
        else:
            raise ValueError(f"Invalid pad_mode {pad_mode}")  # utils.py:86
    num_frames = 1 + (x.shape[0] - n_fft) // hop_length  # utils.py:91
    if num_frames <= 0:
        raise ValueError(
            f"Input is too short (length={x.shape[0]}) for n_fft={n_fft} with "
            f"hop_length={hop_length} and center={center}.")
    frames = np.lib.stride_tricks.as_strided(
        x, shape=(num_frames, n_fft), strides=(hop_length * x.strides[0], x.strides[0]), writeable=False)
    spec = np.fft.rfft(frames * w, axis=-1)  # utils.py:101
    return spec.astype(np.complex64 if dtype == np.float32 else np.complex128)


def istft(x, hop_length: Optional[int] = None, win_length: Optional[int] = None, window: ArrayOrStr = "hann",
          center: bool = True, length: Optional[int] = None, dtype=np.float32):
    """utils.py:104-158.  x (bins, frames) complex -> 1-D signal.  Normaliser is sum(w), NOT sum(w^2)."""
    x = np.asarray(x)
    if win_length is None:
        win_length = (x.shape[1] - 1) * 2  # NB: reference uses shape[1] (utils.py:112-113)
    if hop_length is None:
        hop_length = win_length // 4
    w = _resolve_window(window, win_length, periodic=True, dtype=dtype)
    if w.shape[0] < win_length:
        w = np.concatenate([w, np.zeros(win_length - w.shape[0], dtype)])
    num_frames = x.shape[1]
    t = (num_frames - 1) * hop_length + win_length
    frames_time = np.fft.irfft(x, axis=0).astype(dtype).T  # (frames, n) ; utils.py:135
    idx = (np.arange(num_frames) * hop_length)[:, None] + np.arange(win_length)[None, :]
    rec = np.zeros(t, dtype)
    wsum = np.zeros(t, dtype)
    np.add.at(rec, idx.ravel(), (frames_time * w).astype(dtype).ravel())  # utils.py:142-146
    np.add.at(wsum, idx.ravel(), np.tile(w, num_frames))  # utils.py:147
    nz = wsum != 0
    rec[nz] = rec[nz] / wsum[nz]  # utils.py:150
    if center and length is None:  # utils.py:152-153
        rec = rec[win_length // 2: -win_length // 2]
    if length is not None:  # utils.py:155-156 (no centre trim in this case)
        rec = rec[:length]
    return rec


def _linspace(start, stop, num, dtype):
    # mx.linspace in float32: lerp form start*(1-t) + stop*t with t = i/(num-1).
    t = (np.arange(num, dtype=dtype) / dtype(num - 1)).astype(dtype)
    return (dtype(start) * (dtype(1.0) - t) + dtype(stop) * t).astype(dtype)


def mel_filters(sample_rate: int, n_fft: int, n_mels: int, f_min: float = 0, f_max: Optional[float] = None,
                norm: Optional[str] = None, mel_scale: str = "htk", dtype=np.float32) -> np.ndarray:
    """utils.py:164-237.  Returns (n_mels, n_fft//2+1).  Any mel_scale != "htk" (incl. None) is slaney."""
    dt = np.dtype(dtype).type

    def hz_to_mel(freq):  # python doubles (utils.py:174-186)
        if mel_scale == "htk":
            return 2595.0 * math.log10(1.0 + freq / 700.0)
        f_sp = 200.0 / 3
        mels = (freq - 0.0) / f_sp
        min_log_hz = 1000.0
        min_log_mel = (min_log_hz - 0.0) / f_sp
        logstep = math.log(6.4) / 27.0
        if freq >= min_log_hz:
            mels = min_log_mel + math.log(freq / min_log_hz) / logstep
        return mels

    def mel_to_hz(mels):  # array math in fp32 (utils.py:188-203)
        if mel_scale == "htk":
            return (dt(700.0) * (np.power(dt(10.0), mels / dt(2595.0)) - dt(1.0))).astype(dtype)
        f_sp = 200.0 / 3
        freqs = (dt(0.0) + dt(f_sp) * mels).astype(dtype)
        min_log_hz = 1000.0
        min_log_mel = (min_log_hz - 0.0) / f_sp
        logstep = math.log(6.4) / 27.0
        return np.where(mels >= dt(min_log_mel),
                        dt(min_log_hz) * np.exp(dt(logstep) * (mels - dt(min_log_mel))), freqs).astype(dtype)

    f_max = f_max or sample_rate / 2
    n_freqs = n_fft // 2 + 1
    all_freqs = _linspace(0, sample_rate // 2, n_freqs, dt)  # utils.py:210
    m_pts = _linspace(hz_to_mel(f_min), hz_to_mel(f_max), n_mels + 2, dt)  # utils.py:214-216
    f_pts = mel_to_hz(m_pts)
    f_diff = f_pts[1:] - f_pts[:-1]
    slopes = f_pts[None, :] - all_freqs[:, None]  # (n_freqs, n_mels+2)
    down = (-slopes[:, :-2]) / f_diff[:-1]
    up = slopes[:, 2:] / f_diff[1:]
    fb = np.maximum(np.zeros_like(down), np.minimum(down, up))  # utils.py:226-230
    if norm == "slaney":  # utils.py:232-234
        enorm = dt(2.0) / (f_pts[2:n_mels + 2] - f_pts[:n_mels])
        fb = fb * enorm[None, :]
    return np.ascontiguousarray(fb.T.astype(dtype))  # utils.py:236


# ---- log-mel recipes defined by the reference's callers of utils.stft / mel_filters
def log_mel_vocos(x, sample_rate=24000, n_fft=1024, hop_length=256, n_mels=80, drop_last=False, dtype=np.float32):
    """indextts/mel.py:20-37 and codec/models/vocos/mel.py:22-33 (drop_last=True):
    log(max(abs(stft(x)) @ mel_filters(htk).T, 1e-5)) -> (frames, n_mels)."""
    spec = np.abs(stft(x, n_fft=n_fft, hop_length=hop_length, win_length=n_fft, window="hann", dtype=dtype)).astype(dtype)
    if drop_last:
        spec = spec[:-1]
    fb = mel_filters(sample_rate, n_fft, n_mels, norm=None, mel_scale="htk", dtype=dtype)
    mel = spec @ fb.T
    return np.log(np.maximum(mel, np.dtype(dtype).type(1e-5))).astype(dtype)


def log_mel_whisper(x, n_mels=80, n_fft=400, hop_length=160, sample_rate=16000, dtype=np.float32):
    """stt/models/whisper/audio.py:72-82: power spectrum (last frame dropped), slaney mel,
    log10(max(.,1e-10)), clamp to max-8, (x+4)/4 -> (frames, n_mels)."""
    dt = np.dtype(dtype).type
    freqs = stft(x, n_fft=n_fft, hop_length=hop_length, win_length=n_fft, window="hann", dtype=dtype)
    mag = (np.abs(freqs[:-1, :]).astype(dtype)) ** 2
    fb = mel_filters(sample_rate, n_fft, n_mels, norm="slaney", mel_scale=None, dtype=dtype)
    mel = mag @ fb.T
    log_spec = np.log10(np.maximum(mel, dt(1e-10))).astype(dtype)
    log_spec = np.maximum(log_spec, log_spec.max() - dt(8.0))
    return ((log_spec + dt(4.0)) / dt(4.0)).astype(dtype)


def istft_head_vocos(x, weight, bias, n_fft: int, hop_length: int, dtype=np.float32) -> np.ndarray:
    """codec/models/vocos/vocos.py:119-140 ``ISTFTHead``: x (1, L, dim); Linear(dim, n_fft + 2) -> [log-mag | phase]
    -> min(exp(mag), 1e2) * (cos p + i sin p) -> istft with the SYMMETRIC hanning(n_fft) window (passed as an array, so
    utils.istft does not make it periodic), hop_length, win_length = n_fft.  Returns the 1-D signal."""
    x = np.asarray(x, dtype)
    h = (x @ np.asarray(weight, dtype).T + np.asarray(bias, dtype)).astype(dtype)     # nn.Linear :124
    h = np.swapaxes(h, 1, 2)                                                           # (1, n_fft + 2, L)
    mag, p = np.split(h, 2, axis=1)
    mag = np.minimum(np.exp(mag), dtype(1e2))                                          # :129-130
    S = mag * (np.cos(p) + 1j * np.sin(p))
    return istft(S[0].astype(np.complex64), window=hanning(n_fft, dtype), hop_length=hop_length, win_length=n_fft,
                 dtype=dtype)


def log_mel_parakeet(x, sample_rate=16000, n_fft=512, win_length=400, hop_length=160, features=80, window="hann",
                     normalize="per_feature", dtype=np.float32) -> np.ndarray:
    """stt/models/parakeet/audio.py:38-70: power spectrum, mel_filters(norm=normalize, mel_scale=None), log(x + 1e-5),
    (x - mean) / (std + 1e-5) per feature over time (or over everything) -> (1, frames, features)."""
    fn = STR_TO_WINDOW_FN.get(window)
    w = fn(win_length, dtype) if fn else hanning(win_length, dtype)
    S = stft(x, n_fft, hop_length, win_length, w, dtype=dtype)
    p = (np.abs(S).astype(dtype)) ** 2
    fb = mel_filters(sample_rate, n_fft, features, norm=normalize, mel_scale=None, dtype=dtype)
    m = np.log((fb @ p.T).astype(dtype) + dtype(1e-5))                      # (features, frames)
    if normalize == "per_feature":
        mean, std = m.mean(axis=1, keepdims=True), m.std(axis=1, keepdims=True)
    else:
        mean, std = m.mean(), m.std()
    return ((m - mean) / (std + dtype(1e-5))).T[None].astype(dtype)


def log_mel_s3(x, sample_rate=16000, n_mels=128, n_fft=400, hop_length=160, dtype=np.float32) -> np.ndarray:
    """codec/models/s3/utils.py:8-42: periodic Hann, power spectrum (all frames), slaney mel, log10(max(., 1e-10)),
    clamp to max - 8, (x + 4) / 4 -> (n_mels, frames)."""
    dt = np.dtype(dtype).type
    w = hanning(n_fft + 1, dtype)[:-1]
    S = stft(x, n_fft=n_fft, hop_length=hop_length, win_length=n_fft, window=w, dtype=dtype).T
    mag = (np.abs(S).astype(dtype)) ** 2
    fb = mel_filters(sample_rate, n_fft, n_mels, norm="slaney", mel_scale="slaney", dtype=dtype)
    ls = np.log10(np.maximum((fb @ mag).astype(dtype), dt(1e-10))).astype(dtype)
    ls = np.maximum(ls, ls.max() - dt(8.0))
    return ((ls + dt(4.0)) / dt(4.0)).astype(dtype)


def mel_spark(x, sample_rate=16000, n_mels=128, n_fft=1024, f_min=10, f_max=None, hop_length=320, win_length=640,
              dtype=np.float32) -> np.ndarray:
    """tts/models/spark/bicodec.py:22-52: periodic Hann(win_length) zero-extended to n_fft, |X| @ slaney mel (f_min 10)
    -> (1, frames, n_mels); no logarithm."""
    w = hanning(win_length + 1, dtype)[:-1]
    S = stft(x, n_fft=n_fft, hop_length=hop_length, win_length=win_length, window=w, dtype=dtype)
    fb = mel_filters(sample_rate, n_fft, n_mels, f_min=f_min, f_max=f_max, norm="slaney", mel_scale="slaney", dtype=dtype)
    return (np.abs(S).astype(dtype) @ fb.T).astype(dtype)[None]


def snake(x, alpha, beta=None, logscale=False, channels_first=False, eps=1e-9, dtype=np.float32) -> np.ndarray:
    """codec/models/bigvgan/activation.py:16-24 (Snake: channels-first), :41-51 (SnakeBeta: channels-last, beta),
    codec/models/descript/nn/layers.py:116-119: x + 1 / (b + eps) * sin(a x)^2, a / b = exp(.) in log-scale mode."""
    x = np.asarray(x, dtype)
    shp = (1, -1, 1) if channels_first else (1, 1, -1)
    a = np.asarray(alpha, dtype).reshape(shp)
    b = a if beta is None else np.asarray(beta, dtype).reshape(shp)
    if logscale:
        a, b = np.exp(a), np.exp(b)
    return (x + (dtype(1.0) / (b + dtype(eps))) * np.power(np.sin(x * a), 2)).astype(dtype)
