"""Oracle restatement of mlx_audio/tts/models/kokoro/istftnet.py (Decoder and below).

TEST INFRASTRUCTURE ONLY - see oracle/__init__.py.

Functional float32 restatement over a flat ``{name: ndarray}`` weight dict that
uses the reference's own parameter names and MLX layouts (weight_v is (O,K,I)).
Elementwise / bit-sensitive parts (the harmonic source) are NumPy float32 with
the reference's operation order; the dense convolutions use torch-CPU
``conv1d`` / ``conv_transpose1d`` (float32, or float64 for the shadow run).

Tensors between blocks are channels-first (B, C, L) exactly where the reference
holds them channels-first.
"""
from __future__ import annotations

import math
from typing import Dict, Optional

import numpy as np
import torch
import torch.nn.functional as F

from . import dsp
from .interp import interpolate, interpolate1d

_DEFAULT_CFG = dict(
    dim_in=512, style_dim=128, dim_out=80,
    upsample_rates=[10, 6], upsample_kernel_sizes=[20, 12], gen_istft_n_fft=20, gen_istft_hop_size=5,
    resblock_kernel_sizes=[3, 7, 11], resblock_dilation_sizes=[[1, 3, 5]] * 3, upsample_initial_channel=512,
)


def get_padding(kernel_size: int, dilation: int = 1) -> int:  # istftnet.py:14-15
    return int((kernel_size * dilation - dilation) / 2)


def weight_norm(v: torch.Tensor, g: torch.Tensor) -> torch.Tensor:
    """istftnet.py:53-93 with dim=0 (:130): n = sqrt(sum(v*v, axes=(1,2))); w = v / (n + 1e-7) * g."""
    n = torch.sqrt(torch.sum(v * v, dim=(1, 2), keepdim=True))
    return (v / (n + 1e-7)) * g


def conv1d_mlx(x_blc, w_oki, stride=1, padding=0, dilation=1, groups=1):
    """mx.conv1d: x (B,L,Cin), w (O,K,Cin/groups) -> (B,Lout,O).  Cross-correlation, zero padding."""
    y = F.conv1d(x_blc.transpose(1, 2), w_oki.permute(0, 2, 1), None, stride, padding, dilation, groups)
    return y.transpose(1, 2)


def conv_transpose1d_mlx(x_blc, w_oki, stride=1, padding=0, groups=1):
    """mx.conv_transpose1d: x (B,L,Cin), w (Cout,K,Cin/groups) -> (B,(L-1)s-2p+K,Cout)."""
    if groups == 1:
        wt = w_oki.permute(2, 0, 1)  # torch (Cin, Cout, K)
    else:  # depthwise: (C,K,1) -> torch (C,1,K)
        wt = w_oki.permute(0, 2, 1)
    y = F.conv_transpose1d(x_blc.transpose(1, 2), wt, None, stride, padding, 0, groups)
    return y.transpose(1, 2)


def leaky_relu(x, slope):  # istftnet.py:692-693 / nn.LeakyReLU
    return torch.where(x > 0, x, x * slope)


class DecoderOracle:
    """Restatement of ``Decoder`` (istftnet.py:902-963) and everything it calls."""

    def __init__(self, weights: Dict[str, np.ndarray], cfg: Optional[dict] = None, dtype=torch.float32):
        self.cfg = dict(_DEFAULT_CFG if cfg is None else cfg)
        self.dtype = dtype
        self.w = {k: torch.from_numpy(np.ascontiguousarray(v)).to(dtype) for k, v in weights.items()}
        self.w32 = {k: np.asarray(v, np.float32) for k, v in weights.items()
                    if "m_source" in k}
        self.taps: Dict[str, np.ndarray] = {}
        self.keep_taps = False
        c = self.cfg
        self.upsample_scale = int(np.prod(c["upsample_rates"])) * c["gen_istft_hop_size"]  # :714
        self.n_fft, self.hop = c["gen_istft_n_fft"], c["gen_istft_hop_size"]

    # ------------------------------------------------------------------ helpers
    def _tap(self, name, t):
        if self.keep_taps:
            self.taps[name] = (t.detach().cpu().numpy() if isinstance(t, torch.Tensor) else np.asarray(t)).copy()

    def _wn(self, prefix):
        return weight_norm(self.w[f"{prefix}.weight_v"], self.w[f"{prefix}.weight_g"])

    def _conv_weighted(self, prefix, x_bcl, stride=1, padding=1, dilation=1):
        """ConvWeighted.__call__(x.swapaxes, mx.conv1d).swapaxes (istftnet.py:128-166)."""
        w = self._wn(prefix)
        y = conv1d_mlx(x_bcl.transpose(1, 2), w, stride, padding, dilation)
        b = self.w.get(f"{prefix}.bias")
        if b is not None:
            y = y + b.reshape(1, 1, -1)
        return y.transpose(1, 2)

    def _adain(self, prefix, x_bcl, s):
        """AdaIN1d (istftnet.py:327-338) + InstanceNorm1d (:216-255)."""
        h = s @ self.w[f"{prefix}.fc.weight"].T + self.w[f"{prefix}.fc.bias"]
        h = h.unsqueeze(2)
        gamma, beta = torch.chunk(h, 2, dim=1)
        mean = x_bcl.mean(dim=2, keepdim=True)
        var = x_bcl.var(dim=2, unbiased=False, keepdim=True)
        xn = (x_bcl - mean) / torch.sqrt(var + 1e-5)
        return (1 + gamma) * xn + beta

    # ------------------------------------------------------- AdaINResBlock1 :341
    def _adain_resblock1(self, prefix, x, s, k, dils=(1, 3, 5)):
        for i in range(3):
            a1 = self.w[f"{prefix}.alpha1.{i}"]
            a2 = self.w[f"{prefix}.alpha2.{i}"]
            xt = self._adain(f"{prefix}.adain1.{i}", x, s)
            xt = xt + (1 / a1) * (torch.sin(a1 * xt) ** 2)  # Snake1D :382
            xt = self._conv_weighted(f"{prefix}.convs1.{i}", xt, 1, get_padding(k, dils[i]), dils[i])
            xt = self._adain(f"{prefix}.adain2.{i}", xt, s)
            xt = xt + (1 / a2) * (torch.sin(a2 * xt) ** 2)
            xt = self._conv_weighted(f"{prefix}.convs2.{i}", xt, 1, get_padding(k, 1), 1)
            x = xt + x
        return x

    # --------------------------------------------------------- AdainResBlk1d :825
    def _adain_resblk1d(self, prefix, x, s, upsample=False):
        learned_sc = f"{prefix}.conv1x1.weight_v" in self.w
        # residual (:874-894)
        r = self._adain(f"{prefix}.norm1", x, s)
        r = leaky_relu(r, 0.2)
        if upsample:
            w = self._wn(f"{prefix}.pool")  # (C,3,1), groups=C, stride 2, padding 1
            rb = conv_transpose1d_mlx(r.transpose(1, 2), w, 2, 1, groups=w.shape[0])
            rb = rb + self.w[f"{prefix}.pool.bias"].reshape(1, 1, -1)
            rb = F.pad(rb, (0, 0, 1, 0))  # left zero-pad one time step (:881)
            r = rb.transpose(1, 2)
        r = self._conv_weighted(f"{prefix}.conv1", r, 1, 1, 1)
        r = self._adain(f"{prefix}.norm2", r, s)
        r = leaky_relu(r, 0.2)
        r = self._conv_weighted(f"{prefix}.conv2", r, 1, 1, 1)
        # shortcut (:863-872)
        sc = x
        if upsample:
            sc = torch.repeat_interleave(sc, 2, dim=2)  # nn.Upsample(scale 2, nearest)
        if learned_sc:
            sc = self._conv_weighted(f"{prefix}.conv1x1", sc, 1, 0, 1)
        return (r + sc) / math.sqrt(2)  # :898  (mx.sqrt(2) -> float32 sqrt(2))

    # ------------------------------------------------ SineGen / SourceModuleHnNSF
    def harmonic_source(self, f0_frames: np.ndarray, noise: np.ndarray,
                        rand_ini: Optional[np.ndarray] = None) -> np.ndarray:
        """f0 (B,2T) float32, noise (B,600T,9) -> har_source (B,600T) float32.

        Generator.__call__ :770-772, SourceModuleHnNSF :668-680, SineGen :555-623,
        interpolate.py:82-106.  All float32, every op rounded separately."""
        f32 = np.float32
        f0_frames = np.asarray(f0_frames, f32)
        up = self.upsample_scale
        B, L2 = f0_frames.shape
        f0 = np.repeat(f0_frames, up, axis=1)[:, :, None]  # nn.Upsample nearest x300 (:770) -> (B,600T,1)
        harm = np.arange(1, 10).astype(f32)[None, None, :]
        fn = (f0 * harm).astype(f32)  # :610
        rad = np.mod((fn / f32(24000)).astype(f32), f32(1)).astype(f32)  # :561
        if rand_ini is not None:  # :563-565 (provably dead: sample 0 is never read by the /300 resample)
            ri = np.asarray(rand_ini, f32).copy()
            ri[:, 0] = 0
            rad[:, 0, :] = rad[:, 0, :] + ri
        # :568-572 linear down-sample by 1/upsample_scale (scale is an mx float32 scalar)
        rad_ds = interpolate(rad.transpose(0, 2, 1), scale_factor=f32(1) / f32(up), mode="linear").transpose(0, 2, 1)
        assert rad_ds.shape[1] == L2, (rad_ds.shape, L2)
        csum = np.cumsum(rad_ds, axis=1, dtype=f32)  # sequential float32 (:573)
        phase = ((csum * f32(2)).astype(f32) * f32(np.pi)).astype(f32)
        phase = (phase.transpose(0, 2, 1) * f32(up)).astype(f32)  # :575
        phase = interpolate1d(phase, L2 * up, mode="linear").transpose(0, 2, 1)  # :574-578
        sines = np.sin(phase).astype(f32)  # :579
        sine_waves = (sines * f32(0.1)).astype(f32)  # :613
        uv = (f0 > f32(10)).astype(f32)  # :553, voiced_threshold 10 (:716)
        noise_amp = (uv * f32(0.003) + ((f32(1) - uv) * f32(0.1)) / f32(3)).astype(f32)  # :619
        nz = (noise_amp * np.asarray(noise, f32)).astype(f32)  # :620
        sine_waves = (sine_waves * uv + nz).astype(f32)  # :622
        lw = self.w32["decoder.generator.m_source.l_linear.weight"]
        lb = self.w32["decoder.generator.m_source.l_linear.bias"]
        merged = np.tanh((sine_waves @ lw.T + lb).astype(f32)).astype(f32)  # :677
        return merged[:, :, 0]

    def stft_transform(self, har_source: np.ndarray):
        """MLXSTFT.transform (:463-495): per item stft(n_fft,hop,win,'hann',center,reflect).T -> |X|, angle."""
        mags, phases = [], []
        for b in range(har_source.shape[0]):
            X = dsp.stft(har_source[b], n_fft=self.n_fft, hop_length=self.hop, win_length=self.n_fft,
                         window="hann", center=True, pad_mode="reflect").T  # (bins, frames)
            re = X.real.astype(np.float32)
            im = X.imag.astype(np.float32).copy()
            im[0, :] = 0.0   # r2c: DC / Nyquist imaginary parts are +0
            im[-1, :] = 0.0
            mags.append(np.abs(re + 1j * im).astype(np.float32))  # mx.abs(complex64) :484
            phases.append(np.arctan2(im, re).astype(np.float32))  # mlx_angle :399-414
        return np.stack(mags), np.stack(phases)

    def stft_inverse(self, spec: np.ndarray, phase: np.ndarray) -> np.ndarray:
        """MLXSTFT.inverse (:497-523).  mlx_unwrap is the identity for |phase| <= 1 (:442)."""
        out = []
        for b in range(spec.shape[0]):
            ph = _unwrap(phase[b])
            re = spec[b] * np.cos(ph)
            im = spec[b] * np.sin(ph)
            X = (re + 1j * im)
            out.append(dsp.istft(X, hop_length=self.hop, win_length=self.n_fft, window="hann", center=True,
                                 dtype=spec.dtype.type))
        return np.stack(out)[:, None, :]

    # ----------------------------------------------------------- Generator :769
    def generator(self, x, s, f0_curve: np.ndarray, noise: np.ndarray):
        c = self.cfg
        g = "decoder.generator"
        har_source = self.harmonic_source(f0_curve, noise)
        self._tap("har_source", har_source)
        mag, ph = self.stft_transform(har_source)
        har = np.concatenate([mag, ph], axis=1)  # (B,22,frames) :774
        self._tap("har", har)
        har_t = torch.from_numpy(har).to(self.dtype)
        n_up = len(c["upsample_rates"])
        n_k = len(c["resblock_kernel_sizes"])
        for i in range(n_up):
            x = leaky_relu(x, 0.1)  # :777
            nc_w = self.w[f"{g}.noise_convs.{i}.weight"]
            nc_b = self.w[f"{g}.noise_convs.{i}.bias"]
            if i + 1 < n_up:
                stride_f0 = int(np.prod(c["upsample_rates"][i + 1:]))
                xs_ = conv1d_mlx(har_t.transpose(1, 2), nc_w, stride_f0, (stride_f0 + 1) // 2) + nc_b
                nk = 7
            else:
                xs_ = conv1d_mlx(har_t.transpose(1, 2), nc_w, 1, 0) + nc_b
                nk = 11
            x_source = xs_.transpose(1, 2)
            self._tap(f"noise_conv{i}", x_source)
            x_source = self._adain_resblock1(f"{g}.noise_res.{i}", x_source, s, nk)
            self._tap(f"noise_res{i}", x_source)
            # ups[i]: ConvWeighted(encode=True) driven with mx.conv_transpose1d; weight.T -> (Cout,K,Cin) :161-166, :783
            u, k = c["upsample_rates"][i], c["upsample_kernel_sizes"][i]
            w = self._wn(f"{g}.ups.{i}").permute(2, 1, 0)  # full transpose of (O,K,I)
            xu = conv_transpose1d_mlx(x.transpose(1, 2), w, u, (k - u) // 2)
            xu = xu + self.w[f"{g}.ups.{i}.bias"].reshape(1, 1, -1)
            x = xu.transpose(1, 2)
            if i == n_up - 1:
                x = F.pad(x, (1, 0))  # "reflection_pad" is a ZERO left pad (:683-689, :786-787)
            self._tap(f"ups{i}", x)
            x = x + x_source
            xs = None
            for j in range(n_k):
                r = self._adain_resblock1(f"{g}.resblocks.{i * n_k + j}", x, s, c["resblock_kernel_sizes"][j],
                                          c["resblock_dilation_sizes"][j])
                xs = r if xs is None else xs + r
            x = xs / n_k  # :796
            self._tap(f"mrf{i}", x)
        x = leaky_relu(x, 0.01)  # :798
        x = self._conv_weighted(f"{g}.conv_post", x, 1, 3, 1)
        self._tap("conv_post", x)
        nb = self.n_fft // 2 + 1
        spec = torch.exp(x[:, :nb, :])  # :804
        phase = torch.sin(x[:, nb:, :])  # :805
        audio = self.stft_inverse(spec.numpy(), phase.numpy())
        return audio

    # ------------------------------------------------------------- Decoder :947
    def __call__(self, asr, F0_curve, N, s, noise, keep_taps: bool = False) -> np.ndarray:
        """asr (B,512,T), F0_curve (B,2T), N (B,2T), s (B,128), noise (B,600T,9) -> (B,1,600T)."""
        self.keep_taps, self.taps = keep_taps, {}
        dt = self.dtype
        asr_t = torch.from_numpy(np.asarray(asr, np.float32)).to(dt)
        s_t = torch.from_numpy(np.asarray(s, np.float32)).to(dt)
        f0_t = torch.from_numpy(np.asarray(F0_curve, np.float32)).to(dt)[:, None, :]
        n_t = torch.from_numpy(np.asarray(N, np.float32)).to(dt)[:, None, :]
        with torch.no_grad():
            F0 = self._conv_weighted("decoder.F0_conv", f0_t, 2, 1, 1)  # :949
            Nn = self._conv_weighted("decoder.N_conv", n_t, 2, 1, 1)  # :950
            x = torch.cat([asr_t, F0, Nn], dim=1)  # :951
            x = self._adain_resblk1d("decoder.encode", x, s_t)
            self._tap("encode", x)
            asr_res = self._conv_weighted("decoder.asr_res.0", asr_t, 1, 0, 1)  # :953
            res = True
            for i in range(4):
                if res:
                    x = torch.cat([x, asr_res, F0, Nn], dim=1)  # :957
                up = i == 3
                x = self._adain_resblk1d(f"decoder.decode.{i}", x, s_t, upsample=up)
                self._tap(f"decode{i}", x)
                if up:
                    res = False
            audio = self.generator(x, s_t, np.asarray(F0_curve, np.float32), noise)
        return audio


def _unwrap(p: np.ndarray, axis: int = 1, period: float = 2 * math.pi) -> np.ndarray:
    """mlx_unwrap (istftnet.py:417-450)."""
    discont = period / 2
    dd = np.diff(p, axis=axis)
    ih = period / 2
    il = -ih
    ddmod = dd - period * np.floor((dd - il) / period)
    ddmod = np.where((np.abs(dd - ih) < 1e-10) & (dd > 0), ih, ddmod)
    corr = ddmod - dd
    corr = np.where(np.abs(dd) < discont, 0, corr)
    pad_shape = list(corr.shape)
    pad_shape[axis] = 1
    cum = np.cumsum(np.concatenate([np.zeros(pad_shape, p.dtype), corr.astype(p.dtype)], axis=axis), axis=axis)
    return (p + cum).astype(p.dtype)
