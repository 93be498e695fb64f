"""Seeded synthetic weights and inputs for the Kokoro ISTFTNet decoder path.

The reference ships no checkpoint that can be fetched here (no network) and its
default initialisation is all-ones / all-zeros (istftnet.py:119-126), which is
useless for parity.  This module is the single source of the synthetic recipe
 so that the CUDA path, the CPU oracle, the golden-vector
generator and bench.py all read byte-identical tensors.

numpy only: no torch, no CUDA, nothing from ``oracle/``.
"""
from __future__ import annotations

from typing import Dict, Iterator, List, Tuple

import numpy as np

# Canonical Kokoro-82M decoder configuration
# (reference: mlx_audio/tts/tests/test_models.py:92-122, kokoro.py:107-112).
KOKORO_ISTFTNET = dict(
    upsample_rates=[10, 6],
    upsample_kernel_sizes=[20, 12],
    gen_istft_n_fft=20,
    gen_istft_hop_size=5,
    resblock_kernel_sizes=[3, 7, 11],
    resblock_dilation_sizes=[[1, 3, 5], [1, 3, 5], [1, 3, 5]],
    upsample_initial_channel=512,
)
KOKORO_DECODER = dict(dim_in=512, style_dim=128, dim_out=80, **KOKORO_ISTFTNET)

SAMPLE_RATE = 24000
SAMPLES_PER_FRAME = 600  # one asr frame = 25 ms = 600 samples (10*6*5*2)
NUM_HARMONICS = 9  # harmonic_num=8 -> 9 sine components (istftnet.py:715, 546)


def decoder_param_specs(cfg: dict = None) -> List[Tuple[str, Tuple[int, ...], str]]:
    """(name, shape, kind) for every parameter of ``Decoder`` in the reference's
    own key naming and MLX layout (weight_v is (O, K, I)); mirrors the module
    tree built at istftnet.py:903-945, 697-767, 853-861, 350-375."""
    cfg = dict(KOKORO_DECODER if cfg is None else cfg)
    dim_in, style = cfg["dim_in"], cfg["style_dim"]
    specs: List[Tuple[str, Tuple[int, ...], str]] = []

    def wconv(prefix, cin, cout, k, bias=True, encode=False, groups=1):
        # ConvWeighted(in, out, k): weight_v (out, k, in/groups) (istftnet.py:119-126)
        specs.append((f"{prefix}.weight_g", (cout, 1, 1), "g"))
        specs.append((f"{prefix}.weight_v", (cout, k, cin // groups if groups > 1 else cin), "v"))
        if bias:
            specs.append((f"{prefix}.bias", (cin if encode else cout,), "bias"))

    def adain(prefix, c):
        specs.append((f"{prefix}.fc.weight", (2 * c, style), "fc_w"))
        specs.append((f"{prefix}.fc.bias", (2 * c,), "fc_b"))

    def adain_resblk(prefix, cin, cout, upsample=False):
        wconv(f"{prefix}.conv1", cin, cout, 3)
        wconv(f"{prefix}.conv2", cout, cout, 3)
        adain(f"{prefix}.norm1", cin)
        adain(f"{prefix}.norm2", cout)
        if cin != cout:
            wconv(f"{prefix}.conv1x1", cin, cout, 1, bias=False)
        if upsample:
            # ConvWeighted(1, dim_in, 3, stride 2, padding 1, groups=dim_in) (istftnet.py:849-851)
            specs.append((f"{prefix}.pool.weight_g", (cin, 1, 1), "g"))
            specs.append((f"{prefix}.pool.weight_v", (cin, 3, 1), "v"))
            specs.append((f"{prefix}.pool.bias", (cin,), "bias"))

    def resblock1(prefix, ch, k):
        for i in range(3):
            wconv(f"{prefix}.convs1.{i}", ch, ch, k)
        for i in range(3):
            wconv(f"{prefix}.convs2.{i}", ch, ch, k)
        for i in range(3):
            adain(f"{prefix}.adain1.{i}", ch)
        for i in range(3):
            adain(f"{prefix}.adain2.{i}", ch)
        for i in range(3):
            specs.append((f"{prefix}.alpha1.{i}", (1, ch, 1), "alpha"))
        for i in range(3):
            specs.append((f"{prefix}.alpha2.{i}", (1, ch, 1), "alpha"))

    p = "decoder"
    adain_resblk(f"{p}.encode", dim_in + 2, 1024)
    for i in range(3):
        adain_resblk(f"{p}.decode.{i}", 1024 + 2 + 64, 1024)
    adain_resblk(f"{p}.decode.3", 1024 + 2 + 64, 512, upsample=True)
    wconv(f"{p}.F0_conv", 1, 1, 3)
    wconv(f"{p}.N_conv", 1, 1, 3)
    wconv(f"{p}.asr_res.0", 512, 64, 1)

    g = f"{p}.generator"
    specs.append((f"{g}.m_source.l_linear.weight", (1, NUM_HARMONICS), "l_w"))
    specs.append((f"{g}.m_source.l_linear.bias", (1,), "l_b"))
    uic = cfg["upsample_initial_channel"]
    rates, ksz = cfg["upsample_rates"], cfg["upsample_kernel_sizes"]
    n_fft = cfg["gen_istft_n_fft"]
    for i, (u, k) in enumerate(zip(rates, ksz)):
        # ConvWeighted(small, big, k, u, encode=True): weight_v (big, k, small), bias (small,)
        wconv(f"{g}.ups.{i}", uic // 2 ** (i + 1), uic // 2**i, k, encode=True)
    ch = uic
    for i in range(len(rates)):
        ch = uic // 2 ** (i + 1)
        for j, k in enumerate(cfg["resblock_kernel_sizes"]):
            resblock1(f"{g}.resblocks.{i * len(cfg['resblock_kernel_sizes']) + j}", ch, k)
        if i + 1 < len(rates):
            stride_f0 = int(np.prod(rates[i + 1:]))
            specs.append((f"{g}.noise_convs.{i}.weight", (ch, stride_f0 * 2, n_fft + 2), "noise_w"))
            specs.append((f"{g}.noise_convs.{i}.bias", (ch,), "bias"))
            resblock1(f"{g}.noise_res.{i}", ch, 7)
        else:
            specs.append((f"{g}.noise_convs.{i}.weight", (ch, 1, n_fft + 2), "noise_w"))
            specs.append((f"{g}.noise_convs.{i}.bias", (ch,), "bias"))
            resblock1(f"{g}.noise_res.{i}", ch, 11)
    wconv(f"{g}.conv_post", ch, n_fft + 2, 7)
    return specs


def prosody_param_specs(style_dim: int = 128, d_hid: int = 512) -> List[Tuple[str, Tuple[int, ...], str]]:
    """(name, shape, kind) of the parameters ``ProsodyPredictor.F0Ntrain`` touches (modules.py:295-340), reference key
    naming, MLX layout: ``predictor.{F0,N}.{0,1,2}.*`` (AdainResBlk1d), ``predictor.{F0,N}_proj.{weight,bias}``
    (nn.Conv1d(d_hid/2, 1, 1): weight (1, 1, d_hid/2)) and the ``predictor.shared`` bi-LSTM (modules.py:110-150)."""
    specs: List[Tuple[str, Tuple[int, ...], str]] = []
    h2 = d_hid // 2
    for head in ("F0", "N"):
        for i, (cin, cout, up) in enumerate([(d_hid, d_hid, False), (d_hid, h2, True), (h2, h2, False)]):
            p = f"predictor.{head}.{i}"
            for cname, ci, co, k in (("conv1", cin, cout, 3), ("conv2", cout, cout, 3)):
                specs += [(f"{p}.{cname}.weight_g", (co, 1, 1), "g"), (f"{p}.{cname}.weight_v", (co, k, ci), "v"),
                          (f"{p}.{cname}.bias", (co,), "bias")]
            specs += [(f"{p}.norm1.fc.weight", (2 * cin, style_dim), "fc_w"), (f"{p}.norm1.fc.bias", (2 * cin,), "fc_b"),
                      (f"{p}.norm2.fc.weight", (2 * cout, style_dim), "fc_w"), (f"{p}.norm2.fc.bias", (2 * cout,), "fc_b")]
            if cin != cout:
                specs += [(f"{p}.conv1x1.weight_g", (cout, 1, 1), "g"), (f"{p}.conv1x1.weight_v", (cout, 1, cin), "v")]
            if up:
                specs += [(f"{p}.pool.weight_g", (cin, 1, 1), "g"), (f"{p}.pool.weight_v", (cin, 3, 1), "v"),
                          (f"{p}.pool.bias", (cin,), "bias")]
        specs += [(f"predictor.{head}_proj.weight", (1, 1, h2), "proj_w"), (f"predictor.{head}_proj.bias", (1,), "bias")]
    for d in ("forward", "backward"):
        specs += [(f"predictor.shared.Wx_{d}", (2 * d_hid, d_hid + style_dim), "lstm"),
                  (f"predictor.shared.Wh_{d}", (2 * d_hid, h2), "lstm"),
                  (f"predictor.shared.bias_ih_{d}", (2 * d_hid,), "lstm"),
                  (f"predictor.shared.bias_hh_{d}", (2 * d_hid,), "lstm")]
    return specs


def synth_prosody_weights(seed: int = 2100, style_dim: int = 128, d_hid: int = 512) -> Dict[str, np.ndarray]:
    """Random-init ``predictor.*`` weights with the decoder recipe; projections ~ N(0, 1/C) scaled so that
    F0 lands in a voiced range (x 60 + bias 150 Hz for F0_proj); LSTM ~ U[-1/sqrt(H), 1/sqrt(H)] (modules.py:108)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    out: Dict[str, np.ndarray] = {}
    for name, shape, kind in prosody_param_specs(style_dim, d_hid):
        if kind == "v":
            a = rng.standard_normal(shape, dtype=np.float32)
        elif kind == "g":
            a = rng.uniform(0.7, 1.3, size=shape).astype(np.float32)
        elif kind == "bias":
            a = 0.02 * rng.standard_normal(shape, dtype=np.float32)
        elif kind == "fc_w":
            a = rng.standard_normal(shape, dtype=np.float32) * np.float32(0.1 / np.sqrt(shape[1]))
        elif kind == "fc_b":
            a = np.zeros(shape, np.float32)
        elif kind == "proj_w":
            a = rng.standard_normal(shape, dtype=np.float32) / np.float32(np.sqrt(shape[2]))
            if "F0_proj" in name:
                a = a * np.float32(60.0)
        elif kind == "lstm":
            sc = 1.0 / np.sqrt(d_hid // 2)
            a = rng.uniform(-sc, sc, size=shape).astype(np.float32)
        else:  # pragma: no cover
            raise AssertionError(kind)
        if name.endswith("F0_proj.bias"):
            a = a + np.float32(150.0)
        out[name] = np.ascontiguousarray(a, dtype=np.float32)
    return out


def synth_prosody_inputs(batch: int, frames: int, seed: int = 1100, style_dim: int = 128, d_hid: int = 512
                         ) -> Dict[str, np.ndarray]:
    """x ~ N(0,1) (B, d_hid + style, L) (``en``, kokoro.py:157), s ~ 0.5 N(0,1) (B, style)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    return dict(x=rng.standard_normal((batch, d_hid + style_dim, frames), dtype=np.float32),
                s=(0.5 * rng.standard_normal((batch, style_dim), dtype=np.float32)).astype(np.float32))


def synth_decoder_weights(seed: int = 2000, cfg: dict = None) -> Dict[str, np.ndarray]:
    """Random-init decoder weights (float32, MLX layout) following SURVEY.md 8(d):
    weight_v ~ N(0,1); weight_g ~ U[0.7,1.3]; conv bias ~ 0.02 N(0,1);
    AdaIN fc.weight ~ N(0, 0.1^2/128), fc.bias = 0; alpha ~ U[0.5,2];
    l_linear.weight ~ N(0,1/9), bias 0; noise_convs.weight ~ N(0, 1/(22 k));
    conv_post: weight_g = 0.5 and bias -2 on the 11 log-magnitude channels."""
    rng = np.random.Generator(np.random.PCG64(seed))
    cfg = dict(KOKORO_DECODER if cfg is None else cfg)
    n_bins = cfg["gen_istft_n_fft"] // 2 + 1
    out: Dict[str, np.ndarray] = {}
    for name, shape, kind in decoder_param_specs(cfg):
        if kind == "v":
            a = rng.standard_normal(shape, dtype=np.float32)
        elif kind == "g":
            a = rng.uniform(0.7, 1.3, size=shape).astype(np.float32)
        elif kind == "bias":
            a = (0.02 * rng.standard_normal(shape, dtype=np.float32)).astype(np.float32)
        elif kind == "fc_w":
            a = (rng.standard_normal(shape, dtype=np.float32) * np.float32(0.1 / np.sqrt(shape[1]))).astype(np.float32)
        elif kind == "fc_b":
            a = np.zeros(shape, np.float32)
        elif kind == "alpha":
            a = rng.uniform(0.5, 2.0, size=shape).astype(np.float32)
        elif kind == "l_w":
            a = (rng.standard_normal(shape, dtype=np.float32) / np.float32(3.0)).astype(np.float32)
        elif kind == "l_b":
            a = np.zeros(shape, np.float32)
        elif kind == "noise_w":
            a = (rng.standard_normal(shape, dtype=np.float32)
                 * np.float32(1.0 / np.sqrt(shape[1] * shape[2]))).astype(np.float32)
        else:  # pragma: no cover
            raise AssertionError(kind)
        if name.endswith("conv_post.weight_g"):
            a = np.full(shape, 0.5, np.float32)
        if name.endswith("conv_post.bias"):
            a = a.copy()
            a[:n_bins] -= np.float32(2.0)
        out[name] = np.ascontiguousarray(a, dtype=np.float32)
    return out


def synth_f0(batch: int, frames2: int, seed: int = 1004) -> np.ndarray:
    """F0 curve (B, 2T): voiced segments (about 60 % of frames, length U[10,60])
    with f0 = f_b (1 + 0.15 sin(2 pi k / 97)), f_b ~ U[90,320] Hz; unvoiced = 0."""
    rng = np.random.Generator(np.random.PCG64(seed))
    f0 = np.zeros((batch, frames2), np.float32)
    for b in range(batch):
        k = 0
        while k < frames2:
            seg = int(rng.integers(10, 61))
            voiced = rng.random() < 0.6
            end = min(frames2, k + seg)
            if voiced:
                fb = rng.uniform(90.0, 320.0)
                idx = np.arange(k, end)
                f0[b, k:end] = (fb * (1.0 + 0.15 * np.sin(2.0 * np.pi * idx / 97.0))).astype(np.float32)
            k = end
    return f0


def synth_decoder_inputs(batch: int, frames: int, seed: int = 1000, with_noise: bool = True,
                         dim_in: int = 512, style_dim: int = 128) -> Dict[str, np.ndarray]:
    """asr ~ N(0,1) (B,512,T); s ~ 0.5 N(0,1) (B,128); N ~ N(0,1) (B,2T);
    F0_curve (B,2T) from synth_f0; noise ~ N(0,1) (B,600T,9).  Seeds seed+1..seed+5."""
    def rn(s, shape, scale=1.0):
        r = np.random.Generator(np.random.PCG64(s))
        return (r.standard_normal(shape, dtype=np.float32) * np.float32(scale)).astype(np.float32)

    out = dict(
        asr=rn(seed + 1, (batch, dim_in, frames)),
        s=rn(seed + 2, (batch, style_dim), 0.5),
        N=rn(seed + 3, (batch, 2 * frames)),
        F0_curve=synth_f0(batch, 2 * frames, seed + 4),
    )
    if with_noise:
        out["noise"] = rn(seed + 5, (batch, SAMPLES_PER_FRAME * frames, NUM_HARMONICS))
    return out


def synth_audio(batch: int, samples: int, seed: int = 3001, scale: float = 0.1) -> np.ndarray:
    """Config-2 signal: 0.1 N(0,1), (batch, samples) float32."""
    r = np.random.Generator(np.random.PCG64(seed))
    return (r.standard_normal((batch, samples), dtype=np.float32) * np.float32(scale)).astype(np.float32)


def iter_conv_layers(cfg: dict = None) -> Iterator[Tuple[str, int, int, int]]:
    """(prefix, C_in, C_out, k) of every dense conv for FLOP accounting."""
    for name, shape, kind in decoder_param_specs(cfg):
        if kind == "v" and not name.endswith("pool.weight_v"):
            yield name[: -len(".weight_v")], shape[2], shape[0], shape[1]
        elif kind == "noise_w":
            yield name[: -len(".weight")], shape[2], shape[0], shape[1]


# Algorithmic work of the decoder (convs only), SURVEY.md 8(d) / Appendix A.
MAC_PER_FRAME = 649_006_284
FLOP_PER_AUDIO_SECOND = 2.0 * MAC_PER_FRAME * (SAMPLE_RATE / SAMPLES_PER_FRAME)


def synth_named_weights(shapes: Dict[str, Tuple[int, ...]], seed: int = 0) -> Dict[str, np.ndarray]:
    """Seeded weights for a module tree given as {dotted parameter name: shape} (the codec generators, whose parameter
    names end in weight_v / weight_g / bias / alpha / beta).  Each tensor has its own generator keyed on (seed, name), so the
    values do not depend on the order or on which other parameters exist."""
    import zlib
    out = {}
    for name in sorted(shapes):
        shape = tuple(int(s) for s in shapes[name])
        r = np.random.Generator(np.random.PCG64([seed, zlib.crc32(name.encode())]))
        if name.endswith("weight_v"):
            fan = max(1, int(np.prod(shape[1:])))
            a = r.standard_normal(shape, dtype=np.float32) / np.float32(np.sqrt(fan))
        elif name.endswith("weight_g"):
            a = r.uniform(0.25, 0.5, shape).astype(np.float32)     # small gains: the tanh head must not saturate
        elif name.endswith("alpha") or name.endswith("beta"):
            a = r.uniform(0.5, 2.0, shape).astype(np.float32)
        else:  # bias
            a = (0.1 * r.standard_normal(shape, dtype=np.float32)).astype(np.float32)
        out[name] = a
    return out
