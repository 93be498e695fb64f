"""Import the reference's own hot-path source files, unmodified, on top of oracle/mlx_shim.

TEST INFRASTRUCTURE ONLY.  Works only where /root/reference exists (the build
container); never used on the GPU box.  Used by tests/golden/make_golden.py to
mint golden vectors from the reference's own Python code, and by the CPU tests
(skipped when the reference tree is absent) to re-check the oracle live.
"""
from __future__ import annotations

import importlib.util
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("KK_REFERENCE_ROOT", "/root/reference")
_SHIM_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "mlx_shim")


def reference_available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "mlx_audio", "tts", "models", "kokoro", "istftnet.py"))


def _pkg(name: str, path: str):
    m = types.ModuleType(name)
    m.__path__ = [path]
    m.__package__ = name
    sys.modules[name] = m
    return m


def _load(name: str, path: str):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


_CACHE = {}


def load_reference():
    """Returns a namespace with the reference modules ``utils``, ``interpolate``, ``base``,
    ``istftnet`` (loaded from REFERENCE_ROOT) plus the shim's ``mx`` / ``nn``."""
    if "ns" in _CACHE:
        return _CACHE["ns"]
    if not reference_available():
        raise FileNotFoundError(f"reference tree not found under {REFERENCE_ROOT}")
    real_mlx = None
    try:  # prefer a real MLX if one is ever importable
        import mlx.core  # noqa: F401
        real_mlx = sys.modules["mlx"]
    except Exception:
        if _SHIM_DIR not in sys.path:
            sys.path.insert(0, _SHIM_DIR)
        for k in [k for k in sys.modules if k == "mlx" or k.startswith("mlx.")]:
            del sys.modules[k]
    import mlx.core as mx
    import mlx.nn as nn

    root = os.path.join(REFERENCE_ROOT, "mlx_audio")
    # Stub the package chain so the heavy __init__ files (misaki, mlx_lm, ...) are never executed.
    _pkg("mlx_audio", root)
    _pkg("mlx_audio.tts", os.path.join(root, "tts"))
    _pkg("mlx_audio.tts.models", os.path.join(root, "tts", "models"))
    _pkg("mlx_audio.tts.models.kokoro", os.path.join(root, "tts", "models", "kokoro"))
    ns = types.SimpleNamespace(mx=mx, nn=nn, shim=real_mlx is None)
    ns.utils = _load("mlx_audio.utils", os.path.join(root, "utils.py"))
    ns.base = _load("mlx_audio.tts.models.base", os.path.join(root, "tts", "models", "base.py"))
    ns.interpolate = _load("mlx_audio.tts.models.interpolate", os.path.join(root, "tts", "models", "interpolate.py"))
    ns.istftnet = _load("mlx_audio.tts.models.kokoro.istftnet",
                        os.path.join(root, "tts", "models", "kokoro", "istftnet.py"))
    _CACHE["ns"] = ns
    return ns


def build_reference_decoder(weights, cfg):
    """Instantiate the reference ``Decoder`` and load ``weights`` (reference key names, MLX layout)."""
    ns = load_reference()
    dec = ns.istftnet.Decoder(**cfg)
    for k, v in weights.items():
        assert k.startswith("decoder."), k
        dec.set_parameter(k[len("decoder."):], v)
    return dec


def run_reference_decoder(weights, cfg, asr, F0_curve, N, s, noise):
    """Run the reference Decoder.__call__ with ``mx.random.normal`` replaying ``noise``.

    Draw order inside the reference (istftnet.py:563, :620, :679): rand_ini (B,9),
    noise (B,600T,9), unused (B,600T,1)."""
    import numpy as np
    ns = load_reference()
    mx = ns.mx
    dec = build_reference_decoder(weights, cfg)
    B = asr.shape[0]
    L = noise.shape[1]
    mx.random.queue = [np.zeros((B, noise.shape[2]), np.float32), np.asarray(noise, np.float32),
                       np.zeros((B, L, 1), np.float32)]
    mx.random.log = []
    out = dec(mx.array(asr), mx.array(F0_curve), mx.array(N), mx.array(s))
    assert not mx.random.queue, "reference drew fewer random tensors than expected"
    return np.asarray(out)


def load_reference_modules():
    """The reference's kokoro/modules.py (ProsodyPredictor, LSTM ...) on the same stubbed package chain."""
    ns = load_reference()
    if not hasattr(ns, "modules"):
        ns.modules = _load("mlx_audio.tts.models.kokoro.modules",
                           os.path.join(REFERENCE_ROOT, "mlx_audio", "tts", "models", "kokoro", "modules.py"))
    return ns.modules


def run_reference_f0ntrain(weights, x, s, style_dim=128, d_hid=512):
    """Run the reference ``ProsodyPredictor.F0Ntrain(x, s)`` (modules.py:355-377) with ``weights`` (keys
    ``predictor.*``: the F0 / N blocks, the projections and the shared LSTM).  Returns (F0, N) numpy arrays."""
    import numpy as np
    ns = load_reference()
    mod = load_reference_modules()
    pp = mod.ProsodyPredictor(style_dim, d_hid, 3, 50, 0.2)
    for k, v in weights.items():
        assert k.startswith("predictor."), k
        pp.set_parameter(k[len("predictor."):], v)
    F0, N = pp.F0Ntrain(ns.mx.array(x), ns.mx.array(s))
    return np.asarray(F0), np.asarray(N)


def _function_from_source(path, name, namespace):
    """Compile ONE top-level function of a reference file (unmodified source text, parsed with ``ast``) in a namespace
    that supplies its free names - for files whose module-level imports (omegaconf, safetensors.torch, whole model
    trees) cannot be satisfied here.  The function body that runs is the reference's own."""
    import ast
    import typing
    tree = ast.parse(open(path, encoding="utf-8").read(), filename=path)
    fn = next(n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == name)
    mod = ast.Module(body=[fn], type_ignores=[])
    env = dict(namespace, Optional=typing.Optional, Any=typing.Any, Dict=typing.Dict)
    exec(compile(mod, path, "exec"), env)
    return env[name]


def load_reference_frontends():
    """The reference's log-mel front ends and the Vocos ``ISTFTHead``, loaded from their own source
    files: codec/models/vocos/{mel,vocos}.py, tts/models/indextts/mel.py, stt/models/whisper/audio.py.  Unrelated
    imports of those files (Encodec, load_audio) are stubbed; nothing of them is on the tested path."""
    ns = load_reference()
    if hasattr(ns, "vocos"):
        return ns
    root = os.path.join(REFERENCE_ROOT, "mlx_audio")
    _pkg("mlx_audio.codec", os.path.join(root, "codec"))
    _pkg("mlx_audio.codec.models", os.path.join(root, "codec", "models"))
    _pkg("mlx_audio.codec.models.vocos", os.path.join(root, "codec", "models", "vocos"))
    enc = _pkg("mlx_audio.codec.models.encodec", os.path.join(root, "codec", "models", "encodec"))
    enc.Encodec = type("Encodec", (), {})
    _pkg("mlx_audio.tts.models.indextts", os.path.join(root, "tts", "models", "indextts"))
    stt = _pkg("mlx_audio.stt", os.path.join(root, "stt"))
    stt_utils = types.ModuleType("mlx_audio.stt.utils")
    stt_utils.load_audio = lambda *a, **k: (_ for _ in ()).throw(RuntimeError("load_audio is stubbed"))
    sys.modules["mlx_audio.stt.utils"] = stt_utils
    stt.utils = stt_utils
    _pkg("mlx_audio.stt.models", os.path.join(root, "stt", "models"))
    _pkg("mlx_audio.stt.models.whisper", os.path.join(root, "stt", "models", "whisper"))
    ns.vocos_mel = _load("mlx_audio.codec.models.vocos.mel", os.path.join(root, "codec", "models", "vocos", "mel.py"))
    ns.vocos = _load("mlx_audio.codec.models.vocos.vocos", os.path.join(root, "codec", "models", "vocos", "vocos.py"))
    ns.indextts_mel = _load("mlx_audio.tts.models.indextts.mel", os.path.join(root, "tts", "models", "indextts", "mel.py"))
    ns.whisper_audio = _load("mlx_audio.stt.models.whisper.audio",
                             os.path.join(root, "stt", "models", "whisper", "audio.py"))
    _pkg("mlx_audio.stt.models.parakeet", os.path.join(root, "stt", "models", "parakeet"))
    ns.parakeet_audio = _load("mlx_audio.stt.models.parakeet.audio",
                              os.path.join(root, "stt", "models", "parakeet", "audio.py"))
    _pkg("mlx_audio.codec.models.s3", os.path.join(root, "codec", "models", "s3"))
    ns.s3_utils = _load("mlx_audio.codec.models.s3.utils", os.path.join(root, "codec", "models", "s3", "utils.py"))
    _pkg("mlx_audio.codec.models.bigvgan", os.path.join(root, "codec", "models", "bigvgan"))
    ns.bigvgan_activation = _load("mlx_audio.codec.models.bigvgan.activation",
                                  os.path.join(root, "codec", "models", "bigvgan", "activation.py"))
    _pkg("mlx_audio.codec.models.descript", os.path.join(root, "codec", "models", "descript"))
    _pkg("mlx_audio.codec.models.descript.nn", os.path.join(root, "codec", "models", "descript", "nn"))
    ns.dac_layers = _load("mlx_audio.codec.models.descript.nn.layers",
                          os.path.join(root, "codec", "models", "descript", "nn", "layers.py"))
    ns.spark_mel_spectrogram = _function_from_source(
        os.path.join(root, "tts", "models", "spark", "bicodec.py"), "mel_spectrogram",
        dict(mx=ns.mx, hanning=ns.utils.hanning, mel_filters=ns.utils.mel_filters, stft=ns.utils.stft))
    return ns


def load_reference_snac():
    """The reference's SNAC layers (codec/models/snac/{attention,layers}.py: WNConv1d, WNConvTranspose1d, Snake1d,
    ResidualUnit, NoiseBlock, DecoderBlock, Decoder), unmodified, on the shim - SURVEY 8(f) row 4."""
    ns = load_reference()
    if hasattr(ns, "snac_layers"):
        return ns
    root = os.path.join(REFERENCE_ROOT, "mlx_audio")
    for name, sub in (("mlx_audio.codec", ("codec",)), ("mlx_audio.codec.models", ("codec", "models")),
                      ("mlx_audio.codec.models.snac", ("codec", "models", "snac"))):
        if name not in sys.modules:
            _pkg(name, os.path.join(root, *sub))
    ns.snac_attention = _load("mlx_audio.codec.models.snac.attention",
                              os.path.join(root, "codec", "models", "snac", "attention.py"))
    ns.snac_layers = _load("mlx_audio.codec.models.snac.layers", os.path.join(root, "codec", "models", "snac", "layers.py"))
    return ns


def load_reference_dac():
    """The reference's DAC generator (codec/models/descript/dac.py: ResidualUnit, DecoderBlock, Decoder, unmodified, over
    its own nn/layers.py) on the shim.  ``base.CodecMixin`` (soundfile I/O) and ``nn.quantize`` are stubbed: neither is on
    the decoder's path."""
    ns = load_reference_frontends()
    if hasattr(ns, "dac"):
        return ns
    root = os.path.join(REFERENCE_ROOT, "mlx_audio")
    base = types.ModuleType("mlx_audio.codec.models.descript.base")
    base.CodecMixin = type("CodecMixin", (), {})
    sys.modules[base.__name__] = base
    quant = types.ModuleType("mlx_audio.codec.models.descript.nn.quantize")
    quant.ResidualVectorQuantize = type("ResidualVectorQuantize", (), {})
    sys.modules[quant.__name__] = quant
    if "huggingface_hub" not in sys.modules:
        try:
            import huggingface_hub  # noqa: F401
        except Exception:
            hub = types.ModuleType("huggingface_hub")
            hub.snapshot_download = lambda *a, **k: (_ for _ in ()).throw(RuntimeError("no network"))
            sys.modules["huggingface_hub"] = hub
    ns.dac = _load("mlx_audio.codec.models.descript.dac", os.path.join(root, "codec", "models", "descript", "dac.py"))
    return ns


def load_reference_bigvgan():
    """The reference's BigVGAN generator (codec/models/bigvgan/{activation,conv,resample,amp,bigvgan}.py: Snake /
    SnakeBeta, the weight-normed convolutions, the kaiser-sinc anti-aliased ``Activation1d``, ``AMPBlock1/2``,
    ``BigVGAN``), unmodified, on the shim - SURVEY 8(f) row 4."""
    ns = load_reference()
    if hasattr(ns, "bigvgan"):
        return ns
    root = os.path.join(REFERENCE_ROOT, "mlx_audio")
    for name, sub in (("mlx_audio.codec", ("codec",)), ("mlx_audio.codec.models", ("codec", "models")),
                      ("mlx_audio.codec.models.bigvgan", ("codec", "models", "bigvgan"))):
        if name not in sys.modules:
            _pkg(name, os.path.join(root, *sub))
    d = os.path.join(root, "codec", "models", "bigvgan")
    for mod in ("activation", "conv", "resample", "amp"):
        setattr(ns, f"bigvgan_{mod}", _load(f"mlx_audio.codec.models.bigvgan.{mod}", os.path.join(d, f"{mod}.py")))
    ns.bigvgan = _load("mlx_audio.codec.models.bigvgan.bigvgan", os.path.join(d, "bigvgan.py"))
    return ns
