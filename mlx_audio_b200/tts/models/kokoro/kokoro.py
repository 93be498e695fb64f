"""Boundary shim for mlx_audio/tts/models/kokoro/kokoro.py: ``ModelConfig`` and ``Model`` with the reference's call
surface, wrapping the B200 ``Decoder`` (the hot path).

Scope: the text front end of the reference ``Model`` - ALBERT, TextEncoder, ProsodyPredictor
(kokoro.py:70-106, modules.py) - is NOT part of the rebuilt path; it only produces the decoder inputs
``asr (B,512,T), F0_pred (B,2T), N_pred (B,2T)`` (kokoro.py:142-162).  ``Model`` therefore takes that producer as a
pluggable ``frontend`` callable; everything from ``decoder(asr, F0_pred, N_pred, ref_s[:, :128])`` (kokoro.py:164-165)
on runs here.  ``SyntheticFrontend`` stands in for tests / benchmarks (seeded features of the right shapes).
"""
from __future__ import annotations

from dataclasses import dataclass, field, fields
from numbers import Number
from typing import Any, Callable, Dict, Optional, Tuple

import numpy as np

from ....synth import KOKORO_DECODER, synth_f0
from .istftnet import Decoder
from .modules import ProsodyPredictor


@dataclass
class ModelConfig:
    """kokoro.py:47-63 (only the fields the decoder needs are interpreted; the rest are carried)."""
    istftnet: dict = field(default_factory=lambda: {k: v for k, v in KOKORO_DECODER.items()
                                                    if k not in ("dim_in", "style_dim", "dim_out")})
    dim_in: int = 64
    dropout: float = 0.2
    hidden_dim: int = 512
    max_conv_dim: int = 512
    max_dur: int = 50
    multispeaker: bool = True
    n_layer: int = 3
    n_mels: int = 80
    n_token: int = 178
    style_dim: int = 128
    text_encoder_kernel_size: int = 5
    plbert: dict = field(default_factory=dict)
    vocab: Dict[str, int] = field(default_factory=dict)
    sample_rate: int = 24000
    model_type: str = "kokoro"

    @classmethod
    def from_dict(cls, params: dict) -> "ModelConfig":
        """BaseModelArgs.from_dict (base.py:9-18): unknown keys are ignored."""
        names = {f.name for f in fields(cls)}
        return cls(**{k: v for k, v in params.items() if k in names})


Frontend = Callable[[str, Any, Number], Tuple[Any, Any, Any, Any]]


class SyntheticFrontend:
    """Seeded stand-in for the text front end: every phoneme gets ``frames_per_phoneme`` asr frames; returns
    (asr (1,512,T), F0 (1,2T), N (1,2T), pred_dur (n_phonemes+2,))."""

    def __init__(self, hidden_dim: int = 512, frames_per_phoneme: int = 3, seed: int = 0):
        self.hidden_dim, self.fpp, self.seed = hidden_dim, frames_per_phoneme, seed

    def __call__(self, phonemes: str, ref_s, speed: Number = 1):
        n = len(phonemes) + 2                                     # [0, *ids, 0] (kokoro.py:135)
        dur = np.maximum(np.round(np.full(n, self.fpp) / float(speed)), 1).astype(np.int32)   # kokoro.py:148-150
        T = int(dur.sum())
        rng = np.random.Generator(np.random.PCG64(self.seed + len(phonemes)))
        asr = rng.standard_normal((1, self.hidden_dim, T), dtype=np.float32)
        F0 = synth_f0(1, 2 * T, seed=self.seed + 4)
        N = rng.standard_normal((1, 2 * T), dtype=np.float32)
        return asr, F0, N, dur


class Model:
    """Reference surface: ``Model(config)``; ``model(phonemes, ref_s, speed=1, return_output=False, decoder=None)``
    (kokoro.py:120-170); ``model.decoder`` is the B200 Decoder; ``sanitize`` keeps only what the decoder consumes."""

    @dataclass
    class Output:                      # kokoro.py:115-118
        audio: Any
        pred_dur: Optional[Any] = None

    def __init__(self, config: ModelConfig, repo_id: Optional[str] = None, frontend: Optional[Frontend] = None,
                 dtype: str = "bf16", device: Optional[str] = None):
        self.config = config
        self.repo_id = repo_id
        self.vocab = config.vocab
        self.context_length = int(config.plbert.get("max_position_embeddings", 512)) if config.plbert else 512
        self.frontend = frontend
        self.decoder = Decoder(dim_in=config.hidden_dim, style_dim=config.style_dim, dim_out=config.n_mels,
                               **config.istftnet, dtype=dtype, device=device)    # kokoro.py:107-112
        # F0 / energy heads (kokoro.py:160, modules.py:355-377): used when the front end hands over ``en`` instead of
        # finished F0 / N curves and the ``predictor.*`` weights were loaded
        self.predictor = ProsodyPredictor(style_dim=config.style_dim, d_hid=config.hidden_dim, nlayers=config.n_layer,
                                          max_dur=config.max_dur, dropout=config.dropout, dtype=dtype, device=device)
        self.sample_rate = config.sample_rate
        self._pipelines: Dict[str, Any] = {}
        self._g2p: Optional[Callable[[str], Any]] = None     # pluggable G2P handed to the pipelines (see pipeline.py)

    # ---- weights
    def sanitize(self, weights: Dict[str, Any]) -> Dict[str, Any]:
        """kokoro.py:172-252 restricted to the decoder: ``decoder.*`` keys go through Decoder.sanitize
        (istftnet.py:965-979), every other sub-module's weights belong to the out-of-scope front end."""
        out = {}
        wanted = set("predictor." + n for n in self.predictor.parameter_shapes())
        for k, v in weights.items():
            if k.startswith("decoder."):
                out[k] = self.decoder.sanitize(k, np.asarray(v))
            elif k.startswith("predictor."):                 # kokoro.py:213-247 (F0 / N heads, projections, shared LSTM)
                for nk, nv in self.predictor.sanitize(k, np.asarray(v)).items():
                    if nk in wanted:
                        out[nk] = nv
        return out

    def load_weights(self, weights, strict: bool = True):
        items = list(weights.items() if isinstance(weights, dict) else weights)
        self.decoder.load_weights([(k, v) for k, v in items if k.startswith("decoder.")], strict=strict)
        pred = [(k, v) for k, v in items if k.startswith("predictor.")]
        if pred:
            self.predictor.load_weights(pred, strict=strict)
        return self

    # ---- forward
    def __call__(self, phonemes: str, ref_s, speed: Number = 1, return_output: bool = False, decoder=None):
        n_ids = len(phonemes) if not self.vocab else sum(1 for p in phonemes if self.vocab.get(p) is not None)
        assert n_ids + 2 <= self.context_length, (n_ids + 2, self.context_length)      # kokoro.py:131-134
        if self.frontend is None:
            raise NotImplementedError(
                "Model.__call__ needs a `frontend` (phonemes, ref_s, speed) -> (asr, F0_pred, N_pred, pred_dur): the "
                "reference's text front end (ALBERT / TextEncoder / ProsodyPredictor, kokoro.py:142-162) is outside "
                "the rebuilt hot path; pass your own or mlx_audio_b200...kokoro.SyntheticFrontend for testing")
        ref_s = np.asarray(ref_s, dtype=np.float32) if not hasattr(ref_s, "shape") else ref_s
        feats = self.frontend(phonemes, ref_s, speed)
        if isinstance(feats, dict):
            # front end stops at kokoro.py:157-159: {"asr", "en" (1, d_hid + style, T), "pred_dur"}; the F0 / N heads
            # run here with the prosody half of the style vector, s = ref_s[:, 128:] (kokoro.py:143, 160)
            asr, pred_dur = feats["asr"], feats.get("pred_dur")
            if "en" in feats:
                F0_pred, N_pred = self.predictor.F0Ntrain(feats["en"], ref_s[:, 128:])
            else:
                F0_pred, N_pred = feats["F0_pred"], feats["N_pred"]
        else:
            asr, F0_pred, N_pred, pred_dur = feats
        dec = decoder if decoder is not None else self.decoder
        audio = dec(asr, F0_pred, N_pred, ref_s[:, :128])[0]                            # kokoro.py:165
        return self.Output(audio=audio, pred_dur=pred_dur) if return_output else audio

    # ---- text -> audio segments (the route the reference's CLI and server take: kokoro.py:258-346)
    def _get_pipeline(self, lang_code: str):
        """kokoro.py:258-267: one cached KokoroPipeline per language code."""
        from .pipeline import KokoroPipeline
        if lang_code not in self._pipelines:
            self._pipelines[lang_code] = KokoroPipeline(lang_code=lang_code, model=self, repo_id=self.repo_id,
                                                        g2p=self._g2p)
        return self._pipelines[lang_code]

    def generate(self, text: str, voice: Optional[str] = None, speed: float = 1.0, lang_code: str = "a",
                 split_pattern: str = r"\n+", **kwargs):
        """kokoro.py:269-346: yields one ``GenerationResult`` per segment with duration / real-time-factor statistics.
        Differences, all forced by the scope: voices must be registered (``pipeline.add_voice`` or a pack path - the hub
        download of :280-283 is outside the path, so the voice cache is not reset); ``samples`` is the true sample count
        (the reference reports ``audio.shape[0]``, which is 1 for its (1, S) audio, :297); peak memory is CUDA's."""
        import time

        import torch

        from ..base import GenerationResult
        pipeline = self._get_pipeline(lang_code)
        if voice is None:
            voice = "af_heart"
        start = time.time()
        for segment_idx, (graphemes, phonemes, audio) in enumerate(
                pipeline(text, voice=voice, speed=speed, split_pattern=split_pattern)):
            if hasattr(audio, "is_cuda") and audio.is_cuda:
                torch.cuda.current_stream(audio.device).synchronize()     # mx.eval(audio) of kokoro.py:168
            now = time.time()
            segment_time, start = now - start, now
            samples = int(audio.shape[-1]) if audio is not None else 0
            assert samples > 0, "No audio generated"
            token_count = len(phonemes) if phonemes is not None else 0
            dur = samples / self.sample_rate
            duration_str = f"{int(dur // 3600):02d}:{int(dur // 60):02d}:{int(dur % 60):02d}.{int((dur % 1) * 1000):03d}"
            yield GenerationResult(
                audio=audio[0], samples=samples, sample_rate=self.sample_rate, segment_idx=segment_idx,
                token_count=token_count, audio_duration=duration_str,
                real_time_factor=round(segment_time / dur, 2) if dur > 0 else 0,
                prompt={"tokens": token_count,
                        "tokens-per-sec": round(token_count / segment_time, 2) if segment_time > 0 else 0},
                audio_samples={"samples": samples,
                               "samples-per-sec": round(samples / segment_time, 2) if segment_time > 0 else 0},
                processing_time_seconds=segment_time,
                peak_memory_usage=(torch.cuda.max_memory_allocated() / 1e9) if torch.cuda.is_available() else 0.0)
