"""Boundary shim for mlx_audio/tts/models/kokoro/pipeline.py ``KokoroPipeline``: same constructor / ``__call__`` /
``generate_from_tokens`` / ``Result`` surface, driving the B200 decoder through ``Model``.

G2P (misaki / espeak, pipeline.py:66-127) is CPU text processing outside the rebuilt path: it is imported lazily and
may be replaced by any ``g2p(text) -> (phonemes, tokens)`` callable; the raw-phoneme entry
(``generate_from_tokens``, pipeline.py:238-274) works without it.  Extension over the reference:
``batch_tokens`` decodes many phoneme strings in same-length batches instead of one decoder call per chunk.
"""
from __future__ import annotations

import re
from dataclasses import dataclass
from numbers import Number
from typing import Any, Callable, Dict, Generator, List, Optional, Union

import numpy as np

ALIASES = {"en-us": "a", "en-gb": "b", "es": "e", "fr-fr": "f", "hi": "h", "it": "i", "pt-br": "p", "ja": "j",
           "zh": "z"}                                                                    # pipeline.py:16-27
LANG_CODES = dict(a="American English", b="British English", e="es", f="fr-fr", h="hi", i="it", p="pt-br",
                  j="Japanese", z="Mandarin Chinese")                                    # pipeline.py:29-41
MAX_PHONEMES = 510                                                                       # pipeline.py:270-271


class KokoroPipeline:
    @dataclass
    class Result:                       # pipeline.py:330-356
        graphemes: str
        phonemes: str
        tokens: Optional[List[Any]] = None
        output: Optional[Any] = None
        text_index: Optional[int] = None

        @property
        def audio(self):
            return None if self.output is None else self.output.audio

        @property
        def pred_dur(self):
            return None if self.output is None else self.output.pred_dur

        def __iter__(self):             # backward-compat tuple view (graphemes, phonemes, audio)
            yield self.graphemes
            yield self.phonemes
            yield self.audio

        def __getitem__(self, index):
            return [self.graphemes, self.phonemes, self.audio][index]

        def __len__(self):
            return 3

    def __init__(self, lang_code: str, model=None, repo_id: Optional[str] = None, trf: bool = False,
                 g2p: Optional[Callable[[str], Any]] = None):
        lang_code = ALIASES.get(lang_code.lower(), lang_code.lower())
        assert lang_code in LANG_CODES, (lang_code, LANG_CODES)                          # pipeline.py:84-85
        self.lang_code, self.model, self.repo_id, self.trf = lang_code, model, repo_id, trf
        self.voices: Dict[str, np.ndarray] = {}
        self._g2p = g2p

    # ---- voices (pipeline.py:129-161): packs are (510, 1, 256) arrays; pack[len(ps)-1] -> (1, 256)
    def add_voice(self, name: str, pack) -> None:
        pack = np.asarray(pack, dtype=np.float32)
        if pack.ndim != 3 or pack.shape[1:] != (1, 256):
            raise ValueError(f"voice pack must be (N, 1, 256), got {pack.shape}")
        self.voices[name] = pack

    def load_single_voice(self, voice: str) -> np.ndarray:
        """pipeline.py:129-145 for local files: a registered name or a path to a .pt / .npy / .npz / .safetensors pack."""
        if voice in self.voices:
            return self.voices[voice]
        if voice.endswith((".pt", ".pth", ".npy", ".npz", ".safetensors")):
            from .voice import load_voice_tensor
            self.add_voice(voice, load_voice_tensor(voice))
            return self.voices[voice]
        raise ValueError(f"unknown voice {voice!r}: register it with add_voice(name, pack) or pass a path to a .pt "
                         "pack (downloading voice packs from the hub is outside this path)")

    def load_voice(self, voice: str, delimiter: str = ",") -> np.ndarray:
        """pipeline.py:154-163: ``"a,b"`` averages the packs."""
        if voice in self.voices:
            return self.voices[voice]
        packs = [self.load_single_voice(v) for v in voice.split(delimiter)]
        if len(packs) == 1:
            return packs[0]
        self.voices[voice] = np.mean(np.stack(packs), axis=0).astype(np.float32)
        return self.voices[voice]

    # ---- G2P (lazy; out of scope, pluggable)
    @property
    def g2p(self):
        if self._g2p is None:
            try:
                from misaki import en  # noqa: F401  (pipeline.py:10)
            except ImportError as e:
                raise ImportError("KokoroPipeline text input needs a G2P: install `misaki` or pass g2p=callable; "
                                  "generate_from_tokens(phoneme_string, ...) works without one") from e
            british = self.lang_code == "b"
            self._g2p = en.G2P(trf=self.trf, british=british, fallback=None)
        return self._g2p

    @staticmethod
    def infer(model, ps: str, pack, speed: Number = 1):                                   # pipeline.py:228-236
        return model(ps, pack[len(ps) - 1], speed, return_output=True)

    def generate_from_tokens(self, tokens: Union[str, List[Any]], voice: str, speed: Number = 1, model=None
                             ) -> Generator["KokoroPipeline.Result", None, None]:
        model = model or self.model
        if model and voice is None:
            raise ValueError('Specify a voice: pipeline.generate_from_tokens(..., voice="af_heart")')
        pack = self.load_voice(voice) if model else None
        if not isinstance(tokens, str):
            tokens = "".join(getattr(t, "phonemes", None) or "" for t in tokens)
        if len(tokens) > MAX_PHONEMES:
            raise ValueError(f"Phoneme string too long: {len(tokens)} > {MAX_PHONEMES}")
        output = KokoroPipeline.infer(model, tokens, pack, speed) if model else None
        yield self.Result(graphemes="", phonemes=tokens, output=output)

    def __call__(self, text: Union[str, List[str]], voice: Optional[str] = None, speed: Number = 1,
                 split_pattern: Optional[str] = r"\n+") -> Generator["KokoroPipeline.Result", None, None]:
        if voice is None:
            raise ValueError('Specify a voice: en_us_pipeline(text="Hello world!", voice="af_heart")')
        pack = self.load_voice(voice) if self.model else None
        if isinstance(text, str):
            text = re.split(split_pattern, text.strip()) if split_pattern else [text]      # pipeline.py:371
        for index, graphemes in enumerate(text):
            if not graphemes.strip():
                continue
            res = self.g2p(graphemes)
            ps = res[0] if isinstance(res, (tuple, list)) else res
            tokens = res[1] if isinstance(res, (tuple, list)) and len(res) > 1 else None
            for k in range(0, max(len(ps), 1), MAX_PHONEMES):                             # <= 510 phonemes per call
                chunk = ps[k:k + MAX_PHONEMES]
                if not chunk:
                    continue
                output = KokoroPipeline.infer(self.model, chunk, pack, speed) if self.model else None
                yield self.Result(graphemes=graphemes, phonemes=chunk, tokens=tokens, output=output, text_index=index)

    # ---- extension: sentence batching (the reference decodes one chunk per call, kokoro.py:135-165)
    def batch_tokens(self, phoneme_strings: List[str], voice: str, speed: Number = 1, max_batch: int = 64,
                     bucket: int = 32):
        """Decode many phoneme strings with batched decoder calls: front-end features are produced per string, the
        strings are sorted by frame count and cut into MIXED-length batches of ``max_batch`` (each padded to its longest
        member, rounded up to ``bucket`` frames) that go through ``Decoder.__call__(..., lengths=)`` - one launch plan per
        (batch size, padded length) instead of one per distinct length; every string's audio equals decoding it alone
        (the reference does exactly that, one chunk per call: pipeline.py:381-402).  Returns the waveforms in input order."""
        import torch
        from ....sharding import assign_ragged
        model = self.model
        pack = self.load_voice(voice)
        feats = []
        for ps in phoneme_strings:
            if len(ps) > MAX_PHONEMES:
                raise ValueError(f"Phoneme string too long: {len(ps)} > {MAX_PHONEMES}")
            ref_s = pack[len(ps) - 1]
            asr, F0, N, _ = model.frontend(ps, ref_s, speed)
            feats.append((np.asarray(asr), np.asarray(F0), np.asarray(N), np.asarray(ref_s)[:, :128]))
        out: List[Optional[Any]] = [None] * len(feats)
        frames = [f[0].shape[-1] for f in feats]
        for T, sel in assign_ragged(frames, 1, max_batch, bucket)[0]:
            n = len(sel)
            same = all(frames[i] == frames[sel[0]] for i in sel)
            if same:                        # a batch of one length needs no padding and no masks
                T = frames[sel[0]]
            asr = np.zeros((n, feats[sel[0]][0].shape[1], T), np.float32)
            F0 = np.zeros((n, 2 * T), np.float32)
            N = np.zeros((n, 2 * T), np.float32)
            for r, i in enumerate(sel):
                t = frames[i]
                asr[r, :, :t], F0[r, :2 * t], N[r, :2 * t] = feats[i][0][0], feats[i][1][0], feats[i][2][0]
            s = np.concatenate([feats[i][3] for i in sel])
            y = model.decoder(asr, F0, N, s) if same else model.decoder(asr, F0, N, s, lengths=[frames[i] for i in sel])
            for r, i in enumerate(sel):
                row = y[r, 0, :600 * frames[i]]
                out[i] = row.clone() if isinstance(row, torch.Tensor) else np.array(row)
        return out
