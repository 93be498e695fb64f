"""Reference-facing surface: load_model / Model / KokoroPipeline (host logic on CPU with a stand-in decoder;
the real decoder path under the gpu marker)."""
import json
import os

import numpy as np
import pytest

from mlx_audio_b200 import synth
from mlx_audio_b200.tts.models.kokoro import KokoroPipeline, Model, ModelConfig, SyntheticFrontend
from mlx_audio_b200.tts.utils import load_config, load_model


class FakeDecoder:
    """Records calls; returns (B, 1, 600 T) zeros like Decoder.__call__ (istftnet.py:947-963)."""
    def __init__(self):
        self.calls = []

    def __call__(self, asr, F0_curve, N, s, lengths=None):
        asr, F0_curve, N, s = map(np.asarray, (asr, F0_curve, N, s))
        B, C, T = asr.shape
        assert C == 512 and F0_curve.shape == (B, 2 * T) and N.shape == (B, 2 * T) and s.shape == (B, 128)
        assert lengths is None or (len(lengths) == B and max(lengths) <= T)
        self.calls.append((B, T) if lengths is None else (B, T, tuple(lengths)))
        return np.zeros((B, 1, 600 * T), np.float32)


def _model():
    m = Model(ModelConfig(), frontend=SyntheticFrontend(frames_per_phoneme=2))
    m.decoder = FakeDecoder()
    return m


def test_model_call_contract():
    m = _model()
    ref_s = np.zeros((1, 256), np.float32)
    out = m("hello", ref_s, 1, return_output=True)
    T = 2 * (5 + 2)
    assert out.audio.shape == (1, 600 * T) and out.pred_dur.shape == (7,) and m.decoder.calls == [(1, T)]
    assert m("hello", ref_s).shape == (1, 600 * T)                       # return_output=False -> bare audio
    slow = m("hello", ref_s, 0.5, return_output=True)                  # speed divides the durations (kokoro.py:149)
    assert slow.audio.shape[1] == 2 * out.audio.shape[1]
    with pytest.raises(AssertionError):
        m("x" * 511, ref_s)                                            # context length (kokoro.py:131-134)
    with pytest.raises(NotImplementedError):
        Model(ModelConfig())("hi", ref_s)                              # no front end: out of scope, says so


def test_pipeline_surface_and_result_tuple_compat():
    m = _model()
    p = KokoroPipeline("en-us", model=m, g2p=lambda text: (text.replace(" ", ""), None))
    assert p.lang_code == "a"
    with pytest.raises(ValueError):
        next(p("hello"))                                               # voice required (pipeline.py:365-368)
    with pytest.raises(ValueError):
        p.load_voice("af_unknown")
    p.add_voice("v", np.zeros((510, 1, 256), np.float32))
    res = list(p("ab cd\n\nefg", voice="v"))
    assert [r.phonemes for r in res] == ["abcd", "efg"] and [r.text_index for r in res] == [0, 1]
    g, ps, audio = res[0]                                              # tuple view (pipeline.py:346-356)
    assert (g, ps) == ("ab cd", "abcd") and audio.shape == (1, 600 * 2 * 6) and len(res[0]) == 3
    long = list(p("x" * 600, voice="v"))                               # > 510 phonemes -> chunked
    assert [len(r.phonemes) for r in long] == [510, 90]
    tok = list(p.generate_from_tokens("abc", voice="v"))
    assert tok[0].phonemes == "abc" and tok[0].audio.shape == (1, 600 * 2 * 5)
    with pytest.raises(ValueError):
        list(p.generate_from_tokens("x" * 511, voice="v"))
    with pytest.raises(AssertionError):
        KokoroPipeline("xx")


def test_pipeline_batches_sentences_by_length():
    m = _model()
    p = KokoroPipeline("a", model=m)
    p.add_voice("v", np.zeros((510, 1, 256), np.float32))
    outs = p.batch_tokens(["abc", "xyz", "hello", "pqr"], voice="v", max_batch=2, bucket=4)
    assert [o.shape[0] for o in outs] == [600 * 10, 600 * 10, 600 * 14, 600 * 10]
    # sorted by length, cut into batches of 2: (hello, abc) is a MIXED batch padded to 16 frames with lengths=, the two
    # remaining T=10 strings form a plain batch
    assert m.decoder.calls == [(2, 16, (14, 10)), (2, 10)]


def test_load_model_errors(tmp_path):
    with pytest.raises(FileNotFoundError):
        load_model(tmp_path / "missing")
    with pytest.raises(FileNotFoundError):
        load_config(tmp_path)
    (tmp_path / "config.json").write_text(json.dumps({"model_type": "bark"}))
    with pytest.raises(ValueError):
        load_model(tmp_path)
    (tmp_path / "config.json").write_text(json.dumps({"model_type": "kokoro"}))
    with pytest.raises(FileNotFoundError):
        load_model(tmp_path)                                           # no safetensors


@pytest.mark.gpu
def test_load_model_roundtrip_and_pipeline_on_gpu(tmp_path):
    import torch
    from safetensors.numpy import save_file
    w = synth.synth_decoder_weights(seed=2000)
    # checkpoint in torch layout (O, I, K) for weight_v / noise_convs: sanitize must transpose (istftnet.py:965-979)
    ck = {}
    for k, v in w.items():
        name = "decoder." + k if not k.startswith("decoder.") else k
        if (name.endswith("weight_v") or ("noise_convs" in name and name.endswith(".weight"))) and v.ndim == 3:
            v = np.ascontiguousarray(v.transpose(0, 2, 1))
        ck[name] = np.ascontiguousarray(v)
    ck["bert.dummy"] = np.zeros(3, np.float32)                         # front-end weights are ignored
    save_file(ck, str(tmp_path / "kokoro.safetensors"))
    (tmp_path / "config.json").write_text(json.dumps({"model_type": "kokoro", "hidden_dim": 512, "style_dim": 128,
                                                      "n_mels": 80, "unknown_key": 1}))
    model = load_model(tmp_path, frontend=SyntheticFrontend(frames_per_phoneme=2), dtype="fp32")
    p = KokoroPipeline("a", model=model)
    rng = np.random.Generator(np.random.PCG64(7))
    p.add_voice("v", (0.5 * rng.standard_normal((510, 1, 256))).astype(np.float32))
    r = next(p.generate_from_tokens("hello world", voice="v"))
    T = 2 * (11 + 2)
    assert isinstance(r.audio, torch.Tensor) and r.audio.shape == (1, 600 * T) and torch.isfinite(r.audio).all()
    # same weights straight into a Decoder give the same audio (deterministic Philox noise, same seed)
    from mlx_audio_b200.tts.models.kokoro import build_decoder
    dec = build_decoder(w, dtype="fp32")
    fe = SyntheticFrontend(frames_per_phoneme=2)
    asr, F0, N, _ = fe("hello world", None, 1)
    y = dec(asr, F0, N, p.voices["v"][10][:, :128], seed=0)
    model.decoder.seed = 0
    r2 = next(p.generate_from_tokens("hello world", voice="v"))
    torch.testing.assert_close(r2.audio, y[0], rtol=0, atol=1e-6)
    outs = p.batch_tokens(["abc", "xyz", "hello"], voice="v")
    assert [tuple(o.shape) for o in outs] == [(600 * 10,), (600 * 10,), (600 * 14,)]


def test_model_generate_and_generate_audio_write_wavs(tmp_path):
    """Model.generate (kokoro.py:269-346) and generate_audio (tts/generate.py:203-360) on the host logic."""
    import wave

    from mlx_audio_b200.tts.generate import generate_audio
    m = _model()
    m._g2p = lambda text: (text.replace(" ", ""), None)             # stand-in G2P: letters as phonemes
    pipe = m._get_pipeline("a")
    assert m._get_pipeline("a") is pipe                              # cached per language (kokoro.py:258-267)
    pipe.add_voice("af_heart", np.zeros((510, 1, 256), np.float32))
    res = list(m.generate("hello world\nsecond line"))               # default voice af_heart, split on newlines
    assert [r.segment_idx for r in res] == [0, 1]
    T0 = 2 * (10 + 2)
    assert res[0].samples == 600 * T0 and res[0].sample_rate == 24000 and res[0].token_count == 10
    assert res[0].audio.shape == (600 * T0,) and res[0].audio_duration == "00:00:00.600"
    assert res[0].audio_samples["samples"] == 600 * T0 and res[0].processing_time_seconds >= 0
    files = generate_audio("hello world\nsecond line", m, file_prefix=str(tmp_path / "out"), verbose=False)
    assert [os.path.basename(f) for f in files] == ["out_000.wav", "out_001.wav"]
    with wave.open(files[0], "rb") as f:
        assert f.getframerate() == 24000 and f.getnframes() == 600 * T0 and f.getsampwidth() == 2
    joined = generate_audio("hello world\nsecond line", m, file_prefix=str(tmp_path / "all"), join_audio=True,
                            verbose=False)
    with wave.open(joined[0], "rb") as f:
        assert f.getnframes() == 600 * (T0 + 2 * (10 + 2))
    with pytest.raises(ValueError):
        generate_audio("x", m, audio_format="flac")
