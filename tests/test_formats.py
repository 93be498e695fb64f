"""Wire / disk formats either side of the path: voice packs, safetensors + sanitize, PCM16 WAV."""
import os
import wave

import numpy as np
import pytest
import torch

from mlx_audio_b200 import synth
from mlx_audio_b200.tts import audio_io
from mlx_audio_b200.tts.models.kokoro.pipeline import KokoroPipeline
from mlx_audio_b200.tts.models.kokoro.voice import load_voice_tensor, style_for
from oracle import ref_loader


@pytest.fixture()
def pack():
    rng = np.random.Generator(np.random.PCG64(5))
    return rng.standard_normal((510, 1, 256), dtype=np.float32)


def test_voice_pack_pt_roundtrip_and_reference_unpickler(tmp_path, pack):
    p = str(tmp_path / "af_test.pt")
    torch.save(torch.from_numpy(pack), p)
    got = load_voice_tensor(p)
    assert got.dtype == np.float32 and got.shape == (510, 1, 256)
    np.testing.assert_array_equal(got, pack)
    assert style_for(got, "abc").shape == (1, 256)
    np.testing.assert_array_equal(style_for(got, "abc"), pack[2])
    with pytest.raises(ValueError):
        style_for(got, "")
    if ref_loader.reference_available():     # the reference's torch-free unpickler reads the same bytes
        import importlib.util
        spec = importlib.util.spec_from_file_location(
            "ref_voice", os.path.join(ref_loader.REFERENCE_ROOT, "mlx_audio/tts/models/kokoro/voice.py"))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        np.testing.assert_array_equal(np.asarray(mod.load_voice_tensor(p)), got)


def test_voice_pack_other_containers_and_errors(tmp_path, pack):
    np.save(str(tmp_path / "v.npy"), pack)
    np.testing.assert_array_equal(load_voice_tensor(str(tmp_path / "v.npy")), pack)
    from safetensors.numpy import save_file
    save_file({"pack": pack}, str(tmp_path / "v.safetensors"))
    np.testing.assert_array_equal(load_voice_tensor(str(tmp_path / "v.safetensors")), pack)
    np.save(str(tmp_path / "bad.npy"), np.zeros((3, 256), np.float32))
    with pytest.raises(ValueError):
        load_voice_tensor(str(tmp_path / "bad.npy"))
    with pytest.raises(ValueError):
        load_voice_tensor(str(tmp_path / "v.txt"))


def test_pipeline_voice_blend(tmp_path, pack):
    a, b = str(tmp_path / "a.pt"), str(tmp_path / "b.npy")
    torch.save(torch.from_numpy(pack), a)
    np.save(b, 3 * pack)
    pipe = KokoroPipeline("a", model=None)
    blend = pipe.load_voice(f"{a},{b}")                       # pipeline.py:154-163: mean of the packs
    np.testing.assert_allclose(blend, 2 * pack, rtol=1e-6)
    with pytest.raises(ValueError):
        pipe.load_voice("af_heart")                           # hub download is outside the path


def test_wav_pcm16_roundtrip(tmp_path):
    t = np.arange(2400, dtype=np.float32) / 24000
    x = (0.5 * np.sin(2 * np.pi * 440 * t)).astype(np.float32)
    x[:4] = [1.5, -1.5, 1.0, -1.0]                           # clipping and full scale
    p = str(tmp_path / "a.wav")
    assert audio_io.save_wav(p, torch.from_numpy(x), 24000) == 2400
    with wave.open(p, "rb") as f:
        assert (f.getnchannels(), f.getsampwidth(), f.getframerate(), f.getnframes()) == (1, 2, 24000, 2400)
    y, sr = audio_io.load_wav(p)
    assert sr == 24000 and np.abs(y[4:] - x[4:]).max() <= 1.0 / 32768
    pcm = audio_io.to_pcm16(x)
    assert pcm[0] == 32767 and pcm[1] == -32768 and pcm[2] == 32767 and pcm[3] == -32768


def test_decoder_sanitize_follows_reference_heuristic():
    """Decoder.sanitize (istftnet.py:965-979) with the reference's ``check_array_shape`` heuristic (base.py:21-34,
    pinned by tts/tests/test_base.py:42-62) reproduced as is: a 3-D ``weight_v`` is transposed (O, I, K) -> (O, K, I)
    unless its last two dimensions are equal - i.e. checkpoints are expected in torch layout, which is what the
    published Kokoro safetensors hold.  Host logic only, no device."""
    from mlx_audio_b200.tts.models.base import check_array_shape
    from mlx_audio_b200.tts.models.kokoro.istftnet import Decoder
    assert check_array_shape(np.zeros((64, 3, 3))) and not check_array_shape(np.zeros((64, 3, 4)))
    assert not check_array_shape(np.zeros((2, 3, 3))) and not check_array_shape(np.zeros((64, 3)))
    dec = Decoder.__new__(Decoder)
    w = synth.synth_decoder_weights(seed=3)
    k = "decoder.generator.resblocks.0.convs1.0.weight_v"          # MLX layout (256, 3, 256)
    v = w[k]
    vt = np.ascontiguousarray(v.transpose(0, 2, 1))                # torch layout (O, I, K) = (256, 256, 3)
    np.testing.assert_array_equal(dec.sanitize(k, vt), v)
    sq = np.zeros((64, 3, 3), np.float32)                          # K == I: passes through untouched
    assert dec.sanitize("decoder.x.weight_v", sq) is sq
    nk = "decoder.generator.noise_convs.0.weight"                  # plain Conv1d weight: always transposed (:967-968)
    nw = np.ascontiguousarray(w[nk].transpose(0, 2, 1))
    np.testing.assert_array_equal(dec.sanitize(nk, nw), w[nk])
    b = w["decoder.generator.conv_post.bias"]
    assert dec.sanitize("decoder.generator.conv_post.bias", b) is b


def test_read_checkpoint_decodes_bf16_and_skips_front_end(tmp_path):
    """The reference's default checkpoint is bf16 (tts/generate.py:366): the loader must decode bfloat16 tensors (the
    NumPy safetensors loader cannot) and must not touch the out-of-scope front-end keys."""
    from safetensors.torch import save_file
    from mlx_audio_b200.tts.utils import read_checkpoint
    g = torch.Generator().manual_seed(3)
    w = torch.randn(8, 3, 4, generator=g)
    ck = {"decoder.generator.conv_post.weight_v": w.to(torch.bfloat16),
          "decoder.generator.conv_post.bias": torch.randn(8, generator=g),
          "predictor.F0_proj.bias": torch.ones(1, dtype=torch.float16),
          "bert.embeddings.word_embeddings.weight": torch.zeros(4, 4, dtype=torch.bfloat16)}
    save_file(ck, str(tmp_path / "m.safetensors"))
    got = read_checkpoint([tmp_path / "m.safetensors"])
    assert sorted(got) == ["decoder.generator.conv_post.bias", "decoder.generator.conv_post.weight_v",
                           "predictor.F0_proj.bias"]
    assert all(v.dtype == np.float32 for v in got.values())
    np.testing.assert_array_equal(got["decoder.generator.conv_post.weight_v"], w.to(torch.bfloat16).float().numpy())
