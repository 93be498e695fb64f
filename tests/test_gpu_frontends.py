"""GPU parity of the shared-DSP consumers against goldens minted from the reference's own
codec/models/vocos/{mel,vocos}.py, tts/models/indextts/mel.py and stt/models/whisper/audio.py."""
import os

import numpy as np
import pytest
import torch

from mlx_audio_b200 import synth
from mlx_audio_b200.codec.models.vocos import ISTFTHead
from mlx_audio_b200.codec.models.vocos import log_mel_spectrogram as vocos_mel
from mlx_audio_b200.stt.models.whisper import audio as whisper
from mlx_audio_b200.tts.models.indextts.mel import log_mel_spectrogram as indextts_mel
from oracle import dsp

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def g(golden_dir):
    return np.load(os.path.join(golden_dir, "frontends.npz"))


@pytest.fixture(scope="module")
def sig(g):
    return synth.synth_audio(1, int(g["n"]), seed=int(g["audio_seed"]))[0]


def _cmp(got, ref, mean_tol=1e-4, max_tol=2e-3):
    got = got.cpu().numpy()
    assert got.shape == ref.shape, (got.shape, ref.shape)
    d = np.abs(got - ref)
    assert d.mean() <= mean_tol and d.max() <= max_tol, (d.mean(), d.max())


def test_vocos_and_indextts_log_mel(g, sig):
    _cmp(vocos_mel(sig), g["vocos_mel"])
    _cmp(vocos_mel(sig, n_mels=80, padding=300), g["vocos_mel_pad"])
    _cmp(indextts_mel(sig), g["indextts_mel"])
    # batched extension: rows agree with the single-signal call
    b = vocos_mel(torch.from_numpy(np.stack([sig, sig[::-1].copy()])))
    assert tuple(b.shape) == (2, 93, 100)
    _cmp(b[0:1], g["vocos_mel"])


def test_whisper_log_mel(g, sig):
    _cmp(whisper.log_mel_spectrogram(sig), g["whisper_mel"])
    _cmp(whisper.log_mel_spectrogram(sig, n_mels=128, padding=480), g["whisper_mel_128"])
    x = whisper.pad_or_trim(torch.from_numpy(sig), 16000)
    assert tuple(x.shape) == (16000,)
    x = whisper.pad_or_trim(torch.from_numpy(sig), 30000)
    assert tuple(x.shape) == (30000,) and float(x[24000:].abs().max()) == 0.0
    with pytest.raises(TypeError):
        whisper.log_mel_spectrogram("file.wav")


def test_parakeet_log_mel(g, sig):
    from mlx_audio_b200.stt.models.parakeet.audio import PreprocessArgs, log_mel_spectrogram
    a = PreprocessArgs(16000, "per_feature", 0.025, 0.01, "hann", 80, 512, 0.0)
    _cmp(log_mel_spectrogram(sig, a), g["parakeet_mel"], mean_tol=2e-4, max_tol=5e-3)
    a2 = PreprocessArgs(16000, "all", 0.025, 0.01, "hamming", 128, 512, 0.0, pad_to=30000)
    _cmp(log_mel_spectrogram(sig, a2), g["parakeet_mel_global"], mean_tol=2e-4, max_tol=5e-3)
    b = log_mel_spectrogram(torch.from_numpy(np.stack([sig, 0.5 * sig])), a)
    assert tuple(b.shape) == (2, 151, 80)
    _cmp(b[0:1], g["parakeet_mel"], mean_tol=2e-4, max_tol=5e-3)


def test_s3_and_spark_mel(g, sig):
    from mlx_audio_b200.codec.models.s3.utils import log_mel_spectrogram as s3_mel
    from mlx_audio_b200.tts.models.spark.bicodec import mel_spectrogram as spark_mel
    _cmp(s3_mel(sig), g["s3_mel"])
    b = s3_mel(torch.from_numpy(np.stack([sig, 0.3 * sig])))
    assert tuple(b.shape) == (2, 128, 151)
    _cmp(b[0], g["s3_mel"])
    got = spark_mel(sig).cpu().numpy()
    ref = g["spark_mel"]
    assert got.shape == ref.shape
    assert np.abs(got - ref).max() <= 1e-4 * max(1.0, float(np.abs(ref).max()))


def test_vocos_istft_head(g):
    rng = np.random.Generator(np.random.PCG64(int(g["head_seed"])))
    dim, Lh = int(g["head_dim"]), int(g["head_frames"])
    hw = (rng.standard_normal((1026, dim), dtype=np.float32) * np.float32(0.2)).astype(np.float32)
    hb = (rng.standard_normal((1026,), dtype=np.float32) * np.float32(0.1)).astype(np.float32)
    hb[:513] -= np.float32(1.0)
    hx = rng.standard_normal((1, Lh, dim), dtype=np.float32)
    head = ISTFTHead(dim, 1024, 256).load_weights({"out.weight": hw, "out.bias": hb})
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        y = head(hx)
        torch.cuda.synchronize()
        ref = g["istft_head"]
        assert tuple(y.shape) == ref.shape
        assert np.abs(y.cpu().numpy() - ref).max() <= 1e-4 * max(1.0, float(np.abs(ref).max()))
        # a longer, batched call against the oracle restatement
        hx2 = rng.standard_normal((3, 150, dim), dtype=np.float32)
        y2 = head(hx2)
        for b in range(3):
            r = dsp.istft_head_vocos(hx2[b:b + 1], hw, hb, 1024, 256)
            assert np.abs(y2[b].cpu().numpy() - r).max() <= 1e-4 * max(1.0, float(np.abs(r).max()))
    finally:
        torch.backends.cuda.matmul.allow_tf32 = prev
    with pytest.raises(ValueError):
        head(np.zeros((1, 4, dim + 1), np.float32))
    with pytest.raises(ValueError):
        ISTFTHead(dim, 1024, 256).load_weights({"out.weight": np.zeros((3, 3), np.float32)})


def test_snake_activations(golden_dir):
    from mlx_audio_b200.codec.models.bigvgan.activation import Snake, SnakeBeta
    from mlx_audio_b200.codec.models.descript.nn.layers import Snake1d, snake
    gs = np.load(os.path.join(golden_dir, "snake.npz"))
    rs = np.random.Generator(np.random.PCG64(41))
    xs = (2.0 * rs.standard_normal((2, 12, 50), dtype=np.float32)).astype(np.float32)
    al = rs.uniform(0.3, 2.0, 12).astype(np.float32)
    be = rs.uniform(0.3, 2.0, 50).astype(np.float32)
    al50 = rs.uniform(-1.0, 1.0, 50).astype(np.float32)
    sn = Snake(12); sn.alpha = torch.from_numpy(al)
    assert np.abs(sn(xs).cpu().numpy() - gs["snake"]).max() < 5e-6
    snl = Snake(12, alpha_logscale=True); snl.alpha = torch.from_numpy(al - 1.0)
    assert np.abs(snl(xs).cpu().numpy() - gs["snake_log"]).max() < 5e-6
    sb = SnakeBeta(50, alpha_logscale=True); sb.alpha = torch.from_numpy(al50); sb.beta = torch.from_numpy(be - 1.0)
    assert np.abs(sb(xs).cpu().numpy() - gs["snakebeta_log"]).max() < 5e-6
    assert np.abs(snake(xs, be.reshape(1, 1, 50)).cpu().numpy() - gs["dac_snake"]).max() < 5e-6
    s1 = Snake1d(50)                                        # alpha = 1: x + sin(x)^2 / (1 + 1e-9)
    assert np.abs(s1(xs).cpu().numpy() - dsp.snake(xs, np.ones(50, np.float32))).max() < 5e-6
    with pytest.raises(ValueError):
        sn(np.zeros((2, 13, 5), np.float32))


def test_dac_wnconv1d(golden_dir):
    from mlx_audio_b200.codec.models.descript.nn.layers import WNConv1d
    gs = np.load(os.path.join(golden_dir, "snake.npz"))
    rs = np.random.Generator(np.random.PCG64(41))
    rs.standard_normal((2, 12, 50), dtype=np.float32); rs.uniform(0.3, 2.0, 12); rs.uniform(0.3, 2.0, 50)
    rs.uniform(-1.0, 1.0, 50)                                    # the draws of the Snake cases come first
    wv = rs.standard_normal((10, 5, 6), dtype=np.float32)
    wg = rs.uniform(0.5, 1.5, (10, 1, 1)).astype(np.float32)
    wb = (0.1 * rs.standard_normal((10,), dtype=np.float32)).astype(np.float32)
    xw = rs.standard_normal((2, 37, 6), dtype=np.float32)
    conv = WNConv1d(6, 10, 5, stride=2, padding=3, dilation=2)
    conv.weight_v, conv.weight_g, conv.bias = torch.from_numpy(wv), torch.from_numpy(wg), torch.from_numpy(wb)
    y = conv(xw)
    torch.cuda.synchronize()
    ref = gs["wnconv"]
    assert tuple(y.shape) == ref.shape == (2, 18, 10)
    assert np.abs(y.cpu().numpy() - ref).max() < 1e-5
    with pytest.raises(NotImplementedError):
        WNConv1d(6, 12, 3, groups=2)
    with pytest.raises(ValueError):
        conv(np.zeros((1, 4, 7), np.float32))
