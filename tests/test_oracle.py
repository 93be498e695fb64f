"""CPU tests: the oracle (oracle/) against the golden vectors minted from the reference's
own source (tests/golden/make_golden.py), the reference's KATs and independent libraries."""
import os

import numpy as np
import pytest
import torch

from mlx_audio_b200 import synth
from oracle import dsp, interp, ref_loader
from oracle.istftnet import DecoderOracle


def _g(golden_dir, name):
    return np.load(os.path.join(golden_dir, name), allow_pickle=False)


# ---------------------------------------------------------------- interpolate (reference KATs)
def test_interpolate_kats_from_reference_tests():
    # mlx_audio/tts/tests/test_interpolate.py:40-84
    x = np.array([[[1.0, 2.0, 3.0, 4.0]]], np.float32)
    np.testing.assert_allclose(interp.interpolate1d(x, 8, "nearest"), [[[1, 1, 2, 2, 3, 3, 4, 4]]], rtol=1e-5)
    np.testing.assert_allclose(interp.interpolate1d(x, 2, "nearest"), [[[1, 3]]], rtol=1e-5)
    x = np.array([[[1.0, 3.0, 5.0, 7.0]]], np.float32)
    np.testing.assert_allclose(interp.interpolate1d(x, 7, "linear", True), [[[1, 2, 3, 4, 5, 6, 7]]], rtol=1e-5)
    assert interp.interpolate1d(x, 7, "linear", False).shape == (1, 1, 7)
    np.testing.assert_allclose(interp.interpolate1d(np.array([[[5.0]]], np.float32), 4, "linear"), [[[5, 5, 5, 5]]])
    with pytest.raises(ValueError):
        interp.interpolate(np.zeros((2, 3)), size=4)
    with pytest.raises(ValueError):
        interp.interpolate(np.zeros((2, 3, 4)), size=8, scale_factor=2)
    with pytest.raises(ValueError):
        interp.interpolate(np.zeros((2, 3, 4)))
    with pytest.raises(ValueError):
        interp.interpolate(np.zeros((2, 3, 4, 5)), size=8)
    assert interp.interpolate(np.zeros((2, 3, 4), np.float32), size=8).shape == (2, 3, 8)
    assert interp.interpolate(np.zeros((2, 3, 4), np.float32), scale_factor=2).shape == (2, 3, 8)


def test_interpolate_golden(golden_dir):
    g = _g(golden_dir, "interpolate.npz")
    eq = np.testing.assert_array_equal
    eq(interp.interpolate1d(g["x4"], 8, "nearest"), g["nearest_4_to_8"])
    eq(interp.interpolate1d(g["x4"], 2, "nearest"), g["nearest_4_to_2"])
    eq(interp.interpolate1d(g["x4b"], 7, "linear", True), g["linear_ac_4_to_7"])
    eq(interp.interpolate1d(g["x4b"], 7, "linear", False), g["linear_4_to_7"])
    eq(interp.interpolate(g["xr"], scale_factor=5, mode="linear"), g["linear_up_x5"])
    eq(interp.interpolate(g["xr"], size=9, mode="linear"), g["linear_down_37_to_9"])
    eq(interp.interpolate(g["xr"], size=50, mode="nearest"), g["nearest_37_to_50"])


def test_interpolate_wraps_negative_index_unlike_torch():
    # interpolate.py:97-102: floor(x) == -1 is used as a gather index -> last element (SURVEY 8(a) row 7 quirk ii)
    x = np.arange(1, 9, dtype=np.float32)[None, None, :]
    ours = interp.interpolate(x, scale_factor=4, mode="linear")[0, 0]
    ref = torch.nn.functional.interpolate(torch.from_numpy(x), scale_factor=4, mode="linear", align_corners=False)[0, 0].numpy()
    assert not np.allclose(ours[:2], ref[:2])      # first scale/2 outputs blend in the LAST element
    np.testing.assert_allclose(ours[2:], ref[2:], rtol=1e-5, atol=1e-6)
    assert ours[0] == pytest.approx(8 * (1 - 0.625) + 1 * 0.625)


# ---------------------------------------------------------------- stft / istft / mel
STFT_CASES = {
    "1024_256": dict(n_fft=1024, hop_length=256),
    "400_160": dict(n_fft=400, hop_length=160),
    "20_5": dict(n_fft=20, hop_length=5),
    "800_default": dict(),
    "512_128_win400_hamming": dict(n_fft=512, hop_length=128, win_length=400, window="hamming"),
    "256_64_constant": dict(n_fft=256, hop_length=64, pad_mode="constant"),
    "256_64_nocenter": dict(n_fft=256, hop_length=64, center=False),
}


@pytest.mark.parametrize("tag", list(STFT_CASES))
def test_stft_golden(golden_dir, tag):
    g = _g(golden_dir, "dsp.npz")
    sig = synth.synth_audio(1, int(g["n"]), seed=int(g["seed"]))[0]
    S = dsp.stft(sig, **STFT_CASES[tag])
    ref = g[f"stft_{tag}"]
    assert S.shape == ref.shape
    np.testing.assert_allclose(S, ref, rtol=0, atol=2e-5 * max(1.0, float(np.abs(ref).max())))


@pytest.mark.parametrize("tag,kw", [("1024_256", dict(hop_length=256, win_length=1024)),
                                    ("400_160", dict(hop_length=160, win_length=400)),
                                    ("20_5", dict(hop_length=5, win_length=20))])
def test_istft_golden(golden_dir, tag, kw):
    g = _g(golden_dir, "dsp.npz")
    y = dsp.istft(g[f"stft_{tag}"].T, **kw)
    assert y.shape == g[f"istft_{tag}"].shape
    np.testing.assert_allclose(y, g[f"istft_{tag}"], atol=2e-6)


def test_istft_length_and_explicit_window_golden(golden_dir):
    g = _g(golden_dir, "dsp.npz")
    S = g["stft_1024_256"]
    np.testing.assert_allclose(dsp.istft(S.T, hop_length=256, win_length=1024, length=5000),
                               g["istft_1024_256_length5000"], atol=2e-6)
    np.testing.assert_allclose(dsp.istft(S.T, hop_length=256, win_length=1024, window=dsp.hanning(1024)),
                               g["istft_1024_256_symwin"], atol=2e-6)


def test_windows_and_mel_golden(golden_dir):
    g = _g(golden_dir, "dsp.npz")
    for name in ("hanning", "hamming", "blackman", "bartlett"):
        np.testing.assert_array_equal(getattr(dsp, name)(20), g[f"win_{name}_20"])
    np.testing.assert_allclose(dsp.mel_filters(24000, 1024, 80), g["mel_24000_1024_80_htk"], atol=1e-7)
    np.testing.assert_allclose(dsp.mel_filters(24000, 1024, 100, norm=None, mel_scale="htk"), g["mel_24000_1024_100_htk"], atol=1e-7)
    np.testing.assert_allclose(dsp.mel_filters(16000, 400, 80, norm="slaney", mel_scale=None), g["mel_16000_400_80_slaney"], atol=1e-9)
    np.testing.assert_allclose(dsp.mel_filters(16000, 400, 128, norm="slaney", mel_scale=None), g["mel_16000_400_128_slaney"], atol=1e-9)
    np.testing.assert_allclose(dsp.mel_filters(22050, 512, 40, f_min=50.0, f_max=8000.0), g["mel_22050_512_40_fmin_fmax"], atol=1e-7)


def test_vocos_shape_pin(golden_dir):
    # codec/tests/test_vocos.py:61-73: 120000 samples -> 469 frames; drop last -> istft -> 119552 samples
    g = _g(golden_dir, "dsp.npz")
    S = dsp.stft(np.zeros(120000, np.float32), n_fft=1024, hop_length=256)
    assert S.shape[0] == 469 == int(g["vocos_frames"])
    y = dsp.istft(S[:-1].T, hop_length=256, win_length=1024)
    assert y.shape[0] == 119552 == int(g["vocos_istft_len"])


def test_stft_errors():
    x = np.zeros(100, np.float32)
    with pytest.raises(ValueError):
        dsp.stft(x, window="nope")
    with pytest.raises(ValueError):
        dsp.stft(x, n_fft=64, pad_mode="edge")
    with pytest.raises(ValueError):
        dsp.stft(np.zeros(10, np.float32), n_fft=64, center=False)
    with pytest.raises(ValueError):
        dsp.istft(np.zeros((33, 4), np.complex64), window="nope")


def test_stft_against_torch_and_istft_sum_w_normaliser():
    sig = synth.synth_audio(1, 8000, seed=11)[0]
    S = dsp.stft(sig, n_fft=512, hop_length=128)
    St = torch.stft(torch.from_numpy(sig), 512, 128, 512, window=torch.hann_window(512, periodic=False),
                    center=True, pad_mode="reflect", return_complex=True).T.numpy()
    np.testing.assert_allclose(S, St, atol=2e-4)
    # the reference divides by sum(w), not sum(w^2): 0.75x torch.istft for periodic Hann, hop=n/4 (utils.py:142-150)
    y = dsp.istft(S.T, hop_length=128, win_length=512)
    yt = torch.istft(torch.from_numpy(St.T.copy()), 512, 128, 512, window=torch.hann_window(512, periodic=True),
                     center=True).numpy()
    mid = slice(600, 7000)
    np.testing.assert_allclose(y[mid], 0.75 * yt[mid], atol=2e-5)


def test_mel_filters_against_torchaudio():
    ta = pytest.importorskip("torchaudio")
    fb = ta.functional.melscale_fbanks(513, 0.0, 12000.0, 80, 24000, norm=None, mel_scale="htk").T.numpy()
    np.testing.assert_allclose(dsp.mel_filters(24000, 1024, 80), fb, atol=2e-4)
    fb = ta.functional.melscale_fbanks(201, 0.0, 8000.0, 80, 16000, norm="slaney", mel_scale="slaney").T.numpy()
    np.testing.assert_allclose(dsp.mel_filters(16000, 400, 80, norm="slaney", mel_scale=None), fb, atol=2e-6)


# ---------------------------------------------------------------- decoder
def test_decoder_golden(golden_dir, synth_weights):
    g = _g(golden_dir, "decoder_b2_t16.npz")
    B, T = int(g["batch"]), int(g["frames"])
    inp = synth.synth_decoder_inputs(B, T, seed=int(g["input_seed"]))
    o = DecoderOracle(synth_weights)
    y = o(inp["asr"], inp["F0_curve"], inp["N"], inp["s"], inp["noise"], keep_taps=True)
    assert y.shape == (B, 1, 600 * T) == g["audio"].shape
    # harmonic source is reproduced bit for bit (numpy float32, same op order; rand_ini is dead)
    np.testing.assert_array_equal(o.taps["har_source"], g["har_source"])
    np.testing.assert_allclose(o.taps["har"][:, :11], g["har_spec"], atol=1e-6)
    np.testing.assert_allclose(o.taps["har"][:, 11:], g["har_phase"], atol=1e-5)
    assert np.abs(y - g["audio"]).max() < 2e-5


def test_decoder_shapes_and_determinism(synth_weights):
    inp = synth.synth_decoder_inputs(1, 8, seed=5)
    o = DecoderOracle(synth_weights)
    a = o(inp["asr"], inp["F0_curve"], inp["N"], inp["s"], inp["noise"])
    b = o(inp["asr"], inp["F0_curve"], inp["N"], inp["s"], inp["noise"])
    assert a.shape == (1, 1, 4800) and np.isfinite(a).all()
    np.testing.assert_array_equal(a, b)


def test_harmonic_source_rand_ini_is_dead(synth_weights):
    # istftnet.py:563-565 perturbs only sample 0, which the /300 linear resample never reads
    inp = synth.synth_decoder_inputs(2, 6, seed=9)
    o = DecoderOracle(synth_weights)
    a = o.harmonic_source(inp["F0_curve"], inp["noise"])
    b = o.harmonic_source(inp["F0_curve"], inp["noise"], rand_ini=np.full((2, 9), 0.37, np.float32))
    np.testing.assert_array_equal(a, b)


def test_fp64_shadow_close(synth_weights):
    inp = synth.synth_decoder_inputs(1, 12, seed=21)
    a = DecoderOracle(synth_weights)(inp["asr"], inp["F0_curve"], inp["N"], inp["s"], inp["noise"])
    b = DecoderOracle(synth_weights, dtype=torch.float64)(inp["asr"], inp["F0_curve"], inp["N"], inp["s"], inp["noise"])
    assert np.abs(a - b).max() < 5e-5


@pytest.mark.skipif(not ref_loader.reference_available(), reason="reference tree not present")
def test_oracle_matches_reference_source_live(synth_weights):
    inp = synth.synth_decoder_inputs(1, 10, seed=31)
    y = DecoderOracle(synth_weights)(inp["asr"], inp["F0_curve"], inp["N"], inp["s"], inp["noise"])
    yr = ref_loader.run_reference_decoder(synth_weights, synth.KOKORO_DECODER, inp["asr"], inp["F0_curve"], inp["N"],
                                          inp["s"], inp["noise"])
    assert np.abs(y - yr).max() < 2e-5


def test_param_specs_match_survey_counts(synth_weights):
    n = sum(v.size for v in synth_weights.values())
    assert abs(n - 53.28e6) < 0.05e6  # decoder ~53.28 M parameters
    assert synth_weights["decoder.generator.ups.0.weight_v"].shape == (512, 20, 256)
    assert synth_weights["decoder.generator.ups.0.bias"].shape == (256,)
    assert synth_weights["decoder.decode.3.pool.weight_v"].shape == (1090, 3, 1)


# ------------------------------------------------------------------ ProsodyPredictor.F0Ntrain
def test_f0ntrain_golden(golden_dir):
    from oracle.prosody import F0NOracle
    g = _g(golden_dir, "prosody_b2_l12.npz")
    w = synth.synth_prosody_weights(seed=int(g["weight_seed"]))
    inp = synth.synth_prosody_inputs(int(g["batch"]), int(g["frames"]), seed=int(g["input_seed"]))
    F0, N = F0NOracle(w)(inp["x"], inp["s"])
    assert F0.shape == g["F0"].shape == (2, 24)
    assert np.abs(F0 - g["F0"]).max() <= 1e-5 * np.abs(g["F0"]).max()
    assert np.abs(N - g["N"]).max() <= 1e-5


@pytest.mark.skipif(not ref_loader.reference_available(), reason="reference tree not present")
def test_f0ntrain_oracle_matches_reference_source_live():
    from oracle.prosody import F0NOracle
    w = synth.synth_prosody_weights(seed=7)
    inp = synth.synth_prosody_inputs(1, 9, seed=8)
    F0r, Nr = ref_loader.run_reference_f0ntrain(w, inp["x"], inp["s"])
    F0, N = F0NOracle(w)(inp["x"], inp["s"])
    assert np.abs(F0 - F0r).max() <= 1e-5 * np.abs(F0r).max() and np.abs(N - Nr).max() <= 1e-5


def test_f0ntrain_lstm_matches_cudnn_gate_order():
    # the product runs the shared bi-LSTM through torch (cuDNN on the GPU); gate order i, f, g, o must agree
    from oracle.prosody import F0NOracle
    w = synth.synth_prosody_weights(seed=9)
    inp = synth.synth_prosody_inputs(2, 7, seed=10)
    o = F0NOracle(w)
    lstm = torch.nn.LSTM(640, 256, batch_first=True, bidirectional=True)
    with torch.no_grad():
        for sfx, d in (("", "forward"), ("_reverse", "backward")):
            getattr(lstm, f"weight_ih_l0{sfx}").copy_(o.w[f"shared.Wx_{d}"])
            getattr(lstm, f"weight_hh_l0{sfx}").copy_(o.w[f"shared.Wh_{d}"])
            getattr(lstm, f"bias_ih_l0{sfx}").copy_(o.w[f"shared.bias_ih_{d}"])
            getattr(lstm, f"bias_hh_l0{sfx}").copy_(o.w[f"shared.bias_hh_{d}"])
        got = lstm(torch.from_numpy(inp["x"]).transpose(1, 2))[0].numpy()
    assert np.abs(got - o.hidden_of(inp["x"])).max() <= 2e-6


def test_prosody_sanitize_mirrors_reference():
    from mlx_audio_b200.tts.models.kokoro.modules import ProsodyPredictor
    v = np.zeros((256, 512, 3), np.float32)                      # torch layout (O, I, K)
    assert ProsodyPredictor.sanitize("predictor.F0.1.conv1.weight_v", v)["predictor.F0.1.conv1.weight_v"].shape == (256, 3, 512)
    assert ProsodyPredictor.sanitize("predictor.F0_proj.weight", np.zeros((1, 256, 1), np.float32))[
        "predictor.F0_proj.weight"].shape == (1, 1, 256)
    out = ProsodyPredictor.sanitize("predictor.shared.weight_ih_l0_reverse", np.zeros((1024, 640), np.float32))
    assert list(out) == ["predictor.shared.Wx_backward"]


# ------------------------------------------------------------------ shared-DSP consumers
def _head_inputs(g):
    rng = np.random.Generator(np.random.PCG64(int(g["head_seed"])))
    dim, Lh = int(g["head_dim"]), int(g["head_frames"])
    hw = (rng.standard_normal((1026, dim), dtype=np.float32) * np.float32(0.2)).astype(np.float32)
    hb = (rng.standard_normal((1026,), dtype=np.float32) * np.float32(0.1)).astype(np.float32)
    hb[:513] -= np.float32(1.0)
    hx = rng.standard_normal((1, Lh, dim), dtype=np.float32)
    return hw, hb, hx


def test_frontends_golden(golden_dir):
    g = _g(golden_dir, "frontends.npz")
    sig = synth.synth_audio(1, int(g["n"]), seed=int(g["audio_seed"]))[0]
    v = dsp.log_mel_vocos(sig, n_mels=100, drop_last=True)
    assert g["vocos_mel"].shape == (1,) + v.shape == (1, 93, 100)
    assert np.abs(v - g["vocos_mel"][0]).max() < 2e-4
    i = dsp.log_mel_vocos(sig, n_mels=100)
    assert g["indextts_mel"].shape == (1, 94, 100) and np.abs(i - g["indextts_mel"][0]).max() < 2e-4
    w = dsp.log_mel_whisper(sig)
    assert g["whisper_mel"].shape == w.shape == (150, 80) and np.abs(w - g["whisper_mel"]).max() < 2e-4
    hw, hb, hx = _head_inputs(g)
    y = dsp.istft_head_vocos(hx, hw, hb, 1024, 256)
    assert y.shape == g["istft_head"].shape == ((int(g["head_frames"]) - 1) * 256,)
    assert np.abs(y - g["istft_head"]).max() < 1e-5 * max(1.0, float(np.abs(g["istft_head"]).max()))


def test_parakeet_frontend_golden(golden_dir):
    g = _g(golden_dir, "frontends.npz")
    sig = synth.synth_audio(1, int(g["n"]), seed=int(g["audio_seed"]))[0]
    m = dsp.log_mel_parakeet(sig)
    assert m.shape == g["parakeet_mel"].shape == (1, 151, 80)
    assert np.abs(m - g["parakeet_mel"]).max() < 1e-3          # normalised log values, O(1)
    sp = np.concatenate([sig, np.zeros(6000, np.float32)])
    m2 = dsp.log_mel_parakeet(sp, features=128, window="hamming", normalize="all")
    assert m2.shape == g["parakeet_mel_global"].shape and np.abs(m2 - g["parakeet_mel_global"]).max() < 1e-3


def test_s3_and_spark_frontends_golden(golden_dir):
    g = _g(golden_dir, "frontends.npz")
    sig = synth.synth_audio(1, int(g["n"]), seed=int(g["audio_seed"]))[0]
    m = dsp.log_mel_s3(sig)
    assert m.shape == g["s3_mel"].shape == (128, 151) and np.abs(m - g["s3_mel"]).max() < 2e-4
    sp = dsp.mel_spark(sig)
    assert sp.shape == g["spark_mel"].shape == (1, 76, 128)
    assert np.abs(sp - g["spark_mel"]).max() < 1e-4 * max(1.0, float(np.abs(g["spark_mel"]).max()))


def _snake_inputs():
    rs = np.random.Generator(np.random.PCG64(41))
    xs = (2.0 * rs.standard_normal((2, 12, 50), dtype=np.float32)).astype(np.float32)
    al = rs.uniform(0.3, 2.0, 12).astype(np.float32)
    be = rs.uniform(0.3, 2.0, 50).astype(np.float32)
    al50 = rs.uniform(-1.0, 1.0, 50).astype(np.float32)
    return xs, al, be, al50


def test_snake_activations_golden(golden_dir):
    g = _g(golden_dir, "snake.npz")
    xs, al, be, al50 = _snake_inputs()
    assert np.abs(dsp.snake(xs, al, channels_first=True) - g["snake"]).max() < 2e-6
    assert np.abs(dsp.snake(xs, al - 1.0, logscale=True, channels_first=True) - g["snake_log"]).max() < 2e-6
    assert np.abs(dsp.snake(xs, al50, be - 1.0, logscale=True) - g["snakebeta_log"]).max() < 2e-6
    assert np.abs(dsp.snake(xs, be) - g["dac_snake"]).max() < 2e-6


def test_dac_wnconv1d_golden(golden_dir):
    """WNConv1d (codec/models/descript/nn/layers.py:15-59) restated with the oracle's conv: w = g v / ||v|| (no eps)."""
    from oracle.istftnet import conv1d_mlx
    g = _g(golden_dir, "snake.npz")
    rs = np.random.Generator(np.random.PCG64(41))
    rs.standard_normal((2, 12, 50), dtype=np.float32); rs.uniform(0.3, 2.0, 12); rs.uniform(0.3, 2.0, 50)
    rs.uniform(-1.0, 1.0, 50)
    wv = torch.from_numpy(rs.standard_normal((10, 5, 6), dtype=np.float32))
    wg = torch.from_numpy(rs.uniform(0.5, 1.5, (10, 1, 1)).astype(np.float32))
    wb = torch.from_numpy((0.1 * rs.standard_normal((10,), dtype=np.float32)).astype(np.float32))
    xw = torch.from_numpy(rs.standard_normal((2, 37, 6), dtype=np.float32))
    w = wg * wv / torch.sqrt((wv ** 2).sum(dim=(1, 2), keepdim=True))
    y = conv1d_mlx(xw, w, stride=2, padding=3, dilation=2) + wb
    assert tuple(y.shape) == g["wnconv"].shape and np.abs(y.numpy() - g["wnconv"]).max() < 1e-5


def test_reference_snac_shape_pin_on_the_shim():
    """codec/tests/test_snac.py:25-36 on oracle/mlx_shim: the reference's own SNAC decoder must produce 120 907 samples
    from 236 latent frames - this pins the shim's reading of mx.conv_transpose1d (output_padding is the sixth positional
    argument in the MLX the reference requires, so snac/layers.py:111-118 runs with output_padding = groups = 1)."""
    from oracle import ref_loader
    if not ref_loader.reference_available():
        pytest.skip("reference tree not present")
    ns = ref_loader.load_reference_snac()
    dec = ns.snac_layers.Decoder(24, 64, [8, 8, 4, 2], True, depthwise=True, attn_window_size=None)
    y = np.asarray(dec(ns.mx.array(np.zeros((1, 236, 24), np.float32))))
    assert y.shape == (1, 120907, 1)
    assert [tuple(s) for s in ns.mx.random.log[-4:]] == [(1, 1, 32), (1, 1, 16), (1, 1, 8), (1, 1, 4)]   # noise is (B, 1, C)


def test_bigvgan_activation_closed_form_vs_reference_golden(golden_dir):
    """oracle.codec.aa_activation (the clamped-index closed form the CUDA kernel evaluates) against the output of the
    reference's own UpSample1d / SnakeBeta / DownSample1d (tests/golden/make_golden_bigvgan.py), incl. sequences
    shorter than the filters."""
    from oracle import codec
    from mlx_audio_b200.codec.models.bigvgan.resample import kaiser_sinc_filter1d
    g = np.load(os.path.join(golden_dir, "bigvgan.npz"))
    f = kaiser_sinc_filter1d(0.25, 0.3, 12).numpy().reshape(-1)
    for T in (1, 2, 3, 7, 8, 9, 40):
        y = codec.aa_activation(g[f"act.T{T}.x"], g[f"act.T{T}.alpha"], g[f"act.T{T}.beta"], f, f, True)
        ref = g[f"act.T{T}.y"]
        assert y.shape == ref.shape and np.abs(y - ref).max() < 2e-6 * max(1.0, float(np.abs(ref).max())), T


def test_reference_bigvgan_runs_on_the_shim_and_matches_its_golden(golden_dir):
    """Live re-check: the reference's own BigVGAN on oracle/mlx_shim reproduces the committed golden (and yields
    T * prod(upsample_rates) samples)."""
    from oracle import ref_loader
    if not ref_loader.reference_available():
        pytest.skip("reference tree not present")
    from mlx_audio_b200 import synth
    g = np.load(os.path.join(golden_dir, "bigvgan.npz"))
    ns = ref_loader.load_reference_bigvgan()
    cfg = dict(num_mels=16, upsample_rates=[4, 2], upsample_kernel_sizes=[8, 4], upsample_initial_channel=64, resblock="1",
               resblock_kernel_sizes=[3, 7], resblock_dilation_sizes=[[1, 3, 5], [1, 3, 5]], activation="snakebeta",
               snake_logscale=True)
    model = ns.bigvgan.BigVGAN(ns.bigvgan.BigVGANConfig(**cfg))
    shapes = {k: tuple(v.shape) for k, v in model.named_parameters().items()}
    for k, v in synth.synth_named_weights({k: s for k, s in shapes.items() if not k.endswith("filter")},
                                          seed=int(g["amp1.seed"])).items():
        model.set_parameter(k, v)
    y = np.asarray(model(ns.mx.array(g["amp1.x"])))
    assert y.shape == (2, 1, 20 * 8) and np.abs(y - g["amp1.y"]).max() < 1e-6
