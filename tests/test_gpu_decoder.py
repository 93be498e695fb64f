"""GPU parity tests: source / STFT / iSTFT-head kernels, the general DSP surface and the whole decoder,
all against the CPU oracle and the golden vectors minted from the reference source."""
import os

import numpy as np
import pytest
import torch

from mlx_audio_b200 import _lib as L
from mlx_audio_b200 import synth, utils
from mlx_audio_b200.tts.models.kokoro import Decoder, build_decoder
from oracle import dsp
from oracle.istftnet import DecoderOracle
from parity_gates import BF16_MEL_L1, bf16_ok, record

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def lib():
    L.require_cuda()
    return L.load()


@pytest.fixture(scope="module")
def oracle(synth_weights):
    return DecoderOracle(synth_weights)


@pytest.fixture(scope="module")
def dec_fp32(synth_weights):
    return build_decoder(synth_weights, dtype="fp32")


@pytest.fixture(scope="module")
def dec_bf16(synth_weights, lib):
    if not lib.kk_device_is_sm100():
        pytest.skip("needs sm_100")
    return build_decoder(synth_weights, dtype="bf16")


_KEEP = []


@pytest.fixture(autouse=True)
def _release_device_temporaries():
    yield
    torch.cuda.synchronize()
    _KEEP.clear()


def dev(x):
    """Host array -> CUDA tensor that stays alive until the end of the test (raw pointers are passed to C)."""
    t = torch.as_tensor(np.asarray(x)).to("cuda", torch.float32).contiguous()
    _KEEP.append(t)
    return t


# ------------------------------------------------------------------ harmonic source + STFT(20,5)
def test_hnsf_source_and_stft_n20(lib, oracle, synth_weights):
    B, T = 3, 24
    inp = synth.synth_decoder_inputs(B, T, seed=77)
    ref_src = oracle.harmonic_source(inp["F0_curve"], inp["noise"])
    lw = dev(synth_weights["decoder.generator.m_source.l_linear.weight"].reshape(-1))
    lb = dev(synth_weights["decoder.generator.m_source.l_linear.bias"].reshape(-1))
    ws = torch.empty(B, 9, 2 * T, device="cuda")
    src = torch.empty(B, 600 * T, device="cuda")
    L.check(lib.kk_hnsf_source(L.ptr(dev(inp["F0_curve"])), L.ptr(dev(inp["noise"])), 0, None, L.ptr(lw), L.ptr(lb), L.ptr(ws),
                               L.ptr(src), B, 2 * T, 300, None, L.stream()))
    got = src.cpu().numpy()
    # the phase chain is bit-faithful; sin / tanh / the 9-term dot product differ by a few ulp
    assert np.abs(got - ref_src).max() < 2e-6
    har = torch.empty(B, 120 * T + 1, 24, device="cuda")
    L.check(lib.kk_stft_n20(L.ptr(dev(ref_src)), L.ptr(har), 24, B, 600 * T, None, 0, L.stream()))
    mag, ph = oracle.stft_transform(ref_src)
    h = har.cpu().numpy()
    np.testing.assert_allclose(h[:, :, :11].transpose(0, 2, 1), mag, atol=2e-6)
    # phase: compare on the unit circle (|X| weighted) so that near-zero bins do not dominate
    dphi = np.angle(np.exp(1j * (h[:, :, 11:22].transpose(0, 2, 1) - ph)))
    assert (np.abs(dphi) * np.maximum(mag, 1e-3)).max() < 5e-6
    assert set(np.unique(h[:, :, 11])) <= {np.float32(0.0), np.float32(np.pi)}  # DC phase is exactly 0 or +pi


def test_hnsf_source_philox_noise_statistics(lib, synth_weights):
    B, T = 2, 40
    f0 = np.zeros((B, 2 * T), np.float32)  # unvoiced: output = tanh(sum_k w_k * (0.1/3) n_k + b)
    lw = synth_weights["decoder.generator.m_source.l_linear.weight"].reshape(-1)
    ws = torch.empty(B, 9, 2 * T, device="cuda")
    src = torch.empty(B, 600 * T, device="cuda")
    L.check(lib.kk_hnsf_source(L.ptr(dev(f0)), None, 1234, None, L.ptr(dev(lw)), L.ptr(dev(np.zeros(1, np.float32))), L.ptr(ws),
                               L.ptr(src), B, 2 * T, 300, None, L.stream()))
    x = np.arctanh(src.cpu().numpy().astype(np.float64))
    sigma = (0.1 / 3) * np.sqrt((lw.astype(np.float64) ** 2).sum())
    assert abs(x.mean()) < 4 * sigma / np.sqrt(x.size) + 1e-6
    assert abs(x.std() / sigma - 1) < 0.02
    assert abs(np.corrcoef(x[0, :-1], x[0, 1:])[0, 1]) < 0.02
    src2 = torch.empty_like(src)
    L.check(lib.kk_hnsf_source(L.ptr(dev(f0)), None, 1235, None, L.ptr(dev(lw)), L.ptr(dev(np.zeros(1, np.float32))), L.ptr(ws),
                               L.ptr(src2), B, 2 * T, 300, None, L.stream()))
    assert not torch.equal(src, src2)


def test_istft_head(lib, oracle):
    r = np.random.Generator(np.random.PCG64(5))
    B, F_ = 2, 120 * 7 + 1
    y = r.standard_normal((B, F_, 24), dtype=np.float32)
    y[:, :, :11] = y[:, :, :11] * 0.5 - 2.0
    audio = torch.empty(B, 5 * (F_ - 1), device="cuda")
    L.check(lib.kk_istft_head_n20(L.ptr(dev(y)), 24, L.ptr(audio), B, F_, None, 0, 0, L.stream()))
    spec = np.exp(y[:, :, :11]).transpose(0, 2, 1)
    phase = np.sin(y[:, :, 11:22]).transpose(0, 2, 1)
    ref = oracle.stft_inverse(spec.astype(np.float32), phase.astype(np.float32))[:, 0]
    assert audio.shape == ref.shape
    np.testing.assert_allclose(audio.cpu().numpy(), ref, atol=2e-6)


# ------------------------------------------------------------------ general stft / istft / mel surface
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
def test_stft_matches_reference_golden(golden_dir, tag):
    g = np.load(os.path.join(golden_dir, "dsp.npz"))
    sig = synth.synth_audio(1, int(g["n"]), seed=int(g["seed"]))[0]
    S = utils.stft(torch.from_numpy(sig), **STFT_CASES[tag])
    ref = g[f"stft_{tag}"]
    assert tuple(S.shape) == ref.shape and S.dtype == torch.complex64
    err = np.abs(S.cpu().numpy() - ref).max()
    assert err < 1e-4 * max(1.0, float(np.abs(ref).max())), err


@pytest.mark.parametrize("tag,kw", [("1024_256", dict(hop_length=256, win_length=1024)),
                                    ("400_160", dict(hop_length=160, win_length=400)),
                                    ("20_5", dict(hop_length=5, win_length=20))])
def test_istft_matches_reference_golden(golden_dir, tag, kw):
    g = np.load(os.path.join(golden_dir, "dsp.npz"))
    S = torch.from_numpy(g[f"stft_{tag}"])
    y = utils.istft(S.T, **kw)
    ref = g[f"istft_{tag}"]
    assert tuple(y.shape) == ref.shape
    np.testing.assert_allclose(y.cpu().numpy(), ref, atol=1e-5)
    if tag == "1024_256":
        y2 = utils.istft(S.T, hop_length=256, win_length=1024, length=5000)
        np.testing.assert_allclose(y2.cpu().numpy(), g["istft_1024_256_length5000"], atol=1e-5)
        y3 = utils.istft(S.T, hop_length=256, win_length=1024, window=utils.hanning(1024))
        np.testing.assert_allclose(y3.cpu().numpy(), g["istft_1024_256_symwin"], atol=1e-5)


RAGGED_DSP = [dict(n_fft=1024, hop_length=256), dict(n_fft=400, hop_length=160),
              dict(n_fft=256, hop_length=64, pad_mode="constant"), dict(n_fft=256, hop_length=64, center=False)]


@pytest.mark.parametrize("kw", RAGGED_DSP, ids=lambda k: "_".join(str(v) for v in k.values()))
def test_stft_logmel_istft_ragged_lengths_match_solo_calls(kw):
    """``lengths=``: every row of a zero-padded ragged batch equals the transform of that signal
    alone - its own reflect padding, last frame and overlap-add tail - bit for bit, followed by zeros; the padding
    holds garbage.  One signal is too short to transform at all (all zeros), one ends on a frame boundary."""
    from oracle import dsp
    n_fft, hop = kw["n_fft"], kw["hop_length"]
    n = 9000
    lens = [n, 5000, n_fft + 3 * hop, n_fft // 2 + 5, n_fft // 4, 0, 7777]
    x = torch.from_numpy(synth.synth_audio(len(lens), n, seed=91)).cuda()
    for b, l in enumerate(lens):
        x[b, l:] = 1e6                                               # garbage past the end must not leak in
    for dev_lens in (False, True):
        lengths = torch.tensor(lens, dtype=torch.int32, device="cuda") if dev_lens else lens
        S = utils.stft(x, lengths=lengths, **kw)
        M = utils.log_mel_spectrogram(x, sample_rate=24000, n_mels=40, lengths=lengths, **kw) if n_fft >= 256 else None
        fr = []
        for b, l in enumerate(lens):
            try:
                solo = utils.stft(x[b, :l], **kw) if l else None
            except ValueError:
                solo = None                                          # too short: the batched call writes zeros
            f = 0 if solo is None else solo.shape[0]
            fr.append(f)
            if f:
                assert torch.equal(S[b, :f], solo), (b, l)
                assert torch.equal(M[b, :f], utils.log_mel_spectrogram(x[b, :l], sample_rate=24000, n_mels=40, **kw))
            assert not S[b, f:].abs().any() and not M[b, f:].abs().any()
        ref = dsp.stft(x[1, :lens[1]].cpu().numpy(), **kw)
        assert np.abs(S[1, :fr[1]].cpu().numpy() - ref).max() < 1e-4 * max(1.0, float(np.abs(ref).max()))
        # inverse: frames per signal -> each row is the solo synthesis, zeros after its own end
        for b, f in enumerate(fr):
            S[b, f:] = 1e6
        fl = torch.tensor(fr, dtype=torch.int32, device="cuda") if dev_lens else fr
        for ikw in (dict(), dict(length=4321)) if kw.get("center", True) else (dict(center=False),):
            y = utils.istft(S.transpose(1, 2), hop_length=hop, win_length=n_fft, lengths=fl, **ikw)
            for b, f in enumerate(fr):
                m = 0
                if f and ((f - 1) * hop > 0 or ikw):
                    solo = utils.istft(S[b, :f].T, hop_length=hop, win_length=n_fft, **ikw)
                    m = solo.shape[0]
                    assert torch.equal(y[b, :m], solo), (b, f, ikw)
                assert not y[b, m:].abs().any(), (b, f, ikw)
    with pytest.raises(ValueError):
        utils.stft(x, lengths=[n + 1] * len(lens), **kw)
    with pytest.raises(ValueError):
        utils.stft(x, lengths=[1, 2], **kw)


def test_mel_filters_and_windows_golden(golden_dir):
    g = np.load(os.path.join(golden_dir, "dsp.npz"))
    np.testing.assert_allclose(utils.mel_filters(24000, 1024, 80).cpu().numpy(), g["mel_24000_1024_80_htk"], atol=1e-7)
    np.testing.assert_allclose(utils.mel_filters(16000, 400, 80, norm="slaney", mel_scale=None).cpu().numpy(),
                               g["mel_16000_400_80_slaney"], atol=1e-9)
    for name in ("hanning", "hamming", "blackman", "bartlett"):
        np.testing.assert_array_equal(getattr(utils, name)(20).cpu().numpy(), g[f"win_{name}_20"])


def test_stft_errors_like_reference():
    x = torch.zeros(100)
    with pytest.raises(ValueError):
        utils.stft(x, window="nope")
    with pytest.raises(ValueError):
        utils.stft(x, n_fft=64, pad_mode="edge")
    with pytest.raises(ValueError):
        utils.stft(torch.zeros(10), n_fft=64, center=False)
    with pytest.raises(ValueError):
        utils.istft(torch.zeros(33, 4, dtype=torch.complex64), window="nope")


def test_vocos_shape_pin_and_batched_roundtrip():
    # codec/tests/test_vocos.py:61-73: 120000 samples -> 469 frames; [:-1] -> istft -> 119552
    S = utils.stft(torch.zeros(120000), n_fft=1024, hop_length=256)
    assert tuple(S.shape) == (469, 513)
    y = utils.istft(S[:-1].T, hop_length=256, win_length=1024)
    assert tuple(y.shape) == (119552,)
    x = torch.from_numpy(synth.synth_audio(5, 30000, seed=8)).cuda()
    Sb = utils.stft(x, n_fft=1024, hop_length=256)
    yb = utils.istft(Sb.transpose(1, 2), hop_length=256, win_length=1024)
    # sum(w) normaliser (utils.py:142-150): ~0.75 x the signal away from the edges (Hann, hop = n/4)
    mid = slice(1024, yb.shape[1] - 1024)
    np.testing.assert_allclose(yb[:, mid].cpu().numpy(), 0.75 * x[:, mid].cpu().numpy(), rtol=2e-3, atol=2e-4)
    for b in (0, 4):
        ref = dsp.istft(dsp.stft(x[b].cpu().numpy(), n_fft=1024, hop_length=256).T, hop_length=256, win_length=1024)
        np.testing.assert_allclose(yb[b].cpu().numpy(), ref, atol=1e-5)


@pytest.mark.parametrize("hop", [256, 512, 1024])
@pytest.mark.parametrize("n", [30000, 24577, 2500])
def test_istft_streaming_overlap_add_all_hops(hop, n):
    """The streaming (register-accumulator) iSTFT path: n_fft 1024 with R = n_fft / hop in {4, 2, 1}, ragged lengths
    (several blocks, warm-up frames, tail drain), with and without ``length=`` - against the oracle restatement of
    utils.istft (utils.py:104-158), which sums frames in the reference's order."""
    x = synth.synth_audio(3, n, seed=100 + hop)
    S = np.stack([dsp.stft(x[b], n_fft=1024, hop_length=hop) for b in range(3)])       # (3, frames, 513)
    St = torch.from_numpy(S).cuda().transpose(1, 2)
    y = utils.istft(St, hop_length=hop, win_length=1024)
    y5 = utils.istft(St, hop_length=hop, win_length=1024, length=n - 777)
    for b in range(3):
        ref = dsp.istft(S[b].T, hop_length=hop, win_length=1024)
        assert tuple(y[b].shape) == ref.shape
        np.testing.assert_allclose(y[b].cpu().numpy(), ref, atol=1e-5)
        ref5 = dsp.istft(S[b].T, hop_length=hop, win_length=1024, length=n - 777)
        np.testing.assert_allclose(y5[b].cpu().numpy(), ref5, atol=1e-5)


def test_log_mel_recipes():
    sig = synth.synth_audio(3, 24000, seed=12)
    got = utils.log_mel_spectrogram(torch.from_numpy(sig), n_fft=1024, hop_length=256, n_mels=80)
    for b in range(3):
        ref = dsp.log_mel_vocos(sig[b], n_fft=1024, hop_length=256, n_mels=80)
        assert tuple(got[b].shape) == ref.shape
        assert np.abs(got[b].cpu().numpy() - ref).mean() < 1e-4
        assert np.abs(got[b].cpu().numpy() - ref).max() < 2e-3
    # whisper-style front end: power spectrum, slaney mel, log10 (stt/models/whisper/audio.py:72-82)
    w = utils.log_mel_spectrogram(torch.from_numpy(sig[0]), sample_rate=16000, n_fft=400, hop_length=160, n_mels=80,
                                  power=2, log="log10", floor=1e-10, norm="slaney", mel_scale=None)
    w = w[:-1]
    w = torch.maximum(w, w.max() - 8.0)
    w = (w + 4.0) / 4.0
    ref = dsp.log_mel_whisper(sig[0])
    assert np.abs(w.cpu().numpy() - ref).max() < 2e-3


# ------------------------------------------------------------------ whole decoder
def _run(dec, inp, **kw):
    y = dec(inp["asr"], inp["F0_curve"], inp["N"], inp["s"], noise=inp["noise"], **kw)
    torch.cuda.synchronize()
    return y.cpu().numpy()


def test_decoder_fp32_matches_reference_golden(golden_dir, dec_fp32):
    g = np.load(os.path.join(golden_dir, "decoder_b2_t16.npz"))
    B, T = int(g["batch"]), int(g["frames"])
    inp = synth.synth_decoder_inputs(B, T, seed=int(g["input_seed"]))
    y = _run(dec_fp32, inp, capture_taps=True)
    taps = dec_fp32.taps(B, T)
    assert np.abs(taps["har_source"].cpu().numpy() - g["har_source"]).max() < 2e-6
    assert y.shape == g["audio"].shape == (B, 1, 600 * T)
    err = np.abs(y - g["audio"]).max()
    assert err <= 1e-3, f"fp32 waveform max-abs error {err} vs golden (north-star gate 1e-3)"


def test_decoder_fp32_config1_vs_oracle_waveform_and_mel(dec_fp32, oracle):
    # BASELINE config 1: B=1, T=200 (5 s).  Gates: waveform max-abs <= 1e-3, log-mel-80 L1 <= 1e-3.
    inp = synth.synth_decoder_inputs(1, 200, seed=1000)
    ref = oracle(inp["asr"], inp["F0_curve"], inp["N"], inp["s"], inp["noise"], keep_taps=True)
    y = _run(dec_fp32, inp, capture_taps=True)
    taps = dec_fp32.taps(1, 200)
    report = {k: float(np.abs(taps[k].cpu().numpy() - oracle.taps[k]).max()) for k in oracle.taps if k in taps}
    print("fp32 tap max-abs errors:", report)
    err = np.abs(y - ref).max()
    assert err <= 1e-3, (err, report)
    mel_ref = dsp.log_mel_vocos(ref[0, 0])
    mel_got = dsp.log_mel_vocos(y[0, 0])
    assert np.abs(mel_ref - mel_got).mean() <= 1e-3


@pytest.mark.parametrize("arm", ["ffma", "tf32"])
def test_decoder_fp32_engine_arms(arm, synth_weights, oracle, lib):
    """fp32 mode, non-default convolution arms (engine.DecoderEngine fp32_conv=): "ffma" (CUDA cores, the round-1 path)
    meets the same 1e-3 gate as the default 3xTF32 tensor-core path; plain "tf32" (10-bit operands) is measured and
    reported - SURVEY 7.2 expects it to miss 1e-3 - and only gated loosely."""
    if not lib.kk_device_is_sm100():
        pytest.skip("needs sm_100")
    dec = build_decoder(synth_weights, dtype="fp32", engine_options=dict(fp32_conv=arm))
    inp = synth.synth_decoder_inputs(2, 64, seed=2024)
    ref = oracle(inp["asr"], inp["F0_curve"], inp["N"], inp["s"], inp["noise"])
    y = _run(dec, inp)
    err = np.abs(y - ref).max()
    record(f"fp32 arm {arm} B=2 T=64", max_abs=err, rel_l2=np.linalg.norm(y - ref) / np.linalg.norm(ref),
           peak=np.abs(ref).max())
    if arm == "ffma":
        assert err <= 1e-3, err
    else:
        assert np.isfinite(y).all() and err <= 5e-2 * np.abs(ref).max(), err


def test_decoder_fp32_runs_on_tensor_cores(dec_fp32, lib):
    """The default fp32 mode issues its stride-1 convolutions through kk_conv1d_tf32_tc (tcgen05 kind::tf32, 3xTF32)."""
    if not lib.kk_device_is_sm100():
        pytest.skip("needs sm_100")
    plan = dec_fp32._plan(1, 8)
    names = [fn.__name__ for fn, _ in plan.ops if fn is not None]
    assert names.count("kk_conv1d_tf32_tc") >= 60 and names.count("kk_conv1d_f32") <= 8, \
        (names.count("kk_conv1d_tf32_tc"), names.count("kk_conv1d_f32"))


def test_decoder_bf16_vs_oracle(dec_bf16, oracle):
    # bf16 operands / fp32 accumulate.  Stated tolerance: tests/parity_gates.py (about 3x the measured errors).
    inp = synth.synth_decoder_inputs(2, 64, seed=2024)
    ref = oracle(inp["asr"], inp["F0_curve"], inp["N"], inp["s"], inp["noise"], keep_taps=True)
    y = _run(dec_bf16, inp, capture_taps=True)
    taps = dec_bf16.taps(2, 64)
    report = {k: float(np.abs(taps[k].cpu().numpy() - oracle.taps[k]).max()) for k in oracle.taps if k in taps}
    print("bf16 tap max-abs errors:", report)
    rel = np.linalg.norm(y - ref) / np.linalg.norm(ref)
    mel = np.abs(dsp.log_mel_vocos(ref[0, 0]) - dsp.log_mel_vocos(y[0, 0])).mean()
    record("bf16 B=2 T=64", max_abs=np.abs(y - ref).max(), rel_l2=rel, mel_l1=mel, peak=np.abs(ref).max())
    assert bf16_ok(y, ref) and mel <= BF16_MEL_L1


def test_decoder_bf16_config1_vs_oracle(dec_bf16, oracle):
    # BASELINE config 1's shape (B=1, T=200) in the tensor-core mode
    inp = synth.synth_decoder_inputs(1, 200, seed=1000)
    ref = oracle(inp["asr"], inp["F0_curve"], inp["N"], inp["s"], inp["noise"])
    y = _run(dec_bf16, inp)
    rel = np.linalg.norm(y - ref) / np.linalg.norm(ref)
    mel = np.abs(dsp.log_mel_vocos(ref[0, 0]) - dsp.log_mel_vocos(y[0, 0])).mean()
    record("bf16 C1 B=1 T=200", max_abs=np.abs(y - ref).max(), rel_l2=rel, mel_l1=mel, peak=np.abs(ref).max())
    assert bf16_ok(y, ref) and mel <= BF16_MEL_L1


@pytest.mark.parametrize("opts", [dict(fused=False), dict(res16=0), dict(res16=2)], ids=["unfused", "res16_0", "res16_2"])
def test_decoder_bf16_engine_arms(opts, synth_weights, oracle, lib):
    """Every non-default engine arm (engine.DecoderEngine fused= / res16=) against the oracle, same gates as the default."""
    if not lib.kk_device_is_sm100():
        pytest.skip("needs sm_100")
    dec = build_decoder(synth_weights, dtype="bf16", engine_options=opts)
    inp = synth.synth_decoder_inputs(2, 64, seed=2024)
    ref = oracle(inp["asr"], inp["F0_curve"], inp["N"], inp["s"], inp["noise"])
    y = _run(dec, inp)
    rel = np.linalg.norm(y - ref) / np.linalg.norm(ref)
    record(f"bf16 arm {opts} B=2 T=64", max_abs=np.abs(y - ref).max(), rel_l2=rel)
    assert bf16_ok(y, ref)


def test_decoder_ragged_shapes_and_errors(dec_fp32, oracle):
    for B, T in ((1, 1), (3, 7), (2, 33)):
        inp = synth.synth_decoder_inputs(B, T, seed=B * 100 + T)
        ref = oracle(inp["asr"], inp["F0_curve"], inp["N"], inp["s"], inp["noise"])
        y = _run(dec_fp32, inp)
        assert y.shape == (B, 1, 600 * T)
        assert np.abs(y - ref).max() <= 1e-3
    inp = synth.synth_decoder_inputs(1, 4, seed=1)
    with pytest.raises(ValueError):
        dec_fp32(inp["asr"][:, :100], inp["F0_curve"], inp["N"], inp["s"])
    with pytest.raises(ValueError):
        dec_fp32(inp["asr"], inp["F0_curve"][:, :-1], inp["N"], inp["s"])


def _branch_cut_utterances(oracle):
    """Utterances whose harmonic-source STFT has a bin sitting on atan2's branch cut (phase within 1e-5 of +-pi: the imaginary part is
    below the rounding noise of the transform): rounding decides between +pi and -pi there, the raw phase feeds noise_convs, and the
    two equally valid answers differ by 2 pi (SURVEY 8(a) row 8, 8(d) "branch-cut frames").  Not comparable."""
    ph = oracle.taps["har"][:, 11:]
    return set(np.nonzero((np.abs(np.abs(ph) - np.pi) < 1e-5)[:, 1:10].any(axis=(1, 2)))[0].tolist())


@pytest.mark.parametrize("B,T", [(1, 1), (3, 7), (2, 33), (5, 129)])
def test_decoder_bf16_ragged_shapes_vs_oracle(dec_bf16, dec_fp32, oracle, B, T):
    # the tensor-core path at shapes that do not fill a tile (T = 1: 121 rows at the last stage), straddle tiles, and
    # leave ragged tails in every stage; stated bf16 tolerance: tests/parity_gates.py.
    # (B, T) = (3, 7) with this seed has an utterance on the atan2 branch cut (bf16 AND fp32 CUDA paths deviate from
    # the oracle by 0.046 there): flagged utterances are excluded, and any phase deviation in them is checked to be
    # the 2-pi flip it is claimed to be.
    inp = synth.synth_decoder_inputs(B, T, seed=B * 1000 + T)
    ref = oracle(inp["asr"], inp["F0_curve"], inp["N"], inp["s"], inp["noise"], keep_taps=True)
    cut = _branch_cut_utterances(oracle)
    keep = [b for b in range(B) if b not in cut]
    assert keep, "every utterance of this case sits on the branch cut: pick another seed"
    y = _run(dec_bf16, inp)
    assert y.shape == (B, 1, 600 * T) and np.isfinite(y).all()
    rel = np.linalg.norm(y[keep] - ref[keep]) / np.linalg.norm(ref[keep])
    record(f"bf16 B={B} T={T}", max_abs=np.abs(y[keep] - ref[keep]).max(), rel_l2=rel, peak=np.abs(ref[keep]).max())
    assert bf16_ok(y[keep], ref[keep]), (np.abs(y[keep] - ref[keep]).max(), rel)
    if cut:
        y32 = _run(dec_fp32, inp, capture_taps=True)
        har = dec_fp32.taps(B, T)["har"].cpu().numpy()
        for b in cut:
            d = np.abs(har[b, 11:] - oracle.taps["har"][b, 11:])
            flips = d > 1.0
            assert np.allclose(d[flips], 2 * np.pi, atol=1e-3) and d[~flips].max() < 5e-3
        assert np.abs(y32[keep] - ref[keep]).max() <= 1e-3


def _ragged_inputs(B, T, lens, seed):
    """Padded batch whose padding holds garbage (the kernels must never let it through), plus the per-utterance views."""
    inp = synth.synth_decoder_inputs(B, T, seed=seed)
    lens = np.asarray(lens, np.int32)
    for b, t in enumerate(lens):
        inp["asr"][b, :, t:] = 37.0
        inp["F0_curve"][b, 2 * t:] = 220.0
        inp["N"][b, 2 * t:] = -5.0
        inp["noise"][b, 600 * t:] = 9.0
    singles = [dict(asr=inp["asr"][b:b + 1, :, :t], F0_curve=inp["F0_curve"][b:b + 1, :2 * t], N=inp["N"][b:b + 1, :2 * t],
                    s=inp["s"][b:b + 1], noise=inp["noise"][b:b + 1, :600 * t]) for b, t in enumerate(lens)]
    return inp, lens, singles


@pytest.mark.parametrize("mode", ["fp32", "bf16"])
@pytest.mark.parametrize("T,lens", [(40, [40, 17, 33, 1, 26]), (9, [9, 2, 8, 9, 1, 5, 7]), (130, [130, 129, 65])])
def test_decoder_ragged_lengths_in_one_launch(mode, T, lens, dec_fp32, dec_bf16, oracle):
    """Decoder.__call__(..., lengths=): a padded batch of DIFFERENT lengths decoded by ONE launch plan.  Every utterance
    must come out as if decoded alone - the reference decodes one utterance per call (kokoro.py:135-165) - although
    the padding holds garbage: per-utterance zero padding at the conv edges, instance-norm statistics over the
    utterance's own rows (istftnet.py:221-230), per-utterance interpolation / reflection edges in the harmonic
    source and its STFT, per-utterance iSTFT tail.  Lengths straddle the 128- and 256-row tiles of the tcgen05 kernels."""
    dec = dec_fp32 if mode == "fp32" else dec_bf16
    B = len(lens)
    inp, lens, singles = _ragged_inputs(B, T, lens, seed=4242 + T)
    y = dec(inp["asr"], inp["F0_curve"], inp["N"], inp["s"], noise=inp["noise"], lengths=lens)
    torch.cuda.synchronize()
    y = y.cpu().numpy()
    assert y.shape == (B, 1, 600 * T) and np.isfinite(y).all()
    worst = 0.0
    for b, t in enumerate(lens):
        one = singles[b]
        alone = _run(dec, one)
        assert alone.shape == (1, 1, 600 * t)
        # same kernels, same per-utterance arithmetic: identical up to fp64-atomic ordering in the statistics
        np.testing.assert_allclose(y[b, 0, :600 * t], alone[0, 0], rtol=0, atol=2e-5)
        assert not y[b, 0, 600 * t:].any(), "padding samples must be zero"
        ref = oracle(one["asr"], one["F0_curve"], one["N"], one["s"], one["noise"], keep_taps=True)
        if b in _branch_cut_utterances(oracle) or t < 2:
            continue
        err = np.abs(y[b, 0, :600 * t] - ref[0, 0]).max()
        worst = max(worst, err / max(np.abs(ref).max(), 1e-3))
        if mode == "fp32":
            assert err <= 1e-3, (b, t, err)
        else:
            rel = np.linalg.norm(y[b, 0, :600 * t] - ref[0, 0]) / np.linalg.norm(ref)
            assert bf16_ok(y[b, 0, :600 * t], ref[0, 0]), (b, t, err, rel)
    record(f"{mode} ragged T={T} lens={list(map(int, lens))}", max_abs_over_peak=worst)
    with pytest.raises(ValueError):
        dec(inp["asr"], inp["F0_curve"], inp["N"], inp["s"], lengths=[T + 1] * B)
    with pytest.raises(ValueError):
        dec(inp["asr"], inp["F0_curve"], inp["N"], inp["s"], lengths=[1] * (B + 1))


def test_decoder_ragged_many_utterances(dec_bf16):
    """More utterances than SMs (persistent tile loops, utterance changes inside a CTA's tile range) and CUDA-graph
    replay with new lengths (the lengths live in device memory)."""
    B, T = 170, 6
    r = np.random.Generator(np.random.PCG64(9))
    lens = r.integers(1, T + 1, size=B)
    inp, lens, singles = _ragged_inputs(B, T, lens, seed=77)
    y = dec_bf16(inp["asr"], inp["F0_curve"], inp["N"], inp["s"], noise=inp["noise"], lengths=lens).cpu().numpy()
    for b in (0, 1, 57, 99, 148, 169):
        alone = _run(dec_bf16, singles[b])
        np.testing.assert_allclose(y[b, 0, :600 * lens[b]], alone[0, 0], rtol=0, atol=2e-5)
    dec_bf16.use_cuda_graph = True
    try:
        a = dec_bf16(inp["asr"], inp["F0_curve"], inp["N"], inp["s"], seed=5, lengths=lens).cpu().numpy()
        lens2 = np.maximum(1, lens[::-1]).copy()
        b2 = dec_bf16(inp["asr"], inp["F0_curve"], inp["N"], inp["s"], seed=5, lengths=lens2).cpu().numpy()
    finally:
        dec_bf16.use_cuda_graph = False
    for b in range(0, B, 17):
        assert not a[b, 0, 600 * lens[b]:].any() and not b2[b, 0, 600 * lens2[b]:].any()
        assert a[b, 0, :600 * lens[b]].any() and b2[b, 0, :600 * lens2[b]].any()


@pytest.mark.parametrize("mode", ["fp32", "bf16"])
def test_decoder_batch_independence(mode, dec_fp32, dec_bf16):
    # no op mixes utterances: an utterance decodes identically alone and inside a batch
    dec = dec_fp32 if mode == "fp32" else dec_bf16
    inp = synth.synth_decoder_inputs(3, 20, seed=55)
    full = _run(dec, inp)
    one = {k: v[1:2] for k, v in inp.items()}
    alone = _run(dec, one)
    np.testing.assert_allclose(full[1:2], alone, atol=2e-5)


def test_decoder_default_noise_and_cuda_graph(dec_bf16):
    inp = synth.synth_decoder_inputs(2, 16, seed=3)
    a = dec_bf16(inp["asr"], inp["F0_curve"], inp["N"], inp["s"], seed=7).clone()
    b = dec_bf16(inp["asr"], inp["F0_curve"], inp["N"], inp["s"], seed=7).clone()
    c = dec_bf16(inp["asr"], inp["F0_curve"], inp["N"], inp["s"], seed=8).clone()
    assert torch.allclose(a, b, atol=1e-5) and not torch.allclose(a, c, atol=1e-3)
    assert torch.isfinite(a).all()
    dec_bf16.use_cuda_graph = True
    try:
        g1 = dec_bf16(inp["asr"], inp["F0_curve"], inp["N"], inp["s"], seed=7).clone()
        g2 = dec_bf16(inp["asr"], inp["F0_curve"], inp["N"], inp["s"], seed=7).clone()
        # the seed is read from device memory: a replay with another seed draws other noise (the reference draws fresh
        # mx.random.normal per call, istftnet.py:620) and matches the un-captured run with that seed
        g3 = dec_bf16(inp["asr"], inp["F0_curve"], inp["N"], inp["s"], seed=8).clone()
        dec_bf16.seed = 100
        d1 = dec_bf16(inp["asr"], inp["F0_curve"], inp["N"], inp["s"])       # default: a fresh seed per call
        d2 = dec_bf16(inp["asr"], inp["F0_curve"], inp["N"], inp["s"])
    finally:
        dec_bf16.use_cuda_graph = False
    torch.cuda.synchronize()
    assert torch.allclose(g1, a, atol=1e-5) and torch.allclose(g1, g2, atol=1e-5)
    assert torch.allclose(g3, c, atol=1e-5) and not torch.allclose(g3, g1, atol=1e-3)
    assert not torch.allclose(d1, d2, atol=1e-3)
    # results are fresh tensors (the reference returns a new array per call); borrow=True is the zero-copy opt-in
    assert d1.data_ptr() != d2.data_ptr()
    b1 = dec_bf16(inp["asr"], inp["F0_curve"], inp["N"], inp["s"], seed=7, borrow=True)
    b2 = dec_bf16(inp["asr"], inp["F0_curve"], inp["N"], inp["s"], seed=8, borrow=True)
    assert b1.data_ptr() == b2.data_ptr()


@pytest.mark.gpu
def test_shared_workspace_across_shapes(dec_bf16):
    """Plans of different (B, T) share one workspace (ragged corpora: one plan per length bucket).  Switching
    shapes - including a larger one that re-allocates pool buffers - must not change any result: the zero padding
    a plan relies on is restored when it becomes active again."""
    shapes = [(2, 40), (3, 24), (1, 71), (4, 90), (2, 40), (1, 71)]
    first = {}
    for B, T in shapes:
        inp = synth.synth_decoder_inputs(B, T, seed=B * 1000 + T)
        y = _run(dec_bf16, inp)
        assert y.shape == (B, 1, 600 * T) and np.isfinite(y).all()
        if (B, T) in first:
            np.testing.assert_array_equal(y, first[(B, T)])
        else:
            first[(B, T)] = y.copy()
    eng = dec_bf16.engine
    # one allocation per named buffer, however many shapes were decoded
    assert all(t.dtype == torch.uint8 for t in eng.pool.values())
    for plan in eng.plans.values():
        for name, v in plan.buffers.items():
            assert v.untyped_storage().data_ptr() == eng.pool[name].untyped_storage().data_ptr(), name


@pytest.mark.gpu
def test_streaming_decoder_matches_direct_calls(dec_bf16):
    """StreamingDecoder (copies overlapped with the forward on side streams) returns exactly what one blocking
    Decoder call per batch returns, in order, for batches of different shapes."""
    from mlx_audio_b200.streaming import StreamingDecoder
    shapes = [(2, 30), (2, 30), (1, 55), (3, 18), (2, 30)]
    batches = []
    for i, (B, T) in enumerate(shapes):
        inp = synth.synth_decoder_inputs(B, T, seed=500 + i, with_noise=False)
        batches.append({k: torch.from_numpy(inp[k]).pin_memory() for k in ("asr", "F0_curve", "N", "s")})
    dec_bf16.seed = 1234
    direct = []
    for b in batches:
        y = dec_bf16(b["asr"], b["F0_curve"], b["N"], b["s"])
        torch.cuda.synchronize()
        direct.append(y.cpu().numpy().copy())
    dec_bf16.seed = 1234
    got = [w.numpy().copy() for w in StreamingDecoder(dec_bf16, depth=2).run(iter(batches))]
    assert len(got) == len(direct)
    for a, b in zip(got, direct):
        np.testing.assert_array_equal(a, b)
    with pytest.raises(ValueError):
        StreamingDecoder(dec_bf16, depth=1)


@pytest.mark.parametrize("mode", ["fp32", "bf16"])
def test_decoder_workspace_poisoning(mode, dec_fp32, dec_bf16):
    """Stand-in for compute-sanitizer's initcheck (the tool is closed on this pool, profiles/r02_sanitizer.txt): every
    workspace buffer of the engine - activations, staging copies, padding - is filled with 0xFF bytes (NaN in fp32 and
    bf16) before a forward; the plan's declared zero-padding buffers are then re-cleared the way a plan switch does.
    Any kernel that reads a row, a pad column or a halo it did not write this forward, or that relies on stale data from
    an earlier forward, turns the NaN into the waveform.  Uniform and ragged batches; the result must equal the clean
    run bit for bit (same launches, same inputs)."""
    dec = dec_fp32 if mode == "fp32" else dec_bf16
    eng = dec.engine

    def poison():
        torch.cuda.synchronize()
        for name, t in eng.pool.items():
            if not name.startswith("in."):
                t.fill_(0xFF)
        eng.active_plan = None                    # activate() re-clears the buffers a plan declares as zero padding

    inp = synth.synth_decoder_inputs(3, 21, seed=808)
    clean = _run(dec, inp)
    poison()
    again = _run(dec, inp)
    assert np.isfinite(again).all()
    np.testing.assert_array_equal(again, clean)
    lens = [21, 5, 14]
    rin, lens, _ = _ragged_inputs(3, 21, lens, seed=809)
    y0 = dec(rin["asr"], rin["F0_curve"], rin["N"], rin["s"], noise=rin["noise"], lengths=lens).cpu().numpy()
    poison()
    y1 = dec(rin["asr"], rin["F0_curve"], rin["N"], rin["s"], noise=rin["noise"], lengths=lens).cpu().numpy()
    assert np.isfinite(y1).all()
    np.testing.assert_array_equal(y1, y0)


@pytest.mark.parametrize("mode", ["fp32", "bf16"])
def test_decoder_parallel_noise_branch(mode, dec_fp32, dec_bf16):
    """Decoder.parallel_branches: the generator's noise branch on a side stream (eager) / as a parallel graph branch
    must give the bits of the serial schedule - same launches, same buffers, only the order between independent
    kernels changes.  Repeated, so that a missing join would show as run-to-run differences."""
    dec = dec_fp32 if mode == "fp32" else dec_bf16
    inp = synth.synth_decoder_inputs(2, 30, seed=515)
    rin, lens, _ = _ragged_inputs(3, 24, [24, 9, 17], seed=516)
    old = dec.parallel_branches
    try:
        dec.parallel_branches = False
        ref = _run(dec, inp)
        ref_r = dec(rin["asr"], rin["F0_curve"], rin["N"], rin["s"], noise=rin["noise"], lengths=lens).cpu().numpy()
        refg = dec(inp["asr"], inp["F0_curve"], inp["N"], inp["s"], seed=3).cpu().numpy()
        dec.parallel_branches = True
        for _ in range(5):
            np.testing.assert_array_equal(_run(dec, inp), ref)
            y = dec(rin["asr"], rin["F0_curve"], rin["N"], rin["s"], noise=rin["noise"], lengths=lens).cpu().numpy()
            np.testing.assert_array_equal(y, ref_r)
        dec.use_cuda_graph = True
        for _ in range(3):
            np.testing.assert_array_equal(dec(inp["asr"], inp["F0_curve"], inp["N"], inp["s"], seed=3).cpu().numpy(), refg)
    finally:
        dec.use_cuda_graph = False
        dec.parallel_branches = old
