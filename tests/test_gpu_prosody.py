"""GPU parity of ProsodyPredictor.F0Ntrain against the CPU oracle and the reference golden."""
import os

import numpy as np
import pytest
import torch

from mlx_audio_b200 import _lib as L
from mlx_audio_b200 import synth
from mlx_audio_b200.tts.models.kokoro.modules import build_prosody_predictor
from oracle.prosody import F0NOracle

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def weights():
    return synth.synth_prosody_weights(seed=2100)


@pytest.fixture(scope="module")
def oracle(weights):
    return F0NOracle(weights)


def _np(t):
    torch.cuda.synchronize()
    return t.cpu().numpy().copy()


def test_f0ntrain_fp32_matches_reference_golden(golden_dir, weights):
    g = np.load(os.path.join(golden_dir, "prosody_b2_l12.npz"))
    inp = synth.synth_prosody_inputs(int(g["batch"]), int(g["frames"]), seed=int(g["input_seed"]))
    pp = build_prosody_predictor(weights, dtype="fp32")
    F0, N = pp.F0Ntrain(inp["x"], inp["s"])
    F0, N = _np(F0), _np(N)
    assert F0.shape == g["F0"].shape and N.shape == g["N"].shape
    # F0 is in Hz (~100-300): 1e-3 relative to its scale; N is O(1): 1e-3 absolute (the decoder's fp32 gate)
    assert np.abs(F0 - g["F0"]).max() <= 1e-3 * np.abs(g["F0"]).max(), np.abs(F0 - g["F0"]).max()
    assert np.abs(N - g["N"]).max() <= 1e-3, np.abs(N - g["N"]).max()


@pytest.mark.parametrize("B,Lq", [(1, 200), (3, 37)])
def test_f0ntrain_fp32_vs_oracle(weights, oracle, B, Lq):
    inp = synth.synth_prosody_inputs(B, Lq, seed=1200 + Lq)
    hidden = oracle.hidden_of(inp["x"])
    F0r, Nr = oracle(s=inp["s"], hidden=hidden)
    pp = build_prosody_predictor(weights, dtype="fp32")
    F0, N = pp.F0Ntrain(s=inp["s"], hidden=hidden)           # conv path alone
    assert np.abs(_np(F0) - F0r).max() <= 1e-3 * np.abs(F0r).max()
    assert np.abs(_np(N) - Nr).max() <= 1e-3
    F0b, Nb = pp.F0Ntrain(inp["x"], inp["s"])                # with the native LSTM in front
    assert np.abs(_np(F0b) - F0r).max() <= 1e-3 * np.abs(F0r).max()
    assert np.abs(_np(Nb) - Nr).max() <= 1e-3


def _torch_lstm(weights, device="cuda"):
    """The same layer as a torch module (cuDNN), TF32 off: the independent check of the recurrent kernel."""
    w = {k[len("predictor."):] if k.startswith("predictor.") else k: torch.as_tensor(np.asarray(v)) for k, v in weights.items()}
    H = w["shared.Wh_forward"].shape[1]
    m = torch.nn.LSTM(w["shared.Wx_forward"].shape[1], H, batch_first=True, bidirectional=True).to(device)
    with torch.no_grad():
        for sfx, d in (("", "forward"), ("_reverse", "backward")):
            getattr(m, f"weight_ih_l0{sfx}").copy_(w[f"shared.Wx_{d}"])
            getattr(m, f"weight_hh_l0{sfx}").copy_(w[f"shared.Wh_{d}"])
            getattr(m, f"bias_ih_l0{sfx}").copy_(w[f"shared.bias_ih_{d}"])
            getattr(m, f"bias_hh_l0{sfx}").copy_(w[f"shared.bias_hh_{d}"])
    return m.eval()


@pytest.mark.timeout(120)
@pytest.mark.parametrize("B,Lq", [(1, 1), (3, 37), (9, 200), (64, 320)])
def test_shared_lstm_native_vs_oracle_and_cudnn(weights, oracle, B, Lq):
    """``kk_lstm_bidir_f32`` + the 3xTF32 input GEMM against the numpy restatement of modules.py:93-286 and against
    cuDNN (TF32 off).  h is in (-1, 1): 2e-5 absolute."""
    inp = synth.synth_prosody_inputs(B, Lq, seed=1500 + Lq)
    pp = build_prosody_predictor(weights, dtype="fp32")
    x = torch.as_tensor(inp["x"]).cuda().transpose(1, 2).contiguous()
    got = _np(pp.shared(x))
    assert got.shape == (B, Lq, 512)
    if B * Lq <= 2000:
        assert np.abs(got - oracle.hidden_of(inp["x"])).max() <= 2e-5
    with torch.no_grad(), torch.backends.cudnn.flags(enabled=True, allow_tf32=False):
        ref = _np(_torch_lstm(weights)(x)[0])
    assert np.abs(got - ref).max() <= 2e-5, np.abs(got - ref).max()
    if B == 64:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            pp.shared(x)
        e1.record()
        e1.synchronize()
        print(f"native bi-LSTM B={B} L={Lq}: {e0.elapsed_time(e1) / 5:.3f} ms per call")


@pytest.mark.timeout(120)
@pytest.mark.parametrize("H,D,B,Lq", [(64, 24, 3, 19), (128, 100, 9, 33), (192, 64, 1, 7), (256, 640, 5, 12)])
def test_lstm_kernel_other_sizes_vs_cudnn(H, D, B, Lq):
    """kk_lstm_bidir_f32 through the C ABI at every hidden size it accepts (clusters of 8 CTAs, H / 8 units each), with the
    input projection done in fp64 on the host side of the test; H = 96 must be rejected."""
    lib = L.load()
    g = torch.Generator().manual_seed(H + Lq)
    m = torch.nn.LSTM(D, H, batch_first=True, bidirectional=True).cuda().eval()
    x = torch.randn(B, Lq, D, generator=g).cuda()
    with torch.no_grad(), torch.backends.cudnn.flags(enabled=True, allow_tf32=False):
        ref = m(x)[0]
    xw = torch.cat([(x.double() @ getattr(m, f"weight_ih_l0{s}").double().t()
                     + getattr(m, f"bias_ih_l0{s}").double() + getattr(m, f"bias_hh_l0{s}").double()) for s in ("", "_reverse")],
                   dim=2).float().contiguous()                                     # (B, L, 2 * 4H)
    whT = torch.stack([m.weight_hh_l0.detach().t(), m.weight_hh_l0_reverse.detach().t()]).contiguous()
    out = torch.empty(B, Lq, 2 * H, device="cuda")
    L.check(lib.kk_lstm_bidir_f32(L.ptr(xw), L.ptr(whT), L.ptr(out), B, Lq, H, None, L.stream()))
    torch.cuda.synchronize()
    assert float((out - ref).abs().max()) <= 2e-5
    assert lib.kk_lstm_bidir_f32(L.ptr(xw), L.ptr(whT), L.ptr(out), B, Lq, 96, None, L.stream()) == L.KK_ERR_UNSUPPORTED


@pytest.mark.timeout(120)
@pytest.mark.parametrize("B", [64, 100, 150])
def test_lstm_result_does_not_depend_on_the_batch_partition(B):
    """The kernel picks the number of sequences per cluster from B (8, 10 or 12: all clusters co-resident when they can
    be - 8 per cluster would make B = 64 two waves) and runs more than one wave beyond that.  A sequence's result must not
    depend on it: the big batch equals its slices of 8 bit for bit (ragged lengths included)."""
    H, Lq = 256, 24
    lib = L.load()
    g = torch.Generator(device="cuda").manual_seed(B)
    xw = torch.randn(B, Lq, 2 * 4 * H, device="cuda", generator=g) * 0.7
    whT = torch.randn(2, H, 4 * H, device="cuda", generator=g) / 16
    lens = torch.randint(1, Lq + 1, (B,), device="cuda", generator=g, dtype=torch.int32)
    lens[0] = Lq
    out = torch.empty(B, Lq, 2 * H, device="cuda")
    L.check(lib.kk_lstm_bidir_f32(L.ptr(xw), L.ptr(whT), L.ptr(out), B, Lq, H, L.ptr(lens), L.stream()))
    for b0 in range(0, B, 8):
        n = min(8, B - b0)
        xs, ls = xw[b0:b0 + n].contiguous(), lens[b0:b0 + n].contiguous()
        o = torch.empty(n, Lq, 2 * H, device="cuda")
        L.check(lib.kk_lstm_bidir_f32(L.ptr(xs), L.ptr(whT), L.ptr(o), n, Lq, H, L.ptr(ls), L.stream()))
        assert torch.equal(o, out[b0:b0 + n]), b0
    assert torch.isfinite(out).all()


@pytest.mark.timeout(120)
def test_shared_lstm_ragged_lengths_match_solo(weights):
    """lengths=: every sequence of a padded batch equals running it alone (the backward direction starts at its own last
    step), zeros after its end; the padding holds garbage."""
    B, Lq = 11, 50
    lens = [50, 1, 17, 50, 33, 8, 49, 2, 25, 40, 5]
    inp = synth.synth_prosody_inputs(B, Lq, seed=1600)
    pp = build_prosody_predictor(weights, dtype="fp32")
    x = torch.as_tensor(inp["x"]).cuda().transpose(1, 2).contiguous()
    for b, l in enumerate(lens):
        x[b, l:] = 1e4
    got = pp.shared(x, lengths=lens)
    for b, l in enumerate(lens):
        solo = pp.shared(x[b:b + 1, :l].contiguous())
        assert torch.equal(got[b, :l], solo[0]), b
        assert not got[b, l:].abs().any(), b


@pytest.mark.timeout(120)
@pytest.mark.parametrize("dtype", ["fp32", "bf16"])
def test_f0ntrain_ragged_lengths_match_solo(weights, dtype):
    """F0Ntrain(lengths=): sequences of a padded batch come out as when run alone, zeros after 2 * length."""
    if dtype == "bf16" and not L.load().kk_device_is_sm100():
        pytest.skip("needs sm_100")
    B, Lq = 7, 48
    lens = [48, 5, 31, 48, 16, 40, 1]
    inp = synth.synth_prosody_inputs(B, Lq, seed=1700)
    x = torch.as_tensor(inp["x"]).cuda()
    for b, l in enumerate(lens):
        x[b, :, l:] = 100.0
    pp = build_prosody_predictor(weights, dtype=dtype)
    F0, N = pp.F0Ntrain(x, inp["s"], lengths=lens)
    for b, l in enumerate(lens):
        f0s, ns = pp.F0Ntrain(x[b:b + 1, :, :l].contiguous(), inp["s"][b:b + 1])
        for got, solo in ((F0, f0s), (N, ns)):
            scale = max(1.0, float(solo.abs().max()))
            assert float((got[b, :2 * l] - solo[0]).abs().max()) <= 2e-5 * scale, (b, l)
            assert not got[b, 2 * l:].abs().any(), (b, l)
    with pytest.raises(ValueError):
        pp.F0Ntrain(x, inp["s"], lengths=[49] * B)


def test_f0ntrain_bf16_vs_oracle(weights, oracle):
    if not L.load().kk_device_is_sm100():
        pytest.skip("needs sm_100")
    inp = synth.synth_prosody_inputs(4, 160, seed=1300)
    hidden = oracle.hidden_of(inp["x"])
    F0r, Nr = oracle(s=inp["s"], hidden=hidden)
    pp = build_prosody_predictor(weights, dtype="bf16")
    F0, N = pp.F0Ntrain(s=inp["s"], hidden=hidden)
    F0, N = _np(F0), _np(N)
    # bf16 operands / fp32 accumulate: stated tolerance 3e-2 relative L2 (same as the decoder's bf16 gate)
    rel_f0 = np.linalg.norm(F0 - F0r) / np.linalg.norm(F0r - F0r.mean())
    rel_n = np.linalg.norm(N - Nr) / np.linalg.norm(Nr)
    print("bf16 F0Ntrain rel-l2:", rel_f0, rel_n)
    assert rel_f0 <= 3e-2 and rel_n <= 3e-2


def test_f0ntrain_feeds_decoder(weights):
    """F0Ntrain output has the decoder's F0_curve / N contract: (B, 2L) for asr (B, 512, L)."""
    from mlx_audio_b200.tts.models.kokoro import build_decoder
    B, Lq = 2, 16
    inp = synth.synth_prosody_inputs(B, Lq, seed=1400)
    pp = build_prosody_predictor(weights, dtype="fp32")
    F0, N = pp.F0Ntrain(inp["x"], inp["s"])
    assert tuple(F0.shape) == tuple(N.shape) == (B, 2 * Lq)
    dec = build_decoder(synth.synth_decoder_weights(seed=2000), dtype="fp32")
    di = synth.synth_decoder_inputs(B, Lq, seed=1000)
    y = dec(di["asr"], F0.clone(), N.clone(), di["s"], noise=di["noise"])
    torch.cuda.synchronize()
    assert tuple(y.shape) == (B, 1, 600 * Lq) and bool(torch.isfinite(y).all())


def test_f0ntrain_errors(weights):
    pp = build_prosody_predictor(weights, dtype="fp32")
    with pytest.raises(ValueError):
        pp.F0Ntrain(np.zeros((1, 100, 8), np.float32), np.zeros((1, 128), np.float32))
    with pytest.raises(ValueError):
        pp.F0Ntrain(np.zeros((1, 640, 8), np.float32), np.zeros((2, 128), np.float32))
    with pytest.raises(ValueError):
        build_prosody_predictor({"predictor.bogus": np.zeros(1, np.float32)})


def test_model_runs_f0ntrain_when_frontend_returns_en(weights):
    """Model.__call__ (kokoro.py:157-165): a front end that stops at ``en`` -> F0Ntrain(en, ref_s[:, 128:]) -> decoder."""
    from mlx_audio_b200.tts.models.kokoro import Model, ModelConfig, build_decoder
    T = 10
    pin = synth.synth_prosody_inputs(1, T, seed=1500)
    din = synth.synth_decoder_inputs(1, T, seed=1501)
    dw = synth.synth_decoder_weights(seed=2000)

    def frontend(phonemes, ref_s, speed):
        return {"asr": din["asr"], "en": pin["x"], "pred_dur": np.ones(len(phonemes) + 2, np.int32)}

    m = Model(ModelConfig(), frontend=frontend, dtype="fp32")
    ck = dict(dw)
    ck.update(weights)
    ck["bert.encoder.dummy"] = np.zeros(2, np.float32)
    m.load_weights(ck)          # MLX-layout weights: no sanitize; keys outside decoder.* / predictor.* are ignored
    ref_s = np.concatenate([din["s"], pin["s"]], axis=1)                    # [decoder style | prosody style]
    m.decoder.seed = 3
    out = m("abcdefgh", ref_s, return_output=True)
    torch.cuda.synchronize()
    F0, N = build_prosody_predictor(weights, dtype="fp32").F0Ntrain(pin["x"], pin["s"])
    y = build_decoder(dw, dtype="fp32")(din["asr"], F0.clone(), N.clone(), din["s"], seed=3)
    torch.testing.assert_close(out.audio, y[0], rtol=0, atol=1e-6)
    assert out.pred_dur.shape == (10,)


@pytest.mark.timeout(120)
@pytest.mark.parametrize("dtype", ["fp32", "bf16"])
def test_f0ntrain_parallel_heads_equal_one_stream(weights, dtype):
    """The N head runs as a parallel branch on a second stream (forked after the style projection, joined after F0's
    last launch): the same launches, so the same bits as the one-stream schedule."""
    B, Lq = 5, 37
    inp = synth.synth_prosody_inputs(B, Lq, seed=1700)
    pp = build_prosody_predictor(weights, dtype=dtype)
    x, s = torch.as_tensor(inp["x"]).cuda(), torch.as_tensor(inp["s"]).cuda()
    assert pp.parallel_heads
    f0_p, n_p = pp.F0Ntrain(x, s)
    f0_p2, n_p2 = pp.F0Ntrain(x, s)
    pp.parallel_heads = False
    f0_s, n_s = pp.F0Ntrain(x, s)
    assert torch.equal(f0_p, f0_s) and torch.equal(n_p, n_s)
    assert torch.equal(f0_p, f0_p2) and torch.equal(n_p, n_p2)


@pytest.mark.timeout(120)
def test_lstm_repeated_launches_are_bit_identical():
    """Stand-in for racecheck on the cluster protocol (st.async into eight CTAs' shared memory, per-CTA mbarriers, two h
    buffers): 40 launches of the same ragged batch, with other work and fresh allocations in between, give the same bits."""
    H, B, Lq = 256, 64, 48
    lib = L.load()
    g = torch.Generator(device="cuda").manual_seed(77)
    xw = torch.randn(B, Lq, 2 * 4 * H, device="cuda", generator=g) * 0.7
    whT = torch.randn(2, H, 4 * H, device="cuda", generator=g) / 16
    lens = torch.randint(1, Lq + 1, (B,), device="cuda", generator=g, dtype=torch.int32)
    first = None
    for i in range(40):
        out = torch.full((B, Lq, 2 * H), float("nan"), device="cuda")
        junk = torch.randn(1 << (10 + i % 12), device="cuda")           # shifts the allocator and keeps other kernels in flight
        junk.mul_(2.0)
        L.check(lib.kk_lstm_bidir_f32(L.ptr(xw), L.ptr(whT), L.ptr(out), B, Lq, H, L.ptr(lens), L.stream()))
        if first is None:
            first = out.clone()
            assert torch.isfinite(first).all()
        else:
            assert torch.equal(out, first), i
