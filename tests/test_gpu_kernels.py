"""GPU parity tests of the individual kernels, called through the C ABI, against the CPU oracle."""
import ctypes as C
import math

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from mlx_audio_b200 import _lib as L
from mlx_audio_b200 import synth
from mlx_audio_b200._lib import ConvDesc

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def lib():
    L.require_cuda()
    return L.load()


_KEEP = []


@pytest.fixture(autouse=True)
def _release_device_temporaries():
    yield
    torch.cuda.synchronize()
    _KEEP.clear()


def dev(x, dtype=torch.float32):
    """Host array -> CUDA tensor that stays alive until the end of the test (raw pointers are passed to C)."""
    t = torch.as_tensor(np.asarray(x)).to("cuda", dtype).contiguous()
    _KEEP.append(t)
    return t


def rng(seed):
    return np.random.Generator(np.random.PCG64(seed))


# ------------------------------------------------------------------ weights
def test_fold_weight_norm(lib):
    r = rng(1)
    v = r.standard_normal((37, 7, 29), dtype=np.float32)
    g = r.uniform(0.5, 1.5, (37, 1, 1)).astype(np.float32)
    out = torch.empty(37, 7, 29, device="cuda")
    L.check(lib.kk_fold_weight_norm(L.ptr(dev(v)), L.ptr(dev(g)), L.ptr(out), 37, 7 * 29, L.stream()))
    n = np.sqrt((v.astype(np.float64) ** 2).sum((1, 2), keepdims=True))
    ref = (v / (n + 1e-7)) * g
    np.testing.assert_allclose(out.cpu().numpy(), ref, rtol=2e-6, atol=1e-7)


def test_pack_conv_weight_modes(lib):
    r = rng(2)
    w = r.standard_normal((5, 3, 10), dtype=np.float32)  # (O,K,I)
    out = torch.empty(5, 3 * 64, device="cuda", dtype=torch.bfloat16)
    L.check(lib.kk_pack_conv_weight(L.ptr(dev(w)), L.ptr(out), 1, 0, 5, 3, 10, 1, 64, L.stream()))
    ref = np.zeros((5, 3, 64), np.float32)
    ref[:, :, :10] = w
    np.testing.assert_array_equal(out.float().cpu().numpy().reshape(5, 3, 64),
                                  torch.from_numpy(ref).bfloat16().float().numpy())
    # transposed / polyphase: w (C_in_big=6, K=2*up=4, C_out_small=3), up=2
    wt = r.standard_normal((6, 4, 3), dtype=np.float32)
    out = torch.empty(2 * 3, 2 * 8, device="cuda")
    L.check(lib.kk_pack_conv_weight(L.ptr(dev(wt)), L.ptr(out), 0, 1, 6, 4, 3, 2, 8, L.stream()))
    o = out.cpu().numpy().reshape(2, 3, 2, 8)  # (ph, co, t, ci)
    for ph in range(2):
        for t in range(2):
            np.testing.assert_array_equal(o[ph, :, t, :6], wt[:, ph + 2 * (1 - t), :].T)
    assert (o[..., 6:] == 0).all()


# ------------------------------------------------------------------ elementwise
def test_style_affine(lib):
    r = rng(3)
    B, R, S = 37, 301, 128
    s, W, b = r.standard_normal((B, S), dtype=np.float32), r.standard_normal((R, S), dtype=np.float32), \
        r.standard_normal(R, dtype=np.float32)
    h = torch.empty(B, R, device="cuda")
    L.check(lib.kk_style_affine(L.ptr(dev(s)), L.ptr(dev(W)), L.ptr(dev(b)), L.ptr(h), B, R, S, L.stream()))
    np.testing.assert_allclose(h.cpu().numpy(), s @ W.T + b, rtol=1e-4, atol=1e-4)


@pytest.mark.parametrize("B,Lr,Cc,pitch", [(2, 1000, 70, 72), (3, 77, 1090, 1090), (1, 4801, 128, 128)])
def test_instnorm_stats_and_adain(lib, B, Lr, Cc, pitch):
    r = rng(4)
    x = np.zeros((B, Lr, pitch), np.float32)
    x[:, :, :Cc] = r.standard_normal((B, Lr, Cc), dtype=np.float32) * 2 + 0.5
    xd = dev(x)
    stats = torch.zeros(B, Cc, 2, device="cuda", dtype=torch.float64)
    L.check(lib.kk_instnorm_stats(L.ptr(xd), L.ptr(stats), B, Lr, Cc, pitch, None, 0, 0, L.stream()))
    xs = x[:, :, :Cc].astype(np.float64)
    np.testing.assert_allclose(stats[..., 0].cpu().numpy(), xs.sum(1), rtol=1e-5, atol=1e-3)
    np.testing.assert_allclose(stats[..., 1].cpu().numpy(), (xs ** 2).sum(1), rtol=1e-5)
    h = (r.standard_normal((B, 2 * Cc + 10), dtype=np.float32) * 0.3)
    alpha = r.uniform(0.5, 2.0, Cc).astype(np.float32)
    mean, var = xs.mean(1, keepdims=True), xs.var(1, keepdims=True)
    xn = (xs - mean) / np.sqrt(var + 1e-5)
    y = (1 + h[:, None, 5:5 + Cc]) * xn + h[:, None, 5 + Cc:5 + 2 * Cc]
    hd = dev(h)
    for act, ref in ((L.KK_ACT_SNAKE, y + (1 / alpha) * np.sin(alpha * y) ** 2),
                     (L.KK_ACT_LEAKY, np.where(y > 0, y, 0.2 * y))):
        for is_bf16 in (0, 1):
            out = torch.zeros(B, Lr, pitch, device="cuda", dtype=torch.bfloat16 if is_bf16 else torch.float32)
            L.check(lib.kk_adain_act(L.ptr(xd), pitch, L.ptr(stats), hd.data_ptr() + 5 * 4, h.shape[1], L.ptr(dev(alpha)),
                                     act, 0.2, L.ptr(out), is_bf16, pitch, B, Lr, Cc, 0, None, None, None, 0, 0, L.stream()))
            got = out.float().cpu().numpy()[:, :, :Cc]
            tol = 2e-2 if is_bf16 else 2e-5
            np.testing.assert_allclose(got, ref, rtol=tol, atol=tol)


def test_adain_act_pool_matches_depthwise_conv_transpose(lib):
    r = rng(5)
    B, Lr, Cc = 2, 33, 50
    x = r.standard_normal((B, Lr, Cc), dtype=np.float32)
    pw = r.standard_normal((Cc, 3), dtype=np.float32)
    pb = r.standard_normal(Cc, dtype=np.float32)
    out = torch.empty(B, 2 * Lr, Cc, device="cuda")
    L.check(lib.kk_adain_act(L.ptr(dev(x)), Cc, None, None, 0, None, L.KK_ACT_LEAKY, 0.2, L.ptr(out), 0, Cc, B, Lr, Cc,
                             1, L.ptr(dev(pw)), L.ptr(dev(pb)), None, 0, 0, L.stream()))
    a = torch.from_numpy(np.where(x > 0, x, 0.2 * x)).transpose(1, 2)  # (B,C,L)
    ref = F.conv_transpose1d(a, torch.from_numpy(pw)[:, None, :], torch.from_numpy(pb), stride=2, padding=1, groups=Cc)
    ref = F.pad(ref, (1, 0)).transpose(1, 2).numpy()  # left zero pad (istftnet.py:881)
    assert ref.shape == (B, 2 * Lr, Cc)
    np.testing.assert_allclose(out.cpu().numpy(), ref, rtol=1e-5, atol=1e-5)


def test_ragged_instnorm_and_adain_pool(lib):
    """lens= on the elementwise kernels: statistics over each utterance's own rows, padding rows written as zero, the
    depthwise transposed-conv pool (istftnet.py:880-881) ending at the utterance's last row."""
    r = rng(11)
    B, Lr, Cc = 3, 50, 36
    lens = np.array([50, 7, 31], np.int32)
    x = r.standard_normal((B, Lr, Cc), dtype=np.float32)
    pw = r.standard_normal((Cc, 3), dtype=np.float32)
    pb = r.standard_normal(Cc, dtype=np.float32)
    h = (r.standard_normal((B, 2 * Cc), dtype=np.float32) * 0.3).astype(np.float32)
    ld = torch.from_numpy(lens).cuda()
    xd, hd = dev(x), dev(h)
    stats = torch.zeros(B, Cc, 2, dtype=torch.float64, device="cuda")
    L.check(lib.kk_instnorm_stats(L.ptr(xd), L.ptr(stats), B, Lr, Cc, Cc, L.ptr(ld), 1, 0, L.stream()))
    out = torch.full((B, 2 * Lr, Cc), 5.0, device="cuda")
    L.check(lib.kk_adain_act(L.ptr(xd), Cc, L.ptr(stats), L.ptr(hd), 2 * Cc, None, L.KK_ACT_LEAKY, 0.2, L.ptr(out), 0, Cc,
                             B, Lr, Cc, 1, L.ptr(dev(pw)), L.ptr(dev(pb)), L.ptr(ld), 1, 0, L.stream()))
    flat = torch.full((B, Lr, Cc), 5.0, device="cuda")
    L.check(lib.kk_adain_act(L.ptr(xd), Cc, L.ptr(stats), L.ptr(hd), 2 * Cc, None, L.KK_ACT_LEAKY, 0.2, L.ptr(flat), 0, Cc,
                             B, Lr, Cc, 0, None, None, L.ptr(ld), 1, 0, L.stream()))
    st = stats.cpu().numpy()
    for b, n in enumerate(lens):
        xs = x[b, :n].astype(np.float64)
        np.testing.assert_allclose(st[b, :, 0], xs.sum(0), rtol=1e-12, atol=1e-9)
        np.testing.assert_allclose(st[b, :, 1], (xs * xs).sum(0), rtol=1e-12, atol=1e-9)
        xn = (xs - xs.mean(0)) / np.sqrt(xs.var(0) + 1e-5)
        y = (1 + h[b, None, :Cc]) * xn + h[b, None, Cc:]
        a = np.where(y > 0, y, 0.2 * y).astype(np.float32)
        np.testing.assert_allclose(flat[b, :n].cpu().numpy(), a, rtol=2e-5, atol=2e-5)
        assert not flat[b, n:].any()
        at = torch.from_numpy(a.T[None])
        ref = F.conv_transpose1d(at, torch.from_numpy(pw)[:, None, :], torch.from_numpy(pb), stride=2, padding=1, groups=Cc)
        ref = F.pad(ref, (1, 0))[0].T.numpy()          # (2n, C): left zero pad (istftnet.py:881)
        np.testing.assert_allclose(out[b, :2 * n].cpu().numpy(), ref, rtol=2e-5, atol=2e-5)
        assert not out[b, 2 * n:].any()


def test_front_assemble(lib):
    r = rng(6)
    B, Cc, T = 2, 70, 45
    asr = r.standard_normal((B, Cc, T), dtype=np.float32)
    f0 = r.standard_normal((B, 2 * T), dtype=np.float32)
    n = r.standard_normal((B, 2 * T), dtype=np.float32)
    wf = r.standard_normal(4, dtype=np.float32)
    wn = r.standard_normal(4, dtype=np.float32)
    out = torch.zeros(B, T, Cc + 2, device="cuda")
    L.check(lib.kk_front_assemble(L.ptr(dev(asr)), L.ptr(dev(f0)), L.ptr(dev(n)), L.ptr(dev(wf)), L.ptr(dev(wn)),
                                  L.ptr(out), Cc + 2, 0, Cc, Cc + 1, B, Cc, T, None, L.stream()))
    o = out.cpu().numpy()
    np.testing.assert_array_equal(o[:, :, :Cc], asr.transpose(0, 2, 1))
    for src, w, ch in ((f0, wf, Cc), (n, wn, Cc + 1)):
        ref = F.conv1d(torch.from_numpy(src)[:, None], torch.from_numpy(w[:3])[None, None], torch.from_numpy(w[3:]),
                       stride=2, padding=1)[:, 0].numpy()
        np.testing.assert_allclose(o[:, :, ch], ref, rtol=1e-5, atol=1e-6)


# ------------------------------------------------------------------ convolutions
def _conv_case(lib, tc, B, Lin, Cin, Cout, k, dil, stride=1, transposed_up=0, with_res=True, left_pad=0, seed=0,
               block_n=0, tf32_terms=0, bf16_res=False):
    """tc: bf16 tensor-core kernel; tf32_terms = 3 / 1: fp32 operands through kk_conv1d_tf32_tc (3xTF32 / plain TF32)."""
    tf32 = tf32_terms > 0
    assert not (tc and tf32)
    r = rng(100 + seed)
    bf = torch.bfloat16
    x = r.standard_normal((B, Lin, Cin), dtype=np.float32)
    bias = r.standard_normal(Cout, dtype=np.float32) * 0.1
    if transposed_up:
        u = transposed_up
        K = 2 * u
        wfold = r.standard_normal((Cin, K, Cout), dtype=np.float32) / math.sqrt(Cin * 2)  # (C_in_big, K, C_out_small)
        pad = (K - u) // 2
        Ltrue = Lin * u
        Lout = Ltrue + left_pad
        M, taps, padl, up_off, N, mode = Lin + 1, 2, 1, pad, u * Cout, 1
    else:
        wfold = r.standard_normal((Cout, k, Cin), dtype=np.float32) / math.sqrt(Cin * k)
        padl = (k * dil - dil) // 2
        Ltrue = (Lin + 2 * padl - dil * (k - 1) - 1) // stride + 1
        Lout = Ltrue
        M, taps, up_off, N, mode, u = Ltrue, k, 0, Cout, 0, 1
    xq = torch.from_numpy(x).to(bf).float() if tc else torch.from_numpy(x)
    wq = torch.from_numpy(wfold).to(bf).float() if tc else torch.from_numpy(wfold)
    # reference in float64 on the (rounded) operands
    if transposed_up:
        y = F.conv_transpose1d(xq.double().transpose(1, 2), wq.double().permute(0, 2, 1), torch.from_numpy(bias).double(),
                               stride=u, padding=pad).transpose(1, 2)
        y = F.pad(y, (0, 0, left_pad, 0))
    else:
        y = F.conv1d(xq.double().transpose(1, 2), wq.double().permute(0, 2, 1), torch.from_numpy(bias).double(),
                     stride=stride, padding=padl, dilation=dil).transpose(1, 2)
    assert y.shape == (B, Lout, Cout), (y.shape, (B, Lout, Cout))
    res = r.standard_normal((B, Lout, Cout), dtype=np.float32) if with_res else None
    res2 = r.standard_normal((B, Lout, Cout), dtype=np.float32) if with_res else None
    if bf16_res:      # bf16 residual stream (res_is_bf16): one residual, rounded to bf16, through the TMA epilogue
        res = torch.from_numpy(res).to(bf).float().numpy()
        res2 = np.zeros_like(res)
    scale = 0.7
    ref = y.numpy()
    if with_res:
        ref = ref + res + res2
    ref = ref * scale
    # device side
    Cp = (Cin + 63) // 64 * 64 if tc else ((Cin + 31) // 32 * 32 if tf32 else Cin)
    xp = torch.zeros(B, Lin, Cp, dtype=bf if tc else torch.float32, device="cuda")
    xp[:, :, :Cin] = dev(x).to(xp.dtype)
    wp = torch.empty(N, taps * Cp, dtype=bf if tc else torch.float32, device="cuda")
    if transposed_up:
        L.check(lib.kk_pack_conv_weight(L.ptr(dev(wfold)), L.ptr(wp), int(tc), 1, Cin, 2 * u, Cout, u, Cp, L.stream()))
    else:
        L.check(lib.kk_pack_conv_weight(L.ptr(dev(wfold)), L.ptr(wp), int(tc), 0, Cout, k, Cin, 1, Cp, L.stream()))
    out = torch.full((B, Lout, Cout), float("nan"), device="cuda")
    act = torch.zeros(B, Lout, Cout, dtype=bf if tc else torch.float32, device="cuda")
    stats = torch.zeros(B, Cout, 2, dtype=torch.float64, device="cuda")
    d = ConvDesc()
    d.B, d.L_in, d.C_in, d.x_pitch = B, Lin, Cin, Cp
    d.taps, d.dil, d.pad_left, d.stride, d.Cw = taps, dil, padl, stride, Cp
    d.M, d.N, d.C_out, d.up, d.up_off = M, N, Cout, u, up_off
    d.L_true, d.left_pad, d.L_out = Ltrue, left_pad, Lout
    bd, rd, r2d = dev(bias), (dev(res) if with_res else None), (dev(res2) if with_res else None)
    d.bias = L.ptr(bd)
    if with_res and bf16_res:
        rd = rd.to(bf).contiguous()
        _KEEP.append(rd)
        d.res, d.res_pitch, d.res_shift, d.L_res, d.res_is_bf16 = L.ptr(rd), Cout, 0, Lout, 1
    elif with_res:
        d.res, d.res_pitch, d.res_shift, d.L_res = L.ptr(rd), Cout, 0, Lout
        d.res2, d.res2_pitch = L.ptr(r2d), Cout
    d.scale = scale
    d.out, d.out_pitch, d.stats = L.ptr(out), Cout, L.ptr(stats)
    d.out_act, d.act_pitch, d.act_is_bf16, d.act_slope = L.ptr(act), Cout, int(tc), 0.1
    if tf32:
        d.out_act = None            # the TF32 path serves the fp32 mode: fp32 outputs only
        whi, wlo = torch.empty_like(wp), torch.empty_like(wp)
        L.check(lib.kk_split_tf32(L.ptr(wp), L.ptr(whi), L.ptr(wlo), wp.numel(), L.stream()))
        assert torch.equal(whi.view(torch.int32) & 0x1fff, torch.zeros_like(whi, dtype=torch.int32))
        assert float((whi + wlo - wp).abs().max()) <= 2.0 ** -22 * float(wp.abs().max())
        L.check(lib.kk_conv1d_tf32_tc(L.ptr(xp), L.ptr(whi), L.ptr(wlo), C.byref(d), tf32_terms, L.stream()), "tf32")
    elif tc:
        L.check(lib.kk_conv1d_bf16_tc(L.ptr(xp), L.ptr(wp), C.byref(d), block_n, L.stream()), "tc")
    else:
        L.check(lib.kk_conv1d_f32(L.ptr(xp), L.ptr(wp), C.byref(d), L.stream()), "f32")
    torch.cuda.synchronize()
    got = out.cpu().numpy()
    if left_pad:  # the padded row carries only the residuals
        exp0 = ((res[:, 0] + res2[:, 0]) * scale) if with_res else np.zeros((B, Cout), np.float32)
        np.testing.assert_allclose(got[:, 0], exp0, rtol=1e-5, atol=1e-5)
    tol = 2e-4 if tf32_terms != 1 else 2e-2        # plain TF32: 10-bit operands
    np.testing.assert_allclose(got, ref, rtol=tol, atol=tol)
    if tf32_terms == 3:
        # 3xTF32 is fp32-grade.  What remains is the tensor core's own fp32 accumulation over K / 8 sequential steps
        # (measured 3e-5 of scale at K = 3270, the longest reduction of the decoder; 1e-5 or better at K <= 1408)
        e3 = np.abs(got - ref).max() / max(1.0, np.abs(ref).max())
        print(f"tf32x3 K={taps * Cin}: max error / scale = {e3:.2e}")
        assert e3 < 5e-5, e3
    sref = got.astype(np.float64) if tf32_terms == 1 else ref     # plain TF32: the statistics of what was stored
    np.testing.assert_allclose(stats[..., 0].cpu().numpy(), sref.sum(1), rtol=1e-3, atol=2e-2)
    np.testing.assert_allclose(stats[..., 1].cpu().numpy(), (sref.astype(np.float64) ** 2).sum(1), rtol=1e-3, atol=1e-2)
    a_ref = np.where(ref > 0, ref, 0.1 * ref)
    if tf32:
        return
    np.testing.assert_allclose(act.float().cpu().numpy(), a_ref, rtol=1e-2 if tc else 1e-4, atol=1e-2 if tc else 1e-4)


F32_CASES = [
    dict(B=2, Lin=200, Cin=22, Cout=48, k=12, dil=1, stride=6),     # noise_convs.0 geometry
    dict(B=1, Lin=333, Cin=70, Cout=130, k=7, dil=3),
    dict(B=2, Lin=100, Cin=64, Cout=64, k=1, dil=1, with_res=False),
    dict(B=2, Lin=50, Cin=40, Cout=32, k=0, dil=1, transposed_up=10),
    dict(B=1, Lin=61, Cin=32, Cout=32, k=0, dil=1, transposed_up=6, left_pad=1),
]


@pytest.mark.parametrize("case", F32_CASES)
def test_conv1d_f32(lib, case):
    _conv_case(lib, False, **case)


TC_CASES = [
    dict(B=2, Lin=300, Cin=128, Cout=128, k=3, dil=1),
    dict(B=1, Lin=517, Cin=128, Cout=128, k=11, dil=5),
    dict(B=2, Lin=260, Cin=256, Cout=256, k=7, dil=3),
    dict(B=2, Lin=130, Cin=256, Cout=256, k=3, dil=1, block_n=128),
    dict(B=2, Lin=70, Cin=1090, Cout=512, k=3, dil=1, with_res=False),   # K not a multiple of 64
    dict(B=1, Lin=100, Cin=514, Cout=1024, k=1, dil=1, with_res=False),
    dict(B=2, Lin=200, Cin=128, Cout=22, k=7, dil=1, with_res=False),    # conv_post geometry (N=22 -> BN=32)
    dict(B=2, Lin=40, Cin=512, Cout=256, k=0, dil=1, transposed_up=10),  # ups.0 geometry
    dict(B=1, Lin=150, Cin=256, Cout=128, k=0, dil=1, transposed_up=6, left_pad=1),  # ups.1 geometry
    dict(B=3, Lin=64, Cin=64, Cout=64, k=3, dil=1, block_n=64),
    # bf16 residual through the TMA epilogue: plain conv (edge chunks: Lin not a multiple of 32) and the ups.1 geometry
    dict(B=2, Lin=300, Cin=128, Cout=128, k=3, dil=1, bf16_res=True),
    dict(B=2, Lin=150, Cin=256, Cout=128, k=0, dil=1, transposed_up=6, left_pad=1, bf16_res=True),
]


@pytest.mark.parametrize("case", TC_CASES)
def test_conv1d_bf16_tc(lib, case):
    if not lib.kk_device_is_sm100():
        pytest.skip("needs sm_100")
    _conv_case(lib, True, **case)


TF32_CASES = [
    dict(B=2, Lin=300, Cin=128, Cout=128, k=3, dil=1),
    dict(B=1, Lin=517, Cin=128, Cout=128, k=11, dil=5),
    dict(B=2, Lin=260, Cin=256, Cout=256, k=7, dil=3),
    dict(B=2, Lin=70, Cin=1090, Cout=512, k=3, dil=1, with_res=False),   # K not a multiple of 32
    dict(B=2, Lin=200, Cin=128, Cout=22, k=7, dil=1, with_res=False),    # conv_post geometry
    dict(B=2, Lin=40, Cin=512, Cout=256, k=0, dil=1, transposed_up=10),  # ups.0 geometry
    dict(B=1, Lin=150, Cin=256, Cout=128, k=0, dil=1, transposed_up=6, left_pad=1),  # ups.1 geometry
    dict(B=3, Lin=64, Cin=64, Cout=64, k=3, dil=1),
    dict(B=2, Lin=100, Cin=514, Cout=64, k=1, dil=1, with_res=False),    # asr_res geometry (512 + 2 -> 64)
]


# N a multiple of 32 but not of the 128-column tile, rows not a multiple of 32: the last column tile has chunks past N,
# and the edge row blocks take the per-row store path.  Round 2 found those chunks stored anyway there - a polyphase
# chunk with "phase = up" landed on phase 0 of the next row block (bias only) and raced with its owner (SNAC's stride-7
# upsampler failed once in ~10 full-suite runs); with up = 1 it wrote past the row.
PAST_N_CASES = [
    dict(B=2, Lin=37, Cin=96, Cout=32, k=0, dil=1, transposed_up=6, with_res=False),     # N = 192: chunks 6, 7 of tile 1 are past N
    dict(B=2, Lin=45, Cin=64, Cout=96, k=3, dil=1, with_res=False),                       # N = 96 in a 128-wide tile
    dict(B=2, Lin=45, Cin=64, Cout=96, k=3, dil=1),                                       # ... with residuals
]


@pytest.mark.parametrize("kind", ["bf16", "tf32x3"])
@pytest.mark.parametrize("case", PAST_N_CASES)
def test_conv_tc_chunks_past_n_are_not_stored(lib, case, kind):
    if not lib.kk_device_is_sm100():
        pytest.skip("needs sm_100")
    for rep in range(12):            # a race: repeat with fresh buffers
        if kind == "bf16":
            _conv_case(lib, True, **case)
        else:
            _conv_case(lib, False, tf32_terms=3, **case)


@pytest.mark.parametrize("terms", [3, 1])
@pytest.mark.parametrize("case", TF32_CASES)
def test_conv1d_tf32_tc(lib, case, terms):
    """fp32 operands on tcgen05 kind::tf32: 3xTF32 (hi/lo split, three MMAs) must be fp32-grade, plain TF32 is reported
    separately with its own (loose) tolerance - SURVEY 7.2 expects it to miss the 1e-3 decoder gate."""
    if not lib.kk_device_is_sm100():
        pytest.skip("needs sm_100")
    _conv_case(lib, False, tf32_terms=terms, **case)


# ------------------------------------------------------------------ fused AdaIN + Snake + conv (tensor cores)
def _fused_case(lib, B, Lin, Cch, k, dil, x_bf16=False, act="snake", with_res=True, with_norm=True, seed=0):
    from mlx_audio_b200._lib import AdainIn
    r = rng(300 + seed)
    bf = torch.bfloat16
    x = (r.standard_normal((B, Lin, Cch), dtype=np.float32) * 1.7 + 0.4).astype(np.float32)
    if x_bf16:
        x = torch.from_numpy(x).to(bf).float().numpy()
    S_ = 2 * Cch
    h = (r.standard_normal((B, 2 * S_), dtype=np.float32) * 0.3).astype(np.float32)   # rows wider than 2C: h_stride
    alpha = r.uniform(0.5, 2.0, Cch).astype(np.float32)
    wfold = r.standard_normal((Cch, k, Cch), dtype=np.float32) / math.sqrt(Cch * k)
    bias = r.standard_normal(Cch, dtype=np.float32) * 0.1
    padl = (k * dil - dil) // 2
    # reference: AdaIN (population variance, eps 1e-5) -> activation -> bf16 operand rounding -> conv in float64
    xd = torch.from_numpy(x).double()
    if with_norm:
        mean, var = xd.mean(1, keepdim=True), xd.var(1, unbiased=False, keepdim=True)
        g = torch.from_numpy(h[:, None, :Cch]).double()
        bt = torch.from_numpy(h[:, None, Cch:2 * Cch]).double()
        y = (1 + g) * (xd - mean) / torch.sqrt(var + 1e-5) + bt
    else:
        y = xd
    if act == "snake":
        al = torch.from_numpy(alpha).double()
        y = y + (1 / al) * torch.sin(al * y) ** 2
    elif act == "leaky":
        y = torch.where(y > 0, y, 0.2 * y)
    a_ref = y.float()
    aq = a_ref.to(bf).double()
    wq = torch.from_numpy(wfold).to(bf).double()
    out_ref = F.conv1d(aq.transpose(1, 2), wq.permute(0, 2, 1), torch.from_numpy(bias).double(), padding=padl,
                       dilation=dil).transpose(1, 2).numpy()
    res = r.standard_normal((B, Lin, Cch), dtype=np.float32) if with_res else None
    scale = 0.5
    ref = (out_ref + (res if with_res else 0.0)) * scale
    # device
    xdev = dev(x).to(bf).contiguous() if x_bf16 else dev(x)
    stats = torch.stack([xd.sum(1), (xd * xd).sum(1)], -1).to("cuda").contiguous()      # (B, C, 2) float64
    wp = torch.empty(Cch, k * Cch, dtype=bf, device="cuda")
    L.check(lib.kk_pack_conv_weight(L.ptr(dev(wfold)), L.ptr(wp), 1, 0, Cch, k, Cch, 1, Cch, L.stream()))
    out = torch.full((B, Lin, Cch), float("nan"), device="cuda")
    oact = torch.zeros(B, Lin, Cch, dtype=bf, device="cuda")
    ostats = torch.zeros(B, Cch, 2, dtype=torch.float64, device="cuda")
    hd, ad, bd = dev(h), dev(alpha), dev(bias)
    rd = dev(res) if with_res else None
    a = AdainIn()
    a.x, a.x_is_bf16, a.x_pitch = L.ptr(xdev), int(x_bf16), Cch
    a.stats = L.ptr(stats) if with_norm else None
    a.h, a.h_stride = L.ptr(hd), 2 * S_
    a.alpha = L.ptr(ad)
    a.act = {"snake": L.KK_ACT_SNAKE_FAST, "leaky": L.KK_ACT_LEAKY, "none": L.KK_ACT_NONE}[act]
    a.slope = 0.2
    d = ConvDesc()
    d.B, d.L_in, d.C_in, d.x_pitch = B, Lin, Cch, Cch
    d.taps, d.dil, d.pad_left, d.stride, d.Cw = k, dil, padl, 1, Cch
    d.M, d.N, d.C_out, d.up, d.up_off = Lin, Cch, Cch, 1, 0
    d.L_true, d.left_pad, d.L_out = Lin, 0, Lin
    d.bias = L.ptr(bd)
    if with_res:
        d.res, d.res_pitch, d.res_shift, d.L_res = L.ptr(rd), Cch, 0, Lin
    d.scale = scale
    d.out, d.out_pitch, d.stats = L.ptr(out), Cch, L.ptr(ostats)
    d.out_act, d.act_pitch, d.act_is_bf16, d.act_slope = L.ptr(oact), Cch, 1, 1.0
    # out_add: the result tile is also accumulated (TMA reduce-add) into a wider, pre-filled tensor
    add0 = r.standard_normal((B, Lin, Cch + 4), dtype=np.float32)
    addd = dev(add0)
    d.out_add, d.add_pitch = L.ptr(addd), Cch + 4
    L.check(lib.kk_conv1d_adain_bf16_tc(C.byref(a), L.ptr(wp), C.byref(d), L.stream()), "fused")
    torch.cuda.synchronize()
    got = out.cpu().numpy()
    addg = addd.cpu().numpy()
    np.testing.assert_allclose(addg[..., :Cch], add0[..., :Cch] + got, rtol=0, atol=1e-5)
    np.testing.assert_array_equal(addg[..., Cch:], add0[..., Cch:])
    # ... and alone (no `out`): same values
    addd2 = dev(add0)
    d.out, d.out_add = None, L.ptr(addd2)
    ostats.zero_()
    L.check(lib.kk_conv1d_adain_bf16_tc(C.byref(a), L.ptr(wp), C.byref(d), L.stream()), "fused/out_add only")
    torch.cuda.synchronize()
    np.testing.assert_array_equal(addd2.cpu().numpy(), addg)
    if with_res:
        # bf16 residual (res_is_bf16): (i) with the fp32 output - the result tile overlays the residual tile in
        # shared memory; (ii) bf16 output only - the configuration of the running stream inside AdaINResBlock1
        r16 = rd.to(bf).contiguous()
        exp16 = got + (r16.float().cpu().numpy() - res) * scale
        out2 = torch.full((B, Lin, Cch), float("nan"), device="cuda")
        d.res, d.res_is_bf16 = L.ptr(r16), 1
        d.out, d.out_add = L.ptr(out2), None
        ostats.zero_()
        L.check(lib.kk_conv1d_adain_bf16_tc(C.byref(a), L.ptr(wp), C.byref(d), L.stream()), "fused/res16")
        torch.cuda.synchronize()
        np.testing.assert_allclose(out2.cpu().numpy(), exp16, rtol=0, atol=2e-6)
        oact2 = torch.zeros(B, Lin, Cch, dtype=bf, device="cuda")
        st2 = torch.zeros_like(ostats)
        d.out, d.out_act, d.stats = None, L.ptr(oact2), L.ptr(st2)
        L.check(lib.kk_conv1d_adain_bf16_tc(C.byref(a), L.ptr(wp), C.byref(d), L.stream()), "fused/res16 bf16 out only")
        torch.cuda.synchronize()
        assert torch.equal(oact2, out2.to(bf))
        np.testing.assert_allclose(st2.cpu().numpy(), ostats.cpu().numpy(), rtol=1e-12, atol=1e-9)
        ostats.copy_(st2)
        got_stats = out2.cpu().numpy()
    else:
        got_stats = got
    # tolerance: the kernel's fp32 affine / MUFU sine may round an operand to the neighbouring bf16 value
    err = np.abs(got - ref)
    assert np.isfinite(got).all()
    assert err.max() < 3e-2 and err.mean() < 2e-3, (err.max(), err.mean())
    np.testing.assert_allclose(ostats[..., 0].cpu().numpy(), got_stats.astype(np.float64).sum(1), rtol=1e-4, atol=1e-2)
    np.testing.assert_allclose(ostats[..., 1].cpu().numpy(), (got_stats.astype(np.float64) ** 2).sum(1), rtol=1e-4, atol=1e-2)
    np.testing.assert_allclose(oact.float().cpu().numpy(), got, rtol=1e-2, atol=1e-2)


FUSED_CASES = [
    # fp32 input (conv1 of an AdaINResBlock1 iteration: no residual) / bf16 input (conv2: residual + fp32 output)
    dict(B=2, Lin=700, Cch=128, k=11, dil=5, with_res=False),
    dict(B=1, Lin=1000, Cch=128, k=3, dil=1, x_bf16=True),
    dict(B=3, Lin=257, Cch=128, k=7, dil=3, with_res=False),
    dict(B=2, Lin=300, Cch=256, k=11, dil=5, x_bf16=True),
    dict(B=2, Lin=129, Cch=256, k=7, dil=1, x_bf16=True, with_res=False),
    dict(B=1, Lin=40, Cch=256, k=3, dil=3, with_res=False),
    dict(B=2, Lin=200, Cch=128, k=3, dil=1, act="leaky", with_res=False),
    dict(B=1, Lin=200, Cch=128, k=1, dil=1, act="none", with_norm=False, x_bf16=True),
    dict(B=150, Lin=300, Cch=128, k=7, dil=1, x_bf16=True),   # > 148 tiles: persistent loop, utterance changes per CTA
    dict(B=3, Lin=700, Cch=256, k=7, dil=5, with_res=False),
]


def test_conv1d_adain_rejects_fp32_input_with_residual(lib):
    if not lib.kk_device_is_sm100():
        pytest.skip("needs sm_100")
    with pytest.raises(L.KokoroB200Error, match="unsupported"):
        _fused_case(lib, B=1, Lin=64, Cch=128, k=3, dil=1, x_bf16=False, with_res=True)


@pytest.mark.parametrize("case", FUSED_CASES)
def test_conv1d_adain_bf16_tc(lib, case):
    if not lib.kk_device_is_sm100():
        pytest.skip("needs sm_100")
    _fused_case(lib, **case)


@pytest.mark.parametrize("B,Lin", [(2, 700), (1, 257)])
def test_conv1d_adain_narrow_head(lib, B, Lin):
    """(C_in, N) = (128, 22): conv_post through the row-shifted-descriptor kernel, no normalisation / activation
    (the transform warps pass the TMA-written operand through), `out` rows 24 floats wide, 22 written."""
    if not lib.kk_device_is_sm100():
        pytest.skip("needs sm_100")
    from mlx_audio_b200._lib import AdainIn
    r = rng(77 + Lin)
    bf = torch.bfloat16
    Cch, N, k = 128, 22, 7
    x = torch.from_numpy(r.standard_normal((B, Lin, Cch), dtype=np.float32)).to(bf)
    wfold = r.standard_normal((N, k, Cch), dtype=np.float32) / math.sqrt(Cch * k)
    bias = r.standard_normal(N, dtype=np.float32) * 0.1
    wq = torch.from_numpy(wfold).to(bf).double()
    ref = F.conv1d(x.double().transpose(1, 2), wq.permute(0, 2, 1), torch.from_numpy(bias).double(),
                   padding=3).transpose(1, 2).numpy()
    xdev = x.cuda().contiguous()
    wp = torch.empty(N, k * Cch, dtype=bf, device="cuda")
    L.check(lib.kk_pack_conv_weight(L.ptr(dev(wfold)), L.ptr(wp), 1, 0, N, k, Cch, 1, Cch, L.stream()))
    out = torch.full((B, Lin, 24), 7.0, device="cuda")
    bd = dev(bias)
    a = AdainIn()
    a.x, a.x_is_bf16, a.x_pitch = L.ptr(xdev), 1, Cch
    a.act = L.KK_ACT_NONE
    d = ConvDesc()
    d.B, d.L_in, d.C_in, d.x_pitch = B, Lin, Cch, Cch
    d.taps, d.dil, d.pad_left, d.stride, d.Cw = k, 1, 3, 1, Cch
    d.M, d.N, d.C_out, d.up, d.up_off = Lin, N, N, 1, 0
    d.L_true, d.left_pad, d.L_out = Lin, 0, Lin
    d.bias = L.ptr(bd)
    d.scale = 1.0
    d.out, d.out_pitch = L.ptr(out), 24
    L.check(lib.kk_conv1d_adain_bf16_tc(C.byref(a), L.ptr(wp), C.byref(d), L.stream()), "fused narrow")
    torch.cuda.synchronize()
    got = out.cpu().numpy()
    np.testing.assert_allclose(got[..., :N], ref, rtol=0, atol=2e-4)
    # pad columns: untouched, or zero where TMA's 16-byte store granularity reaches past channel 21
    assert np.isin(got[..., N:], (0.0, 7.0)).all()
