"""SNAC generator on the decoder's convolution / Snake kernels, against goldens minted by executing
the reference's own codec/models/snac/layers.py (tests/golden/make_golden_snac.py), plus kernel-level checks of the
pieces that are new here: WNConvTranspose1d on the polyphase implicit GEMM (even and odd strides, the reference's
effective output_padding = 1) and the depthwise WNConv1d."""
import os

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from mlx_audio_b200 import _lib as L
from mlx_audio_b200 import synth
from mlx_audio_b200.codec.models.snac import layers as S
from parity_gates import record

pytestmark = pytest.mark.gpu

CASES = {"dw": (24, 256, [4, 3, 2], True, True), "dense": (16, 64, [2, 3], False, False)}


@pytest.mark.parametrize("name", sorted(CASES))
def test_snac_decoder_matches_reference_golden(golden_dir, name):
    g = np.load(os.path.join(golden_dir, "snac_decoder.npz"))
    cin, ch, rates, noise, dw = CASES[name]
    dec = S.Decoder(cin, ch, rates, noise, depthwise=dw, attn_window_size=None)
    shapes = {k: tuple(v.shape) for k, v in dec.named_parameters().items()}
    assert sorted(shapes) == [str(s) for s in g[f"{name}.param_names"]]          # the reference's parameter names
    dec.load_weights(synth.synth_named_weights(shapes, seed=int(g[f"{name}.seed"])))
    noises = [g[f"{name}.noise{i}"] for i in range(len(rates))] if noise else None
    y = dec(g[f"{name}.z"], noise=noises)
    h = torch.as_tensor(g[f"{name}.z"]).cuda()
    for nb, n in zip(dec.noise_blocks(), noises or []):
        nb._noise = n
    for layer in dec.model.layers[:-1]:
        h = layer(h)
    torch.cuda.synchronize()
    ref, pre = g[f"{name}.y"], g[f"{name}.pre_tanh"]
    assert tuple(y.shape) == ref.shape
    e_pre = np.abs(h.cpu().numpy() - pre).max() / np.abs(pre).max()
    e_y = np.abs(y.cpu().numpy() - ref).max()
    record(f"snac decoder {name}", pre_tanh_rel=e_pre, y_abs=e_y, peak=np.abs(ref).max())
    assert e_pre < 1e-4 and e_y < 1e-5, (e_pre, e_y)


def test_snac_decoder_shape_pin_of_the_reference_test():
    """codec/tests/test_snac.py:8-36: rates [8, 8, 4, 2], 236 latent frames -> 120 907 samples (512 * 236 + 75: one
    extra sample per transposed convolution)."""
    dec = S.Decoder(24, 64, [8, 8, 4, 2], True, depthwise=True, attn_window_size=None)
    y = dec(torch.zeros(1, 236, 24))
    assert tuple(y.shape) == (1, 120907, 1) and bool(torch.isfinite(y).all())
    with pytest.raises(NotImplementedError):
        S.Decoder(24, 64, [2], attn_window_size=32)


@pytest.mark.parametrize("stride,cin,cout", [(8, 64, 32), (7, 96, 64), (3, 40, 20), (2, 32, 32)])
def test_wnconv_transpose1d_polyphase(stride, cin, cout):
    r = np.random.Generator(np.random.PCG64(stride))
    m = S.WNConvTranspose1d(cin, cout, 2 * stride, stride=stride, padding=-(-stride // 2), output_padding=stride % 2)
    v = r.standard_normal((cin, 2 * stride, cout), dtype=np.float32)
    gq = r.uniform(0.5, 1.5, (cin, 1, 1)).astype(np.float32)
    b = (0.1 * r.standard_normal(cout, dtype=np.float32)).astype(np.float32)
    m.load_weights({"weight_v": v, "weight_g": gq, "bias": b})
    x = r.standard_normal((2, 37, cin), dtype=np.float32)
    y = m(x).cpu().numpy()
    w = torch.from_numpy(gq * v / np.sqrt((v.astype(np.float64) ** 2).sum(axis=(1, 2), keepdims=True))).double()
    ref = F.conv_transpose1d(torch.from_numpy(x).double().transpose(1, 2), w.permute(0, 2, 1), torch.from_numpy(b).double(),
                             stride=stride, padding=-(-stride // 2), output_padding=1).transpose(1, 2).numpy()
    assert y.shape == ref.shape == (2, 36 * stride - 2 * (-(-stride // 2)) + 2 * stride + 1, cout)
    np.testing.assert_allclose(y, ref, rtol=0, atol=2e-5 * np.abs(ref).max())


@pytest.mark.parametrize("C,K,dil", [(64, 7, 1), (50, 7, 9), (128, 7, 3)])
def test_depthwise_wnconv1d(C, K, dil):
    r = np.random.Generator(np.random.PCG64(C + dil))
    pad = (K - 1) * dil // 2
    m = S.WNConv1d(C, C, K, padding=pad, dilation=dil, groups=C)
    v = r.standard_normal((C, K, 1), dtype=np.float32)
    gq = r.uniform(0.5, 1.5, (C, 1, 1)).astype(np.float32)
    b = (0.1 * r.standard_normal(C, dtype=np.float32)).astype(np.float32)
    m.load_weights({"weight_v": v, "weight_g": gq, "bias": b})
    x = r.standard_normal((3, 41, C), dtype=np.float32)
    res = r.standard_normal((3, 41, C), dtype=np.float32)
    y = m(x, res=torch.from_numpy(res).cuda()).cpu().numpy()
    w = torch.from_numpy(gq * v / np.sqrt((v.astype(np.float64) ** 2).sum(axis=(1, 2), keepdims=True))).double()
    ref = F.conv1d(torch.from_numpy(x).double().transpose(1, 2), w.permute(0, 2, 1), torch.from_numpy(b).double(), padding=pad,
                   dilation=dil, groups=C).transpose(1, 2).numpy() + res
    np.testing.assert_allclose(y, ref, rtol=0, atol=1e-5)
    lib = L.load()
    xd = torch.zeros(1, 8, 4, device="cuda")
    assert lib.kk_dwconv1d_f32(L.ptr(xd), L.ptr(xd), None, None, L.ptr(xd), 1, 8, 4, 7, 1, 2, L.stream()) == L.KK_ERR_INVALID_ARG


DAC_CASES = {"wide": (32, 256, [4, 3, 2]), "small": (16, 64, [2, 5])}


@pytest.mark.parametrize("name", sorted(DAC_CASES))
def test_dac_decoder_matches_reference_golden(golden_dir, name):
    """DAC's generator (codec/models/descript/dac.py:83-131): same layer types as SNAC's dense one, transposed-conv weights
    in MLX's (out, K, in) layout normalised per input channel; golden from tests/golden/make_golden_dac.py."""
    from mlx_audio_b200.codec.models.descript import dac as D
    g = np.load(os.path.join(golden_dir, "dac_decoder.npz"))
    cin, ch, rates = DAC_CASES[name]
    dec = D.Decoder(cin, ch, rates)
    shapes = {k: tuple(v.shape) for k, v in dec.named_parameters().items()}
    assert sorted(shapes) == [str(s) for s in g[f"{name}.param_names"]]
    dec.load_weights(synth.synth_named_weights(shapes, seed=int(g[f"{name}.seed"])))
    y = dec(g[f"{name}.z"])
    h = torch.as_tensor(g[f"{name}.z"]).cuda()
    for layer in dec.model.layers[:-1]:
        h = layer(h)
    torch.cuda.synchronize()
    ref, pre = g[f"{name}.y"], g[f"{name}.pre_tanh"]
    assert tuple(y.shape) == ref.shape
    e_pre = np.abs(h.cpu().numpy() - pre).max() / np.abs(pre).max()
    e_y = np.abs(y.cpu().numpy() - ref).max()
    record(f"dac decoder {name}", pre_tanh_rel=e_pre, y_abs=e_y, peak=np.abs(ref).max())
    assert e_pre < 1e-4 and e_y < 1e-5, (e_pre, e_y)
