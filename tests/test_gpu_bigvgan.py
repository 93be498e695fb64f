"""BigVGAN generator on the decoder's convolution kernels + the fused anti-aliased activation, against
goldens minted by executing the reference's own codec/models/bigvgan/*.py (tests/golden/make_golden_bigvgan.py)."""
import os

import numpy as np
import pytest
import torch

from mlx_audio_b200 import synth
from mlx_audio_b200.codec.models.bigvgan.activation import Snake, SnakeBeta
from mlx_audio_b200.codec.models.bigvgan.bigvgan import BigVGAN, BigVGANConfig
from mlx_audio_b200.codec.models.bigvgan.conv import WNConvTranspose1d
from mlx_audio_b200.codec.models.bigvgan.resample import Activation1d, kaiser_sinc_filter1d
from parity_gates import record

pytestmark = pytest.mark.gpu

CASES = {
    "amp1": dict(num_mels=16, upsample_rates=[4, 2], upsample_kernel_sizes=[8, 4], upsample_initial_channel=64, resblock="1",
                 resblock_kernel_sizes=[3, 7], resblock_dilation_sizes=[[1, 3, 5], [1, 3, 5]], activation="snakebeta",
                 snake_logscale=True),
    "amp2": dict(num_mels=20, upsample_rates=[2, 2], upsample_kernel_sizes=[4, 4], upsample_initial_channel=128, resblock="2",
                 resblock_kernel_sizes=[3, 5, 11], resblock_dilation_sizes=[[1, 3], [1, 3], [1, 3]], activation="snakebeta",
                 snake_logscale=False, use_bias_at_final=False, use_tanh_at_final=False),
}


@pytest.mark.parametrize("T", [1, 2, 3, 7, 8, 9, 40])
def test_activation1d_matches_reference(golden_dir, T):
    """UpSample1d -> SnakeBeta -> DownSample1d (resample.py:152-177) as one kernel, incl. sequences shorter than the
    filters (the edge replication of both paddings)."""
    g = np.load(os.path.join(golden_dir, "bigvgan.npz"))
    act = SnakeBeta(6, alpha_logscale=True)
    act.alpha, act.beta = torch.from_numpy(g[f"act.T{T}.alpha"]), torch.from_numpy(g[f"act.T{T}.beta"])
    y = Activation1d(act)(g[f"act.T{T}.x"]).cpu().numpy()
    ref = g[f"act.T{T}.y"]
    assert y.shape == ref.shape
    assert np.abs(y - ref).max() <= 2e-6 * max(1.0, float(np.abs(ref).max())), np.abs(y - ref).max()


def test_activation1d_long_sequence_matches_unfused_torch():
    """Interior tiles, tile tails and many channels against the three reference steps written with torch ops."""
    import torch.nn.functional as F
    r = np.random.Generator(np.random.PCG64(5))
    B, T, C = 2, 1003, 160
    x = torch.from_numpy(r.standard_normal((B, T, C), dtype=np.float32)).cuda()
    act = SnakeBeta(C, alpha_logscale=False)
    act.alpha = torch.from_numpy(r.uniform(0.5, 2.0, C).astype(np.float32))
    act.beta = torch.from_numpy(r.uniform(0.5, 2.0, C).astype(np.float32))
    y = Activation1d(act)(x)
    f = kaiser_sinc_filter1d(0.25, 0.3, 12).reshape(1, 1, 12).double().cuda().expand(C, 1, 12)
    xd = x.double().transpose(1, 2)
    u = 2 * F.conv_transpose1d(F.pad(xd, (5, 5), mode="replicate"), f, stride=2, groups=C)[..., 15:-15]
    a, b = act.alpha.double().cuda()[None, :, None], act.beta.double().cuda()[None, :, None]
    v = u + torch.sin(u * a) ** 2 / (b + 1e-9)
    ref = F.conv1d(F.pad(v, (5, 6), mode="replicate"), f, stride=2, groups=C).transpose(1, 2)
    assert tuple(y.shape) == tuple(ref.shape) == (B, T, C)
    assert float((y.double() - ref).abs().max()) <= 2e-6 * float(ref.abs().max())


@pytest.mark.parametrize("name", sorted(CASES))
def test_bigvgan_matches_reference_golden(golden_dir, name):
    g = np.load(os.path.join(golden_dir, "bigvgan.npz"))
    model = BigVGAN(BigVGANConfig(**CASES[name]))
    shapes = {k: tuple(v.shape) for k, v in model.named_parameters().items()}
    assert sorted(shapes) == [str(s) for s in g[f"{name}.param_names"]]          # the reference's parameter names
    w = synth.synth_named_weights({k: s for k, s in shapes.items() if not k.endswith("filter")}, seed=int(g[f"{name}.seed"]))
    model.load_weights(w, strict=False)
    y = model(g[f"{name}.x"])
    ref, pre_ref = g[f"{name}.y"], g[f"{name}.pre"]
    assert tuple(y.shape) == ref.shape
    e_y = np.abs(y.cpu().numpy() - ref).max()
    record(f"bigvgan {name}", y_abs=e_y, peak=np.abs(ref).max())
    assert e_y <= 1e-5 * max(1.0, float(np.abs(pre_ref).max())) + 2e-6, e_y


def test_bigvgan_sanitize_and_errors():
    cfg = CASES["amp1"]
    model = BigVGAN(BigVGANConfig(**cfg))
    cur = model.named_parameters()
    k = "ups.0.0.weight_v"                                                        # ours: (out, K, in); torch: (in, out, K)
    torch_layout = cur[k].permute(2, 0, 1).contiguous().numpy()
    kc = "conv_pre.weight_v"                                                      # torch conv: (out, in, K)
    s = model.sanitize({k: torch_layout, kc: cur[kc].permute(0, 2, 1).contiguous().numpy(), "x.num_batches_tracked": np.zeros(1)})
    assert tuple(s[k].shape) == tuple(cur[k].shape) and np.array_equal(s[k], cur[k].numpy())
    assert tuple(s[kc].shape) == tuple(cur[kc].shape) and "x.num_batches_tracked" not in s
    with pytest.raises(NotImplementedError):
        BigVGAN(BigVGANConfig(**{**cfg, "activation": "snake"}))
    with pytest.raises(NotImplementedError):
        WNConvTranspose1d(8, 4, 7, 2)
    with pytest.raises(NotImplementedError):
        Activation1d(Snake(4))
