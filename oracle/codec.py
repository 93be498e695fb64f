"""TEST INFRASTRUCTURE (not product code): CPU restatement of the BigVGAN anti-aliased activation, in closed form.

``Activation1d`` (mlx_audio/codec/models/bigvgan/resample.py:152-177) with the default ratios (2, 2) and 12-tap filters:
  UpSample1d   :104-134   edge-pad 5; y = conv_transpose1d(x, f, stride 2) (y[n] = sum_i xp[i] f[n - 2 i]); u = 2 y[15:-15]
  SnakeBeta    activation.py:27-51   v = u + sin(alpha u)^2 / (beta + 1e-9), alpha / beta through exp() when log-scale
  DownSample1d :137-149 (LowPassFilter1d :53-101)   edge-pad (5, 6); out[t] = sum_j vp[2 t + j] g[j]
Written out per output sample with clamped indices (what the CUDA kernel evaluates); pinned against the output of the
reference's own classes in tests/golden/bigvgan.npz (tests/test_oracle.py).
"""
import numpy as np


def aa_activation(x, alpha, beta, up_filter, down_filter, logscale: bool, eps: float = 1e-9):
    """x (B, T, C) float32 -> (B, T, C); alpha / beta (C,), filters (12,)."""
    x = np.asarray(x, np.float64)
    B, T, C = x.shape
    f, g = np.asarray(up_filter, np.float64).reshape(-1), np.asarray(down_filter, np.float64).reshape(-1)
    assert f.size == 12 and g.size == 12
    a, b = np.asarray(alpha, np.float64), np.asarray(beta, np.float64)
    if logscale:
        a, b = np.exp(a), np.exp(b)
    clampT = lambda i: min(max(i, 0), T - 1)
    u = np.zeros((B, 2 * T, C))
    for m in range(2 * T):
        q = m // 2
        if m % 2 == 0:      # u[2q] = 2 sum_r f[11 - 2r] x[q - 3 + r]
            u[:, m] = 2 * sum(f[11 - 2 * r] * x[:, clampT(q - 3 + r)] for r in range(6))
        else:               # u[2q+1] = 2 sum_r f[10 - 2r] x[q - 2 + r]
            u[:, m] = 2 * sum(f[10 - 2 * r] * x[:, clampT(q - 2 + r)] for r in range(6))
    v = u + np.sin(u * a) ** 2 / (b + eps)
    out = np.zeros((B, T, C))
    for t in range(T):
        out[:, t] = sum(g[j] * v[:, min(max(2 * t + j - 5, 0), 2 * T - 1)] for j in range(12))
    return out.astype(np.float32)
