#!/usr/bin/env python
"""Mint tests/golden/bigvgan.npz by executing the reference's own BigVGAN generator
(mlx_audio/codec/models/bigvgan/{bigvgan,amp,conv,resample,activation}.py, unmodified, on oracle/mlx_shim) on seeded
weights (mlx_audio_b200.synth.synth_named_weights; the kaiser-sinc filters keep the values the reference designs) and
inputs.  Also stores one ``Activation1d`` on its own (edge handling of the fused kernel).  Runs only where
/root/reference exists."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from mlx_audio_b200 import synth  # noqa: E402
from oracle import ref_loader  # noqa: E402

CASES = {
    # AMPBlock1, log-scale SnakeBeta, tanh head
    "amp1": (dict(num_mels=16, upsample_rates=[4, 2], upsample_kernel_sizes=[8, 4], upsample_initial_channel=64, resblock="1",
                  resblock_kernel_sizes=[3, 7], resblock_dilation_sizes=[[1, 3, 5], [1, 3, 5]], activation="snakebeta",
                  snake_logscale=True), 2, 20),
    # AMPBlock2, linear-scale SnakeBeta, clip head without bias, three kernels, tcgen05-sized channels
    "amp2": (dict(num_mels=20, upsample_rates=[2, 2], upsample_kernel_sizes=[4, 4], upsample_initial_channel=128, resblock="2",
                  resblock_kernel_sizes=[3, 5, 11], resblock_dilation_sizes=[[1, 3], [1, 3], [1, 3]], activation="snakebeta",
                  snake_logscale=False, use_bias_at_final=False, use_tanh_at_final=False), 1, 33),
}


def run_case(ns, cfg, B, T, seed):
    mx = ns.mx
    model = ns.bigvgan.BigVGAN(ns.bigvgan.BigVGANConfig(**cfg))
    shapes = {k: tuple(v.shape) for k, v in model.named_parameters().items()}
    w = synth.synth_named_weights({k: s for k, s in shapes.items() if not k.endswith("filter")}, seed=seed)
    for k, v in w.items():
        model.set_parameter(k, v)
    r = np.random.Generator(np.random.PCG64(seed + 1))
    x = r.standard_normal((B, cfg["num_mels"], T), dtype=np.float32)
    y = np.asarray(model(mx.array(x))).astype(np.float32)
    # the signal before the tanh / clip head
    model.use_tanh_at_final = False
    clip = mx.clip
    mx.clip = lambda a, lo, hi: a
    try:
        pre = np.asarray(model(mx.array(x))).astype(np.float32)
    finally:
        mx.clip = clip
    return shapes, x, y, pre


def main():
    ns = ref_loader.load_reference_bigvgan()
    out = {"backend": "shim" if ns.shim else "mlx"}
    for name, (cfg, B, T) in CASES.items():
        seed = 7100 + len(name)
        shapes, x, y, pre = run_case(ns, cfg, B, T, seed)
        out[f"{name}.x"], out[f"{name}.y"], out[f"{name}.pre"], out[f"{name}.seed"] = x, y, pre, seed
        out[f"{name}.param_names"] = np.array(sorted(shapes))
        print(name, x.shape, "->", y.shape, "peak", float(np.abs(y).max()), "pre std", float(pre.std()),
              "clipped", float((np.abs(pre) > 1).mean()))
    # Activation1d alone: short and long sequences (every edge case of the replicate padding)
    A = ns.bigvgan_resample.Activation1d
    r = np.random.Generator(np.random.PCG64(7177))
    for T in (1, 2, 3, 7, 8, 9, 40):
        act = ns.bigvgan_activation.SnakeBeta(6, alpha_logscale=True)
        act.alpha = ns.mx.array(r.uniform(-0.5, 1.0, 6).astype(np.float32))
        act.beta = ns.mx.array(r.uniform(-0.5, 1.0, 6).astype(np.float32))
        x = r.standard_normal((2, T, 6), dtype=np.float32)
        out[f"act.T{T}.x"], out[f"act.T{T}.alpha"], out[f"act.T{T}.beta"] = x, np.asarray(act.alpha), np.asarray(act.beta)
        out[f"act.T{T}.y"] = np.asarray(A(act)(ns.mx.array(x.copy()))).astype(np.float32)
    np.savez_compressed(os.path.join(HERE, "bigvgan.npz"), **out)


if __name__ == "__main__":
    main()
