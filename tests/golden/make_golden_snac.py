#!/usr/bin/env python
"""Mint tests/golden/snac_decoder.npz by executing the reference's own SNAC generator
(mlx_audio/codec/models/snac/layers.py: Decoder / DecoderBlock / ResidualUnit / NoiseBlock / WNConv1d /
WNConvTranspose1d / Snake1d, unmodified, on oracle/mlx_shim) on seeded weights (mlx_audio_b200.synth.synth_named_weights)
and inputs.  Runs only where /root/reference exists.  Also re-checks the reference's own shape pin
(codec/tests/test_snac.py:36) on the shim."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from mlx_audio_b200 import synth  # noqa: E402
from oracle import ref_loader  # noqa: E402

CASES = {
    # name: (input_channel, channels, rates, noise, depthwise, batch, frames)
    "dw": (24, 256, [4, 3, 2], True, True, 2, 20),        # depthwise + noise, even and odd strides, tcgen05-sized channels
    "dense": (16, 64, [2, 3], False, False, 1, 33),       # dense k=7 convolutions, no noise, small channels
}


def run_case(ns, cfg, seed):
    cin, ch, rates, noise, dw, B, T = cfg
    mx = ns.mx
    dec = ns.snac_layers.Decoder(cin, ch, rates, noise, depthwise=dw, attn_window_size=None)
    shapes = {k: tuple(v.shape) for k, v in dec.named_parameters().items()}
    w = synth.synth_named_weights(shapes, seed=seed)
    for k, v in w.items():
        dec.set_parameter(k, v)
    r = np.random.Generator(np.random.PCG64(seed + 1))
    z = r.standard_normal((B, T, cin), dtype=np.float32)
    noises = [r.standard_normal((B, 1, ch // 2 ** (i + 1)), dtype=np.float32) for i in range(len(rates))] if noise else []
    mx.random.queue = [n.copy() for n in noises]
    mx.random.log = []
    h = mx.array(z)
    for layer in dec.model.layers[:-1]:          # everything but the Tanh head: the unsaturated signal
        h = layer(h)
    pre = np.asarray(h).astype(np.float32)
    assert not mx.random.queue
    mx.random.queue = [n.copy() for n in noises]
    y = np.asarray(dec(mx.array(z))).astype(np.float32)
    assert not mx.random.queue
    return shapes, z, noises, y, pre


def main():
    ns = ref_loader.load_reference_snac()
    # the reference's own shape pin: 236 frames -> 120 907 samples with rates [8, 8, 4, 2] (test_snac.py:8-36)
    pin = ns.snac_layers.Decoder(24, 64, [8, 8, 4, 2], True, depthwise=True, attn_window_size=None)
    assert np.asarray(pin(ns.mx.array(np.zeros((1, 236, 24), np.float32)))).shape == (1, 120907, 1)
    out = {"backend": "shim" if ns.shim else "mlx"}
    for name, cfg in CASES.items():
        seed = 7000 + len(name)
        shapes, z, noises, y, pre = run_case(ns, cfg, seed)
        out[f"{name}.z"], out[f"{name}.y"], out[f"{name}.pre_tanh"], out[f"{name}.seed"] = z, y, pre, seed
        for i, n in enumerate(noises):
            out[f"{name}.noise{i}"] = n
        out[f"{name}.param_names"] = np.array(sorted(shapes))
        print(name, "z", z.shape, "->", y.shape, "peak", float(np.abs(y).max()), "saturated", float((np.abs(y) > 0.99).mean()),
              "pre-tanh std", float(pre.std()))
    np.savez_compressed(os.path.join(HERE, "snac_decoder.npz"), **out)


if __name__ == "__main__":
    main()
