#!/usr/bin/env python
"""Mint tests/golden/dac_decoder.npz by executing the reference's own DAC generator (mlx_audio/codec/models/descript/dac.py
Decoder / DecoderBlock / ResidualUnit over its nn/layers.py, unmodified, on oracle/mlx_shim) on seeded weights
(mlx_audio_b200.synth.synth_named_weights) and inputs.  Runs only where /root/reference exists."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from mlx_audio_b200 import synth  # noqa: E402
from oracle import ref_loader  # noqa: E402

CASES = {"wide": (32, 256, [4, 3, 2], 2, 18),      # (input_channel, channels, rates, batch, frames): tcgen05-sized channels
         "small": (16, 64, [2, 5], 1, 29)}


def main():
    ns = ref_loader.load_reference_dac()
    mx = ns.mx
    out = {"backend": "shim" if ns.shim else "mlx"}
    for name, (cin, ch, rates, B, T) in CASES.items():
        seed = 7100 + len(name)
        dec = ns.dac.Decoder(cin, ch, rates)
        shapes = {k: tuple(v.shape) for k, v in dec.named_parameters().items()}
        for k, v in synth.synth_named_weights(shapes, seed=seed).items():
            dec.set_parameter(k, v)
        z = np.random.Generator(np.random.PCG64(seed + 1)).standard_normal((B, T, cin), dtype=np.float32)
        h = mx.array(z)
        for layer in dec.model.layers[:-1]:
            h = layer(h)
        pre = np.asarray(h).astype(np.float32)
        y = np.asarray(dec(mx.array(z))).astype(np.float32)
        out[f"{name}.z"], out[f"{name}.y"], out[f"{name}.pre_tanh"], out[f"{name}.seed"] = z, y, pre, seed
        out[f"{name}.param_names"] = np.array(sorted(shapes))
        print(name, z.shape, "->", y.shape, "peak", float(np.abs(y).max()), "saturated", float((np.abs(y) > 0.99).mean()))
    np.savez_compressed(os.path.join(HERE, "dac_decoder.npz"), **out)


if __name__ == "__main__":
    main()
