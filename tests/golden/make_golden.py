#!/usr/bin/env python
"""Mint golden vectors by executing the REFERENCE'S OWN source files.

Run in the build container only (needs /root/reference):

    python tests/golden/make_golden.py

It imports mlx_audio/utils.py, tts/models/interpolate.py and
tts/models/kokoro/istftnet.py unmodified from /root/reference (via
oracle/ref_loader.py; real MLX if importable, else oracle/mlx_shim) and stores
their outputs on seeded synthetic inputs (mlx_audio_b200/synth.py) as small
.npz fixtures next to this script.  Weights/inputs are NOT stored - they are
regenerated from the seeds recorded in each file.
"""
from __future__ import annotations

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from mlx_audio_b200 import synth  # noqa: E402
from oracle import ref_loader  # noqa: E402


def main():
    ns = ref_loader.load_reference()
    mx = ns.mx
    backend = "mlx_shim(numpy/torch-cpu)" if ns.shim else "mlx"
    print("reference backend:", backend)

    # ---- 1. interpolate: the reference's own KAT inputs (tts/tests/test_interpolate.py:40-84) + a ragged case
    x4 = np.array([[[1.0, 2.0, 3.0, 4.0]]], np.float32)
    x4b = np.array([[[1.0, 3.0, 5.0, 7.0]]], np.float32)
    rng = np.random.Generator(np.random.PCG64(77))
    xr = rng.standard_normal((2, 3, 37), dtype=np.float32)
    np.savez_compressed(
        os.path.join(HERE, "interpolate.npz"),
        backend=backend,
        x4=x4, x4b=x4b, xr=xr,
        nearest_4_to_8=np.asarray(ns.interpolate.interpolate1d(mx.array(x4), size=8, mode="nearest")),
        nearest_4_to_2=np.asarray(ns.interpolate.interpolate1d(mx.array(x4), size=2, mode="nearest")),
        linear_ac_4_to_7=np.asarray(ns.interpolate.interpolate1d(mx.array(x4b), size=7, mode="linear", align_corners=True)),
        linear_4_to_7=np.asarray(ns.interpolate.interpolate1d(mx.array(x4b), size=7, mode="linear", align_corners=False)),
        linear_up_x5=np.asarray(ns.interpolate.interpolate(mx.array(xr), scale_factor=5, mode="linear")),
        linear_down_37_to_9=np.asarray(ns.interpolate.interpolate(mx.array(xr), size=9, mode="linear")),
        nearest_37_to_50=np.asarray(ns.interpolate.interpolate(mx.array(xr), size=50, mode="nearest")),
    )

    # ---- 2. utils.stft / istft / mel_filters
    sig = synth.synth_audio(1, 6000, seed=3001)[0]
    out = dict(backend=backend, seed=3001, n=6000)
    for tag, kw in {
        "1024_256": dict(n_fft=1024, hop_length=256),
        "400_160": dict(n_fft=400, hop_length=160),
        "20_5": dict(n_fft=20, hop_length=5),
        "800_default": dict(),
        "512_128_win400_hamming": dict(n_fft=512, hop_length=128, win_length=400, window="hamming"),
        "256_64_constant": dict(n_fft=256, hop_length=64, pad_mode="constant"),
        "256_64_nocenter": dict(n_fft=256, hop_length=64, center=False),
    }.items():
        S = np.asarray(ns.utils.stft(mx.array(sig), **kw))
        out[f"stft_{tag}"] = S.astype(np.complex64)
    for tag, kw in {
        "1024_256": dict(hop_length=256, win_length=1024),
        "400_160": dict(hop_length=160, win_length=400),
        "20_5": dict(hop_length=5, win_length=20),
    }.items():
        S = out[f"stft_{tag}"]
        out[f"istft_{tag}"] = np.asarray(ns.utils.istft(mx.array(S.T), **kw)).astype(np.float32)
    S = out["stft_1024_256"]
    out["istft_1024_256_length5000"] = np.asarray(
        ns.utils.istft(mx.array(S.T), hop_length=256, win_length=1024, length=5000)).astype(np.float32)
    out["istft_1024_256_symwin"] = np.asarray(  # Vocos passes an explicit symmetric window (vocos.py:134-139)
        ns.utils.istft(mx.array(S.T), hop_length=256, win_length=1024, window=ns.utils.hanning(1024))).astype(np.float32)
    out["mel_24000_1024_80_htk"] = np.asarray(ns.utils.mel_filters(24000, 1024, 80))
    out["mel_24000_1024_100_htk"] = np.asarray(ns.utils.mel_filters(24000, 1024, 100, norm=None, mel_scale="htk"))
    out["mel_16000_400_80_slaney"] = np.asarray(ns.utils.mel_filters(16000, 400, 80, norm="slaney", mel_scale=None))
    out["mel_16000_400_128_slaney"] = np.asarray(ns.utils.mel_filters(16000, 400, 128, norm="slaney", mel_scale=None))
    out["mel_22050_512_40_fmin_fmax"] = np.asarray(ns.utils.mel_filters(22050, 512, 40, f_min=50.0, f_max=8000.0))
    for name in ("hanning", "hamming", "blackman", "bartlett"):
        out[f"win_{name}_20"] = np.asarray(getattr(ns.utils, name)(20))
    # vocos shape pin (codec/tests/test_vocos.py:61-73): 120000 -> 469 frames -> [:-1] -> 119552 after istft
    z = mx.array(np.zeros(120000, np.float32))
    Sz = ns.utils.stft(z, n_fft=1024, hop_length=256)
    out["vocos_frames"] = np.int64(Sz.shape[0])
    yz = ns.utils.istft(mx.array(np.asarray(Sz)[:-1].T), hop_length=256, win_length=1024)
    out["vocos_istft_len"] = np.int64(yz.shape[0])
    np.savez_compressed(os.path.join(HERE, "dsp.npz"), **out)

    # ---- 3. Decoder (+ intermediate taps of the harmonic source) at a tiny shape
    B, T = 2, 16
    w = synth.synth_decoder_weights(seed=2000)
    inp = synth.synth_decoder_inputs(B, T, seed=1000)
    audio = ref_loader.run_reference_decoder(w, synth.KOKORO_DECODER, inp["asr"], inp["F0_curve"], inp["N"], inp["s"],
                                             inp["noise"])
    # harmonic source + har tensor straight from the reference generator's sub-modules
    dec = ref_loader.build_reference_decoder(w, synth.KOKORO_DECODER)
    gen = dec.generator
    mx.random.queue = [np.full((B, 9), 0.37, np.float32), inp["noise"], np.zeros((B, 600 * T, 1), np.float32)]
    f0 = gen.f0_upsamp(mx.array(inp["F0_curve"])[:, None].transpose(0, 2, 1))
    har_source, _, uv = gen.m_source(f0)
    har_source = mx.squeeze(har_source.transpose(0, 2, 1), axis=1)
    spec, phase = gen.stft.transform(har_source)
    np.savez_compressed(
        os.path.join(HERE, "decoder_b2_t16.npz"),
        backend=backend, weight_seed=2000, input_seed=1000, batch=B, frames=T,
        audio=np.asarray(audio).astype(np.float32),
        har_source=np.asarray(har_source).astype(np.float32),
        har_spec=np.asarray(spec).astype(np.float32),
        har_phase=np.asarray(phase).astype(np.float32),
    )
    # ---- 4. ProsodyPredictor.F0Ntrain (modules.py:355-377): reference modules.py + istftnet.AdainResBlk1d, unmodified
    Bp, Lp = 2, 12
    wp = synth.synth_prosody_weights(seed=2100)
    ip = synth.synth_prosody_inputs(Bp, Lp, seed=1100)
    F0, N = ref_loader.run_reference_f0ntrain(wp, ip["x"], ip["s"])
    np.savez_compressed(os.path.join(HERE, "prosody_b2_l12.npz"), backend=backend, weight_seed=2100, input_seed=1100,
                        batch=Bp, frames=Lp, F0=F0.astype(np.float32), N=N.astype(np.float32))
    # ---- 5. consumers of the shared DSP: the reference's own front-end files
    fe = ref_loader.load_reference_frontends()
    sig24 = synth.synth_audio(1, 24000, seed=12)[0]
    rngh = np.random.Generator(np.random.PCG64(31))
    dim, Lh = 64, 20
    hw = (rngh.standard_normal((1026, dim), dtype=np.float32) * np.float32(0.2)).astype(np.float32)
    hb = (rngh.standard_normal((1026,), dtype=np.float32) * np.float32(0.1)).astype(np.float32)
    hb[:513] -= np.float32(1.0)
    hx = rngh.standard_normal((1, Lh, dim), dtype=np.float32)
    head = fe.vocos.ISTFTHead(dim, 1024, 256)
    head.set_parameter("out.weight", hw)
    head.set_parameter("out.bias", hb)
    np.savez_compressed(
        os.path.join(HERE, "frontends.npz"), backend=backend, audio_seed=12, n=24000, head_seed=31, head_dim=dim,
        head_frames=Lh,
        vocos_mel=np.asarray(fe.vocos_mel.log_mel_spectrogram(mx.array(sig24))).astype(np.float32),
        vocos_mel_pad=np.asarray(fe.vocos_mel.log_mel_spectrogram(mx.array(sig24), n_mels=80, padding=300)).astype(np.float32),
        indextts_mel=np.asarray(fe.indextts_mel.log_mel_spectrogram(mx.array(sig24))).astype(np.float32),
        whisper_mel=np.asarray(fe.whisper_audio.log_mel_spectrogram(sig24)).astype(np.float32),
        whisper_mel_128=np.asarray(fe.whisper_audio.log_mel_spectrogram(sig24, n_mels=128, padding=480)).astype(np.float32),
        istft_head=np.asarray(head(mx.array(hx))).astype(np.float32),
        parakeet_mel=np.asarray(fe.parakeet_audio.log_mel_spectrogram(mx.array(sig24), fe.parakeet_audio.PreprocessArgs(
            16000, "per_feature", 0.025, 0.01, "hann", 80, 512, 0.0))).astype(np.float32),
        s3_mel=np.asarray(fe.s3_utils.log_mel_spectrogram(mx.array(sig24))).astype(np.float32),
        spark_mel=np.asarray(fe.spark_mel_spectrogram(mx.array(sig24))).astype(np.float32),
        parakeet_mel_global=np.asarray(fe.parakeet_audio.log_mel_spectrogram(mx.array(sig24), fe.parakeet_audio.PreprocessArgs(
            16000, "all", 0.025, 0.01, "hamming", 128, 512, 0.0, pad_to=30000))).astype(np.float32),
    )
    # ---- 6. Snake activations of the codec generators
    rs = np.random.Generator(np.random.PCG64(41))
    xs = (2.0 * rs.standard_normal((2, 12, 50), dtype=np.float32)).astype(np.float32)
    al = rs.uniform(0.3, 2.0, 12).astype(np.float32)
    be = rs.uniform(0.3, 2.0, 50).astype(np.float32)
    al50 = rs.uniform(-1.0, 1.0, 50).astype(np.float32)
    sn = fe.bigvgan_activation.Snake(12); sn.alpha = mx.array(al)
    snl = fe.bigvgan_activation.Snake(12, alpha_logscale=True); snl.alpha = mx.array(al - 1.0)
    sb = fe.bigvgan_activation.SnakeBeta(50, alpha_logscale=True); sb.alpha = mx.array(al50); sb.beta = mx.array(be - 1.0)
    wn = fe.dac_layers.WNConv1d(6, 10, 5, stride=2, padding=3, dilation=2)
    wv = rs.standard_normal((10, 5, 6), dtype=np.float32)
    wg = rs.uniform(0.5, 1.5, (10, 1, 1)).astype(np.float32)
    wb = (0.1 * rs.standard_normal((10,), dtype=np.float32)).astype(np.float32)
    xw = rs.standard_normal((2, 37, 6), dtype=np.float32)
    wn.weight_v, wn.weight_g, wn.bias = mx.array(wv), mx.array(wg), mx.array(wb)
    np.savez_compressed(
        os.path.join(HERE, "snake.npz"), backend=backend, seed=41,
        wnconv=np.asarray(wn(mx.array(xw))).astype(np.float32),
        snake=np.asarray(sn(mx.array(xs.copy()))).astype(np.float32),
        snake_log=np.asarray(snl(mx.array(xs.copy()))).astype(np.float32),
        snakebeta_log=np.asarray(sb(mx.array(xs.copy()))).astype(np.float32),
        dac_snake=np.asarray(fe.dac_layers.snake(mx.array(xs.copy()), mx.array(be.reshape(1, 1, 50)))).astype(np.float32))
    print("wrote", sorted(f for f in os.listdir(HERE) if f.endswith(".npz")))


if __name__ == "__main__":
    main()
