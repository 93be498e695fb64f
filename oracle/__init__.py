"""CPU oracle for the Kokoro ISTFTNet decoder + STFT/iSTFT/mel path.

TEST INFRASTRUCTURE ONLY.  Nothing in ``mlx_audio_b200`` may import, call, link
or execute anything in this package.  The only permitted users are ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl
reference`` legs, and there only as the checker / the timed CPU baseline.

What it is: a float32 NumPy / torch-CPU restatement of the reference's
arithmetic for this path (each function cites the reference file:line it
follows):

* ``oracle.dsp``      <- mlx_audio/utils.py            (windows, stft, istft, mel_filters)
* ``oracle.interp``   <- mlx_audio/tts/models/interpolate.py
* ``oracle.istftnet`` <- mlx_audio/tts/models/kokoro/istftnet.py (Decoder and everything below it)

Parity pinning (see DESIGN.md "Oracle"):  the reference's arithmetic lives in
the third-party MLX library (requirements.txt:7 ``mlx>=0.25.2``), which is NOT
installed here and cannot be fetched, so the real reference cannot execute in
this environment.  The oracle is pinned in three ways instead:

1. against the reference's only numeric known-answer tests for this path
   (mlx_audio/tts/tests/test_interpolate.py:40-84) and its shape pins
   (codec/tests/test_vocos.py:61-73: 120000 samples -> 469 frames -> 119552);
2. against the reference's OWN Python source files (istftnet.py, utils.py,
   interpolate.py), imported unmodified from /root/reference and executed on
   ``oracle/mlx_shim`` - a small NumPy/torch-CPU stand-in for the ``mlx.core`` /
   ``mlx.nn`` API those three files use.  ``tests/golden/make_golden.py`` does
   this in the build container and commits the resulting vectors under
   ``tests/golden/``; the restated oracle must reproduce them;
3. block-by-block against independent library implementations
   (torch.stft/istft geometry, torchaudio melscale_fbanks, F.instance_norm,
   F.interpolate) with the documented deliberate deviations.

MLX's own kernels never ran, so parity with *MLX numerics* is "pinned to the
reference source over a shim", not to MLX binaries; DESIGN.md says so.
"""
