"""NumPy/torch-CPU stand-in for the subset of the ``mlx`` API that the reference's
hot-path source files use (mlx_audio/utils.py, tts/models/interpolate.py,
tts/models/kokoro/istftnet.py, tts/models/base.py).

TEST INFRASTRUCTURE ONLY (part of ``oracle/``).  Real MLX is not installable in
this environment (no wheel, no network); this shim exists so that the
reference's OWN, unmodified Python source can be executed to mint golden
vectors (tests/golden/make_golden.py).  Semantics follow MLX's documented
behaviour: float32 default dtype, weakly-typed Python scalars, channels-last
conv1d / conv_transpose1d with (C_out, K, C_in) weights, population variance,
negative gather indices wrap, remainder with the sign of the divisor.
"""
