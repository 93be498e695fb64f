"""One launch each of the config-2 STFT / iSTFT / log-mel kernels (for an ncu capture)."""
import importlib.util
import os
import sys
sys.path.insert(0, ".")
from mlx_audio_b200 import _lib as L   # noqa: E402
spec = importlib.util.spec_from_file_location("fft_ab", os.path.join(os.path.dirname(__file__), "fft_ab.py"))
m = importlib.util.module_from_spec(spec)
spec.loader.exec_module(m)
print(m.run(L.load(), 256, 720000))
