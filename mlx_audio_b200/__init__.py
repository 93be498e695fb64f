"""mlx_audio_b200: B200-native (sm_100a) implementation of mlx-audio's Kokoro ISTFTNet decoder hot path and
the shared STFT / iSTFT / mel transforms, behind the reference's own Python surface.

    from mlx_audio_b200.utils import stft, istft, mel_filters
    from mlx_audio_b200.tts.utils import load_model
    from mlx_audio_b200.tts.models.kokoro import Decoder, KokoroPipeline, Model

All arithmetic runs in hand-written CUDA kernels (libkokoro_b200.so, C ABI in include/kokoro_b200.h).
There is no CPU fallback.
"""
__version__ = "0.1.0"
