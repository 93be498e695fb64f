"""Boundary shim for ``generate_audio`` of mlx_audio/tts/generate.py:203-360 restricted to the rebuilt path: drive
``model.generate`` and write the segments as PCM16 WAV files (``{file_prefix}_{i:03d}.wav``, or one joined
``{file_prefix}.wav``).  Model download, reference-audio cloning, playback and the argparse CLI (:361-452) are outside
the path; the model is passed in (``load_model`` result) instead of a hub id."""
from __future__ import annotations

from typing import List, Optional

import numpy as np

from .audio_io import save_wav


def generate_audio(text: str, model, voice: str = "af_heart", speed: float = 1.0, lang_code: str = "a",
                   file_prefix: str = "audio", audio_format: str = "wav", join_audio: bool = False,
                   verbose: bool = True, **kwargs) -> List[str]:
    """Returns the list of files written (the reference returns None and prints them, generate.py:313-314)."""
    if audio_format != "wav":
        raise ValueError("only wav output is available here (soundfile is not installed)")
    written: List[str] = []
    parts = []
    for i, result in enumerate(model.generate(text=text, voice=voice, speed=speed, lang_code=lang_code, **kwargs)):
        audio = result.audio.detach().float().cpu().numpy() if hasattr(result.audio, "detach") else np.asarray(result.audio)
        if join_audio:
            parts.append(audio)
        else:
            name = f"{file_prefix}_{i:03d}.{audio_format}"                    # generate.py:312
            save_wav(name, audio, result.sample_rate)
            written.append(name)
        if verbose:                                                           # generate.py:316-331
            print(f"Duration:              {result.audio_duration}")
            print(f"Audio:                 {result.audio_samples['samples']} samples, "
                  f"{result.audio_samples['samples-per-sec']:.1f} samples-per-sec")
            print(f"Real-time factor:      {result.real_time_factor:.2f}x")
            print(f"Processing time:       {result.processing_time_seconds:.2f}s")
            print(f"Peak memory usage:     {result.peak_memory_usage:.2f}GB")
    if join_audio and parts:                                                  # generate.py:333-343
        name = f"{file_prefix}.{audio_format}"
        save_wav(name, np.concatenate(parts, axis=0), model.sample_rate)
        written.append(name)
    return written
