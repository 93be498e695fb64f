"""Voice-pack I/O for the Kokoro pipeline: a pack is a (510, 1, 256) float tensor; row
``len(phonemes) - 1`` is the style of an utterance, its first 128 values feed the decoder (pipeline.py:236,
kokoro.py:165).  The reference ships a torch-free unpickler for the ``.pt`` files (kokoro/voice.py:9-81) because MLX
hosts have no torch; this host has, so ``.pt`` goes through ``torch.load(weights_only=True)`` (tensor-only unpickling).
``.npy`` / ``.npz`` / ``.safetensors`` are accepted too."""
from __future__ import annotations

import numpy as np

PACK_SHAPE_TAIL = (1, 256)


def load_voice_tensor(path: str) -> np.ndarray:
    """-> float32 array (N, 1, 256).  Raises ValueError for a file that is not a voice pack."""
    if path.endswith(".pt") or path.endswith(".pth"):
        import torch
        obj = torch.load(path, map_location="cpu", weights_only=True)
        if isinstance(obj, dict):            # tolerate {"pack": tensor}-style files
            tensors = [v for v in obj.values() if hasattr(v, "shape")]
            if len(tensors) != 1:
                raise ValueError(f"{path}: expected one tensor, found {len(tensors)}")
            obj = tensors[0]
        arr = obj.detach().to(torch.float32).numpy()
    elif path.endswith(".npy"):
        arr = np.load(path)
    elif path.endswith(".npz"):
        z = np.load(path)
        arr = z[z.files[0]]
    elif path.endswith(".safetensors"):
        from safetensors.numpy import load_file
        d = load_file(path)
        arr = d[sorted(d)[0]]
    else:
        raise ValueError(f"unsupported voice file {path!r} (.pt, .npy, .npz, .safetensors)")
    arr = np.asarray(arr, dtype=np.float32)
    if arr.ndim != 3 or arr.shape[1:] != PACK_SHAPE_TAIL:
        raise ValueError(f"{path}: voice pack must be (N, 1, 256), got {arr.shape}")
    return np.ascontiguousarray(arr)


def style_for(pack: np.ndarray, phonemes: str) -> np.ndarray:
    """``pack[len(ps) - 1]`` (pipeline.py:236) -> (1, 256)."""
    if not 1 <= len(phonemes) <= pack.shape[0]:
        raise ValueError(f"phoneme count {len(phonemes)} outside the pack's range 1..{pack.shape[0]}")
    return pack[len(phonemes) - 1]
