"""Boundary shim for mlx_audio/tts/utils.py ``load_model`` (tts/utils.py:150-268) restricted to the rebuilt path:
local checkpoints only (config.json + *.safetensors in a directory), model type "kokoro".  Hub download, conversion
and quantisation (tts/utils.py:23-55, 241-260, 326-421) are packaging, outside the hot path."""
from __future__ import annotations

import glob
import json
import os
from pathlib import Path
from typing import Union

MODEL_REMAPPING = {"kokoro": "kokoro"}


def load_config(model_path: Union[str, Path]) -> dict:
    """tts/utils.py:124-147."""
    cfg = Path(model_path) / "config.json"
    if not cfg.exists():
        raise FileNotFoundError(f"Config not found at {model_path}")
    with open(cfg, encoding="utf-8") as f:
        return json.load(f)


def read_checkpoint(files, prefixes=("decoder.", "predictor.")) -> dict:
    """Read the tensors of the rebuilt sub-modules from safetensors files as float32 NumPy arrays.  Goes through the
    "pt" framework: safetensors' NumPy loader has no bfloat16, and the reference's default checkpoint
    (mlx-community/Kokoro-82M-bf16, tts/generate.py:366) is bf16.  Keys outside ``prefixes`` (the out-of-scope text
    front end: bert.*, bert_encoder.*, text_encoder.*) are never decoded."""
    from safetensors import safe_open
    weights = {}
    for wf in files:
        with safe_open(str(wf), framework="pt", device="cpu") as f:
            for k in f.keys():
                if k.startswith(tuple(prefixes)):
                    weights[k] = f.get_tensor(k).float().numpy()
    return weights


def load_model(model_path: Union[str, Path], lazy: bool = False, strict: bool = True, **kwargs):
    """Load a Kokoro checkpoint directory into ``mlx_audio_b200...kokoro.Model``.  ``kwargs``: ``frontend`` (see
    kokoro.Model), ``dtype`` ("bf16" | "fp32"), ``device``.  Raises like the reference: FileNotFoundError for a
    missing path / weights (tts/utils.py:206-213), ValueError for an unsupported model type (:108-113)."""
    from .models.kokoro import Model, ModelConfig

    model_path = Path(model_path)
    if not model_path.exists():
        raise FileNotFoundError(f"Model path {model_path} does not exist (hub download is outside this path)")
    config = load_config(model_path)
    config.update({k: v for k, v in kwargs.items() if k not in ("frontend", "dtype", "device")})
    model_type = MODEL_REMAPPING.get(config.get("model_type", "kokoro"), config.get("model_type"))
    if model_type != "kokoro":
        raise ValueError(f"Model type {model_type} not supported.")
    files = sorted(glob.glob(str(model_path / "*.safetensors")))
    if not files:
        raise FileNotFoundError(f"No safetensors found in {model_path}")
    weights = read_checkpoint(files)
    model = Model(ModelConfig.from_dict(config), repo_id=str(model_path), frontend=kwargs.get("frontend"),
                  dtype=kwargs.get("dtype", "bf16"), device=kwargs.get("device"))
    weights = model.sanitize(weights)
    model.load_weights(weights, strict=strict)
    return model
