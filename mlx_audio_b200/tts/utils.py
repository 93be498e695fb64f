"""Boundary shim for mlx_audio/tts/utils.py ``load_model`` (tts/utils.py:150-268) restricted to the rebuilt path:
local checkpoints only (config.json + *.safetensors in a directory), model type "kokoro".  Hub download, conversion
and quantisation (tts/utils.py:23-55, 241-260, 326-421) are packaging, outside the hot path."""
from __future__ import annotations

import glob
import json
import os
from pathlib import Path
from typing import Union

import numpy as np

MODEL_REMAPPING = {"kokoro": "kokoro"}


def load_config(model_path: Union[str, Path]) -> dict:
    """tts/utils.py:124-147."""
    cfg = Path(model_path) / "config.json"
    if not cfg.exists():
        raise FileNotFoundError(f"Config not found at {model_path}")
    with open(cfg, encoding="utf-8") as f:
        return json.load(f)


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
    from safetensors.numpy import load_file
    weights = {}
    for wf in files:
        weights.update(load_file(wf))
    model = Model(ModelConfig.from_dict(config), repo_id=str(model_path), frontend=kwargs.get("frontend"),
                  dtype=kwargs.get("dtype", "bf16"), device=kwargs.get("device"))
    weights = model.sanitize({k: np.asarray(v, dtype=np.float32) if v.dtype != np.float32 else v for k, v in weights.items()})
    model.load_weights(weights, strict=strict)
    return model
