"""Checkpoint side of the reference's export script (host logic only, no device work):

  rave/core.py:132-139          search_for_run: newest `*.ckpt` under a run directory whose name holds `mode`
  scripts/export.py:543-545     gin.parse_config_file(run/config.gin), search_for_run(run)
  scripts/export.py:557-568     torch.load(ckpt); `callbacks.EMA` (scripts/train.py:64-103: a flat
                                {parameter name: tensor} dict) when --ema_weights, else `state_dict`;
                                load_state_dict(..., strict=False)
  scripts/export.py:598-600     remove_weight_norm: folded at load here (weights.py:_folded)

`split_state_dict` separates the `speaker_encoder.*` tensors (model.py:164-174) from the model's own.
"""
import os
from pathlib import Path


def search_for_run(run_path, mode="last"):
    """rave/core.py:132-139."""
    if run_path is None:
        return None
    if ".ckpt" in run_path:
        return run_path
    ckpts = map(str, Path(run_path).rglob("*.ckpt"))
    ckpts = filter(lambda e: mode in os.path.basename(str(e)), ckpts)
    ckpts = sorted(ckpts)
    if len(ckpts):
        return ckpts[-1]
    return None


def load_checkpoint(path, ema_weights=False):
    """scripts/export.py:557-568 -> the flat state_dict the reference would hand to load_state_dict.
    With ema_weights the EMA callback's dict holds parameters only (buffers keep their constructed
    values upstream, hence strict=False there and here)."""
    import torch
    try:
        ckpt = torch.load(path, map_location="cpu", weights_only=False)
    except TypeError:                      # torch < 1.13 has no weights_only
        ckpt = torch.load(path, map_location="cpu")
    if ema_weights and "EMA" in ckpt.get("callbacks", {}):
        return dict(ckpt["callbacks"]["EMA"])
    return dict(ckpt["state_dict"])


def split_state_dict(sd, prefix="speaker_encoder."):
    """-> (model tensors, speaker-encoder tensors with the prefix stripped)."""
    own = {k: v for k, v in sd.items() if not k.startswith(prefix)}
    spk = {k[len(prefix):]: v for k, v in sd.items() if k.startswith(prefix)}
    return own, spk
