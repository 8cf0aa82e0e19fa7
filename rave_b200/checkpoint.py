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


def search_for_run(run_path, mode="last"):
    """What rave/core.py:132-139 returns: `run_path` itself when it already names a checkpoint, else the
    lexicographically last `*.ckpt` anywhere below it whose file name contains `mode`, else None."""
    if run_path is None:
        return None
    run_path = str(run_path)
    if ".ckpt" in run_path:
        return run_path
    best = None
    for root, _dirs, files in os.walk(run_path):
        for name in files:
            if name.endswith(".ckpt") and mode in name:
                full = os.path.join(root, name)
                if best is None or full > best:
                    best = full
    return best


def load_checkpoint(path, ema_weights=False):
    """scripts/export.py:557-568 -> the flat state_dict the reference would hand to load_state_dict.
    With ema_weights the EMA callback's dict holds parameters only (buffers keep their constructed
    values upstream, hence strict=False there and here)."""
    import torch
    try:
        # tensors-only unpickling first: a checkpoint may come from a downloaded run directory
        ckpt = torch.load(path, map_location="cpu", weights_only=True)
    except Exception:
        # Lightning checkpoints pickle callback / hyper-parameter objects next to the tensors
        ckpt = torch.load(path, map_location="cpu", weights_only=False)
    if ema_weights and "EMA" in ckpt.get("callbacks", {}):
        return dict(ckpt["callbacks"]["EMA"])
    return dict(ckpt["state_dict"])


def split_state_dict(sd, prefix="speaker_encoder."):
    """-> (model tensors, speaker-encoder tensors with the prefix stripped)."""
    own = {k: v for k, v in sd.items() if not k.startswith(prefix)}
    spk = {k[len(prefix):]: v for k, v in sd.items() if k.startswith(prefix)}
    return own, spk
