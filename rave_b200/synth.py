"""Deterministic, name-keyed synthetic weights and inputs (no pretrained checkpoints exist and there is
no network): a tensor's values depend only on (seed, name, shape) -- numpy PCG64 seeded with
crc32(name) ^ seed -- so a fixture stores just the (name, shape) list and any machine rebuilds the
identical state_dict.  Used by bench.py / bench_configs.py for their synthetic workloads and, through
oracle/weights_recipe.py, by the tests and the golden-vector generator.  Pure numpy: no device work.
"""
import zlib

import numpy as np


def _rng(name: str, seed: int):
    return np.random.Generator(np.random.PCG64((zlib.crc32(name.encode()) ^ (seed * 0x9E3779B1)) & 0xFFFFFFFF))


def recipe_tensor(name: str, shape, seed: int = 0):
    """Return float32 ndarray for tensor `name`, or None to leave the tensor untouched."""
    leaf = name.rsplit(".", 1)[-1]
    g = _rng(name, seed)
    n = lambda: g.standard_normal(shape)  # noqa: E731
    if name.startswith("pqmf.") or ".pqmf." in name:
        return None                      # designed filters stay as designed
    if leaf in ("pad", "cache") or ".paddings." in name or ".downsampling_delay." in name:
        return None                      # streaming state starts at zero
    if leaf in ("target_size", "warmed_up", "enabled", "gru_state", "inited",
                "cluster_size", "embed_avg", "num_batches_tracked", "learn_x", "learn_y"):
        return None
    if leaf == "weight_g":
        v = 0.7 + 0.4 * g.random(shape)
    elif leaf in ("weight_v", "weight"):
        if len(shape) == 1:              # BatchNorm affine weight
            v = 1.0 + 0.1 * n()
        else:
            fan = int(np.prod(shape[1:]))
            v = n() / np.sqrt(max(fan, 1))
    elif leaf == "bias":
        v = 0.05 * n()
    elif leaf == "alpha":                # Snake
        v = 1.0 + 0.1 * n()
    elif leaf in ("mean_x", "mean_y", "running_mean"):
        v = 0.1 * n()
    elif leaf in ("std_x", "std_y"):
        v = 1.0 + 0.1 * np.abs(n())
    elif leaf == "running_var":
        v = 0.5 + g.random(shape)
    elif leaf in ("num_update_x", "num_update_y"):
        v = np.ones(shape)
    elif leaf == "embed":                # RVQ codebooks
        v = n()
    elif leaf == "speaker":
        v = n()
    else:
        return None
    return np.ascontiguousarray(v, dtype=np.float32)


def recipe_input(shape, seed: int, scale: float = 0.1, uniform: bool = False):
    g = np.random.Generator(np.random.PCG64(seed))
    if uniform:
        return np.ascontiguousarray(g.random(shape) * 2 - 1, dtype=np.float32)
    return np.ascontiguousarray(scale * g.standard_normal(shape), dtype=np.float32)


def build_state_dict(names_shapes, seed: int = 0, base=None):
    """{name: ndarray} for every (name, shape) the recipe covers; `base` supplies the rest."""
    out = dict(base or {})
    for name, shape in names_shapes:
        v = recipe_tensor(name, tuple(shape), seed)
        if v is not None:
            out[name] = v
    return out


def golden_names_shapes(name, root=None):
    """(meta, [(tensor name, shape)]) of tests/golden/<name>.npz -- the list only, no reference outputs."""
    import json
    import os
    root = root or os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
    d = np.load(os.path.join(root, name + ".npz"), allow_pickle=False)
    meta = json.loads(str(d["meta"]))
    return meta, [(n, tuple(s)) for n, s in meta["names_shapes"]]


def golden_state_dict(name):
    meta, ns = golden_names_shapes(name)
    return meta, build_state_dict(ns, meta["seed"])
