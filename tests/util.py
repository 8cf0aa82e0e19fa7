"""Shared helpers for the parity tests (test infrastructure)."""
import json
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")


def load_golden(name):
    d = np.load(os.path.join(GOLDEN, name + ".npz"), allow_pickle=False)
    meta = json.loads(str(d["meta"]))
    arrays = {k: d[k] for k in d.files if k != "meta"}
    return meta, arrays


def golden_state_dict(meta):
    """Rebuild the reference state_dict of a golden fixture from the weight recipe."""
    from oracle.weights_recipe import build_state_dict
    return build_state_dict([(n, tuple(s)) for n, s in meta["names_shapes"]], meta["seed"])


def golden_input(arrays):
    from oracle.weights_recipe import recipe_input
    return recipe_input(tuple(int(v) for v in arrays["x_shape"]), int(arrays["x_seed"].item()))


def rel_l2(a, b):
    a = np.asarray(a, dtype=np.float64).ravel()
    b = np.asarray(b, dtype=np.float64).ravel()
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))
