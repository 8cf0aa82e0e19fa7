"""TEST INFRASTRUCTURE ONLY -- deterministic, name-keyed synthetic weights.

There are no pretrained checkpoints (the reference's blobs are missing and there is no
network), so every parity test runs on synthetic weights.  A tensor's values depend only
on (seed, name, shape): numpy PCG64 seeded with crc32(name) ^ seed.  This lets a golden
file store just the (name, shape) list and the outputs, and lets the GPU box rebuild the
identical state_dict without torch RNG or module-construction-order dependence.

The recipe itself lives in rave_b200/synth.py (pure numpy, shared with the benchmarks' synthetic
workloads); this module re-exports it so test infrastructure keeps one import path.
"""
from rave_b200.synth import _rng, build_state_dict, recipe_input, recipe_tensor  # noqa: F401
