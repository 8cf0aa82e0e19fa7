"""TEST INFRASTRUCTURE ONLY -- flat architecture dicts for the config stacks under test.

Each dict is the result of resolving a gin stack of the reference by hand (file:line
given), so the oracle does not depend on the product's gin reader.  `capacity` is left
as a parameter: golden fixtures use small capacities (like the reference's own
tests/test_configs.py:58 `CAPACITY=2`), the benchmark uses the real one.
"""

V2_DILATIONS = [[1, 3, 9], [1, 3, 9], [1, 3, 9], [1, 3]]      # v2.gin:16-21


def v2(capacity=64, causal=False, **over):
    """rave/configs/v2.gin (+ v1.gin it includes): this fork's non-variational v2."""
    cfg = dict(
        arch="v2", n_band=16, attenuation=100,                 # v1.gin:15,37-39
        enc_data_size=6,                                       # v2.gin:40, model.py:613
        capacity=capacity, ratios=[4, 4, 2, 2],                # v2.gin:22-23
        latent_size=64,                                        # v1.gin:16
        speaker_size=256, dec_latent_size=64 + 0 + 256,        # v2.gin:24-36, core.py:78-79
        kernel_size=3, dilations=V2_DILATIONS,                 # v2.gin:15-21
        activation="leaky_relu", adain=False, noise=False,
        encoder_type="variational", causal=causal, sr=44100)
    cfg.update(over)
    return cfg


def v3(capacity=64, causal=False, **over):
    """configs/v3.gin = v2 + adain.gin + snake.gin."""
    return v2(capacity, causal, activation="snake", adain=True, **over)


def discrete(capacity=96, causal=False, **over):
    """configs/discrete.gin:13-19,26-42."""
    cfg = v2(capacity, causal, latent_size=128, dec_latent_size=128 + 0 + 256,
             encoder_type="discrete", num_quantizers=16, codebook_size=1024)
    cfg.update(over)
    return cfg


def v1(capacity=64, causal=False, **over):
    """configs/v1.gin:13-17,44-72; decoder latent 320 hard-wired (v1.gin:59)."""
    cfg = dict(
        arch="v1", n_band=16, attenuation=100, enc_data_size=6, capacity=capacity,
        ratios=[4, 2, 2, 2], latent_size=64, speaker_size=256, dec_latent_size=320,
        activation="leaky_relu", adain=False, noise=False, encoder_type="variational",
        causal=causal, sr=44100)
    cfg.update(over)
    return cfg


def raspberry(capacity=16, causal=False, **over):
    """configs/raspberry.gin -> onnx.gin -> v1.gin, use_noise False (onnx.gin:12)."""
    return v1(capacity, causal, **over)


NAMED = {"v2": v2, "v3": v3, "discrete": discrete, "v1": v1, "raspberry": raspberry}


def total_ratio(cfg):
    r = cfg["n_band"]
    for x in cfg["ratios"]:
        r *= x
    return r
