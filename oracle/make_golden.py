"""TEST INFRASTRUCTURE ONLY -- writes tests/golden/*.npz from the REAL reference modules.

Run in the build container (needs /root/reference):  python -m oracle.make_golden
Each fixture stores: the flat config, the (name, shape) list of the reference
state_dict (weights are rebuilt from oracle/weights_recipe.py, so none are stored), the
input recipe and the reference outputs.  The reference holds no golden vectors of its
own, so these are outputs of the reference code itself run here.
"""
import json
import os
import sys
from unittest import mock

import numpy as np
import torch

from . import configs as C
from . import ref_loader as R
from .weights_recipe import recipe_input

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
SEED = 0


def _names_shapes(model):
    return [(k, list(v.shape)) for k, v in model.state_dict().items()]


def _save(name, cfg, model, arrays):
    os.makedirs(OUT, exist_ok=True)
    meta = dict(cfg=cfg, names_shapes=_names_shapes(model), seed=SEED)
    np.savez_compressed(os.path.join(OUT, name + ".npz"), meta=json.dumps(meta),
                        **{k: np.ascontiguousarray(v) for k, v in arrays.items()})
    print("wrote", name, {k: tuple(np.shape(v)) for k, v in arrays.items()})


def _sd(model):
    return {k: v.detach().clone() for k, v in model.state_dict().items()}


@torch.no_grad()
def offline_and_stream(name, cfg, B, T, chunk=None, x_seed=1234, with_stream=True):
    model = R.build_reference(cfg)
    R.fill_deterministic(model, SEED)
    x = torch.from_numpy(recipe_input((B, 1, T), x_seed))
    arrays = dict(x_seed=np.int64(x_seed), x_shape=np.array([B, 1, T]))
    noise = None
    if cfg.get("noise", False):
        L = T // cfg["n_band"]
        noise = torch.from_numpy(recipe_input((B, L // 8, cfg["n_band"], 8), 7, uniform=True))
        arrays["noise_seed"] = np.int64(7)
        with mock.patch("torch.rand_like", side_effect=lambda t: (noise + 1) / 2):
            z = model.encode(x)
            y = model.decode(z)
    else:
        z = model.encode(x)
        y = model.decode(z)
    arrays.update(z=z.numpy(), y=y.numpy())
    if with_stream:
        ms = R.build_reference(cfg, streaming=True)
        ms.load_state_dict(_sd(model), strict=False)
        zs, ys = [], []
        for c in range(T // chunk):
            xc = x[..., c * chunk:(c + 1) * chunk]
            zc = ms.encode(xc)
            zs.append(zc)
            if noise is not None:
                # the chunk's slice of the injected draw (blocks.py:287 rand_like inside the cached module)
                q = chunk // cfg["n_band"] // 8
                nc = noise[:, c * q:(c + 1) * q]
                with mock.patch("torch.rand_like", side_effect=lambda t: (nc + 1) / 2):
                    ys.append(ms.decode(zc))
            else:
                ys.append(ms.decode(zc))
        arrays.update(chunk=np.int64(chunk), z_stream=torch.cat(zs, -1).numpy(),
                      y_stream=torch.cat(ys, -1).numpy())
    if cfg.get("encoder_type") == "discrete":
        zq = torch.from_numpy(recipe_input((4, cfg["latent_size"], 64), 99, scale=1.0))
        idx = model.encoder.rvq.encode(zq)
        acc = model.encoder.rvq.decode(idx)
        arrays.update(rvq_z_seed=np.int64(99), rvq_idx=idx.numpy(), rvq_dec=acc.numpy())
    _save(name, cfg, model, arrays)


def pqmf_fixture():
    pq = R.load("pqmf").CachedPQMF(attenuation=100, n_band=16)
    os.makedirs(OUT, exist_ok=True)
    np.savez_compressed(os.path.join(OUT, "pqmf_bank.npz"), h=pq.h.numpy(), hk=pq.hk.numpy(),
                        wf=pq.forward_conv.weight.detach().numpy(),
                        wi=pq.inverse_conv.weight.detach().numpy())
    print("wrote pqmf_bank", tuple(pq.hk.shape))


@torch.no_grad()
def speaker_fixture():
    """SpeakerRAVE (CombinedRave.py:200-328) on recipe weights: bands -> embedding, and the
    model.py:246-247 chain utterance -> CachedPQMF -> embedding."""
    cc = R.cc
    cc.use_cached_conv(False)
    cc.set_padding_mode("centered")
    spk = R.load("CombinedRave").SpeakerRAVE().eval()
    R.fill_deterministic(spk, SEED)
    pq = R.load("pqmf").CachedPQMF(attenuation=100, n_band=16)
    bands = torch.from_numpy(recipe_input((2, 16, 2048), 41, scale=1.0))
    audio = torch.from_numpy(recipe_input((1, 1, 32768), 42, scale=0.3))
    _save("speaker", dict(kind="speaker"), spk,
          dict(bands_seed=np.int64(41), bands_shape=np.array([2, 16, 2048]), emb=spk(bands).numpy(),
               audio_seed=np.int64(42), audio_shape=np.array([1, 1, 32768]),
               emb_audio=spk(pq(audio)).numpy()))


@torch.no_grad()
def resampler_fixture():
    """rave/resampler.py on the ratios of the reference's tests/test_resampler.py: offline for 2 and 3,
    cached_conv streaming (chunk by chunk) for 2."""
    res = R.load("resampler")
    cc = R.cc
    arrays = dict(x_seed=np.int64(51), x_shape=np.array([2, 1, 6144]), chunk=np.int64(1536))
    x = torch.from_numpy(recipe_input((2, 1, 6144), 51, scale=0.5))
    try:
        for ratio, (tsr, msr) in ((2, (44100, 22050)), (3, (48000, 16000))):
            cc.use_cached_conv(False)
            r = res.Resampler(tsr, msr)
            y = r.to_model_sampling_rate(x)
            arrays[f"down{ratio}"] = y.numpy()
            arrays[f"up{ratio}"] = r.from_model_sampling_rate(y).numpy()
            arrays[f"w_down{ratio}"] = r.downsample.weight.detach().numpy()
            arrays[f"w_up{ratio}"] = r.upsample.weight.detach().numpy()
        cc.use_cached_conv(True)
        r = res.Resampler(44100, 22050)
        ys, zs = [], []
        for c in range(4):
            y = r.to_model_sampling_rate(x[..., c * 1536:(c + 1) * 1536])
            ys.append(y)
            zs.append(r.from_model_sampling_rate(y))
        arrays["down2_stream"] = torch.cat(ys, -1).numpy()
        arrays["up2_stream"] = torch.cat(zs, -1).numpy()
    finally:
        cc.use_cached_conv(False)
    os.makedirs(OUT, exist_ok=True)
    np.savez_compressed(os.path.join(OUT, "resampler.npz"), **arrays)
    print("wrote resampler", {k: tuple(np.shape(v)) for k, v in arrays.items()})


@torch.no_grad()
def pitch_fixture():
    """rave/pitch_utils.py get_pitch / estimate / get_f0_norm (FCPE replaced by the YIN output) on the
    deterministic signals of tests/util.py:pitch_signals."""
    sys.path.insert(0, os.path.dirname(OUT))
    from tests.util import pitch_signals
    pu = R.load("pitch_utils")
    x = torch.from_numpy(pitch_signals())
    f0 = pu.get_pitch(x, 1024)
    f0_wide = pu.estimate(x, 44100, 50., 800., 0.01, 0.15)
    pu.extract_utterance_fcpe = lambda y, sr, fl: f0.unsqueeze(-1).clone()
    oh, ln = pu.get_f0_norm(x, torch.zeros(4, 1), torch.ones(4, 1), 44100, 1024)
    os.makedirs(OUT, exist_ok=True)
    np.savez_compressed(os.path.join(OUT, "pitch.npz"), f0=f0.numpy(), f0_wide=f0_wide.numpy(),
                        one_hot_idx=oh.argmax(-1).numpy().astype(np.int16), log_f0_norm=ln.numpy())
    print("wrote pitch", tuple(f0.shape), tuple(f0_wide.shape), tuple(oh.shape))


def main():
    if not R.available():
        sys.exit("reference tree not available; golden files can only be made in the build container")
    torch.manual_seed(0)
    only = set(sys.argv[1:])
    if only:                      # python -m oracle.make_golden v2_c64 discrete_c96 ...: just these
        for name, args in SIZED.items():
            if name in only:
                offline_and_stream(name, *args)
        return
    pqmf_fixture()
    speaker_fixture()
    resampler_fixture()
    pitch_fixture()
    offline_and_stream("v2_c8", C.v2(8), 2, 16384, 2048)
    offline_and_stream("v2causal_c8", C.v2(8, causal=True), 2, 16384, 2048)
    offline_and_stream("v3_c8", C.v3(8), 2, 16384, 2048)
    offline_and_stream("v3causal_c8", C.v3(8, causal=True), 1, 8192, 2048)
    offline_and_stream("discrete_c8", C.discrete(8), 1, 8192, 2048)
    offline_and_stream("v2noise_c8", C.v2(8, noise=True), 2, 8192, 2048)
    offline_and_stream("v1_c8", C.v1(8), 1, 8192, 2048)
    offline_and_stream("raspberry_c16", C.raspberry(16), 1, 8192, 1024)
    offline_and_stream("raspberry_causal_c16", C.raspberry(16, causal=True), 1, 8192, 1024)
    for name, args in SIZED.items():
        offline_and_stream(name, *args)


# The configurations BASELINE.json quotes numbers on, at their REAL channel widths (shortened in time):
# configs[0]/[1] v2 CAPACITY 64, configs[2] causal v2 in 2048-sample chunks, configs[3] discrete
# (discrete.gin:13-19: CAPACITY 96, LATENT_SIZE 128), configs[4] v3 (Snake + AdaIN) CAPACITY 64.
SIZED = {
    "v2_c64": (C.v2(64), 1, 32768, 2048),
    "v2causal_c64": (C.v2(64, causal=True), 2, 8192, 2048),
    "discrete_c96": (C.discrete(96), 1, 8192, 2048),
    "v3_c64": (C.v3(64), 1, 8192, 2048),
    "v3causal_c64": (C.v3(64, causal=True), 1, 8192, 2048),
}


if __name__ == "__main__":
    main()
