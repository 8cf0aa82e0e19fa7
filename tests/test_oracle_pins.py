"""Pins the CPU oracle: (1) against the committed golden fixtures (outputs of the real
reference modules, made by oracle/make_golden.py); (2) where /root/reference exists,
against the real modules live, including the reference's own tests/test_residual.py
identities that pin the restated cached_conv."""
import os

import numpy as np
import pytest
import torch

from oracle import configs as C
from oracle import rave_oracle as O
from oracle import ref_loader as R
from oracle.model import OracleRAVE
from oracle.weights_recipe import recipe_input

from .util import golden_input, golden_state_dict, load_golden, rel_l2

FIXTURES = ["v2_c8", "v2causal_c8", "v3_c8", "v3causal_c8", "discrete_c8", "v2noise_c8",
            "v1_c8", "raspberry_c16", "raspberry_causal_c16",
            # BASELINE.json configs at their real channel widths (oracle/make_golden.py SIZED)
            "v2_c64", "v2causal_c64", "discrete_c96", "v3_c64", "v3causal_c64"]
needs_ref = pytest.mark.skipif(not R.available(), reason="reference tree not present")


def test_pqmf_design_matches_reference_bank():
    d = np.load(os.path.join(os.path.dirname(__file__), "golden", "pqmf_bank.npz"))
    h, hk = O.pqmf_bank(100, 16)
    wf, wi = O.pqmf_conv_weights(hk)
    assert hk.shape == (16, 512) and wf.shape == (16, 1, 513) and wi.shape == (16, 16, 33)
    assert rel_l2(h.numpy(), d["h"]) < 1e-6
    assert rel_l2(hk.numpy(), d["hk"]) < 1e-6
    assert rel_l2(wf.numpy(), d["wf"]) < 1e-6
    assert rel_l2(wi.numpy(), d["wi"]) < 1e-6


@pytest.mark.parametrize("name", FIXTURES)
def test_oracle_offline_matches_golden(name):
    meta, a = load_golden(name)
    cfg = meta["cfg"]
    o = OracleRAVE(cfg, golden_state_dict(meta))
    x = torch.from_numpy(golden_input(a))
    z = o.encode(x)
    assert rel_l2(z.numpy(), a["z"]) < 2e-5
    noise = None
    if cfg.get("noise"):
        B, _, T = x.shape
        noise = torch.from_numpy(recipe_input((B, T // 16 // 8, 16, 8), int(a["noise_seed"].item()), uniform=True))
    y = o.decode(torch.from_numpy(a["z"]), noise)
    assert y.shape == x.shape                      # reference tests/test_configs.py:68
    assert rel_l2(y.numpy(), a["y"]) < 5e-5


@pytest.mark.parametrize("name", FIXTURES)
def test_oracle_streaming_matches_golden(name):
    meta, a = load_golden(name)
    o = OracleRAVE(meta["cfg"], golden_state_dict(meta), streaming=True)
    x = torch.from_numpy(golden_input(a))
    n = int(a["chunk"].item())
    noise = None
    if meta["cfg"].get("noise"):
        B, _, T = x.shape
        noise = torch.from_numpy(recipe_input((B, T // 16 // 8, 16, 8), int(a["noise_seed"].item()), uniform=True))
    zs, ys = [], []
    for c in range(x.shape[-1] // n):
        z = o.encode(x[..., c * n:(c + 1) * n])
        zs.append(z)
        q = n // 16 // 8
        ys.append(o.decode(z, None if noise is None else noise[:, c * q:(c + 1) * q]))
    assert rel_l2(torch.cat(zs, -1).numpy(), a["z_stream"]) < 2e-5
    assert rel_l2(torch.cat(ys, -1).numpy(), a["y_stream"]) < 5e-5


def test_causal_stream_equals_offline_after_delay():
    """causal v2 -- encoder stream == offline; decoder stream == offline
    delayed by 58 sub-band samples = 928 audio samples, past the start-up transient."""
    meta, a = load_golden("v2causal_c8")
    zs, ys, z, y = a["z_stream"], a["y_stream"], a["z"], a["y"]
    assert rel_l2(zs, z) < 1e-5
    d = 928
    assert rel_l2(ys[..., 14336:], y[..., 14336 - d:-d]) < 1e-5   # transient: 4e-2 @4096, 1e-4 @12288


@pytest.mark.parametrize("name", ["discrete_c8", "discrete_c96"])
def test_rvq_oracle_matches_golden_bit_exact(name):
    meta, a = load_golden(name)
    o = OracleRAVE(meta["cfg"], golden_state_dict(meta))
    zq = torch.from_numpy(recipe_input((4, 128, 64), int(a["rvq_z_seed"].item()), scale=1.0))
    idx = o.rvq_encode(zq)
    assert idx.dtype == torch.int64 and tuple(idx.shape) == (4, 16, 64)
    assert np.array_equal(idx.numpy(), a["rvq_idx"])
    dec = o.rvq_decode(torch.from_numpy(a["rvq_idx"]))
    assert rel_l2(dec.numpy(), a["rvq_dec"]) < 1e-6


# ---------------------------------------------------------------- live reference (container only)

@needs_ref
@pytest.mark.reference
@pytest.mark.parametrize("ratio", [2, 4, 8])
def test_ref_upsample_layer_stream_identity(ratio):
    """reference tests/test_residual.py:98-122 on the restated cached_conv."""
    from oracle import cached_conv_restated as cc
    blocks = R.load("blocks")
    torch.manual_seed(0)
    x = torch.randn(1, 16, 32)
    cc.use_cached_conv(False)
    reg = blocks.UpsampleLayer(16, 16, ratio)
    cc.use_cached_conv(True)
    stream = blocks.UpsampleLayer(16, 16, ratio)
    cc.use_cached_conv(False)
    for p1, p2 in zip(reg.parameters(), stream.parameters()):
        p2.data.copy_(p1.data)
    delay = stream.cumulative_delay
    with torch.no_grad():
        y_r, y_s = reg(x), stream(x)
    if delay:
        y_r, y_s = y_r[..., delay:-delay], y_s[..., 2 * delay:]
    assert torch.allclose(y_r, y_s, 1e-3, 1e-4)


@needs_ref
@pytest.mark.reference
@pytest.mark.parametrize("k,dil", [(1, [1, 1]), (3, [1, 1]), (1, [3, 1]), (3, [3, 1])])
def test_ref_residual_layer_stream_identity(k, dil):
    """reference tests/test_residual.py:71-95."""
    from oracle import cached_conv_restated as cc
    blocks = R.load("blocks")
    torch.manual_seed(0)
    x = torch.randn(1, 16, 32)
    cc.use_cached_conv(False)
    reg = blocks.ResidualLayer(16, k, dil)
    cc.use_cached_conv(True)
    stream = blocks.ResidualLayer(16, k, dil)
    cc.use_cached_conv(False)
    for p1, p2 in zip(reg.parameters(), stream.parameters()):
        p2.data.copy_(p1.data)
    delay = stream.cumulative_delay
    with torch.no_grad():
        y_r, y_s = reg(x), stream(x)
    if delay:
        y_r, y_s = y_r[..., delay:-delay], y_s[..., 2 * delay:]
    assert torch.allclose(y_r, y_s, 1e-4, 1e-4)


@needs_ref
@pytest.mark.reference
def test_ref_residual_stack_stream_identity():
    """reference tests/test_residual.py:37-68."""
    from oracle import cached_conv_restated as cc
    blocks = R.load("blocks")
    torch.manual_seed(0)
    x = torch.randn(1, 16, 32)
    kw = dict(dim=16, kernel_sizes=[3], dilations_list=[[1, 1], [3, 1], [5, 1]])
    cc.use_cached_conv(False)
    reg = blocks.ResidualStack(**kw)
    cc.use_cached_conv(True)
    stream = blocks.ResidualStack(**kw)
    cc.use_cached_conv(False)
    for p1, p2 in zip(reg.parameters(), stream.parameters()):
        p2.data.copy_(p1.data)
    delay = stream.cumulative_delay
    with torch.no_grad():
        y_r, y_s = reg(x), stream(x)
    y_r, y_s = y_r[..., delay:-delay], y_s[..., 2 * delay:]
    assert torch.allclose(y_r, y_s, 1e-4, 1e-4)


@needs_ref
@pytest.mark.reference
@pytest.mark.parametrize("name", ["v2", "v3", "discrete", "v1"])
def test_oracle_matches_live_reference(name):
    cfg = C.NAMED[name](4)
    m = R.build_reference(cfg)
    R.fill_deterministic(m, 3)
    o = OracleRAVE(cfg, {k: v.detach().clone() for k, v in m.state_dict().items()})
    x = torch.from_numpy(recipe_input((1, 1, 4096), 11))
    with torch.no_grad():
        z = m.encode(x)
        y = m.decode(z)
    assert rel_l2(o.encode(x).numpy(), z.numpy()) < 2e-5
    assert rel_l2(o.decode(z).numpy(), y.numpy()) < 5e-5


def _ref_update_adain(m, blocks, learn_source=False, learn_target=False, reset_source=False, reset_target=False):
    """scripts/export.py:248-265 (ScriptedRAVE.update_adain) applied to the live reference modules."""
    for mod in m.modules():
        if isinstance(mod, blocks.AdaptiveInstanceNormalization):
            mod.learn_x.zero_()
            mod.learn_y.zero_()
            if learn_target:
                mod.learn_y.add_(1)
            if learn_source:
                mod.learn_x.add_(1)
            if reset_target:
                mod.reset_y()
            if reset_source:
                mod.reset_x()


@needs_ref
@pytest.mark.reference
def test_oracle_adain_statistic_learning_matches_live_reference():
    """AdaIN learn_y / learn_x running statistics and the transfer that follows (blocks.py:893-919)."""
    cfg = C.NAMED["v3"](4)
    m = R.build_reference(cfg)
    R.fill_deterministic(m, 5)
    blocks = R.load("blocks")
    o = OracleRAVE(cfg, {k: v.detach().clone() for k, v in m.state_dict().items()})
    xs = [torch.from_numpy(recipe_input((2, 1, 4096), 20 + i)) for i in range(5)]
    plan = [dict(reset_source=True, reset_target=True, learn_target=True), dict(learn_target=True),
            dict(learn_source=True), dict(learn_source=True), dict()]
    for x, kw in zip(xs, plan):
        _ref_update_adain(m, blocks, **kw)
        o.update_adain(**kw)
        with torch.no_grad():
            y = m.decode(m.encode(x))
        assert rel_l2(o.forward(x).numpy(), y.numpy()) < 3e-4      # random-init Snake net: fp32 noise grows
    for k, v in m.state_dict().items():
        if any(t in k for t in ("mean_x", "std_x", "mean_y", "std_y", "num_update")):
            assert torch.allclose(o.sd[k], v, rtol=1e-3, atol=1e-3), k


# ----------------------------------------------------------------------------- speaker encoder
def _speaker_golden():
    meta, a = load_golden("speaker")
    sd = {k: torch.from_numpy(v) for k, v in golden_state_dict(meta).items()}
    bands = torch.from_numpy(recipe_input(tuple(int(v) for v in a["bands_shape"]), int(a["bands_seed"].item()), scale=1.0))
    audio = torch.from_numpy(recipe_input(tuple(int(v) for v in a["audio_shape"]), int(a["audio_seed"].item()), scale=0.3))
    return sd, a, bands, audio


def test_speaker_oracle_matches_golden():
    """oracle/speaker_oracle.py vs the outputs of the real SpeakerRAVE / CachedPQMF (make_golden)."""
    from oracle import speaker_oracle as S
    sd, a, bands, audio = _speaker_golden()
    assert rel_l2(S.speaker_embed(sd, bands).numpy(), a["emb"]) < 1e-5
    _, hk = O.pqmf_bank(100, 16)
    wf, _ = O.pqmf_conv_weights(hk)
    assert rel_l2(S.speaker_from_audio(sd, wf, audio).numpy(), a["emb_audio"]) < 1e-5


@needs_ref
@pytest.mark.reference
def test_speaker_oracle_matches_live_reference():
    from oracle import speaker_oracle as S
    R.cc.use_cached_conv(False)
    spk = R.load("CombinedRave").SpeakerRAVE().eval()
    R.fill_deterministic(spk, 3)
    x = torch.from_numpy(recipe_input((3, 16, 1024), 8, scale=1.0))
    with torch.no_grad():
        ref = spk(x)
    got = S.speaker_embed(spk.state_dict(), x)
    assert ref.shape == (3, 256)
    assert rel_l2(got.numpy(), ref.numpy()) < 1e-5


# ----------------------------------------------------------------------------- resampler
def _resampler_golden():
    d = np.load(os.path.join(os.path.dirname(__file__), "golden", "resampler.npz"))
    x = torch.from_numpy(recipe_input(tuple(int(v) for v in d["x_shape"]), int(d["x_seed"].item()), scale=0.5))
    return d, x


def test_resampler_oracle_matches_golden():
    """oracle/resampler_oracle.py vs outputs of the real rave.resampler.Resampler (make_golden)."""
    from oracle.resampler_oracle import ResamplerOracle, design
    d, x = _resampler_golden()
    for ratio in (2, 3):
        down, up = design(ratio)
        assert np.array_equal(down.numpy(), d[f"w_down{ratio}"].reshape(-1))
        assert np.array_equal(up.numpy(), d[f"w_up{ratio}"].reshape(ratio, -1))
        o = ResamplerOracle(ratio)
        y = o.to_model_sampling_rate(x)
        assert rel_l2(y.numpy(), d[f"down{ratio}"]) < 1e-6
        assert rel_l2(o.from_model_sampling_rate(y).numpy(), d[f"up{ratio}"]) < 1e-6
    o = ResamplerOracle(2, streaming=True)
    c = int(d["chunk"].item())
    ys = [o.to_model_sampling_rate(x[..., i:i + c]) for i in range(0, x.shape[-1], c)]
    zs = [o.from_model_sampling_rate(y) for y in ys]
    assert rel_l2(torch.cat(ys, -1).numpy(), d["down2_stream"]) < 1e-6
    assert rel_l2(torch.cat(zs, -1).numpy(), d["up2_stream"]) < 1e-6


@needs_ref
@pytest.mark.reference
@pytest.mark.parametrize("target_sr,model_sr", [(44100, 22050), (48000, 16000)])   # tests/test_resampler.py:8
def test_resampler_oracle_matches_live_reference(target_sr, model_sr):
    from oracle.resampler_oracle import ResamplerOracle
    R.load("pqmf")
    res = R.load("resampler")
    x = torch.from_numpy(recipe_input((1, 1, 2 ** 12 * 3), 9, scale=1.0))
    for streaming in (False, True):
        R.cc.use_cached_conv(streaming)
        try:
            try:
                ref = res.Resampler(target_sr, model_sr)
            except ValueError:
                assert streaming and (target_sr // model_sr) % 2
                continue
            o = ResamplerOracle(target_sr // model_sr, streaming)
            for _ in range(2):
                with torch.no_grad():
                    y = ref.to_model_sampling_rate(x)
                    z = ref.from_model_sampling_rate(y)
                yo = o.to_model_sampling_rate(x)
                assert z.shape == x.shape
                assert torch.equal(y, yo) and torch.equal(z, o.from_model_sampling_rate(yo))
        finally:
            R.cc.use_cached_conv(False)


# ----------------------------------------------------------------------------- pitch front-end
def test_pitch_oracle_matches_golden():
    """oracle/pitch_oracle.py vs outputs of the real rave/pitch_utils.py functions (make_golden)."""
    from oracle import pitch_oracle as P
    from .util import pitch_signals
    d = np.load(os.path.join(os.path.dirname(__file__), "golden", "pitch.npz"))
    x = torch.from_numpy(pitch_signals())
    f0 = P.get_pitch(x, 1024)
    assert np.array_equal(f0.numpy(), d["f0"])
    assert np.array_equal(P.estimate(x, 44100, 50., 800., 0.01, 0.15).numpy(), d["f0_wide"])
    oh, ln = P.f0_norm_one_hot(f0)
    assert np.array_equal(oh.argmax(-1).numpy(), d["one_hot_idx"])
    assert np.array_equal(np.nan_to_num(ln.numpy(), nan=-9.), np.nan_to_num(d["log_f0_norm"], nan=-9.))
    assert (d["f0"][0] > 0).all() and (d["f0"][1] == 0).all()       # voiced tone / white noise


@needs_ref
@pytest.mark.reference
def test_pitch_oracle_matches_live_reference():
    from oracle import pitch_oracle as P
    from .util import pitch_signals
    pu = R.load("pitch_utils")
    x = torch.from_numpy(pitch_signals(32768))
    assert torch.equal(pu.get_pitch(x, 512), P.get_pitch(x, 512))
    assert torch.equal(pu.estimate(x, 44100, 30., 1000., 0.02, 0.2), P.estimate(x, 44100, 30., 1000., 0.02, 0.2))
    f0 = P.get_pitch(x, 1024)
    pu.extract_utterance_fcpe = lambda y, sr, fl: f0.unsqueeze(-1).clone()
    oh, ln = pu.get_f0_norm(x, torch.zeros(4, 1), torch.ones(4, 1), 44100, 1024)
    oh2, ln2 = P.f0_norm_one_hot(f0)
    assert torch.equal(oh, oh2) and torch.equal(torch.nan_to_num(ln, nan=-9.), torch.nan_to_num(ln2, nan=-9.))


# ----------------------------------------------------------------------------- latent processing
@needs_ref
@pytest.mark.reference
def test_latent_oracle_angles_match_live_reference():
    from oracle import latent_oracle as L
    blocks = R.load("blocks")
    x = torch.from_numpy(recipe_input((3, 16, 40), 21, scale=1.0))
    x = x / x.norm(dim=1, keepdim=True)
    a_ref = blocks.unit_norm_vector_to_angles(x.clone())
    assert torch.equal(a_ref, L.unit_norm_vector_to_angles(x.clone()))
    assert torch.equal(blocks.angles_to_unit_norm_vector(a_ref.clone()), L.angles_to_unit_norm_vector(a_ref.clone()))
    assert rel_l2(L.angles_to_unit_norm_vector(a_ref.clone()).numpy(), x.numpy()) < 1e-5     # round trip


@needs_ref
def test_latent_oracle_pca_matches_live_export_methods():
    """Pins oracle/latent_oracle.py's two PCA methods against the reference's OWN method bodies
    (scripts/export.py:484-499, cut out of the file's syntax tree by ref_loader.load_export_latent_classes),
    with the reference's torch.randn draw reproduced by seeding."""
    from types import SimpleNamespace
    from oracle import latent_oracle as L
    cls = R.load_export_latent_classes()["VariationalScriptedRAVE"]
    g = torch.Generator().manual_seed(3)
    C_, n = 32, 8
    q, _ = torch.linalg.qr(torch.randn(C_, C_, generator=g))
    mean = 0.2 * torch.randn(C_, generator=g)
    m = cls()
    m.latent_pca, m.latent_mean, m.latent_size, m.full_latent_size = q, mean, n, C_
    m.encoder = SimpleNamespace(reparametrize=lambda z: (z, None))       # blocks.py:720-729 returns (z, kl)
    z = torch.randn(3, C_, 11, generator=g)
    want = m.post_process_latent(z)
    got = L.post_process_latent(z, q, mean, n)
    assert want.shape == (3, n, 11) and torch.equal(got, want)
    torch.manual_seed(77)
    back = m.pre_process_latent(want)
    torch.manual_seed(77)
    noise = torch.randn(3, C_ - n, 11)
    assert torch.equal(L.pre_process_latent(want, noise, q, mean), back)
    # discrete wrapper (export.py:503-517): clamp + .long() in front of rvq.decode, encode -> float
    dcls = R.load_export_latent_classes()["DiscreteScriptedRAVE"]
    quant = R.load("quantization")
    torch.manual_seed(0)
    rvq = quant.ResidualVectorQuantization(num_quantizers=4, dim=16, codebook_size=32, kmeans_init=False)
    d = dcls()
    d.encoder = SimpleNamespace(rvq=rvq, noise_augmentation=0)
    zz = torch.randn(2, 16, 5)
    idx = d.post_process_latent(zz)
    assert idx.dtype == torch.float32 and torch.equal(idx.long(), rvq.encode(zz))
    assert torch.equal(d.pre_process_latent(idx + 0.25), rvq.decode(idx.long()))


@needs_ref
def test_plain_pqmf_alias_matches_live_class():
    """pqmf.PQMF (pqmf.py:179-231, bound by no shipped config): analysis identical to CachedPQMF, synthesis =
    CachedPQMF.inverse advanced by one sub-band row (16 samples).  Pins the oracle's `pqmf="PQMF"` branch."""
    pm = R.load("pqmf")
    pq = pm.PQMF(100, 16)
    o = OracleRAVE(dict(C.v2(8), pqmf="PQMF"), {"speaker": torch.zeros(1, 256, 1)})
    x = torch.from_numpy(recipe_input((2, 1, 4096), 3))
    yb = torch.from_numpy(recipe_input((2, 16, 256), 4, scale=1.0))
    assert rel_l2(o.pqmf_forward(x).numpy(), pq(x).numpy()) < 1e-6
    assert rel_l2(o.pqmf_inverse(yb).numpy(), pq.inverse(yb).numpy()) < 1e-5


def test_latent_oracle_pca_round_trip():
    from oracle import latent_oracle as L
    g = torch.Generator().manual_seed(5)
    q, _ = torch.linalg.qr(torch.randn(16, 16, generator=g))
    mean = torch.randn(16, generator=g)
    z = torch.randn(2, 16, 30, generator=g)
    y = L.post_process_latent(z, q, mean, 16)
    back = L.pre_process_latent(y, z.new_zeros(2, 0, 30), q, mean)
    assert rel_l2(back.numpy(), z.numpy()) < 1e-5                                          # orthonormal PCA
    assert L.post_process_latent(z, q, mean, 4).shape == (2, 4, 30)
