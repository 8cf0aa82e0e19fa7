"""CPU-side tests: gin reader, netspec builder, weight folding, C-ABI library exports."""
import ctypes
import os
import re

import numpy as np
import pytest

from rave_b200 import capi, gin_lite, netspec
from rave_b200.weights import _phase_pack, collect

from .util import ROOT, golden_state_dict, load_golden

REF = "/root/reference/rave/configs"


def _spec(files, bindings=(), streaming=False, search=None):
    cfg = gin_lite.parse_config_files_and_bindings(list(files), list(bindings), search_paths=search)
    return netspec.build_netspec(cfg, streaming=streaming)


def test_get_padding_matches_cached_conv_rule():
    assert netspec.get_padding(1) == (0, 0)
    assert netspec.get_padding(3) == (1, 1)
    assert netspec.get_padding(3, dilation=9) == (9, 9)
    assert netspec.get_padding(8, 4) == (3, 4)            # stride ignored
    assert netspec.get_padding(8, 4, mode="causal") == (7, 0)
    assert netspec.get_padding(513) == (256, 256)
    assert netspec.get_padding(513, mode="causal") == (512, 0)
    with pytest.raises(ValueError):
        netspec.get_padding(3, mode="bogus")


def test_gin_lite_syntax():
    cfg = gin_lite.Config().parse_string('''
A = 3  # comment
B = [%A, 2,
     1]
mod.Cls.x = %B
scope/mod.Cls.y = @other.Thing
mod.Fn:
    p = {'a': 1,
         'b': (5, 1)}
    q = @mod.fn()
''')
    assert cfg.macro("A") == 3
    assert cfg.params("Cls")["x"] == [3, 2, 1]
    assert "y" not in cfg.params("Cls")
    assert cfg.params("Cls", "scope")["y"].name == "Thing"
    assert cfg.params("Fn")["p"] == {"a": 1, "b": (5, 1)}
    assert cfg.params("Fn")["q"].evaluate


def test_v2_macs_and_state_match_survey():
    s = _spec(["v2.gin"])
    m = s.macs_per_frame()
    # enc 19.243 M + 0.197 M PQMF(6 bands); dec 20.775 M + 0.541 M
    assert m["enc"] == 19_440_000 and m["dec"] == 21_315_584
    assert s.samples_per_frame == 1024 and s.slots[s.enc_out].channels == 320
    st = _spec(["v2.gin"], streaming=True)
    assert 40_000 < sum(sl.hist * sl.channels for sl in st.slots) < 43_000    # ~42 K elements


@pytest.mark.parametrize("stack", [["v1"], ["v2"], ["v2", "adain"], ["v2", "wasserstein"], ["v2", "spherical"],
                                   ["discrete"], ["discrete", "snake"], ["discrete", "snake", "adain"],
                                   ["discrete", "noise"], ["v3"], ["raspberry"], ["onnx"], ["discrete_v3"]])
@pytest.mark.parametrize("causal", [False, True])
def test_config_stacks_build(stack, causal):
    """The config stacks of the reference's tests/test_configs.py:15-31 (+causal) all build,
    offline and streaming, with CAPACITY=2 as in test_configs.py:58."""
    files = [s + ".gin" for s in stack] + (["causal.gin"] if causal else [])
    for streaming in (False, True):
        s = _spec(files, ["CAPACITY=2"], streaming)
        assert s.causal == causal and s.slots[s.dec_out].per_frame == s.samples_per_frame


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference tree not present")
@pytest.mark.parametrize("name", ["v1", "v2", "v3", "discrete", "raspberry", "discrete_v3"])
def test_reference_gin_files_give_same_program(name):
    """Parsing the reference's own .gin files yields the same program as the in-repo subset."""
    sp = ["/root/reference", "/root/reference/rave", REF]
    a = _spec([f"{REF}/{name}.gin"], search=sp)
    b = _spec([f"{name}.gin"])
    assert [(o.kind, o.c_in, o.c_out, o.k, o.stride, o.dilation, o.off, o.pre_act, o.name) for o in a.ops] == \
           [(o.kind, o.c_in, o.c_out, o.k, o.stride, o.dilation, o.off, o.pre_act, o.name) for o in b.ops]


def test_unsupported_and_invalid_configs_raise():
    with pytest.raises(ValueError):
        _spec(["v2.gin"], ["rave.RAVE.encoder=@blocks.NoSuchEncoder"])
    with pytest.raises(Exception):
        _spec(["v2.gin"], ["blocks.normalization.mode='spectral'"])
    with pytest.raises(ValueError):
        _spec(["v2.gin"], ["cc.get_padding.mode='reflect'"])


def test_phase_pack_equals_conv_transpose():
    import torch
    import torch.nn.functional as F
    torch.manual_seed(0)
    for r in (2, 4, 8):
        w = torch.randn(5, 3, 2 * r)
        x = torch.randn(2, 5, 11)
        ref = F.conv_transpose1d(x, w, None, r, r // 2)
        weq = torch.from_numpy(_phase_pack(w.numpy(), r, r // 2, 3))
        y = F.conv1d(F.pad(x, (1, 1)), weq)                       # [B, r*3, 11]
        y = y.reshape(2, r, 3, 11).permute(0, 2, 3, 1).reshape(2, 3, 11 * r)
        assert torch.allclose(y, ref, atol=1e-5)


def test_weight_collection_covers_every_op():
    meta, _ = load_golden("v3_c8")
    cfg = gin_lite.parse_config_files_and_bindings(["v3.gin"], ["CAPACITY=8"])
    s = netspec.build_netspec(cfg)
    items = collect(s, golden_state_dict(meta), pqmf_weights=(np.zeros((16, 1, 513)), np.zeros((16, 16, 33))))
    have = {(i, r) for i, r, _ in items}
    for i, op in enumerate(s.ops):
        if op.kind == netspec.OP_CONV:
            assert (i, capi.RB_W_CONV) in have
            if op.has_bias:
                assert (i, capi.RB_W_BIAS) in have
            if op.pre_act == netspec.ACT_SNAKE:
                assert (i, capi.RB_W_ALPHA) in have
    with pytest.raises(KeyError):
        collect(s, {}, None)


def test_library_exports_every_symbol_of_the_header():
    """The C-ABI library loads and exports every function include/rave_b200.h declares."""
    hdr = open(os.path.join(ROOT, "include", "rave_b200.h")).read()
    declared = set(re.findall(r"\b(rb_[a-z_0-9]+)\s*\(", hdr))
    assert declared == set(capi.EXPORTS), declared ^ set(capi.EXPORTS)
    assert os.path.exists(capi.LIB_PATH), "build with python -m rave_b200.build"
    lib = ctypes.CDLL(capi.LIB_PATH)
    for name in declared:
        assert hasattr(lib, name), name
    lib.rb_abi_version.restype = ctypes.c_int
    assert lib.rb_abi_version() == capi.ABI_VERSION


def test_struct_layouts_match_header_sizes():
    assert ctypes.sizeof(capi.RbSlot) == 16
    assert ctypes.sizeof(capi.RbOp) == 4 * 20
    assert ctypes.sizeof(capi.RbTensorDesc) == 24


def test_no_cuda_device_fails_loudly():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from rave_b200 import RAVE
    with pytest.raises(RuntimeError):
        RAVE.from_gin(["v2.gin"], ["CAPACITY=2"])


def test_checkpoint_search_and_load(tmp_path):
    """rave/core.py:132-139 + scripts/export.py:557-568 on a Lightning-shaped checkpoint file."""
    import torch
    from rave_b200 import checkpoint as ck
    run = tmp_path / "run" / "version_0" / "checkpoints"
    run.mkdir(parents=True)
    sd = {"encoder.encoder.net.0.weight": torch.ones(2, 3, 7), "speaker_encoder.in_layer.weight": torch.zeros(128, 16, 7),
          "speaker_encoder.bn5.running_var": torch.ones(1536), "decoder.net.1.weight_g": torch.ones(4, 1, 1)}
    ema = {"encoder.encoder.net.0.weight": torch.full((2, 3, 7), 2.0)}
    torch.save({"state_dict": sd, "callbacks": {"EMA": ema}, "epoch": 3}, run / "epoch3-last.ckpt")
    torch.save({"state_dict": sd, "callbacks": {}}, run / "best.ckpt")
    torch.save({"state_dict": {}, "callbacks": {}}, run / "epoch1-last.ckpt")
    root = str(tmp_path / "run")
    assert ck.search_for_run(None) is None
    assert ck.search_for_run(str(run / "best.ckpt")) == str(run / "best.ckpt")
    assert ck.search_for_run(root).endswith("epoch3-last.ckpt")          # sorted, last one wins
    assert ck.search_for_run(root, mode="best").endswith("best.ckpt")
    assert ck.search_for_run(str(tmp_path / "run" / "version_0"), mode="nothing") is None
    got = ck.load_checkpoint(ck.search_for_run(root))
    assert set(got) == set(sd) and torch.equal(got["encoder.encoder.net.0.weight"], sd["encoder.encoder.net.0.weight"])
    got = ck.load_checkpoint(ck.search_for_run(root), ema_weights=True)
    assert set(got) == set(ema) and float(got["encoder.encoder.net.0.weight"].mean()) == 2.0
    got = ck.load_checkpoint(str(run / "best.ckpt"), ema_weights=True)   # no EMA callback: falls back to state_dict
    assert set(got) == set(sd)
    own, spk = ck.split_state_dict(sd)
    assert set(spk) == {"in_layer.weight", "bn5.running_var"} and "decoder.net.1.weight_g" in own and len(own) == 2


def test_resampler_design_matches_oracle_and_rejects_what_upstream_rejects():
    from oracle.resampler_oracle import design as oracle_design
    from rave_b200.resampler import design
    for ratio in (2, 3):
        d, u = design(ratio)
        do, uo = oracle_design(ratio)
        assert np.array_equal(d, do.numpy()) and np.array_equal(u, uo.numpy())
    with pytest.raises(RuntimeError):
        design(4)                        # upstream's reshape(-1, ratio) fails for this filter length too


OPERATIVE = '''import rave
import cached_conv

# Macros:
# ==============================================================================
CAPACITY = 8
N_BAND = 16

# Parameters for variational/blocks.EncoderV2:
# ==============================================================================
variational/blocks.EncoderV2.n_out = 1

# Parameters for blocks.EncoderV2:
# ==============================================================================
blocks.EncoderV2.dilations = \\
    [[1, 3, 9], [1, 3, 9], [1, 3, 9], [1, 3]]
blocks.EncoderV2.capacity = %CAPACITY

# Parameters for pqmf.CachedPQMF:
# ==============================================================================
pqmf.CachedPQMF.n_band = \\
    %N_BAND
'''


def test_gin_lite_reads_operative_config_dumps_and_never_evaluates_code():
    """run/config.gin is gin.operative_config_str() (train_rave.py:287-288): `# Parameters for ...` headers,
    scoped selectors and long bindings wrapped with a trailing backslash.  Values are literals only."""
    cfg = gin_lite.Config().parse_string('include "v2.gin"\n' + OPERATIVE)
    assert cfg.params("EncoderV2")["dilations"] == [[1, 3, 9], [1, 3, 9], [1, 3, 9], [1, 3]]
    assert cfg.params("EncoderV2")["capacity"] == 8
    assert cfg.params("EncoderV2", "variational")["n_out"] == 1
    assert cfg.params("CachedPQMF")["n_band"] == 16
    spec = netspec.build_netspec(cfg)
    assert spec.samples_per_frame == 1024
    for bad in ("().__class__.__base__.__subclasses__()", "__import__('os').system('true')", "[1, 2][0]", "1 + 2",
                "open('/etc/passwd')", "{**{}}", "-'a'"):
        with pytest.raises(ValueError):
            gin_lite.parse_value(bad)
    assert gin_lite.parse_value("{'a': (-1.5e-3, [True, None]), 2: 'x'}") == {"a": (-1.5e-3, [True, None]), 2: "x"}


def test_plain_pqmf_alias_and_leaky_slope():
    """pqmf.PQMF (pqmf.py:179-231): symmetric padding whatever the conv padding mode, synthesis one row early,
    no streaming twin; blocks.leaky_relu.alpha other than 0.2 is refused instead of silently ignored."""
    from rave_b200.netspec import OP_PQMF_FWD, OP_PQMF_INV
    a = _spec(["v2.gin"], ["CAPACITY=8"])
    b = _spec(["v2.gin", "causal.gin"], ["CAPACITY=8", "rave.RAVE.pqmf = @pqmf.PQMF", "pqmf.PQMF.n_band = 16",
                                          "pqmf.PQMF.attenuation = 100"])
    fa = [o for o in a.ops if o.kind == OP_PQMF_FWD][0]
    fb = [o for o in b.ops if o.kind == OP_PQMF_FWD][0]
    ia = [o for o in a.ops if o.kind == OP_PQMF_INV][0]
    ib = [o for o in b.ops if o.kind == OP_PQMF_INV][0]
    assert fb.off == fa.off and ib.off == ia.off - 1
    with pytest.raises(NotImplementedError):
        _spec(["v2.gin"], ["rave.RAVE.pqmf = @pqmf.PQMF", "pqmf.PQMF.n_band = 16"], streaming=True)
    with pytest.raises(NotImplementedError):
        _spec(["v2.gin"], ["blocks.leaky_relu.alpha = 0.1"])


def test_ema_state_dict_defaults_batchnorm_buffers():
    """EMA dicts hold parameters only (train_rave.py:85): with strict=False the v1 encoder's convs still load,
    their BatchNorm buffers default to the constructed mean 0 / var 1, and the caller sees what was skipped."""
    from rave_b200 import capi
    from rave_b200.synth import golden_state_dict
    from rave_b200.weights import collect
    spec = _spec(["v1.gin"], ["CAPACITY=8"])
    _, sd = golden_state_dict("v1_c8")
    ema = {k: v for k, v in sd.items() if "running_" not in k and "num_batches" not in k}
    from rave_b200.pqmf_design import design
    _, _, wa, ws = design(100, 16)
    skipped = []
    got = collect(spec, ema, (wa, ws), strict=False, skipped=skipped)
    convs = {i for i, role, _ in got if role == capi.RB_W_CONV}
    full = {i for i, role, _ in collect(spec, sd, (wa, ws)) if role == capi.RB_W_CONV}
    assert convs == full and any("defaulted" in s for s in skipped)
    with pytest.raises(KeyError):
        collect(spec, ema, (wa, ws), strict=True)
