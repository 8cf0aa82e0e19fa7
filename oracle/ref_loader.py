"""TEST INFRASTRUCTURE ONLY -- loads the REAL reference modules by file path.

Works only where /root/reference exists (the build container).  Used to (a) pin the
restated `cached_conv` with the reference's own tests/test_residual.py identities,
(b) pin oracle/rave_oracle.py, (c) generate tests/golden/*.npz (oracle/make_golden.py).
Never used on the GPU box, never imported by rave_b200/.

Why by path: `import rave` fails here -- rave/model.py:29-30 loads an
audio file at import time, rave/pitch_utils.py:12 spawns a model on cuda:2, and gin,
pytorch_lightning, librosa, lmdb, GPUtil and cached_conv are not installed.  The four
files on the hot path (quantization.py, core.py, pqmf.py, blocks.py) import cleanly once
stub modules for those names are registered and two removed scipy names are shimmed.
"""
import importlib.util
import os
import sys
import types

import numpy as np
import torch

from . import cached_conv_restated as cc

REFERENCE_ROOT = os.environ.get("RAVE_REFERENCE_ROOT", "/root/reference")

# gin bindings that the reference's configs apply to functions defined in blocks.py
# (v1.gin:41 `blocks.normalization.mode = 'weight_norm'`).  name -> kwargs
GIN_BINDINGS = {"normalization": {"mode": "weight_norm"}}

_loaded = {}


def available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "rave", "blocks.py"))


def _make_gin_stub():
    gin = types.ModuleType("gin")

    def configurable(*dargs, **dkwargs):
        def wrap(fn):
            if not isinstance(fn, type):
                name = fn.__name__

                def bound(*a, **k):
                    for kk, vv in GIN_BINDINGS.get(name, {}).items():
                        k.setdefault(kk, vv)
                    return fn(*a, **k)
                bound.__name__ = name
                bound.__wrapped__ = fn
                return bound
            return fn
        if len(dargs) == 1 and callable(dargs[0]) and not dkwargs:
            return wrap(dargs[0])
        return wrap

    gin.configurable = configurable
    gin.enter_interactive_mode = lambda: None
    gin.external_configurable = lambda f, module=None, name=None: f
    gin.add_config_file_search_path = lambda p: None
    return gin


def _install_stubs():
    if "gin" not in sys.modules:
        sys.modules["gin"] = _make_gin_stub()
    for name in ("GPUtil", "librosa", "lmdb"):
        if name not in sys.modules:
            try:
                __import__(name)
            except Exception:
                sys.modules[name] = types.ModuleType(name)
    if "pytorch_lightning" not in sys.modules:
        try:
            import pytorch_lightning  # noqa: F401
        except Exception:
            pl = types.ModuleType("pytorch_lightning")
            pl.Callback = object
            pl.LightningModule = torch.nn.Module
            sys.modules["pytorch_lightning"] = pl
    if "torchaudio" not in sys.modules:            # CombinedRave.py:16 imports Spectrogram (unused by SpeakerRAVE)
        try:
            import torchaudio  # noqa: F401
        except Exception:
            ta, tt = types.ModuleType("torchaudio"), types.ModuleType("torchaudio.transforms")
            tt.Spectrogram = object
            ta.transforms = tt
            sys.modules["torchaudio"], sys.modules["torchaudio.transforms"] = ta, tt
    if "torchfcpe" not in sys.modules:             # pitch_utils.py:10-12 spawns the FCPE model on cuda:2 at import
        try:
            import torchfcpe  # noqa: F401
        except Exception:
            tf = types.ModuleType("torchfcpe")
            tf.spawn_bundled_infer_model = lambda device=None: None
            sys.modules["torchfcpe"] = tf
    sys.modules["cached_conv"] = cc
    # scipy shim: pqmf.py:10 imports scipy.signal.kaiser (moved to scipy.signal.windows) and
    # pqmf.py:69 calls firwin(..., nyq=pi) (keyword replaced by fs=2*nyq).
    import scipy.signal as ss
    import scipy.signal.windows as ssw
    if not hasattr(ss, "kaiser"):
        ss.kaiser = ssw.kaiser
    if not getattr(ss.firwin, "_rave_shim", False):
        _firwin = ss.firwin

        def firwin(*a, nyq=None, **k):
            if nyq is not None:
                k["fs"] = 2 * nyq
            return _firwin(*a, **k)
        firwin._rave_shim = True
        ss.firwin = firwin


def load(name: str):
    """Return the reference module rave/<name>.py loaded by path (cached)."""
    if name in _loaded:
        return _loaded[name]
    if not available():
        raise RuntimeError("reference tree not present at " + REFERENCE_ROOT)
    _install_stubs()
    if "rave" not in sys.modules or not getattr(sys.modules["rave"], "_oracle_pkg", False):
        pkg = types.ModuleType("rave")
        pkg.__path__ = [os.path.join(REFERENCE_ROOT, "rave")]
        pkg._oracle_pkg = True
        sys.modules["rave"] = pkg
    if name == "blocks":
        load("core")
    if name in ("CombinedRave", "resampler"):
        load("pqmf")
    path = os.path.join(REFERENCE_ROOT, "rave", name + ".py")
    spec = importlib.util.spec_from_file_location("rave." + name, path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules["rave." + name] = mod
    spec.loader.exec_module(mod)
    setattr(sys.modules["rave"], name, mod)
    _loaded[name] = mod
    return mod


class RefRAVE(torch.nn.Module):
    """rave/model.py:594-634 restated around the real reference blocks.

    `RAVE` itself cannot be constructed here (model.py:198,240-247 need cuda:2, a torch.hub
    download and files under /home/jupyter-arbu).  encode/decode/forward are 12 lines.
    `speaker` stands for `self.speaker` (model.py:247), a fixed [1, S, 1] embedding.
    """

    def __init__(self, pqmf, encoder, decoder, speaker, n_enc_bands=6):
        super().__init__()
        self.pqmf = pqmf
        self.encoder = encoder
        self.decoder = decoder
        self.register_buffer("speaker", speaker)
        self.n_enc_bands = n_enc_bands

    def encode(self, x):
        if self.pqmf is not None:
            x = self.pqmf(x)                                   # model.py:610-611
        z = self.encoder(x[:, :self.n_enc_bands, :])           # model.py:613
        emb = self.speaker.repeat(z.shape[0], 1, z.shape[-1])  # model.py:618
        return torch.cat((z, emb), dim=1)                      # model.py:620

    def decode(self, z):
        y = self.decoder(z)                                    # model.py:626
        if self.pqmf is not None:
            y = self.pqmf.inverse(y)                           # model.py:627-628
        return y

    def forward(self, x):
        return self.decode(self.encode(x))                     # model.py:634


def build_reference(cfg: dict, streaming: bool = False) -> RefRAVE:
    """Build the real reference modules for a flat config dict (see oracle/configs.py)."""
    blocks = load("blocks")
    pqmf_mod = load("pqmf")
    cc.use_cached_conv(streaming)
    cc.set_padding_mode("causal" if cfg.get("causal", False) else "centered")
    try:
        act = cfg.get("activation", "leaky_relu")
        if act == "snake":
            activation = blocks.Snake
        else:
            activation = lambda dim: torch.nn.LeakyReLU(.2)  # noqa: E731
        adain = blocks.AdaptiveInstanceNormalization if cfg.get("adain", False) else None
        pq = pqmf_mod.CachedPQMF(attenuation=cfg.get("attenuation", 100), n_band=cfg["n_band"])
        if cfg["arch"] == "v2":
            # DilatedUnit takes its activation from a gin binding (snake.gin:10-11), not from
            # its parent: patch the class default the way gin would.
            du_init = blocks.DilatedUnit.__init__
            if act == "snake":
                def patched(self, dim, kernel_size, dilation, activation=blocks.Snake):
                    du_init(self, dim, kernel_size, dilation, activation=activation)
                blocks.DilatedUnit.__init__ = patched
            try:
                enc_core = blocks.EncoderV2(
                    data_size=cfg["enc_data_size"], capacity=cfg["capacity"],
                    ratios=list(cfg["ratios"]), latent_size=cfg["latent_size"], n_out=1,
                    kernel_size=cfg["kernel_size"], dilations=cfg["dilations"],
                    activation=activation, adain=adain)
                noise_module = None
                if cfg.get("noise", False):
                    def noise_module(in_size):
                        return blocks.NoiseGeneratorV2(
                            in_size=in_size, hidden_size=128, data_size=cfg["n_band"],
                            ratios=[2, 2, 2], noise_bands=5, activation=activation)
                dec = blocks.GeneratorV2(
                    data_size=cfg["n_band"], capacity=cfg["capacity"],
                    ratios=list(cfg["ratios"]), latent_size=cfg["dec_latent_size"],
                    kernel_size=cfg["kernel_size"], dilations=cfg["dilations"],
                    amplitude_modulation=True, noise_module=noise_module,
                    activation=activation, adain=adain)
            finally:
                blocks.DilatedUnit.__init__ = du_init
            if cfg.get("encoder_type", "variational") == "discrete":
                quant = load("quantization")
                enc = blocks.DiscreteEncoder(
                    lambda: enc_core,
                    lambda: quant.ResidualVectorQuantization(
                        num_quantizers=cfg["num_quantizers"], dim=cfg["latent_size"],
                        codebook_size=cfg["codebook_size"], kmeans_init=False),
                    cfg["num_quantizers"])
            else:
                enc = blocks.VariationalEncoder(lambda: enc_core)
        elif cfg["arch"] == "v1":
            old = dict(blocks.ResidualStack.__init__.__kwdefaults__ or {})
            enc_core = blocks.Encoder(
                data_size=cfg["enc_data_size"], capacity=cfg["capacity"],
                latent_size=cfg["latent_size"], ratios=list(cfg["ratios"]), n_out=1,
                sample_norm=False, repeat_layers=1)
            enc = blocks.VariationalEncoder(lambda: enc_core)
            # blocks.ResidualStack / NoiseGenerator get their args from gin (v1.gin:64-70)
            rs_init = blocks.ResidualStack.__init__

            def rs_patched(self, dim, kernel_sizes=(3,), dilations_list=((1, 1), (3, 1), (5, 1)),
                           cumulative_delay=0):
                rs_init(self, dim, list(kernel_sizes), [list(d) for d in dilations_list],
                        cumulative_delay=cumulative_delay)
            blocks.ResidualStack.__init__ = rs_patched
            try:
                dec = blocks.Generator(
                    latent_size=cfg["dec_latent_size"], capacity=cfg["capacity"],
                    data_size=cfg["n_band"], ratios=list(cfg["ratios"]), loud_stride=1,
                    use_noise=False)
            finally:
                blocks.ResidualStack.__init__ = rs_init
            del old
        else:
            raise ValueError(cfg["arch"])
    finally:
        cc.use_cached_conv(False)
        cc.set_padding_mode("centered")
    spk = torch.zeros(1, cfg["speaker_size"], 1)
    model = RefRAVE(pq, enc, dec, spk, n_enc_bands=cfg["enc_data_size"])
    model.eval()
    return model


def fill_deterministic(model: torch.nn.Module, seed: int = 0):
    """Overwrite every parameter / float buffer with name-keyed deterministic values.

    Values depend only on (seed, tensor name, shape) -- not on torch's RNG or on module
    construction order -- so a test can rebuild the identical state_dict from the names
    and shapes stored in a golden file.  See oracle/weights_recipe.py.
    """
    from .weights_recipe import recipe_tensor
    sd = model.state_dict()
    new = {}
    for name, t in sd.items():
        v = recipe_tensor(name, tuple(t.shape), seed)
        if v is not None and t.dtype.is_floating_point:
            new[name] = torch.from_numpy(v).to(t.dtype)
    model.load_state_dict(new, strict=False)
    return model


def load_export_latent_classes():
    """The REAL `post_process_latent` / `pre_process_latent` methods of scripts/export.py:482-539.

    scripts/export.py cannot be imported (absl flags at import, nn_tilde, cached_conv, librosa), so the
    four `*ScriptedRAVE` class definitions are cut out of its syntax tree and executed against a stub
    base class: the method bodies that run are the reference's own lines, not a restatement.  Returns
    {class name: class}; instances need the attributes the methods read (latent_pca, latent_mean,
    latent_size, full_latent_size, encoder)."""
    import ast
    import torch.nn.functional as F
    path = os.path.join(REFERENCE_ROOT, "scripts", "export.py")
    tree = ast.parse(open(path).read(), path)
    want = ("VariationalScriptedRAVE", "DiscreteScriptedRAVE", "WasserteinScriptedRAVE", "SphericalScriptedRAVE")
    body = [n for n in tree.body if isinstance(n, ast.ClassDef) and n.name in want]
    if len(body) != len(want):
        raise RuntimeError("scripts/export.py no longer defines the four *ScriptedRAVE classes")

    class ScriptedRAVE:                                    # stub base: the methods only touch attributes
        pass

    rave_ns = types.SimpleNamespace(blocks=load("blocks"))
    ns = {"ScriptedRAVE": ScriptedRAVE, "torch": torch, "F": F, "rave": rave_ns, "np": np}
    exec(compile(ast.Module(body=body, type_ignores=[]), path, "exec"), ns)
    return {k: ns[k] for k in want}
