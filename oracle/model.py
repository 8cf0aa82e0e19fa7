"""TEST INFRASTRUCTURE ONLY -- `OracleRAVE`: RAVE.encode/decode/forward on the CPU oracle.

Restates rave/model.py:594-634 (encode / decode / forward) and CachedPQMF.forward /
inverse (rave/pqmf.py:269-284) over the functional stacks in oracle/rave_oracle.py.
Streaming mode carries explicit per-layer state (oracle.rave_oracle.StreamState) exactly
as cached_conv's module buffers would (scripts/export.py:542-594).
"""
import numpy as np
import torch
import torch.nn.functional as F

from . import rave_oracle as O
from .weights_recipe import build_state_dict


def _to_torch(sd):
    return {k: (torch.from_numpy(np.asarray(v)) if not torch.is_tensor(v) else v) for k, v in sd.items()}


class OracleRAVE:
    """CPU fp32 oracle with the reference's public surface for the hot path."""

    def __init__(self, cfg, state_dict, streaming=False):
        self.cfg = dict(cfg)
        self.sd = _to_torch(state_dict)
        self.causal = bool(cfg.get("causal", False))
        self.streaming = streaming
        self.stream = O.StreamState() if streaming else None
        if "pqmf.forward_conv.weight" not in self.sd:
            _, hk = O.pqmf_bank(cfg.get("attenuation", 100), cfg["n_band"])
            wf, wi = O.pqmf_conv_weights(hk)
            self.sd["pqmf.hk"] = hk
            self.sd["pqmf.forward_conv.weight"] = wf
            self.sd["pqmf.inverse_conv.weight"] = wi
        self.sr = cfg.get("sr", 44100)
        self.latent_size = cfg["latent_size"]

    # -- construction helpers -------------------------------------------------------
    @classmethod
    def from_recipe(cls, cfg, names_shapes, seed=0, streaming=False, base=None):
        return cls(cfg, build_state_dict(names_shapes, seed, base=base), streaming)

    def reset(self):
        self.stream = O.StreamState() if self.streaming else None

    def _ctx(self):
        return O._Ctx(self.sd, self.causal, self.stream)

    def update_adain(self, learn_source=False, learn_target=False, reset_source=False, reset_target=False):
        """ScriptedRAVE.update_adain (scripts/export.py:248-265) on every AdaIN module's buffers:
        learn_source -> learn_x, learn_target -> learn_y; resets as blocks.py:875-883."""
        import torch
        for key in [k for k in self.sd if k.endswith("num_update_x")]:
            p = key[:-len("num_update_x")]
            self.sd[p + "learn_x"] = torch.full((1,), 1.0 if learn_source else 0.0)
            self.sd[p + "learn_y"] = torch.full((1,), 1.0 if learn_target else 0.0)
            if reset_target:
                self.sd[p + "mean_y"].zero_()
                self.sd[p + "std_y"].zero_().add_(1)
                self.sd[p + "num_update_y"].zero_()
            if reset_source:
                self.sd[p + "mean_x"].zero_()
                self.sd[p + "std_x"].zero_().add_(1)
                self.sd[p + "num_update_x"].zero_()

    # -- PQMF -------------------------------------------------------------------------
    def pqmf_forward(self, x):
        """CachedPQMF.forward, pqmf.py:269-273."""
        w = self.sd["pqmf.forward_conv.weight"]
        M, k = w.shape[0], w.shape[-1]
        pl, pr = O.get_padding(k, 1, self.causal and self.cfg.get("pqmf", "CachedPQMF") != "PQMF")
        if self.stream is None:
            x = F.pad(x, (pl, pr))
        else:
            x = self.stream.pad("pqmf.forward_conv.cache", x, pl + pr)
        return O._reverse_half(F.conv1d(x, w, None, M))

    def pqmf_inverse(self, y):
        """CachedPQMF.inverse, pqmf.py:275-284."""
        w = self.sd["pqmf.inverse_conv.weight"]
        M, k = w.shape[0], w.shape[-1]
        pl, pr = O.get_padding(k, 1, self.causal and self.cfg.get("pqmf", "CachedPQMF") != "PQMF")
        y = O._reverse_half(y)
        if self.stream is None:
            y = F.pad(y, (pl, pr))
        else:
            y = self.stream.pad("pqmf.inverse_conv.cache", y, pl + pr)
        if self.cfg.get("pqmf", "CachedPQMF") == "PQMF" and self.stream is None:
            # pqmf.PQMF.inverse (pqmf.py:111-135): the same polyphase synthesis one sub-band row earlier
            # (pinned against the live class in tests/test_oracle_pins.py)
            y = F.pad(y[..., 1:], (0, 1))
        y = F.conv1d(y, w) * M
        y = y.flip(1)                                          # [B, M, L]
        B, _, L = y.shape
        return y.permute(0, 2, 1).reshape(B, 1, L * M)         # time-major interleave

    # -- stacks -----------------------------------------------------------------------
    def encoder(self, xb):
        cfg = self.cfg
        if cfg["arch"] == "v2":
            return O.encoder_v2(self._ctx(), "encoder.encoder.", xb, cfg)
        return O.encoder_v1(self._ctx(), "encoder.encoder.", xb, cfg)

    def decoder(self, z, noise=None):
        cfg = self.cfg
        if cfg["arch"] == "v2":
            return O.generator_v2(self._ctx(), "decoder.", z, cfg, noise)
        return O.generator_v1(self._ctx(), "decoder.", z, cfg)

    # -- public surface (rave/model.py:594-634) ------------------------------------------
    @torch.no_grad()
    def encode(self, x, speaker=None, f0=None):
        """model.py:594-622.  `speaker` [B, S, 1]: one embedding per stream instead of the module's single
        buffer (the export wrapper swaps `target_emb` per call, scripts/export.py:381-394); `f0` [B, C, F]:
        the pitch channels of scripts/export.py:397 `z = cat(z, emb, f0_norm)`."""
        xb = self.pqmf_forward(x)
        z = self.encoder(xb[:, :self.cfg["enc_data_size"], :])
        if speaker is None:
            emb = self.sd["speaker"].repeat(z.shape[0], 1, z.shape[-1])
        else:
            emb = speaker.reshape(z.shape[0], -1, 1).repeat(1, 1, z.shape[-1])
        parts = (z, emb) if f0 is None else (z, emb, f0)
        return torch.cat(parts, dim=1)

    @torch.no_grad()
    def decode(self, z, noise=None):
        return self.pqmf_inverse(self.decoder(z, noise))

    @torch.no_grad()
    def forward(self, x, noise=None, speaker=None, f0=None):
        return self.decode(self.encode(x, speaker, f0), noise)

    __call__ = forward

    # -- discrete boundary (scripts/export.py:503-517) --------------------------------------
    @torch.no_grad()
    def rvq_encode(self, z):
        return O.rvq_encode(self.sd, "encoder.rvq.", z, self.cfg["num_quantizers"])

    @torch.no_grad()
    def rvq_decode(self, idx):
        return O.rvq_decode(self.sd, "encoder.rvq.", idx)


def rel_l2(a, b):
    a = torch.as_tensor(a).double().flatten()
    b = torch.as_tensor(b).double().flatten()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))
