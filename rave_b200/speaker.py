"""Host-side mirror of the reference's speaker encoder `SpeakerRAVE`
(rave/CombinedRave.py:200-328), the step before RAVE.encode: model.py:246-247 computes
`speaker = speaker_encoder(pqmf(utterance)).unsqueeze(2)`, which encode() appends to every
latent frame (model.py:618-620).  All compute runs in librave_b200.so (csrc/speaker.cu); there is
no CPU or PyTorch fallback.
"""
import ctypes as C

import numpy as np
import torch

from . import capi
from .weights import _folded

# state_dict prefix of each rb_speaker_layer, in the header's order
LAYER_KEYS = (
    "in_layer",
    "layer2.0.aligned.branches.0.net.1", "layer2.0.aligned.branches.0.net.3", "layer2.2",
    "layer3.0.aligned.branches.0.net.1", "layer3.0.aligned.branches.0.net.3", "layer3.2",
    "layer4.0.aligned.branches.0.net.1", "layer4.0.aligned.branches.0.net.3", "layer4.2",
    "cat_layer", "out_layer", "attention.0", "attention.3", "fc6",
)
BN_KEYS = ("attention.2", "bn5")
N_BANDS = 16
TIME_MULTIPLE = 32           # product of the three strides (4, 4, 2)


def _f32(t):
    if isinstance(t, torch.Tensor):
        t = t.detach().cpu().numpy()
    return np.ascontiguousarray(t, dtype=np.float32)


def _ptr(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


class SpeakerRAVE:
    """`SpeakerRAVE.forward` on the GPU: [B, 16, T] PQMF bands -> [B, 256] (eval mode).

    state_dict: the module's own state_dict (plain `weight` / `bias`; weight-normalised
    `weight_g` / `weight_v` pairs are folded).  pqmf_weight: CachedPQMF.forward_conv.weight
    [16, 1, k], needed only by `embed_audio`.
    """

    def __init__(self, state_dict, pqmf_weight=None, device=0):
        if not torch.cuda.is_available():
            raise RuntimeError("rave_b200.SpeakerRAVE needs a CUDA device (there is no CPU fallback)")
        self._lib = capi.load_library()
        dev = torch.device("cuda", device) if isinstance(device, int) else torch.device(device)
        self.device = torch.device("cuda", dev.index if dev.index is not None else torch.cuda.current_device())
        h = C.c_void_p()
        capi.check(self._lib.rb_speaker_create(self.device.index, C.byref(h)))
        self._h = h
        for i, key in enumerate(LAYER_KEYS):
            w, b = _folded(state_dict, key + ".")
            capi.check(self._lib.rb_speaker_load(self._h, i, _ptr(w), w.size, _ptr(b), b.size))
        for i, key in enumerate(BN_KEYS):
            g, b, m, v = (_f32(state_dict[f"{key}.{n}"]) for n in ("weight", "bias", "running_mean", "running_var"))
            capi.check(self._lib.rb_speaker_load_bn(self._h, i, _ptr(g), _ptr(b), _ptr(m), _ptr(v), g.size, 1e-5))
        self.has_pqmf = pqmf_weight is not None
        if self.has_pqmf:
            w = _f32(pqmf_weight)
            if w.ndim != 3 or w.shape[0] != N_BANDS or w.shape[1] != 1:
                raise ValueError(f"pqmf_weight must be [16, 1, k], got {tuple(w.shape)}")
            capi.check(self._lib.rb_speaker_load_pqmf(self._h, _ptr(w), w.shape[0], w.shape[2]))

    def _run(self, fn, x, channels, multiple):
        if x.dim() != 3 or x.shape[1] != channels:
            raise ValueError(f"expected [B, {channels}, T], got {tuple(x.shape)}")
        if x.shape[2] % multiple or x.shape[2] < 2 * multiple:
            raise ValueError(f"T = {x.shape[2]} must be a multiple of {multiple} and at least {2 * multiple}")
        x = x.to(self.device, torch.float32).contiguous()
        emb = torch.empty(x.shape[0], 256, device=self.device, dtype=torch.float32)
        with torch.cuda.device(self.device):
            stream = torch.cuda.current_stream().cuda_stream
            capi.check(fn(self._h, x.data_ptr(), x.shape[0], x.shape[2], emb.data_ptr(), stream))
        return emb

    def forward(self, x):
        """CombinedRave.py:297-328.  x: [B, 16, T] (T % 32 == 0) -> [B, 256]."""
        return self._run(self._lib.rb_speaker_embed, x, N_BANDS, TIME_MULTIPLE)

    __call__ = forward

    def embed_audio(self, audio):
        """model.py:246-247: speaker_encoder(pqmf(audio)).  audio: [B, 1, T] (T % 512 == 0) -> [B, 256]."""
        if not self.has_pqmf:
            raise RuntimeError("embed_audio needs pqmf_weight at construction")
        return self._run(self._lib.rb_speaker_embed_audio, audio, 1, N_BANDS * TIME_MULTIPLE)

    @property
    def kernel_launches(self):
        n = C.c_int64()
        capi.check(self._lib.rb_speaker_launches(self._h, C.byref(n)))
        return n.value

    def close(self):
        if getattr(self, "_h", None):
            self._lib.rb_speaker_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
