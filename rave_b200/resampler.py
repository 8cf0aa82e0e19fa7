"""Host-side mirror of `rave.resampler.Resampler` (rave/resampler.py:9-66): the Kaiser FIR pair the
exported model wraps around encode / decode (scripts/export.py:377-378, :407-408).  Filter design on
the host exactly as upstream; both directions run in librave_b200.so (csrc/resampler.cu).  No CPU
or PyTorch fallback.
"""
import ctypes as C

import numpy as np
import torch

from . import capi
from .model import MAX_BATCH_SIZE
from .pqmf_design import _lowpass


def design(ratio):
    """resampler.py:26-56 -> (down [k], up [ratio, k_up]) float32."""
    filt = _lowpass(np.pi / ratio, 140).astype(np.float32)          # kaiser_filter(wc, 140), pqmf.py:55-70
    down = filt.copy()
    pad = len(filt) % ratio
    f = np.concatenate([np.zeros(pad, np.float32), filt])
    if len(f) % ratio:
        # upstream's reshape(-1, ratio) raises here too (e.g. ratio 4)
        raise RuntimeError(f"shape '[-1, {ratio}]' is invalid for input of size {len(f)}")
    f = f.reshape(-1, ratio).T
    pad = (f.shape[-1] + 1) % 2
    up = np.concatenate([np.zeros((ratio, pad), np.float32), f], axis=1)
    return np.ascontiguousarray(down), np.ascontiguousarray(up)


class Resampler:
    def __init__(self, target_sr, model_sr, streaming=False, device="cuda:0", max_streams=MAX_BATCH_SIZE):
        assert target_sr != model_sr, "identical source and target rates"
        if not torch.cuda.is_available():
            raise RuntimeError("rave_b200.Resampler needs a CUDA device (there is no CPU fallback)")
        self.model_sr = model_sr
        self.taget_sr = target_sr          # (sic) upstream attribute name, resampler.py:16
        ratio = target_sr // model_sr
        assert int(ratio) == ratio
        if ratio % 2 and streaming:
            raise ValueError(f"When using streaming mode, resampling ratio must be a power of 2, got {ratio}")
        self.ratio = ratio
        self.streaming = bool(streaming)
        self.down_weight, self.up_weight = design(ratio)
        self._lib = capi.load_library()
        dev = torch.device(device)
        self.device = torch.device("cuda", dev.index if dev.index is not None else torch.cuda.current_device())
        h = C.c_void_p()
        fp = lambda a: a.ctypes.data_as(C.POINTER(C.c_float))  # noqa: E731
        capi.check(self._lib.rb_resampler_create(self.device.index, ratio, fp(self.down_weight), self.down_weight.size,
                                                 fp(self.up_weight), self.up_weight.shape[1], int(self.streaming),
                                                 int(max_streams), C.byref(h)))
        self._h = h

    def _run(self, fn, x, out_len):
        if x.dim() != 3 or x.shape[1] != 1:
            raise ValueError(f"expected [B, 1, T], got {tuple(x.shape)}")
        if x.dtype not in (torch.float32, torch.bfloat16):
            raise ValueError("dtype must be torch.float32 or torch.bfloat16")
        x = x.to(self.device).contiguous()
        y = torch.empty(x.shape[0], 1, out_len, device=self.device, dtype=x.dtype)
        with torch.cuda.device(self.device):
            stream = torch.cuda.current_stream().cuda_stream
            capi.check(fn(self._h, x.data_ptr(), x.shape[0], x.shape[2],
                          capi.RB_F32 if x.dtype == torch.float32 else capi.RB_BF16, y.data_ptr(), stream))
        return y

    def to_model_sampling_rate(self, x):
        """resampler.py:58-59."""
        if x.shape[-1] % self.ratio:
            raise ValueError(f"T = {x.shape[-1]} is not a multiple of the ratio {self.ratio}")
        return self._run(self._lib.rb_resampler_to_model, x, x.shape[-1] // self.ratio)

    def from_model_sampling_rate(self, x):
        """resampler.py:61-66."""
        return self._run(self._lib.rb_resampler_from_model, x, x.shape[-1] * self.ratio)

    def reset(self):
        with torch.cuda.device(self.device):
            capi.check(self._lib.rb_resampler_reset(self._h, torch.cuda.current_stream().cuda_stream))

    @property
    def kernel_launches(self):
        n = C.c_int64()
        capi.check(self._lib.rb_resampler_launches(self._h, C.byref(n)))
        return n.value

    def close(self):
        if getattr(self, "_h", None):
            self._lib.rb_resampler_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
