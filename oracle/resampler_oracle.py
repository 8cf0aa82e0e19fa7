"""TEST INFRASTRUCTURE ONLY -- CPU restatement of `rave.resampler.Resampler` (rave/resampler.py:9-66),
the Kaiser FIR pair the exported model wraps around encode / decode (scripts/export.py:377-378,
407-408).

  design        resampler.py:26-56   kaiser_filter(pi / ratio, 140) (pqmf.py:55-70); the up-sampler is
                                     the same filter split into `ratio` polyphase rows (left-padded
                                     to a multiple of the ratio, then to an odd length)
  to_model      resampler.py:58-59   strided conv, centred padding (offline) / cached conv (streaming)
  from_model    resampler.py:61-66   `ratio`-channel conv, channels interleaved into time

Streaming follows cached_conv.CachedConv1d (oracle/cached_conv_restated.py): the input is
preceded by `total padding` (+ a stride-alignment delay) samples of history, zero at first.
Written here as ONE continuous FIR over the concatenated stream, which is what the caches compute.

PINNING: tests/test_oracle_pins.py::test_resampler_oracle_matches_live_reference (both modes, the
ratios of the reference's tests/test_resampler.py) and tests/golden/resampler.npz.
"""
import numpy as np
import torch
import torch.nn.functional as F

from .rave_oracle import _kaiser_lowpass


def design(ratio):
    """-> (down [k], up [ratio, k_up]) float32 tensors.  resampler.py:26-56."""
    filt = torch.from_numpy(_kaiser_lowpass(np.pi / ratio, 140)).float()
    down = filt.clone()
    pad = len(filt) % ratio
    f = F.pad(filt, (pad, 0))
    f = f.reshape(-1, ratio).permute(1, 0)
    pad = (f.shape[-1] + 1) % 2
    up = F.pad(f, (pad, 0)).contiguous()
    return down, up


def _pads(k):
    return (k - 1) // 2, k // 2            # cached_conv.get_padding, centred; stride ignored


class ResamplerOracle:
    def __init__(self, ratio, streaming=False):
        self.ratio = ratio
        self.streaming = streaming
        self.down, self.up = design(ratio)
        self.hist_down = None
        self.hist_up = None

    def _fir(self, x, w, stride, hist_name):
        """w: [C, k]; continuous FIR.  offline: centred zero padding; streaming: history on the left."""
        k = w.shape[-1]
        pl, pr = _pads(k)
        if not self.streaming:
            xp = F.pad(x, (pl, pr))
        else:
            total = pl + pr
            delay = (stride - (pr % stride)) % stride          # CachedConv1d stride_delay (cumulative_delay 0)
            H = total + delay
            hist = getattr(self, hist_name)
            if hist is None:
                hist = x.new_zeros(x.shape[0], 1, H)
            full = torch.cat([hist, x], -1)
            setattr(self, hist_name, full[..., full.shape[-1] - H:].clone())
            xp = full[..., :full.shape[-1] - delay] if delay else full
        return F.conv1d(xp, w.unsqueeze(1), stride=stride)

    def to_model_sampling_rate(self, x):
        return self._fir(x, self.down.unsqueeze(0), self.ratio, "hist_down")

    def from_model_sampling_rate(self, x):
        y = self._fir(x, self.up, 1, "hist_up")                 # [B, ratio, T]
        return y.permute(0, 2, 1).reshape(y.shape[0], -1).unsqueeze(1)
