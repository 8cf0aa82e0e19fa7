"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the un-vendored `cached_conv` package.

The reference (abargum/RAVE) takes all of its convolution arithmetic and all of its
streaming semantics from the third-party package ``cached-conv>=2.5.0``
(/root/reference/requirements.txt:14; upstream acids-ircam/cached_conv).  Its source is
not in the reference tree and cannot be installed here (no network), so its published
behaviour is restated below.  Call sites this must satisfy:
rave/blocks.py (30x cc.Conv1d, 26x cc.get_padding, 9x cc.CachedSequential,
3x cc.AlignBranches, 2x cc.ConvTranspose1d, 4x cc.MAX_BATCH_SIZE), rave/pqmf.py:245-263,
scripts/export.py:543 (cc.use_cached_conv) and scripts/export_onnx.py:35 (`_pad`).

It is pinned by the reference's own tests/test_residual.py:37-122 (stream == offline after
`cumulative_delay`), which tests/test_oracle_pins.py re-runs against this module.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
may import this file.  The product (rave_b200/) never does.
"""
import torch
import torch.nn as nn
import torch.nn.functional as F

MAX_BATCH_SIZE = 64
USE_BUFFER_CONV = False
_PADDING_MODE = "centered"   # gin-bindable default of get_padding.mode (configs/causal.gin:5)


def use_cached_conv(state: bool):
    global USE_BUFFER_CONV
    USE_BUFFER_CONV = bool(state)


def set_padding_mode(mode: str):
    """Stand-in for the gin binding `cc.get_padding.mode = 'causal'`."""
    global _PADDING_MODE
    assert mode in ("centered", "causal")
    _PADDING_MODE = mode


def chunk_process(f, x, N):
    x = torch.split(x, x.shape[-1] // N, -1)
    return torch.cat([f(_x) for _x in x], -1)


def get_padding(kernel_size, stride=1, dilation=1, mode=None):
    """'same' padding; `stride` is accepted and ignored, as upstream."""
    if mode is None:
        mode = _PADDING_MODE
    if kernel_size == 1:
        return (0, 0)
    p = (kernel_size - 1) * dilation + 1
    if mode == "centered":
        return ((p - 1) // 2, p // 2)
    if mode == "causal":
        return (p // 2 + (p - 1) // 2, 0)
    raise ValueError(f"padding mode {mode} is not valid")


class CachedSequential(nn.Sequential):
    """Sequential that forwards the delay of its last delay-bearing child."""

    def __init__(self, *args, **kwargs):
        cumulative_delay = kwargs.pop("cumulative_delay", 0)
        stride = kwargs.pop("stride", 1)
        super().__init__(*args, **kwargs)
        self.cumulative_delay = int(cumulative_delay) * stride
        last = 0
        for i in range(1, len(self) + 1):
            d = getattr(self[-i], "cumulative_delay", None)
            if d is not None:
                last = d
                break
        self.cumulative_delay += last


class Sequential(CachedSequential):
    pass


class CachedPadding1d(nn.Module):
    """Left-pads with the tail of the previous call (zeros at first call)."""

    def __init__(self, padding, crop=False):
        super().__init__()
        self.initialized = 0
        self.padding = padding
        self.crop = crop

    @torch.no_grad()
    def init_cache(self, x):
        b, c, _ = x.shape
        self.register_buffer("pad", torch.zeros(MAX_BATCH_SIZE, c, self.padding).to(x))
        self.initialized += 1

    def forward(self, x):
        if not self.initialized:
            self.init_cache(x)
        if self.padding:
            x = torch.cat([self.pad[:x.shape[0]], x], -1)
            self.pad[:x.shape[0]].copy_(x[..., -self.padding:])
            if self.crop:
                x = x[..., :-self.padding]
        return x


def _split_padding(padding):
    if isinstance(padding, int):
        return padding, 2 * padding
    return padding[1], padding[0] + padding[1]


class CachedConv1d(nn.Conv1d):

    def __init__(self, *args, **kwargs):
        padding = kwargs.get("padding", 0)
        cd = kwargs.pop("cumulative_delay", 0)
        kwargs["padding"] = 0
        super().__init__(*args, **kwargs)
        r_pad, total = _split_padding(padding)
        s = self.stride[0]
        stride_delay = (s - ((r_pad + cd) % s)) % s
        self.cumulative_delay = (r_pad + stride_delay + cd) // s
        self.cache = CachedPadding1d(total)
        self.downsampling_delay = CachedPadding1d(stride_delay, crop=True)

    def forward(self, x):
        x = self.downsampling_delay(x)
        x = self.cache(x)
        return F.conv1d(x, self.weight, self.bias, self.stride, 0, self.dilation, self.groups)


class CachedConvTranspose1d(nn.ConvTranspose1d):

    def __init__(self, *args, **kwargs):
        cd = kwargs.pop("cumulative_delay", 0)
        super().__init__(*args, **kwargs)
        self.initialized = 0
        self.cumulative_delay = self.padding[0] + cd * self.stride[0]

    @torch.no_grad()
    def init_cache(self, x):
        b, c, _ = x.shape
        self.register_buffer("cache", torch.zeros(MAX_BATCH_SIZE, c, 2 * self.padding[0]).to(x))
        self.initialized += 1

    def forward(self, x):
        x = F.conv_transpose1d(x, self.weight, None, self.stride, 0,
                               self.output_padding, self.groups, self.dilation)
        if not self.initialized:
            self.init_cache(x)
        p2 = 2 * self.padding[0]
        x[..., :p2] += self.cache[:x.shape[0]]
        self.cache[:x.shape[0]].copy_(x[..., -p2:])
        x = x[..., :-p2]
        if self.bias is not None:
            x = x + self.bias.unsqueeze(-1)
        return x


class _PlainConvTranspose1d(nn.ConvTranspose1d):

    def __init__(self, *args, **kwargs):
        kwargs.pop("cumulative_delay", 0)
        super().__init__(*args, **kwargs)
        self.cumulative_delay = 0


class _PlainConv1d(nn.Conv1d):

    def __init__(self, *args, **kwargs):
        self._pad = kwargs.get("padding", (0, 0))
        kwargs.pop("cumulative_delay", 0)
        kwargs["padding"] = 0
        super().__init__(*args, **kwargs)
        self.cumulative_delay = 0

    def forward(self, x):
        pad = self._pad
        if isinstance(pad, int):
            pad = (pad, pad)
        x = F.pad(x, tuple(pad))
        return F.conv1d(x, self.weight, self.bias, self.stride, 0, self.dilation, self.groups)


class _AlignBranchesCached(nn.Module):

    def __init__(self, *branches, delays=None, cumulative_delay=0, stride=1):
        super().__init__()
        self.branches = nn.ModuleList(branches)
        if delays is None:
            delays = [b.cumulative_delay for b in self.branches]
        max_delay = max(delays)
        self.paddings = nn.ModuleList(
            [CachedPadding1d(max_delay - d, crop=True) for d in delays])
        self.cumulative_delay = int(cumulative_delay * stride) + max_delay

    def forward(self, x):
        return [b(p(x)) for b, p in zip(self.branches, self.paddings)]


class _Branches(nn.Module):

    def __init__(self, *branches, delays=None, cumulative_delay=0, stride=1):
        super().__init__()
        self.branches = nn.ModuleList(branches)
        if delays is None:
            delays = [b.cumulative_delay for b in self.branches]
        self.cumulative_delay = int(cumulative_delay * stride) + max(delays)

    def forward(self, x):
        return [b(x) for b in self.branches]


# gin-bindable defaults `cc.Conv1d.bias = True`, `cc.ConvTranspose1d.bias = False`
# (rave/configs/v1.gin:33-34) are injected by the factories below.
_CONV_DEFAULTS = {"bias": True}
_CONVT_DEFAULTS = {"bias": False}


def Conv1d(*args, **kwargs):
    for k, v in _CONV_DEFAULTS.items():
        kwargs.setdefault(k, v)
    return (CachedConv1d if USE_BUFFER_CONV else _PlainConv1d)(*args, **kwargs)


def ConvTranspose1d(*args, **kwargs):
    for k, v in _CONVT_DEFAULTS.items():
        kwargs.setdefault(k, v)
    return (CachedConvTranspose1d if USE_BUFFER_CONV else _PlainConvTranspose1d)(*args, **kwargs)


def AlignBranches(*args, **kwargs):
    return (_AlignBranchesCached if USE_BUFFER_CONV else _Branches)(*args, **kwargs)
