"""PQMF filter-bank design (host side, once per model).

Same design procedure as the reference's rave/pqmf.py:20-89,196-263: a Kaiser low-pass
prototype whose cutoff minimises the band-overlap objective, cosine-modulated into M
bands, centre-padded to a power of two, then laid out as the analysis conv weight
[M,1,K+1] and the polyphase synthesis conv weight [M,M,K/M+1].  Uses today's scipy
names (`firwin(fs=2*pi)` for the removed `nyq=pi`).  A loaded state_dict overrides these
buffers (`pqmf.forward_conv.weight`, `pqmf.inverse_conv.weight`).
"""
import functools
import math

import numpy as np


def _lowpass(wc, atten):
    from scipy.signal import firwin, kaiserord
    n, beta = kaiserord(atten, wc / np.pi)
    n = 2 * (n // 2) + 1
    return firwin(n, wc, window=("kaiser", beta), scale=False, fs=2 * np.pi)


def _overlap_objective(wc, atten, n_band):
    h = _lowpass(wc, atten)
    g = np.convolve(h, h[::-1], "full")
    g = np.abs(g[g.shape[-1] // 2::2 * n_band][1:])
    return np.max(g)


@functools.lru_cache(maxsize=8)
def design(attenuation=100, n_band=16):
    """Returns float32 arrays (h, hk[M,K], w_analysis[M,1,K+1], w_synthesis[M,M,K/M+1])."""
    from scipy.optimize import fmin
    power = math.log2(n_band)
    if power != math.floor(power):
        raise ValueError("when using the polyphase algorithm, n_band must be a power of 2")
    wc = fmin(lambda w: _overlap_objective(w, attenuation, n_band), 1 / n_band, disp=0)[0]
    # the reference casts the prototype to float32 BEFORE modulating (pqmf.py:201-202) and
    # evaluates the cosine in float32 torch arithmetic; do the same with torch so the bank
    # is bit-identical where possible.
    import torch
    h = torch.from_numpy(_lowpass(wc, attenuation)).float()
    k = torch.arange(n_band).reshape(-1, 1)
    N = h.shape[-1]
    t = torch.arange(-(N // 2), N // 2 + 1)
    hk = 2 * h * torch.cos((2 * k + 1) * math.pi / (2 * n_band) * t + (-1) ** k * math.pi / 4)
    size = 2 ** math.ceil(math.log2(N))
    pad = size - N
    hk = torch.nn.functional.pad(hk, (pad // 2, pad // 2 + pad % 2))
    wa = hk
    if wa.shape[-1] % 2 == 0:
        wa = torch.nn.functional.pad(wa, (0, 1))
    wa = wa.unsqueeze(1)
    ws = hk.flip(-1).reshape(n_band, -1, n_band).permute(2, 0, 1)
    if ws.shape[-1] % 2 == 0:
        ws = torch.nn.functional.pad(ws, (0, 1))
    return (h.numpy().copy(), hk.numpy().copy(),
            np.ascontiguousarray(wa.numpy()), np.ascontiguousarray(ws.numpy()))
