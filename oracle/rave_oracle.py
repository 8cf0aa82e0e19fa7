"""TEST INFRASTRUCTURE ONLY -- standalone CPU restatement of RAVE's encode/decode path.

This is the oracle the CUDA path is checked against.  It is a functional restatement
(plain torch CPU ops on a flat {name: tensor} state_dict) of:

  rave/model.py:594-634        RAVE.encode / decode / forward
  rave/pqmf.py:13-89,234-284   PQMF design, CachedPQMF.forward / inverse
  rave/blocks.py:508-710       EncoderV2 / GeneratorV2 (+ DilatedUnit :84-113, Residual :32-46)
  rave/blocks.py:321-497       v1 Generator / Encoder (raspberry / onnx configs)
  rave/blocks.py:244-291       NoiseGeneratorV2 (+ core.py:66-67,95-129)
  rave/blocks.py:845-919       Snake, AdaptiveInstanceNormalization (eval transfer)
  rave/quantization.py:131-154,239-249,302-318   RVQ encode / decode
  cached_conv (un-vendored)    see oracle/cached_conv_restated.py for the module form

PINNING: tests/test_oracle_pins.py compares this file with the REAL reference modules
loaded by path (oracle/ref_loader.py) whenever /root/reference exists, and with the
committed tests/golden/*.npz (made by oracle/make_golden.py from the real modules)
everywhere.  The reference itself holds no numeric golden vectors.

It runs on the GPU box (no /root/reference needed) for smoke(), the -m gpu parity tests
and bench.py's cpu_baseline / --impl reference legs.  rave_b200/ never imports it.
"""
import math

import numpy as np
import torch
import torch.nn.functional as F

# ----------------------------------------------------------------------------- padding


def get_padding(k, dilation=1, causal=False):
    """cached_conv.get_padding (stride is ignored upstream)."""
    if k == 1:
        return (0, 0)
    p = (k - 1) * dilation + 1
    if causal:
        return (p // 2 + (p - 1) // 2, 0)
    return ((p - 1) // 2, p // 2)


# ----------------------------------------------------------------------------- PQMF design


def _kaiser_lowpass(wc, atten, N=None):
    """pqmf.py:55-70 with today's scipy names (firwin nyq=pi -> fs=2pi)."""
    from scipy.signal import firwin, kaiserord
    n_, beta = kaiserord(atten, wc / np.pi)
    n_ = 2 * (n_ // 2) + 1
    N = N if N is not None else n_
    return firwin(N, wc, window=("kaiser", beta), scale=False, fs=2 * np.pi)


def pqmf_prototype(atten, n_band):
    """pqmf.py:73-89: Kaiser prototype whose cutoff minimises the overlap objective."""
    from scipy.optimize import fmin

    def objective(wc):
        h = _kaiser_lowpass(wc, atten, None)
        g = np.convolve(h, h[::-1], "full")
        g = abs(g[g.shape[-1] // 2::2 * n_band][1:])
        return np.max(g)

    wc = fmin(objective, 1 / n_band, disp=0)[0]
    return _kaiser_lowpass(wc, atten, None)


def pqmf_bank(atten=100, n_band=16):
    """Returns (h, hk): prototype and cosine-modulated bank padded to a power of two.

    pqmf.py:32-52 (modulation), :20-23 (center pad), :201-203 (float32 cast first).
    """
    h = torch.from_numpy(pqmf_prototype(atten, n_band)).float()
    k = torch.arange(n_band).reshape(-1, 1)
    N = h.shape[-1]
    t = torch.arange(-(N // 2), N // 2 + 1)
    phase = (-1) ** k * math.pi / 4
    hk = 2 * h * torch.cos((2 * k + 1) * math.pi / (2 * n_band) * t + phase)
    nxt = 2 ** math.ceil(math.log2(hk.shape[-1]))
    pad = nxt - hk.shape[-1]
    hk = F.pad(hk, (pad // 2, pad // 2 + pad % 2))
    return h, hk


def pqmf_conv_weights(hk):
    """CachedPQMF.__init__ pqmf.py:239-263: analysis [M,1,513] and synthesis [M,M,33]."""
    M = hk.shape[0]
    wf = F.pad(hk, (0, 1)).unsqueeze(1) if hk.shape[-1] % 2 == 0 else hk.unsqueeze(1)
    hki = hk.flip(-1)
    hki = hki.reshape(M, -1, M).permute(2, 0, 1)          # "c (t m) -> m c t"
    if hki.shape[-1] % 2 == 0:
        hki = F.pad(hki, (0, 1))
    return wf.contiguous(), hki.contiguous()


def _reverse_half(x):
    """pqmf.py:13-17: flip the sign of odd bands at even sub-band times."""
    x = x.clone()
    x[..., 1::2, ::2] *= -1
    return x


# ----------------------------------------------------------------------------- stream state


class StreamState:
    """Per-layer left context, zero at first use (cached_conv.CachedPadding1d)."""

    def __init__(self):
        self.buf = {}

    def pad(self, key, x, P, crop=False):
        if P == 0:
            return x
        prev = self.buf.get(key)
        if prev is None:
            prev = x.new_zeros(x.shape[0], x.shape[1], P)
        x = torch.cat([prev, x], -1)
        self.buf[key] = x[..., -P:].clone()
        return x[..., :-P] if crop else x

    def overlap_add(self, key, y, P):
        prev = self.buf.get(key)
        if prev is None:
            prev = y.new_zeros(y.shape[0], y.shape[1], P)
        y = y.clone()
        y[..., :P] += prev
        self.buf[key] = y[..., -P:].clone()
        return y[..., :-P]

    def num_elements_per_stream(self):
        return sum(int(v.shape[1] * v.shape[2]) for v in self.buf.values())


# ----------------------------------------------------------------------------- primitives


def _folded(sd, prefix):
    """weight (weight-norm folded: w = g * v / ||v||, norm over dims != 0) and bias."""
    if prefix + "weight_g" in sd:
        v, g = sd[prefix + "weight_v"], sd[prefix + "weight_g"]
        nrm = v.reshape(v.shape[0], -1).norm(dim=1).reshape(-1, *([1] * (v.dim() - 1)))
        w = v * (g / nrm)
    else:
        w = sd[prefix + "weight"]
    return w, sd.get(prefix + "bias")


class _Ctx:
    def __init__(self, sd, causal, stream):
        self.sd, self.causal, self.stream = sd, causal, stream


def _conv(ctx, prefix, x, k, stride=1, dilation=1, pad=None, cd=0, groups=1):
    """cc.Conv1d in either mode.  Returns (y, cumulative_delay)."""
    w, b = _folded(ctx.sd, prefix)
    assert w.shape[-1] == k, (prefix, w.shape, k)
    pl, pr = pad if pad is not None else get_padding(k, dilation, ctx.causal)
    if ctx.stream is None:
        return F.conv1d(F.pad(x, (pl, pr)), w, b, stride, 0, dilation, groups), 0
    sdel = (stride - ((pr + cd) % stride)) % stride
    x = ctx.stream.pad(prefix + "downsampling_delay", x, sdel, crop=True)
    x = ctx.stream.pad(prefix + "cache", x, pl + pr)
    return F.conv1d(x, w, b, stride, 0, dilation, groups), (pr + sdel + cd) // stride


def _conv_transpose(ctx, prefix, x, r, cd=0):
    """cc.ConvTranspose1d(C, C', 2r, stride=r, padding=r//2), bias after the overlap-add."""
    w, b = _folded(ctx.sd, prefix)
    p = r // 2
    if ctx.stream is None:
        return F.conv_transpose1d(x, w, b, r, p), 0
    y = F.conv_transpose1d(x, w, None, r, 0)
    y = ctx.stream.overlap_add(prefix + "cache", y, 2 * p)
    if b is not None:
        y = y + b.unsqueeze(-1)
    return y, p + cd * r


def _act(ctx, prefix, x, kind):
    if kind == "snake":                                        # blocks.py:845-853
        a = ctx.sd[prefix + "alpha"]
        return x + (a + 1e-9).reciprocal() * (a * x).sin().pow(2)
    return F.leaky_relu(x, 0.2)


def _adain_update(target, source, num_updates):
    """AdaptiveInstanceNormalization.update (blocks.py:870-873): running mean over calls."""
    bs = source.shape[0]
    target[:bs] += (source - target[:bs]) / (num_updates + 1)


def _adain(ctx, prefix, x):
    """AdaptiveInstanceNormalization.forward in eval mode (blocks.py:893-919).  The statistics
    buffers and the learn_x / learn_y flags live in the state dict and are updated IN PLACE,
    like the module's registered buffers (driven by OracleRAVE.update_adain = export.py:248-265)."""
    sd = ctx.sd
    learn_x = float(sd[prefix + "learn_x"]) if prefix + "learn_x" in sd else 0.0
    learn_y = float(sd[prefix + "learn_y"]) if prefix + "learn_y" in sd else 0.0
    if learn_y:
        mean = x.mean(-1, keepdim=True)
        std = x.std(-1, keepdim=True)
        _adain_update(sd[prefix + "mean_y"], mean, sd[prefix + "num_update_y"])
        _adain_update(sd[prefix + "std_y"], std, sd[prefix + "num_update_y"])
        sd[prefix + "num_update_y"] += 1
        return x
    if learn_x:
        mean = x.mean(-1, keepdim=True)
        std = x.std(-1, keepdim=True)
        _adain_update(sd[prefix + "mean_x"], mean, sd[prefix + "num_update_x"])
        _adain_update(sd[prefix + "std_x"], std, sd[prefix + "num_update_x"])
        sd[prefix + "num_update_x"] += 1
    if float(sd[prefix + "num_update_x"]) and float(sd[prefix + "num_update_y"]):
        bs = x.shape[0]                                        # transfer, blocks.py:885-891
        x = (x - sd[prefix + "mean_x"][:bs]) / (sd[prefix + "std_x"][:bs] + 1e-5)
        x = x * sd[prefix + "std_y"][:bs] + sd[prefix + "mean_y"][:bs]
    return x


def _dilated_residual(ctx, prefix, x, k, d, act):
    """Residual(DilatedUnit): x + conv1(act(conv_k,d(act(x)))); identity delayed in stream."""
    p = prefix + "aligned.branches.0.net."
    h = _act(ctx, p + "0.", x, act)
    h, cd = _conv(ctx, p + "1.", h, k, dilation=d)
    h = _act(ctx, p + "2.", h, act)
    h, _ = _conv(ctx, p + "3.", h, 1)
    if ctx.stream is not None:
        x = ctx.stream.pad(prefix + "aligned.paddings.1", x, cd, crop=True)
    return h + x


# ----------------------------------------------------------------------------- v2 stacks


def _norm_dilations(dilations, ratios):
    if isinstance(dilations[0], int):
        return [list(dilations) for _ in ratios]
    return [list(d) for d in dilations]


def encoder_v2(ctx, prefix, x, cfg):
    """blocks.EncoderV2 (blocks.py:533-597); module indices follow its `net` list."""
    k, act = cfg["kernel_size"], cfg["activation"]
    i = 0
    x, _ = _conv(ctx, f"{prefix}net.{i}.", x, 2 * k + 1)
    i += 1
    for r, dil in zip(cfg["ratios"], _norm_dilations(cfg["dilations"], cfg["ratios"])):
        for d in dil:
            if cfg["adain"]:
                x = _adain(ctx, f"{prefix}net.{i}.", x)
                i += 1
            x = _dilated_residual(ctx, f"{prefix}net.{i}.", x, k, d, act)
            i += 1
        x = _act(ctx, f"{prefix}net.{i}.", x, act)
        i += 1
        x, _ = _conv(ctx, f"{prefix}net.{i}.", x, 2 * r, stride=r,
                     pad=get_padding(2 * r, 1, ctx.causal))
        i += 1
    x = _act(ctx, f"{prefix}net.{i}.", x, act)
    i += 1
    x, _ = _conv(ctx, f"{prefix}net.{i}.", x, k)
    return x


def mod_sigmoid(x):
    return 2 * torch.sigmoid(x) ** 2.3 + 1e-7                  # core.py:66-67


def noise_filter_taps(amp, target_size):
    """core.amp_to_impulse_response (core.py:95-116) in closed form.

    irfft of n real bins -> 2(n-1) taps, rolled to centre, Hann-windowed, zero-padded to
    target_size and rolled back.  Evaluated here as an explicit cosine sum instead of FFT.
    """
    n = amp.shape[-1]
    N = 2 * (n - 1)
    t = torch.arange(N, dtype=amp.dtype)
    kk = torch.arange(n, dtype=amp.dtype)
    wgt = torch.full((n,), 2.0, dtype=amp.dtype)
    wgt[0] = 1.0
    wgt[-1] = 1.0
    basis = torch.cos(2 * math.pi * kk[:, None] * t[None, :] / N) * wgt[:, None] / N
    ir = amp @ basis                                           # real irfft, imag = 0
    ir = torch.roll(ir, N // 2, -1)
    ir = ir * torch.hann_window(N, dtype=amp.dtype)
    ir = F.pad(ir, (0, int(target_size) - N))
    return torch.roll(ir, -(N // 2), -1)


def causal_block_convolve(signal, kernel):
    """core.fft_convolve (core.py:119-129): per-block linear convolution kept to the block.

    out[i] = sum_j signal[j] * K[n + i - j] with K = zero-left-padded kernel, i.e.
    out[i] = sum_{j<=i} signal[j] * kernel[i - j].
    """
    n = signal.shape[-1]
    out = torch.zeros_like(signal)                             # explicit loops: n is 8
    for i in range(n):
        for j in range(i + 1):
            out[..., i] += signal[..., j] * kernel[..., i - j]
    return out


def noise_generator_v2(ctx, prefix, x, cfg, noise):
    """blocks.NoiseGeneratorV2 (blocks.py:244-291) with the uniform noise injected.

    `noise` is U(-1,1) of shape [B, L/8, data_size, 8] (the reference draws it with an
    unseeded torch.rand_like, blocks.py:287, so parity needs it passed in).
    """
    ratios, bands, data = [2, 2, 2], 5, cfg["n_band"]
    act = cfg["activation"]
    h = x
    i = 0
    for n_, r in enumerate(ratios):
        h, _ = _conv(ctx, f"{prefix}net.{i}.", h, 2 * r, stride=r, pad=(r, 0))
        i += 1
        if n_ != len(ratios) - 1:
            h = _act(ctx, f"{prefix}net.{i}.", h, act)
            i += 1
    amp = mod_sigmoid(h - 5)
    amp = amp.permute(0, 2, 1)
    amp = amp.reshape(amp.shape[0], amp.shape[1], data, bands)
    target = int(np.prod(ratios))
    ir = noise_filter_taps(amp, target)
    out = causal_block_convolve(noise, ir).permute(0, 2, 1, 3)
    return out.reshape(out.shape[0], out.shape[1], -1)


def generator_v2(ctx, prefix, z, cfg, noise=None):
    """blocks.GeneratorV2 (blocks.py:631-707)."""
    k, act = cfg["kernel_size"], cfg["activation"]
    ratios = list(cfg["ratios"])[::-1]
    dils = _norm_dilations(cfg["dilations"], cfg["ratios"])[::-1]
    i = 0
    x, _ = _conv(ctx, f"{prefix}net.{i}.", z, k)
    i += 1
    for r, dil in zip(ratios, dils):
        x = _act(ctx, f"{prefix}net.{i}.", x, act)
        i += 1
        x, _ = _conv_transpose(ctx, f"{prefix}net.{i}.", x, r)
        i += 1
        for d in dil:
            if cfg["adain"]:
                x = _adain(ctx, f"{prefix}net.{i}.", x)
                i += 1
            x = _dilated_residual(ctx, f"{prefix}net.{i}.", x, k, d, act)
            i += 1
    x = _act(ctx, f"{prefix}net.{i}.", x, act)
    i += 1
    nz = 0.
    if cfg.get("noise", False):
        nz = noise_generator_v2(ctx, prefix + "noise_module.", x, cfg, noise)
        x, _ = _conv(ctx, prefix + "waveform_module.", x, 2 * k + 1)
    else:
        x, _ = _conv(ctx, f"{prefix}net.{i}.", x, 2 * k + 1)
    x, amp = x.split(x.shape[1] // 2, 1)                       # amplitude_modulation
    x = x * torch.sigmoid(amp)
    return torch.tanh(x + nz)


# ----------------------------------------------------------------------------- v1 stacks


def encoder_v1(ctx, prefix, x, cfg):
    """blocks.Encoder (blocks.py:436-497): conv7, [BN, LReLU, strided conv]*, LReLU, conv5."""
    sd = ctx.sd
    x, cd = _conv(ctx, f"{prefix}net.0.", x, 7)
    i = 1
    for r in cfg["ratios"]:
        p = f"{prefix}net.{i}."
        x = (x - sd[p + "running_mean"][None, :, None]) / torch.sqrt(
            sd[p + "running_var"][None, :, None] + 1e-5)
        x = x * sd[p + "weight"][None, :, None] + sd[p + "bias"][None, :, None]
        x = F.leaky_relu(x, 0.2)
        x, cd = _conv(ctx, f"{prefix}net.{i + 2}.", x, 2 * r + 1, stride=r,
                      pad=get_padding(2 * r + 1, 1, ctx.causal), cd=cd)
        i += 3
    x = F.leaky_relu(x, 0.2)
    x, _ = _conv(ctx, f"{prefix}net.{i + 1}.", x, 5, cd=cd)
    return x


V1_STACK_DILATIONS = [[1, 1], [3, 1], [5, 1]]                  # v1.gin:64-66


def _residual_stack_v1(ctx, prefix, x, k=3):
    """blocks.ResidualStack with kernel_sizes=[3]: one ResidualBlock of three ResidualLayers."""
    for l, dil in enumerate(V1_STACK_DILATIONS):
        p = f"{prefix}net.branches.0.net.{l}.net.aligned."
        h, cd = x, 0
        for j, d in enumerate(dil):
            h = F.leaky_relu(h, 0.2)
            h, cd = _conv(ctx, f"{p}branches.0.{2 * j + 1}.", h, k, dilation=d, cd=cd)
        if ctx.stream is not None:
            x = ctx.stream.pad(p + "paddings.1", x, cd, crop=True)
        x = h + x
    return x


def generator_v1(ctx, prefix, z, cfg):
    """blocks.Generator (blocks.py:321-419) with use_noise=False, loud_stride=1."""
    x, _ = _conv(ctx, f"{prefix}net.0.", z, 7)
    for i, r in enumerate(cfg["ratios"]):
        x = F.leaky_relu(x, 0.2)
        x, _ = _conv_transpose(ctx, f"{prefix}net.{1 + 2 * i}.net.1.", x, r)
        x = _residual_stack_v1(ctx, f"{prefix}net.{2 + 2 * i}.", x)
    xw, xl = x, x
    if ctx.stream is not None:
        dw = get_padding(7, 1, ctx.causal)[1]
        dl = get_padding(3, 1, ctx.causal)[1]
        m = max(dw, dl)
        xw = ctx.stream.pad(prefix + "synth.paddings.0", x, m - dw, crop=True)
        xl = ctx.stream.pad(prefix + "synth.paddings.1", x, m - dl, crop=True)
    wave, _ = _conv(ctx, prefix + "synth.branches.0.", xw, 7)
    loud, _ = _conv(ctx, prefix + "synth.branches.1.", xl, 3)
    return torch.tanh(wave) * mod_sigmoid(loud)


# ----------------------------------------------------------------------------- RVQ


def rvq_encode(sd, prefix, z, num_quantizers):
    """ResidualVectorQuantization.encode (quantization.py:302-311): int64 [B, Q, F].

    Per stage: dist = -(|x|^2 - 2 x E^T + |E|^2) in fp32, first maximal index
    (quantization.py:131-136), residual -= E[ind].
    """
    B, D, Fr = z.shape
    res = z.permute(0, 2, 1).reshape(-1, D)
    out = []
    for q in range(num_quantizers):
        E = sd[f"{prefix}layers.{q}._codebook.embed"]
        Et = E.t()
        dist = -(res.pow(2).sum(1, keepdim=True) - 2 * res @ Et + Et.pow(2).sum(0, keepdim=True))
        ind = dist.max(dim=-1).indices
        res = res - F.embedding(ind, E)
        out.append(ind.reshape(B, Fr))
    return torch.stack(out, 1)


def rvq_decode(sd, prefix, idx):
    """ResidualVectorQuantization.decode (quantization.py:313-318): [B,Q,F] -> [B,D,F]."""
    acc = 0.
    for q in range(idx.shape[1]):
        E = sd[f"{prefix}layers.{q}._codebook.embed"]
        acc = acc + F.embedding(idx[:, q], E).permute(0, 2, 1)
    return acc
