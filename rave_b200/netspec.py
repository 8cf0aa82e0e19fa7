"""gin bindings -> flat op program for the CUDA engine.

Mirrors what `rave.RAVE()` does when gin injects its constructor arguments
(reference rave/model.py:157-164, scripts/export.py:547-551): the functions below walk
the same constructor logic as blocks.EncoderV2.__init__ (blocks.py:533-589),
blocks.GeneratorV2.__init__ (:631-690), blocks.Encoder.__init__ (:436-493),
blocks.Generator.__init__ (:336-395), NoiseGeneratorV2.__init__ (:255-279) and
CachedPQMF.__init__ (pqmf.py:236-263), but instead of nn.Modules they emit `rb_op`
records (include/rave_b200.h) plus, per op, where its parameters live in the reference
state_dict (names as in SURVEY.md 3.4) and how they are folded.

Streaming (`cc.use_cached_conv(True)` before construction, export.py:543) is a
constructor argument here; it changes `off`, `hist` and `res_delay` of the ops exactly
as cached_conv's CachedConv1d / CachedConvTranspose1d / AlignBranches do.
"""
from dataclasses import dataclass, field
from typing import Dict, List, Optional

from . import gin_lite

OP_CONV, OP_PQMF_FWD, OP_PQMF_INV, OP_AFFINE, OP_HEAD_V2, OP_HEAD_V1, OP_NOISE_V2, OP_CONCAT = range(1, 9)
ACT_NONE, ACT_LEAKY, ACT_SNAKE = 0, 1, 2


def get_padding(kernel_size, stride=1, dilation=1, mode="centered"):
    """cached_conv.get_padding: 'same' padding; `stride` is accepted and ignored."""
    if kernel_size == 1:
        return (0, 0)
    p = (kernel_size - 1) * dilation + 1
    if mode == "centered":
        return ((p - 1) // 2, p // 2)
    if mode == "causal":
        return (p // 2 + (p - 1) // 2, 0)
    raise ValueError(f"padding mode {mode} is not valid")


@dataclass
class Slot:
    channels: int
    per_frame: int
    hist: int = 0


@dataclass
class Op:
    kind: int
    src: int
    dst: int
    res: int = -1
    src2: int = -1
    c_in: int = 0
    c_out: int = 0
    k: int = 1
    stride: int = 1
    dilation: int = 1
    off: int = 0
    ileave: int = 1
    pre_act: int = ACT_NONE
    res_delay: int = 0
    has_bias: int = 0
    n_band: int = 0
    src_ch_off: int = 0
    # where the parameters come from (consumed by rave_b200.weights)
    params: Dict[str, object] = field(default_factory=dict)
    name: str = ""


@dataclass
class NetSpec:
    slots: List[Slot]
    ops: List[Op]
    n_enc_ops: int
    enc_in: int
    enc_out: int
    dec_in: int
    dec_out: int
    samples_per_frame: int
    streaming: bool
    causal: bool
    cond_channels: int
    latent_size: int
    dec_latent_size: int
    n_band: int
    enc_bands: int
    sr: int
    noise_slot: int = -1
    rvq: Optional[dict] = None
    encoder_type: str = "variational"
    arch: str = "v2"
    attenuation: int = 100
    activation: str = "leaky_relu"
    adain: bool = False
    f0_channels: int = 0      # pitch channels the decoder takes behind the speaker channels (export.py:397)

    def macs_per_frame(self):
        """Algorithmic multiply-accumulates per latent frame."""
        total = {"enc": 0, "dec": 0}
        for i, op in enumerate(self.ops):
            side = "enc" if i < self.n_enc_ops else "dec"
            pf = self.slots[op.dst].per_frame
            if op.kind == OP_CONV:
                if op.ileave > 1:          # transposed conv: 2*Cin*Cout per output position
                    total[side] += pf * 2 * op.c_in * (op.c_out // op.ileave)
                else:
                    total[side] += pf * op.c_in * op.c_out * op.k
            elif op.kind == OP_PQMF_FWD:
                total[side] += self.slots[op.dst].per_frame * op.c_out * op.k
            elif op.kind == OP_PQMF_INV:
                total[side] += self.slots[op.src].per_frame * op.n_band * op.n_band * op.k
        return total


class _Builder:
    def __init__(self, streaming, mode):
        self.slots: List[Slot] = []
        self.ops: List[Op] = []
        self.streaming = streaming
        self.mode = mode

    def slot(self, channels, per_frame):
        self.slots.append(Slot(int(channels), int(per_frame)))
        return len(self.slots) - 1

    def need_hist(self, slot, n):
        if self.streaming:
            self.slots[slot].hist = max(self.slots[slot].hist, int(n))

    def conv(self, src, c_in, c_out, k, prefix, *, stride=1, dilation=1, pad=None, cd=0,
             pre_act=ACT_NONE, alpha=None, res=-1, res_delay=0, bias=True, bn=None,
             norm=True, dst=None, src_ch_off=0, name=""):
        """cc.Conv1d (offline `_pad` + conv, or CachedConv1d). Returns (dst, cumulative_delay)."""
        pl, pr = pad if pad is not None else get_padding(k, dilation=dilation, mode=self.mode)
        pf = self.slots[src].per_frame
        assert pf % stride == 0, "chunk would not stay integral through a strided layer"
        if dst is None:
            dst = self.slot(c_out, pf // stride)
        if self.streaming:
            sdel = (stride - ((pr + cd) % stride)) % stride
            off = pl + pr + sdel
            delay = (pr + sdel + cd) // stride
            self.need_hist(src, off)
        else:
            off, delay = pl, 0
        if res >= 0 and self.streaming:
            self.need_hist(res, res_delay)
        params = {"conv": prefix, "norm": norm, "bias": bias}
        if alpha is not None:
            params["alpha"] = alpha
        if bn is not None:
            params["bn"] = bn
        self.ops.append(Op(OP_CONV, src, dst, res=res, c_in=c_in, c_out=c_out, k=k, stride=stride,
                           dilation=dilation, off=off, pre_act=pre_act,
                           res_delay=res_delay if self.streaming else 0,
                           has_bias=int(bias), params=params, src_ch_off=src_ch_off,
                           name=name or prefix))
        return dst, delay

    def conv_transpose(self, src, c_in, c_out, r, prefix, *, pre_act=ACT_NONE, alpha=None,
                       bias=False, name=""):
        """cc.ConvTranspose1d(c_in, c_out, 2r, stride=r, padding=r//2) as r phase-convs.

        Offline: y[t*r+ph] = sum_j W[:, :, j] x[(t*r + ph + p - j)/r]  -> taps at input
        offsets m in {-1,0,+1}.  Streaming (CachedConvTranspose1d): unpadded transposed conv
        with an overlap-add of the previous call's tail == taps m in {-1,0} over an input
        with one sample of history.  Weight phase-packing happens in rave_b200.weights.
        """
        pf = self.slots[src].per_frame
        dst = self.slot(c_out, pf * r)
        p = r // 2
        if self.streaming:
            k, off = 2, 1
            self.need_hist(src, 1)
        else:
            k, off = 3, 1
        params = {"convt": prefix, "r": r, "p": 0 if self.streaming else p, "norm": True, "bias": bias}
        if alpha is not None:
            params["alpha"] = alpha
        self.ops.append(Op(OP_CONV, src, dst, c_in=c_in, c_out=c_out * r, k=k, off=off, ileave=r,
                           pre_act=pre_act, has_bias=int(bias), params=params, name=name or prefix))
        return dst, (p if self.streaming else 0)


def _act_kind(ref):
    if ref is None:
        return ACT_LEAKY                      # default `lambda dim: nn.LeakyReLU(.2)`
    name = ref.name if isinstance(ref, gin_lite.Ref) else str(ref)
    if name == "Snake":
        return ACT_SNAKE
    if name in ("LeakyReLU", "leaky_relu"):
        return ACT_LEAKY
    raise ValueError(f"activation {name} not supported")


def _norm_dilations(dilations, ratios):
    if isinstance(dilations[0], int):
        return [list(dilations) for _ in ratios]
    return [list(d) for d in dilations]


def _alpha(act, prefix):
    return prefix + "alpha" if act == ACT_SNAKE else None


def _residual_dilated_unit(b, x, dim, k, d, prefix, act):
    """Residual(DilatedUnit(dim, k, d)) -- blocks.py:32-46,84-113."""
    p = prefix + "aligned.branches.0.net."
    h, cd = b.conv(x, dim, dim, k, p + "1.", dilation=d, pre_act=act, alpha=_alpha(act, p + "0."))
    y, _ = b.conv(h, dim, dim, 1, p + "3.", pre_act=act, alpha=_alpha(act, p + "2."),
                  res=x, res_delay=cd)
    return y


def _build_encoder_v2(b, x, prefix, P, du_act):
    """blocks.EncoderV2.__init__ (blocks.py:533-589)."""
    k = P["kernel_size"]
    act = _act_kind(P.get("activation"))
    adain = P.get("adain") is not None
    if P.get("spectrogram") is not None or P.get("recurrent_layer") is not None:
        raise NotImplementedError("spectrogram / recurrent encoders (hybrid.gin) are outside the hot path")
    ratios = list(P["ratios"])
    i = 0
    x, _ = b.conv(x, P["data_size"], P["capacity"], 2 * k + 1, f"{prefix}net.{i}.")
    i += 1
    ch = P["capacity"]
    for r, dil in zip(ratios, _norm_dilations(P["dilations"], ratios)):
        for d in dil:
            if adain:
                b.ops.append(Op(OP_AFFINE, x, x, c_in=ch, c_out=ch,
                                params={"adain": f"{prefix}net.{i}."}, name=f"{prefix}net.{i}"))
                i += 1
            x = _residual_dilated_unit(b, x, ch, k, d, f"{prefix}net.{i}.", du_act)
            i += 1
        a_prefix = f"{prefix}net.{i}."
        i += 1
        out = ch * r if P.get("keep_dim", False) else ch * 2
        x, _ = b.conv(x, ch, out, 2 * r, f"{prefix}net.{i}.", stride=r,
                      pad=get_padding(2 * r, r, mode=b.mode), pre_act=act, alpha=_alpha(act, a_prefix))
        i += 1
        ch = out
    a_prefix = f"{prefix}net.{i}."
    i += 1
    x, _ = b.conv(x, ch, P["latent_size"] * P["n_out"], k, f"{prefix}net.{i}.",
                  pre_act=act, alpha=_alpha(act, a_prefix))
    return x


def _build_generator_v2(b, z, prefix, P, du_act, noise_P):
    """blocks.GeneratorV2.__init__ / forward (blocks.py:631-707)."""
    k = P["kernel_size"]
    act = _act_kind(P.get("activation"))
    adain = P.get("adain") is not None
    if P.get("recurrent_layer") is not None:
        raise NotImplementedError("recurrent generators (hybrid.gin) are outside the hot path")
    ratios = list(P["ratios"])[::-1]
    dils = _norm_dilations(P["dilations"], P["ratios"])[::-1]
    keep = P.get("keep_dim", False)
    ch = P["capacity"]
    for r in ratios:
        ch = ch * r if keep else ch * 2
    i = 0
    x, _ = b.conv(z, P["latent_size"], ch, k, f"{prefix}net.{i}.")
    i += 1
    for r, dil in zip(ratios, dils):
        out = ch // r if keep else ch // 2
        a_prefix = f"{prefix}net.{i}."
        i += 1
        x, _ = b.conv_transpose(x, ch, out, r, f"{prefix}net.{i}.", pre_act=act,
                                alpha=_alpha(act, a_prefix), bias=P["_convt_bias"])
        i += 1
        ch = out
        for d in dil:
            if adain:
                b.ops.append(Op(OP_AFFINE, x, x, c_in=ch, c_out=ch,
                                params={"adain": f"{prefix}net.{i}."}, name=f"{prefix}net.{i}"))
                i += 1
            x = _residual_dilated_unit(b, x, ch, k, d, f"{prefix}net.{i}.", du_act)
            i += 1
    a_prefix = f"{prefix}net.{i}."
    i += 1
    am = bool(P.get("amplitude_modulation", False))
    data = P["data_size"]
    wave_ch = data * 2 if am else data
    noise_slot, noise_out = -1, -1
    if P.get("noise_module") is not None:
        wprefix = prefix + "waveform_module."
        # NoiseGeneratorV2.__init__ (blocks.py:255-279): plain cc.Conv1d, no weight norm
        nact = _act_kind(noise_P.get("activation"))
        nr = list(noise_P["ratios"])
        chans = [ch] + (len(nr) - 1) * [noise_P["hidden_size"]] + [noise_P["data_size"] * noise_P["noise_bands"]]
        h, j = x, 0
        pa, al = act, _alpha(act, a_prefix)
        for n_, r in enumerate(nr):
            h, _ = b.conv(h, chans[n_], chans[n_ + 1], 2 * r, f"{prefix}noise_module.net.{j}.",
                          stride=r, pad=(r, 0), pre_act=pa, alpha=al, norm=False)
            j += 1
            if n_ != len(nr) - 1:
                pa, al = nact, _alpha(nact, f"{prefix}noise_module.net.{j}.")
                j += 1
        total = 1
        for r in nr:
            total *= r
        if noise_P["noise_bands"] != 5 or total != 8:
            raise NotImplementedError("noise synth kernel is specialised to 5 bands -> 8 taps (noise.gin)")
        pf = b.slots[x].per_frame
        noise_slot = b.slot(noise_P["data_size"], pf)          # injected U(-1,1), [B, L/8, data, 8]
        noise_out = b.slot(noise_P["data_size"], pf)
        b.ops.append(Op(OP_NOISE_V2, h, noise_out, src2=noise_slot, c_in=chans[-1],
                        c_out=noise_P["data_size"], k=total, name=prefix + "noise_module"))
    else:
        wprefix = f"{prefix}net.{i}."
    w, _ = b.conv(x, ch, wave_ch, 2 * k + 1, wprefix, pre_act=act, alpha=_alpha(act, a_prefix))
    y = b.slot(data, b.slots[w].per_frame)
    b.ops.append(Op(OP_HEAD_V2, w, y, src2=noise_out, c_in=wave_ch, c_out=data,
                    k=int(am), name=prefix + "head"))
    return y, noise_slot


def _build_encoder_v1(b, x, prefix, P):
    """blocks.Encoder.__init__ (blocks.py:436-493); BatchNorm (eval) folds into the conv before it."""
    if P.get("sample_norm", False) or P.get("repeat_layers", 1) != 1 or P.get("recurrent_layer") is not None:
        raise NotImplementedError("v1 encoder variants other than BatchNorm / repeat_layers=1")
    ratios = list(P["ratios"])
    cap = P["capacity"]
    x, cd = b.conv(x, P["data_size"], cap, 7, f"{prefix}net.0.", norm=False, bn=f"{prefix}net.1.")
    i = 1
    for n_, r in enumerate(ratios):
        cin, cout = 2 ** n_ * cap, 2 ** (n_ + 1) * cap
        bn = f"{prefix}net.{i + 3}." if n_ != len(ratios) - 1 else None
        x, cd = b.conv(x, cin, cout, 2 * r + 1, f"{prefix}net.{i + 2}.", stride=r,
                       pad=get_padding(2 * r + 1, r, mode=b.mode), cd=cd, pre_act=ACT_LEAKY,
                       norm=False, bn=bn)
        i += 3
    n_out = P["n_out"]
    if n_out != 1:
        raise NotImplementedError("grouped latent conv (n_out > 1)")
    x, _ = b.conv(x, cout, P["latent_size"] * n_out, 5, f"{prefix}net.{i + 1}.", cd=cd,
                  pre_act=ACT_LEAKY, norm=False)
    return x


def _build_generator_v1(b, z, prefix, P, stack_P, convt_bias):
    """blocks.Generator.__init__ / forward (blocks.py:336-419), use_noise False."""
    # use_noise=True (v1.gin:62): the noise branch only acts once the `warmed_up` buffer is set
    # (blocks.py:414); at construction it is 0, which is what tests/test_configs.py runs.  The
    # v1 NoiseGenerator itself is not implemented: weights.collect refuses a checkpoint whose
    # decoder.warmed_up is non-zero.
    if P.get("loud_stride", 1) != 1 or P.get("recurrent_layer") is not None:
        raise NotImplementedError("loud_stride != 1 / recurrent layer")
    ratios = list(P["ratios"])
    cap = P["capacity"]
    ks = list(stack_P.get("kernel_sizes", [3]))
    dl = [list(d) for d in stack_P.get("dilations_list", [[1, 1], [3, 1], [5, 1]])]
    if len(ks) != 1:
        raise NotImplementedError("ResidualStack with several kernel sizes")
    x, _ = b.conv(z, P["latent_size"], 2 ** len(ratios) * cap, 7, f"{prefix}net.0.")
    for i, r in enumerate(ratios):
        cin = 2 ** (len(ratios) - i) * cap
        cout = 2 ** (len(ratios) - i - 1) * cap
        x, _ = b.conv_transpose(x, cin, cout, r, f"{prefix}net.{1 + 2 * i}.net.1.",
                                pre_act=ACT_LEAKY, bias=convt_bias)
        for l, dil in enumerate(dl):
            p = f"{prefix}net.{2 + 2 * i}.net.branches.0.net.{l}.net.aligned.branches.0."
            h, cd = x, 0
            for j, d in enumerate(dil):
                last = j == len(dil) - 1
                if not last:
                    h, cd = b.conv(h, cout, cout, ks[0], f"{p}{2 * j + 1}.", dilation=d, cd=cd,
                                   pre_act=ACT_LEAKY)
                else:
                    pr = get_padding(ks[0], dilation=d, mode=b.mode)[1]
                    h, cd = b.conv(h, cout, cout, ks[0], f"{p}{2 * j + 1}.", dilation=d, cd=cd,
                                   pre_act=ACT_LEAKY, res=x, res_delay=cd + pr)
            x = h
    data = P["data_size"]
    dw = get_padding(7, mode=b.mode)[1]
    dlo = get_padding(3, mode=b.mode)[1]
    m = max(dw, dlo)
    # synth = AlignBranches(wave_gen, loud_gen): each branch input delayed by (max - own delay)
    w, _ = b.conv(x, cout, data, 7, prefix + "synth.branches.0.", cd=m - dw)
    lo, _ = b.conv(x, cout, 1, 3, prefix + "synth.branches.1.", cd=m - dlo)
    if b.streaming:
        # cd above only models the branch-alignment delay: add it to the read offset
        b.ops[-2].off += m - dw
        b.ops[-1].off += m - dlo
        b.need_hist(x, max(b.ops[-2].off, b.ops[-1].off))
    y = b.slot(data, b.slots[w].per_frame)
    b.ops.append(Op(OP_HEAD_V1, w, y, src2=lo, c_in=data, c_out=data, name=prefix + "head"))
    return y


# ----------------------------------------------------------------------------- top level

_ENCODER_WRAPPERS = {
    "VariationalEncoder": ("encoder", "variational"),
    "DiscreteEncoder": ("encoder_cls", "discrete"),
    "WasserteinEncoder": ("encoder_cls", "wasserstein"),
    "SphericalEncoder": ("encoder_cls", "spherical"),
}


def _eval_ref(cfg, value):
    """Evaluate `@core.get_augmented_latent_size()` (core.py:78-79: pitch term is commented out)."""
    if isinstance(value, gin_lite.Ref) and value.evaluate:
        if value.name == "get_augmented_latent_size":
            p = cfg.params(value.name, value.scope)
            return p["latent_size"] + p["noise_augmentation"] + p["speaker_embedding_size"]
        raise ValueError(f"cannot evaluate {value!r}")
    return value


def build_netspec(cfg: "gin_lite.Config", streaming: bool = False) -> NetSpec:
    """What `rave.RAVE()` would build under these bindings, as a flat program."""
    R = cfg.params("RAVE")
    for need in ("pqmf", "encoder", "decoder", "latent_size"):
        if need not in R:
            raise ValueError(f"rave.RAVE.{need} is not bound")
    if not R.get("enable_pqmf_encode", True) or not R.get("enable_pqmf_decode", True):
        raise NotImplementedError("enable_pqmf_encode/decode = False (hybrid.gin) is outside the hot path")
    mode = cfg.params("get_padding").get("mode", "centered")
    if mode not in ("centered", "causal"):
        raise ValueError(f"padding mode {mode} is not valid")
    norm_mode = cfg.params("normalization").get("mode", "identity")
    if norm_mode not in ("identity", "weight_norm"):
        raise Exception(f"Normalization mode {norm_mode} not supported")     # blocks.py:24
    # blocks.leaky_relu(dim, alpha) takes its slope from gin; the kernels (and the activated-storage inverse
    # x = min(z, 5z)) are written for the 0.2 every shipped config uses
    slope = cfg.params("leaky_relu").get("alpha", 0.2)
    if abs(float(slope) - 0.2) > 1e-9:
        raise NotImplementedError(f"blocks.leaky_relu.alpha = {slope}: only the default slope 0.2 is implemented")
    conv_bias = bool(cfg.params("Conv1d").get("bias", True))
    convt_bias = bool(cfg.params("ConvTranspose1d").get("bias", True))
    if not conv_bias:
        raise NotImplementedError("cc.Conv1d.bias = False")

    pq_name = R["pqmf"].name
    if pq_name not in ("CachedPQMF", "PQMF"):
        raise ValueError(f"rave.RAVE.pqmf = @{pq_name}: only pqmf.CachedPQMF and pqmf.PQMF exist (pqmf.py:179-284)")
    pq = cfg.params(pq_name, R["pqmf"].scope)
    n_band = int(pq["n_band"])
    atten = pq.get("attenuation", 100)
    # pqmf.PQMF (pqmf.py:179-231), the non-cached class no shipped config binds: its own symmetric padding
    # whatever cc.get_padding.mode says, the same analysis as CachedPQMF, and a synthesis that is
    # CachedPQMF.inverse advanced by one sub-band row (n_band samples; SURVEY 8(c) probe)
    plain_pqmf = pq_name == "PQMF"
    if plain_pqmf and streaming:
        raise NotImplementedError("pqmf.PQMF has no cached_conv mode: bind pqmf.CachedPQMF for streaming")
    if plain_pqmf and not pq.get("polyphase", True):
        raise NotImplementedError("pqmf.PQMF(polyphase=False): the classic_* path is not implemented")
    pq_mode = "centered" if plain_pqmf else mode

    enc_ref = R["encoder"]
    if enc_ref.name not in _ENCODER_WRAPPERS:
        raise ValueError(f"Encoder type {enc_ref.name} not supported")         # export.py:151-154
    key, enc_type = _ENCODER_WRAPPERS[enc_ref.name]
    W = cfg.params(enc_ref.name, enc_ref.scope)
    inner = W[key]
    EP = cfg.params(inner.name, inner.scope)
    dec_ref = R["decoder"]
    DP = cfg.params(dec_ref.name, dec_ref.scope)
    DP["latent_size"] = _eval_ref(cfg, DP["latent_size"])
    DP["_convt_bias"] = convt_bias
    ratios = list(EP["ratios"])
    spf = n_band
    for r in ratios:
        spf *= r
    enc_bands = 6                                   # model.py:613 `x[:, :6, :]`
    if EP["data_size"] != enc_bands:
        raise ValueError("encoder data_size must be 6: RAVE.encode feeds x[:, :6] (model.py:613)")
    cond = int(cfg.macro("SPEAKER_EMB_SIZE", 256))

    from .pqmf_design import design
    _, hk, wa, ws = design(atten, n_band)
    kf, ki = wa.shape[-1], ws.shape[-1]

    b = _Builder(streaming, mode)
    audio_in = b.slot(1, spf)
    xb = b.slot(enc_bands, spf // n_band)
    pl, pr = get_padding(kf, mode=pq_mode)
    off = pl + pr if streaming else pl
    b.need_hist(audio_in, off)
    b.ops.append(Op(OP_PQMF_FWD, audio_in, xb, c_in=1, c_out=enc_bands, k=kf, stride=n_band, off=off,
                    n_band=n_band, params={"pqmf_fwd": "pqmf.forward_conv."}, name="pqmf.forward_conv"))
    du_act = _act_kind(cfg.params("DilatedUnit").get("activation"))
    arch = "v2" if inner.name == "EncoderV2" else "v1"
    if inner.name == "EncoderV2":
        z = _build_encoder_v2(b, xb, "encoder.encoder.", EP, du_act)
    elif inner.name == "Encoder":
        z = _build_encoder_v1(b, xb, "encoder.encoder.", EP)
    else:
        raise ValueError(f"encoder class {inner.name} not supported")
    latent = int(EP["latent_size"]) * int(EP["n_out"])
    # scripts/export.py:393-397 `cat(z, emb, f0_norm)`: a decoder bound wider than latent + speaker takes
    # pitch channels behind the speaker channels (this fork's get_augmented_latent_size, core.py:78-79,
    # leaves pitch_embedding_size out, so the shipped configs have none)
    f0_ch = max(0, int(DP["latent_size"]) - latent - cond)
    zc = b.slot(latent + cond + f0_ch, 1)
    b.ops.append(Op(OP_CONCAT, z, zc, c_in=latent, c_out=latent + cond + f0_ch, params={"cond": "speaker"},
                    name="speaker_concat"))
    n_enc = len(b.ops)

    dec_in = b.slot(int(DP["latent_size"]), 1)
    noise_slot = -1
    if dec_ref.name == "GeneratorV2":
        NP = cfg.params("NoiseGeneratorV2") if DP.get("noise_module") is not None else {}
        y, noise_slot = _build_generator_v2(b, dec_in, "decoder.", DP, du_act, NP)
    elif dec_ref.name == "Generator":
        y = _build_generator_v1(b, dec_in, "decoder.", DP, cfg.params("ResidualStack"), convt_bias)
    else:
        raise ValueError(f"decoder class {dec_ref.name} not supported")
    if b.slots[y].channels != n_band:
        raise ValueError("decoder data_size must equal the number of PQMF bands")
    audio_out = b.slot(1, spf)
    pl, pr = get_padding(ki, mode=pq_mode)
    off = pl + pr if streaming else (pl - 1 if plain_pqmf else pl)
    b.need_hist(y, off)
    b.ops.append(Op(OP_PQMF_INV, y, audio_out, c_in=n_band, c_out=n_band, k=ki, off=off, n_band=n_band,
                    params={"pqmf_inv": "pqmf.inverse_conv."}, name="pqmf.inverse_conv"))

    rvq = None
    if enc_type == "discrete":
        Q = cfg.params(W["vq_cls"].name, W["vq_cls"].scope)
        rvq = dict(num_quantizers=int(Q["num_quantizers"]), dim=int(Q["dim"]),
                   codebook_size=int(Q["codebook_size"]), prefix="encoder.rvq.")
    act_name = "snake" if _act_kind(EP.get("activation")) == ACT_SNAKE else "leaky_relu"
    return NetSpec(slots=b.slots, ops=b.ops, n_enc_ops=n_enc, enc_in=audio_in, enc_out=zc,
                   dec_in=dec_in, dec_out=audio_out, samples_per_frame=spf, streaming=streaming,
                   causal=(mode == "causal"), cond_channels=cond, latent_size=int(R["latent_size"]),
                   dec_latent_size=int(DP["latent_size"]), n_band=n_band, enc_bands=enc_bands,
                   sr=int(R.get("sampling_rate", 44100)), noise_slot=noise_slot, rvq=rvq,
                   encoder_type=enc_type, arch=arch, attenuation=atten, activation=act_name,
                   adain=EP.get("adain") is not None, f0_channels=f0_ch)
