"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the speaker encoder `SpeakerRAVE`.

  rave/CombinedRave.py:200-328   SpeakerRAVE (the class v2.gin:98 / onnx.gin:20 bind)
  rave/CombinedRave.py:26-41     Residual, :77-108 DilatedUnit (LeakyReLU(.2), identity normalisation)
  rave/model.py:246-247          speaker = speaker_encoder(pqmf(utterance)).unsqueeze(2)

Functional form on a flat {name: tensor} state_dict, eval mode (BatchNorm uses its running
statistics; `bn6` is constructed upstream but never applied).  The encoder runs offline on a
whole utterance (cc.use_cached_conv is False when model.py:247 runs), so every conv is
zero-padded with cached_conv.get_padding's centred split -- `stride` is ignored there, so the
k = 2r strided convs pad (k - 1)//2 on the left and k//2 on the right like any other conv.

PINNING: tests/test_oracle_pins.py::test_speaker_oracle_matches_live_reference compares this
file with the real class loaded by path, and tests/golden/speaker.npz holds the reference's
output on recipe weights (oracle/make_golden.py).  rave_b200/ never imports this file.
"""
import torch
import torch.nn.functional as F

from .rave_oracle import _reverse_half, get_padding

# (prefix, dilation, stride ratio) of layer2..layer4, CombinedRave.py:216-263
_STAGES = (("layer2", 1, 4), ("layer3", 3, 4), ("layer4", 5, 2))


def _conv(sd, name, x, k, dilation=1, stride=1):
    pad = get_padding(k, dilation)
    return F.conv1d(F.pad(x, pad), sd[name + ".weight"], sd[name + ".bias"], stride=stride,
                    dilation=dilation)


def _bn(sd, name, x, eps=1e-5):
    shape = (1, -1) + (1,) * (x.dim() - 2)
    inv = torch.rsqrt(sd[name + ".running_var"] + eps)
    return (x - sd[name + ".running_mean"].view(shape)) * (inv * sd[name + ".weight"]).view(shape) \
        + sd[name + ".bias"].view(shape)


def speaker_embed(sd, x):
    """x: [B, 16, T] PQMF bands of the utterance -> [B, 256].  CombinedRave.py:297-328."""
    x = _conv(sd, "in_layer", x, 7)
    feats = []
    for prefix, d, r in _STAGES:
        # Residual(DilatedUnit): act, k3 dilated conv, act, 1x1 conv, + input      (:26-41, :77-108)
        h = _conv(sd, prefix + ".0.aligned.branches.0.net.1", F.leaky_relu(x, .2), 3, dilation=d)
        h = _conv(sd, prefix + ".0.aligned.branches.0.net.3", F.leaky_relu(h, .2), 1)
        x = x + h
        x = _conv(sd, prefix + ".2", F.leaky_relu(x, .2), 2 * r, stride=r)
        feats.append(x)
    x1, x2, x3 = feats
    x4 = _conv(sd, "cat_layer", x3, 1)
    x = torch.cat((F.max_pool1d(x2, 2), x3, x4), dim=1)                               # :305
    x = F.leaky_relu(_conv(sd, "out_layer", x, 3), .2)
    t = x.shape[-1]
    mean = x.mean(dim=2, keepdim=True)
    std = torch.sqrt(x.var(dim=2, keepdim=True).clamp(min=1e-4, max=1e4))            # unbiased
    g = torch.cat((x, mean.repeat(1, 1, t), std.repeat(1, 1, t)), dim=1)               # :312-315
    a = F.conv1d(g, sd["attention.0.weight"], sd["attention.0.bias"])
    a = _bn(sd, "attention.2", F.relu(a))
    a = F.conv1d(a, sd["attention.3.weight"], sd["attention.3.bias"])
    w = torch.softmax(a, dim=2)
    mu = torch.sum(x * w, dim=2)
    sg = torch.sqrt((torch.sum((x ** 2) * w, dim=2) - mu ** 2).clamp(min=1e-4, max=1e4))
    v = _bn(sd, "bn5", torch.cat((mu, sg), 1))
    return F.linear(v, sd["fc6.weight"], sd["fc6.bias"])


def pqmf_analysis(wf, audio):
    """CachedPQMF.forward offline (pqmf.py:269-273): strided conv with the analysis bank
    wf [M, 1, k], centred padding, then reverse_half.  audio: [B, 1, T] -> [B, M, T / M]."""
    pad = get_padding(wf.shape[-1])
    return _reverse_half(F.conv1d(F.pad(audio, pad), wf, stride=wf.shape[0]))


def speaker_from_audio(sd, wf, audio):
    """model.py:246-247: speaker_encoder(pqmf(utterance)) -> [B, 256]."""
    return speaker_embed(sd, pqmf_analysis(wf, audio))
