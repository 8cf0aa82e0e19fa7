"""Reference state_dict -> folded fp32 tensors for rb_load_weights.

Accepts the key names of a reference checkpoint, with or without
weight-norm pairs (`weight_g`/`weight_v` vs `weight`; scripts/export.py:598-600 removes
them before export).  Folding done here, once, at load time:
  * weight norm  w = g * v / ||v||, norm over every dim but 0 (torch weight_norm default;
    for ConvTranspose1d dim 0 is the INPUT channel) -- rave/blocks.py:17-24;
  * eval-mode BatchNorm1d of the v1 encoder (blocks.py:445-446) into the conv before it;
  * ConvTranspose1d(C, C', 2r, stride r, padding r//2) into r phase-convs (netspec.py);
  * AdaIN transfer (blocks.py:885-891) into a per-(row, channel) scale/shift.
"""
import numpy as np

from . import capi
from .netspec import OP_AFFINE, OP_CONCAT, OP_CONV, OP_PQMF_FWD, OP_PQMF_INV, NetSpec


def _np(t):
    if hasattr(t, "detach"):
        t = t.detach().to("cpu").float().numpy()
    return np.ascontiguousarray(np.asarray(t), dtype=np.float32)


def _folded(sd, prefix, norm=True):
    if prefix + "weight_g" in sd:
        # fp32 like the reference: weight_norm computes g * v / ||v|| in the parameter dtype
        v, g = _np(sd[prefix + "weight_v"]), _np(sd[prefix + "weight_g"])
        nrm = np.sqrt((v.reshape(v.shape[0], -1) ** 2).sum(1, dtype=np.float32))
        w = v * (g / nrm.reshape((-1,) + (1,) * (v.ndim - 1)))
    elif prefix + "weight" in sd:
        w = _np(sd[prefix + "weight"])
    else:
        raise KeyError(f"state_dict has neither {prefix}weight nor {prefix}weight_g/_v")
    b = _np(sd[prefix + "bias"]) if prefix + "bias" in sd else None
    return np.ascontiguousarray(w, dtype=np.float32), b


def _phase_pack(w, r, p, k_taps):
    """ConvTranspose1d weight [c_in][c_out][2r] -> equivalent conv weight [r*c_out][c_in][k_taps].

    Output phase ph of y[t*r + ph] reads input offsets m (tap index m + 1) with kernel tap
    j = ph + p - r*m (netspec._Builder.conv_transpose)."""
    c_in, c_out, k = w.shape
    assert k == 2 * r
    out = np.zeros((r * c_out, c_in, k_taps), dtype=np.float32)
    for ph in range(r):
        for tap in range(k_taps):
            m = tap - 1
            j = ph + p - r * m
            if 0 <= j < k:
                out[ph * c_out:(ph + 1) * c_out, :, tap] = w[:, :, j].T
    return out


def collect(spec: NetSpec, sd, pqmf_weights=None, strict=True, skipped=None, max_batch_size=None):
    """Returns a list of (op_index, role, float32 ndarray) for everything found in `sd`.
    `max_batch_size`: rows of the AdaIN statistics tables (cc.MAX_BATCH_SIZE at construction, blocks.py:860-866);
    tables with fewer rows are extended with the constructor's values (mean 0, std 1)."""
    out = []
    missing = []
    defaulted = []
    if spec.arch == "v1" and "decoder.warmed_up" in sd and "decoder.synth.branches.2.net.0.weight" in sd \
            and int(np.asarray(_np(sd["decoder.warmed_up"])).reshape(-1)[0]) != 0:
        raise NotImplementedError("v1 Generator with an active NoiseGenerator branch (warmed_up=1)")
    for i, op in enumerate(spec.ops):
        P = op.params
        try:
            if op.kind == OP_CONV and "conv" in P:
                w, b = _folded(sd, P["conv"])
                if "bn" in P:
                    bn = P["bn"]
                    gamma = _np(sd[bn + "weight"])
                    if strict or bn + "running_var" in sd:
                        var, mean = _np(sd[bn + "running_var"]), _np(sd[bn + "running_mean"])
                    else:
                        # parameter-only dicts (the EMA callback, train_rave.py:85): upstream's
                        # load_state_dict(strict=False) leaves the buffers as constructed (mean 0, var 1)
                        var, mean = np.ones_like(gamma), np.zeros_like(gamma)
                        defaulted.append(bn + "running_mean/var")
                    scale = gamma / np.sqrt(var + np.float32(1e-5))
                    w = w * scale[:, None, None]
                    b = (b - mean) * scale + _np(sd[bn + "bias"])
                out.append((i, capi.RB_W_CONV, w))
                if op.has_bias:
                    if b is None:
                        raise KeyError(P["conv"] + "bias")
                    out.append((i, capi.RB_W_BIAS, b))
            elif op.kind == OP_CONV and "convt" in P:
                w, b = _folded(sd, P["convt"])
                out.append((i, capi.RB_W_CONV, _phase_pack(w, P["r"], P["p"], op.k)))
                if op.has_bias:
                    if b is None:
                        raise KeyError(P["convt"] + "bias")
                    out.append((i, capi.RB_W_BIAS, b))
            elif op.kind in (OP_PQMF_FWD, OP_PQMF_INV):
                key = ("pqmf.forward_conv.weight" if op.kind == OP_PQMF_FWD else "pqmf.inverse_conv.weight")
                if key in sd:
                    w = _np(sd[key])
                elif pqmf_weights is not None:
                    w = pqmf_weights[0 if op.kind == OP_PQMF_FWD else 1]
                else:
                    raise KeyError(key)
                out.append((i, capi.RB_W_CONV, w))
            elif op.kind == OP_AFFINE:
                p = P["adain"]
                mx, sx = _np(sd[p + "mean_x"])[..., 0], _np(sd[p + "std_x"])[..., 0]
                my, sy = _np(sd[p + "mean_y"])[..., 0], _np(sd[p + "std_y"])[..., 0]
                if max_batch_size and mx.shape[0] < max_batch_size:
                    n = max_batch_size - mx.shape[0]
                    ext = lambda t, v: np.concatenate([t, np.full((n, t.shape[1]), v, np.float32)], 0)  # noqa: E731
                    mx, sx, my, sy = ext(mx, 0.), ext(sx, 1.), ext(my, 0.), ext(sy, 1.)
                # the four buffers + the two counters travel as they are (blocks.py:860-868): the engine
                # derives a = std_y / (std_x + 1e-5), b = mean_y - mean_x * a and can go on learning
                nx = float(_np(sd[p + "num_update_x"]).reshape(-1)[0])
                ny = float(_np(sd[p + "num_update_y"]).reshape(-1)[0])
                stats = np.concatenate([np.ascontiguousarray(t, np.float32).reshape(-1) for t in (mx, sx, my, sy)] +
                                       [np.array([nx, ny], np.float32)])
                out.append((i, capi.RB_W_ADAIN_STATS, stats))
            elif op.kind == OP_CONCAT:
                if "speaker" in sd:
                    out.append((i, capi.RB_W_COND, _np(sd["speaker"]).reshape(-1)))
            if "alpha" in P:
                out.append((i, capi.RB_W_ALPHA, _np(sd[P["alpha"]]).reshape(-1)))
        except KeyError as e:
            missing.append(str(e))
    if spec.rvq is not None:
        keys = [f"{spec.rvq['prefix']}layers.{q}._codebook.embed" for q in range(spec.rvq["num_quantizers"])]
        if all(k in sd for k in keys):
            out.append((0, capi.RB_W_CODEBOOK, np.stack([_np(sd[k]) for k in keys], 0)))
        else:
            missing.append("RVQ codebooks")
    if missing and strict:
        raise KeyError("missing keys in state_dict: " + ", ".join(missing[:8]) +
                       (" ..." if len(missing) > 8 else ""))
    if skipped is not None:                 # strict=False callers can see what was not loaded
        skipped.extend(missing + ["(defaulted) " + d for d in defaulted])
    return [(i, r, np.ascontiguousarray(a, dtype=np.float32)) for i, r, a in out]
