#!/usr/bin/env python
"""Secondary measurements for BASELINE.json configs[2..4] (bench.py carries the headline line):

  C3  v2 causal streaming, 2048-sample chunks, 1024 concurrent streams, state resident across chunks
  C4  discrete config: RVQ encode (bit-exact indices are a test) + decode-from-indices throughput
  C5  v3 (Snake + AdaIN) batched decode-only from latents

One JSON line per leg on stdout.  Synthetic weights (name-keyed recipe), CUDA-event timing.
Every leg runs at the stated size of its config: C3 v2 causal CAPACITY 64, C4 discrete CAPACITY 96 /
LATENT_SIZE 128 (discrete.gin:13-19), C5 v3 CAPACITY 64 with 512 streams (parameter lists of the sized
golden fixtures).
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
SR = 44100


def _sd(name):
    from rave_b200.synth import golden_state_dict
    return golden_state_dict(name)


def _time(fn, warmup, steps):
    import torch
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


def c0(args):
    """BASELINE configs[0]: the reference's own CPU-runnable case, one stream x 2^17 samples (latency-bound)."""
    import torch
    from rave_b200 import RAVE
    _, sd = _sd("v2_c64")
    out = []
    for dtype, tag in ((torch.bfloat16, "bf16"), (torch.float32, "f32")):
        m = RAVE.from_gin(["v2.gin"], dtype=dtype)
        m.load_state_dict(sd)
        x = (torch.rand(1, 1, 2 ** 17, device="cuda") - 0.5).to(dtype)
        l0 = m.kernel_launches()
        ms = _time(lambda: m(x), 5, args.steps)
        out.append({"config": "C0 v2 offline, 1 stream x 2^17 samples, CAPACITY=64 (BASELINE configs[0])", "dtype": tag,
                    "ms_per_step": ms, "value": 2 ** 17 / SR / (ms / 1e3), "unit": "audio-sec/sec",
                    "launches_per_step": (m.kernel_launches() - l0) // (5 + args.steps)})
        del m
    return out


def _c3_parity(sd, meta):
    """Outside every timed region: streams 0 and 1 over 8 chunks of 2048 samples on the bf16 engine against the
    CPU oracle's streaming run of the same chunks (the checker, as in bench.py's parity_check)."""
    import torch
    from oracle.model import OracleRAVE
    from rave_b200 import RAVE
    from rave_b200.synth import recipe_input
    m = RAVE.from_gin(["v2.gin", "causal.gin"], streaming=True, dtype=torch.bfloat16)
    m.load_state_dict(sd)
    m.alloc_streams(64, 2048)
    o = OracleRAVE(dict(meta["cfg"], causal=True), sd, streaming=True)
    x = torch.from_numpy(recipe_input((2, 1, 8 * 2048), 98))
    ys, yo = [], []
    for c in range(8):
        xc = x[..., c * 2048:(c + 1) * 2048].contiguous()
        ys.append(m(xc.cuda().bfloat16()).float().cpu())
        yo.append(o.forward(xc))
    ys, yo = torch.cat(ys, -1).double(), torch.cat(yo, -1).double()
    err = [float((ys[b] - yo[b]).norm() / yo[b].norm()) for b in range(2)]
    return {"streams": [0, 1], "chunks": 8, "rel_l2": err, "tolerance": 1e-2, "ok": max(err) < 1e-2,
            "against": "OracleRAVE streaming (fp32 CPU port of the reference), chunk by chunk"}


def c3(args):
    import torch
    from rave_b200 import RAVE
    from rave_b200.synth import recipe_input
    meta, sd = _sd("v2_c64")
    out = []
    for dtype, tag in ((torch.bfloat16, "bf16"), (torch.float32, "f32")):
        m = RAVE.from_gin(["v2.gin", "causal.gin"], streaming=True, dtype=dtype)
        m.load_state_dict(sd)
        B, n = args.streams, 2048
        m.alloc_streams(B, n)
        x = torch.from_numpy(recipe_input((64, 1, n), 99)).cuda().repeat(B // 64, 1, 1).contiguous().to(dtype)
        l0 = m.kernel_launches()
        ms = _time(lambda: m(x), 8, args.steps)
        # the same step replayed from a CUDA graph (what a serving loop would do)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            m(x)
        ms_graph = _time(g.replay, 8, args.steps)
        out.append({"config": "C3 v2 causal streaming, %d streams x 2048-sample chunks, CAPACITY=64" % B,
                    "ms_per_step_cuda_graph": ms_graph,
                    "dtype": tag, "ms_per_step": ms, "value": B * n / SR / (ms / 1e3), "unit": "audio-sec/sec",
                    "real_time_factor_per_stream": (n / SR) / (ms / 1e3),
                    "state_elements_per_stream": m.stream_state_elements(),
                    "launches_per_step": (m.kernel_launches() - l0) // (8 + args.steps)})
        del m
    out[0]["parity_check"] = _c3_parity(sd, meta)
    return out


def c4(args):
    import torch
    from rave_b200 import RAVE
    from rave_b200.synth import recipe_input
    meta, sd = _sd("discrete_c96")        # BASELINE configs[3] at its stated size (discrete.gin:13-19)
    m = RAVE.from_gin(["discrete.gin"], [], dtype=torch.bfloat16)
    m.load_state_dict(sd)
    B, T = args.batch, 2 ** 18
    x = torch.from_numpy(recipe_input((64, 1, T), 7)).cuda().repeat(B // 64, 1, 1).contiguous().bfloat16()
    z = m.encode(x)
    nl = m.latent_size
    zl = z[:, :nl].float().contiguous()
    idx = m.encoder.rvq.encode(zl)
    ms_enc = _time(lambda: m.encoder.rvq.encode(zl), 3, args.steps)

    def dec():
        zq = m.encoder.rvq.decode(idx)
        zz = z.clone()
        zz[:, :nl] = zq.to(z.dtype)
        return m.decode(zz)

    ms_dec = _time(dec, 3, args.steps)
    ms_fwd = _time(lambda: m(x), 3, args.steps)
    mf = m.spec.macs_per_frame()
    flops = 2.0 * (mf["enc"] + mf["dec"]) * B * (T // m.spec.samples_per_frame)
    return [{"config": "C4 discrete (RVQ %dx%d, dim %d), CAPACITY=96, %d streams x 2^18" %
                       (idx.shape[1], m.spec.rvq["codebook_size"], nl, B),
             "dtype": "bf16 net / fp32 RVQ", "rvq_encode_ms": ms_enc, "vectors": int(B * idx.shape[2]),
             "rvq_encode_Mvec_per_s": B * idx.shape[2] / ms_enc / 1e3,
             "decode_from_indices_ms": ms_dec, "value": B * T / SR / (ms_dec / 1e3), "unit": "audio-sec/sec (decode)",
             "forward_ms": ms_fwd, "forward_audio_sec_per_sec": B * T / SR / (ms_fwd / 1e3),
             "forward_TFLOPs": flops / ms_fwd / 1e9}]


def c5(args):
    import torch
    from rave_b200 import RAVE
    meta, sd = _sd("v3_c64")              # BASELINE configs[4] at its stated size (v3.gin: CAPACITY 64)
    m = RAVE.from_gin(["v3.gin"], [], dtype=torch.bfloat16)
    B, F = args.batch, 256
    # the checkpoint's AdaIN tables hold cc.MAX_BATCH_SIZE = 64 rows; rows 64.. get the constructor's values
    m.load_state_dict(sd, max_batch_size=B)
    g = torch.Generator(device="cuda").manual_seed(5)
    z = torch.randn(B, m.spec.dec_latent_size, F, generator=g, device="cuda").bfloat16()
    ms = _time(lambda: m.decode(z), 3, args.steps)
    T = F * m.spec.samples_per_frame
    flops = 2.0 * m.spec.macs_per_frame()["dec"] * B * F
    m.profile_ops(True)                    # per-op CUDA-event times of one more pass (outside the timed region)
    m.decode(z)
    t = m.op_times_ms()
    m.profile_ops(False)
    top = sorted(((t[i], m.spec.ops[i].name, m.spec.ops[i].kind) for i in range(len(t)) if t[i] > 0), reverse=True)
    by_kind = {}
    for ms_i, _, k in top:
        by_kind[k] = by_kind.get(k, 0.0) + ms_i
    print(json.dumps({"c5_ops_ms_top": [(round(a, 3), n) for a, n, _ in top[:14]],
                      "c5_ms_by_op_kind": {str(k): round(v, 3) for k, v in by_kind.items()}}), file=sys.stderr)
    return [{"config": "C5 v3 (Snake + AdaIN) decode-only, CAPACITY=64, %d streams x %d frames" % (B, F),
             "dtype": "bf16", "ms_per_step": ms, "value": B * T / SR / (ms / 1e3), "unit": "audio-sec/sec (decode)",
             "TFLOPs": flops / ms / 1e9}]


def c6(args):
    """Resampler pair (rave/resampler.py) and the speaker encoder (CombinedRave.py:200-328): the steps
    either side of the path.  Resampler: 512 streams x 2^19 samples at 2x the model rate; algorithmic
    bytes = input + output once."""
    import torch
    from rave_b200.resampler import Resampler
    from rave_b200.speaker import SpeakerRAVE
    rows = []
    B, T = args.batch, 2 ** 19
    for dtype, name, eb in ((torch.bfloat16, "bf16", 2), (torch.float32, "f32", 4)):
        r = Resampler(88200, 44100)
        x = (torch.rand(B, 1, T, device="cuda") - 0.5).to(dtype)
        y = r.to_model_sampling_rate(x)
        ms_d = _time(lambda: r.to_model_sampling_rate(x), 3, args.steps)
        ms_u = _time(lambda: r.from_model_sampling_rate(y), 3, args.steps)
        nbytes = B * (T + T // 2) * eb
        rows.append({"config": "C6 resampler ratio 2 (k=39 / 2x21), %d streams x 2^19 samples" % B, "dtype": name,
                     "to_model_ms": ms_d, "to_model_GBps": nbytes / ms_d / 1e6, "from_model_ms": ms_u,
                     "from_model_GBps": nbytes / ms_u / 1e6, "note": "times include the output allocation of the shim"})
        del x, y
    meta, sd = _sd("speaker")
    enc = SpeakerRAVE(sd)
    xb = torch.randn(1, 16, 8192, device="cuda")          # model.py:30: 131072 samples = 8192 band samples
    ms = _time(lambda: enc(xb), 3, args.steps)
    rows.append({"config": "C6 speaker encoder, 1 utterance x 131072 samples (16 x 8192 band samples)", "dtype": "f32",
                 "ms_per_step": ms, "launches_per_step": 20})
    return rows


def c7(args):
    """YIN pitch front-end (rave/pitch_utils.py get_pitch, 70-400 Hz, one frame per 1024-sample block) and the
    f0 -> one-hot channels, on the streams of the headline workload."""
    import torch
    from rave_b200 import pitch
    B, T = args.batch, 2 ** 18
    x = torch.rand(B, T, device="cuda") - 0.5
    f0 = pitch.get_pitch(x, 1024)
    ms = _time(lambda: pitch.get_pitch(x, 1024), 2, max(2, args.steps // 4))
    ms_oh = _time(lambda: pitch.f0_norm_one_hot(f0, channels_first=True), 3, args.steps)
    frames = f0.numel()
    return [{"config": "C7 YIN get_pitch (tau_max 630, frame 1260) + f0 one-hot, %d streams x 2^18 samples" % B,
             "dtype": "f32", "yin_ms": ms, "frames": frames, "Mframes_per_s": frames / ms / 1e3,
             "GFMA_per_s": frames * 630 * 1260 / ms / 1e6, "value": B * T / SR / (ms / 1e3), "unit": "audio-sec/sec (pitch)",
             "one_hot_ms": ms_oh}]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--legs", default="c0,c3,c4,c5,c6,c7")
    ap.add_argument("--steps", type=int, default=32)
    ap.add_argument("--streams", type=int, default=1024)
    ap.add_argument("--batch", type=int, default=512)
    args = ap.parse_args()
    for leg in args.legs.split(","):
        for row in {"c0": c0, "c3": c3, "c4": c4, "c5": c5, "c6": c6, "c7": c7}[leg](args):
            print(json.dumps(row), flush=True)


if __name__ == "__main__":
    main()
