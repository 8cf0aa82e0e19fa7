#!/usr/bin/env python
"""bench.py -- audio-seconds per wall-second of v2 encode+decode (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W [--impl reference] [--dtype bf16|f32]

A step = one pass of RAVE.forward (encode -> decode) over the synthetic batch of
BASELINE.json configs[1]: 512 streams x 2^18 samples @44.1 kHz, random-init v2 weights
(name-keyed recipe, rave_b200/synth.py).  Default (`--scaling strong`) is the literal
configs[1]: the 512 streams are SHARDED over the N GPUs (contiguous blocks) and
every step ends with the NCCL all-gather of the output waveforms, inside the timed region.
With N > 1 the same line also carries a `weak` record (512 streams per GPU, no data-path
collective).  Ranks bind to the CPU cores NVML reports as local to their GPU before any pinned
buffer is allocated.  Prints ONE JSON line on rank 0.

--impl reference: the reference's own CPU path (the oracle port of its PyTorch modules --
the reference itself is not importable) on the host cores, each step a
bounded sample (REF_STREAMS streams x 2^18 samples) of the same workload.  `cpu_baseline` of
the product arm is the same function on the same sample: one CPU number, not two.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

SR = 44100
STREAMS = 512
SAMPLES = 2 ** 18
METRIC = "audio-sec/sec v2 encode+decode"
REF_STREAMS = 32          # CPU sample: 32 of the 512 streams at the full 2^18 length


_OUT = None


def _emit(record):
    out = _OUT if _OUT is not None else sys.stdout
    out.write(json.dumps(record) + "\n")
    out.flush()


def _workload(B, T, streams, gathered, how=None):
    return (f"v2 encode+decode (BASELINE configs[1]), {streams} streams x {T} samples @44.1 kHz in all, {B} per GPU, " +
            (f"all-gather of every step's outputs inside the timed region, overlapped with the next step's compute "
             f"({how or 'NCCL'})" if gathered else "no data-path collective"))


def _peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=d["hbm_gbs"], tf_burst=d["bf16_tflops"], tf_sus=d["bf16_tflops_sustained"],
                    src="measured (MEASURED_PEAKS.json)")
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sus=1400.0, src="fallback (B200_PROFILING.md)")


def _state_dict():
    from rave_b200.synth import golden_state_dict
    return golden_state_dict("v2_c64")         # only the (name, shape) list is read from the fixture


def _synthetic_rows(rows, T, seed=1234):
    """Rows `rows` (global stream ids) of the synthetic batch, fp32 numpy [len(rows), 1, T]: one seeded
    pattern per id modulo 64 plus a per-stream offset, identical on every rank and for every N."""
    import numpy as np
    from rave_b200.synth import recipe_input
    base = recipe_input((64, 1, T), seed)
    rows = np.asarray(rows)
    return base[rows % 64] + (0.001 * rows).astype(np.float32).reshape(-1, 1, 1)


def bind_to_gpu_numa(local):
    """Pin this process to the CPU cores local to its GPU (NVML affinity mask) so that pinned host
    buffers allocated afterwards live on that socket.  Returns the number of cores (None: unchanged)."""
    try:
        import pynvml
        import torch
        pynvml.nvmlInit()
        pr = torch.cuda.get_device_properties(local)
        h = pynvml.nvmlDeviceGetHandleByPciBusId(
            f"{pr.pci_domain_id:08x}:{pr.pci_bus_id:02x}:{pr.pci_device_id:02x}.0".encode())
        n = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (n + 63) // 64)
        cpus = {64 * w + b for w, word in enumerate(words) for b in range(64) if (word >> b) & 1}
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            return len(cpus)
    except Exception:
        pass
    return None


class ClockSampler(threading.Thread):
    """Samples nvidia-smi clocks / throttle reasons while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop = index, [], False

    def run(self):
        while not self.stop:
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                      "-i", str(self.index)], capture_output=True, text=True, timeout=5).stdout
                self.rows.append([c.strip() for c in out.strip().split(",")])
            except Exception:
                pass
            time.sleep(0.1)

    def summary(self):
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 6 for n, v in zip(names, r[2:6]) if v == "Active"})
        mx = max((int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()), default=None)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": reasons,
                "samples": len(sm)}


# ------------------------------------------------------- CPU legs (the oracle as checker / baseline)
def _cpu_sample(T):
    return f"{REF_STREAMS} of the {STREAMS} streams x {T} samples (full length), fp32, torch CPU, all host threads"


def cpu_reference_pass(steps, warmup, T, budget_s=None):
    """The reference's CPU path (oracle port) on REF_STREAMS streams of the workload.  Returns
    (audio-s/s, seconds per pass, cores, passes timed)."""
    import torch
    from oracle.model import OracleRAVE
    meta, sd = _state_dict()
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    torch.set_num_threads(cores)
    o = OracleRAVE(meta["cfg"], sd)
    x = torch.from_numpy(_synthetic_rows(range(REF_STREAMS), T))
    for _ in range(max(warmup, 1)):
        o.forward(x)
    times = []
    t_end = time.time() + (budget_s or 1e9)
    for _ in range(steps):
        t0 = time.perf_counter()
        o.forward(x)
        times.append(time.perf_counter() - t0)
        if budget_s and time.time() > t_end and len(times) >= 3:
            break
    dt = sum(times) / len(times)
    return REF_STREAMS * T / SR / dt, dt, cores, len(times)


def parity_check(m, x, rows, first_stream, T):
    """Outside the timed region: rows `rows` of this rank's step output against OracleRAVE.forward (fp32
    CPU) on the same rows, held to BASELINE.json's stated bars: rel-L2 <= 1e-2 (bf16) / 1e-3 (fp32)."""
    import numpy as np
    import torch
    from oracle.model import OracleRAVE
    meta, sd = _state_dict()
    y = m(x)
    torch.cuda.synchronize()
    got = y[rows].float().cpu().numpy()
    xin = x[rows].float().cpu()                     # the exact (bf16-rounded) input rows the device saw
    want = OracleRAVE(meta["cfg"], sd).forward(xin).numpy()
    rel = [float(np.linalg.norm(got[i].astype(np.float64) - want[i]) / np.linalg.norm(want[i]))
           for i in range(len(rows))]
    tol = 1e-2 if x.dtype == torch.bfloat16 else 1e-3
    return {"rows": [int(first_stream + r) for r in rows], "rel_l2": rel, "tolerance": tol,
            "against": "OracleRAVE.forward (fp32 CPU port of the reference) on the same input rows, full length",
            "ok": bool(max(rel) < tol)}


def eager_gpu(T, streams=32):
    """Informational: 'what you get without this repo' -- the reference's PyTorch modules (oracle port, plain
    F.conv1d / F.conv_transpose1d -> cuDNN) on this GPU, fp32 and bf16 autocast, `streams` streams."""
    import torch
    from oracle.model import OracleRAVE
    meta, sd = _state_dict()
    out = {}
    try:
        o = OracleRAVE(meta["cfg"], sd)
        o.sd = {k: (v.cuda() if torch.is_tensor(v) else v) for k, v in o.sd.items()}
        x = torch.from_numpy(_synthetic_rows(range(streams), T)).cuda()
        for tag, ctx in (("f32", torch.autocast("cuda", enabled=False)),
                         ("bf16_autocast", torch.autocast("cuda", dtype=torch.bfloat16))):
            with ctx:
                for _ in range(2):
                    o.forward(x)
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(3):
                    o.forward(x)
                e1.record()
                torch.cuda.synchronize()
            out[tag] = streams * T / SR / (e0.elapsed_time(e1) / 3e3)
        out["unit"] = "audio-sec/sec"
        out["sample"] = f"{streams} streams x {T} samples, eager PyTorch (cuDNN) on the same GPU, oracle modules"
    except Exception as e:                      # informational leg: never fails the bench
        out["error"] = repr(e)[:200]
    torch.cuda.empty_cache()
    return out


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    val, dt, cores, n = cpu_reference_pass(args.steps, args.warmup, args.samples)
    sample = "each step = " + _cpu_sample(args.samples)
    strong = args.scaling == "strong"
    _emit(({
        "impl": "reference", "metric": METRIC, "value": val, "unit": "audio-sec/sec", "n_gpus": args.gpus,
        "steps": n, "warmup": args.warmup, "ms_per_step": 1e3 * dt,
        "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": _workload(args.streams // args.gpus if strong else args.streams, args.samples,
                                         args.streams if strong else args.streams * args.gpus,
                                         strong and args.gpus > 1),
                   "sample": sample, "same_config": False,
                   "note": "one CPU process (rank 0) whatever N is: a baseline, not a scaling curve"},
        "cpu_baseline": {"value": val, "unit": "audio-sec/sec", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": "audio-sec/sec", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


def run_ours(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    from rave_b200 import RAVE

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    numa_cores = bind_to_gpu_numa(local)          # before any pinned allocation
    if world > 1:
        # stdout carries exactly one JSON line: NCCL's own banner ("NCCL version ...") goes to stderr
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dtype = torch.bfloat16 if args.dtype == "bf16" else torch.float32
    esz = 2 if args.dtype == "bf16" else 4
    T = args.samples
    from rave_b200.shard import OverlappedGather, PeerCopyGather, gather_outputs, max_over_ranks, stream_shard
    gather_how = [None]
    m = RAVE.from_gin(["v2.gin"], device=f"cuda:{local}", dtype=dtype)
    m.load_state_dict(_state_dict()[1])
    if args.microbatch:
        m.set_microbatch(args.microbatch)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def make_input(first, B):
        return torch.from_numpy(_synthetic_rows(np.arange(first, first + B), T)).cuda().to(dtype)

    def timed(scaling, sampler=None):
        """K timed steps of one scaling mode (ms per step = max over ranks of the device time)."""
        first, B = stream_shard(args.streams, world, rank, scaling)
        x = make_input(first, B)
        gathered = world > 1 and scaling == "strong"
        gather = None
        if gathered:
            gather_how[0] = "nccl all_gather_into_tensor on a side stream"
            if args.gather == "p2p":
                try:
                    gather = PeerCopyGather(world, rank, (B, 1, T), dtype, torch.device("cuda", local))
                    gather_how[0] = "peer copies into symmetric buffers (copy engines) + one barrier per step, side stream"
                except Exception as e:                       # no symmetric memory on this box: the NCCL collective
                    sys.stderr.write(f"PeerCopyGather unavailable ({e!r}); using the NCCL all-gather\n")
            if gather is None:
                gather = OverlappedGather(world)

        def step():
            y = m(x)
            return gather(y) if gathered else y

        for _ in range(args.warmup):
            step()
        if gathered:
            gather.finish()
        barrier()
        if sampler is not None:
            sampler.start()
        l0 = m.kernel_launches()
        ev0.record()
        for _ in range(args.steps):
            step()
        if gathered:
            gather.finish()                    # the last step's gather is inside the timed region
        ev1.record()
        barrier()
        if sampler is not None:
            sampler.stop = True
        if gathered:
            # outside the timed region: what the gather delivered against the NCCL collective on the same outputs
            got = gather(m(x))
            gather.finish()
            torch.cuda.synchronize()
            want = gather_outputs(m(x), world)
            gather_how.append(bool(torch.equal(got, want)))
            del got, want
        ms = max_over_ranks(ev0.elapsed_time(ev1), "cuda") / args.steps
        return ms, m.kernel_launches() - l0, first, B, B * world, x, gathered

    sampler = ClockSampler(local) if rank == 0 else None
    ms, launches, first, B, streams, x, gathered = timed(args.scaling, sampler)
    audio_s = streams * T / SR
    value = audio_s / (ms / 1e3)
    weak = None
    if world > 1 and args.scaling == "strong" and not args.no_weak:
        del x
        wms, _, _, wB, wstreams, wx, _ = timed("weak")
        weak = {"value": wstreams * T / SR / (wms / 1e3), "unit": "audio-sec/sec", "ms_per_step": wms,
                "streams_per_gpu": wB, "streams": wstreams, "collective": "none"}
        del wx
        x = make_input(first, B)

    # ---- per-op device times (CUDA events around every op on the launching stream)
    macs = m.spec.macs_per_frame()
    frames = B * T // m.spec.samples_per_frame
    flops_step = 2.0 * (macs["enc"] + macs["dec"]) * frames
    m.profile_ops(True)
    m(x)
    torch.cuda.synchronize()
    op_ms = m.op_times_ms()
    m.profile_ops(False)
    # every conv / PQMF op is one launch of tc_conv_kernel (a fused DilatedUnit: one launch for two
    # ops -- the second op then reports 0 ms): algorithmic flops and HBM bytes of those launches
    slots = m.spec.slots

    def slot_bytes(sl):
        return slots[sl].per_frame * frames * slots[sl].channels * esz

    conv_ms = conv_flops = conv_bytes = 0.0
    n_launch = 0
    rows = []
    ops = list(m.spec.ops)
    for i, (op, t_ms) in enumerate(zip(ops, op_ms)):
        mac = 0
        if op.kind == 1:
            pf = slots[op.dst].per_frame
            mac = pf * (2 * op.c_in * (op.c_out // op.ileave) if op.ileave > 1 else op.c_in * op.c_out * op.k)
        elif op.kind in (2, 3):
            mac = slots[op.dst if op.kind == 2 else op.src].per_frame * op.k * op.c_out * (1 if op.kind == 2 else op.c_in)
        if op.kind in (1, 2, 3):
            conv_ms += t_ms
            conv_flops += 2.0 * mac * frames
            if t_ms > 0:
                n_launch += 1
                fused_next = i + 1 < len(ops) and ops[i + 1].kind in (1, 5) and op_ms[i + 1] == 0.0
                dst = ops[i + 1].dst if fused_next else op.dst
                conv_bytes += slot_bytes(op.src) + slot_bytes(dst)
                if op.res >= 0 and not fused_next:
                    conv_bytes += slot_bytes(op.res)
        rows.append({"name": op.name, "kind": op.kind, "c_in": op.c_in, "c_out": op.c_out, "k": op.k,
                     "stride": op.stride, "dil": op.dilation, "ileave": op.ileave, "ms": t_ms,
                     "gflop": 2e-9 * mac * frames, "out_mb": slot_bytes(op.dst) / 1e6})
    if args.dump_ops and rank == 0:
        os.makedirs(os.path.dirname(os.path.abspath(args.dump_ops)), exist_ok=True)
        json.dump(rows, open(args.dump_ops, "w"), indent=1)
    pk = _peaks()
    # achieved = algorithmic flops of one step (81.51 MFLOP per latent frame) / the measured step
    achieved = flops_step / (ms / 1e3) / 1e12
    traffic, traffic_src = None, None
    for name in ("r02_tc_conv_traffic.json", "r01_tc_conv_traffic.json"):
        tp = os.path.join(ROOT, "profiles", name)
        if os.path.exists(tp):
            traffic = json.load(open(tp)).get("dram_bytes_per_step")
            traffic_src = f"profiles/{name} (ncu dram__bytes_read + dram__bytes_write over the tc_conv / tc_du_pair launches of one step of this command, tools/traffic.py; an earlier run, not measured in this one)"
            break
    per_op_tf = conv_flops / (conv_ms / 1e3) / 1e12 if conv_ms > 0 else 0.0
    hbm_gbs = conv_bytes / (conv_ms / 1e3) / 1e9 if conv_ms > 0 else 0.0
    roofline = {"bound": "tensor", "achieved": achieved, "peak": pk["tf_sus"], "unit": "TFLOP/s",
                "frac": achieved / pk["tf_sus"], "traffic": traffic, "traffic_source": traffic_src,
                "kernel": "tc_conv_kernel / tc_du_pair_kernel: the %d launches of one step (%.2f of %.2f ms of per-op CUDA-event time); "
                          "achieved = algorithmic flops of the step / ms_per_step" % (n_launch, conv_ms, sum(op_ms)),
                "peak_source": pk["src"] + ", sustained bf16",
                "algorithmic_flops_per_step_per_gpu": flops_step,
                "layer_by_layer_bytes": conv_bytes, "fused_path_bytes": 2.0 * B * T * esz,
                "hbm": {"achieved": hbm_gbs, "peak": pk["hbm"], "unit": "GB/s", "frac": hbm_gbs / pk["hbm"],
                        "note": "layer-by-layer algorithmic bytes / summed per-op time"},
                "per_op_profile_frac": per_op_tf / pk["tf_sus"]}

    # ---- end to end through the public API with HOST buffers (pinned), H2D + D2H inside
    xh = x.float().cpu().pin_memory() if esz == 4 else x.cpu().pin_memory()
    yh = torch.empty_like(xh).pin_memory()
    e2e_steps = max(2, min(args.steps, 10))

    # raw copy bandwidth of this rank's pinned buffers, all ranks copying at once (names the e2e limiter)
    def copy_gbs(dst, src):
        dst.copy_(src, non_blocking=True)
        torch.cuda.synchronize()
        ev0.record()
        for _ in range(3):
            dst.copy_(src, non_blocking=True)
        ev1.record()
        torch.cuda.synchronize()
        return 3 * src.numel() * src.element_size() / (ev0.elapsed_time(ev1) / 1e3) / 1e9

    xd = torch.empty_like(x)
    barrier()
    h2d = copy_gbs(xd, xh)
    barrier()
    d2h = copy_gbs(yh, xd)
    del xd
    h2d_min = -max_over_ranks(-h2d, "cuda")         # the slowest rank sets the pace
    d2h_min = -max_over_ranks(-d2h, "cuda")

    # every step copies its input shard from pinned host memory and its output back; the
    # public host-buffer call pipelines blocks of streams over three streams (H2D | forward | D2H)
    chunks = args.host_chunks if B % args.host_chunks == 0 else 1

    def e2e_step():
        m.forward_host_async(xh, yh, chunks)

    e2e_step()
    m.host_sync()
    barrier()
    ev0.record()
    for _ in range(e2e_steps):
        e2e_step()
    m.host_join()                       # the timing stream waits for the last device->host copy
    ev1.record()
    barrier()
    e2e_val = e2e_steps * audio_s / (max_over_ranks(ev0.elapsed_time(ev1), "cuda") / 1e3)

    if rank == 0:
        sampler.join(timeout=2)
        out = {
            "metric": METRIC, "value": value, "unit": "audio-sec/sec", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
            "scaling": args.scaling, "vs_baseline": None, "dtype": args.dtype, "data": "synthetic",
            "config": {"workload": _workload(B, T, streams, gathered, gather_how[0]),
                       "l2": f"input {B * T * esz >> 20} MiB + layer activations (up to {B * T * esz * 4 >> 20} MiB each) "
                             "per GPU and step > 126 MB L2: no flush needed",
                       "weights": "random-init (name-keyed recipe), CAPACITY=64",
                       "numa": (f"rank bound to its GPU's {numa_cores} local cores before pinned allocation"
                                if numa_cores else "no NUMA binding (NVML affinity unavailable)")},
            "clocks": sampler.summary(),
            "e2e": {"value": e2e_val, "unit": "audio-sec/sec", "h2d_bytes_per_step": B * T * esz * world,
                    "d2h_bytes_per_step": B * T * esz * world, "steps": e2e_steps,
                    "h2d_gbs_per_rank_min": h2d_min, "d2h_gbs_per_rank_min": d2h_min},
            "gpu_launches": int(launches),
            "roofline": roofline,
        }
        if weak is not None:
            out["weak"] = weak
        if gathered:
            out["gather"] = {"how": gather_how[0], "equals_nccl_all_gather": gather_how[1] if len(gather_how) > 1 else None}
        if not args.no_cpu:
            # the oracle as the CHECKER: two rows of this rank's step output, outside the timed region
            out["parity_check"] = parity_check(m, x, [0, B - 1], first, T)
        if world == 1 and not args.no_cpu:
            val, dt, cores, n = cpu_reference_pass(20, 1, T, budget_s=20.0)
            out["cpu_baseline"] = {"value": val, "unit": "audio-sec/sec", "cores": cores, "kind": "port",
                                   "sample": _cpu_sample(T) + f", mean of {n} passes after 1 warm-up"}
            del x
            torch.cuda.empty_cache()
            out["eager_gpu"] = eager_gpu(T)
        _emit(out)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--dtype", default=os.environ.get("RAVE_BENCH_DTYPE", "bf16"), choices=["bf16", "f32"])
    ap.add_argument("--streams", type=int, default=STREAMS)
    ap.add_argument("--samples", type=int, default=SAMPLES)
    ap.add_argument("--microbatch", type=int, default=0)
    ap.add_argument("--host-chunks", type=int, default=2, help="stream blocks per step of the host-buffer (e2e) call")
    ap.add_argument("--scaling", default="strong", choices=["weak", "strong"],
                    help="strong (default, the literal BASELINE configs[1]): --streams in all, sharded over the ranks, "
                         "NCCL all-gather of every step's outputs inside the timed region; weak: --streams per GPU, "
                         "no data-path collective (identical at N = 1)")
    ap.add_argument("--gather", default="p2p", choices=["p2p", "nccl"],
                    help="strong scaling, N > 1: how every step's outputs reach all ranks (p2p: peer copies into "
                         "symmetric buffers on the copy engines; nccl: all_gather_into_tensor)")
    ap.add_argument("--no-weak", action="store_true", help="N > 1: skip the additional weak-scaling record")
    ap.add_argument("--no-cpu", action="store_true", help="skip the oracle legs (parity check, CPU baseline, eager GPU)")
    ap.add_argument("--dump-ops", default="", help="write per-op device times (JSON) to this file")
    args = ap.parse_args()
    # stdout carries exactly ONE line, the JSON record: whatever native libraries write to file descriptor 1
    # (NCCL prints its "NCCL version ..." banner there whatever NCCL_DEBUG_FILE says) is sent to stderr, and the
    # record goes out through a private copy of the original descriptor (_emit)
    global _OUT
    sys.stdout.flush()
    _OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
