"""Multi-GPU plumbing of the offline path: streams are independent, so N ranks run N
shards with no data-path exchange.  Used by bench.py; covered on CPU by a world_size-2 gloo test."""
import torch
import torch.distributed as dist


def stream_shard(streams, world, rank, scaling="weak"):
    """(first stream, stream count) of `rank`.  weak: every rank runs `streams` streams of its own
    (global ids rank*streams ..); strong: `streams` in all, split into contiguous blocks."""
    if scaling == "weak":
        return rank * streams, streams
    if streams % world:
        raise ValueError(f"{streams} streams do not split evenly over {world} ranks")
    n = streams // world
    return rank * n, n


def max_over_ranks(value, device="cpu"):
    """Device time of a step = the slowest rank's (all ranks get the same number back)."""
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def gather_outputs(y, world):
    """strong-scaling variant only: every rank receives all waveforms [world*B, 1, T]."""
    out = torch.empty((world * y.shape[0],) + tuple(y.shape[1:]), dtype=y.dtype, device=y.device)
    dist.all_gather_into_tensor(out, y.contiguous())
    return out


class OverlappedGather:
    """Final gather of the strong-scaling variant, taken off the critical path: the all-gather of
    step i runs on a side stream (NVLink, a few SMs) while step i+1 computes; two receive
    buffers alternate.  `finish()` joins the side stream (call before stopping the clock)."""

    def __init__(self, world):
        self.world = world
        self.side = torch.cuda.Stream()
        self.bufs = [None, None]
        self.i = 0

    def __call__(self, y):
        b = self.i & 1
        self.i += 1
        if self.bufs[b] is None:
            self.bufs[b] = torch.empty((self.world * y.shape[0],) + tuple(y.shape[1:]), dtype=y.dtype, device=y.device)
        self.side.wait_stream(torch.cuda.current_stream())
        y.record_stream(self.side)
        with torch.cuda.stream(self.side):
            dist.all_gather_into_tensor(self.bufs[b], y)
        return self.bufs[b]

    def finish(self):
        torch.cuda.current_stream().wait_stream(self.side)
