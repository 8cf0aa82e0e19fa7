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


class PeerCopyGather:
    """The same gather without a collective kernel: every rank owns two symmetric receive buffers
    (torch.distributed._symmetric_memory: peer-mapped over NVLink), and pushes its block of output rows
    into every rank's buffer with plain device-to-device copies on a side stream -- copy engines, no SMs.
    The model's kernels are persistent grids of one CTA per SM: a collective kernel that holds even a few
    SMs while they run forces a second wave on every launch it overlaps (8 GPUs, 64 streams each: 2.00 ms
    per step with the NCCL all-gather against 1.61 ms for the same 64 streams without it).  One symmetric-memory
    barrier per step (a one-block kernel) tells every rank that all blocks have landed."""

    def __init__(self, world, rank, block_shape, dtype, device):
        import torch.distributed._symmetric_memory as symm
        self.world, self.rank = world, rank
        self.side = torch.cuda.Stream()
        rows = block_shape[0]
        full = (world * rows,) + tuple(block_shape[1:])
        self.bufs, self.hdls, self.mine = [], [], []
        for _ in range(2):
            t = symm.empty(full, dtype=dtype, device=device)
            h = symm.rendezvous(t, dist.group.WORLD)
            self.bufs.append(t)
            self.hdls.append(h)
            # this rank's rows inside every rank's buffer
            self.mine.append([h.get_buffer(p, full, dtype, 0)[rank * rows:(rank + 1) * rows] for p in range(world)])
        self.i = 0

    def __call__(self, y):
        b = self.i & 1
        self.i += 1
        self.side.wait_stream(torch.cuda.current_stream())
        y.record_stream(self.side)
        with torch.cuda.stream(self.side):
            for k in range(self.world):                      # staggered targets: no two ranks start on the same peer
                self.mine[b][(self.rank + k) % self.world].copy_(y, non_blocking=True)
            self.hdls[b].barrier(channel=0)
        return self.bufs[b]

    def finish(self):
        torch.cuda.current_stream().wait_stream(self.side)
