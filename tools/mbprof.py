"""Step time and per-op device time of the C2 forward (512 streams x 2^18 samples, bf16) for several stream
micro-batch sizes.  usage: python tools/mbprof.py 0 64 32 16   (0 = all streams at once)"""
import os as _os, sys as _sys; _sys.path.insert(0, _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))))  # repo root
import sys

import numpy as np
import torch

import bench
from rave_b200 import RAVE

B, T = int(_os.environ.get("RB_MB_STREAMS", 512)), 1 << 18
m = RAVE.from_gin(["v2.gin"], dtype=torch.bfloat16)
m.load_state_dict(bench._state_dict()[1])
x = torch.from_numpy(bench._synthetic_rows(np.arange(B), T)).cuda().bfloat16()
cols = []
for mb in [int(a) for a in sys.argv[1:]] or [0]:
    m.set_microbatch(mb)
    for _ in range(3):
        m(x)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        m(x)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    m.profile_ops(True)
    m(x)
    torch.cuda.synchronize()
    t = np.array(m.op_times_ms())
    m.profile_ops(False)
    cols.append((mb, ms, t))
print("microbatch   " + " ".join("%8d" % c[0] for c in cols))
print("step ms      " + " ".join("%8.3f" % c[1] for c in cols))
print("sum op ms    " + " ".join("%8.3f" % c[2].sum() for c in cols))
for i, o in enumerate(m.spec.ops):
    if any(c[2][i] > 0 for c in cols):
        print("%2d %4d->%4d k%d s%d il%d  " % (i, o.c_in, o.c_out, o.k, o.stride, o.ileave) + " ".join("%8.3f" % c[2][i] for c in cols))
