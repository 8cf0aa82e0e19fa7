import os as _os, sys as _sys; _sys.path.insert(0, _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))))  # repo root
import torch, numpy as np
from rave_b200 import RAVE
from rave_b200.synth import golden_state_dict, recipe_input
_, sd = golden_state_dict("v2_c64")
m = RAVE.from_gin(["v2.gin", "causal.gin"], streaming=True, dtype=torch.bfloat16)
m.load_state_dict(sd)
B, n = 1024, 2048
m.alloc_streams(B, n)
x = torch.from_numpy(recipe_input((64, 1, n), 99)).cuda().repeat(B // 64, 1, 1).contiguous().bfloat16()
for _ in range(5): m(x)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20): m(x)
e1.record(); torch.cuda.synchronize()
print("step ms", e0.elapsed_time(e1) / 20)
m.profile_ops(True)
for _ in range(10): m(x)
t = np.array(m.op_times_ms()) / 10
print("sum of op times ms", t.sum(), "ops", len(t))
order = np.argsort(-t)[:14]
for i in order:
    o = m.spec.ops[i]
    print(i, o.name if hasattr(o, "name") else "", o.kind, o.c_in, o.c_out, o.k, o.stride, o.ileave, "%.3f ms" % t[i])
