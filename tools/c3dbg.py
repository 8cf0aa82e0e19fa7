import os as _os, sys as _sys; _sys.path.insert(0, _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))))  # repo root
import os, torch
from rave_b200 import RAVE
from rave_b200.synth import golden_state_dict, recipe_input
_, sd = golden_state_dict("v2_c64")
m = RAVE.from_gin(["v2.gin", "causal.gin"], streaming=True, dtype=torch.bfloat16)
m.load_state_dict(sd)
B, n = 1024, 2048
m.alloc_streams(B, n)
x = torch.from_numpy(recipe_input((64, 1, n), 99)).cuda().repeat(B // 64, 1, 1).contiguous().bfloat16()
m(x); torch.cuda.synchronize()
os.environ["RB_TC_SR_DEBUG"] = "1"
m(x); torch.cuda.synchronize()
