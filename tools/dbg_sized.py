"""Debug: rel-L2 of the sized fixtures on both engines (python tools/dbg_sized.py [names...])."""
import os as _os, sys as _sys; _sys.path.insert(0, _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))))  # repo root
import sys
import torch
import tests.test_gpu_parity as T
from tests.util import bf16_floor, golden_input, rel_l2

for name in (sys.argv[1:] or ["v3_c64", "v3causal_c64", "discrete_c96", "v2_c64", "v2causal_c64"]):
    for dtype in (torch.float32, torch.bfloat16):
        try:
            m, meta, a = T._model(name, dtype=dtype)
            x = torch.from_numpy(golden_input(a)).cuda().to(dtype)
            z = m.encode(x)
            nl = m.latent_size
            y = m.decode(torch.from_numpy(a["z"]).cuda().to(dtype))
            y2 = m(x)
            print(name, dtype, "z %.3e  zlat %.3e  y %.3e  fwd %.3e" % (
                rel_l2(z.float().cpu().numpy(), a["z"]), rel_l2(z[:, :nl].float().cpu().numpy(), a["z"][:, :nl]),
                rel_l2(y.float().cpu().numpy(), a["y"]), rel_l2(y2.float().cpu().numpy(), a["y"])), flush=True)
        except Exception as e:
            print(name, dtype, "ERROR", repr(e)[:300], flush=True)
    if name.startswith("v3"):
        print(name, "bf16 floor (z, y, fwd):", bf16_floor(name), flush=True)
