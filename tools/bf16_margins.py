import sys, torch, numpy as np
import os; R = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, 'tests'))
import tests.test_gpu_parity as T
for name in T.GIN:
    m, meta, a = T._model(name, dtype=torch.bfloat16)
    x = torch.from_numpy(T.golden_input(a)).cuda().bfloat16()
    if meta["cfg"].get("noise"):
        from oracle.weights_recipe import recipe_input
        B, _, TT = x.shape
        m.set_noise(torch.from_numpy(recipe_input((B, TT // 16 // 8, 16, 8), int(a["noise_seed"].item()), uniform=True)).cuda())
    z = m.encode(x); y = m.decode(torch.from_numpy(a["z"]).cuda().bfloat16()); y2 = m(x)
    print("%-22s z %.4f  dec %.4f  fwd %.4f" % (name, T.rel_l2(z.float().cpu().numpy(), a["z"]), T.rel_l2(y.float().cpu().numpy(), a["y"]), T.rel_l2(y2.float().cpu().numpy(), a["y"])))
