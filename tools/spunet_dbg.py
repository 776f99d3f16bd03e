import os, sys, json
import numpy as np, torch
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
from ptv3_b200 import spunet as S, synth, ops
g = np.load(os.path.join(REPO, "tests/golden/spunet.npz"))
shapes = json.load(open(os.path.join(REPO, "tests/golden/param_shapes_spunet.json")))
W = {k: torch.from_numpy(v) for k, v in synth.make_weights(shapes, seed=0).items()}
m = S.SpUNetBase(4, 13, compute_dtype=torch.bfloat16)
m.load_state_dict(W, strict=False)
m = m.eval().cuda()
orig_p, orig_g = ops.subm_conv_packed, ops.subm_conv
def wrap(name, fn):
    def f(feat, *a, **k):
        nbr = a[2]
        print(name, "feat", tuple(feat.shape), feat.dtype, "nbr", tuple(nbr.shape), "max idx", int(nbr.max()), flush=True)
        out = fn(feat, *a, **k)
        torch.cuda.synchronize()
        return out
    return f
ops.subm_conv_packed = wrap("packed", orig_p)
ops.subm_conv = wrap("general", orig_g)
out = m(dict(grid_coord=torch.from_numpy(g["grid_coord"]).cuda(), feat=torch.from_numpy(g["feat"]).cuda(),
             offset=torch.from_numpy(g["offset"]).cuda()))
print("ok", out.shape)
