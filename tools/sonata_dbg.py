"""Debug aid: per-module rel-L2 between the fp32 and the bf16 PT-v3m2 forward (same inputs, same permutations)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch  # noqa: E402

import sonata_common as SC  # noqa: E402
from ptv3_b200 import model as M, ops  # noqa: E402

tag = sys.argv[1] if len(sys.argv) > 1 else "dec"
cfg = SC.CFGS[tag]


def run(dtype, native=True):
    m = M.DefaultSegmentorV2(cfg["num_classes"], cfg["out_channels"],
                             backbone=dict(type="PT-v3m2", enable_flash=False, drop_path=0.0, **cfg["backbone"]))
    m.load_state_dict(SC.weights(tag), strict=False)
    m.backbone.set_compute_dtype(dtype)
    m = m.cuda().eval()
    feats = {}
    for name, mod in m.backbone.named_modules():
        if name.count(".") in (1, 2) and (name.startswith("enc.") or name.startswith("dec.")):
            mod.register_forward_hook(lambda mo, i, o, name=name: feats.__setitem__(name, o.feat.detach().float().clone())
                                      if hasattr(o, "feat") else None)
    c, g, f, off = SC.inputs(cfg["backbone"]["in_channels"])
    d = dict(coord=torch.from_numpy(c).cuda(), grid_coord=torch.from_numpy(g).cuda(), feat=torch.from_numpy(f).cuda(),
             offset=torch.from_numpy(off).cuda(), grid_size=0.02)
    ops.NATIVE_BLOCK = native
    torch.manual_seed(400)
    m(d)
    ops.NATIVE_BLOCK = True
    return feats


a = run(torch.float32)
for native in (True, False):
    b = run(torch.bfloat16, native)
    print("native" if native else "op-by-op")
    for k in a:
        print(f"  {k:28s} {((a[k] - b[k]).norm() / a[k].norm()).item():.3e}  rows {a[k].shape[0]}")
