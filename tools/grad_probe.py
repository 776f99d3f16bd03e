"""Where does the bf16 training step's gradient error come from?  (run on the GPU box)
Compares parameter gradients of the bf16 step with the frozen reference, with selected module kinds kept in fp32."""
import json, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from ptv3_b200 import model as M, synth
from oracle import ptv3 as P
from make_golden_names import TRAIN_GRADS

G = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
g = np.load(os.path.join(G, "train.npz"))
shapes = json.load(open(os.path.join(G, "param_shapes.json")))
W = {k: torch.from_numpy(v) for k, v in synth.make_weights(shapes, seed=0).items()}


def run(tag, fp32_kinds=()):
    m = M.OffsetKeypointPTv3(backbone_conf=dict(P.BASE_CFG, drop_path=0.0, compute_dtype=torch.bfloat16))
    m.load_state_dict(W, strict=False)
    m = m.cuda().train()
    for mod in m.modules():
        if isinstance(mod, fp32_kinds):
            mod.compute_dtype = torch.float32
    n = len(g["coord"])
    torch.manual_seed(400)
    out = m(dict(coord=torch.from_numpy(g["coord"]).cuda(), grid_coord=torch.from_numpy(g["grid_coord"]).cuda(),
                 feat=torch.from_numpy(g["feat"]).cuda(), offset=torch.tensor([n]).cuda(), grid_size=0.02,
                 target=torch.from_numpy(g["target"]).cuda()))
    out["loss"].backward()
    p = dict(m.named_parameters())
    errs = []
    for i, k in enumerate(TRAIN_GRADS):
        ref = torch.from_numpy(g[f"grad_{i}"])
        errs.append(((p[k].grad.cpu() - ref).norm() / ref.norm()).item())
    print(tag, "loss", f"{out['loss'].item():.5f}", " ".join(f"{e:.3f}" for e in errs))


run("all bf16         ")
run("pool fp32        ", (M.SerializedPooling,))
run("pool+unpool fp32 ", (M.SerializedPooling, M.SerializedUnpooling))
run("blocks fp32      ", (M.Block,))
run("all fp32         ", (M.SerializedPooling, M.SerializedUnpooling, M.Block))
