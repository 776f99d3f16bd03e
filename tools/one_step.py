"""Three warm-up forwards + one forward of the bench workload, nothing else (the command ncu wraps)."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import bench  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--scenes", type=int, default=16)
ap.add_argument("--points", type=int, default=100000)
ap.add_argument("--scale", type=float, default=4.1)
ap.add_argument("--steps", type=int, default=4)
a = ap.parse_args()
dev = torch.device("cuda", 0)
model, _ = bench.build_model(torch.bfloat16, dev)
raw = bench.make_raw_scenes(a.scenes, a.points, a.scale, seed0=0)
host, counts = bench.voxelize_on_device(raw, dev)
res = {k: v.to(dev) for k, v in host.items()}
for _ in range(a.steps):
    out = model(dict(res))["pred"]
torch.cuda.synchronize()
print("ok", sum(counts), float(out.float().abs().mean()))
