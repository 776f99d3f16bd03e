"""A/B of host-side / debug-slot switches on the bench workload (CUDA events; development aid)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import bench  # noqa: E402
from ptv3_b200 import _lib, ops  # noqa: E402

dev = torch.device("cuda", 0)
lib = _lib.load()
model, _ = bench.build_model(torch.bfloat16, dev)
raw = bench.make_raw_scenes(16, 100000, 4.1, seed0=0)
host, counts = bench.voxelize_on_device(raw, dev)
res = {k: v.to(dev) for k, v in host.items()}


def run(label, steps=8):
    for _ in range(3):
        model(dict(res))
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(steps):
        model(dict(res))
    b.record()
    torch.cuda.synchronize()
    ops.profile_start()
    model(dict(res))
    recs = ops.profile_stop()
    tot = {}
    for name, t, _ in recs:
        tot[name] = tot.get(name, 0.0) + t
    print(f"{label:40s} {a.elapsed_time(b) / steps:8.3f} ms/step  " + " ".join(f"{k}={v:.2f}" for k, v in tot.items()),
          flush=True)


cfgs = [a for a in sys.argv[1:]] or ["base"]
for c in cfgs:
    lib.ptv3_debug_set(2, 1 if "split1" in c else (2 if "split2" in c else (4 if "split4" in c else 0)))
    run(c)
