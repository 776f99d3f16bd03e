"""SpUNet-v1m1 forward on LiDAR-shaped sweeps (BASELINE configs[3]); run on the GPU box.
python tools/spunet_bench.py [sweeps] [points] [steps]"""
import os
import sys
import time

import numpy as np
import torch

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
from ptv3_b200 import ops, spunet as S, synth  # noqa: E402


def main():
    sweeps = int(sys.argv[1]) if len(sys.argv) > 1 else 8
    points = int(sys.argv[2]) if len(sys.argv) > 2 else 120000
    steps = int(sys.argv[3]) if len(sys.argv) > 3 else 10
    dev = torch.device("cuda", 0)
    grids, feats, counts = [], [], []
    for i in range(sweeps):
        c, f = synth.lidar_sweep(points, seed=i)
        ct = torch.from_numpy(c).to(dev)
        r = ops.gridsample(ct, 0.05, "fnv")
        iu = r["idx_unique"]
        grids.append(r["grid_coord"][iu])
        feats.append(torch.from_numpy(f).to(dev)[iu])
        counts.append(int(iu.numel()))
    batch = dict(grid_coord=torch.cat(grids), feat=torch.cat(feats),
                 offset=torch.tensor(np.cumsum(counts), device=dev))
    n = sum(counts)
    for dtype in (torch.bfloat16, torch.float32):
        m = S.SpUNetBase(4, 19, compute_dtype=dtype).eval().to(dev)
        for _ in range(3):
            out = m(batch)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(steps):
            out = m(batch)
        b.record()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b) / steps
        print(f"SpUNet-v1m1 {dtype}: {sweeps} sweeps x {points} rays -> {n} voxels (grid 0.05 m): {ms:.2f} ms/forward, "
              f"{n / ms * 1e3 / 1e6:.2f} M voxels/s, logits {tuple(out.shape)}")
        if dtype == torch.bfloat16:
            from torch.profiler import ProfilerActivity, profile
            with profile(activities=[ProfilerActivity.CUDA]) as prof:
                m(batch)
                torch.cuda.synchronize()
            tot = sum(r.device_time_total for r in prof.key_averages())
            print(f"  device total {tot / 1e3:.2f} ms")
            for r in sorted(prof.key_averages(), key=lambda r: -r.device_time_total)[:10]:
                print(f"  {r.device_time_total / 1e3:8.3f} ms {100 * r.device_time_total / tot:5.1f}% n={r.count:4d} {r.key[:90]}")


main()
