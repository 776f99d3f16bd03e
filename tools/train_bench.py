"""Training step (forward in train() mode + loss + backward + AdamW) of OffsetKeypointPTv3 on synthetic scenes
(run on the GPU box).  python tools/train_bench.py [scenes] [points] [steps]
Prints ms per step and the device-time split of the backward kernels (torch profiler)."""
import os
import sys
import time

import torch

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
sys.path.insert(0, os.path.join(REPO, "tests"))
import bench  # noqa: E402
from make_golden_names import make_target  # noqa: E402


def main():
    scenes = int(sys.argv[1]) if len(sys.argv) > 1 else 4
    points = int(sys.argv[2]) if len(sys.argv) > 2 else 100000
    steps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
    dev = torch.device("cuda", 0)
    raw = bench.make_raw_scenes(scenes, points, 4.1, seed0=0)
    host, counts = bench.voxelize_on_device(raw, dev)
    batch = {k: v.to(dev) for k, v in host.items()}
    n = sum(counts)
    batch["target"] = torch.from_numpy(make_target(n, 1)).to(dev)
    model, _ = bench.build_model(torch.bfloat16, dev)
    model.train()
    opt = torch.optim.AdamW(model.parameters(), lr=1e-4, fused=True)

    def step():
        opt.zero_grad(set_to_none=True)
        out = model(dict(batch))
        out["loss"].backward()
        opt.step()
        return out["loss"]

    for _ in range(2):
        loss = step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        loss = step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    print(f"voxels={n} train step {ms:.1f} ms ({n / ms * 1e3 / 1e6:.2f} M voxels/s) loss {loss.item():.4f} "
          f"peak mem {torch.cuda.max_memory_allocated() / 2**30:.1f} GiB")
    from torch.profiler import ProfilerActivity, profile
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        step()
        torch.cuda.synchronize()
    rows = sorted(prof.key_averages(), key=lambda r: -r.device_time_total)[:40]
    tot = sum(r.device_time_total for r in prof.key_averages())
    print(f"device total {tot / 1e3:.1f} ms")
    for r in rows:
        print(f"  {r.device_time_total / 1e3:9.2f} ms {100 * r.device_time_total / tot:5.1f}% n={r.count:4d} {r.key[:100]}")


if __name__ == "__main__":
    main()
