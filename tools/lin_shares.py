"""Per-kernel device-time shares of one BASELINE configs[4] step (PTv3-large linear probe, bench.py --config 5) via
torch.profiler.  Development aid only: numbers taken under a profiler are never bench values."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

import bench  # noqa: E402
from ptv3_b200 import model as M, synth  # noqa: E402

dev = torch.device("cuda", 0)
scenes = int(sys.argv[1]) if len(sys.argv) > 1 else 8
model = M.DefaultSegmentorV2(num_classes=22, backbone_out_channels=1728, backbone=dict(bench.LARGE_M2), freeze_backbone=True,
                             criteria=[dict(type="CrossEntropyLoss", loss_weight=1.0, ignore_index=-1),
                                       dict(type="LovaszLoss", mode="multiclass", loss_weight=1.0, ignore_index=-1)])
shapes = {k: tuple(v.shape) for k, v in model.state_dict().items()}
model.load_state_dict({k: torch.from_numpy(v) for k, v in synth.make_weights(shapes, seed=0).items()}, strict=False)
model = model.to(dev).train()
raw = []
for i in range(scenes):
    c, f = synth.lidar_sweep(int(180000 * 2.2), seed=i, in_channels=9, beams=64, max_range=75.0)
    keep = np.random.default_rng(i).permutation(len(c))[:180000]
    raw.append(((c[keep] * 0.2).astype(np.float32), f[keep]))
host, counts = bench.voxelize_on_device(raw, dev, 0.01)
n = int(sum(counts))
rng = np.random.default_rng(7)
host["segment"] = torch.from_numpy(rng.integers(0, 22, n).astype(np.int64))
res = {k: v.to(dev) for k, v in host.items()}
opt = torch.optim.AdamW([p for p in model.parameters() if p.requires_grad], lr=0.002, fused=True)


def step():
    opt.zero_grad(set_to_none=True)
    out = model(dict(res))
    out["loss"].backward()
    opt.step()


for _ in range(3):
    step()
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(5):
    step()
b.record()
torch.cuda.synchronize()
print(f"voxels {n}  step {a.elapsed_time(b) / 5:.3f} ms  peak mem {torch.cuda.max_memory_allocated() / 2 ** 30:.2f} GiB")
from torch.profiler import profile, ProfilerActivity  # noqa: E402
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    step()
    torch.cuda.synchronize()
ev = [e for e in prof.key_averages() if e.device_time_total > 0 and e.device_type.name == "CUDA"]
tot = sum(e.device_time_total for e in ev)
print(f"device total {tot / 1e3:.3f} ms over {sum(e.count for e in ev)} launches")
for e in sorted(ev, key=lambda e: -e.device_time_total)[:40]:
    print(f"{e.device_time_total / 1e3:9.3f} ms {100 * e.device_time_total / tot:5.1f}% n={e.count:4d} {e.key[:110]}")
ks = sorted([e for e in prof.events() if e.device_type.name == "CUDA" and e.time_range.end > e.time_range.start],
            key=lambda e: e.time_range.start)
span = ks[-1].time_range.end - ks[0].time_range.start
idle = sum(max(0, b_.time_range.start - a_.time_range.end) for a_, b_ in zip(ks[:-1], ks[1:]))
print(f"first-to-last kernel span {span / 1e3:.3f} ms, idle inside {idle / 1e3:.3f} ms")
gc = res["grid_coord"].long()
bt = torch.repeat_interleave(torch.arange(len(counts), device=dev), torch.tensor(counts, device=dev))
print("rows per level:", [int(torch.unique((bt << 48) | ((gc[:, 0] >> s_) << 32) | ((gc[:, 1] >> s_) << 16) | (gc[:, 2] >> s_)).numel())
                          for s_ in range(5)])
convs = [e for e in ks if "subm_conv_tc_kernel" in e.name]
print("conv launches in order (us):", " ".join(f"{e.name.split('<')[1].split('>')[0].replace(' ', '')}:{e.time_range.end - e.time_range.start:.0f}" for e in convs))
