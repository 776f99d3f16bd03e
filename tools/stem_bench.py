"""Stem conv timing: hash-probe search vs block-occupancy search on the bench batch.  Development aid."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import bench  # noqa: E402
from ptv3_b200 import ops  # noqa: E402
from ptv3_b200.structure import Point  # noqa: E402

dev = torch.device("cuda", 0)
raw = bench.make_raw_scenes(16, 100000, 4.1, seed0=0)
host, counts = bench.voxelize_on_device(raw, dev)
res = {k: v.to(dev) for k, v in host.items()}
p = Point(dict(res))
p.serialization(order=["z", "z-trans", "hilbert", "hilbert-trans"], shuffle_orders=False)
p.sparsify()
n = p.feat.shape[0]
w = torch.randn(125, 32, 4, device=dev) * 0.05
b = torch.randn(32, device=dev)
z = (p.serialized_code[0], p.serialized_order[0], int(p.serialized_depth))


def t(fn, reps=10):
    fn()
    torch.cuda.synchronize()
    a, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    e.record()
    torch.cuda.synchronize()
    return a.elapsed_time(e) / reps


idx = p.sparse_conv_feat.indices
print(f"n = {n}, depth {z[2]}")
print(f"hash probes   {t(lambda: ops.subm_stem_fused(idx, p.feat, w, b, 5)):.3f} ms")
print(f"block search  {t(lambda: ops.subm_stem_fused(idx, p.feat, w, b, 5, z=z)):.3f} ms")
a = ops.subm_stem_fused(idx, p.feat, w, b, 5)
c = ops.subm_stem_fused(idx, p.feat, w, b, 5, z=z)
print("identical", torch.equal(a[0], c[0]))
from torch.profiler import profile, ProfilerActivity  # noqa: E402
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    ops.subm_stem_fused(idx, p.feat, w, b, 5)
    ops.subm_stem_fused(idx, p.feat, w, b, 5, z=z)
    torch.cuda.synchronize()
for e in sorted(prof.key_averages(), key=lambda e: -e.device_time_total)[:8]:
    print(f"{e.device_time_total / 1e3:8.3f} ms n={e.count} {e.key[:90]}")
