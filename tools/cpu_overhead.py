"""Host-side cost of one forward: enqueue time without waiting for the GPU, and the top Python frames (cProfile)."""
import cProfile
import os
import pstats
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import bench  # noqa: E402

dev = torch.device("cuda", 0)
model, _ = bench.build_model(torch.bfloat16, dev)
raw = bench.make_raw_scenes(int(sys.argv[1]) if len(sys.argv) > 1 else 16, 100000, 4.1, seed0=0)
host, counts = bench.voxelize_on_device(raw, dev)
res = {k: v.to(dev) for k, v in host.items()}
for _ in range(3):
    model(dict(res))
torch.cuda.synchronize()
ts = []
for _ in range(5):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    model(dict(res))
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    ts.append(((t1 - t0) * 1e3, (t2 - t0) * 1e3))
print("enqueue ms / total ms per forward:", [(round(a, 2), round(b, 2)) for a, b in ts])
pr = cProfile.Profile()
pr.enable()
for _ in range(3):
    model(dict(res))
torch.cuda.synchronize()
pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(18)
