"""Per-kernel device-time shares of one bench step via torch.profiler (CUPTI).  Development aid only: numbers
taken under a profiler are never bench values."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scenes", type=int, default=16)
    ap.add_argument("--points", type=int, default=100000)
    ap.add_argument("--scale", type=float, default=4.1)
    ap.add_argument("--dtype", default="bf16")
    ap.add_argument("--top", type=int, default=45)
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    dtype = torch.bfloat16 if a.dtype == "bf16" else torch.float32
    model, _ = bench.build_model(dtype, dev)
    raw = bench.make_raw_scenes(a.scenes, a.points, a.scale, seed0=0)
    host, counts = bench.voxelize_on_device(raw, dev)
    res = {k: v.to(dev) for k, v in host.items()}
    for _ in range(3):
        model(dict(res))
    torch.cuda.synchronize()
    from torch.profiler import profile, ProfilerActivity
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        model(dict(res))
        torch.cuda.synchronize()
    ev = [e for e in prof.key_averages() if e.device_time_total > 0 and e.device_type.name == "CUDA"]
    tot = sum(e.device_time_total for e in ev)
    print(f"voxels={sum(counts)} device total {tot/1e3:.3f} ms over {sum(e.count for e in ev)} launches")
    for e in sorted(ev, key=lambda e: -e.device_time_total)[: a.top]:
        print(f"{e.device_time_total/1e3:9.3f} ms {100*e.device_time_total/tot:5.1f}% n={e.count:4d} {e.key[:100]}")
    # idle intervals of the device inside the step: where the host does not keep the queue full
    ks = sorted([e for e in prof.events() if e.device_type.name == "CUDA" and e.time_range.end > e.time_range.start],
                key=lambda e: e.time_range.start)
    if ks:
        span = ks[-1].time_range.end - ks[0].time_range.start
        gaps = [(b.time_range.start - a.time_range.end, a.name, b.name) for a, b in zip(ks[:-1], ks[1:])]
        idle = sum(g[0] for g in gaps if g[0] > 0)
        print(f"first-to-last kernel span {span/1e3:.3f} ms, idle inside {idle/1e3:.3f} ms in {sum(g[0] > 2 for g in gaps)} gaps > 2 us")
        for g, an, bn in sorted(gaps, key=lambda g: -g[0])[:14]:
            print(f"  gap {g:8.1f} us after {an[:60]:60s} before {bn[:50]}")


if __name__ == "__main__":
    main()
