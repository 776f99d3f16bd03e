"""Micro-benchmark of the tcgen05 submanifold conv on bench-shaped neighbour maps (CUDA events; development aid).
Stage s of the 16-scene workload: voxel grid >> s, C = 32 * 2**s (plus C = 64 at stage 0 for the decoder)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import bench  # noqa: E402
from ptv3_b200 import ops, _lib  # noqa: E402

dev = torch.device("cuda", 0)
raw = bench.make_raw_scenes(16, 100000, 4.1, seed0=0)
host, counts = bench.voxelize_on_device(raw, dev)
grid = host["grid_coord"].to(dev)
batch = torch.repeat_interleave(torch.arange(16, device=dev), torch.tensor(counts, device=dev))
_lib.load().ptv3_debug_set(3, int(os.environ.get('CONV_ASYNC', '1')))   # 1 = gather copies arrive on the stage barrier themselves (default)
stages = int(sys.argv[1]) if len(sys.argv) > 1 else 3
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
for s in range(stages):
    g = grid >> s
    key = (batch << 48) | (g[:, 0] << 32) | (g[:, 1] << 16) | g[:, 2]
    uk = torch.unique(key)
    idx = torch.stack([uk >> 48, (uk >> 32) & 0xFFFF, (uk >> 16) & 0xFFFF, uk & 0xFFFF], 1).int()
    if s == 0:
        idx = idx[torch.randperm(idx.shape[0], device=dev)]      # stage 0 rows are in hash (random) order
    n = idx.shape[0]
    nbr = ops.subm_build_map(idx, 3)
    present = float((nbr >= 0).sum())
    for c in ([32, 64] if s == 0 else [32 << s]):
        x = torch.randn(n, c, device=dev).bfloat16()
        w = (torch.randn(27, c, c, device=dev) / (c * 27) ** 0.5).bfloat16()
        packed = ops.subm_pack_weight(w)
        bias = torch.zeros(c, device=dev)
        for _ in range(3):
            y = ops.subm_conv_packed(x, packed, bias, nbr, c)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            y = ops.subm_conv_packed(x, packed, bias, nbr, c)
        b.record()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b) / reps
        fl = 2.0 * c * c * present
        print(f"stage {s} n={n} C={c} pbar={present / n:.2f}: {ms:.3f} ms  {fl / ms / 1e9:.1f} TFLOP/s useful  "
              f"gather {present * c * 2 / ms / 1e9:.2f} TB/s  chk={float(y.float().abs().mean()):.4f}", flush=True)
