"""Micro-benchmark of the fused MLP kernel at the bench's stage-0/1 shapes (CUDA events; development aid)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from ptv3_b200 import ops  # noqa: E402

dev = torch.device("cuda", 0)
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 10
shapes = ((1165350, 32), (1165350, 64), (588256, 64), (181696, 128), (588256, 128), (47616, 256), (181696, 256))
if len(sys.argv) > 2:
    shapes = ((int(sys.argv[2]), int(sys.argv[3])),)
for n, c in shapes:
    x = torch.randn(n, c, device=dev).bfloat16()
    res = torch.randn(n, c, device=dev)
    w1 = (torch.randn(4 * c, c, device=dev) / c ** 0.5).bfloat16()
    w2 = (torch.randn(c, 4 * c, device=dev) / (4 * c) ** 0.5).bfloat16()
    b1, b2 = torch.randn(4 * c, device=dev) * 0.1, torch.randn(c, device=dev) * 0.1
    w1p, w2p = ops.pack_kmajor(w1, ops.mlp_fc1_tile_rows(c)), ops.pack_kmajor(w2, c)
    for _ in range(3):
        r, o = ops.mlp_fused(x, w1p, b1, w2p, b2, 4 * c, residual=res)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        r, o = ops.mlp_fused(x, w1p, b1, w2p, b2, 4 * c, residual=res)
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / reps
    ref = res + torch.nn.functional.gelu(x[:4096].float() @ w1.float().T + b1).bfloat16().float() @ w2.float().T + b2 \
        if False else None
    hbm = n * c * (2 + 4 + 4 + 2)
    print(f"n={n} C={c}: {ms:.3f} ms  {hbm / ms / 1e9:.2f} TB/s algorithmic  {16.0 * n * c * c / ms / 1e9:.1f} TFLOP/s  "
          f"chk={float(r.abs().mean()):.4f}", flush=True)
