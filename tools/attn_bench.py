"""Micro-benchmark of the window-attention kernels on bench-shaped inputs (CUDA events, L2-sized inputs)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from ptv3_b200 import ops, _lib
lib = _lib.load()
variants = [int(x) for x in (sys.argv[1].split(",") if len(sys.argv) > 1 else ["0"])]
shapes = [(16 * 72934, 2), (16 * 36838, 4), (16 * 11356, 8), (16 * 2976, 16), (16 * 746, 32)]
for n, H in shapes:
    per = n // 16
    offset = torch.arange(1, 17, device="cuda") * per
    C = H * 16
    qkv = torch.randn(n, 3 * C, device="cuda").bfloat16()
    order = torch.cat([torch.randperm(per, device="cuda") + i * per for i in range(16)])
    w, nw = ops.build_windows(offset, 1024, n)
    def run(algo):
        for _ in range(2):
            ops.attn_window(qkv, order, w, nw, H, 1024, algo=algo)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(5):
            ops.attn_window(qkv, order, w, nw, H, 1024, algo=algo)
        b.record(); torch.cuda.synchronize()
        return a.elapsed_time(b) / 5
    exps = float((w[: int(nw)][:, 1] * w[: int(nw)][:, 3]).sum()) * H
    line = f"n={n} H={H} exps={exps/1e9:.2f}G  mma.sync {run(2):.3f} ms"
    for v in variants:
        lib.ptv3_debug_set(1, v)
        t = run(3)
        line += f" | tc05[v{v}] {t:.3f} ms ({exps/t/1e9:.2f} Texp/s)"
    lib.ptv3_debug_set(1, 0)
    print(line)
