"""Offset-keypoint validation metric (A10, ptv3_keypoint_eval) at the bench batch shape: device time per batch with CUDA
events, achieved HBM GB/s on the algorithmic bytes (32 B per row and keypoint + 12 B per masked row), and beside it the
reference's formulation (hooks/offset_keypoint_evaluator.py:46-91: a host loop over scenes and keypoints of torch ops
with .item() reads) run on the same device tensors.  python tools/eval_bench.py [scenes] [rows_per_scene]"""
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ptv3_b200 import ops  # noqa: E402


def host_loop(coord, pred, target, offset, scale):
    tot = 0.0
    for b in range(len(offset)):
        s, e = (0 if b == 0 else offset[b - 1]), offset[b]
        c, p, t = coord[s:e], pred[s:e], target[s:e]
        for k in range(pred.shape[1]):
            idx = torch.nonzero(t[:, k, 3] > 0).squeeze(-1)
            if len(idx) == 0:
                continue
            gt = (c[idx] + t[idx, k, :3]).mean(dim=0)
            best = torch.argmax(p[:, k, 3])
            tot += (torch.norm(c[best] + p[best, k, :3] - gt, p=2) * scale[b].item()).item()
    return tot


def main():
    scenes = int(sys.argv[1]) if len(sys.argv) > 1 else 16
    rows = int(sys.argv[2]) if len(sys.argv) > 2 else 72834
    K, n = 6, scenes * rows
    g = torch.Generator(device="cuda").manual_seed(0)
    coord = torch.randn(n, 3, device="cuda", generator=g)
    pred = torch.randn(n, K, 4, device="cuda", generator=g)
    target = torch.randn(n, K, 4, device="cuda", generator=g)
    target[..., 3] = (torch.rand(n, K, device="cuda", generator=g) < 0.05).float()
    offset = torch.arange(1, scenes + 1, device="cuda") * rows
    scale = torch.full((scenes,), 873.69, device="cuda")
    acc = torch.zeros(2 + 2 * K, dtype=torch.float64, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    def timed(split):
        for _ in range(3):
            ops.keypoint_eval(coord, pred, target, offset, scale, acc, split=split)
        ms = []
        for _ in range(20):
            flush.zero_()                                  # L2 flush between timed launches
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            ops.keypoint_eval(coord, pred, target, offset, scale, acc, split=split)
            e1.record()
            torch.cuda.synchronize()
            ms.append(e0.elapsed_time(e1))
        ms.sort()
        return ms[len(ms) // 2]
    single = timed(False)
    med = timed(True)
    masked = int((target[..., 3] > 0).sum().item())
    alg = n * K * 32 + masked * 12
    off_host = offset.cpu()
    host_loop(coord, pred, target, off_host, scale)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    host_loop(coord, pred, target, off_host, scale)
    torch.cuda.synchronize()
    loop_ms = (time.perf_counter() - t0) * 1e3
    print(json.dumps(dict(kernel="keypoint_eval", scenes=scenes, rows=n, keypoints=K, ms=round(med, 4),
                          algorithmic_bytes=alg, achieved_gb_s=round(alg / med / 1e6, 1),
                          single_pass_ms=round(single, 4), single_pass_gb_s=round(alg / single / 1e6, 1),
                          reference_host_loop_ms=round(loop_ms, 3), speedup=round(loop_ms / med, 1))))


if __name__ == "__main__":
    main()
