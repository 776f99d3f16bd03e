"""Attention backward kernel alone on the stage-0 shapes of the bench batch (run on the GPU box)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ptv3_b200 import _lib, ops

def main():
    dev = torch.device("cuda", 0)
    n_scene, scenes, H, K = 72900, 8, 4, 1024
    n = n_scene * scenes
    g = torch.Generator(device="cuda").manual_seed(0)
    qkv = torch.randn(n, 3 * H * 16, device=dev, generator=g).bfloat16()
    dout = torch.randn(n, H * 16, device=dev, generator=g).bfloat16()
    offset = torch.arange(1, scenes + 1, device=dev) * n_scene
    order = torch.cat([torch.randperm(n_scene, device=dev, generator=g) + i * n_scene for i in range(scenes)])
    windows, nwin = ops.build_windows(offset, K, n)
    out = ops.attn_window(qkv, order, windows, nwin, H, K)
    lib = _lib.load()
    out2, lse = ops.attn_window_lse(qkv, order, windows, nwin, H, K)
    assert torch.equal(out, out2)
    for warps, use_lse in ((8, False), (16, False), (16, True)):
        lib.ptv3_debug_set(6, warps)
        kw = dict(lse=lse) if use_lse else {}
        for _ in range(2):
            ops.attn_window_bwd(qkv, out, dout, order, windows, nwin, H, K, **kw)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(5):
            ops.attn_window_bwd(qkv, out, dout, order, windows, nwin, H, K, **kw)
        b.record(); torch.cuda.synchronize()
        print(f"warps={warps} lse={use_lse}: bwd {a.elapsed_time(b) / 5:.3f} ms (incl. the zero-fill of dqkv)")
    lib.ptv3_debug_set(6, 16)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(5):
        ops.attn_window(qkv, order, windows, nwin, H, K)
    b.record(); torch.cuda.synchronize()
    print(f"fwd {a.elapsed_time(b) / 5:.3f} ms   n={n} H={H}")

main()
