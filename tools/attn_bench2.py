"""Round-2 micro-benchmark of the window-attention kernels on the bench shapes (CUDA events on the launching stream,
inputs larger than L2 at the big stages).  Columns: first-generation mma.sync kernel (algo 2), the default f16-staged
kernel (algo 4) in both chunk / occupancy variants, and -- as the GPU comparator BASELINE.md names -- FlashAttention-2
`flash_attn_varlen_qkvpacked_func` (flash-attn 2.8 from the image, LIBRARY code, never on the product path) on the SAME
windows, fed the already gathered + padded qkv the reference would build (:184-190), i.e. without the gather / scatter
our kernels fuse.  Development aid: prints a markdown table for profiles/."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from ptv3_b200 import ops  # noqa: E402

SM_CLK = 1965e6 * 148


def timed(fn, iters=5, warm=2):
    for _ in range(warm):
        fn()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters


def main():
    shapes = [(16 * 72934, 2), (16 * 36838, 4), (16 * 11356, 8), (16 * 2976, 16), (16 * 746, 32)]
    try:
        from flash_attn import flash_attn_varlen_qkvpacked_func as fa2
    except Exception as e:  # noqa: BLE001
        fa2 = None
        print("flash_attn unavailable:", e)
    print("| rows | H | scores | algo 2 (r01) ms | algo 4 / 64-key chunks ms | scores/clk/SM | algo 4 / 32-key chunks ms | "
          "FA2 2.8 (pre-gathered input) ms |")
    print("|---|---|---|---|---|---|---|---|")
    tot = {}
    for n, H in shapes:
        per = n // 16
        offset = torch.arange(1, 17, device="cuda") * per
        C = H * 16
        qkv = torch.randn(n, 3 * C, device="cuda").bfloat16()
        order = torch.cat([torch.randperm(per, device="cuda") + i * per for i in range(16)])
        w, nw = ops.build_windows(offset, 1024, n)
        scores = float((w[: int(nw)][:, 1] * w[: int(nw)][:, 3]).sum()) * H
        t2 = timed(lambda: ops.attn_window(qkv, order, w, nw, H, 1024, algo=2))
        ops.debug_set(7, 1)
        t4a = timed(lambda: ops.attn_window(qkv, order, w, nw, H, 1024, algo=4))
        ops.debug_set(7, 2)
        t4b = timed(lambda: ops.attn_window(qkv, order, w, nw, H, 1024, algo=4))
        ops.debug_set(7, 0)
        tfa = float("nan")
        if fa2 is not None:
            # the reference's padded layout: every scene padded to a multiple of 1024 by repeating rows (:141-164)
            from oracle import ops as O
            pad, unpad, cu = O.padding_and_inverse(offset.cpu().numpy(), 1024)
            g = qkv[order[torch.from_numpy(pad).cuda()]].reshape(-1, 3, H, 16).contiguous()
            cu_t = torch.from_numpy(cu).cuda()
            tfa = timed(lambda: fa2(g, cu_t, 1024, 0.0, softmax_scale=0.25))
        for k, v in (("algo2", t2), ("algo4_64", t4a), ("algo4_32", t4b), ("fa2", tfa)):
            tot[k] = tot.get(k, 0.0) + v
        print(f"| {n} | {H} | {scores / 1e9:.2f} G | {t2:.3f} | {t4a:.3f} | {scores / (t4a * 1e-3) / SM_CLK:.2f} | {t4b:.3f} | "
              f"{tfa:.3f} |")
    print("| sum | | | " + " | ".join(f"{tot[k]:.3f}" for k in ("algo2", "algo4_64")) + " | | " +
          " | ".join(f"{tot[k]:.3f}" for k in ("algo4_32", "fa2")) + " |")


if __name__ == "__main__":
    main()
