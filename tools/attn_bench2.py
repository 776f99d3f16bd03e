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


def profile_one(stage):
    """One shape, default kernel only, few launches: the command line for `ncu --set full`."""
    n, H = [(16 * 72934, 2), (16 * 36838, 4), (16 * 11356, 8), (16 * 2976, 16), (16 * 746, 32)][stage]
    per = n // 16
    offset = torch.arange(1, 17, device="cuda") * per
    qkv = torch.randn(n, 3 * H * 16, device="cuda").bfloat16()
    order = torch.cat([torch.randperm(per, device="cuda") + i * per for i in range(16)])
    w, nw = ops.build_windows(offset, 1024, n)
    algo = 0
    if len(sys.argv) > 3:
        if int(sys.argv[3]) < 10:
            algo = int(sys.argv[3])
        else:
            ops.debug_set(7, int(sys.argv[3]))
    if len(sys.argv) > 4:
        ops.debug_set(9, int(sys.argv[4]))
    for _ in range(3):
        ops.attn_window(qkv, order, w, nw, H, 1024, algo=algo)
    torch.cuda.synchronize()
    print("ok")


def main():
    if len(sys.argv) > 2 and sys.argv[1] == "--profile":
        return profile_one(int(sys.argv[2]))
    shapes = [(16 * 72934, 2), (16 * 36838, 4), (16 * 11356, 8), (16 * 2976, 16), (16 * 746, 32)]
    # debug slot 7: hundreds = ping-pong of the exponential phase (1 off, 2 on), tens = 1 + eighths of the exponentials
    # on the FMA-pipe polynomial, units = 1: 64-key chunks, 2 CTAs / SM; 2: 32-key chunks, 3 CTAs / SM
    variants = [int(v) for v in sys.argv[1].split(",")] if len(sys.argv) > 1 else [111, 211, 221, 231, 241, 112, 212, 232]
    try:
        from flash_attn import flash_attn_varlen_qkvpacked_func as fa2
    except Exception as e:  # noqa: BLE001
        fa2 = None
        print("flash_attn unavailable:", e)
    names = ["algo 2 (r01)"] + [f"algo 5 (tcgen05), poly {pn}/8" for pn in range(3)] + ["algo 5 ABLATION no exp"] + [(f"v{v}: {'pp' if v // 100 == 2 else 'rr'}, poly {(v // 10) % 10 - 1}/8, {['', '2 tiles/warp 64-key 2 CTAs', '2 tiles/warp 32-key 3 CTAs', '1 tile/warp 64-key 3 CTAs', '1 tile/warp 64-key 4 CTAs'][v % 10]}" if v < 1000 else
              "ABLATION without " + "+".join(n for b, n in ((1, "exp"), (2, "PV MMAs"), (4, "QK MMAs"), (8, "max")) if (v - 1000) & b))
             for v in variants] + \
            ["FA2 2.8 (pre-gathered)"]
    print("| rows | H | scores | " + " | ".join(names) + " | best scores/clk/SM |")
    print("|" + "---|" * (len(names) + 4))
    tot = [0.0] * len(names)
    for n, H in shapes:
        per = n // 16
        offset = torch.arange(1, 17, device="cuda") * per
        C = H * 16
        qkv = torch.randn(n, 3 * C, device="cuda").bfloat16()
        order = torch.cat([torch.randperm(per, device="cuda") + i * per for i in range(16)])
        w, nw = ops.build_windows(offset, 1024, n)
        scores = float((w[: int(nw)][:, 1] * w[: int(nw)][:, 3]).sum()) * H
        ts = [timed(lambda: ops.attn_window(qkv, order, w, nw, H, 1024, algo=2))]
        for pn in (0, 1, 2, 9):
            ops.debug_set(9, pn)
            ts.append(timed(lambda: ops.attn_window(qkv, order, w, nw, H, 1024, algo=5)))
        ops.debug_set(9, 0)
        ref = ops.attn_window(qkv, order, w, nw, H, 1024, algo=1).float() if n < 100000 else None
        errs = []
        for v in variants:
            ops.debug_set(7, v)
            ts.append(timed(lambda: ops.attn_window(qkv, order, w, nw, H, 1024, algo=4)))
            if ref is not None:
                got = ops.attn_window(qkv, order, w, nw, H, 1024, algo=4).float()
                errs.append(f"{((got - ref).norm() / ref.norm()).item():.1e}")
        ops.debug_set(7, 0)
        tfa = float("nan")
        if fa2 is not None:
            # the reference's padded layout: every scene padded to a multiple of 1024 by repeating rows (:141-164)
            from oracle import ops as O
            pad, unpad, cu = O.padding_and_inverse(offset.cpu().numpy(), 1024)
            g = qkv[order[torch.from_numpy(pad).cuda()]].reshape(-1, 3, H, 16).contiguous()
            cu_t = torch.from_numpy(cu).cuda()
            tfa = timed(lambda: fa2(g, cu_t, 1024, 0.0, softmax_scale=0.25))
        ts.append(tfa)
        tot = [a + b for a, b in zip(tot, ts)]
        if ref is not None:
            for pn in (9, 2, 1, 0):
                ops.debug_set(9, pn)
                got = ops.attn_window(qkv, order, w, nw, H, 1024, algo=5).float()
                errs.insert(0, f"{((got - ref).norm() / ref.norm()).item():.1e}")
            ops.debug_set(9, 0)
        best = min(ts[1:-1])
        print(f"| {n} | {H} | {scores / 1e9:.2f} G | " + " | ".join(f"{t:.3f}" for t in ts) +
              f" | {scores / (best * 1e-3) / SM_CLK:.2f} |")
        if errs:
            print(f"|  |  | rel-L2 vs the fp32-math kernel (algo 1) | | " + " | ".join(errs) + " | | |")
    print("| sum | | | " + " | ".join(f"{t:.3f}" for t in tot) + " | |")


if __name__ == "__main__":
    main()
