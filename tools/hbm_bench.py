"""Achieved HBM bandwidth of every memory-bound kernel on the bench shapes: algorithmic bytes
 / CUDA-event duration / the measured copy peak (MEASURED_PEAKS.json).  One op per row, timed alone
on the launching stream, L2 flushed (a 512 MB write) before every timed launch, median of 7.  `ncu` rows for the same
command line (dram bytes, duration) go to profiles/r02_hbm_kernels_ncu.csv.  Development aid: prints a markdown table.
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

import bench  # noqa: E402
from ptv3_b200 import ops  # noqa: E402
from ptv3_b200.structure import Point  # noqa: E402

PEAK = bench.peaks()["hbm"]
FLUSH = None


def timed(fn, reps=7):
    global FLUSH
    if FLUSH is None:
        FLUSH = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
    fn()
    ts = []
    for _ in range(reps):
        FLUSH.fill_(1)                       # evict L2 (126 MB)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return float(np.median(ts))


ROWS = []


def row(name, what, nbytes, ms, launches=1):
    gbs = nbytes / (ms * 1e-3) / 1e9
    ROWS.append((name, what, nbytes, ms, gbs, gbs / PEAK, launches))
    print(f"| `{name}` | {what} | {nbytes / 1e6:.1f} | {launches} | {ms * 1e3:.1f} | {gbs:.0f} | {gbs / PEAK:.2f} |", flush=True)


def main():
    dev = torch.device("cuda", 0)
    quick = "--quick" in sys.argv
    scenes = 4 if quick else 16
    raw = bench.make_raw_scenes(scenes, 100000, 4.1, seed0=0)
    host, counts = bench.voxelize_on_device(raw, dev)
    res = {k: v.to(dev) for k, v in host.items()}
    n = int(sum(counts))
    print(f"# HBM-bound kernels on the bench batch ({scenes} scenes, {n} voxels), peak {PEAK:.1f} GB/s (measured copy)\n")
    print("| kernel(s) | what | algorithmic MB | launches | us | GB/s | of measured peak |")
    print("|---|---|---|---|---|---|---|")
    p = Point(dict(res))
    depth = 9
    orders = ["z", "z-trans", "hilbert", "hilbert-trans"]
    # ---- A2 encode: 24 B grid + 8 B batch in, 4 x 8 B out per point
    ms = timed(lambda: ops.encode_orders(p.grid_coord, p.batch, depth, orders))
    row("encode_kernel", "Morton + Hilbert codes, 4 orders, one pass", n * (24 + 8 + 32), ms)
    code = ops.encode_orders(p.grid_coord, p.batch, depth, orders)
    key_bits = depth * 3 + max(1, (scenes - 1).bit_length())
    passes = (key_bits + 7) // 8
    ms = timed(lambda: ops.argsort_rows(code, key_bits=key_bits))
    # per 8-bit pass and row: histogram read (8 B) + scatter read (8 + 4 B) + write (8 + 4 B); + inverse 8 B once
    row("rs_hist / rs_scatter", f"radix argsort of (4, N) int64 codes, {key_bits} key bits = {passes} passes, + inverse",
        4 * n * (passes * (8 + 12 + 12) + 16), ms, launches=2 * passes * 1 + 2)
    order, inverse = ops.argsort_rows(code, key_bits=key_bits)
    # ---- A0 GridSample, per scene (bench leg) and the whole batch in one chain
    c0 = torch.from_numpy(raw[0][0]).to(dev)
    r = c0.shape[0]
    ms = timed(lambda: ops.gridsample(c0, 0.02, "fnv"))
    gs_bytes = lambda r_: r_ * (12 + 24 * 2 + 8 + 8 * (8 + 12 + 12) + 8 * 6)
    row("gridsample_* + radix (one scene)", "floor/min, FNV key, 64-bit stable sort (8 passes), unique, count, inverse, select; "
        "includes the host read of the voxel count", gs_bytes(r), ms, launches=26)
    if hasattr(ops, "gridsample_batch"):
        call = [torch.from_numpy(c).to(dev) for c, _ in raw]
        allc = torch.cat(call)
        off = torch.tensor(np.cumsum([c.shape[0] for c in call]), device=dev)
        ms = timed(lambda: ops.gridsample_batch(allc, off, 0.02, "fnv"))
        row("gridsample batch (all scenes, one chain)", "same, scene id in the sort key, one host read per batch",
            gs_bytes(allc.shape[0]), ms, launches=30)
    # ---- A6 pooling front end + gathers
    ms = timed(lambda: ops.pool_front(code, order, 1))
    row("pool_front (flags, scan, orders)", "run heads of code >> 3 over 4 sorted orders -> cluster, CSR, pooled code / order / inverse; "
        "one host read", n * (4 * 16 + 4 * 24 + 8 * 3), ms, launches=9)
    pf = ops.pool_front(code, order, 1)
    m = pf["n_pooled"]
    ms = timed(lambda: ops.pool_gather_ints(p.grid_coord, p.batch, pf["head"], 1))
    row("pool_gather_ints_kernel", "grid >> 1, batch, spconv indices of the pooled rows", m * (8 + 24 + 8 + 24 + 8 + 16), ms)
    for cch in (64,):
        y = torch.randn(n, cch, device=dev).bfloat16()
        ms = timed(lambda: ops.segment_csr(y, pf["idx_ptr"], "max", gather=pf["indices"]))
        row("seg_short_vec4_kernel", f"segment max over pooled clusters, gather fused, C = {cch} bf16",
            n * (cch * 2 + 8) + m * (cch * 2 + 8), ms)
    coordf = p.coord.float()
    ms = timed(lambda: ops.segment_csr(coordf, pf["idx_ptr"], "mean", gather=pf["indices"]))
    row("seg_short_scalar_kernel", "segment mean of coord (3 x fp32), gather fused", n * (12 + 8) + m * (12 + 8), ms)
    # ---- A7 unpool, A8 glue
    for cch in (64,):
        skip = torch.randn(n, cch, device=dev).bfloat16()
        up = torch.randn(m, cch, device=dev).bfloat16()
        ms = timed(lambda: ops.unpool_add(skip, up, pf["cluster"], gelu=True))
        row("unpool_add_kernel", f"GELU(skip) + GELU(up)[inverse] -> fp32 + bf16, C = {cch}", n * (cch * (2 + 2 + 4 + 2) + 8), ms)
    for cch in (32, 64):
        x = torch.randn(n, cch, device=dev).bfloat16()
        resid = torch.randn(n, cch, device=dev)
        g1, b1, g2, b2 = (torch.randn(cch, device=dev) for _ in range(4))
        ms = timed(lambda: ops.ln_residual(x, g1, b1, 1e-5, residual=resid, gamma2=g2, beta2=b2, out2_dtype=torch.bfloat16))
        row(f"ln_residual_vec_kernel<{cch // 8} lanes>", f"LN(x) + residual -> fp32, LN2 -> bf16, C = {cch}", n * cch * (2 + 4 + 4 + 2), ms)
    x = torch.randn(n, 128, device=dev).bfloat16()
    ms = timed(lambda: ops.affine_act(x, None, None, "gelu", want_f32=False, want_bf16=True))
    row("gelu_bf16_kernel", "exact GELU of the MLP hidden activation (between the two library GEMMs), 128 channels", n * 128 * 4, ms)
    pooled = torch.randn(m, 64, device=dev).bfloat16()
    s_, t_ = torch.randn(64, device=dev), torch.randn(64, device=dev)
    ms = timed(lambda: ops.affine_act(pooled, s_, t_, "gelu", want_f32=True, want_bf16=True))
    row("affine_act_kernel", "eval BatchNorm + GELU after pooling -> fp32 + bf16, C = 64", m * 64 * (2 + 4 + 2), ms)
    ms = timed(lambda: ops.offset_to_batch(p.offset, n))
    row("offset_to_batch_kernel", "batch id per row (binary search over the offsets)", n * 8, ms)
    idx4 = ops.pool_gather_ints(p.grid_coord, p.batch, None, 0, want_grid=False, want_batch=False)[2]
    ms = timed(lambda: ops.subm_build_map(idx4, 3))
    row("subm_insert + subm_lookup (k = 3)", "hash of [b,x,y,z] -> (N, 27) int32 neighbour table (latency-bound probes)",
        n * (16 + 27 * 4 + 16), ms, launches=4)
    pred = torch.randn(n, 6, 4, device=dev)
    ms = timed(lambda: ops.keypoint_aggregate(p.coord.float(), pred, p.offset))
    row("keypoint_aggregate_kernel", "per (scene, keypoint) argmax vote of coord + offset", n * (6 * 16 + 12), ms)
    out = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out", "r2_hbm_kernels.json")
    with open(out, "w") as f:
        json.dump([dict(kernel=a, what=b, bytes=c, ms=d, gbs=e, frac=g, launches=h) for a, b, c, d, e, g, h in ROWS], f, indent=0)


if __name__ == "__main__":
    main()
