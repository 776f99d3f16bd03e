"""bench.py -- PTv3 (PT-v3m1 + offset-keypoint head) forward throughput on B200.

  python bench.py --gpus N --steps K --warmup W            our arm (CUDA kernels through the C ABI)
  python bench.py --impl reference --gpus N ...             the reference's CPU path (oracle port) on the host cores

Workload (BASELINE.json configs[1]): 16 synthetic 100k-point pig scans per GPU (SURVEY.md 8d generator, scale 4.1,
GridSample 0.02 m -> ~72.9k voxels per scene), bf16 inference, keypoint-offset head.  A "step" is one forward of
the batch.  `value` = input voxels / s with inputs resident in HBM; `e2e` = the same through the public call with
pinned HOST inputs (H2D inside the timed region) and a D2H read of the per-scene keypoints.  One JSON line on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

REPO = os.path.dirname(os.path.abspath(__file__))
if REPO not in sys.path:
    sys.path.insert(0, REPO)

import numpy as np  # noqa: E402
import torch  # noqa: E402

CFG = dict(in_channels=4, order=("z", "z-trans", "hilbert", "hilbert-trans"), stride=(2, 2, 2, 2),
           enc_depths=(2, 2, 2, 6, 2), enc_channels=(32, 64, 128, 256, 512), enc_num_head=(2, 4, 8, 16, 32),
           enc_patch_size=(1024,) * 5, dec_depths=(2, 2, 2, 2), dec_channels=(64, 64, 128, 256),
           dec_num_head=(4, 4, 8, 16), dec_patch_size=(1024,) * 4, shuffle_orders=True)
GRID = 0.02
METRIC = "ptv3_forward_voxels_per_sec"
UNIT = "voxels/s"

# BASELINE.json configs[0..2] as bench workloads (--config 1 | 2 | 3; 2 is the headline the metric is quoted on)
WORKLOADS = {
    1: dict(scenes=1, points=40000, scale=1.0, in_channels=4, head="offset",
            name="configs[0]: one 40k-pt pig scan, batch 1, keypoint-offset head"),
    2: dict(scenes=16, points=100000, scale=4.1, in_channels=4, head="offset",
            name="configs[1]: 16 x 100k-pt pig scans, keypoint-offset head"),
    3: dict(scenes=8, points=150000, scale=5.0, in_channels=6, head="segmentor",
            name="configs[2]: semseg-pt-v3m1-0-base shape, 8 ScanNet-sized scenes (150k pts, C_in 6, 20 classes)"),
    # configs[4]: configs/concerto/semseg-ptv3-large-v1m1-waymo-6a-lin.py -- PT-v3m2 encoder (3,3,3,12,3) x
    # (64,128,256,512,768), frozen, + Linear(1728, 22) head trained with CE + Lovasz; batch 64 over 8 GPUs = 8 sweeps per
    # GPU; Waymo-shaped 64-beam sweeps of ~180 k returns, scaled by 0.2 and voxelised at 0.01 (the config's RandomScale /
    # GridSample: 5 cm cells of the original sweep).  --mode train only.
    5: dict(scenes=8, points=180000, scale=0.2, in_channels=9, head="lin", grid=0.01,
            name="configs[4]: PTv3-large (PT-v3m2) linear-probe TRAINING step, 8 Waymo-shaped 180k-pt sweeps per GPU"),
}

LARGE_M2 = dict(type="PT-v3m2", in_channels=9, order=("z", "z-trans", "hilbert", "hilbert-trans"), stride=(2, 2, 2, 2),
                enc_depths=(3, 3, 3, 12, 3), enc_channels=(64, 128, 256, 512, 768), enc_num_head=(4, 8, 16, 32, 48),
                enc_patch_size=(1024,) * 5, mlp_ratio=4, qkv_bias=True, drop_path=0.3, shuffle_orders=True,
                pre_norm=True, enable_rpe=False, enable_flash=True, traceable=False, mask_token=False, enc_mode=True,
                freeze_encoder=False)


def peaks():
    try:
        with open(os.path.join(REPO, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return dict(hbm=p["hbm_gbs"], tf_burst=p["bf16_tflops"], tf_sust=p["bf16_tflops_sustained"], src="measured")
    except Exception:
        return dict(hbm=6650.0, tf_burst=1590.0, tf_sust=1400.0, src="fallback")


def ncu_traffic(kernel_substr):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernel, from the committed summary of
    one `ncu --set full` capture (profiles/r02_ncu_attention_summary.json for the attention kernel of this round, else
    profiles/r01_ncu_full_summary.json); None if that kernel was not captured."""
    try:
        for name in ("r02_ncu_attention_summary.json", "r01_ncu_full_summary.json"):
            with open(os.path.join(REPO, "profiles", name)) as f:
                rows = [r for r in json.load(f) if kernel_substr in r["kernel"] and r.get("dram_read_MB") is not None]
            if rows:
                break
        if not rows:
            return None, "kernel not in the committed ncu capture"
        if name.startswith("r02"):
            rows = rows[:1]          # the default kernel's own capture (the other rows are ablations / other designs)
        b = sum((r["dram_read_MB"] + r["dram_write_MB"]) * 1e6 for r in rows) / len(rows)
        return round(b), f"bytes per launch, mean of {len(rows)} captured launch(es) (profiles/{name})"
    except Exception as e:  # noqa: BLE001
        return None, f"no ncu summary ({type(e).__name__})"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=["nvidia-smi unavailable"])
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            parts = [p.strip() for p in ln.split(",")]
            if len(parts) < 6:
                continue
            try:
                sm.append(float(parts[0]))
                mx.append(float(parts[1]))
            except ValueError:
                continue
            for nm, v in zip(names, parts[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return dict(sm_mhz=float(np.median(sm)) if sm else None, sm_max_mhz=max(mx) if mx else None,
                    reasons=sorted(reasons), samples=len(sm))


def make_raw_scenes(n_scenes, points, scale, seed0, in_channels=4):
    from ptv3_b200 import synth
    return [synth.pig_scan(points, scale=scale, seed=seed0 + i, in_channels=in_channels) for i in range(n_scenes)]


def build_model(dtype, device, in_channels=4, head="offset"):
    from ptv3_b200 import model as M, synth
    conf = dict(CFG, in_channels=in_channels, compute_dtype=dtype)
    if head == "segmentor":      # configs/scannet/semseg-pt-v3m1-0-base.py: DefaultSegmentorV2, 20 classes
        m = M.DefaultSegmentorV2(num_classes=20, backbone_out_channels=64, backbone=conf).eval()
    else:
        m = M.OffsetKeypointPTv3(backbone_conf=conf).eval()
    shapes = {k: tuple(v.shape) for k, v in m.state_dict().items() if not k.endswith("num_batches_tracked")}
    W = {k: torch.from_numpy(v) for k, v in synth.make_weights(shapes, seed=0).items()}
    m.load_state_dict(W, strict=False)
    return m.to(device), W


def shard_scenes(n_scenes, rank, world, scaling):
    """Scene ids of this rank.  weak: every rank owns n_scenes scenes of its own; strong: the n_scenes of the job are
    dealt round-robin (BASELINE configs[2]: batch 8 at 1 / 2 / 4 / 8 GPUs = 8 ... 1 scene per GPU)."""
    if scaling == "strong":
        return [i for i in range(n_scenes) if i % world == rank]
    return [1000 * rank + i for i in range(n_scenes)]


def voxelize_on_device(raw, device, grid=GRID):
    """GridSample (our CUDA kernel, first point per voxel) per scene -> host numpy batch dict."""
    from ptv3_b200 import ops
    coords, feats, grids, counts = [], [], [], []
    for coord, feat in raw:
        c = torch.from_numpy(coord).to(device)
        r = ops.gridsample(c, grid, "fnv")
        iu = r["idx_unique"]
        coords.append(c[iu].cpu())
        feats.append(torch.from_numpy(feat).to(device)[iu].cpu())
        grids.append(r["grid_coord"][iu].cpu())
        counts.append(int(iu.numel()))
    return dict(coord=torch.cat(coords), feat=torch.cat(feats), grid_coord=torch.cat(grids),
                offset=torch.tensor(np.cumsum(counts), dtype=torch.int64)), counts


def attn_flops(rec_meta):
    """Algorithmic FLOPs of one window-attention launch: 4 * C * sum_w q_len * k_len (QK^T + PV)."""
    w = rec_meta["windows"][: int(rec_meta["n_windows"].item())].cpu().numpy()
    pairs = float((w[:, 1] * w[:, 3]).sum())
    return 4.0 * rec_meta["h"] * rec_meta["d"] * pairs, pairs * rec_meta["h"]


def conv_flops(rec_meta):
    """Algorithmic FLOPs of one submanifold conv launch: 2 * Cin * Cout * (number of present neighbours)."""
    present = float((rec_meta["nbr"] >= 0).sum().item())
    return 2.0 * rec_meta["cin"] * rec_meta["cout"] * present, present


def roofline_from_records(recs, step_ms, pk, clocks):
    """recs = ops.profile_stop(): (op, ms, meta) per instrumented launch of one step -> (roofline dict of the op with
    the largest share, per-op ms)."""
    tot = {}
    for name, t, meta in recs:
        if name == "attn_window":
            fl, units = attn_flops(meta)
        elif name == "mlp_fused":      # fc1 + fc2: 2 * (2 * n * c * hidden)
            fl, units = 4.0 * meta["n"] * meta["c"] * meta["hidden"], float(meta["n"])
        elif name == "proj_mlp_fused":  # + the attention projection
            fl, units = 4.0 * meta["n"] * meta["c"] * meta["hidden"] + 2.0 * meta["n"] * meta["c"] ** 2, float(meta["n"])
        elif name == "subm_conv":
            fl, units = conv_flops(meta)
        else:
            continue
        e = tot.setdefault(name, dict(ms=0.0, flops=0.0, units=0.0, launches=0))
        e["ms"] += t
        e["flops"] += fl
        e["units"] += units
        e["launches"] += 1
    ops_ms = {k: round(v["ms"], 3) for k, v in tot.items()}
    ops_ms["step_instrumented"] = round(step_ms, 3)
    top = max(tot, key=lambda k: tot[k]["ms"])
    e = tot[top]
    ach = e["flops"] / (e["ms"] * 1e-3) / 1e12
    roof = dict(kernel=top, bound="tensor", achieved=round(ach, 2), peak=pk["tf_sust"], unit="TFLOP/s",
                frac=round(ach / pk["tf_sust"], 4), traffic=None, peak_source=pk["src"] + " (sustained bf16 GEMM)",
                launches_per_step=e["launches"], ms_per_step=round(e["ms"], 3),
                share_of_step=round(e["ms"] / step_ms, 3))
    roof["traffic"], roof["traffic_note"] = ncu_traffic({"attn_window": "attn_window_f16", "subm_conv": "subm_conv_tc",
                                                         "mlp_fused": "mlp_fused"}.get(top, top))
    if top == "attn_window":   # head_dim 16: the exp pipe, not the tensor pipe, bounds this kernel
        roof["exp_per_s"] = round(e["units"] / (e["ms"] * 1e-3) / 1e12, 3)
        roof["exp_peak_mufu_per_s"] = round(16 * 148 * (clocks.get("sm_mhz") or 1965.0) * 1e6 / 1e12, 3)
        roof["limiter"] = ("mufu: 16 ex2 / clk / SM measured (tools/ubench/pipes.cu; the packed f16x2 / bf16x2 forms "
                           "are two MUFU ops each); frac_of_limiter = exp_per_s / exp_peak_mufu_per_s")
        roof["frac_of_limiter"] = round(roof["exp_per_s"] / roof["exp_peak_mufu_per_s"], 4)
    return roof, ops_ms


def run_ours(args):
    from ptv3_b200 import _lib, ops
    from ptv3_b200.model import aggregate_keypoints
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=device)
    dtype = torch.bfloat16 if args.dtype == "bf16" else torch.float32
    model, _ = build_model(dtype, device, args.in_channels, args.head)
    from ptv3_b200 import synth
    ids = shard_scenes(args.scenes, rank, world, args.scaling)
    if not ids:
        raise SystemExit(f"--scaling strong needs at least one scene per rank ({args.scenes} scenes, {world} ranks)")
    raw = [synth.pig_scan(args.points, scale=args.scale, seed=i, in_channels=args.in_channels) for i in ids]
    host, counts = voxelize_on_device(raw, device)
    n_vox = int(sum(counts))
    pinned = {k: v.pin_memory() for k, v in host.items()}
    resident = {k: v.to(device) for k, v in host.items()}
    h2d_bytes = int(sum(v.numel() * v.element_size() for v in pinned.values()))

    out_key = "seg_logits" if args.head == "segmentor" else "pred"

    def step_resident():
        return model(dict(resident))[out_key]

    # e2e: every step copies ITS inputs from pinned host memory and reads ITS keypoints back.  The copy of step i+1
    # runs on a side stream while step i computes (double-buffered device inputs); all of it is inside the timed region.
    copy_stream = torch.cuda.Stream(device=device)
    staged = {}

    def stage_inputs():
        with torch.cuda.stream(copy_stream):
            d = {k: v.to(device, non_blocking=True) for k, v in pinned.items()}
            ev = torch.cuda.Event()
            ev.record(copy_stream)
        staged["next"] = (d, ev)

    def step_e2e():
        if "next" not in staged:
            stage_inputs()
        d, ev = staged.pop("next")
        torch.cuda.current_stream().wait_event(ev)
        for v in d.values():
            v.record_stream(torch.cuda.current_stream())
        stage_inputs()                                                    # H2D of the next step overlaps this one
        pred = model(d)[out_key]
        if args.head == "segmentor":
            return pred.argmax(1).to(torch.uint8).cpu()                   # D2H read: predicted class of every voxel
        kp = aggregate_keypoints(d["coord"], pred, d["offset"])            # (scenes, 6, 3) on device, one kernel
        return kp.cpu()                                                   # D2H read of the step's result

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            import torch.distributed as dist
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        a.record()
        for _ in range(steps):
            out = fn()
        b.record()
        torch.cuda.synchronize()
        wall = (time.perf_counter() - t0) * 1e3
        barrier()
        ms = max(a.elapsed_time(b), 0.0)
        t = torch.tensor([ms, wall], device=device, dtype=torch.float64)
        if world > 1:
            import torch.distributed as dist
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0]), float(t[1]), out

    torch.manual_seed(0)
    for _ in range(max(args.warmup, 3)):
        step_resident()
    sampler = ClockSampler(local)
    sampler.start()
    l0 = _lib.launch_count()
    ms, _, _ = timed(step_resident, args.steps)
    launches = _lib.launch_count() - l0
    clocks = sampler.stop()
    for _ in range(2):
        step_e2e()
    _, e2e_wall, kp = timed(step_e2e, args.steps)
    staged.clear()
    d2h_bytes = int(kp.numel() * kp.element_size())

    total_vox = torch.tensor([n_vox], device=device, dtype=torch.float64)
    if world > 1:
        import torch.distributed as dist
        dist.all_reduce(total_vox)
    total_vox = float(total_vox.item())
    value = total_vox * args.steps / (ms * 1e-3)
    e2e = total_vox * args.steps / (e2e_wall * 1e-3)

    # ---- roofline leg: per-op CUDA events over instrumented steps (rank 0) -------------------------------
    roof, ops_ms = None, {}
    if rank == 0:
        pk = peaks()
        # per-op CUDA events need the op-by-op path: same kernels on the same stream, without the one-call-per-Block
        # wrapper (ops.block_fwd) that hides them from the host
        ops.NATIVE_BLOCK = False
        step_resident()
        torch.cuda.synchronize()
        ops.profile_start()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        step_resident()
        b.record()
        recs = ops.profile_stop()
        ops.NATIVE_BLOCK = True
        step_ms = a.elapsed_time(b)
        roof, ops_ms = roofline_from_records(recs, step_ms, pk, clocks)
    # ---- GridSample leg (BASELINE metric: "GridSample reported separately in raw points/s"), raw scans resident ----
    gs_leg = None
    if rank == 0:
        raw_dev = [torch.from_numpy(c).to(device) for c, _ in raw]
        for c in raw_dev[:2]:
            ops.gridsample(c, GRID, "fnv")
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for c in raw_dev:
            ops.gridsample(c, GRID, "fnv")
        b.record()
        torch.cuda.synchronize()
        gs_ms = a.elapsed_time(b)
        n_raw = int(sum(c.shape[0] for c in raw_dev))
        allc = torch.cat(raw_dev)
        off = torch.tensor(np.cumsum([c.shape[0] for c in raw_dev]), device=device)
        for _ in range(2):
            ops.gridsample_batch(allc, off, GRID, "fnv")
        a2, b2 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a2.record()
        for _ in range(3):
            ops.gridsample_batch(allc, off, GRID, "fnv")
        b2.record()
        torch.cuda.synchronize()
        gsb_ms = a2.elapsed_time(b2) / 3
        gs_leg = dict(raw_points_per_s=round(n_raw / (gsb_ms * 1e-3), 1), ms_per_batch=round(gsb_ms, 3), raw_points=n_raw,
                      note="ptv3_gridsample_batch: every scene of the step in one launch chain (floor, per-scene min, FNV "
                           "key, 64-bit stable radix sort + scene pass, unique, count, inverse, select), one host read",
                      per_scene_calls=dict(raw_points_per_s=round(n_raw / (gs_ms * 1e-3), 1), ms_per_batch=round(gs_ms, 3),
                                           note="ptv3_gridsample per scene, one host read each (round 1 path)"))
        del raw_dev, allc
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:      # reported at N = 1 only
        cpu = cpu_baseline(args, seconds=20.0)

    if rank == 0:
        seg = args.head == "segmentor"
        head_name = "20-class segmentor head (DefaultSegmentorV2)" if seg else "offset-keypoint head"
        result_name = ("per-voxel class ids (uint8 argmax of the logits, taken on the device)" if seg
                       else "per-scene keypoints (scenes,6,3) fp32 after on-device aggregation")
        line = dict(metric=METRIC, value=round(value, 1), unit=UNIT, n_gpus=world, steps=args.steps,
                    warmup=max(args.warmup, 3), ms_per_step=round(ms / args.steps, 3), higher_is_better=True,
                    scaling=args.scaling, vs_baseline=None, dtype=args.dtype, data="synthetic",
                    config=dict(workload=f"PTv3-v1m1 + {head_name} forward, {args.scenes} scenes x "
                                         f"{args.points} raw pts (scale {args.scale}, grid {GRID} m) "
                                         f"{'in total, split over the ranks' if args.scaling == 'strong' else 'per GPU'}"
                                         f" ({args.workload_name})",
                                voxels_per_gpu=n_vox, scenes_per_gpu=len(ids), raw_points_per_scene=args.points,
                                patch=1024, parallelism=f"scene-sharded x{world} (no collective in forward)",
                                l2="per-step working set (qkv 224 MB at stage 0) exceeds the 126 MB L2"),
                    e2e=dict(value=round(e2e, 1), unit=UNIT, h2d_bytes_per_step=h2d_bytes,
                             d2h_bytes_per_step=d2h_bytes, ms_per_step=round(e2e_wall / args.steps, 3),
                             result=result_name),
                    gpu_launches=int(launches), clocks=clocks, roofline=roof, cpu_baseline=cpu, op_ms=ops_ms,
                    gridsample=gs_leg)
        print(json.dumps(line))
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


def cpu_sample_scene(args, seconds, steps=1):
    """One scene of the same workload, scaled so that `steps` oracle forwards take about `seconds` on this host."""
    from oracle import gridsample as G
    from ptv3_b200 import synth
    cores = os.cpu_count() or 1
    est_rate = 9000.0 * max(1.0, (cores / 8.0) ** 0.5)          # voxels/s, indicative
    target = min(72900.0, max(4000.0, est_rate * seconds / steps))
    full = 72900.0 * (args.scale / 4.1) ** 2 * min(1.0, args.points / 100000.0) ** 0.5     # voxels of one full scene
    target = min(full, target)
    scale = args.scale * (target / full) ** 0.5
    if getattr(args, "head", "") == "lin":       # a bounded part of one Waymo-shaped sweep: fewer rays, same geometry
        target = max(3000.0, min(150000.0, 1500.0 * max(1.0, (cores / 8.0) ** 0.5) * seconds / steps))
        coord, feat = synth.lidar_sweep(int(target * 1.25), seed=0, in_channels=9, beams=64, max_range=75.0)
        gs = G.grid_sample((coord * args.scale).astype(np.float32), WORKLOADS[5]["grid"])
        iu = gs["idx_unique"]
        return (coord * args.scale).astype(np.float32)[iu], feat[iu], gs["grid_coord"][iu], args.scale
    coord, feat = synth.pig_scan(args.points, scale=scale, seed=0, in_channels=getattr(args, "in_channels", 4))
    gs = G.grid_sample(coord, GRID)
    iu = gs["idx_unique"]
    return coord[iu], feat[iu], gs["grid_coord"][iu], scale


def cpu_forward(W, scene, head="offset"):
    from oracle import ptv3 as P
    c, f, g, _ = scene
    perms = [torch.randperm(4).numpy() for _ in range(5)]
    if head == "lin":      # configs[4]: frozen PT-v3m2 encoder (no graph), up-cast, Linear head with CE loss + head grads
        from oracle import sonata as SO
        cfg = {k: v for k, v in LARGE_M2.items() if k != "type"}
        pt = SO.backbone_forward(W, c, g, f, [len(c)], cfg=cfg, perms=perms)
        feat = SO.upcast_features(pt)
        w = W["seg_head.weight"].clone().requires_grad_(True)
        b = W["seg_head.bias"].clone().requires_grad_(True)
        with torch.enable_grad():
            loss = torch.nn.functional.cross_entropy(torch.nn.functional.linear(feat, w, b),
                                                     torch.zeros(len(c), dtype=torch.long))
            loss.backward()
        return loss.detach()
    pt = P.backbone_forward(W, c, g, f, [len(c)], cfg=dict(CFG, in_channels=f.shape[1]), perms=perms)
    return P.seg_head(W, pt.feat) if head == "segmentor" else P.offset_head(W, pt.feat)


def cpu_weights(args):
    from ptv3_b200 import model as M, synth
    conf = dict(CFG, in_channels=args.in_channels)
    if args.head == "lin":
        m = M.DefaultSegmentorV2(num_classes=22, backbone_out_channels=1728, backbone=dict(LARGE_M2))
    else:
        m = (M.DefaultSegmentorV2(num_classes=20, backbone_out_channels=64, backbone=conf) if args.head == "segmentor"
             else M.OffsetKeypointPTv3(backbone_conf=conf))
    shapes = {k: tuple(v.shape) for k, v in m.state_dict().items() if not k.endswith("num_batches_tracked")}
    return {k: torch.from_numpy(v) for k, v in synth.make_weights(shapes, seed=0).items()}


def cpu_baseline(args, seconds=20.0):
    """The oracle port of the reference forward (fp32, dense-branch attention, restated spconv) on the host cores."""
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    W = cpu_weights(args)
    scene = cpu_sample_scene(args, seconds / 4.0)                 # 1 warm-up + 3 timed forwards inside the budget
    cpu_forward(W, scene, args.head)                              # warm-up (first-call allocations, thread pool)
    ts = []
    for _ in range(3):
        t0 = time.perf_counter()
        cpu_forward(W, scene, args.head)
        ts.append(time.perf_counter() - t0)
    dt = float(np.median(ts))
    n = len(scene[0])
    return dict(value=round(n / dt, 1), unit=UNIT, cores=cores, kind="port", dtype="f32",
                sample=f"1 scene of the workload at scale {scene[3]:.2f} -> {n} voxels, median of 3 forwards after a "
                       f"warm-up, fp32 (the GPU arm computes in {args.dtype}), {dt:.2f} s each; the oracle PORT of the "
                       f"reference forward, not the reference's own Python (which cannot travel to the GPU box)")


def run_reference(args):
    """--impl reference: the reference's CPU path (oracle port; the reference itself is Python that cannot travel to
    the GPU box) timed on all host cores; each step = one bounded-sample scene forward."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    W = cpu_weights(args)
    total = args.steps + args.warmup
    scene = cpu_sample_scene(args, seconds=150.0, steps=total)
    n = len(scene[0])
    torch.manual_seed(0)
    for _ in range(args.warmup):
        cpu_forward(W, scene, args.head)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_forward(W, scene, args.head)
    dt = time.perf_counter() - t0
    v = n * args.steps / dt
    sample = f"1 scene per step at scale {scene[3]:.2f} -> {n} voxels (of the {args.scenes}-scene batch), fp32"
    lin = args.head == "lin"
    line = dict(impl="reference", metric="ptv3_train_step_voxels_per_sec" if lin else METRIC, value=round(v, 1), unit=UNIT, n_gpus=int(os.environ.get("WORLD_SIZE", "1")),
                steps=args.steps, warmup=args.warmup, ms_per_step=round(dt / args.steps * 1e3, 3),
                higher_is_better=True, scaling=args.scaling, vs_baseline=None, dtype="f32", data="synthetic",
                config=dict(workload=(f"PTv3-large (PT-v3m2) frozen-encoder forward + Linear(1728, 22) head loss / gradient on "
                                      f"the host CPU ({args.workload_name}; oracle port, restated spconv), {sample}") if lin
                            else f"PTv3-v1m1 + {'20-class segmentor' if args.head == 'segmentor' else 'offset-keypoint'} "
                                 f"head forward on the host CPU ({args.workload_name}; oracle port of the "
                                 f"reference, enable_flash=False semantics one scene per forward, restated spconv), {sample}"),
                cpu_baseline=dict(value=round(v, 1), unit=UNIT, cores=cores, kind="port", sample=sample),
                e2e=dict(value=round(v, 1), unit=UNIT, h2d_bytes_per_step=0, d2h_bytes_per_step=0),
                gpu_launches=0)
    print(json.dumps(line))


def run_train(args):
    """SURVEY 8f rank 1 / BASELINE configs[4] shape of work: one training step = forward in train() mode (BatchNorm
    batch statistics, DropPath), the offset-keypoint loss, backward through every kernel, gradient all-reduce over
    NCCL when N > 1 (ptv3_b200/ddp.py, overlapped with backward), fused AdamW.  Same scenes per GPU as the forward
    bench; not the headline line (run with --mode train)."""
    from ptv3_b200 import _lib
    from ptv3_b200.ddp import GradSync
    sys.path.insert(0, os.path.join(REPO, "tests"))
    from make_golden_names import make_target
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    import torch.distributed as dist
    if world > 1:
        dist.init_process_group("nccl", device_id=device)
    dtype = torch.bfloat16 if args.dtype == "bf16" else torch.float32
    lin = args.head == "lin"
    if lin:       # BASELINE configs[4]
        from ptv3_b200 import model as M, synth
        model = M.DefaultSegmentorV2(
            num_classes=22, backbone_out_channels=1728, backbone=dict(LARGE_M2, compute_dtype=dtype), freeze_backbone=True,
            criteria=[dict(type="CrossEntropyLoss", loss_weight=1.0, ignore_index=-1),
                      dict(type="LovaszLoss", mode="multiclass", loss_weight=1.0, ignore_index=-1)])
        shapes = {k: tuple(v.shape) for k, v in model.state_dict().items()}
        model.load_state_dict({k: torch.from_numpy(v) for k, v in synth.make_weights(shapes, seed=0).items()},
                              strict=False)
        model = model.to(device)
        raw = []
        for i in range(args.scenes):      # rays without a return (sky) are dropped by the generator: oversample, keep `points`
            c, f = synth.lidar_sweep(int(args.points * 2.2), seed=1000 * rank + i, in_channels=9, beams=64, max_range=75.0)
            keep = np.random.default_rng(i).permutation(len(c))[: args.points]
            raw.append(((c[keep] * args.scale).astype(np.float32), f[keep]))
        grid = WORKLOADS[5]["grid"]
    else:
        model, _ = build_model(dtype, device)
        raw = make_raw_scenes(args.scenes, args.points, args.scale, seed0=1000 * rank)
        grid = GRID
    model.train()
    if getattr(args, "sync_bn", False) and world > 1:
        from ptv3_b200.ddp import convert_sync_batchnorm
        convert_sync_batchnorm(model)
    host, counts = voxelize_on_device(raw, device, grid)
    n_vox = int(sum(counts))
    if lin:
        rng = np.random.default_rng(7 + rank)
        seg = rng.integers(0, 22, n_vox).astype(np.int64)
        seg[rng.random(n_vox) < 0.05] = -1
        host["segment"] = torch.from_numpy(seg)
    else:
        host["target"] = torch.from_numpy(make_target(n_vox, 7 + rank))
    pinned = {k: v.pin_memory() for k, v in host.items()}
    resident = {k: v.to(device) for k, v in host.items()}
    h2d_bytes = int(sum(v.numel() * v.element_size() for v in pinned.values()))
    trainable = [p for p in model.parameters() if p.requires_grad]
    if lin:       # the config's optimizer: AdamW(lr 0.002, weight_decay 0.005) over what is not frozen (the head)
        opt = torch.optim.AdamW(trainable, lr=0.002, weight_decay=0.005, fused=True)
    else:
        opt = torch.optim.AdamW(trainable, lr=1e-4, weight_decay=0.05, fused=True)
    sync = GradSync(trainable) if world > 1 else None

    def step(batch):
        if sync is not None:
            sync.zero_grad()
        else:
            opt.zero_grad(set_to_none=True)
        out = model(dict(batch))
        out["loss"].backward()
        if sync is not None:
            sync.finish()
        opt.step()
        return out["loss"]

    def step_e2e():
        d = {k: v.to(device, non_blocking=True) for k, v in pinned.items()}
        return float(step(d).item())                                  # D2H read of the step's loss

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        a.record()
        for _ in range(steps):
            out = fn()
        b.record()
        torch.cuda.synchronize()
        wall = (time.perf_counter() - t0) * 1e3
        barrier()
        t = torch.tensor([a.elapsed_time(b), wall], device=device, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0]), float(t[1]), out

    torch.manual_seed(rank)
    for _ in range(max(args.warmup, 3)):
        step(resident)
    sampler = ClockSampler(local)
    sampler.start()
    l0 = _lib.launch_count()
    ms, _, loss = timed(lambda: step(resident), args.steps)
    launches = _lib.launch_count() - l0
    clocks = sampler.stop()
    step_e2e()
    _, e2e_wall, loss_e2e = timed(step_e2e, args.steps)
    total_vox = torch.tensor([n_vox], device=device, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(total_vox)
    total_vox = float(total_vox.item())
    n_params = sum(p.numel() for p in model.parameters())
    roof, ops_ms = None, {}
    if rank == 0 and lin:      # roofline leg, as in run_ours: one instrumented step on the op-by-op path (frozen backbone)
        from ptv3_b200 import ops
        ops.NATIVE_BLOCK = False
        step(resident)
        torch.cuda.synchronize()
        ops.profile_start()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        step(resident)
        b.record()
        recs = ops.profile_stop()
        ops.NATIVE_BLOCK = True
        if recs:
            roof, ops_ms = roofline_from_records(recs, a.elapsed_time(b), peaks(), clocks)
    if lin:
        what = (f"PTv3-large (PT-v3m2 encoder 3-3-3-12-3 x 64-128-256-512-768, frozen, train() mode with DropPath 0.3) + "
                f"Linear(1728, 22) head TRAINING step (fwd + CE + Lovasz + head bwd + AdamW), {args.scenes} Waymo-shaped "
                f"sweeps x {args.points} returns (x {args.scale}, grid {grid} m) per GPU ({args.workload_name})")
    else:
        what = (f"PTv3-v1m1 + offset-keypoint head TRAINING step (fwd + loss + bwd + AdamW), {args.scenes} scenes x "
                f"{args.points} raw pts (scale {args.scale}, grid {GRID} m) per GPU")
    if rank == 0:
        print(json.dumps(dict(
            metric="ptv3_train_step_voxels_per_sec", value=round(total_vox * args.steps / (ms * 1e-3), 1), unit=UNIT,
            n_gpus=world, steps=args.steps, warmup=max(args.warmup, 3), ms_per_step=round(ms / args.steps, 3),
            higher_is_better=True, scaling="weak", vs_baseline=None, dtype=args.dtype, data="synthetic", mode="train",
            config=dict(workload=what, voxels_per_gpu=n_vox, params=n_params,
                        trainable_params=sum(p.numel() for p in trainable),
                        parallelism=f"replicas x{world}, scenes sharded per rank, bucketed gradient all-reduce "
                                    f"({'NCCL AVG, overlapped with backward' if world > 1 else 'none at N=1'})"),
            e2e=dict(value=round(total_vox * args.steps / (e2e_wall * 1e-3), 1), unit=UNIT, h2d_bytes_per_step=h2d_bytes,
                     d2h_bytes_per_step=4, ms_per_step=round(e2e_wall / args.steps, 3), result="loss (fp32 scalar)"),
            gpu_launches=int(launches), clocks=clocks, roofline=roof, op_ms=ops_ms, loss=round(float(loss.item()), 5),
            peak_mem_gib=round(torch.cuda.max_memory_allocated() / 2 ** 30, 2))))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--dtype", default="bf16", choices=["bf16", "f32"])
    ap.add_argument("--config", type=int, default=2, choices=[1, 2, 3, 5],
                    help="BASELINE.json workload: 1 = configs[0] (one 40k-pt scan), 2 = configs[1] (16 x 100k-pt scans, the "
                         "headline), 3 = configs[2] (8 ScanNet-sized scenes, C_in 6, 20-class segmentor), 5 = configs[4] (PTv3-large "
                         "linear-probe training step on Waymo-shaped sweeps; implies --mode train)")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: --scenes per GPU (default); strong: --scenes in total, dealt over the ranks")
    ap.add_argument("--scenes", type=int, default=None)
    ap.add_argument("--points", type=int, default=None)
    ap.add_argument("--scale", type=float, default=None)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--sync-bn", action="store_true",
                    help="--mode train: BatchNorm statistics over every rank's rows (the reference's cfg.sync_bn)")
    ap.add_argument("--mode", default="forward", choices=["forward", "train"],
                    help="forward = the headline metric (default); train = training step (SURVEY 8f rank 1)")
    args = ap.parse_args()
    wl = WORKLOADS[args.config]
    for k in ("scenes", "points", "scale"):
        if getattr(args, k) is None:
            setattr(args, k, wl[k])
    args.in_channels, args.head, args.workload_name = wl["in_channels"], wl["head"], wl["name"]
    if args.config == 5:
        args.mode = "train"
    if args.impl == "reference":
        run_reference(args)
    elif args.mode == "train":
        run_train(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
