"""Latency of the device-resident serving path (infer.detect_keypoints) for ONE raw scan: host scan -> H2D ->
normalise -> first point per voxel -> OffsetKeypointPTv3 -> vote -> keypoints on the host.  Run on the GPU box."""
import os
import sys
import time

import numpy as np
import torch

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
import bench  # noqa: E402
from ptv3_b200 import infer, synth  # noqa: E402


def main():
    dev = torch.device("cuda", 0)
    model, _ = bench.build_model(torch.bfloat16, dev)
    for points, scale in ((40000, 1.0), (100000, 4.1), (150000, 5.0)):
        coord, feat = synth.pig_scan(points, scale=scale, seed=3)
        coord = (coord * 1000.0).astype(np.float32)                      # millimetres, like the Qt client sends
        c_pin, f_pin = torch.from_numpy(coord).pin_memory(), torch.from_numpy(feat).pin_memory()
        # grid chosen so that the normalised scan has the same voxel count as the bench scenes (0.02 m on the scan)
        grid = 0.02 / float(np.linalg.norm(coord / 1000.0 - (coord / 1000.0).mean(0), axis=1).max())

        def once():
            c, f = c_pin.to(dev, non_blocking=True), f_pin.to(dev, non_blocking=True)
            return infer.detect_keypoints(model, c, f, grid, mode="weighted").cpu()

        for _ in range(5):
            kp = once()
        torch.cuda.synchronize()
        ts = []
        for _ in range(30):
            t0 = time.perf_counter()
            kp = once()
            ts.append((time.perf_counter() - t0) * 1e3)
        ts = np.sort(np.asarray(ts))
        print(f"{points} raw points (scale {scale}): median {np.median(ts):.2f} ms, p90 {ts[int(0.9 * len(ts))]:.2f} ms per scan "
              f"-> {points / np.median(ts) * 1e3 / 1e6:.1f} M raw points/s; keypoints finite: {bool(torch.isfinite(kp).all())}")


main()
