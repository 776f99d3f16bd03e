"""Synthetic inputs: seedable "pig" surface scans and
deterministic weights keyed by parameter name.  numpy only."""
import zlib

import numpy as np

_PARTS = (  # centre, semi-axes, share of points
    ((0.0, 0.0, 0.55), (0.80, 0.28, 0.30), 0.62),
    ((0.95, 0.0, 0.62), (0.22, 0.16, 0.17), 0.10),
    ((0.5, 0.17, 0.18), (0.07, 0.07, 0.20), 0.07),
    ((0.5, -0.17, 0.18), (0.07, 0.07, 0.20), 0.07),
    ((-0.5, 0.17, 0.18), (0.07, 0.07, 0.20), 0.07),
    ((-0.5, -0.17, 0.18), (0.07, 0.07, 0.20), 0.07),
)


def pig_scan(n_points, scale=1.0, seed=0, in_channels=4):
    """Raw scan: coord float32 (R,3) in metres, feat float32 (R,in_channels)."""
    rng = np.random.default_rng(seed)
    counts = [int(round(n_points * p[2])) for p in _PARTS]
    counts[0] += n_points - sum(counts)
    pts = []
    for (c, ax, _), m in zip(_PARTS, counts):
        d = rng.standard_normal((m, 3))
        d /= np.linalg.norm(d, axis=1, keepdims=True) + 1e-12
        pts.append(d * np.asarray(ax) + np.asarray(c))
    coord = np.concatenate(pts) * scale + rng.normal(0.0, 0.002, (n_points, 3))
    feat = rng.standard_normal((n_points, in_channels))
    perm = rng.permutation(n_points)
    return coord[perm].astype(np.float32), feat[perm].astype(np.float32)


def _rng_for(name, seed):
    return np.random.default_rng([zlib.crc32(name.encode()), seed])


def make_weights(shapes, seed=0):
    """Deterministic float32 weights for a {name: shape} dict using reference
    state_dict names.  Scales mimic torch default init so activations stay O(1)."""
    out = {}
    for name, shape in shapes.items():
        shape = tuple(shape)
        rng = _rng_for(name, seed)
        leaf = name.rsplit(".", 1)[-1]
        if leaf == "num_batches_tracked":
            continue
        if leaf == "running_mean":
            w = rng.normal(0.0, 0.1, shape)
        elif leaf == "running_var":
            w = rng.uniform(0.5, 1.5, shape)
        elif leaf == "gamma":                            # PT-v3m2 LayerScale
            w = rng.uniform(0.5, 1.5, shape)
        elif len(shape) == 1 and leaf == "weight":      # LN / BN gamma
            w = rng.uniform(0.8, 1.2, shape)
        elif len(shape) == 1:                             # biases, beta
            w = rng.normal(0.0, 0.05, shape)
        else:
            fan_in = int(np.prod(shape[1:])) if len(shape) == 2 else int(np.prod(shape[1:]))
            w = rng.uniform(-1.0, 1.0, shape) * (3.0 / fan_in) ** 0.5
        out[name] = w.astype(np.float32)
    return out


def lidar_sweep(n_points, seed=0, in_channels=4, beams=64, max_range=50.0):
    """LiDAR-shaped sweep (BASELINE configs[3]: SemanticKITTI-like 120 k-pt sweeps, grid 0.05 m):
    a 64-beam spinning sensor 1.8 m above a ground plane with random boxes (cars / walls) within `max_range`;
    every return is the nearest hit of its ray.  coord float32 (R,3) metres, feat float32 (R,in_channels)
    (feat[:, 0] = an intensity-like value, the rest N(0,1))."""
    rng = np.random.default_rng(seed)
    per_beam = n_points // beams
    elev = np.deg2rad(np.linspace(-24.8, 2.0, beams))
    az = rng.uniform(0.0, 2 * np.pi, (beams, per_beam))
    d = np.stack([np.cos(elev)[:, None] * np.cos(az), np.cos(elev)[:, None] * np.sin(az),
                  np.sin(elev)[:, None] * np.ones_like(az)], -1).reshape(-1, 3)
    origin = np.array([0.0, 0.0, 1.8])
    t = np.full(len(d), max_range)
    down = d[:, 2] < -1e-3
    t[down] = np.minimum(t[down], -origin[2] / d[down, 2])                    # ground plane z = 0
    n_boxes = 40
    centre = np.concatenate([rng.uniform(-40, 40, (n_boxes, 2)), np.zeros((n_boxes, 1))], 1)
    half = np.concatenate([rng.uniform(0.8, 4.0, (n_boxes, 2)), rng.uniform(0.7, 2.5, (n_boxes, 1))], 1)
    centre[:, 2] = half[:, 2]
    for c, h in zip(centre, half):                                           # slab test per box
        lo, hi = (c - h - origin) / d, (c + h - origin) / d
        tn, tf = np.minimum(lo, hi).max(1), np.maximum(lo, hi).min(1)
        hit = (tn < tf) & (tn > 0.5)
        t[hit] = np.minimum(t[hit], tn[hit])
    keep = t < max_range
    coord = (origin + d * t[:, None])[keep] + rng.normal(0.0, 0.01, (int(keep.sum()), 3))
    feat = rng.standard_normal((len(coord), in_channels))
    feat[:, 0] = np.exp(-t[keep] / 20.0)
    return coord.astype(np.float32), feat.astype(np.float32)
