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
        elif len(shape) == 1 and leaf == "weight":      # LN / BN gamma
            w = rng.uniform(0.8, 1.2, shape)
        elif len(shape) == 1:                             # biases, beta
            w = rng.normal(0.0, 0.05, shape)
        else:
            fan_in = int(np.prod(shape[1:])) if len(shape) == 2 else int(np.prod(shape[1:]))
            w = rng.uniform(-1.0, 1.0, shape) * (3.0 / fan_in) ** 0.5
        out[name] = w.astype(np.float32)
    return out
