"""Shared by the PT-v3m2 tests: the two golden configurations of tests/golden/make_golden.py (gold_sonata) and their
seeded two-scene input batch."""
import json
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))

_ORDER = ["z", "z-trans", "hilbert", "hilbert-trans"]
CFGS = dict(
    dec=dict(num_classes=13, out_channels=64,
             backbone=dict(in_channels=6, order=_ORDER, stride=(2, 2, 2, 2), enc_depths=(2, 2, 2, 2, 2),
                           enc_channels=(32, 64, 128, 256, 512), enc_num_head=(2, 4, 8, 16, 32),
                           enc_patch_size=(1024,) * 5, dec_depths=(2, 2, 2, 2), dec_channels=(64, 64, 128, 256),
                           dec_num_head=(4, 4, 8, 16), dec_patch_size=(1024,) * 4, shuffle_orders=True,
                           layer_scale=None, enc_mode=False)),
    enc=dict(num_classes=22, out_channels=32 + 64 + 128 + 256 + 768,
             backbone=dict(in_channels=9, order=_ORDER, stride=(2, 2, 2, 2), enc_depths=(1, 1, 1, 2, 1),
                           enc_channels=(32, 64, 128, 256, 768), enc_num_head=(2, 4, 8, 16, 48),
                           enc_patch_size=(1024,) * 5, shuffle_orders=True, layer_scale=1e-5, enc_mode=True)),
)


def inputs(in_channels):
    from oracle import gridsample as G
    from ptv3_b200 import synth
    cs, gs_, fs, off = [], [], [], []
    for r, scale, seed in ((3600, 0.55, 31), (1500, 0.3, 32)):
        coord, feat = synth.pig_scan(r, scale=scale, seed=seed, in_channels=in_channels)
        gs = G.grid_sample(coord, 0.02)
        iu = gs["idx_unique"]
        cs.append(coord[iu]); gs_.append(gs["grid_coord"][iu]); fs.append(feat[iu]); off.append(len(iu))
    return np.concatenate(cs), np.concatenate(gs_), np.concatenate(fs), np.cumsum(off).astype(np.int64)


def weights(tag):
    import torch
    from ptv3_b200 import synth
    with open(os.path.join(HERE, "golden", "param_shapes_sonata.json")) as f:
        shapes = json.load(f)[tag]
    return {k: torch.from_numpy(v) for k, v in synth.make_weights(shapes, seed=0).items()}


def golden():
    return np.load(os.path.join(HERE, "golden", "sonata.npz"))
