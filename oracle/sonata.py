"""Oracle: PT-v3m2 ("Sonata" PTv3: GridPooling, LayerScale, Linear stem, LayerNorm everywhere) forward, functional
torch-CPU fp32 over a flat weight dict with the reference's state_dict names.  Restates

  PointTransformerV3.forward   point_transformer_v3m2_sonata.py:720-732
  Embedding                    :507-540 (Linear -> LayerNorm -> GELU, optional mask token)
  Block                        :262-361 (= PT-v3m1's plus LayerScale on both branches; oracle/ptv3.py:_block)
  GridPooling                  :365-463 (torch.unique(dim=0) clusters in (batch, x, y, z) order, segment max / mean,
                               LayerNorm + GELU, then a FRESH Point.serialization of the pooled grid)
  GridUnpooling                :466-504 (both branches Linear -> LayerNorm -> GELU; the sparse tensor is refreshed)
  DefaultSegmentorV2.forward   default.py:63-95 (enc_mode: features of every level concatenated down to level 0)

Parity pinned on the unmodified reference: tests/golden/sonata.npz (made by tests/golden/make_golden.py sonata).
Test infrastructure only (see oracle/__init__.py).
"""
import numpy as np
import torch
import torch.nn.functional as F

from . import ops as O
from . import serialization as S
from .ptv3 import _Pt, _block, _lin, _ln, _offset_from_batch, _sparse_indices

BASE_CFG = dict(
    in_channels=6,
    order=("z", "z-trans", "hilbert", "hilbert-trans"),
    stride=(2, 2, 2, 2),
    enc_depths=(2, 2, 2, 6, 2),
    enc_channels=(32, 64, 128, 256, 512),
    enc_num_head=(2, 4, 8, 16, 32),
    enc_patch_size=(1024,) * 5,
    dec_depths=(2, 2, 2, 2),
    dec_channels=(64, 64, 128, 256),
    dec_num_head=(4, 4, 8, 16),
    dec_patch_size=(1024,) * 4,
    shuffle_orders=True,
    enc_mode=False,
)


def _serialize(pt, orders, perm):
    depth, code, order, inverse = S.serialization(pt.grid_coord.numpy(), pt.batch.numpy(), orders,
                                                  perm=None if perm is None else np.asarray(perm))
    pt.depth, pt.code = depth, torch.from_numpy(code)
    pt.order, pt.inverse = torch.from_numpy(order), torch.from_numpy(inverse)
    return pt


def _grid_pool(pt, W, p, stride, orders, perm):
    """GridPooling.forward :393-463."""
    g = torch.div(pt.grid_coord, stride, rounding_mode="trunc")
    rows = g.clone()
    rows[:, 0] |= pt.batch << 48
    uniq, cluster, counts = torch.unique(rows, sorted=True, return_inverse=True, return_counts=True, dim=0)
    grid = uniq & ((1 << 48) - 1)
    indices = torch.sort(cluster, stable=True)[1]
    idx_ptr = torch.cat([counts.new_zeros(1), torch.cumsum(counts, 0)])
    head = indices[idx_ptr[:-1]]
    feat = O.segment_csr(_lin(pt.feat, W, p + ".proj")[indices], idx_ptr, "max")
    coord = O.segment_csr(pt.coord[indices], idx_ptr, "mean")
    q = _Pt(feat=feat, coord=coord, grid_coord=grid, batch=pt.batch[head], pooling_inverse=cluster, parent=pt)
    q.offset = _offset_from_batch(q.batch)
    q.feat = F.gelu(_ln(q.feat, W, p + ".norm.0"))
    return _serialize(q, orders, perm)


def _grid_unpool(pt, W, p):
    """GridUnpooling.forward :491-504."""
    parent, inv = pt.pop("parent"), pt.pooling_inverse
    a = F.gelu(_ln(_lin(pt.feat, W, p + ".proj.0"), W, p + ".proj.1"))
    b = F.gelu(_ln(_lin(parent.feat, W, p + ".proj_skip.0"), W, p + ".proj_skip.1"))
    parent.feat = b + a[inv]
    return parent


@torch.no_grad()
def backbone_forward(W, coord, grid_coord, feat, offset, cfg=BASE_CFG, perms=None, prefix="backbone.", trace=None,
                     mask=None):
    """perms: (1 + n_pool) row permutations in the order torch.randperm draws them (the first serialization, then one
    per GridPooling), or None for shuffle_orders=False.  Returns the last point of the network: level 0 with a decoder,
    the coarsest level (with .parent / .pooling_inverse links) in enc_mode."""
    W = {k[len(prefix):]: v for k, v in W.items() if k.startswith(prefix)} if prefix else W
    orders = list(cfg["order"])
    offset = torch.as_tensor(offset, dtype=torch.int64)
    counts = torch.diff(offset, prepend=torch.zeros(1, dtype=torch.int64))
    batch = torch.repeat_interleave(torch.arange(len(offset)), counts)
    perms = list(perms) if perms is not None else None
    nxt = (lambda: perms.pop(0)) if perms is not None else (lambda: None)
    pt = _Pt(coord=torch.as_tensor(coord, dtype=torch.float32), grid_coord=torch.as_tensor(grid_coord).to(torch.int64),
             feat=torch.as_tensor(feat, dtype=torch.float32), batch=batch, offset=offset)
    # Embedding :532-540 (before the serialization in m2; it does not depend on it)
    pt.feat = F.gelu(_ln(_lin(pt.feat, W, "embedding.stem.linear"), W, "embedding.stem.norm"))
    if mask is not None:
        pt.feat = torch.where(torch.as_tensor(mask)[:, None], W["embedding.mask_token"], pt.feat)
    if trace is not None:
        trace["stem"] = pt.feat.clone()
    pt = _serialize(pt, orders, nxt())
    ns = len(cfg["enc_depths"])
    for s in range(ns):
        if s > 0:
            pt = _grid_pool(pt, W, f"enc.enc{s}.down", cfg["stride"][s - 1], orders, nxt())
            if trace is not None:
                trace[f"pool{s}"] = pt.feat.clone()
        pt.nbr = O.subm_neighbor_table(_sparse_indices(pt), 3)
        for i in range(cfg["enc_depths"][s]):
            pt = _block(pt, W, f"enc.enc{s}.block{i}", cfg["enc_num_head"][s], cfg["enc_patch_size"][s],
                        i % len(orders), pt.nbr, cfg.get("enable_flash", True))
        if trace is not None:
            trace[f"enc{s}"] = pt.feat.clone()
    if not cfg.get("enc_mode", False):
        for s in reversed(range(ns - 1)):
            pt = _grid_unpool(pt, W, f"dec.dec{s}.up")
            for i in range(cfg["dec_depths"][s]):
                pt = _block(pt, W, f"dec.dec{s}.block{i}", cfg["dec_num_head"][s], cfg["dec_patch_size"][s],
                            i % len(orders), pt.nbr, cfg.get("enable_flash", True))
            if trace is not None:
                trace[f"dec{s}"] = pt.feat.clone()
    return pt


@torch.no_grad()
def upcast_features(pt):
    """default.py:70-82: walk the pooling_parent chain, concatenating each level's features onto its parent's rows."""
    while "parent" in pt:
        parent, inv = pt.pop("parent"), pt.pop("pooling_inverse")
        parent.feat = torch.cat([parent.feat, pt.feat[inv]], dim=-1)
        pt = parent
    return pt.feat


@torch.no_grad()
def seg_head(W, feat):
    return _lin(feat, W, "seg_head")
