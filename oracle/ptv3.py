"""Oracle: PTv3 (PT-v3m1) forward + fork heads, functional torch-CPU fp32.

A restatement of the reference forward as one straight-line function over a
flat weight dict that uses the reference's state_dict names, so a reference
checkpoint (or the deterministic test weights) drives both sides:
  PointTransformerV3.forward   point_transformer_v3m1_base.py:699-714
  Embedding                    :485-516      Block            :251-338
  SerializedAttention          :51-222       MLP              :225-248
  SerializedPooling            :341-444      SerializedUnpooling :447-482
  Point.serialization/sparsify structure.py:52-146
  OffsetKeypointPTv3.forward   offset_keypoint_ptv3.py:37-107
  KeypointPTv3.forward         keypoint_ptv3.py:37-98
  DefaultSegmentorV2.forward   default.py:63-95
Eval mode by default (BatchNorm running stats, no dropout / drop-path); BN_TRAIN / offset_keypoint_grads switch to
batch statistics and run torch.autograd through the same functions for the training-step tests.  Attention
uses flash window semantics (fixed K, short scenes = one short window) by default, which
equals the reference's enable_flash=False branch when run one scene per
forward; cfg["enable_flash"] = False switches to the dense branch's batch-wide
K = min(smallest scene, patch) (:173-176), pinned on a two-scene run of the reference (dense_batch.npz).  Test infrastructure only (see oracle/__init__.py).
"""
import math

import numpy as np
import torch
import torch.nn.functional as F

from . import ops as O
from . import serialization as S

BASE_CFG = dict(
    in_channels=4,
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
)


BN_TRAIN = False     # tests of the training step set this: nn.BatchNorm1d in train() uses batch statistics


def _bn(x, W, p, eps=1e-3):
    if BN_TRAIN:
        return F.batch_norm(x, None, None, W[p + ".weight"], W[p + ".bias"], True, 0.0, eps)
    return (x - W[p + ".running_mean"]) / torch.sqrt(W[p + ".running_var"] + eps) * W[p + ".weight"] + W[p + ".bias"]


def _ln(x, W, p):
    return F.layer_norm(x, (x.shape[1],), W[p + ".weight"], W[p + ".bias"], 1e-5)


def _lin(x, W, p):
    return F.linear(x, W[p + ".weight"], W.get(p + ".bias"))


def _offset_from_batch(batch):
    return torch.cumsum(torch.bincount(batch), 0)


class _Pt(dict):
    __getattr__ = dict.__getitem__
    __setattr__ = dict.__setitem__


def _sparse_indices(pt):
    return torch.cat([pt.batch[:, None].to(torch.int32), pt.grid_coord.to(torch.int32)], 1)


def _block(pt, W, p, H, K, order_index, nbr, enable_flash=True):
    x = pt.feat
    # The conv reads point.sparse_conv_feat.features (modules.py:84-87), which
    # SerializedUnpooling leaves stale: right after an unpool it still holds the
    # proj_skip branch only (:476-477 update parent.feat but not the sparse
    # tensor).  The first decoder block of every stage therefore convolves the
    # skip branch while the shortcut is the full sum.  Reproduced on purpose.
    conv_in = pt.pop("stale_conv_in", x)
    y = O.subm_conv(conv_in, nbr, W[p + ".cpe.0.weight"], W[p + ".cpe.0.bias"])
    y = _ln(_lin(y, W, p + ".cpe.1"), W, p + ".cpe.2")
    x = x + y
    y = _ln(x, W, p + ".norm1.0")
    qkv = _lin(y, W, p + ".attn.qkv")
    if not enable_flash:      # dense branch :173-176: one patch size for the whole batch
        K = min(int(torch.diff(pt.offset, prepend=pt.offset.new_zeros(1)).min()), K)
    y = O.window_attention(qkv, pt.order[order_index], pt.inverse[order_index], pt.offset.numpy(), H, K)
    y = _lin(y, W, p + ".attn.proj")
    if p + ".ls1.0.gamma" in W:      # PT-v3m2 LayerScale (point_transformer_v3m2_sonata.py:27-39, :349)
        y = y * W[p + ".ls1.0.gamma"]
    x = x + y
    y = _ln(x, W, p + ".norm2.0")
    y = _lin(F.gelu(_lin(y, W, p + ".mlp.0.fc1")), W, p + ".mlp.0.fc2")
    if p + ".ls2.0.gamma" in W:
        y = y * W[p + ".ls2.0.gamma"]
    pt.feat = x + y
    return pt


def _pool(pt, W, p, stride, perm):
    """SerializedPooling.forward :371-444."""
    pd = (math.ceil(stride) - 1).bit_length()
    if pd > pt.depth:
        pd = 0
    code = pt.code >> (pd * 3)
    _, cluster, counts = torch.unique(code[0], sorted=True, return_inverse=True, return_counts=True)
    indices = torch.sort(cluster, stable=True)[1]
    idx_ptr = torch.cat([counts.new_zeros(1), torch.cumsum(counts, 0)])
    head = indices[idx_ptr[:-1]]
    code = code[:, head]
    order, inverse = S.argsort_rows(code.numpy())
    order, inverse = torch.from_numpy(order), torch.from_numpy(inverse)
    if perm is not None:
        code, order, inverse = code[perm], order[perm], inverse[perm]
    feat = O.segment_csr(_lin(pt.feat, W, p + ".proj")[indices], idx_ptr, "max")
    coord = O.segment_csr(pt.coord[indices], idx_ptr, "mean")
    q = _Pt(feat=feat, coord=coord, grid_coord=pt.grid_coord[head] >> pd, code=code, order=order,
            inverse=inverse, depth=pt.depth - pd, batch=pt.batch[head], pooling_inverse=cluster, parent=pt)
    q.offset = _offset_from_batch(q.batch)
    q.feat = F.gelu(_bn(q.feat, W, p + ".norm.0"))
    return q


def _unpool(pt, W, p):
    """SerializedUnpooling.forward :471-482."""
    parent, inv = pt.parent, pt.pooling_inverse
    a = F.gelu(_bn(_lin(pt.feat, W, p + ".proj.0"), W, p + ".proj.1"))
    b = F.gelu(_bn(_lin(parent.feat, W, p + ".proj_skip.0"), W, p + ".proj_skip.1"))
    parent.feat = b + a[inv]
    parent["stale_conv_in"] = b
    return parent


@torch.no_grad()
def backbone_forward(W, coord, grid_coord, feat, offset, cfg=BASE_CFG, perms=None, prefix="backbone.",
                     trace=None):
    """perms: list of (1 + n_pool) row permutations (what torch.randperm would
    have drawn: one in Point.serialization, one per SerializedPooling) or None
    for shuffle_orders=False.  Returns the stage-0 point (feat (N, dec_channels[0]))."""
    W = {k[len(prefix):]: v for k, v in W.items() if k.startswith(prefix)} if prefix else W
    orders = list(cfg["order"])
    offset = torch.as_tensor(offset, dtype=torch.int64)
    counts = torch.diff(offset, prepend=torch.zeros(1, dtype=torch.int64))
    batch = torch.repeat_interleave(torch.arange(len(offset)), counts)
    grid_coord = torch.as_tensor(grid_coord).to(torch.int64)
    perms = list(perms) if perms is not None else None
    nxt = (lambda: torch.as_tensor(perms.pop(0))) if perms is not None else (lambda: None)
    depth, code, order, inverse = S.serialization(grid_coord.numpy(), batch.numpy(), orders,
                                                  perm=None if perms is None else np.asarray(perms.pop(0)))
    pt = _Pt(coord=torch.as_tensor(coord, dtype=torch.float32), grid_coord=grid_coord,
             feat=torch.as_tensor(feat, dtype=torch.float32), batch=batch, offset=offset, depth=depth,
             code=torch.from_numpy(code), order=torch.from_numpy(order), inverse=torch.from_numpy(inverse))
    # stem (:499-511); SubM ignores padding=1 -> centred 5^3 window
    nbr5 = O.subm_neighbor_table(_sparse_indices(pt), 5)
    pt.feat = F.gelu(_bn(O.subm_conv(pt.feat, nbr5, W["embedding.stem.conv.weight"], None), W, "embedding.stem.norm"))
    if trace is not None:
        trace["stem"] = pt.feat.clone()
    ns = len(cfg["enc_depths"])
    nbrs = {}
    for s in range(ns):
        if s > 0:
            pt = _pool(pt, W, f"enc.enc{s}.down", cfg["stride"][s - 1], nxt())
            if trace is not None:
                trace[f"pool{s}"] = pt.feat.clone()
        nbrs[s] = O.subm_neighbor_table(_sparse_indices(pt), 3)
        pt.nbr = nbrs[s]
        for i in range(cfg["enc_depths"][s]):
            pt = _block(pt, W, f"enc.enc{s}.block{i}", cfg["enc_num_head"][s], cfg["enc_patch_size"][s],
                        i % len(orders), pt.nbr, cfg.get("enable_flash", True))
        if trace is not None:
            trace[f"enc{s}"] = pt.feat.clone()
    if not cfg.get("enc_mode", False):
        for s in reversed(range(ns - 1)):
            pt = _unpool(pt, W, f"dec.dec{s}.up")
            for i in range(cfg["dec_depths"][s]):
                pt = _block(pt, W, f"dec.dec{s}.block{i}", cfg["dec_num_head"][s], cfg["dec_patch_size"][s],
                            i % len(orders), pt.nbr, cfg.get("enable_flash", True))
            if trace is not None:
                trace[f"dec{s}"] = pt.feat.clone()
    return pt


@torch.no_grad()
def offset_head(W, feat, num_keypoints=6):
    """OffsetKeypointPTv3 head, eval (:44-46, :101-105)."""
    h = F.relu(_bn(_lin(feat, W, "head.0"), W, "head.1", eps=1e-5))
    pred = _lin(h, W, "head.3").view(-1, num_keypoints, 4).clone()
    pred[..., 3] = torch.sigmoid(pred[..., 3])
    return pred


@torch.no_grad()
def keypoint_head(W, feat, batch, num_keypoints=6):
    """KeypointPTv3 head, eval (keypoint_ptv3.py:44-48)."""
    g = O.scatter_mean(feat, batch)
    h = F.relu(_bn(_lin(g, W, "reg_head.0"), W, "reg_head.1", eps=1e-5))
    h = F.relu(_lin(h, W, "reg_head.4"))
    return _lin(h, W, "reg_head.6").view(-1, num_keypoints, 3)


@torch.no_grad()
def seg_head(W, feat):
    """DefaultSegmentorV2 (default.py:63-95): PTv3 with decoder has no pooling_parent left."""
    return _lin(feat, W, "seg_head")


def offset_keypoint_loss(W, feat, target, num_keypoints=6):
    """OffsetKeypointPTv3 training branch (offset_keypoint_ptv3.py:44-86): BCE-with-logits on every point's mask
    channel + 2 x masked L1 on the offsets.  Differentiable (call the un-decorated backbone_forward.__wrapped__
    under torch.enable_grad() for gradients)."""
    h = F.relu(_bn(_lin(feat, W, "head.0"), W, "head.1", eps=1e-5))
    pred = _lin(h, W, "head.3").view(-1, num_keypoints, 4)
    offset_gt, mask_gt = target[..., :3], target[..., 3]
    cls_loss = F.binary_cross_entropy_with_logits(pred[..., 3], mask_gt, reduction="none").mean()
    valid = (mask_gt > 0.5).float().unsqueeze(-1)
    reg_loss = (F.l1_loss(pred[..., :3], offset_gt, reduction="none") * valid).sum() / (valid.sum() * 3 + 1e-6)
    return cls_loss + reg_loss * 2.0


def offset_keypoint_grads(W, coord, grid_coord, feat, offset, target, perms=None, cfg=BASE_CFG, bn_train=True):
    """loss and d loss / d W for every floating weight (engines/train.py:184-245: forward in train(), loss.backward()).
    DropPath is not modelled (tests build the reference with drop_path = 0)."""
    global BN_TRAIN
    Wg = {k: (v.clone().requires_grad_(True) if v.is_floating_point() and "running_" not in k else v)
          for k, v in W.items()}
    prev, BN_TRAIN = BN_TRAIN, bool(bn_train)
    try:
        with torch.enable_grad():
            pt = backbone_forward.__wrapped__(Wg, coord, grid_coord, feat, offset, cfg=cfg, perms=perms)
            loss = offset_keypoint_loss(Wg, pt.feat, torch.as_tensor(target))
            loss.backward()
    finally:
        BN_TRAIN = prev
    return loss.detach(), {k: v.grad for k, v in Wg.items() if v.requires_grad and v.grad is not None}


def aggregate_keypoints(coord, pred, mode="argmax", scale=1.0, centroid=0.0, mask_thresh=0.5):
    """tools/infer_offset.py:366-399: kp_k = (coord[i*] + offset[i*,k])*scale +
    centroid with i* = argmax_i mask[i,k]; 'weighted' = confidence-weighted mean
    over points with mask > thresh, falling back to argmax if none."""
    votes = (coord[:, None, :] + pred[..., :3]) * scale + centroid      # (N, K, 3)
    conf = pred[..., 3]
    K = pred.shape[1]
    out = torch.zeros(K, 3, dtype=votes.dtype)
    for k in range(K):
        valid = conf[:, k] > mask_thresh
        if mode == "weighted" and bool(valid.any()):
            w = conf[valid, k] / conf[valid, k].sum()
            out[k] = (votes[valid, k] * w[:, None]).sum(0)
        else:
            out[k] = votes[conf[:, k].argmax(), k]
    return out


def offset_keypoint_eval(coord, pred, target, offset, scale=None, acc=None):
    """hooks/offset_keypoint_evaluator.py:46-91 for one batch (test infrastructure, CPU): per scene b / keypoint k the
    ground truth is the mean of coord + target[:, k, :3] over the rows with target[:, k, 3] > 0 (keypoint skipped when
    there are none), the prediction is the vote of the row with the highest predicted mask, the distance is their
    L2 norm times scale[b].  Returns dist (B,K) fp32, valid (B,K) bool and `acc`, the evaluator's running sums
    [total_dist, total_samples, dist_per_kp[K], samples_per_kp[K]] as Python floats, updated like lines 83-91."""
    coord, pred, target = torch.as_tensor(coord), torch.as_tensor(pred), torch.as_tensor(target)
    offset = [int(o) for o in offset]
    K = pred.shape[1]
    acc = [0.0] * (2 + 2 * K) if acc is None else list(acc)
    dist = torch.zeros(len(offset), K, dtype=torch.float32)
    valid = torch.zeros(len(offset), K, dtype=torch.bool)
    for b, end in enumerate(offset):
        start = offset[b - 1] if b else 0
        c, p, t = coord[start:end], pred[start:end], target[start:end]
        s = float(scale[b]) if scale is not None else 1.0
        b_sum, b_n = 0.0, 0
        for k in range(K):
            idx = torch.nonzero(t[:, k, 3] > 0).squeeze(-1)
            if len(idx) == 0:
                continue
            gt = (c[idx] + t[idx, k, :3]).mean(dim=0)
            best = torch.argmax(p[:, k, 3])
            d = torch.norm(c[best] + p[best, k, :3] - gt, p=2) * s
            dist[b, k], valid[b, k] = d, True
            acc[2 + k] += d.item()
            acc[2 + K + k] += 1
            b_sum += d.item()
            b_n += 1
        if b_n > 0:
            acc[0] += b_sum / b_n
            acc[1] += 1
    return dist, valid, acc


def offset_keypoint_eval_summary(acc, num_keypoints):
    """mean_dist and the per-keypoint means the evaluator logs (offset_keypoint_evaluator.py:105-109)."""
    K = num_keypoints
    return acc[0] / (acc[1] + 1e-6), [acc[2 + k] / (acc[2 + K + k] + 1e-6) for k in range(K)]
