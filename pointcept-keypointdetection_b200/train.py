"""Training-mode forward of the same module tree (backward + AMP-style mixed precision + DDP).

`model.py` is the fused inference path; this file walks the SAME nn.Modules (same parameters / state_dict) with
autograd enabled: the custom kernels enter the graph through `autograd.py`, the dense glue (Linear = cuBLAS,
LayerNorm, BatchNorm in batch-statistics mode, GELU) is torch, exactly the split the reference has between its
third-party CUDA ops and torch.  Semantics follow the reference in train() mode:

  * BatchNorm1d uses batch statistics and updates the running ones (:505 `momentum=0.01, eps=1e-3`);
  * DropPath (timm) drops whole rows of the residual branch with the per-block rate of :603-661;
  * `compute_dtype` plays the role of torch.autocast (engines/train.py:184): GEMM / conv / attention operands are
    cast to it, the residual stream, the norms, the head and the loss stay fp32;
  * the loss of OffsetKeypointPTv3 is offset_keypoint_ptv3.py:50-86, of KeypointPTv3 keypoint_ptv3.py:52-63, of
    DefaultSegmentorV2 the caller's criterion on the logits (default.py:86-88).
"""
import math

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import autograd as A
from . import ops
from .structure import Point, SparseConvTensor, z_info


def _lin(x, lin, cd):
    return A.linear(x, lin, cd)


def _ln(x, ln, out_dtype=torch.float32):
    return A.layer_norm(x, ln, out_dtype)


def _bn(x, bn):
    from . import ddp
    if ddp._sync_active(bn):        # cfg.sync_bn (engines/train.py:256-257): statistics over every rank's rows
        return ddp.sync_batch_norm(x.float(), bn)
    return bn(x.float())            # nn.BatchNorm1d: batch statistics in train(), running ones in eval()


def _residual(x, y, rate, training):
    """x + DropPath(y) (timm.layers.DropPath: one Bernoulli draw per row, kept rows scaled by 1 / keep) in ONE
    elementwise kernel: x fp32, y in the compute dtype, promoted inside the kernel instead of y.float() first."""
    if rate == 0.0 or not training:
        return x + y
    keep = 1.0 - rate
    mask = x.new_empty((x.shape[0], 1)).bernoulli_(keep).div_(keep)
    return torch.addcmul(x, y, mask)


def embedding_forward(mod, point):
    conv = mod.stem.conv
    sp = point.sparse_conv_feat
    y = A.subm_conv(point.feat.float(), conv.weight, conv.bias, conv.neighbor_map(sp), torch.float32)
    if hasattr(mod.stem, "norm"):
        y = _bn(y, mod.stem.norm)
    if hasattr(mod.stem, "act"):
        y = mod.stem.act(y)
    point.feat = y
    point.sparse_conv_feat = sp.replace_feature(y)
    return point


def block_forward(mod, point):
    cd = mod.compute_dtype
    sp = point.sparse_conv_feat
    x = point.feat.float()
    conv = mod.cpe[0]
    y = A.subm_conv(sp.features, conv.weight, conv.bias, conv.neighbor_map(sp), cd)
    y = _ln(_lin(y, mod.cpe[1], cd), mod.cpe[2])
    x = x + y
    att = mod.attn
    qkv = _lin(_ln(x, mod.norm1[0], cd), att.qkv, cd)
    windows, nwin = att.get_windows(point)
    a = A.attn_window(qkv, point.serialized_order[att.order_index], windows, nwin, att.num_heads, att.patch_size,
                      att.scale, att.algo)
    g1, g2 = mod._gammas()                    # PT-v3m2 LayerScale (None for PT-v3m1)
    p = _lin(a, att.proj, cd)
    x = _residual(x, p if g1 is None else p * g1, mod.drop_path_rate, mod.training)
    mlp = mod.mlp[0]
    f = _lin(mlp.act(_lin(_ln(x, mod.norm2[0], cd), mlp.fc1, cd)), mlp.fc2, cd)
    x = _residual(x, f if g2 is None else f * g2, mod.drop_path_rate, mod.training)
    point.feat = x
    point["_feat_cd"] = None
    point.sparse_conv_feat = sp.replace_feature(x)
    return point


def pooling_forward(mod, point):
    cd = mod.compute_dtype
    depth = mod.pooling_depth(point.serialized_depth)
    n_scenes = len(point.offset)
    with torch.no_grad():
        plans = point.pop("_pool_plans", None)
        plan = plans.pop(0) if plans else mod.plan(point.serialized_code, point.serialized_order, point.grid_coord,
                                                   point.batch, depth, n_scenes, point.get("_z_row"))
        pf, grid_coord, batch, indices4 = plan["pf"], plan["grid_coord"], plan["batch"], plan["indices4"]
        code, order, inverse, offset = plan["code"], plan["order"], plan["inverse"], plan["offset"]
        coord = ops.segment_csr(point.coord.float(), pf["idx_ptr"], "mean", gather=pf["indices"])
    feat = A.segment_csr(_lin(point.feat, mod.proj, cd), pf["idx_ptr"], mod.reduce, pf["indices"]).float()
    if hasattr(mod, "norm"):
        feat = _bn(feat, mod.norm[0])
    if hasattr(mod, "act"):
        feat = mod.act[0](feat)
    new = Point(feat=feat, coord=coord, grid_coord=grid_coord, serialized_code=code, serialized_order=order,
                serialized_inverse=inverse, serialized_depth=point.serialized_depth - depth, batch=batch, offset=offset)
    for k in ("condition", "context", "grid_size", "order"):
        if k in point:
            new[k] = point[k]
    if plans:
        new["_pool_plans"] = plans
    new["pooling_inverse"] = pf["cluster"]
    new["pooling_parent"] = point
    new["_pooling_csr"] = (pf["indices"], pf["idx_ptr"])
    new["sparse_shape"] = [(1 << max(int(new.serialized_depth), 0)) + 96] * 3
    new["sparse_conv_feat"] = SparseConvTensor(feat, indices4, new["sparse_shape"], n_scenes)
    new["_z_row"] = plan.get("z_row")
    z = z_info(new)
    if z is not None:
        new["sparse_conv_feat"].indice_dict["_z"] = z
    return new


def _proj_seq(seq, x, cd):
    for m in seq:
        if isinstance(m, nn.Linear):
            x = _lin(x, m, cd)
        elif isinstance(m, nn.BatchNorm1d):
            x = _bn(x, m)
        else:
            x = m(x)
    return x


def unpooling_forward(mod, point):
    cd = mod.compute_dtype
    parent = point.pop("pooling_parent")
    inverse = point.pop("pooling_inverse")
    indices, idx_ptr = point.pop("_pooling_csr")
    up = _proj_seq(mod.proj, point.feat, cd).float()
    skip = _proj_seq(mod.proj_skip, parent.feat, cd).float()
    parent.feat = skip + A.gather_rows(up, inverse, indices, idx_ptr)
    parent["_feat_cd"] = None
    # the parent's sparse tensor keeps the proj_skip branch only (reference :476-477, see model.SerializedUnpooling)
    parent.sparse_conv_feat = parent.sparse_conv_feat.replace_feature(skip)
    if mod.traceable:
        parent["unpooling_parent"] = point
    return parent


def _seq_tail(x, mods):
    """norm / activation modules after a Linear, on fp32 rows (torch: LayerNorm / BatchNorm / GELU)."""
    for m in mods:
        x = _bn(x, m) if isinstance(m, nn.BatchNorm1d) else m(x.float())
    return x


def embedding_forward_m2(mod, point):
    """point_transformer_v3m2_sonata.py:532-540."""
    mods = list(mod.stem._modules.values())
    y = _seq_tail(F.linear(point.feat.float(), mods[0].weight, mods[0].bias), mods[1:])
    if "mask" in point:
        y = torch.where(point.mask.unsqueeze(-1), mod.mask_token.to(y.dtype), y)
    point.feat = y
    point["_feat_cd"] = None
    if "sparse_conv_feat" in point:
        point.sparse_conv_feat = point.sparse_conv_feat.replace_feature(y)
    return point


def grid_pooling_forward(mod, point):
    """GridPooling :393-463 with autograd through the projection and the segment reduction."""
    cd = mod.compute_dtype
    n_scenes = len(point.offset)
    with torch.no_grad():
        plans = point.pop("_pool_plans", None)
        plan = plans.pop(0) if plans else mod.plan(point.grid_coord, point.batch, mod._depth_of(point), n_scenes,
                                                   point.order)
        coord = ops.segment_csr(point.coord.float(), plan["idx_ptr"], "mean", gather=plan["indices"])
    feat = A.segment_csr(_lin(point.feat, mod.proj, cd), plan["idx_ptr"], mod.reduce, plan["indices"]).float()
    if hasattr(mod, "norm"):
        feat = _seq_tail(feat, [mod.norm[0]])
    if hasattr(mod, "act"):
        feat = mod.act[0](feat)
    new = Point(feat=feat, coord=coord, grid_coord=plan["grid_coord"], batch=plan["batch"], offset=plan["offset"],
                serialized_depth=plan["depth"], serialized_code=plan["code"], serialized_order=plan["order"],
                serialized_inverse=plan["inverse"], order=point.order)
    for k in ("condition", "context", "name", "split"):
        if k in point:
            new[k] = point[k]
    if "grid_size" in point:
        new["grid_size"] = point.grid_size * mod.stride
    if plans:
        new["_pool_plans"] = plans
    new["pooling_inverse"] = plan["cluster"]
    new["pooling_parent"] = point
    new["idx_ptr"] = plan["idx_ptr"]
    new["_pooling_csr"] = (plan["indices"], plan["idx_ptr"])
    new["sparse_shape"] = [(1 << max(int(plan["depth"]), 0)) + 96] * 3
    new["sparse_conv_feat"] = SparseConvTensor(feat, plan["indices4"], new["sparse_shape"], n_scenes)
    new["_z_row"] = plan.get("z_row")
    z = z_info(new)
    if z is not None:
        new["sparse_conv_feat"].indice_dict["_z"] = z
    return new


def grid_unpooling_forward(mod, point):
    """GridUnpooling :491-504: parent.feat = proj_skip(parent) + proj(point)[inverse]; the sparse tensor is refreshed."""
    cd = mod.compute_dtype
    parent = point.pop("pooling_parent")
    inverse = point.pooling_inverse
    indices, idx_ptr = point["_pooling_csr"]
    feat = point.feat

    def branch(seq, x):
        mods = list(seq._modules.values())
        return _seq_tail(_lin(x, mods[0], cd), mods[1:]).float()
    up, skip = branch(mod.proj, point.feat), branch(mod.proj_skip, parent.feat)
    parent.feat = skip + A.gather_rows(up, inverse, indices, idx_ptr)
    parent["_feat_cd"] = None
    parent.sparse_conv_feat = parent.sparse_conv_feat.replace_feature(parent.feat)
    if mod.traceable:
        point.feat = feat
        parent["unpooling_parent"] = point
    return parent


def backbone_forward_m2(bb, data_dict):
    """PT-v3m2 forward in train() mode (point_transformer_v3m2_sonata.py:720-732).  A backbone without any trainable
    parameter (DefaultSegmentorV2(freeze_backbone=True), the "lin" configs) runs without building a graph; DropPath
    stays active like in the reference, which leaves the frozen backbone in train() mode."""
    from . import model as M
    from . import sonata as S2
    point = Point(data_dict)
    if not hasattr(bb, "_param_cast"):
        bb._param_cast = A.ParamCast()
    A.set_param_cast(bb._param_cast.refresh(bb, bb.compute_dtype))
    frozen = not any(p.requires_grad for p in bb.parameters()) and not point.feat.requires_grad
    with torch.no_grad():
        point = bb.plan_levels(point)
        if frozen and bb.fused_when_frozen:
            # no graph needed and no BatchNorm anywhere in PT-v3m2: the fused inference modules ARE the train() forward,
            # with DropPath applied inside the residual kernels (model.Block.forward)
            point = bb.enc(bb.embedding(point))
            return point if bb.enc_mode else bb.dec(point)
    with torch.set_grad_enabled(not frozen and torch.is_grad_enabled()):
        point = embedding_forward_m2(bb.embedding, point)
        for seq in [bb.enc] + ([] if bb.enc_mode else [bb.dec]):
            for stage in seq._modules.values():
                for m in stage._modules.values():
                    if isinstance(m, M.Block):
                        point = block_forward(m, point)
                    elif isinstance(m, S2.GridPooling):
                        point = grid_pooling_forward(m, point)
                    elif isinstance(m, S2.GridUnpooling):
                        point = grid_unpooling_forward(m, point)
                    else:
                        raise NotImplementedError(type(m))
    return point


def backbone_forward(bb, data_dict):
    from . import model as M
    point = Point(data_dict)
    if not hasattr(bb, "_param_cast"):
        bb._param_cast = A.ParamCast()
    A.set_param_cast(bb._param_cast.refresh(bb, bb.compute_dtype))
    with torch.no_grad():
        point.serialization(order=bb.order, shuffle_orders=bb.shuffle_orders)
        point.sparsify()
        if bb.plan_ahead:        # integer front ends of all pooling levels first (see model.SerializedPooling.plan)
            plans, code, order = [], point.serialized_code, point.serialized_order
            grid, batch, depth = point.grid_coord, point.batch, point.serialized_depth
            z_row = point.get("_z_row")
            for stage in bb.enc._modules.values():
                down = stage._modules.get("down")
                if down is None:
                    continue
                pd = down.pooling_depth(depth)
                plans.append(down.plan(code, order, grid, batch, pd, len(point.offset), z_row))
                code, order, grid, batch = plans[-1]["code"], plans[-1]["order"], plans[-1]["grid_coord"], plans[-1]["batch"]
                z_row = plans[-1]["z_row"]
                depth -= pd
            point["_pool_plans"] = plans
    point = embedding_forward(bb.embedding, point)
    stages = [bb.enc] + ([] if bb.enc_mode else [bb.dec])
    for seq in stages:
        for stage in seq._modules.values():
            for m in stage._modules.values():
                if isinstance(m, M.Block):
                    point = block_forward(m, point)
                elif isinstance(m, M.SerializedPooling):
                    point = pooling_forward(m, point)
                elif isinstance(m, M.SerializedUnpooling):
                    point = unpooling_forward(m, point)
                else:
                    raise NotImplementedError(type(m))
    return point


def offset_keypoint_loss(pred, target):
    """offset_keypoint_ptv3.py:50-86 on (N, K, 4) LOGITS: BCE-with-logits on the mask channel of every point + 2 x L1 on
    the offsets of the points whose target mask is set.  Returns (loss, cls_loss, reg_loss)."""
    offset_gt, mask_gt = target[..., :3], target[..., 3]
    cls_loss = F.binary_cross_entropy_with_logits(pred[..., 3], mask_gt, reduction="none").mean()
    valid = (mask_gt > 0.5).float().unsqueeze(-1)
    reg_loss = (F.l1_loss(pred[..., :3], offset_gt, reduction="none") * valid).sum() / (valid.sum() * 3 + 1e-6)
    return cls_loss + reg_loss * 2.0, cls_loss, reg_loss


def offset_keypoint_forward(model, data_dict):
    """offset_keypoint_ptv3.py:37-107: per-point (dx,dy,dz,mask logit); BCE on every point + L1 on the points whose
    target mask is set (weight 2)."""
    point = backbone_forward(model.backbone, data_dict)
    pred = model.head(point.feat.float()).view(-1, model.num_keypoints, 4)
    out = {}
    if "target" in data_dict:
        out["loss"], cls_loss, reg_loss = offset_keypoint_loss(pred, data_dict["target"])
        out["train/cls_loss"] = cls_loss.detach()
        out["train/reg_loss"] = reg_loss.detach()
    if not model.training:
        final = pred.clone()
        final[..., 3] = torch.sigmoid(pred[..., 3])
        out["pred"] = final
    else:
        out["pred_logits"] = pred
    return out


def keypoint_forward(model, data_dict):
    """keypoint_ptv3.py:37-98: per-scene mean pool -> regression MLP; MSE on (B,K,3) targets."""
    point = backbone_forward(model.backbone, data_dict)
    feat = point.feat.float()                                        # torch_scatter.scatter_mean over a few long scenes
    cnt = torch.diff(F.pad(point.offset, (1, 0))).clamp_min(1).to(feat.dtype)
    g = feat.new_zeros((len(point.offset), feat.shape[1])).index_add_(0, point.batch, feat) / cnt[:, None]
    pred = model.reg_head(g).view(-1, model.num_keypoints, 3)
    out = {}
    if "target" in data_dict:
        target = data_dict["target"]
        out["loss"] = F.mse_loss(pred if pred.shape == target.shape else pred.view(-1, 3), target)
    if not model.training:
        out["pred"] = pred
    else:
        out["pred_train"] = pred
    return out


def segmentor_forward(model, input_dict, criterion=None):
    """default.py:63-95 with a decoder backbone: logits = seg_head(feat); loss = criterion(logits, segment)."""
    point = model.backbone(Point(input_dict))           # train() mode: backbone_forward / backbone_forward_m2
    if "pooling_parent" in point and isinstance(model.seg_head, nn.Linear):
        logits, _ = model.upcast_logits(point)          # enc_mode: default.py:70-84 without the (N, sum C) concatenation
    else:
        logits = model.seg_head(point.feat.float())
    out = dict(seg_logits=logits)
    if "segment" in input_dict:
        crit = criterion or getattr(model, "criteria", None) or nn.CrossEntropyLoss(ignore_index=-1)
        out["loss"] = crit(logits, input_dict["segment"])
    return out


# ---------------------------------------------------------------------------------------------------------------
# The trainer's step (engines/train.py:184-245 `Trainer.run_step`, :346-354 `build_scaler`): gradient accumulation,
# GradScaler when AMP is on, gradient clipping after unscale_, and the scheduler advanced only when the scaler did not
# skip the optimizer step.  `torch.autocast` has no role here: mixed precision is the model's `compute_dtype` (bf16
# operands, fp32 residual stream / norms / loss), so `amp_dtype` only selects it.  The kernels compute in bf16 or fp32;
# an fp16 request is refused instead of silently computing in another type.  The reference builds its GradScaler for
# bf16 as well (enable_amp alone decides), so it is kept: with bf16 it never finds an overflow and is a pass-through.
# ---------------------------------------------------------------------------------------------------------------
class TrainStep:
    def __init__(self, model, optimizer, scheduler=None, enable_amp=False, amp_dtype="bfloat16", clip_grad=None,
                 gradient_accumulation_steps=1, grad_sync=None, scaler=None):
        if enable_amp and amp_dtype not in ("bfloat16", torch.bfloat16):
            raise NotImplementedError("ptv3_b200 kernels compute in bf16 (or fp32); amp_dtype='float16' is not built -- "
                                      "bf16 has fp32's exponent range and needs no loss scaling")
        self.model, self.optimizer, self.scheduler = model, optimizer, scheduler
        self.enable_amp, self.clip_grad = bool(enable_amp), clip_grad
        self.accum = int(gradient_accumulation_steps)
        self.sync = grad_sync                      # ddp.GradSync or None (one process)
        if scaler is None and self.enable_amp:     # build_scaler (:346-354)
            dev = next(model.parameters()).device.type
            scaler = torch.amp.GradScaler(device=dev)
        self.scaler = scaler if self.enable_amp else None
        self._count = 0
        backbone = getattr(model, "backbone", model)
        if hasattr(backbone, "set_compute_dtype"):
            backbone.set_compute_dtype(torch.bfloat16 if self.enable_amp else torch.float32)

    def _zero(self):
        if self.sync is not None:
            self.sync.zero_grad()
        else:
            self.optimizer.zero_grad()

    def __call__(self, input_dict):
        """One run_step: returns the model's output dict; `stepped` says whether the optimizer advanced."""
        if self._count == 0:
            self._zero()
        out = self.model(input_dict)
        loss = out["loss"] / self.accum
        if self.scaler is not None:
            self.scaler.scale(loss).backward()
        else:
            loss.backward()
        self._count += 1
        self.stepped = False
        if self._count >= self.accum:
            if self.sync is not None:
                self.sync.finish()                 # the averaged gradients are in place before unscale / clip
            params = list(self.model.parameters())          # clip_grad_norm_(self.model.parameters(), ...) (:220, :234)
            if self.scaler is not None:
                self.scaler.unscale_(self.optimizer)
                if self.clip_grad is not None:
                    torch.nn.utils.clip_grad_norm_(params, self.clip_grad)
                self.scaler.step(self.optimizer)
                scale = self.scaler.get_scale()
                self.scaler.update()
                self.stepped = scale <= self.scaler.get_scale()      # skipped steps shrink the scale (:226-231)
                if self.stepped and self.scheduler is not None:
                    self.scheduler.step()
            else:
                if self.clip_grad is not None:
                    torch.nn.utils.clip_grad_norm_(params, self.clip_grad)
                self.optimizer.step()
                self.stepped = True
                if self.scheduler is not None:
                    self.scheduler.step()
            self._count = 0
        return out
