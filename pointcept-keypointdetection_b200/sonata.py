"""PT-v3m2 ("Sonata" / Concerto PTv3) over the sm_100a kernels: the reference's second PTv3 variant
(pointcept/models/point_transformer_v3/point_transformer_v3m2_sonata.py), registered there as "PT-v3m2" and used by
configs/concerto/semseg-ptv3-large-v1m1-waymo-6a-lin.py (BASELINE configs[4]).

Same module tree, constructor arguments and state_dict names as the reference.  What differs from PT-v3m1 (model.py):

  * Embedding is Linear -> LayerNorm -> GELU (+ optional mask token), :507-540, and runs BEFORE the serialization;
  * Block carries LayerScale on both residual branches (:304-329); in inference gamma is folded into attn.proj / mlp.fc2
    (model.Block._lin), so the Block runs the same fused kernels / the same one-call native path as PT-v3m1;
  * GridPooling (:365-463) clusters by torch.unique(dim=0) over (batch, x/stride, y/stride, z/stride) -- pooled rows
    come out in ascending (batch, x, y, z) order, not in the order of the first serialized code -- and RE-SERIALIZES
    the pooled grid from scratch (new depth, new codes, new randperm draw).  Here: one radix sort + run-length pass
    (ops.grid_pool, ONE host read per level for the row count and the next depth) and the usual encode + argsort;
  * GridUnpooling (:466-504): both branches Linear -> LayerNorm -> GELU; the parent's sparse tensor is refreshed with
    the sum (no stale-feature quirk as in PT-v3m1's SerializedUnpooling);
  * every norm is LayerNorm: nothing to fold, LN runs in ops.ln_residual, GELU + the bf16 copy in ops.affine_act.

All integer planning (clusters, pooled grids, serializations of every level) depends on grid_coord only and runs before
the first feature kernel, in the reference's order of torch.randperm draws.
Training: train() dispatches to train.py (autograd through the kernels), like PT-v3m1.
"""
import torch
import torch.nn as nn
import torch.nn.functional as F

from . import ops
from .model import (Block, LayerScale, PointModule, PointSequential, SubMConv3d, _cached,  # noqa: F401
                    _feat_cd)
from .structure import Point, SparseConvTensor, device_perm, z_info


def _linear_f32_out(x, w, b):
    """x W^T + b with bf16 operands and an fp32 RESULT (cuBLAS writes the fp32 accumulator, no bf16 rounding of the
    output): for the Linear layers whose output feeds a LayerNorm directly (GridPooling after the segment max,
    GridUnpooling), where LN's mean subtraction would amplify the output rounding."""
    if x.dtype != torch.bfloat16:
        return F.linear(x, w, b)
    return torch.addmm(b.float(), x, w.t(), out_dtype=torch.float32)


def _ln_gelu(x, ln, cd):
    """LayerNorm -> exact GELU; returns (fp32, compute-dtype copy or None)."""
    y, _ = ops.ln_residual(x, ln.weight, ln.bias, ln.eps)
    return ops.affine_act(y, None, None, "gelu", want_f32=True, want_bf16=cd == torch.bfloat16)


@torch.no_grad()
def serialize_level(grid_coord, batch, n_scenes, orders, shuffle, grid_max=None):
    """Point.serialization (structure.py:52-109) of one level as plain tensors: (depth, code, order, inverse, row of
    the plain Morton codes or None).
    `grid_max`: per-axis maximum already on the host (saves the sync of :73)."""
    if grid_max is None:
        grid_max = [int(v) for v in ops.grid_max(grid_coord).tolist()]
    depth = int(max(grid_max) + 1).bit_length()
    assert depth * 3 + int(n_scenes).bit_length() <= 63 and depth <= 16
    code = ops.encode_orders(grid_coord, batch, depth, orders)
    key_bits = depth * 3 + max(1, (int(n_scenes) - 1).bit_length())
    order, inverse = ops.argsort_rows(code, key_bits=key_bits)
    z_row = list(orders).index("z") if "z" in orders else None
    if shuffle:
        perm = torch.randperm(code.shape[0])
        if z_row is not None:
            z_row = perm.tolist().index(z_row)
        if not torch.equal(perm, torch.arange(code.shape[0])):
            perm = device_perm(perm, code.device)
            code, order, inverse = code[perm], order[perm], inverse[perm]
    return depth, code, order, inverse, z_row


class GridPooling(PointModule):
    """:365-463."""

    def __init__(self, in_channels, out_channels, stride=2, norm_layer=None, act_layer=None, reduce="max",
                 shuffle_orders=True, traceable=True):
        super().__init__()
        assert reduce in ["sum", "mean", "min", "max"]
        self.in_channels, self.out_channels, self.stride = in_channels, out_channels, stride
        self.reduce, self.shuffle_orders, self.traceable = reduce, shuffle_orders, traceable
        self.proj = nn.Linear(in_channels, out_channels)
        if norm_layer is not None:
            self.norm = PointSequential(norm_layer(out_channels))
        if act_layer is not None:
            self.act = PointSequential(act_layer())
        self.compute_dtype = torch.bfloat16

    @torch.no_grad()
    def plan(self, grid_coord, batch, depth, n_scenes, orders):
        """The integer half (:393-417 + the re-serialization of :460): clusters, CSR, pooled grid / batch / spconv
        indices, and the pooled level's codes / orders.  Depends on the grid only."""
        gp = ops.grid_pool(grid_coord, batch, self.stride, depth, n_scenes)
        d, code, order, inverse, z_row = serialize_level(gp["grid_coord"], gp["batch"], n_scenes, orders,
                                                         self.shuffle_orders, grid_max=gp["grid_max"])
        gp.update(depth=d, code=code, order=order, inverse=inverse, offset=ops.batch_to_offset(gp["batch"], n_scenes),
                  z_row=z_row)
        return gp

    @torch.no_grad()
    def forward(self, point):
        cd = self.compute_dtype
        if "grid_coord" not in point:
            assert {"coord", "grid_size"}.issubset(point.keys()), \
                "[gird_coord] or [coord, grid_size] should be include in the Point"
            point["grid_coord"] = torch.div(point.coord - point.coord.min(0)[0], point.grid_size,
                                            rounding_mode="trunc").int()
        n_scenes = len(point.offset)
        plans = point.pop("_pool_plans", None)
        plan = plans.pop(0) if plans else self.plan(point.grid_coord, point.batch, self._depth_of(point), n_scenes,
                                                    point.order)
        if cd == torch.bfloat16:
            # The pooled rows go max -> LayerNorm: LN removes the (large, max-biased) common mode, which amplifies any
            # operand rounding of this GEMM.  It therefore reads the fp32 residual stream with TF32 products (10-bit
            # mantissa, fp32 accumulate and output) instead of the bf16 copy: 4 launches per forward, N x C_in x 2 more bytes
            prev = torch.backends.cuda.matmul.allow_tf32
            torch.backends.cuda.matmul.allow_tf32 = True
            try:
                y = F.linear(point.feat.float(), self.proj.weight.float(), self.proj.bias.float())
            finally:
                torch.backends.cuda.matmul.allow_tf32 = prev
        else:
            y = F.linear(point.feat.float(), self.proj.weight.float(), self.proj.bias.float())
        pooled = ops.segment_csr(y, plan["idx_ptr"], self.reduce, gather=plan["indices"])
        coord = ops.segment_csr(point.coord.float(), plan["idx_ptr"], "mean", gather=plan["indices"])
        norm = self.norm[0] if hasattr(self, "norm") else None
        act = self.act[0] if hasattr(self, "act") else None
        feat_cd = None
        if isinstance(norm, nn.LayerNorm) and isinstance(act, nn.GELU) and act.approximate == "none" \
                and pooled.shape[1] % 4 == 0:
            feat, feat_cd = _ln_gelu(pooled, norm, cd)
        else:
            feat = pooled.float()
            if norm is not None:
                feat = norm(feat)
            if act is not None:
                feat = act(feat)
        new = Point(feat=feat, coord=coord, grid_coord=plan["grid_coord"], batch=plan["batch"], offset=plan["offset"],
                    serialized_depth=plan["depth"], serialized_code=plan["code"], serialized_order=plan["order"],
                    serialized_inverse=plan["inverse"], order=point.order)
        for k in ("condition", "context", "name", "split"):
            if k in point:
                new[k] = point[k]
        for k in ("origin_coord", "color"):
            if k in point:
                new[k] = ops.segment_csr(point[k].float(), plan["idx_ptr"], "mean", gather=plan["indices"])
        if "grid_size" in point:
            new["grid_size"] = point.grid_size * self.stride
        if self.traceable:
            new["pooling_inverse"] = plan["cluster"]
            new["pooling_parent"] = point
            new["idx_ptr"] = plan["idx_ptr"]
        new["_pooling_csr"] = (plan["indices"], plan["idx_ptr"])
        if plans:
            new["_pool_plans"] = plans
        new["_feat_cd"] = feat_cd if feat_cd is not None else (feat if cd == torch.float32 else feat.to(cd))
        new["sparse_shape"] = [(1 << max(int(plan["depth"]), 0)) + 96] * 3
        new["sparse_conv_feat"] = SparseConvTensor(new["_feat_cd"], plan["indices4"], new["sparse_shape"], n_scenes)
        new["_z_row"] = plan.get("z_row")
        z = z_info(new)
        if z is not None:
            new["sparse_conv_feat"].indice_dict["_z"] = z
        return new

    @staticmethod
    def _depth_of(point):
        if "serialized_depth" in point:
            return int(point.serialized_depth)
        return int(int(ops.grid_max(point.grid_coord).max().item()) + 1).bit_length()


class GridUnpooling(PointModule):
    """:466-504."""

    def __init__(self, in_channels, skip_channels, out_channels, norm_layer=None, act_layer=None, traceable=False):
        super().__init__()
        self.proj = PointSequential(nn.Linear(in_channels, out_channels))
        self.proj_skip = PointSequential(nn.Linear(skip_channels, out_channels))
        if norm_layer is not None:
            self.proj.add(norm_layer(out_channels))
            self.proj_skip.add(norm_layer(out_channels))
        if act_layer is not None:
            self.proj.add(act_layer())
            self.proj_skip.add(act_layer())
        self.traceable = traceable
        self.compute_dtype = torch.bfloat16

    def _branch(self, seq, x, cd):
        """Linear (library GEMM) -> LayerNorm (kernel); returns (tensor, gelu_pending)."""
        mods = list(seq._modules.values())
        lin = mods[0]
        w, b = _cached(self, (id(lin), cd), [lin.weight, lin.bias], lambda: (lin.weight.to(cd), lin.bias.to(cd)))
        rest = mods[1:]
        y = _linear_f32_out(x, w, b) if rest and isinstance(rest[0], nn.LayerNorm) else F.linear(x, w, b)
        if rest and isinstance(rest[0], nn.LayerNorm):
            ln = rest.pop(0)
            _, y = ops.ln_residual(y, ln.weight, ln.bias, ln.eps, want_res=False, out2_dtype=cd)
        gelu = bool(rest) and isinstance(rest[0], nn.GELU) and rest[0].approximate == "none"
        if gelu:
            rest.pop(0)
        for m in rest:          # anything else: torch
            y = m(y.float())
        return y, gelu

    @torch.no_grad()
    def forward(self, point):
        assert "pooling_parent" in point and "pooling_inverse" in point
        cd = self.compute_dtype
        parent = point.pop("pooling_parent")
        inverse = point.pooling_inverse
        feat = point.feat
        up, g1 = self._branch(self.proj, _feat_cd(point, cd), cd)
        skip, g2 = self._branch(self.proj_skip, _feat_cd(parent, cd), cd)
        if g1 != g2:
            raise NotImplementedError("GridUnpooling: both branches end in the same activation in every config")
        # GELU, GELU, gather-add and the compute-dtype copy of the sum: one kernel
        out, out_cd = ops.unpool_add(skip.to(cd), up.to(cd), inverse, gelu=g1, want_skip_act=cd != torch.float32,
                                     second="sum")
        parent.feat = out
        parent["_feat_cd"] = out if out_cd is None else out_cd
        parent.sparse_conv_feat = parent.sparse_conv_feat.replace_feature(parent["_feat_cd"])        # :499
        if self.traceable:
            point.feat = feat
            parent["unpooling_parent"] = point
        return parent


class Embedding(PointModule):
    """:507-540.  state_dict: stem.linear.*, stem.norm.*, mask_token."""

    def __init__(self, in_channels, embed_channels, norm_layer=None, act_layer=None, mask_token=False):
        super().__init__()
        self.in_channels, self.embed_channels = in_channels, embed_channels
        self.stem = PointSequential(linear=nn.Linear(in_channels, embed_channels))
        if norm_layer is not None:
            self.stem.add(norm_layer(embed_channels), name="norm")
        if act_layer is not None:
            self.stem.add(act_layer(), name="act")
        self.mask_token = nn.Parameter(torch.zeros(1, embed_channels)) if mask_token else None
        self.compute_dtype = torch.bfloat16

    @torch.no_grad()
    def forward(self, point):
        cd = self.compute_dtype
        lin = self.stem.linear
        norm, act = getattr(self.stem, "norm", None), getattr(self.stem, "act", None)
        y = F.linear(point.feat.float(), lin.weight.float(), lin.bias.float())      # K = in_channels (6 / 9): fp32 GEMM
        ycd = None
        masked = "mask" in point
        if isinstance(norm, nn.LayerNorm) and isinstance(act, nn.GELU) and act.approximate == "none" \
                and y.shape[1] % 4 == 0:
            y, ycd = _ln_gelu(y, norm, torch.float32 if masked else cd)
        else:
            if norm is not None:
                y = norm(y)
            if act is not None:
                y = act(y)
        if masked:
            y = torch.where(point.mask.unsqueeze(-1), self.mask_token.to(y.dtype), y)
            ycd = None
        point.feat = y
        point["_feat_cd"] = ycd if ycd is not None else (y if cd == torch.float32 else y.to(cd))
        if "sparse_conv_feat" in point:
            point.sparse_conv_feat = point.sparse_conv_feat.replace_feature(point["_feat_cd"])
        return point


class PointTransformerV3(PointModule):
    """:543-732 ("PT-v3m2").  Same constructor and state_dict as the reference; `compute_dtype` as in model.py."""

    def __init__(self, in_channels=6, order=("z", "z-trans"), stride=(2, 2, 2, 2), enc_depths=(2, 2, 2, 6, 2),
                 enc_channels=(32, 64, 128, 256, 512), enc_num_head=(2, 4, 8, 16, 32),
                 enc_patch_size=(48, 48, 48, 48, 48), dec_depths=(2, 2, 2, 2), dec_channels=(64, 64, 128, 256),
                 dec_num_head=(4, 4, 8, 16), dec_patch_size=(48, 48, 48, 48), mlp_ratio=4, qkv_bias=True,
                 qk_scale=None, attn_drop=0.0, proj_drop=0.0, drop_path=0.3, layer_scale=None, pre_norm=True,
                 shuffle_orders=True, enable_rpe=False, enable_flash=True, upcast_attention=False,
                 upcast_softmax=False, traceable=False, mask_token=False, enc_mode=False, freeze_encoder=False,
                 compute_dtype=torch.bfloat16):
        super().__init__()
        self.num_stages = len(enc_depths)
        self.order = [order] if isinstance(order, str) else list(order)
        self.shuffle_orders = shuffle_orders
        self.enc_mode = enc_mode
        self.freeze_encoder = freeze_encoder
        assert self.num_stages == len(stride) + 1
        assert self.num_stages == len(enc_depths) == len(enc_channels) == len(enc_num_head) == len(enc_patch_size)
        assert self.enc_mode or self.num_stages == len(dec_depths) + 1
        assert self.enc_mode or self.num_stages == len(dec_channels) + 1
        assert self.enc_mode or self.num_stages == len(dec_num_head) + 1
        assert self.enc_mode or self.num_stages == len(dec_patch_size) + 1
        ln_layer, act_layer = nn.LayerNorm, nn.GELU
        self.embedding = Embedding(in_channels=in_channels, embed_channels=enc_channels[0], norm_layer=ln_layer,
                                   act_layer=act_layer, mask_token=mask_token)
        common = dict(mlp_ratio=mlp_ratio, qkv_bias=qkv_bias, qk_scale=qk_scale, attn_drop=attn_drop,
                      proj_drop=proj_drop, layer_scale=layer_scale, norm_layer=ln_layer, act_layer=act_layer,
                      pre_norm=pre_norm, enable_rpe=enable_rpe, enable_flash=enable_flash,
                      upcast_attention=upcast_attention, upcast_softmax=upcast_softmax)
        enc_dp = [x.item() for x in torch.linspace(0, drop_path, sum(enc_depths))]
        self.enc = PointSequential()
        for s in range(self.num_stages):
            enc_dp_ = enc_dp[sum(enc_depths[:s]):sum(enc_depths[:s + 1])]
            enc = PointSequential()
            if s > 0:
                enc.add(GridPooling(in_channels=enc_channels[s - 1], out_channels=enc_channels[s],
                                    stride=stride[s - 1], norm_layer=ln_layer, act_layer=act_layer), name="down")
            for i in range(enc_depths[s]):
                enc.add(Block(channels=enc_channels[s], num_heads=enc_num_head[s], patch_size=enc_patch_size[s],
                              drop_path=enc_dp_[i], order_index=i % len(self.order), cpe_indice_key=f"stage{s}",
                              **common), name=f"block{i}")
            if len(enc) != 0:
                self.enc.add(module=enc, name=f"enc{s}")
        if not self.enc_mode:
            dec_dp = [x.item() for x in torch.linspace(0, drop_path, sum(dec_depths))]
            self.dec = PointSequential()
            dec_channels = list(dec_channels) + [enc_channels[-1]]
            for s in reversed(range(self.num_stages - 1)):
                dec_dp_ = dec_dp[sum(dec_depths[:s]):sum(dec_depths[:s + 1])]
                dec_dp_.reverse()
                dec = PointSequential()
                dec.add(GridUnpooling(in_channels=dec_channels[s + 1], skip_channels=enc_channels[s],
                                      out_channels=dec_channels[s], norm_layer=ln_layer, act_layer=act_layer,
                                      traceable=traceable), name="up")
                for i in range(dec_depths[s]):
                    dec.add(Block(channels=dec_channels[s], num_heads=dec_num_head[s], patch_size=dec_patch_size[s],
                                  drop_path=dec_dp_[i], order_index=i % len(self.order), cpe_indice_key=f"stage{s}",
                                  **common), name=f"block{i}")
                self.dec.add(module=dec, name=f"dec{s}")
        if self.freeze_encoder:
            for p in self.embedding.parameters():
                p.requires_grad = False
            for p in self.enc.parameters():
                p.requires_grad = False
        self.apply(self._init_weights)
        self.plan_ahead = True
        self.fused_when_frozen = True      # A/B switch: frozen backbone in train() runs the fused kernels (train.py)
        self.set_compute_dtype(compute_dtype)

    @staticmethod
    def _init_weights(module):
        if isinstance(module, (nn.Linear, SubMConv3d)):      # :706-716
            nn.init.trunc_normal_(module.weight, std=0.02)
            if module.bias is not None:
                nn.init.zeros_(module.bias)

    def set_compute_dtype(self, dtype):
        assert dtype in (torch.bfloat16, torch.float32)
        self.compute_dtype = dtype
        for m in self.modules():
            if hasattr(m, "compute_dtype"):
                m.compute_dtype = dtype
        return self

    def set_algos(self, conv=0, attn=0):
        from .model import SerializedAttention
        for m in self.modules():
            if isinstance(m, Block):
                m.conv_algo = conv
            if isinstance(m, SerializedAttention):
                m.algo = attn
        return self

    @torch.no_grad()
    def plan_levels(self, point):
        """Serialization of level 0 and the integer plans of every GridPooling, in the reference's order of
        torch.randperm draws (forward :724, then one per GridPooling :460)."""
        point.serialization(order=self.order, shuffle_orders=self.shuffle_orders)
        point.sparsify()
        if not self.plan_ahead:
            return point
        plans = []
        grid, batch, depth, n_scenes = point.grid_coord, point.batch, int(point.serialized_depth), len(point.offset)
        for stage in self.enc._modules.values():
            down = stage._modules.get("down")
            if down is None:
                continue
            plans.append(down.plan(grid, batch, depth, n_scenes, self.order))
            grid, batch, depth = plans[-1]["grid_coord"], plans[-1]["batch"], plans[-1]["depth"]
        point["_pool_plans"] = plans
        return point

    def forward(self, data_dict):
        if self.training:
            from . import train
            return train.backbone_forward_m2(self, data_dict)
        return self._forward_eval(data_dict)

    @torch.no_grad()
    def _forward_eval(self, data_dict):
        point = Point(data_dict)
        point = self.plan_levels(point)          # integer work first (the embedding does not depend on it)
        point = self.embedding(point)
        point = self.enc(point)
        if not self.enc_mode:
            point = self.dec(point)
        return point
