"""PT-v3m1 backbone + the fork's heads over the sm_100a kernels (inference).

Same module tree, constructor arguments and state_dict names as the reference
(pointcept/models/point_transformer_v3/point_transformer_v3m1_base.py,
offset_keypoint_ptv3.py, keypoint_ptv3.py, default.py) so its checkpoints load
unchanged; the forward of every module is re-designed around our fused ops:

  * the residual stream (`point.feat`) is fp32; GEMM / conv / attention operands are
    `compute_dtype` (bf16 by default, fp32 for the 1e-4 parity path) -- what the
    reference gets from torch.autocast;
  * eval-mode BatchNorm is folded into the adjacent Linear / conv weights;
  * residual adds, LayerNorms and the casts between them are one kernel
    (ops.ln_residual); the two GELUs + gather-add of unpooling are one kernel;
  * attention reads / writes rows in storage order through the serialized order
    (no [order] gather, no [inverse] scatter, no padding copies);
  * pooling derives clusters and the pooled orders from the parent's orders
    without sorting.
Attention follows BOTH window semantics of the reference: enable_flash=True keeps the
fixed patch size (short scenes form one short window, :208-215); enable_flash=False
(the setting of both fork configs) uses K = min(smallest scene of the batch, patch)
for every scene (:173-176), which differs from the flash windows as soon as a batch
holds a scene shorter than the patch.
Training: in train() mode every entry point dispatches to train.py (autograd through the kernels, autograd.py).
"""
import math

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import ops
from .structure import Point, SparseConvTensor, device_perm, offset2batch, z_info  # noqa: F401


def _cached(module, slot, params, builder):
    """Per-module cache of prepared (folded / re-laid / cast) weights keyed on parameter versions."""
    store = module.__dict__.setdefault("_prep_cache", {})
    hit = store.get(slot)
    if hit is not None:     # fast path: same tensors (storage + version); dtype / device changes move data_ptr
        for p, (ptr, ver) in zip(params, hit[2]):
            if p is not None and (p._version != ver or p.data_ptr() != ptr):
                break
        else:
            return hit[1]
    key = tuple((p.data_ptr(), p._version, p.dtype, p.device) for p in params if p is not None)
    if hit is not None and hit[0] == key:
        return hit[1]
    with torch.no_grad():
        val = builder()
    store[slot] = (key, val, [(p.data_ptr(), p._version) if p is not None else (0, 0) for p in params])
    return val


def _bn_scale_shift(bn):
    s = bn.weight.float() / torch.sqrt(bn.running_var.float() + bn.eps)
    return s, bn.bias.float() - bn.running_mean.float() * s


def _feat_cd(point, dtype):
    """point.feat in the compute dtype (cached copy written by the producing kernel when there is one)."""
    f = point.get("_feat_cd")
    if f is not None and f.dtype == dtype and f.shape == point.feat.shape:
        return f
    return point.feat if point.feat.dtype == dtype else point.feat.to(dtype)


class PointModule(nn.Module):
    """modules.py:28-34."""


class SparseModule(nn.Module):
    """spconv.pytorch.modules.SparseModule stand-in (modules.py:84 tests isinstance)."""


def pad_out_channels(c):
    """Smallest output width the tcgen05 conv kernel tiles (32 ... 512) that holds c channels."""
    for p in (32, 64, 128, 256, 512):
        if c <= p:
            return p
    return c


def is_spconv_module(m):
    return isinstance(m, SparseModule)


class PointSequential(PointModule):
    """modules.py:37-111: named container that dispatches on the module kind."""

    def __init__(self, *args, **kwargs):
        super().__init__()
        for i, m in enumerate(args):
            self.add_module(str(i), m)
        for k, m in kwargs.items():
            self.add_module(k, m)

    def __getitem__(self, idx):
        return list(self._modules.values())[idx]

    def __len__(self):
        return len(self._modules)

    def add(self, module, name=None):
        self.add_module(str(len(self._modules)) if name is None else name, module)
        return self

    def forward(self, input):
        for module in self._modules.values():
            if isinstance(module, PointModule):
                input = module(input)
            elif is_spconv_module(module):
                if isinstance(input, Point):
                    input.sparse_conv_feat = module(input.sparse_conv_feat)
                    input.feat = input.sparse_conv_feat.features
                else:
                    input = module(input)
            else:
                if isinstance(input, Point):
                    input.feat = module(input.feat)
                    input["_feat_cd"] = None
                    if "sparse_conv_feat" in input:
                        input.sparse_conv_feat = input.sparse_conv_feat.replace_feature(input.feat)
                elif isinstance(input, SparseConvTensor):
                    if input.indices.shape[0] != 0:
                        input = input.replace_feature(module(input.features))
                else:
                    input = module(input)
        return input


class SubMConv3d(SparseModule):
    """spconv.pytorch.SubMConv3d surface (call sites :278-284, :499-506): weight (out,kD,kH,kW,in), bias (out),
    forward(SparseConvTensor) -> SparseConvTensor on the same sites; the neighbour map is cached in
    x.indice_dict[(indice_key, k)] and shared through replace_feature.  `padding` is ignored like in spconv 2.x."""

    def __init__(self, in_channels, out_channels, kernel_size, stride=1, padding=0, dilation=1, groups=1, bias=True,
                 indice_key=None):
        super().__init__()
        if stride != 1 or dilation != 1 or groups != 1:
            raise NotImplementedError("SubMConv3d: stride/dilation/groups other than 1 are not on the PTv3 path")
        if not isinstance(kernel_size, int):
            assert len(set(kernel_size)) == 1
            kernel_size = kernel_size[0]
        self.in_channels, self.out_channels, self.kernel_size = in_channels, out_channels, kernel_size
        self.indice_key = indice_key
        k = kernel_size
        self.weight = nn.Parameter(torch.empty(out_channels, k, k, k, in_channels))
        self.bias = nn.Parameter(torch.empty(out_channels)) if bias else None
        bound = 1.0 / math.sqrt(in_channels * k ** 3)
        nn.init.uniform_(self.weight, -bound, bound)
        if bias:
            nn.init.uniform_(self.bias, -bound, bound)

    def neighbor_map(self, x):
        if self.kernel_size == 1:          # a 1x1x1 SubM conv is a per-site Linear: the table is the identity
            n = x.indices.shape[0]
            return torch.arange(n, dtype=torch.int32, device=x.indices.device).view(n, 1)
        key = (self.indice_key, self.kernel_size)
        nbr = x.indice_dict.get(key) if self.indice_key is not None else None
        if nbr is None:
            nbr = ops.subm_build_map(x.indices, self.kernel_size, z=x.indice_dict.get("_z") if self.block_search else None)
            if self.indice_key is not None:
                x.indice_dict[key] = nbr
        return nbr

    def _weights(self, dtype, scale=None, shift=None, cin_p=None, cout_p=None):
        cin_p, cout_p = cin_p or self.in_channels, cout_p or self.out_channels

        def build():
            w = self.weight.float()
            b = self.bias.float() if self.bias is not None else None
            if scale is not None:      # fold an eval-mode BatchNorm that follows the conv
                w = w * scale.view(-1, 1, 1, 1, 1)
                b = shift if b is None else b * scale + shift
            k3 = self.kernel_size ** 3
            w = w.reshape(self.out_channels, k3, self.in_channels).permute(1, 0, 2)
            if (cin_p, cout_p) != (self.in_channels, self.out_channels):     # zero rows / columns (pad_channels)
                w = F.pad(w, (0, cin_p - self.in_channels, 0, cout_p - self.out_channels))
                if b is not None:
                    b = F.pad(b, (0, cout_p - self.out_channels))
            return w.contiguous().to(dtype), (b.contiguous() if b is not None else None)
        return _cached(self, ("w", dtype, scale is not None, cin_p, cout_p), [self.weight, self.bias, scale, shift],
                       build)

    # A/B switch: (N, k^3) neighbour tables through 4^3-block occupancy masks when the level is serialized
    # (ops.subm_build_map(z=...)).  Off by default: measured equal to the hash-probe kernel for k = 3 (0.44 ms per step
    # either way -- 27 probes per row were never the bound there, unlike the stem's 125; profiles/r02_fusions.md)
    block_search = False
    pad_channels = False     # SpUNet sets this: 96-channel levels run as 128-wide tensors whose last columns are zero

    def conv(self, features, x, scale=None, shift=None, algo=0):
        k3 = self.kernel_size ** 3
        cin_p, cout_p = self.in_channels, self.out_channels
        if self.pad_channels and features.dtype == torch.bfloat16:
            cin_p, cout_p = features.shape[1], pad_out_channels(self.out_channels)
        w, b = self._weights(features.dtype, scale, shift, cin_p, cout_p)
        if k3 == 1:      # 1x1x1 SubM conv = a plain per-site GEMM: library call (cuBLAS), no gather
            return F.linear(features, w[0], b.to(features.dtype) if b is not None else None)
        if features.dtype == torch.bfloat16 and algo != 1 and ops.subm_conv_tc_supported(cin_p, cout_p, k3):
            # tcgen05 path: the K-blocked / swizzled weight image is built once per parameter version
            packed = _cached(self, ("packed", scale is not None, cin_p, cout_p),
                             [self.weight, self.bias, scale, shift], lambda: ops.subm_pack_weight(w))
            return ops.subm_conv_packed(features, packed, b, self.neighbor_map(x), cout_p)
        return ops.subm_conv(features, w, b, self.neighbor_map(x), algo=algo)

    def forward(self, x):
        """spconv.pytorch.SubMConv3d.forward as the unmodified reference calls it (modules.py:84-87).  Under
        torch.autocast(bfloat16) -- how the reference's trainer / tester run the model (engines/train.py:202) -- the
        features are taken in bf16 like spconv's autocast-aware ops do, which puts the call on the tcgen05 kernel."""
        f = x.features
        if (f.is_cuda and f.dtype == torch.float32 and torch.is_autocast_enabled()
                and torch.get_autocast_gpu_dtype() == torch.bfloat16 and self.kernel_size > 1
                and ops.subm_conv_tc_supported(self.in_channels, self.out_channels, self.kernel_size ** 3)):
            f = f.to(torch.bfloat16)
        return x.replace_feature(self.conv(f, x))


class SerializedAttention(PointModule):
    """:51-222.  qkv Linear -> fused window attention kernel -> proj Linear."""

    def __init__(self, channels, num_heads, patch_size, qkv_bias=True, qk_scale=None, attn_drop=0.0, proj_drop=0.0,
                 order_index=0, enable_rpe=False, enable_flash=True, upcast_attention=True, upcast_softmax=True):
        super().__init__()
        assert channels % num_heads == 0
        if enable_rpe:
            raise NotImplementedError("enable_rpe=True is not used by any target config (SURVEY.md 8a A4)")
        if channels // num_heads != 16:
            raise NotImplementedError("head_dim must be 16 (true for every PTv3 config, SURVEY.md fact 1)")
        if attn_drop != 0.0 or proj_drop != 0.0:
            raise NotImplementedError("attn_drop / proj_drop are 0 in every target config; the kernels have no dropout")
        self.channels, self.num_heads = channels, num_heads
        self.scale = qk_scale or (channels // num_heads) ** -0.5
        self.order_index = order_index
        self.enable_flash = bool(enable_flash)
        self.patch_size_max = patch_size
        self.patch_size = patch_size
        self.qkv = nn.Linear(channels, channels * 3, bias=qkv_bias)
        self.proj = nn.Linear(channels, channels)
        self.compute_dtype = torch.bfloat16
        self.algo = 0

    @torch.no_grad()
    def get_windows(self, point):
        """Replaces get_padding_and_inverse (:115-170): per-window (q_start,q_len,k_start,k_len) records, cached in
        the Point like the reference caches pad/unpad/cu_seqlens."""
        if not self.enable_flash:
            # dense branch (:173-176): ONE patch size for the whole batch, min(smallest scene, patch_size_max).  Like
            # the reference this reads the scene sizes on the host (once per Point: every block of a stage shares it)
            if "attn_patch_dense" not in point:
                off = point.offset
                counts = torch.diff(off, prepend=off.new_zeros(1))
                point["attn_patch_dense"] = int(counts.min().item())
            self.patch_size = max(1, min(point["attn_patch_dense"], self.patch_size_max))
        key = f"windows_{self.patch_size}"
        if key not in point:
            point[key] = ops.build_windows(point.offset, self.patch_size, point.feat.shape[0])
        return point[key]

    def attend(self, point, qkv):
        windows, nwin = self.get_windows(point)
        order = point.serialized_order[self.order_index]
        return ops.attn_window(qkv, order, windows, nwin, self.num_heads, self.patch_size, self.scale, self.algo)

    def _w(self, lin, dtype):
        return _cached(self, (id(lin), dtype), [lin.weight, lin.bias],
                       lambda: (lin.weight.to(dtype), lin.bias.to(dtype) if lin.bias is not None else None))

    def forward(self, point):
        cd = self.compute_dtype
        wq, bq = self._w(self.qkv, cd)
        wp, bp = self._w(self.proj, cd)
        a = self.attend(point, F.linear(_feat_cd(point, cd), wq, bq))
        point.feat = F.linear(a, wp, bp).float()
        point["_feat_cd"] = None
        return point


class MLP(nn.Module):
    """:225-248."""

    def __init__(self, in_channels, hidden_channels=None, out_channels=None, act_layer=nn.GELU, drop=0.0):
        super().__init__()
        out_channels = out_channels or in_channels
        hidden_channels = hidden_channels or in_channels
        self.fc1 = nn.Linear(in_channels, hidden_channels)
        self.act = act_layer()
        self.fc2 = nn.Linear(hidden_channels, out_channels)
        self.drop = nn.Dropout(drop)

    def forward(self, x):
        return self.drop(self.fc2(self.drop(self.act(self.fc1(x)))))


class LayerScale(nn.Module):
    """point_transformer_v3m2_sonata.py:27-39: per-channel scale of a residual branch.  In the inference path it never
    runs as a kernel: gamma is folded into the weights of the Linear that ends the branch (attn.proj / mlp.fc2)."""

    def __init__(self, dim, init_values=1e-5, inplace=False):
        super().__init__()
        self.inplace = inplace
        self.gamma = nn.Parameter(init_values * torch.ones(dim))

    def forward(self, x):
        return x.mul_(self.gamma) if self.inplace else x * self.gamma


class Block(PointModule):
    """:251-338.  xCPE (SubMConv3d -> Linear -> LN) + attention + MLP with pre-norm; three fused
    residual/LayerNorm kernels replace the six elementwise passes of the reference."""

    def __init__(self, channels, num_heads, patch_size=48, mlp_ratio=4.0, qkv_bias=True, qk_scale=None, attn_drop=0.0,
                 proj_drop=0.0, drop_path=0.0, norm_layer=nn.LayerNorm, act_layer=nn.GELU, pre_norm=True,
                 order_index=0, cpe_indice_key=None, enable_rpe=False, enable_flash=True, upcast_attention=True,
                 upcast_softmax=True, layer_scale=None):
        super().__init__()
        if not pre_norm:
            raise NotImplementedError("pre_norm=False is not used by any target config")
        self.channels, self.pre_norm = channels, pre_norm
        if layer_scale is not None:      # PT-v3m2 (point_transformer_v3m2_sonata.py:304-308, :325-329): state_dict ls1.0.gamma
            self.ls1 = PointSequential(LayerScale(channels, init_values=layer_scale))
            self.ls2 = PointSequential(LayerScale(channels, init_values=layer_scale))
        self.cpe = PointSequential(
            SubMConv3d(channels, channels, kernel_size=3, bias=True, indice_key=cpe_indice_key),
            nn.Linear(channels, channels), norm_layer(channels))
        self.norm1 = PointSequential(norm_layer(channels))
        self.attn = SerializedAttention(channels=channels, patch_size=patch_size, num_heads=num_heads,
                                        qkv_bias=qkv_bias, qk_scale=qk_scale, attn_drop=attn_drop,
                                        proj_drop=proj_drop, order_index=order_index, enable_rpe=enable_rpe,
                                        enable_flash=enable_flash, upcast_attention=upcast_attention,
                                        upcast_softmax=upcast_softmax)
        self.norm2 = PointSequential(norm_layer(channels))
        self.mlp = PointSequential(MLP(in_channels=channels, hidden_channels=int(channels * mlp_ratio),
                                       out_channels=channels, act_layer=act_layer, drop=proj_drop))
        self.drop_path = PointSequential(nn.Identity())   # DropPath is identity in eval (timm); no parameters
        self.drop_path_rate = float(drop_path)            # used by the training forward (train.block_forward)
        self.compute_dtype = torch.bfloat16
        self.conv_algo = 0

    def _lin(self, lin, dtype, ls=None):
        """Weights of a Linear in the compute dtype; `ls` (a LayerScale gamma) is folded in: gamma * (W x + b)."""
        if ls is None:
            return _cached(self, (id(lin), dtype), [lin.weight, lin.bias],
                           lambda: (lin.weight.to(dtype), lin.bias.to(dtype) if lin.bias is not None else None))
        return _cached(self, (id(lin), dtype, "ls"), [lin.weight, lin.bias, ls],
                       lambda: ((lin.weight.float() * ls.float()[:, None]).to(dtype).contiguous(),
                                (lin.bias.float() * ls.float()).to(dtype) if lin.bias is not None else None))

    def _gammas(self):
        ls1, ls2 = self._modules.get("ls1"), self._modules.get("ls2")
        return (ls1[0].gamma if ls1 is not None else None), (ls2[0].gamma if ls2 is not None else None)

    @staticmethod
    def _ln(ln):
        return ln.weight, ln.bias, ln.eps

    native = True      # one C-ABI call per Block (ops.block_fwd) when the shapes allow; False = the op-by-op path
    # A/B switch: attn.proj + LayerNorm + MLP in one kernel (ptv3_proj_mlp_fused, C = 32 / 64) inside the native call.
    # OFF by default: once the fused MLP got its own output warps and 256-bit accesses, the separate launches (cuBLASLt proj,
    # ln_residual, mlp_fused) became faster than the one kernel at both widths (27.45 vs 27.70 ms per step with it on at
    # C = 32; profiles/r02_fusions.md).  True = at C = 32, "all" = at C = 32 and 64; PTV3_FUSE_PROJ=1 / all sets it.
    fuse_proj = {"1": True, "all": "all"}.get(__import__("os").environ.get("PTV3_FUSE_PROJ", "0"), False)

    def _native_params(self):
        """ptv3_block_params of this Block (bf16 inference), rebuilt when a parameter changes."""
        attn, mlp, conv = self.attn, self.mlp[0], self.cpe[0]
        c, hidden = self.channels, mlp.fc1.out_features
        cd = torch.bfloat16
        lins = [self.cpe[1], attn.qkv, attn.proj, mlp.fc1, mlp.fc2]
        lns = [self.cpe[2], self.norm1[0], self.norm2[0]]
        g1, g2 = self._gammas()
        params = [conv.weight, conv.bias] + [q for l in lins for q in (l.weight, l.bias)] + \
                 [q for l in lns for q in (l.weight, l.bias)] + [g1, g2]

        def build():
            fused = ops.mlp_fused_supported(c, hidden) and mlp.fc2.out_features == c
            dev = conv.weight.device
            zeros = lambda n, dt: torch.zeros(n, dtype=dt, device=dev)
            sc = lambda t, g: t.float() if g is None else t.float() * (g.float()[:, None] if t.dim() == 2 else g.float())
            lw = lambda l, g=None: sc(l.weight, g).to(cd).contiguous()
            lb = lambda l, g=None: sc(l.bias, g).to(cd).contiguous() if l.bias is not None else zeros(l.out_features, cd)
            w, b = conv._weights(cd)
            t = dict(conv_w_packed=ops.subm_pack_weight(w), conv_bias=b if b is not None else zeros(c, torch.float32),
                     cpe_lin_w=lw(self.cpe[1]), cpe_lin_b=lb(self.cpe[1]),
                     cpe_ln_g=self.cpe[2].weight.float().contiguous(), cpe_ln_b=self.cpe[2].bias.float().contiguous(),
                     norm1_g=self.norm1[0].weight.float().contiguous(), norm1_b=self.norm1[0].bias.float().contiguous(),
                     norm2_g=self.norm2[0].weight.float().contiguous(), norm2_b=self.norm2[0].bias.float().contiguous(),
                     qkv_w=lw(attn.qkv), qkv_b=lb(attn.qkv), proj_w=lw(attn.proj, g1), proj_b=lb(attn.proj, g1))
            if fused:
                t.update(fc1_w=ops.pack_kmajor(mlp.fc1.weight.to(cd), ops.mlp_fc1_tile_rows(c)), fc1_b=mlp.fc1.bias.float().contiguous(),
                         fc2_w=ops.pack_kmajor(lw(mlp.fc2, g2), c), fc2_b=sc(mlp.fc2.bias, g2).contiguous())
                # proj + LN + MLP as one kernel (see fuse_proj above)
                if ((self.fuse_proj == "all" or (self.fuse_proj and c == 32)) and attn.proj.bias is not None):
                    t.update(proj_w_packed=ops.pack_kmajor(lw(attn.proj, g1), c),
                             proj_b_f32=sc(attn.proj.bias, g1).contiguous())
            else:
                t.update(fc1_w=lw(mlp.fc1), fc1_b=lb(mlp.fc1), fc2_w=lw(mlp.fc2, g2), fc2_b=lb(mlp.fc2, g2))
            return ops.block_params(c=c, heads=attn.num_heads, hidden=hidden, k3=conv.kernel_size ** 3,
                                    patch=attn.patch_size, mlp_fused=int(fused), attn_algo=attn.algo,
                                    scale=float(attn.scale), eps_cpe=float(self.cpe[2].eps),
                                    eps_norm2=float(self.norm2[0].eps), **t)
        return _cached(self, ("native", self.fuse_proj), params, build)

    def _fast(self):
        """Sub-module references and the static half of the native-path test, resolved once (nn.Module attribute
        lookups go through __getattr__; a Block runs 22 times per forward)."""
        f = self.__dict__.get("_fast_refs")
        if f is None:
            mlp = self.mlp[0]
            ok = (isinstance(mlp.act, nn.GELU) and mlp.act.approximate == "none"
                  and all(isinstance(m, nn.LayerNorm) for m in (self.cpe[2], self.norm1[0], self.norm2[0]))
                  and float(self.cpe[2].eps) == float(self.norm1[0].eps)
                  and ops.subm_conv_tc_supported(self.channels, self.channels, self.cpe[0].kernel_size ** 3)
                  and mlp.fc1.out_features % 8 == 0)
            f = self.__dict__["_fast_refs"] = dict(conv=self.cpe[0], attn=self.attn, static_ok=ok)
        return f

    def _native_ok(self, cd):
        return (self.native and ops.NATIVE_BLOCK and cd == torch.bfloat16 and self.conv_algo != 1
                and self._fast()["static_ok"])

    def _drop_scale(self, n, device):
        """timm DropPath in train() mode: one Bernoulli(keep) draw per row, kept rows scaled by 1 / keep (the same draw,
        in the same order, as train._residual); None in eval() or with rate 0."""
        if not self.training or self.drop_path_rate == 0.0:
            return None
        keep = 1.0 - self.drop_path_rate
        return torch.empty((n, 1), dtype=torch.float32, device=device).bernoulli_(keep).div_(keep)

    def forward(self, point):
        """Fused forward WITHOUT a graph.  eval(): the inference path.  train() under torch.no_grad(): a frozen backbone
        in a training step (DefaultSegmentorV2(freeze_backbone=True) leaves it in train() mode, default.py:58-60), i.e.
        the same kernels with DropPath applied inside the residual kernels; anything that needs gradients goes
        through train.block_forward."""
        if self.training and torch.is_grad_enabled():
            raise NotImplementedError("ptv3_b200 Block.forward builds no graph: use train.block_forward (the model's "
                                      "forward dispatches to it in train() mode) or call .eval()")
        with torch.no_grad():
            return self._forward(point)

    def _forward(self, point):
        cd = self.compute_dtype
        x = point.feat if point.feat.dtype == torch.float32 else point.feat.float()
        sp = point.sparse_conv_feat
        conv_in = sp.features if sp.features.dtype == cd else sp.features.to(cd)
        if self._native_ok(cd):
            f = self._fast()
            attn = f["attn"]
            bp, _keep = self._native_params()
            windows, nwin = attn.get_windows(point)
            bp.patch, bp.attn_algo = int(attn.patch_size), int(attn.algo)
            d1 = self._drop_scale(x.shape[0], x.device)         # train() + no graph: DropPath inside the residual kernels
            d2 = self._drop_scale(x.shape[0], x.device)
            bp.drop_attn, bp.drop_mlp = (d1.data_ptr() if d1 is not None else None), \
                (d2.data_ptr() if d2 is not None else None)
            x, xcd = ops.block_fwd(bp, x.contiguous(), conv_in.contiguous(), f["conv"].neighbor_map(sp),
                                   point.serialized_order[attn.order_index], windows, nwin)
            point.feat = x
            point["_feat_cd"] = xcd
            point.sparse_conv_feat = sp.replace_feature(xcd)
            return point
        # cpe: SubMConv3d -> Linear -> LayerNorm, + shortcut, then pre-norm of the attention
        y = self.cpe[0].conv(conv_in, sp, algo=self.conv_algo)
        y = F.linear(y, *self._lin(self.cpe[1], cd))
        g1, b1, eps1 = self._ln(self.cpe[2])
        gn1, bn1, _ = self._ln(self.norm1[0])
        x, h = ops.ln_residual(y, g1, b1, eps1, residual=x, gamma2=gn1, beta2=bn1, out2_dtype=cd)
        # attention
        qkv = F.linear(h, *self.attn._w(self.attn.qkv, cd))
        a = self.attn.attend(point, qkv)
        ls1, ls2 = self._gammas()
        p = F.linear(a, *self._lin(self.attn.proj, cd, ls1))
        gn2, bn2, eps2 = self._ln(self.norm2[0])
        x, h = ops.ln_residual(p, None, None, eps2, residual=x, gamma2=gn2, beta2=bn2, out2_dtype=cd,
                               row_scale=self._drop_scale(x.shape[0], x.device))
        # mlp
        mlp = self.mlp[0]
        hidden = mlp.fc1.out_features
        if (cd == torch.bfloat16 and isinstance(mlp.act, nn.GELU) and mlp.act.approximate == "none"
                and ops.mlp_fused_supported(self.channels, hidden) and mlp.fc2.out_features == self.channels):
            # fc1 -> GELU -> fc2 -> + residual in one tcgen05 kernel: the (N, 4C) hidden activation stays on the SM
            w1p, b1, w2p, b2 = _cached(
                self, "mlp_packed", [mlp.fc1.weight, mlp.fc1.bias, mlp.fc2.weight, mlp.fc2.bias, ls2],
                lambda: (ops.pack_kmajor(mlp.fc1.weight.to(cd), ops.mlp_fc1_tile_rows(self.channels)), mlp.fc1.bias.float().contiguous(),
                         ops.pack_kmajor(self._lin(mlp.fc2, cd, ls2)[0], self.channels),
                         (mlp.fc2.bias.float() * (ls2.float() if ls2 is not None else 1.0)).contiguous()))
            x, xcd = ops.mlp_fused(h, w1p, b1, w2p, b2, hidden, residual=x,
                                   row_scale=self._drop_scale(x.shape[0], x.device))
        else:
            f = F.linear(h, *self._lin(mlp.fc1, cd))
            if cd == torch.bfloat16 and isinstance(mlp.act, nn.GELU) and mlp.act.approximate == "none":
                _, f = ops.affine_act(f, None, None, "gelu", want_f32=False, want_bf16=True)    # 16-byte vectorised GELU
            else:
                f = mlp.act(f)
            f = F.linear(f, *self._lin(mlp.fc2, cd, ls2))
            x, xcd = ops.ln_residual(f, None, None, 1e-5, residual=x, out2_dtype=None if cd == torch.float32 else cd,
                                     row_scale=self._drop_scale(x.shape[0], x.device))
        point.feat = x
        point["_feat_cd"] = xcd if xcd is not None else x
        point.sparse_conv_feat = sp.replace_feature(point["_feat_cd"])
        return point


class SerializedPooling(PointModule):
    """:341-444.  Integer front end without sorts (ops.pool_front), gather-fused segment max / mean."""

    def __init__(self, in_channels, out_channels, stride=2, norm_layer=None, act_layer=None, reduce="max",
                 shuffle_orders=True, traceable=True):
        super().__init__()
        assert stride == 2 ** (math.ceil(stride) - 1).bit_length()
        assert reduce in ["sum", "mean", "min", "max"]
        self.in_channels, self.out_channels, self.stride = in_channels, out_channels, stride
        self.reduce, self.shuffle_orders, self.traceable = reduce, shuffle_orders, traceable
        self.proj = nn.Linear(in_channels, out_channels)
        if norm_layer is not None:
            self.norm = PointSequential(norm_layer(out_channels))
        if act_layer is not None:
            self.act = PointSequential(act_layer())
        self.compute_dtype = torch.bfloat16

    def pooling_depth(self, serialized_depth):
        d = (math.ceil(self.stride) - 1).bit_length()
        return 0 if d > serialized_depth else d

    @torch.no_grad()
    def plan(self, code, order, grid_coord, batch, pooling_depth, n_scenes, z_row=None):
        """The integer half of the pooling (:385-412, :426-433): clusters, CSR, heads, pooled codes / orders / grid /
        batch / offset.  It depends on the serialization only, so PointTransformerV3.forward runs it for ALL levels
        before the first feature kernel: the one host read per level (pooled row count) then waits for a few tiny
        integer kernels instead of draining the whole previous stage, and the feature pipeline has no sync at all."""
        pf = ops.pool_front(code, order, pooling_depth)
        g, b, indices4 = ops.pool_gather_ints(grid_coord, batch, pf["head"], pooling_depth)
        c, o, inv = pf["code"], pf["order"], pf["inverse"]
        if self.shuffle_orders:
            perm = torch.randperm(c.shape[0])
            if z_row is not None:
                z_row = perm.tolist().index(z_row)
            if not torch.equal(perm, torch.arange(c.shape[0])):
                perm = device_perm(perm, c.device)
                c, o, inv = c[perm], o[perm], inv[perm]
        return dict(pf=pf, grid_coord=g, batch=b, indices4=indices4, code=c, order=o, inverse=inv,
                    offset=ops.batch_to_offset(b, n_scenes), z_row=z_row)

    @torch.no_grad()
    def forward(self, point):
        cd = self.compute_dtype
        pooling_depth = self.pooling_depth(point.serialized_depth)
        assert {"serialized_code", "serialized_order", "serialized_inverse", "serialized_depth"}.issubset(
            point.keys()), "Run point.serialization() point cloud before SerializedPooling"
        plans = point.pop("_pool_plans", None)
        plan = plans.pop(0) if plans else self.plan(point.serialized_code, point.serialized_order, point.grid_coord,
                                                    point.batch, pooling_depth, len(point.offset), point.get("_z_row"))
        pf, grid_coord, batch, indices4 = plan["pf"], plan["grid_coord"], plan["batch"], plan["indices4"]
        code, order, inverse = plan["code"], plan["order"], plan["inverse"]
        w, b = _cached(self, ("proj", cd), [self.proj.weight, self.proj.bias],
                       lambda: (self.proj.weight.to(cd), self.proj.bias.to(cd)))
        y = F.linear(_feat_cd(point, cd), w, b)
        pooled = ops.segment_csr(y, pf["idx_ptr"], self.reduce, gather=pf["indices"])
        coord = ops.segment_csr(point.coord.float(), pf["idx_ptr"], "mean", gather=pf["indices"])
        bn = self.norm[0] if hasattr(self, "norm") else None
        act = self.act[0] if hasattr(self, "act") else None
        gelu = isinstance(act, nn.GELU) and act.approximate == "none"
        feat_cd = None
        if (bn is None or (isinstance(bn, nn.BatchNorm1d) and not bn.training)) and (act is None or gelu) \
                and pooled.shape[1] % 4 == 0:
            # eval BatchNorm + GELU + the bf16 copy for the next conv / GEMM in one pass over the pooled rows
            s, t = (None, None) if bn is None else _cached(
                self, "bn", [bn.weight, bn.bias, bn.running_mean, bn.running_var], lambda: _bn_scale_shift(bn))
            feat, feat_cd = ops.affine_act(pooled, s, t, "gelu" if gelu else "none", want_f32=True,
                                           want_bf16=cd == torch.bfloat16)
        else:
            feat = pooled.float()
            if bn is not None:
                feat = bn(feat)
            if act is not None:
                feat = act(feat)
        n_scenes = len(point.offset)
        new = Point(feat=feat, coord=coord, grid_coord=grid_coord, serialized_code=code, serialized_order=order,
                    serialized_inverse=inverse, serialized_depth=point.serialized_depth - pooling_depth, batch=batch,
                    offset=plan["offset"])
        if plans:
            new["_pool_plans"] = plans
        for k in ("condition", "context", "grid_size", "order"):
            if k in point:
                new[k] = point[k]
        if self.traceable:
            new["pooling_inverse"] = pf["cluster"]
            new["pooling_parent"] = point
        new["_feat_cd"] = feat_cd if feat_cd is not None else (feat if cd == torch.float32 else feat.to(cd))
        new["sparse_shape"] = [(1 << max(int(new.serialized_depth), 0)) + 96] * 3
        new["sparse_conv_feat"] = SparseConvTensor(new["_feat_cd"], indices4, new["sparse_shape"], n_scenes)
        new["_z_row"] = plan.get("z_row")
        z = z_info(new)
        if z is not None:
            new["sparse_conv_feat"].indice_dict["_z"] = z
        return new


class SerializedUnpooling(PointModule):
    """:447-482.  Both Linear+BN pairs folded; GELU, GELU and the gather-add are one kernel."""

    def __init__(self, in_channels, skip_channels, out_channels, norm_layer=None, act_layer=None, traceable=False):
        super().__init__()
        self.proj = PointSequential(nn.Linear(in_channels, out_channels))
        self.proj_skip = PointSequential(nn.Linear(skip_channels, out_channels))
        if norm_layer is not None:
            self.proj.add(norm_layer(out_channels))
            self.proj_skip.add(norm_layer(out_channels))
        self.gelu = act_layer is not None and issubclass(act_layer, nn.GELU)
        if act_layer is not None:
            if not self.gelu:
                raise NotImplementedError("SerializedUnpooling: only GELU is fused")
            self.proj.add(act_layer())
            self.proj_skip.add(act_layer())
        self.traceable = traceable
        self.compute_dtype = torch.bfloat16

    def _folded(self, seq, dtype):
        lin = seq[0]
        bn = seq[1] if len(seq) > 1 and isinstance(seq[1], nn.BatchNorm1d) else None
        params = [lin.weight, lin.bias] + ([bn.weight, bn.bias, bn.running_mean, bn.running_var] if bn else [])

        def build():
            w, b = lin.weight.float(), lin.bias.float()
            if bn is not None:
                s, t = _bn_scale_shift(bn)
                w, b = w * s[:, None], b * s + t
            return w.to(dtype).contiguous(), b.to(dtype).contiguous()
        return _cached(self, (id(seq), dtype), params, build)

    @torch.no_grad()
    def forward(self, point):
        assert "pooling_parent" in point and "pooling_inverse" in point
        cd = self.compute_dtype
        parent = point.pop("pooling_parent")
        inverse = point.pop("pooling_inverse")
        up = F.linear(_feat_cd(point, cd), *self._folded(self.proj, cd))
        skip = F.linear(_feat_cd(parent, cd), *self._folded(self.proj_skip, cd))
        out, skip_act = ops.unpool_add(skip, up, inverse, gelu=self.gelu)
        parent.feat = out
        parent["_feat_cd"] = None
        # the reference leaves the parent's sparse tensor holding the proj_skip branch only (PointSequential updates
        # it before `parent.feat = parent.feat + point.feat[inverse]`, :476-477): the next conv reads that.
        parent.sparse_conv_feat = parent.sparse_conv_feat.replace_feature(skip_act)
        if self.traceable:
            parent["unpooling_parent"] = point
        return parent


class Embedding(PointModule):
    """:485-516.  Stem SubMConv3d(k5) with the eval BatchNorm folded, then GELU."""

    def __init__(self, in_channels, embed_channels, norm_layer=None, act_layer=None):
        super().__init__()
        self.in_channels, self.embed_channels = in_channels, embed_channels
        self.stem = PointSequential(conv=SubMConv3d(in_channels, embed_channels, kernel_size=5, padding=1, bias=False,
                                                    indice_key="stem"))
        if norm_layer is not None:
            self.stem.add(norm_layer(embed_channels), name="norm")
        if act_layer is not None:
            self.stem.add(act_layer(), name="act")
        self.compute_dtype = torch.bfloat16
        self.fuse_lookup = True        # A/B switch for the fused stem kernel (tests compare both paths)
        self.block_search = True       # A/B switch: block-occupancy neighbour search over the z-order serialization

    @torch.no_grad()
    def forward(self, point):
        cd = self.compute_dtype
        conv = self.stem.conv
        sp = point.sparse_conv_feat
        bn = getattr(self.stem, "norm", None)
        feat_in = point.feat.float()
        act = getattr(self.stem, "act", None)
        gelu = isinstance(act, nn.GELU) and act.approximate == "none"
        if (self.fuse_lookup and conv.in_channels == 4 and conv.out_channels == 32 and conv.kernel_size in (3, 5)
                and (bn is None or (isinstance(bn, nn.BatchNorm1d) and not bn.training)) and (act is None or gelu)):
            # neighbour search inside the conv (no (N, 125) table), folded BatchNorm, GELU and the bf16 copy in one kernel
            s, t = (None, None) if bn is None else _cached(
                self, "bn", [bn.weight, bn.bias, bn.running_mean, bn.running_var], lambda: _bn_scale_shift(bn))
            w, b = conv._weights(torch.float32, s, t)
            z = None
            zr = point.get("_z_row") if self.block_search else None
            if zr is not None and "serialized_code" in point and point.serialized_code.shape[1] == feat_in.shape[0]:
                # the z-order serialization is already there: 4^3-block occupancy search instead of 125 hash probes
                z = (point.serialized_code[zr], point.serialized_order[zr], int(point.serialized_depth))
            y, ycd = ops.subm_stem_fused(sp.indices, feat_in, w, b, conv.kernel_size, gelu=gelu, want_f32=True,
                                         want_bf16=cd == torch.bfloat16, z=z)
            point.feat = y
            point["_feat_cd"] = ycd if ycd is not None else y
            point.sparse_conv_feat = sp.replace_feature(point["_feat_cd"])
            return point
        if isinstance(bn, nn.BatchNorm1d) and not bn.training:
            s, t = _cached(self, "bn", [bn.weight, bn.bias, bn.running_mean, bn.running_var],
                           lambda: _bn_scale_shift(bn))
            y = conv.conv(feat_in, sp, scale=s, shift=t)
        else:
            y = conv.conv(feat_in, sp)
            if bn is not None:
                y = bn(y)
        ycd = None
        if gelu and y.shape[1] % 4 == 0:
            y, ycd = ops.affine_act(y, None, None, "gelu", want_f32=True, want_bf16=cd == torch.bfloat16)
        elif act is not None:
            y = act(y)
        point.feat = y
        point["_feat_cd"] = ycd if ycd is not None else (y if cd == torch.float32 else y.to(cd))
        point.sparse_conv_feat = sp.replace_feature(point["_feat_cd"])
        return point


class PointTransformerV3(PointModule):
    """:519-714 (PT-v3m1).  Same constructor and state_dict as the reference.
    `compute_dtype`: torch.bfloat16 (default, = the reference under autocast) or torch.float32."""

    def __init__(self, in_channels=6, order=("z", "z-trans"), stride=(2, 2, 2, 2), enc_depths=(2, 2, 2, 6, 2),
                 enc_channels=(32, 64, 128, 256, 512), enc_num_head=(2, 4, 8, 16, 32),
                 enc_patch_size=(48, 48, 48, 48, 48), dec_depths=(2, 2, 2, 2), dec_channels=(64, 64, 128, 256),
                 dec_num_head=(4, 4, 8, 16), dec_patch_size=(48, 48, 48, 48), mlp_ratio=4, qkv_bias=True,
                 qk_scale=None, attn_drop=0.0, proj_drop=0.0, drop_path=0.3, pre_norm=True, shuffle_orders=True,
                 enable_rpe=False, enable_flash=True, upcast_attention=False, upcast_softmax=False, enc_mode=False,
                 pdnorm_bn=False, pdnorm_ln=False, pdnorm_decouple=True, pdnorm_adaptive=False, pdnorm_affine=True,
                 pdnorm_conditions=("ScanNet", "S3DIS", "Structured3D"), compute_dtype=torch.bfloat16):
        super().__init__()
        from functools import partial
        self.num_stages = len(enc_depths)
        self.order = [order] if isinstance(order, str) else list(order)
        self.enc_mode = enc_mode
        self.shuffle_orders = shuffle_orders
        assert self.num_stages == len(stride) + 1
        assert self.num_stages == len(enc_depths) == len(enc_channels) == len(enc_num_head) == len(enc_patch_size)
        assert self.enc_mode or self.num_stages == len(dec_depths) + 1
        assert self.enc_mode or self.num_stages == len(dec_channels) + 1
        assert self.enc_mode or self.num_stages == len(dec_num_head) + 1
        assert self.enc_mode or self.num_stages == len(dec_patch_size) + 1
        if pdnorm_bn or pdnorm_ln:
            raise NotImplementedError("PDNorm is disabled in every target config (SURVEY.md section 2 row 11)")
        bn_layer = partial(nn.BatchNorm1d, eps=1e-3, momentum=0.01)
        ln_layer = nn.LayerNorm
        act_layer = nn.GELU

        self.embedding = Embedding(in_channels=in_channels, embed_channels=enc_channels[0], norm_layer=bn_layer,
                                   act_layer=act_layer)
        common = dict(mlp_ratio=mlp_ratio, qkv_bias=qkv_bias, qk_scale=qk_scale, attn_drop=attn_drop,
                      proj_drop=proj_drop, norm_layer=ln_layer, act_layer=act_layer, pre_norm=pre_norm,
                      enable_rpe=enable_rpe, enable_flash=enable_flash, upcast_attention=upcast_attention,
                      upcast_softmax=upcast_softmax)
        enc_dp = [x.item() for x in torch.linspace(0, drop_path, sum(enc_depths))]                  # :603-610
        self.enc = PointSequential()
        for s in range(self.num_stages):
            enc_dp_ = enc_dp[sum(enc_depths[:s]):sum(enc_depths[:s + 1])]
            enc = PointSequential()
            if s > 0:
                enc.add(SerializedPooling(in_channels=enc_channels[s - 1], out_channels=enc_channels[s],
                                          stride=stride[s - 1], norm_layer=bn_layer, act_layer=act_layer,
                                          shuffle_orders=shuffle_orders), name="down")
            for i in range(enc_depths[s]):
                enc.add(Block(channels=enc_channels[s], num_heads=enc_num_head[s], patch_size=enc_patch_size[s],
                              order_index=i % len(self.order), cpe_indice_key=f"stage{s}", drop_path=enc_dp_[i],
                              **common), name=f"block{i}")
            if len(enc) != 0:
                self.enc.add(module=enc, name=f"enc{s}")
        if not self.enc_mode:
            self.dec = PointSequential()
            dec_channels = list(dec_channels) + [enc_channels[-1]]
            dec_dp = [x.item() for x in torch.linspace(0, drop_path, sum(dec_depths))]              # :652-661
            for s in reversed(range(self.num_stages - 1)):
                dec_dp_ = dec_dp[sum(dec_depths[:s]):sum(dec_depths[:s + 1])]
                dec_dp_.reverse()
                dec = PointSequential()
                dec.add(SerializedUnpooling(in_channels=dec_channels[s + 1], skip_channels=enc_channels[s],
                                            out_channels=dec_channels[s], norm_layer=bn_layer, act_layer=act_layer),
                        name="up")
                for i in range(dec_depths[s]):
                    dec.add(Block(channels=dec_channels[s], num_heads=dec_num_head[s], patch_size=dec_patch_size[s],
                                  order_index=i % len(self.order), cpe_indice_key=f"stage{s}",
                                  drop_path=dec_dp_[i], **common), name=f"block{i}")
                self.dec.add(module=dec, name=f"dec{s}")
        self.plan_ahead = True
        self.plan_overlap = True       # A/B switch: stem + first stage enqueued before the plan phase
        self.set_compute_dtype(compute_dtype)

    def set_compute_dtype(self, dtype):
        assert dtype in (torch.bfloat16, torch.float32)
        self.compute_dtype = dtype
        for m in self.modules():
            if hasattr(m, "compute_dtype"):
                m.compute_dtype = dtype
        return self

    def set_algos(self, conv=0, attn=0):
        """Kernel selection for A/B tests: 0 auto, 1 CUDA-core reference kernels, 2 tensor-core kernels."""
        for m in self.modules():
            if isinstance(m, Block):
                m.conv_algo = conv
            if isinstance(m, SerializedAttention):
                m.algo = attn
        return self

    def forward(self, data_dict):
        if self.training:                     # autograd path over the same modules
            from . import train
            return train.backbone_forward(self, data_dict)
        return self._forward_eval(data_dict)

    @torch.no_grad()
    def _forward_eval(self, data_dict):
        point = Point(data_dict)
        point.serialization(order=self.order, shuffle_orders=self.shuffle_orders)
        point.sparsify()
        stages = list(self.enc._modules.values())
        first_done = False
        if self.plan_ahead:
            # Integer front ends of every pooling level (same order of torch.randperm draws as the reference:
            # serialization, then pooling 1, 2, ...; see SerializedPooling.plan).  They run on a side stream while the
            # main stream already works on the stem and the first encoder stage (which need no plan): the four host
            # reads of the pooled row counts synchronise the side stream only.
            main = torch.cuda.current_stream()
            side = self.__dict__.get("_plan_stream")
            if side is None or side.device != main.device:
                side = self.__dict__["_plan_stream"] = torch.cuda.Stream(device=main.device)
            ready = torch.cuda.Event()
            ready.record(main)                                   # codes / orders / grid / batch are final here
            code, order = point.serialized_code, point.serialized_order
            grid, batch, depth, n_scenes = point.grid_coord, point.batch, point.serialized_depth, len(point.offset)
            z_row = point.get("_z_row")
            if self.plan_overlap and "down" not in stages[0]._modules:
                point = stages[0](self.embedding(point))
                first_done = True
            side.wait_event(ready)
            plans = []
            with torch.cuda.stream(side):
                for stage in stages:
                    down = stage._modules.get("down")
                    if down is None:
                        continue
                    pd = down.pooling_depth(depth)
                    plans.append(down.plan(code, order, grid, batch, pd, n_scenes, z_row))
                    code, order, grid, batch, z_row = (plans[-1][k] for k in ("code", "order", "grid_coord", "batch", "z_row"))
                    depth -= pd
            main.wait_stream(side)
            for plan in plans:                                   # allocated on the side stream, consumed on the main one
                for v in list(plan.values()) + list(plan["pf"].values()):
                    if torch.is_tensor(v):
                        v.record_stream(main)
            point["_pool_plans"] = plans
        if not first_done:
            point = stages[0](self.embedding(point))
        for stage in stages[1:]:
            point = stage(point)
        if not self.enc_mode:
            point = self.dec(point)
        return point


class OffsetKeypointPTv3(nn.Module):
    """offset_keypoint_ptv3.py:7-107 (inference branch): backbone -> Linear/BN/ReLU/Dropout/Linear -> (N,K,4),
    sigmoid on the mask channel.  The head runs in fp32 (offsets need ~1e-3 absolute accuracy)."""

    def __init__(self, backbone_conf, num_keypoints=6, hidden_dim=256, backbone=None):
        super().__init__()
        in_channels = backbone_conf["dec_channels"][0] if "dec_channels" in backbone_conf else 64
        if backbone is None:
            conf = dict(backbone_conf)
            conf.pop("type", None)
            backbone = PointTransformerV3(**conf)
        self.backbone = backbone
        self.num_keypoints = num_keypoints
        self.head = nn.Sequential(nn.Linear(in_channels, hidden_dim), nn.BatchNorm1d(hidden_dim),
                                  nn.ReLU(inplace=True), nn.Linear(hidden_dim, num_keypoints * 4))

    def forward(self, data_dict):
        if self.training:
            from . import train
            return train.offset_keypoint_forward(self, data_dict)
        return self._forward_eval(data_dict)

    @torch.no_grad()
    def _forward_eval(self, data_dict):
        point = self.backbone(data_dict)
        feat = point.feat if isinstance(point, Point) else point
        lin, bn = self.head[0], self.head[1]

        def build():
            s, t = _bn_scale_shift(bn)
            return (lin.weight.float() * s[:, None]).contiguous(), (lin.bias.float() * s + t).contiguous()
        w, b = _cached(self, "head0", [lin.weight, lin.bias, bn.weight, bn.bias, bn.running_mean, bn.running_var],
                       build)
        # fp32 parity mode keeps strict fp32 GEMMs; with a bf16 backbone the features already carry bf16 rounding,
        # so the two head GEMMs use TF32 tensor-core math (fp32 in/out, 10-bit mantissa products, fp32 accumulate)
        tf32 = getattr(self.backbone, "compute_dtype", torch.float32) != torch.float32
        prev = torch.backends.cuda.matmul.allow_tf32
        torch.backends.cuda.matmul.allow_tf32 = bool(tf32)
        try:
            h = torch._addmm_activation(b, feat.float(), w.t())      # GEMM + bias + ReLU epilogue (cuBLASLt), exact
            pred = F.linear(h, self.head[3].weight.float(), self.head[3].bias.float())
        finally:
            torch.backends.cuda.matmul.allow_tf32 = prev
        pred = pred.view(-1, self.num_keypoints, 4)
        out = {}
        if "target" in data_dict:        # validation: the reference returns the loss next to the prediction (:50-86)
            from .train import offset_keypoint_loss
            out["loss"] = offset_keypoint_loss(pred, data_dict["target"])[0]
        pred[..., 3] = torch.sigmoid(pred[..., 3])
        out["pred"] = pred
        return out


class KeypointPTv3(nn.Module):
    """keypoint_ptv3.py:8-98 (inference): per-scene mean pool (torch_scatter.scatter_mean, :44) -> 3-layer MLP."""

    def __init__(self, backbone_conf, num_keypoints=6, hidden_dim=256, backbone=None):
        super().__init__()
        in_channels = backbone_conf["dec_channels"][0] if "dec_channels" in backbone_conf else 64
        if backbone is None:
            conf = dict(backbone_conf)
            conf.pop("type", None)
            backbone = PointTransformerV3(**conf)
        self.backbone = backbone
        self.num_keypoints = num_keypoints
        self.reg_head = nn.Sequential(nn.Linear(in_channels, hidden_dim), nn.BatchNorm1d(hidden_dim),
                                      nn.ReLU(inplace=True), nn.Dropout(0.3), nn.Linear(hidden_dim, hidden_dim),
                                      nn.ReLU(inplace=True), nn.Linear(hidden_dim, num_keypoints * 3))

    def forward(self, data_dict):
        if self.training:
            from . import train
            return train.keypoint_forward(self, data_dict)
        return self._forward_eval(data_dict)

    @torch.no_grad()
    def _forward_eval(self, data_dict):
        point = self.backbone(data_dict)
        indptr = F.pad(point.offset, (1, 0))
        g = ops.segment_csr(point.feat.float(), indptr, "mean", long_segments=True)
        return dict(pred=self.reg_head(g).view(-1, self.num_keypoints, 3))


class DefaultSegmentorV2(nn.Module):
    """default.py:41-95 (inference): backbone -> Linear(backbone_out_channels, num_classes)."""

    def __init__(self, num_classes, backbone_out_channels, backbone=None, criteria=None, freeze_backbone=False):
        super().__init__()
        self.seg_head = nn.Linear(backbone_out_channels, num_classes) if num_classes > 0 else nn.Identity()
        if isinstance(backbone, dict):
            backbone = build_backbone(backbone)
        self.backbone = backbone
        if criteria is not None and not callable(criteria):     # the reference's list-of-dicts config (losses/builder.py)
            from .losses import Criteria
            criteria = Criteria(criteria)
        self.criteria = criteria
        self.freeze_backbone = freeze_backbone
        if freeze_backbone:                                  # default.py:58-60
            for p in self.backbone.parameters():
                p.requires_grad = False

    def forward(self, input_dict, return_point=False):
        if self.training:
            from . import train
            return train.segmentor_forward(self, input_dict)
        return self._forward_eval(input_dict, return_point)

    def upcast_logits(self, point):
        """seg_head(cat(feat_0, feat_1[inv_1], feat_2[inv_2][inv_1], ...)) of default.py:70-84 WITHOUT materialising
        the concatenation: the head is linear, so each level's slice of the weight is applied at that level's own
        (2x / 4x / ... fewer) rows and the partial logits are gathered down the pooling_inverse chain:
            acc_L = feat_L W_L^T ;  acc_l = feat_l W_l^T + acc_{l+1}[inv_{l+1}] ;  logits = acc_0 + bias.
        For the PTv3-large lin config this replaces an (N, 1728) fp32 tensor by (N, 22) ones."""
        levels, invs = [point], []
        while "pooling_parent" in levels[-1]:
            invs.append(levels[-1]["pooling_inverse"])
            levels.append(levels[-1]["pooling_parent"])
        w, b = self.seg_head.weight.float(), self.seg_head.bias
        lo = sum(lv.feat.shape[1] for lv in levels)
        assert lo == w.shape[1], "backbone_out_channels must equal the sum of the level widths"
        acc = None
        # fp32 parity mode keeps strict fp32 GEMMs; with a bf16 backbone (the reference runs this Linear under autocast
        # in bf16) the slices use TF32 tensor-core products with fp32 accumulation, like the offset head
        tf32 = getattr(self.backbone, "compute_dtype", torch.float32) != torch.float32
        prev = torch.backends.cuda.matmul.allow_tf32
        torch.backends.cuda.matmul.allow_tf32 = bool(tf32)
        try:
            # training with a bf16 backbone: bf16 operands like torch.autocast gives the reference's nn.Linear, so that
            # the weight-gradient GEMMs of the backward run on the tensor cores too (the TF32 switch above only covers
            # the forward; fp32 SIMT GEMMs cost 1.5 ms per step of the PTv3-large linear probe)
            amp = tf32 and self.training and torch.is_grad_enabled()
            for lv, inv in zip(levels, invs + [None]):           # coarsest first; its columns are the LAST slice
                c = lv.feat.shape[1]
                if amp:
                    part = F.linear(lv.feat.to(torch.bfloat16), w[:, lo - c:lo].to(torch.bfloat16)).float()
                else:
                    part = F.linear(lv.feat.float(), w[:, lo - c:lo])
                lo -= c
                acc = part if acc is None else part + acc
                if inv is not None:
                    acc = acc.index_select(0, inv)
        finally:
            torch.backends.cuda.matmul.allow_tf32 = prev
        if acc.shape[0] != levels[-1].feat.shape[0]:
            raise RuntimeError("upcast_logits: broken pooling chain")
        return (acc + b.float()) if b is not None else acc, levels[-1]

    @torch.no_grad()
    def _forward_eval(self, input_dict, return_point=False):
        point = self.backbone(Point(input_dict))
        if (isinstance(point, Point) and "pooling_parent" in point and isinstance(self.seg_head, nn.Linear)
                and not return_point):
            logits, _ = self.upcast_logits(point)
            return dict(seg_logits=logits)
        while isinstance(point, Point) and "pooling_parent" in point:     # enc_mode backbones (:74-82)
            parent = point.pop("pooling_parent")
            inverse = point.pop("pooling_inverse")
            parent.feat = torch.cat([parent.feat, point.feat[inverse]], dim=-1)
            point = parent
        feat = point.feat if isinstance(point, Point) else point
        out = dict(seg_logits=self.seg_head(feat.float()))
        if return_point:
            out["point"] = point
        return out


def build_backbone(conf):
    """builder.MODELS.build for the two registered PTv3 variants: type "PT-v3m1" (default) or "PT-v3m2"."""
    conf = dict(conf)
    kind = conf.pop("type", "PT-v3m1")
    if kind == "PT-v3m2":
        from .sonata import PointTransformerV3 as M2
        return M2(**conf)
    if kind not in ("PT-v3m1", "PointTransformerV3"):
        raise NotImplementedError(f"backbone type {kind!r}")
    return PointTransformerV3(**conf)


def aggregate_keypoints(coord, pred, offset, mode="argmax", mask_thresh=0.5):
    """tools/infer_offset.py:366-399 per scene, one kernel, no host sync:
    kp[b,k] = coord[i*] + pred[i*,k,:3], i* = argmax_i pred[i,k,3] over the scene (votes in model units);
    mode "weighted" = confidence-weighted mean of the votes above mask_thresh (argmax vote if none)."""
    return ops.keypoint_aggregate(coord, pred, offset, mode=mode, mask_thresh=mask_thresh)
