"""SpUNet-v1m1 over the sm_100a kernels (BASELINE configs[3]); inference.

Same module tree and state_dict names as pointcept/models/sparse_unet/spconv_unet_v1m1_base.py (BasicBlock :23-86,
SpUNetBase :89-281) so the reference's checkpoints load unchanged.  What runs where:

  * SubMConv3d k = 3 / 5 / 1: the PTv3 conv kernels (tcgen05 for bf16 with C_in = 32 or a multiple of 64 and
    C_out in {32..512}, the CUDA-core kernels otherwise, e.g. the 96 / 224-channel decoder);
  * spconv.SparseConv3d(kernel 2, stride 2): the coarse sites are the run heads of `code >> 3` over a z-order
    serialization (ptv3_encode / ptv3_argsort_rows once, then ptv3_pool_front per level -- no further sorts), the
    convolution is the same gather-GEMM kernel over an (M, 8) table from ptv3_sparse_pair_maps;
  * spconv.SparseInverseConv3d: the same kernel over the (N, 8) inverse table cached under the shared indice_key;
  * eval BatchNorm is folded into the conv weights; ReLU / residual add / concatenation are torch elementwise ops
    in this first version.
"""
from collections import OrderedDict
from functools import partial

import torch
import torch.nn as nn

from . import ops
from .model import SparseModule, SubMConv3d, _bn_scale_shift, _cached, pad_out_channels
from .structure import SparseConvTensor


class SparseSequential(SparseModule):
    """spconv.SparseSequential: sparse modules get the tensor, torch modules its features."""

    def __init__(self, *args, **kwargs):
        super().__init__()
        if len(args) == 1 and isinstance(args[0], OrderedDict):
            for k, m in args[0].items():
                self.add_module(k, m)
        else:
            for i, m in enumerate(args):
                self.add_module(str(i), m)
        for k, m in kwargs.items():
            self.add_module(k, m)

    def __getitem__(self, i):
        return list(self._modules.values())[i]

    def __len__(self):
        return len(self._modules)

    def forward(self, x):
        for m in self._modules.values():
            if isinstance(m, SparseModule):
                x = m(x)
            elif isinstance(x, SparseConvTensor):
                if x.indices.shape[0] != 0:
                    x = x.replace_feature(m(x.features))
            else:
                x = m(x)
        return x


class _PairConv(SparseModule):
    """Common part of SparseConv3d(k2,s2) / SparseInverseConv3d(k2): weight (out, 2,2,2, in), gather-GEMM over an
    8-column table."""

    def __init__(self, in_channels, out_channels, bias, indice_key):
        super().__init__()
        self.in_channels, self.out_channels, self.indice_key = in_channels, out_channels, indice_key
        self.weight = nn.Parameter(torch.empty(out_channels, 2, 2, 2, in_channels))
        self.bias = nn.Parameter(torch.zeros(out_channels)) if bias else None
        nn.init.kaiming_uniform_(self.weight, a=5 ** 0.5)

    def _weights(self, dtype, scale, shift, cin_p, cout_p):
        def build():
            w = self.weight.float()
            b = self.bias.float() if self.bias is not None else None
            if scale is not None:
                w = w * scale.view(-1, 1, 1, 1, 1)
                b = shift if b is None else b * scale + shift
            w = w.reshape(self.out_channels, 8, self.in_channels).permute(1, 0, 2)
            if (cin_p, cout_p) != (self.in_channels, self.out_channels):
                w = torch.nn.functional.pad(w, (0, cin_p - self.in_channels, 0, cout_p - self.out_channels))
                if b is not None:
                    b = torch.nn.functional.pad(b, (0, cout_p - self.out_channels))
            return w.contiguous().to(dtype), (b.contiguous() if b is not None else None)
        return _cached(self, ("w", dtype, scale is not None, cin_p, cout_p), [self.weight, self.bias, scale, shift],
                       build)

    def gemm(self, feat, nbr, scale=None, shift=None):
        """bf16: channel counts the tcgen05 kernel does not tile (96) run zero-padded to the next width it does."""
        cin_p, cout_p = self.in_channels, self.out_channels
        if feat.dtype == torch.bfloat16:
            cin_p, cout_p = feat.shape[1], pad_out_channels(self.out_channels)
        w, b = self._weights(feat.dtype, scale, shift, cin_p, cout_p)
        if feat.dtype == torch.bfloat16 and ops.subm_conv_tc_supported(cin_p, cout_p, 8):
            packed = _cached(self, ("packed", scale is not None, cin_p, cout_p),
                             [self.weight, self.bias, scale, shift], lambda: ops.subm_pack_weight(w))
            return ops.subm_conv_packed(feat, packed, b, nbr, cout_p)
        return ops.subm_conv(feat, w, b, nbr)


def serialize_sites(indices):
    """z-order code + order of the sites of a SparseConvTensor built by somebody else (the unmodified reference's
    SpUNetBase.forward, :240-255): what SparseConv3d needs to find its coarse sites without a hash table."""
    batch = indices[:, 0].long().contiguous()
    grid = indices[:, 1:].long().contiguous()
    depth = max(int(int(ops.grid_max(grid).max().item()) + 1).bit_length(), 1)
    n_scenes = int(batch[-1].item()) + 1 if batch.numel() else 1
    code = ops.encode_orders(grid, batch, depth, ["z"])
    order, _ = ops.argsort_rows(code, key_bits=depth * 3 + max(1, (n_scenes - 1).bit_length()), want_inverse=False)
    return code, order, grid, batch


class SparseConv3d(_PairConv):
    def __init__(self, in_channels, out_channels, kernel_size, stride=1, padding=0, dilation=1, groups=1, bias=True,
                 indice_key=None):
        if kernel_size != 2 or stride != 2 or padding != 0 or dilation != 1 or groups != 1:
            raise NotImplementedError("SparseConv3d: only kernel 2 / stride 2 (SpUNet's down-sampling) is built")
        super().__init__(in_channels, out_channels, bias, indice_key)

    def conv(self, x, scale=None, shift=None):
        """x must carry the z-order serialization of its sites in x.indice_dict['_serial'] (SpUNetBase.forward)."""
        if "_serial" not in x.indice_dict:
            x.indice_dict["_serial"] = serialize_sites(x.indices)
        code, order, grid, batch = x.indice_dict["_serial"]
        pf = ops.pool_front(code, order, 1)
        new_grid, new_batch, indices4 = ops.pool_gather_ints(grid, batch, pf["head"], 1)
        fwd, inv = ops.sparse_pair_maps(grid, pf["cluster"], pf["n_pooled"])
        x.indice_dict[("pairs", self.indice_key)] = (inv, x)
        feat = self.gemm(x.features, fwd, scale, shift)
        out = SparseConvTensor(feat, indices4, [(d + 1) // 2 for d in x.spatial_shape], x.batch_size, dict(x.indice_dict))
        out.indice_dict["_serial"] = (pf["code"], pf["order"], new_grid, new_batch)
        return out

    def forward(self, x):
        return self.conv(x)


class SparseInverseConv3d(_PairConv):
    def __init__(self, in_channels, out_channels, kernel_size, indice_key=None, bias=True):
        if kernel_size != 2:
            raise NotImplementedError("SparseInverseConv3d: only kernel 2 is built")
        super().__init__(in_channels, out_channels, bias, indice_key)

    def conv(self, x, scale=None, shift=None):
        inv, fine = x.indice_dict[("pairs", self.indice_key)]
        return fine.replace_feature(self.gemm(x.features, inv, scale, shift))

    def forward(self, x):
        return self.conv(x)


def _fold(bn):
    return _cached(bn, "fold", [bn.weight, bn.bias, bn.running_mean, bn.running_var], lambda: _bn_scale_shift(bn))


class BasicBlock(SparseModule):
    """:23-86."""
    expansion = 1

    def __init__(self, in_channels, embed_channels, stride=1, norm_fn=None, indice_key=None, bias=False):
        super().__init__()
        assert norm_fn is not None
        if in_channels == embed_channels:
            self.proj = SparseSequential(nn.Identity())
        else:
            self.proj = SparseSequential(SubMConv3d(in_channels, embed_channels, kernel_size=1, bias=False),
                                         norm_fn(embed_channels))
        self.conv1 = SubMConv3d(in_channels, embed_channels, kernel_size=3, stride=stride, padding=1, bias=bias,
                                indice_key=indice_key)
        self.bn1 = norm_fn(embed_channels)
        self.relu = nn.ReLU()
        self.conv2 = SubMConv3d(embed_channels, embed_channels, kernel_size=3, stride=stride, padding=1, bias=bias,
                                indice_key=indice_key)
        self.bn2 = norm_fn(embed_channels)
        self.stride = stride

    @torch.no_grad()
    def forward(self, x):
        if self.training:
            raise NotImplementedError("ptv3_b200 SpUNet: inference only (call .eval())")
        f = x.features
        if isinstance(self.proj[0], nn.Identity):
            res = f
        else:
            res = self.proj[0].conv(f, x, *_fold(self.proj[1]))
        y = torch.relu_(self.conv1.conv(f, x, *_fold(self.bn1)))
        y = self.conv2.conv(y, x, *_fold(self.bn2))
        return x.replace_feature(torch.relu_(y.add_(res)))


class SpUNetBase(nn.Module):
    """:89-281 ("SpUNet-v1m1").  `compute_dtype`: torch.bfloat16 (default) or torch.float32 (1e-4 parity)."""

    def __init__(self, in_channels, num_classes, base_channels=32, channels=(32, 64, 128, 256, 256, 128, 96, 96),
                 layers=(2, 3, 4, 6, 2, 2, 2, 2), enc_mode=False, compute_dtype=torch.bfloat16):
        super().__init__()
        assert len(layers) % 2 == 0 and len(layers) == len(channels)
        if enc_mode:
            raise NotImplementedError("SpUNetBase: enc_mode is not used by the target config")
        self.in_channels, self.num_classes, self.base_channels = in_channels, num_classes, base_channels
        self.channels, self.layers, self.num_stages, self.enc_mode = channels, layers, len(layers) // 2, enc_mode
        self.compute_dtype = compute_dtype
        norm_fn = partial(nn.BatchNorm1d, eps=1e-3, momentum=0.01)
        self.conv_input = SparseSequential(
            SubMConv3d(in_channels, base_channels, kernel_size=5, padding=1, bias=False, indice_key="stem"),
            norm_fn(base_channels), nn.ReLU())
        enc_channels, dec_channels = base_channels, channels[-1]
        self.down, self.up, self.enc, self.dec = nn.ModuleList(), nn.ModuleList(), nn.ModuleList(), nn.ModuleList()
        for s in range(self.num_stages):
            self.down.append(SparseSequential(
                SparseConv3d(enc_channels, channels[s], kernel_size=2, stride=2, bias=False, indice_key=f"spconv{s + 1}"),
                norm_fn(channels[s]), nn.ReLU()))
            self.enc.append(SparseSequential(OrderedDict(
                [(f"block{i}", BasicBlock(channels[s], channels[s], norm_fn=norm_fn, indice_key=f"subm{s + 1}"))
                 for i in range(layers[s])])))
            self.up.append(SparseSequential(
                SparseInverseConv3d(channels[len(channels) - s - 2], dec_channels, kernel_size=2, bias=False,
                                    indice_key=f"spconv{s + 1}"),
                norm_fn(dec_channels), nn.ReLU()))
            self.dec.append(SparseSequential(OrderedDict(
                [(f"block{i}", BasicBlock(dec_channels + enc_channels if i == 0 else dec_channels, dec_channels,
                                          norm_fn=norm_fn, indice_key=f"subm{s}"))
                 for i in range(layers[len(channels) - s - 1])])))
            enc_channels = channels[s]
            dec_channels = channels[len(channels) - s - 2]
        self.final = SubMConv3d(channels[-1], num_classes, kernel_size=1, padding=1, bias=True) \
            if num_classes > 0 else nn.Identity()
        for mod in self.modules():
            if isinstance(mod, SubMConv3d):
                mod.pad_channels = True

    @torch.no_grad()
    def forward(self, input_dict):
        if self.training:
            raise NotImplementedError("ptv3_b200 SpUNet: inference only (call .eval())")
        from .structure import Point
        cd = self.compute_dtype
        point = Point(grid_coord=input_dict["grid_coord"], feat=input_dict["feat"], offset=input_dict["offset"])
        point.serialization(order="z", shuffle_orders=False)        # one encode + one radix argsort for all levels
        point.sparsify()
        assert point.serialized_depth >= self.num_stages, "scene too small for the number of stride-2 stages"
        x = point.sparse_conv_feat
        x.indice_dict["_serial"] = (point.serialized_code, point.serialized_order, point.grid_coord.long(), point.batch)
        stem, bn = self.conv_input[0], self.conv_input[1]
        x = x.replace_feature(torch.relu_(stem.conv(x.features.float(), x, *_fold(bn))).to(cd))
        skips = [x]
        for s in range(self.num_stages):
            x = self.down[s][0].conv(x, *_fold(self.down[s][1]))
            x = x.replace_feature(torch.relu_(x.features))
            x = self.enc[s](x)
            skips.append(x)
        x = skips.pop(-1)
        for s in reversed(range(self.num_stages)):
            x = self.up[s][0].conv(x, *_fold(self.up[s][1]))
            skip = skips.pop(-1)
            c_up, c_skip = self.up[s][0].out_channels, self.down[s][0].in_channels     # logical widths (no padding)
            x = x.replace_feature(torch.cat((torch.relu_(x.features)[:, :c_up], skip.features[:, :c_skip]), dim=1))
            x = self.dec[s](x)
        if isinstance(self.final, nn.Identity):
            return x.features[:, :self.channels[-1]]
        return self.final.conv(x.features, x)[:, :self.num_classes].float()
