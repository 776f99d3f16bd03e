"""Drop-in module names for the UNMODIFIED reference.

`install()` registers, under the names the reference imports, thin modules whose
bodies are our kernels:

    spconv.pytorch            SubMConv3d, SparseConv3d (k2,s2), SparseInverseConv3d (k2), SparseConvTensor,
                              SparseModule, SparseSequential, Identity
    spconv.pytorch.modules    is_spconv_module
    torch_scatter             segment_csr, scatter_mean
    flash_attn                flash_attn_varlen_qkvpacked_func

and `patch_reference()` swaps `pointcept.models.utils.serialization.encode` (as
imported by structure.py:11) for the CUDA encoder.  After that
`pointcept.models.point_transformer_v3` and the keypoint heads run unchanged on
a B200 (inference; no autograd through the kernels yet).  Real packages are
never shadowed unless force=True.
"""
import importlib.util
import sys
import types

import torch
import torch.nn as nn

from . import ops
from .model import SparseModule, SubMConv3d, is_spconv_module
from .structure import SparseConvTensor


class SparseSequential(SparseModule):
    def __init__(self, *args, **kwargs):
        super().__init__()
        for i, m in enumerate(args):
            self.add_module(str(i), m)
        for k, m in kwargs.items():
            self.add_module(k, m)

    def forward(self, x):
        for m in self._modules.values():
            if is_spconv_module(m):
                x = m(x)
            elif isinstance(x, SparseConvTensor):
                if x.indices.shape[0] != 0:
                    x = x.replace_feature(m(x.features))
            else:
                x = m(x)
        return x


# ---------------------------------------------------------------- torch_scatter
def segment_csr(src, indptr, out=None, reduce="sum"):
    """torch_scatter.segment_csr (call site point_transformer_v3m1_base.py:416-421)."""
    r = ops.segment_csr(src, indptr, reduce)
    if out is not None:
        out.copy_(r)
        return out
    return r


def scatter_mean(src, index, dim=0, out=None, dim_size=None):
    """torch_scatter.scatter_mean for a NON-DECREASING index along dim 0 (call site keypoint_ptv3.py:44 passes
    Point.batch, which is sorted by construction)."""
    if dim not in (0, -src.dim()):
        raise NotImplementedError("scatter_mean: dim 0 only")
    n = int(dim_size) if dim_size is not None else int(index.max().item()) + 1
    offset = ops.batch_to_offset(index, n)
    indptr = torch.nn.functional.pad(offset, (1, 0))
    r = ops.segment_csr(src, indptr, "mean", long_segments=True)
    if out is not None:
        out.copy_(r)
        return out
    return r


# ---------------------------------------------------------------- flash_attn
def flash_attn_varlen_qkvpacked_func(qkv, cu_seqlens, max_seqlen, dropout_p=0.0, softmax_scale=None, causal=False,
                                     **unused):
    """flash_attn.flash_attn_varlen_qkvpacked_func (call site point_transformer_v3m1_base.py:208-215):
    qkv (T,3,H,16) bf16 packed, cu_seqlens int32 (S+1) -> (T,H,16)."""
    if causal or dropout_p != 0.0:
        raise NotImplementedError("flash_attn shim: non-causal, dropout-free inference only")
    t, three, h, d = qkv.shape
    assert three == 3
    cu = cu_seqlens.long()
    starts, lens = cu[:-1], cu[1:] - cu[:-1]
    windows = torch.stack([starts, lens, starts, lens], dim=1).contiguous()
    nwin = torch.full((1,), windows.shape[0], dtype=torch.int64, device=qkv.device)
    out = ops.attn_window(qkv.reshape(t, 3 * h * d), None, windows, nwin, h, int(max_seqlen), softmax_scale)
    return out.view(t, h, d)


def _module(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    return m


def _absent(name):
    if name in sys.modules:
        return False
    try:
        return importlib.util.find_spec(name) is None
    except (ImportError, ValueError):
        return True


def install(force=False):
    """Register the stand-in modules.  Returns the list of names that were installed."""
    done = []
    if force or _absent("spconv"):
        from . import spunet        # strided / inverse convs of SpUNet-v1m1 (sparse_unet/spconv_unet_v1m1_base.py:23)
        mods = _module("spconv.pytorch.modules", is_spconv_module=is_spconv_module, SparseModule=SparseModule,
                       SparseSequential=SparseSequential)
        pt = _module("spconv.pytorch", SubMConv3d=SubMConv3d, SparseConvTensor=SparseConvTensor,
                     SparseModule=SparseModule, SparseSequential=spunet.SparseSequential, modules=mods,
                     SparseConv3d=spunet.SparseConv3d, SparseInverseConv3d=spunet.SparseInverseConv3d,
                     Identity=nn.Identity)
        root = _module("spconv", pytorch=pt, __path__=[])
        pt.__path__ = []
        sys.modules.update({"spconv": root, "spconv.pytorch": pt, "spconv.pytorch.modules": mods})
        done.append("spconv")
    if force or _absent("torch_scatter"):
        sys.modules["torch_scatter"] = _module("torch_scatter", segment_csr=segment_csr, scatter_mean=scatter_mean)
        done.append("torch_scatter")
    if force or _absent("flash_attn"):
        sys.modules["flash_attn"] = _module("flash_attn",
                                            flash_attn_varlen_qkvpacked_func=flash_attn_varlen_qkvpacked_func)
        done.append("flash_attn")
    return done


def patch_reference():
    """Point the reference's `encode` (structure.py:11,89) at the CUDA encoder, if pointcept is imported."""
    patched = []
    for name in ("pointcept.models.utils.structure", "pointcept.models.utils.serialization",
                 "pointcept.models.utils.serialization.default"):
        mod = sys.modules.get(name)
        if mod is not None and hasattr(mod, "encode"):
            mod.encode = ops.encode
            patched.append(name)
    return patched
