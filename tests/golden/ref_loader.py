"""Import the UNMODIFIED reference hot path from /root/reference on CPU.

Only usable in the build container (the GPU box has no /root/reference).  Used
by make_golden.py to freeze reference outputs into tests/golden/*.npz and by
the `not gpu` tests that compare the oracle with the live reference.

Four packages the reference imports are absent here (addict, timm,
torch_scatter, spconv).  They are replaced by the thin stand-ins below; the
arithmetic ones delegate to oracle.ops (the restated third-party semantics).
Package __init__ files of pointcept that import the whole model zoo are
bypassed with empty package objects.
"""
import importlib
import importlib.util
import os
import sys
import types

import torch
import torch.nn as nn

REF = os.environ.get("PTV3_REFERENCE", "/root/reference")
_REPO = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def available():
    return os.path.isdir(os.path.join(REF, "pointcept"))


def _install_third_party_standins():
    if _REPO not in sys.path:
        sys.path.insert(0, _REPO)
    from oracle import ops as O

    # ---- addict -----------------------------------------------------------
    if "addict" not in sys.modules:
        m = types.ModuleType("addict")

        class Dict(dict):
            def __init__(self, *a, **k):
                super().__init__()
                for arg in a:
                    if arg:
                        for key, v in (arg.items() if hasattr(arg, "items") else arg):
                            self[key] = v
                for key, v in k.items():
                    self[key] = v

            def __getattr__(self, k):
                try:
                    return self[k]
                except KeyError:
                    raise AttributeError(k)

            def __setattr__(self, k, v):
                self[k] = v

            def __delattr__(self, k):
                del self[k]

        m.Dict = Dict
        sys.modules["addict"] = m

    # ---- timm.layers.DropPath ----------------------------------------------
    if "timm" not in sys.modules:
        timm = types.ModuleType("timm")
        layers = types.ModuleType("timm.layers")

        class DropPath(nn.Module):
            def __init__(self, drop_prob=0.0, scale_by_keep=True):
                super().__init__()
                self.drop_prob, self.scale_by_keep = drop_prob, scale_by_keep

            def forward(self, x):
                if self.drop_prob == 0.0 or not self.training:
                    return x
                keep = 1 - self.drop_prob
                mask = x.new_empty((x.shape[0],) + (1,) * (x.ndim - 1)).bernoulli_(keep)
                if keep > 0 and self.scale_by_keep:
                    mask.div_(keep)
                return x * mask

        layers.DropPath = DropPath
        timm.layers = layers
        sys.modules["timm"] = timm
        sys.modules["timm.layers"] = layers

    # ---- torch_scatter -------------------------------------------------------
    if "torch_scatter" not in sys.modules:
        ts = types.ModuleType("torch_scatter")
        ts.segment_csr = lambda src, indptr, out=None, reduce="sum": O.segment_csr(src, indptr, reduce)
        ts.scatter_mean = lambda src, index, dim=0, out=None, dim_size=None: O.scatter_mean(src, index, dim, dim_size)
        sys.modules["torch_scatter"] = ts

    # ---- spconv.pytorch --------------------------------------------------------
    if "spconv" not in sys.modules:
        sp = types.ModuleType("spconv")
        spt = types.ModuleType("spconv.pytorch")
        spm = types.ModuleType("spconv.pytorch.modules")

        class SparseModule(nn.Module):
            pass

        class SparseConvTensor:
            def __init__(self, features, indices, spatial_shape, batch_size, indice_dict=None):
                self.features, self.indices = features, indices
                self.spatial_shape, self.batch_size = spatial_shape, batch_size
                self.indice_dict = {} if indice_dict is None else indice_dict

            def replace_feature(self, f):
                return SparseConvTensor(f, self.indices, self.spatial_shape, self.batch_size, self.indice_dict)

        class SubMConv3d(SparseModule):
            def __init__(self, in_channels, out_channels, kernel_size, stride=1, padding=0, dilation=1,
                         groups=1, bias=True, indice_key=None, **kw):
                super().__init__()
                k = kernel_size
                self.kernel_size, self.indice_key = k, indice_key
                self.weight = nn.Parameter(torch.empty(out_channels, k, k, k, in_channels))
                self.bias = nn.Parameter(torch.zeros(out_channels)) if bias else None
                nn.init.kaiming_uniform_(self.weight, a=5 ** 0.5)

            def forward(self, x):
                key = (self.indice_key, self.kernel_size)
                if self.indice_key is None or key not in x.indice_dict:
                    nbr = O.subm_neighbor_table(x.indices, self.kernel_size)
                    if self.indice_key is not None:
                        x.indice_dict[key] = nbr
                else:
                    nbr = x.indice_dict[key]
                return x.replace_feature(O.subm_conv(x.features, nbr, self.weight, self.bias))

        class SparseConv3d(SparseModule):
            """k = 2, s = 2 only (SpUNet's down-sampling, spconv_unet_v1m1_base.py:137-151) over oracle/spunet.py."""

            def __init__(self, in_channels, out_channels, kernel_size, stride=1, padding=0, dilation=1, groups=1,
                         bias=True, indice_key=None, **kw):
                super().__init__()
                assert kernel_size == 2 and stride == 2 and padding == 0
                self.indice_key = indice_key
                self.weight = nn.Parameter(torch.empty(out_channels, 2, 2, 2, in_channels))
                self.bias = nn.Parameter(torch.zeros(out_channels)) if bias else None
                nn.init.kaiming_uniform_(self.weight, a=5 ** 0.5)

            def forward(self, x):
                from oracle import spunet as U
                out_idx, cluster, koff = U.down_map(x.indices)
                x.indice_dict[("pairs", self.indice_key)] = (x.indices, cluster, koff, x.spatial_shape)
                f = U.sparse_conv_k2s2(x.features, cluster, koff, out_idx.shape[0], self.weight)
                if self.bias is not None:
                    f = f + self.bias
                # same indice_dict object: the inverse conv finds the pairs under its indice_key like in spconv
                return SparseConvTensor(f, out_idx, [(d + 1) // 2 for d in x.spatial_shape], x.batch_size,
                                        x.indice_dict)

        class SparseInverseConv3d(SparseModule):
            def __init__(self, in_channels, out_channels, kernel_size, indice_key=None, bias=True, **kw):
                super().__init__()
                assert kernel_size == 2
                self.indice_key = indice_key
                self.weight = nn.Parameter(torch.empty(out_channels, 2, 2, 2, in_channels))
                self.bias = nn.Parameter(torch.zeros(out_channels)) if bias else None
                nn.init.kaiming_uniform_(self.weight, a=5 ** 0.5)

            def forward(self, x):
                from oracle import spunet as U
                fine_idx, cluster, koff, shape = x.indice_dict[("pairs", self.indice_key)]
                f = U.sparse_inverse_conv_k2(x.features, cluster, koff, self.weight)
                if self.bias is not None:
                    f = f + self.bias
                return SparseConvTensor(f, fine_idx, shape, x.batch_size, x.indice_dict)

        class SparseSequential(SparseModule):
            def __init__(self, *args, **kwargs):
                super().__init__()
                if len(args) == 1 and isinstance(args[0], dict):
                    for k, m in args[0].items():
                        self.add_module(k, m)
                else:
                    for i, m in enumerate(args):
                        self.add_module(str(i), m)
                for k, m in kwargs.items():
                    self.add_module(k, m)

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

        class Identity(nn.Identity):
            pass

        spt.SparseConv3d, spt.SparseInverseConv3d = SparseConv3d, SparseInverseConv3d
        spt.SparseSequential, spt.Identity = SparseSequential, Identity
        spm.is_spconv_module = lambda m: isinstance(m, SparseModule)
        spm.SparseModule = SparseModule
        spt.SparseModule, spt.SparseConvTensor, spt.SubMConv3d, spt.modules = SparseModule, SparseConvTensor, SubMConv3d, spm
        sp.pytorch = spt
        sys.modules["spconv"], sys.modules["spconv.pytorch"], sys.modules["spconv.pytorch.modules"] = sp, spt, spm


def _stub_package(name, path):
    if name in sys.modules:
        return sys.modules[name]
    m = types.ModuleType(name)
    m.__path__ = [path]
    m.__package__ = name
    sys.modules[name] = m
    return m


_loaded = {}


def load():
    """Returns a namespace with the reference classes/functions of the hot path."""
    if _loaded:
        return types.SimpleNamespace(**_loaded)
    assert available(), "reference tree not present"
    _install_third_party_standins()
    if REF not in sys.path:
        sys.path.insert(0, REF)
    pc = os.path.join(REF, "pointcept")
    _stub_package("pointcept.models", os.path.join(pc, "models"))
    eng = _stub_package("pointcept.engines", os.path.join(pc, "engines"))
    hooks = _stub_package("pointcept.engines.hooks", os.path.join(pc, "engines", "hooks"))
    ppt = _stub_package("pointcept.models.point_prompt_training", os.path.join(pc, "models", "point_prompt_training"))
    _stub_package("pointcept.models.point_transformer_v3", os.path.join(pc, "models", "point_transformer_v3"))
    hooks.HookBase = importlib.import_module("pointcept.engines.hooks.default").HookBase
    eng.hooks = hooks
    ppt.PDNorm = importlib.import_module(
        "pointcept.models.point_prompt_training.prompt_driven_normalization").PDNorm
    ptv3 = importlib.import_module("pointcept.models.point_transformer_v3.point_transformer_v3m1_base")
    try:       # PT-v3m2 ("Sonata" PTv3: GridPooling / LayerScale / Linear stem), registered as "PT-v3m2"
        ptv3m2 = importlib.import_module("pointcept.models.point_transformer_v3.point_transformer_v3m2_sonata")
    except Exception as e:                                 # noqa: BLE001
        print("PT-v3m2 not importable:", repr(e))
        ptv3m2 = None
    okp = importlib.import_module("pointcept.models.offset_keypoint_ptv3")
    kp = importlib.import_module("pointcept.models.keypoint_ptv3")
    # default.py (DefaultSegmentorV2) imports torch_cluster and peft at module level without using them on this path
    for absent, names in (("torch_cluster", ()), ("peft", ("LoraConfig", "get_peft_model"))):
        if absent not in sys.modules:
            stub = types.ModuleType(absent)
            for nm in names:
                setattr(stub, nm, None)
            sys.modules[absent] = stub
    # SpUNet-v1m1 imports torch_geometric.utils.scatter (enc_mode only) and timm's trunc_normal_
    if "torch_geometric" not in sys.modules:
        tg, tgu = types.ModuleType("torch_geometric"), types.ModuleType("torch_geometric.utils")
        tgu.scatter = None
        tg.utils = tgu
        sys.modules["torch_geometric"], sys.modules["torch_geometric.utils"] = tg, tgu
    if not hasattr(sys.modules["timm.layers"], "trunc_normal_"):
        sys.modules["timm.layers"].trunc_normal_ = torch.nn.init.trunc_normal_
    try:
        _stub_package("pointcept.models.sparse_unet", os.path.join(pc, "models", "sparse_unet"))
        spunet = importlib.import_module("pointcept.models.sparse_unet.spconv_unet_v1m1_base")
    except Exception as e:                                 # noqa: BLE001
        print("SpUNet-v1m1 not importable:", repr(e))
        spunet = None
    try:
        seg = importlib.import_module("pointcept.models.default")
    except Exception as e:                                 # noqa: BLE001 -- report, the other fixtures do not need it
        print("pointcept.models.default not importable:", repr(e))
        seg = None
    structure = importlib.import_module("pointcept.models.utils.structure")
    ser = importlib.import_module("pointcept.models.utils.serialization")
    builder = importlib.import_module("pointcept.models.builder")
    # GridSample: load transform.py by path, registered in sys.modules first
    name = "pointcept_ref_transform"
    spec = importlib.util.spec_from_file_location(name, os.path.join(pc, "datasets", "transform.py"))
    tr = importlib.util.module_from_spec(spec)
    sys.modules[name] = tr
    spec.loader.exec_module(tr)
    _loaded.update(ptv3=ptv3, PointTransformerV3=ptv3.PointTransformerV3, SerializedAttention=ptv3.SerializedAttention,
                   OffsetKeypointPTv3=okp.OffsetKeypointPTv3, KeypointPTv3=kp.KeypointPTv3,
                   DefaultSegmentorV2=getattr(seg, "DefaultSegmentorV2", None),
                   SpUNetBase=getattr(spunet, "SpUNetBase", None), ptv3m2=ptv3m2,
                   Point=structure.Point, encode=ser.encode, GridSample=tr.GridSample, builder=builder)
    return types.SimpleNamespace(**_loaded)
