"""Boundary proof on the B200: the third-party calls of one forward of the UNMODIFIED reference
(tests/golden/make_shim_trace.py froze their inputs and outputs at the reference's own call sites: first call of each
kind per width) replayed through the drop-in entry points of ptv3_b200.shims -- spconv.pytorch.SubMConv3d.forward,
torch_scatter.segment_csr, flash_attn.flash_attn_varlen_qkvpacked_func, serialization.encode -- with the reference's
argument layouts (point_transformer_v3m1_base.py:208-215, :278-284, :416-421; structure.py:89)."""
import json
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def rel_l2(a, b):
    a, b = a.detach().float().cpu(), b.detach().float().cpu()
    return ((a - b).norm() / b.norm().clamp_min(1e-30)).item()


@pytest.fixture(scope="module")
def trace(golden):
    return golden("shim_trace")


def _tags(tr, kind):
    return [t for t in tr["kept_tags"].tolist() if t.startswith(kind + "__")]


def test_encode_calls(trace):
    from ptv3_b200 import ops
    tags = _tags(trace, "encode")
    assert len(tags) == 4
    for t in tags:
        got = ops.encode(torch.from_numpy(trace[t + "__grid_coord"]).cuda(), torch.from_numpy(trace[t + "__batch"]).cuda(),
                         int(trace[t + "__depth"]), str(trace[t + "__order"]))
        assert np.array_equal(got.cpu().numpy(), trace[t + "__out"]), t


def test_segment_csr_calls(trace):
    from ptv3_b200 import shims
    tags = _tags(trace, "segment_csr")
    assert len(tags) >= 5                                     # max at 4 widths + the coordinate mean
    for t in tags:
        got = shims.segment_csr(torch.from_numpy(trace[t + "__src"]).cuda(), torch.from_numpy(trace[t + "__indptr"]).cuda(),
                                reduce=str(trace[t + "__reduce"]))
        assert rel_l2(got, torch.from_numpy(trace[t + "__out"])) < 1e-6, t


@pytest.mark.parametrize("autocast", [False, True])
def test_submconv_calls(trace, autocast):
    """Our SubMConv3d module as the reference constructs it (:278-284, :499-506), weights regenerated from the parameter
    name.  fp32 call = the CUDA-core parity kernel; under torch.autocast(bf16) -- how the reference trainer / tester runs
    the model -- the same call lands on the tcgen05 implicit-GEMM kernel."""
    from conftest import GOLDEN
    from ptv3_b200 import shims, synth
    from ptv3_b200.structure import SparseConvTensor
    with open(os.path.join(GOLDEN, "param_shapes.json")) as f:
        shapes = json.load(f)
    tags = _tags(trace, "SubMConv3d.forward")
    assert len(tags) == 6                                     # stem + 5 widths
    for t in tags:
        name = "backbone." + str(trace[t + "__module"]) if not str(trace[t + "__module"]).startswith("backbone.") \
            else str(trace[t + "__module"])
        k = int(trace[t + "__kernel_size"])
        has_bias = bool(trace[t + "__has_bias"])
        sub = {name + ".weight": shapes[name + ".weight"]}
        if has_bias:
            sub[name + ".bias"] = shapes[name + ".bias"]
        W = synth.make_weights(sub, seed=0)
        cout, cin = shapes[name + ".weight"][0], shapes[name + ".weight"][-1]
        conv = shims.SubMConv3d(cin, cout, kernel_size=k, bias=has_bias, indice_key="t").cuda()
        conv.weight.data.copy_(torch.from_numpy(W[name + ".weight"]))
        if has_bias:
            conv.bias.data.copy_(torch.from_numpy(W[name + ".bias"]))
        feats = torch.from_numpy(trace[t + "__features"]).cuda()
        idx = torch.from_numpy(trace[t + "__indices"]).cuda().int()
        x = SparseConvTensor(feats, idx, [4096] * 3, 1)
        ref = torch.from_numpy(trace[t + "__out"])
        with torch.no_grad():
            if autocast:
                with torch.autocast("cuda", dtype=torch.bfloat16):
                    y = conv(x)
                assert y.features.dtype == torch.bfloat16 or cin < 32      # the 4-channel stem stays on the fp32 kernel
                assert rel_l2(y.features, ref) < 1e-2, (t, rel_l2(y.features, ref))
            else:
                y = conv(x)
                assert rel_l2(y.features, ref) < 1e-5, (t, rel_l2(y.features, ref))
        assert ("t", k) in x.indice_dict and y.indice_dict is x.indice_dict      # spconv-style indice_key cache


def test_flash_attn_calls(trace):
    from ptv3_b200 import ops, shims
    tags = _tags(trace, "flash_attn_varlen_qkvpacked_func")
    assert len(tags) == 5                                     # one per head count
    for t in tags:
        qkv = torch.from_numpy(trace[t + "__qkv"]).cuda().bfloat16()
        out = shims.flash_attn_varlen_qkvpacked_func(qkv, torch.from_numpy(trace[t + "__cu_seqlens"]).cuda(),
                                                     max_seqlen=int(trace[t + "__max_seqlen"]), dropout_p=0.0,
                                                     softmax_scale=float(trace[t + "__softmax_scale"]))
        assert out.dtype == torch.bfloat16 and out.shape == trace[t + "__out"].shape
        assert rel_l2(out, torch.from_numpy(trace[t + "__out"])) < 6e-3, t
    assert ops.device_error() == 0
