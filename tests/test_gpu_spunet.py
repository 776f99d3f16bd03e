"""SpUNet-v1m1 (BASELINE configs[3]) on the GPU against the reference class's frozen output and
the oracle's strided / inverse convolutions."""
import json
import os

import numpy as np
import pytest
import torch

from oracle import ops as O
from oracle import spunet as U

pytestmark = pytest.mark.gpu


def rel_l2(a, b):
    a, b = a.detach().float().cpu(), b.detach().float().cpu()
    return ((a - b).norm() / b.norm().clamp_min(1e-30)).item()


def _weights():
    from conftest import GOLDEN
    from ptv3_b200 import synth
    with open(os.path.join(GOLDEN, "param_shapes_spunet.json")) as f:
        shapes = json.load(f)
    return {k: torch.from_numpy(v) for k, v in synth.make_weights(shapes, seed=0).items()}


def _batch_indices(g):
    off = torch.from_numpy(g["offset"])
    batch = torch.repeat_interleave(torch.arange(len(off)), torch.diff(off, prepend=torch.zeros(1, dtype=off.dtype)))
    return batch, torch.cat([batch[:, None], torch.from_numpy(g["grid_coord"])], 1)


def test_strided_and_inverse_conv_vs_oracle(golden):
    """SparseConv3d(k2,s2) + SparseInverseConv3d sharing an indice_key: coarse sites (as a set), features per coarse
    site and the up-sampled features against oracle/spunet.py, fp32."""
    from ptv3_b200 import spunet as S
    from ptv3_b200.structure import Point
    g = golden("spunet")
    batch, idx = _batch_indices(g)
    n = idx.shape[0]
    gen = torch.Generator().manual_seed(3)
    cin, cmid = 32, 64
    feat = torch.randn(n, cin, generator=gen)
    down = S.SparseConv3d(cin, cmid, 2, stride=2, bias=False, indice_key="spconv1")
    up = S.SparseInverseConv3d(cmid, cin, 2, indice_key="spconv1", bias=False)
    out_idx, cluster, koff = U.down_map(idx)
    ref_c = U.sparse_conv_k2s2(feat, cluster, koff, out_idx.shape[0], down.weight.detach())
    ref_u = U.sparse_inverse_conv_k2(ref_c, cluster, koff, up.weight.detach())
    down, up = down.cuda(), up.cuda()
    p = Point(grid_coord=torch.from_numpy(g["grid_coord"]).cuda(), feat=feat.cuda(), offset=torch.from_numpy(g["offset"]).cuda())
    p.serialization(order="z")
    p.sparsify()
    x = p.sparse_conv_feat
    x.indice_dict["_serial"] = (p.serialized_code, p.serialized_order, p.grid_coord.long(), p.batch)
    c = down(x)
    u = up(c)
    # a tensor built by somebody else (no serialization attached, like the reference's SpUNetBase.forward builds it)
    from ptv3_b200.structure import SparseConvTensor
    c2 = down(SparseConvTensor(feat.cuda(), x.indices, x.spatial_shape, x.batch_size))
    assert torch.equal(c2.indices, c.indices) and torch.equal(c2.features, c.features)
    # coarse rows come in z-order, the oracle's in (b,x,y,z) order: compare through the site keys
    key = lambda t: (t[:, 0].long() << 48) | (t[:, 1].long() << 32) | (t[:, 2].long() << 16) | t[:, 3].long()
    ko, kr = key(c.indices.cpu()), key(out_idx)
    assert torch.equal(torch.sort(ko)[0], torch.sort(kr)[0]) and kr.unique().numel() == kr.numel()   # same site set
    assert rel_l2(c.features.cpu()[torch.argsort(ko)], ref_c[torch.argsort(kr)]) < 1e-5
    assert torch.equal(u.indices.cpu(), x.indices.cpu()) and rel_l2(u.features, ref_u) < 1e-5


@pytest.mark.parametrize("cin,cout,k3", [(192, 128, 27), (384, 256, 27), (256, 128, 8), (320, 64, 8)])
def test_tc_conv_channel_counts_of_the_decoder(golden, cin, cout, k3):
    """The tcgen05 conv with 3 / 5 / 6 K-blocks per kernel offset (concatenated skip channels) and 8-column pair
    tables, against the CPU oracle (fp32 math on the same bf16-rounded operands; only the bf16 rounding of the output
    differs) and, as a cross-check, the CUDA-core kernel."""
    from ptv3_b200 import ops
    g = golden("spunet")
    _, idx = _batch_indices(g)
    n = idx.shape[0]
    gen = torch.Generator().manual_seed(cin + k3)
    nbr = ops.subm_build_map(idx.int().cuda(), 3)
    if k3 == 8:
        nbr = nbr[:, :8].contiguous()
    x = torch.randn(n, cin, generator=gen).bfloat16().cuda()
    w = (torch.randn(k3, cout, cin, generator=gen) / (k3 * cin) ** 0.5).bfloat16().cuda()
    b = torch.randn(cout, generator=gen).cuda()
    assert ops.subm_conv_tc_supported(cin, cout, k3)
    got = ops.subm_conv_packed(x, ops.subm_pack_weight(w), b, nbr, cout)
    wo = w.float().cpu().permute(1, 0, 2).reshape(cout, k3, 1, 1, cin)          # oracle layout (out, k..., in)
    ref = O.subm_conv(x.float().cpu(), nbr.cpu().long(), wo, b.cpu())
    assert rel_l2(got, ref) < 4e-3
    assert rel_l2(ops.subm_conv(x, w, b, nbr, algo=1), ref) < 4e-3


@pytest.mark.parametrize("dtype,tol", [(torch.float32, 1e-4), (torch.bfloat16, 1e-2)])
def test_spunet_forward_golden(golden, dtype, tol):
    from ptv3_b200 import spunet as S
    g = golden("spunet")
    m = S.SpUNetBase(in_channels=4, num_classes=13, compute_dtype=dtype)
    missing, unexpected = m.load_state_dict(_weights(), strict=False)
    assert not unexpected and all(k.endswith("num_batches_tracked") for k in missing), (missing, unexpected)
    m = m.eval().cuda()
    out = m(dict(grid_coord=torch.from_numpy(g["grid_coord"]).cuda(), feat=torch.from_numpy(g["feat"]).cuda(),
                 offset=torch.from_numpy(g["offset"]).cuda()))
    ref = torch.from_numpy(g["logits"])
    assert out.shape == ref.shape and torch.isfinite(out).all()
    err = rel_l2(out, ref)
    print(f"SpUNet {dtype}: logits rel-L2 {err:.2e}")
    assert err < tol
    if dtype == torch.float32:
        assert (out.argmax(1).cpu() == ref.argmax(1)).float().mean().item() > 0.999
