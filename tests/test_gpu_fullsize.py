"""Full-size checks (BASELINE.json configs[1]: 16 scenes x 100 000 raw points -> ~1.17 M voxels) through
size-independent properties that can be checked exactly on the whole batch.  The floating-point comparison with the
oracle at these sizes (one scene = ~1.5 s of oracle time) lives in tests/test_gpu_parity_fullsize.py.

  GridSample     keys strictly increasing, counts sum to R, inverse consistent, idempotent on its own output
  serialization  code[order] non-decreasing per row, inverse o order = identity, scenes stay contiguous
  pooling        code >> 3 partitions rows exactly like (batch, grid >> 1); CSR covers every row once
  neighbour map  symmetric (j = nbr[i,k]  <=>  i = nbr[j,k3-1-k]), centre column = identity
  conv           a one-hot weight at offset k is a pure gather (bit-exact in bf16); linear in the features
  attention      a per-window constant V is reproduced (softmax rows sum to one); outputs stay inside V's range
  forward        finite, and bit-identical for two runs with the same seed
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

SCENES, POINTS, SCALE, GRID = 16, 100000, 4.1, 0.02


@pytest.fixture(scope="module")
def batch():
    import bench
    dev = torch.device("cuda", 0)
    raw = bench.make_raw_scenes(SCENES, POINTS, SCALE, seed0=0)
    host, counts = bench.voxelize_on_device(raw, dev)
    return raw, {k: v.to(dev) for k, v in host.items()}, counts


def test_gridsample_full_size(batch):
    from ptv3_b200 import ops
    raw, res, counts = batch
    coord = torch.from_numpy(raw[3][0]).cuda()
    r = ops.gridsample(coord, GRID, "fnv")
    key, cnt, inv, iu = r["key"], r["count"], r["inverse"], r["idx_unique"]
    m = iu.numel()
    assert m == counts[3] == r["n_voxels"]
    ks = key[r["idx_sort"]].view(torch.int64)                                          # uint64 bit pattern, sorted as
    ku = key[iu].view(torch.int64)                                                     # unsigned by the kernel
    flip = torch.tensor(-2 ** 63, device="cuda", dtype=torch.int64)
    assert bool(((ks[1:] ^ flip) >= (ks[:-1] ^ flip)).all())                           # argsort(key) is sorted
    assert bool(((ku[1:] ^ flip) > (ku[:-1] ^ flip)).all())                            # voxel keys strictly increase
    assert torch.equal(key[iu][inv], key)                                              # inverse -> own voxel
    assert int(cnt.sum()) == POINTS and int(cnt.min()) >= 1
    assert int(inv.min()) == 0 and int(inv.max()) == m - 1
    # every raw point falls into the voxel its representative defines
    g = r["grid_coord"]
    assert torch.equal(g[iu][inv], g)
    # idempotence: sampling the representatives again keeps all of them
    r2 = ops.gridsample(coord[iu], GRID, "fnv")
    assert r2["idx_unique"].numel() == m and int(r2["count"].max()) == 1


def test_serialization_full_size(batch):
    from ptv3_b200.structure import Point
    _, res, counts = batch
    p = Point(dict(res))
    p.serialization(order=("z", "z-trans", "hilbert", "hilbert-trans"), shuffle_orders=False)
    n = p.coord.shape[0]
    ar = torch.arange(n, device="cuda")
    for k in range(4):
        code, order, inverse = p.serialized_code[k], p.serialized_order[k], p.serialized_inverse[k]
        sc = code[order]
        assert bool((sc[1:] > sc[:-1]).all())                       # codes are unique after GridSample -> strict
        assert torch.equal(inverse[order], ar)
        assert torch.equal(p.batch[order], p.batch)                 # batch id in the top bits keeps scenes contiguous
    # pooling prefix property on the real batch
    from ptv3_b200 import ops
    pf = ops.pool_front(p.serialized_code, p.serialized_order, 1)
    cluster = pf["cluster"]
    key = (p.batch << 48) | ((p.grid_coord[:, 0] >> 1) << 32) | ((p.grid_coord[:, 1] >> 1) << 16) | (p.grid_coord[:, 2] >> 1)
    uk, inv = torch.unique(key, return_inverse=True)
    m = int(cluster.max()) + 1
    assert m == uk.numel()
    # same partition: the map cluster id -> unique-key id is a bijection
    pair = torch.unique(torch.stack([cluster, inv], 1), dim=0)
    assert pair.shape[0] == m
    idx_ptr = pf["idx_ptr"]
    assert int(idx_ptr[0]) == 0 and int(idx_ptr[-1]) == n and bool((idx_ptr[1:] > idx_ptr[:-1]).all())
    assert torch.equal(torch.sort(pf["indices"])[0], ar)


@pytest.fixture(scope="module")
def nbr_map(batch):
    from ptv3_b200 import ops
    _, res, counts = batch
    b = torch.repeat_interleave(torch.arange(SCENES, device="cuda"), torch.tensor(counts, device="cuda"))
    idx = torch.cat([b[:, None], res["grid_coord"]], 1).int()
    return ops.subm_build_map(idx, 3)


def test_neighbour_map_full_size(nbr_map):
    nbr = nbr_map
    n, k3 = nbr.shape
    ar = torch.arange(n, device="cuda", dtype=torch.int32)
    assert torch.equal(nbr[:, k3 // 2], ar)
    for k in (0, 5, 12):
        j = nbr[:, k].long()
        ok = j >= 0
        assert torch.equal(nbr[j[ok], k3 - 1 - k], ar[ok])          # symmetry
    assert int(nbr.max()) < n and int(nbr.min()) >= -1


@pytest.mark.parametrize("c", [32, 64])
def test_conv_full_size_gather_and_linearity(nbr_map, c):
    from ptv3_b200 import ops
    nbr = nbr_map
    n = nbr.shape[0]
    g = torch.Generator(device="cuda").manual_seed(c)
    x = torch.randn(n, c, device="cuda", generator=g).bfloat16()
    for k in (13, 4):                                               # one-hot weight: out[i] = x[nbr[i,k]] exactly
        w = torch.zeros(27, c, c, device="cuda", dtype=torch.bfloat16)
        w[k] = torch.eye(c, device="cuda", dtype=torch.bfloat16)
        out = ops.subm_conv_packed(x, ops.subm_pack_weight(w), None, nbr, c)
        j = nbr[:, k].long()
        ref = torch.where((j >= 0)[:, None], x[j.clamp_min(0)], torch.zeros_like(x))
        assert torch.equal(out, ref), k
    w = (torch.randn(27, c, c, device="cuda", generator=g) / (27 * c) ** 0.5).bfloat16()
    packed = ops.subm_pack_weight(w)
    y = torch.randn(n, c, device="cuda", generator=g).bfloat16()
    lhs = ops.subm_conv_packed((x.float() + y.float()).bfloat16(), packed, None, nbr, c).float()
    rhs = ops.subm_conv_packed(x, packed, None, nbr, c).float() + ops.subm_conv_packed(y, packed, None, nbr, c).float()
    assert ((lhs - rhs).norm() / rhs.norm()).item() < 1e-2          # bf16 rounding of x + y and of the outputs


def test_attention_full_size_properties(batch):
    from ptv3_b200 import ops
    from ptv3_b200.structure import Point
    _, res, counts = batch
    p = Point(dict(res))
    p.serialization(order=("z", "hilbert"), shuffle_orders=False)
    n, h, d, K = p.coord.shape[0], 4, 16, 1024
    order = p.serialized_order[1]
    windows, nwin = ops.build_windows(p.offset, K, n)
    g = torch.Generator(device="cuda").manual_seed(7)
    qkv = torch.randn(n, 3, h, d, device="cuda", generator=g)
    # V constant per scene and head: every window lies inside one scene, so the output must reproduce it
    vconst = torch.randn(SCENES, h, d, device="cuda", generator=g).bfloat16().float()
    qkv[:, 2] = vconst[p.batch]
    out = ops.attn_window(qkv.reshape(n, -1).bfloat16(), order, windows, nwin, h, K).float().view(n, h, d)
    assert torch.isfinite(out).all()
    assert (out - vconst[p.batch]).abs().max().item() < 2e-2        # bf16 P, fp32 accumulation, bf16 output
    # random V: the output is a convex combination of the scene's values
    qkv[:, 2] = torch.randn(n, h, d, device="cuda", generator=g)
    out = ops.attn_window(qkv.reshape(n, -1).bfloat16(), order, windows, nwin, h, K).float().view(n, h, d)
    v = qkv[:, 2].bfloat16().float()
    assert out.max().item() <= v.max().item() * 1.01 and out.min().item() >= v.min().item() * 1.01


def test_forward_full_size_deterministic(batch):
    import bench
    _, res, counts = batch
    model, _ = bench.build_model(torch.bfloat16, torch.device("cuda", 0))
    torch.manual_seed(0)
    a = model(dict(res))["pred"].clone()
    torch.manual_seed(0)
    b = model(dict(res))["pred"]
    assert a.shape == (sum(counts), 6, 4) and torch.isfinite(a).all()
    assert torch.equal(a, b)
    assert float(a[..., 3].min()) >= 0.0 and float(a[..., 3].max()) <= 1.0
