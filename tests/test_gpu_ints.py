"""GPU parity, integer kernels: bit-exact against the oracle and the frozen reference outputs."""
import numpy as np
import pytest
import torch

from oracle import gridsample as G
from oracle import ops as O
from oracle import serialization as S

pytestmark = pytest.mark.gpu


def _ops():
    from ptv3_b200 import ops
    return ops


def dev(a, dtype=None):
    t = torch.from_numpy(np.ascontiguousarray(a))
    if dtype is not None:
        t = t.to(dtype)
    return t.cuda()


def test_library_loaded():
    from ptv3_b200 import _lib
    assert _lib.load().ptv3_abi_version() == 1


def test_encode_golden(golden):
    ops = _ops()
    g = golden("serialization")
    for o in S.ORDERS:
        got = ops.encode(dev(g["kat_grid"]), dev(g["kat_batch"]), 10, o).cpu().numpy()
        assert np.array_equal(got, g[f"kat_{o}"]), o
    for depth in range(1, 17):
        code = ops.encode_orders(dev(g[f"grid_d{depth}"]), dev(g[f"batch_d{depth}"]), depth, list(S.ORDERS))
        code = code.cpu().numpy()
        for r, o in enumerate(S.ORDERS):
            assert np.array_equal(code[r], g[f"code_d{depth}_{o}"]), (depth, o)
    # int32 grid, no batch
    got = ops.encode(dev(g["grid_d9"], torch.int32), None, 9, "hilbert").cpu().numpy()
    assert np.array_equal(got, S.encode(g["grid_d9"], None, 9, "hilbert"))


@pytest.mark.parametrize("n", [1, 31, 2047, 2048, 2049, 70001])
def test_encode_random_vs_oracle(n):
    ops = _ops()
    rng = np.random.default_rng(n)
    for depth in (3, 9, 16):
        grid = rng.integers(0, 1 << depth, size=(n, 3))
        batch = np.sort(rng.integers(0, 16, size=n))
        code = ops.encode_orders(dev(grid), dev(batch), depth, list(S.ORDERS)).cpu().numpy()
        for r, o in enumerate(S.ORDERS):
            assert np.array_equal(code[r], S.encode(grid, batch, depth, o))


@pytest.mark.parametrize("rows,n,bits", [(1, 1, 8), (4, 2047, 31), (4, 2048, 31), (3, 2049, 63), (4, 150001, 31),
                                         (1, 5000, 5), (2, 300000, 40)])
def test_argsort_rows(rows, n, bits):
    ops = _ops()
    rng = np.random.default_rng(rows * 1000 + n)
    hi = (1 << bits) - 1
    code = rng.integers(0, hi, size=(rows, n), dtype=np.int64, endpoint=True)
    if n > 100:
        code[:, n // 2:] = code[:, : n - n // 2]          # duplicates: stability must define the order
    order, inverse = ops.argsort_rows(dev(code), key_bits=bits)
    ro, ri = S.argsort_rows(code)
    assert np.array_equal(order.cpu().numpy(), ro)
    assert np.array_equal(inverse.cpu().numpy(), ri)


def test_argsort_empty():
    ops = _ops()
    order, inverse = ops.argsort_rows(torch.empty((4, 0), dtype=torch.int64, device="cuda"), key_bits=10)
    assert order.shape == (4, 0) and inverse.shape == (4, 0)


def test_point_serialization_vs_oracle():
    from ptv3_b200.structure import Point
    rng = np.random.default_rng(5)
    n = 30000
    grid = np.unique(rng.integers(0, 300, size=(n, 3)), axis=0)
    rng.shuffle(grid)
    counts = [len(grid) // 3, len(grid) // 5, len(grid) - len(grid) // 3 - len(grid) // 5]
    offset = np.cumsum(counts)
    p = Point(grid_coord=dev(grid), offset=dev(offset), feat=torch.zeros(len(grid), 1, device="cuda"))
    p.serialization(order=list(S.ORDERS), shuffle_orders=False)
    batch = np.repeat(np.arange(3), counts)
    depth, code, order, inverse = S.serialization(grid, batch, S.ORDERS)
    assert p.serialized_depth == depth
    assert np.array_equal(p.serialized_code.cpu().numpy(), code)
    assert np.array_equal(p.serialized_order.cpu().numpy(), order)
    assert np.array_equal(p.serialized_inverse.cpu().numpy(), inverse)
    assert np.array_equal(p.batch.cpu().numpy(), batch)


@pytest.mark.parametrize("hash_type", ["fnv", "ravel"])
def test_gridsample_golden(golden, hash_type):
    ops = _ops()
    g = golden("gridsample")
    r = ops.gridsample(dev(g["coord"]), float(g["grid_size"]), hash_type, rand=dev(g[f"{hash_type}_rand"]),
                       quotient="f64")
    assert np.array_equal(r["key"].cpu().numpy().view(np.uint64), g[f"{hash_type}_key"])
    iu = r["idx_unique"].cpu().numpy()
    assert np.array_equal(r["grid_coord"].cpu().numpy()[iu], g[f"{hash_type}_grid_coord"])
    assert np.array_equal(g["coord"][iu], g[f"{hash_type}_coord"])
    assert np.array_equal(r["inverse"].cpu().numpy(), g[f"{hash_type}_inverse"])
    assert np.allclose(r["min_grid"].cpu().numpy() * float(g["grid_size"]), g[f"{hash_type}_min_coord"].ravel())


@pytest.mark.parametrize("hash_type", ["fnv", "ravel"])
@pytest.mark.parametrize("quotient", ["f32", "f64"])
def test_gridsample_quotient_golden(golden, hash_type, quotient):
    """Both quotient precisions of transform.py:850 (numpy 1.26.4 = float32, numpy >= 2 = float64), bit-exact against
    the unmodified reference class on a scan with points on voxel faces (tests/golden/gridsample.npz *_q32 / *_q64)."""
    ops = _ops()
    from ptv3_b200.transform import GridSample
    g = golden("gridsample")
    tag = "q32" if quotient == "f32" else "q64"
    r = ops.gridsample(dev(g["coord_q"]), float(g["grid_size"]), hash_type, rand=dev(g[f"{hash_type}_{tag}_rand"]),
                       quotient=quotient)
    iu = r["idx_unique"].cpu().numpy()
    assert np.array_equal(r["grid_coord"].cpu().numpy()[iu], g[f"{hash_type}_{tag}_grid_coord"])
    assert np.array_equal(g["coord_q"][iu], g[f"{hash_type}_{tag}_coord"])
    assert np.array_equal(r["inverse"].cpu().numpy(), g[f"{hash_type}_{tag}_inverse"])
    gs = GridSample(grid_size=float(g["grid_size"]), hash_type=hash_type, mode="train", return_grid_coord=True,
                    return_inverse=True, quotient=quotient)
    out = gs(dict(coord=dev(g["coord_q"]), index_valid_keys=["coord"], gridsample_rand=dev(g[f"{hash_type}_{tag}_rand"])))
    assert np.array_equal(out["coord"].cpu().numpy(), g[f"{hash_type}_{tag}_coord"])
    assert np.array_equal(out["inverse"].cpu().numpy(), g[f"{hash_type}_{tag}_inverse"])
    assert GridSample().quotient == "f32"          # default = the reference's own environment


def test_gridsample_transform_modes(golden):
    from ptv3_b200.transform import GridSample
    g = golden("gridsample")
    coord = dev(g["coord"])
    feat = torch.arange(len(g["coord"]), device="cuda", dtype=torch.float32)[:, None]
    # train mode, recorded draw
    gs = GridSample(grid_size=float(g["grid_size"]), hash_type="fnv", mode="train", return_grid_coord=True,
                    return_inverse=True, return_min_coord=True, quotient="f64")
    out = gs(dict(coord=coord.clone(), feat=feat.clone(), index_valid_keys=["coord", "feat"],
                  gridsample_rand=dev(g["fnv_rand"])))
    assert np.array_equal(out["coord"].cpu().numpy(), g["fnv_coord"])
    assert np.array_equal(out["grid_coord"].cpu().numpy(), g["fnv_grid_coord"])
    assert np.array_equal(out["inverse"].cpu().numpy(), g["fnv_inverse"])
    assert np.allclose(out["min_coord"].cpu().numpy(), g["fnv_min_coord"])
    # test mode: every raw point is covered, parts equal the reference's
    gs = GridSample(grid_size=float(g["grid_size"]), hash_type="fnv", mode="test", return_grid_coord=True,
                    quotient="f64")
    parts = gs(dict(coord=coord.clone(), feat=feat.clone(), index_valid_keys=["coord", "feat"]))
    assert len(parts) == int(g["test_num_parts"])
    for i, p in enumerate(parts):
        assert np.array_equal(p["index"].cpu().numpy(), g[f"test_index_{i}"])


@pytest.mark.parametrize("r,scale", [(1, 1.0), (2500, 0.2), (100000, 4.1)])
def test_gridsample_random_vs_oracle(r, scale):
    ops = _ops()
    from ptv3_b200 import synth
    coord, _ = synth.pig_scan(r, scale=scale, seed=r)
    coord = coord - 1.7
    rng = np.random.default_rng(0)
    rand = rng.integers(0, 1000, size=r)
    for hash_type, q in (("fnv", "f32"), ("ravel", "f32"), ("fnv", "f64"), ("ravel", "f64")):
        m = len(G.grid_sample(coord, 0.02, hash_type, quotient=q)["count"])
        ref = G.grid_sample(coord, 0.02, hash_type, rand=rand[:m], quotient=q)
        got = ops.gridsample(dev(coord), 0.02, hash_type, rand=dev(rand), quotient=q)
        assert got["n_voxels"] == m
        assert np.array_equal(got["grid_coord"].cpu().numpy(), ref["grid_coord"])
        assert np.array_equal(got["key"].cpu().numpy().view(np.uint64), ref["key"])
        assert np.array_equal(got["idx_sort"].cpu().numpy(), ref["idx_sort"])
        assert np.array_equal(got["count"].cpu().numpy(), ref["count"])
        assert np.array_equal(got["inverse"].cpu().numpy(), ref["inverse"])
        assert np.array_equal(got["idx_unique"].cpu().numpy(), ref["idx_unique"])
        assert np.array_equal(got["min_grid"].cpu().numpy(), ref["min_grid"])


@pytest.mark.parametrize("hash_type,quotient", [("fnv", "f32"), ("ravel", "f64")])
def test_gridsample_batch_equals_per_scene(hash_type, quotient):
    """ptv3_gridsample_batch (every scene of a step in one launch chain, one host read) = the per-scene results of
    ptv3_gridsample concatenated: ragged scenes incl. an empty one, recorded random picks, both hashes / quotients; and
    against the numpy oracle per scene."""
    ops = _ops()
    from ptv3_b200 import synth
    sizes = [5000, 1, 0, 12345, 777]
    coords = [synth.pig_scan(max(r, 1), scale=0.3 + 0.2 * i, seed=40 + i)[0][:r] - 0.4 * i for i, r in enumerate(sizes)]
    allc = dev(np.concatenate(coords))
    offset = dev(np.cumsum(sizes))
    rng = np.random.default_rng(1)
    rand = rng.integers(0, 1000, size=sum(sizes))
    got = ops.gridsample_batch(allc, offset, 0.02, hash_type, rand=dev(rand), quotient=quotient)
    vo = got["voxel_offset_host"].numpy()
    start_r, start_v = 0, 0
    for i, (c, r) in enumerate(zip(coords, sizes)):
        if r == 0:
            assert vo[i] == start_v
            continue
        m = len(G.grid_sample(c, 0.02, hash_type, quotient=quotient)["count"])
        ref = G.grid_sample(c, 0.02, hash_type, rand=rand[start_v:start_v + m], quotient=quotient)
        assert vo[i] - start_v == m, i
        sl_r, sl_v = slice(start_r, start_r + r), slice(start_v, start_v + m)
        assert np.array_equal(got["grid_coord"][sl_r].cpu().numpy(), ref["grid_coord"])
        assert np.array_equal(got["key"][sl_r].cpu().numpy().view(np.uint64), ref["key"])
        assert np.array_equal(got["idx_sort"][sl_r].cpu().numpy() - start_r, ref["idx_sort"])
        assert np.array_equal(got["inverse"][sl_r].cpu().numpy() - start_v, ref["inverse"])
        assert np.array_equal(got["count"][sl_v].cpu().numpy(), ref["count"])
        assert np.array_equal(got["idx_unique"][sl_v].cpu().numpy() - start_r, ref["idx_unique"])
        assert np.array_equal(got["min_grid"][i].cpu().numpy(), ref["min_grid"])
        one = ops.gridsample(dev(c), 0.02, hash_type, rand=dev(rand[start_v:start_v + m]), quotient=quotient)
        assert np.array_equal(one["idx_unique"].cpu().numpy(), ref["idx_unique"])
        start_r += r
        start_v += m
    assert got["n_voxels"] == start_v


def _scene_grid(seed, n_scenes=3):
    from ptv3_b200 import synth
    grids, counts = [], []
    for s in range(n_scenes):
        coord, _ = synth.pig_scan(6000 + 2500 * s, scale=0.5 + 0.4 * s, seed=seed + s)
        gs = G.grid_sample(coord, 0.02)
        grids.append(gs["grid_coord"][gs["idx_unique"]])
        counts.append(len(gs["idx_unique"]))
    return np.concatenate(grids), np.asarray(counts)


def test_pool_front_vs_oracle():
    """Cluster ids, CSR, heads and the pooled code/order/inverse equal torch.unique/argsort of the reference."""
    ops = _ops()
    grid, counts = _scene_grid(3)
    batch = np.repeat(np.arange(len(counts)), counts)
    depth, code, order, inverse = S.serialization(grid, batch, S.ORDERS, perm=[2, 0, 3, 1])
    for pd in (1, 2):
        pf = ops.pool_front(dev(code), dev(order), pd)
        c = code >> (3 * pd)
        uniq, cluster, cnt = np.unique(c[0], return_inverse=True, return_counts=True)
        m = len(uniq)
        assert pf["n_pooled"] == m
        assert np.array_equal(pf["cluster"].cpu().numpy(), cluster)
        idx_ptr = pf["idx_ptr"].cpu().numpy()
        assert np.array_equal(np.diff(idx_ptr), cnt) and idx_ptr[0] == 0
        indices = pf["indices"].cpu().numpy()
        assert np.array_equal(cluster[indices], np.repeat(np.arange(m), cnt))      # a valid sort of `cluster`
        head = pf["head"].cpu().numpy()
        assert np.array_equal(head, indices[idx_ptr[:-1]])
        new_code = c[:, head]
        assert np.array_equal(pf["code"].cpu().numpy(), new_code)
        ro, ri = S.argsort_rows(new_code)
        assert np.array_equal(pf["order"].cpu().numpy(), ro)
        assert np.array_equal(pf["inverse"].cpu().numpy(), ri)
        ng, nb, ni = ops.pool_gather_ints(dev(grid), dev(batch), pf["head"], pd)
        assert np.array_equal(ng.cpu().numpy(), grid[head] >> pd)
        assert np.array_equal(nb.cpu().numpy(), batch[head])
        assert np.array_equal(ni.cpu().numpy(), np.concatenate([batch[head][:, None], grid[head] >> pd], 1))
        off = ops.batch_to_offset(nb, len(counts)).cpu().numpy()
        assert np.array_equal(off, np.cumsum(np.bincount(batch[head], minlength=len(counts))))


def test_batch_to_offset_with_empty_scenes():
    ops = _ops()
    batch = np.array([1, 1, 1, 4, 4, 6])
    off = ops.batch_to_offset(dev(batch), 8).cpu().numpy()
    assert np.array_equal(off, np.cumsum(np.bincount(batch, minlength=8)))
    # and back: offset2batch (utils/misc.py:13-22) skips the empty scenes exactly like repeat_interleave
    got = ops.offset_to_batch(dev(off), len(batch)).cpu().numpy()
    assert np.array_equal(got, batch)
    sizes = np.array([70001, 1, 0, 33333, 5])
    ref = np.repeat(np.arange(len(sizes)), sizes)
    assert np.array_equal(ops.offset_to_batch(dev(np.cumsum(sizes)), int(sizes.sum())).cpu().numpy(), ref)


def test_windows_golden(golden):
    ops = _ops()
    g = golden("padding")
    for i in range(int(g["n_cases"])):
        off, K = g[f"offset_{i}"], int(g[f"K_{i}"])
        ref = O.windows_from_offset(off, K)
        w, nw = ops.build_windows(dev(off), K, int(off[-1]))
        n = int(nw.item())
        assert n == len(ref), i
        assert np.array_equal(w.cpu().numpy()[:n], ref), i
        assert (w.cpu().numpy()[n:] == 0).all()


def test_subm_map_vs_oracle():
    ops = _ops()
    grid, counts = _scene_grid(11, n_scenes=2)
    batch = np.repeat(np.arange(len(counts)), counts)
    idx = np.concatenate([batch[:, None], grid], 1).astype(np.int32)
    for k in (3, 5):
        ref = O.subm_neighbor_table(torch.from_numpy(idx), k).numpy()
        got = ops.subm_build_map(dev(idx), k).cpu().numpy()
        assert np.array_equal(got, ref)
        # duplicate sites (not produced by GridSample, but allowed by the ABI): the lowest row of a site is the one
        # every neighbour refers to, and a duplicate row gets the same neighbour row as its canonical row
        dup = np.concatenate([idx, idx[:100]], 0)
        got2 = ops.subm_build_map(dev(dup), k).cpu().numpy()
        assert np.array_equal(got2[: len(idx)], ref)
        assert np.array_equal(got2[len(idx):], ref[:100])
    one = ops.subm_build_map(dev(idx), 1).cpu().numpy()
    assert np.array_equal(one[:, 0], np.arange(len(idx)))


@pytest.mark.parametrize("n_scenes,seed", [(1, 3), (2, 11), (5, 12)])
def test_subm_map_block_search_vs_oracle(n_scenes, seed):
    """ptv3_subm_build_map_z (4^3-block occupancy masks over the z-order serialization, 8 probes per row) gives the same
    table as the oracle for k = 3 and k = 5, at the original depth and with spare depth (sites at the grid border, rows
    in scrambled storage order)."""
    ops = _ops()
    grid, counts = _scene_grid(seed, n_scenes=n_scenes)
    batch = np.repeat(np.arange(len(counts)), counts)
    rng = np.random.default_rng(seed)
    perm = np.concatenate([o + rng.permutation(c) for o, c in zip(np.cumsum([0] + list(counts[:-1])), counts)])
    grid = grid[perm]                                           # storage order != any spatial order
    idx = np.concatenate([batch[:, None], grid], 1).astype(np.int32)
    depth0 = int(int(grid.max()) + 1).bit_length()
    for depth in (depth0, depth0 + 2):
        code = ops.encode_orders(dev(grid), dev(batch), depth, ["z"])
        order, _ = ops.argsort_rows(code, key_bits=3 * depth + max(1, (n_scenes - 1).bit_length()))
        for k in (3, 5):
            ref = O.subm_neighbor_table(torch.from_numpy(idx), k).numpy()
            got = ops.subm_build_map(dev(idx), k, z=(code[0], order[0], depth)).cpu().numpy()
            assert np.array_equal(got, ref), (depth, k)
    assert ops.device_error() == 0


def test_voxelize_first_matches_reference(golden):
    """A0': first point per voxel with numpy's float32 quotient, both index orders, bit-exact against the reference's
    `voxelize_points` (frozen) and the oracle on a fresh scan; normalisation to the unit sphere."""
    from oracle import gridsample as G
    from ptv3_b200 import synth, transform
    g = golden("voxelize")
    coord = torch.from_numpy(g["coord"]).cuda()
    idx, grid = transform.voxelize_first(coord, 0.02, keep_order=True)
    assert torch.equal(idx.cpu(), torch.from_numpy(g["idx_sorted"]))
    assert torch.equal(grid.cpu(), torch.from_numpy(g["grid"]).long())
    idx, _ = transform.voxelize_first(coord, 0.02, keep_order=False)
    assert torch.equal(idx.cpu(), torch.from_numpy(g["idx_lex"]))
    c, _ = synth.pig_scan(50000, scale=2.0, seed=91)
    for keep in (True, False):
        ri, rg = G.voxelize_first(c, 0.03, keep_order=keep)
        idx, grid = transform.voxelize_first(torch.from_numpy(c).cuda(), 0.03, keep_order=keep)
        assert torch.equal(idx.cpu(), torch.from_numpy(ri)) and torch.equal(grid.cpu(), torch.from_numpy(rg))
    n, cen, sc = transform.normalize_unit_sphere(torch.from_numpy(c).cuda())
    centered = c - c.mean(0)
    ref = centered / np.linalg.norm(centered, axis=1).max()
    # floating point: the fp32 centroid depends on the summation order (numpy pairwise vs a device tree)
    assert np.allclose(n.cpu().numpy(), ref, atol=1e-5) and abs(float(n.norm(dim=1).max()) - 1.0) < 1e-6
