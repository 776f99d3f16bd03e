"""CPU: the oracle against the frozen outputs of the reference (tests/golden)."""
import numpy as np
import pytest
import torch

from oracle import gridsample as G
from oracle import ops as O
from oracle import ptv3 as P
from oracle import serialization as S
from ptv3_b200 import synth


def test_encode_kat_and_all_depths(golden):
    g = golden("serialization")
    for o in S.ORDERS:
        assert np.array_equal(S.encode(g["kat_grid"], g["kat_batch"], 10, o), g[f"kat_{o}"])
    # SURVEY.md 8c known answers
    assert S.encode(g["kat_grid"], g["kat_batch"], 10, "hilbert").tolist() == \
        [0, 7, 3, 1, 36, 1073743272, 1174405120, 3094547079]
    for depth in range(1, 17):
        for o in S.ORDERS:
            got = S.encode(g[f"grid_d{depth}"], g[f"batch_d{depth}"], depth, o)
            assert np.array_equal(got, g[f"code_d{depth}_{o}"]), (depth, o)


def test_code_prefix_property():
    """code >> 3 partitions points exactly as (batch, grid >> 1): pooling relies on it."""
    rng = np.random.default_rng(0)
    g = rng.integers(0, 256, size=(20000, 3))
    b = rng.integers(0, 3, size=20000)
    ref = (b.astype(np.int64) << 40) | ((g[:, 0] >> 1) << 20) | ((g[:, 1] >> 1) << 10) | (g[:, 2] >> 1)
    for o in S.ORDERS:
        c = S.encode(g, b, 8, o) >> 3
        _, inv_c = np.unique(c, return_inverse=True)
        _, inv_r = np.unique(ref, return_inverse=True)
        # same partition <=> bijection between labels
        pairs = np.unique(np.stack([inv_c, inv_r]), axis=1)
        assert pairs.shape[1] == inv_c.max() + 1 == inv_r.max() + 1


def test_hash_kat():
    a = np.array([[0, 0, 0], [1, 2, 3], [100, 50, 25]])
    assert [hex(x) for x in G.fnv_hash_vec(a)] == ["0xd94d12186c0f2fb7", "0xd949aa186c0c492b", "0xd8d2ec186ba792f0"]
    assert G.ravel_hash_vec(a).tolist() == [0, 1381, 133925]


@pytest.mark.parametrize("hash_type", ["fnv", "ravel"])
def test_gridsample_vs_reference(golden, hash_type):
    g = golden("gridsample")
    r = G.grid_sample(g["coord"], float(g["grid_size"]), hash_type, rand=g[f"{hash_type}_rand"], quotient="f64")
    assert np.array_equal(r["key"], g[f"{hash_type}_key"])
    iu = r["idx_unique"]
    assert np.array_equal(r["grid_coord"][iu], g[f"{hash_type}_grid_coord"])
    assert np.array_equal(g["coord"][iu], g[f"{hash_type}_coord"])
    assert np.array_equal(r["inverse"], g[f"{hash_type}_inverse"])
    assert np.allclose(r["min_grid"] * float(g["grid_size"]), g[f"{hash_type}_min_coord"].ravel())


@pytest.mark.parametrize("hash_type", ["fnv", "ravel"])
@pytest.mark.parametrize("quotient", ["f32", "f64"])
def test_gridsample_quotient_vs_reference(golden, hash_type, quotient):
    """transform.py:850 under the reference's pinned numpy 1.26.4 (float32 quotient) and under numpy >= 2 (float64),
    both frozen from the unmodified class on a scan with points on voxel faces; the two really differ there."""
    g = golden("gridsample")
    tag = "q32" if quotient == "f32" else "q64"
    r = G.grid_sample(g["coord_q"], float(g["grid_size"]), hash_type, rand=g[f"{hash_type}_{tag}_rand"], quotient=quotient)
    iu = r["idx_unique"]
    assert np.array_equal(r["grid_coord"][iu], g[f"{hash_type}_{tag}_grid_coord"])
    assert np.array_equal(g["coord_q"][iu], g[f"{hash_type}_{tag}_coord"])
    assert np.array_equal(r["inverse"], g[f"{hash_type}_{tag}_inverse"])
    other = G.grid_sample(g["coord_q"], float(g["grid_size"]), hash_type, quotient="f64" if quotient == "f32" else "f32")
    assert len(other["count"]) != len(r["count"])


def test_gridsample_test_mode(golden):
    g = golden("gridsample")
    r = G.grid_sample(g["coord"], float(g["grid_size"]), "fnv", quotient="f64")
    parts = G.grid_sample_test_parts(r["count"], r["idx_sort"])
    assert len(parts) == int(g["test_num_parts"])
    for i, p in enumerate(parts):
        assert np.array_equal(p, g[f"test_index_{i}"])
    covered = np.zeros(len(g["coord"]), bool)
    for p in parts:
        covered[p] = True
    assert covered.all()


def test_padding_vs_reference(golden):
    g = golden("padding")
    for i in range(int(g["n_cases"])):
        pad, unpad, cu = O.padding_and_inverse(g[f"offset_{i}"], int(g[f"K_{i}"]))
        assert np.array_equal(pad, g[f"pad_{i}"]), i
        assert np.array_equal(unpad, g[f"unpad_{i}"]), i
        assert np.array_equal(cu, g[f"cu_{i}"]), i


def test_windows_equal_padding_semantics(golden):
    """The (q,k) window records describe the same key sets as the padded layout."""
    g = golden("padding")
    for i in range(int(g["n_cases"])):
        off, K = g[f"offset_{i}"], int(g[f"K_{i}"])
        pad, unpad, cu = O.padding_and_inverse(off, K)
        win = O.windows_from_offset(off, K)
        seen = np.zeros(off[-1], int)
        for q0, ql, k0, kl in win:
            # padded slot of the first query of this window
            slot = unpad[q0]
            w = np.searchsorted(cu, slot, side="right") - 1
            keys = np.sort(pad[cu[w]:cu[w + 1]])
            assert np.array_equal(keys, np.arange(k0, k0 + kl)), (i, q0)
            seen[q0:q0 + ql] += 1
        assert (seen == 1).all()


def test_attention_vs_reference(golden):
    g = golden("attention")
    out = O.window_attention(torch.from_numpy(g["qkv"]), torch.from_numpy(g["order"]), torch.from_numpy(g["inverse"]),
                             g["offset"], int(g["H"]), int(g["K"]))
    out = out @ torch.from_numpy(g["proj_w"]).t() + torch.from_numpy(g["proj_b"])
    ref = torch.from_numpy(g["out"])
    assert ((out - ref).norm() / ref.norm()).item() < 1e-5


def test_subm_conv_vs_dense_conv3d():
    """Pins the restated spconv semantics on torch's dense cross-correlation."""
    torch.manual_seed(0)
    for k in (3, 5):
        B, S_, Cin, Cout = 2, 9, 5, 7
        occ = torch.rand(B, S_, S_, S_) < 0.3
        idx = torch.nonzero(occ).to(torch.int32)
        feat = torch.randn(idx.shape[0], Cin)
        w = torch.randn(Cout, k, k, k, Cin)
        bias = torch.randn(Cout)
        dense = torch.zeros(B, Cin, S_, S_, S_)
        dense[idx[:, 0].long(), :, idx[:, 1].long(), idx[:, 2].long(), idx[:, 3].long()] = feat
        ref = torch.nn.functional.conv3d(dense, w.permute(0, 4, 1, 2, 3), bias, padding=k // 2)
        ref = ref[idx[:, 0].long(), :, idx[:, 1].long(), idx[:, 2].long(), idx[:, 3].long()]
        got = O.subm_conv(feat, O.subm_neighbor_table(idx, k), w, bias)
        assert torch.allclose(got, ref, atol=1e-4, rtol=1e-4)


def test_segment_csr_and_scatter_mean():
    torch.manual_seed(0)
    src = torch.randn(10, 3)
    ptr = torch.tensor([0, 3, 3, 4, 10])
    mx = O.segment_csr(src, ptr, "max")
    assert torch.equal(mx[0], src[0:3].max(0).values) and torch.equal(mx[1], torch.zeros(3))
    assert torch.allclose(O.segment_csr(src, ptr, "mean")[3], src[4:10].mean(0))
    idx = torch.tensor([0, 0, 2, 2, 2, 1, 0, 1, 1, 2])
    sm = O.scatter_mean(src, idx)
    assert torch.allclose(sm[2], src[idx == 2].mean(0))


def _weights():
    import json, os
    from conftest import GOLDEN
    with open(os.path.join(GOLDEN, "param_shapes.json")) as f:
        shapes = json.load(f)
    return {k: torch.from_numpy(v) for k, v in synth.make_weights(shapes, seed=0).items()}


def test_forward_vs_reference(golden):
    g = golden("forward")
    W = _weights()
    for i in range(int(g["n_scenes"])):
        tr = {}
        pt = P.backbone_forward(W, g[f"coord_{i}"], g[f"grid_coord_{i}"], g[f"feat_{i}"], [len(g[f"coord_{i}"])],
                                perms=list(g[f"perms_{i}"]), trace=tr)
        for k in ("stem", "enc0", "pool1", "enc4", "dec3"):
            ref = torch.from_numpy(g[f"{k}_{i}"])
            assert ((tr[k] - ref).norm() / ref.norm()).item() < 1e-5, (i, k)
        pred = P.offset_head(W, pt.feat)
        ref = torch.from_numpy(g[f"pred_{i}"])
        assert ((pred - ref).norm() / ref.norm()).item() < 1e-5


def _head_weights(tag):
    import json, os
    from conftest import GOLDEN
    with open(os.path.join(GOLDEN, "param_shapes_heads.json")) as f:
        shapes = json.load(f)[tag]
    return {k: torch.from_numpy(v) for k, v in synth.make_weights(shapes, seed=0).items()}


def test_keypoint_and_segmentor_heads_vs_reference(golden):
    """KeypointPTv3 (scatter_mean + MLP) and DefaultSegmentorV2 (C_in = 6, 20 classes), frozen from
    the unmodified reference classes (tests/golden/make_golden.py::gold_heads)."""
    g = golden("heads")
    n = len(g["coord"])
    W = _head_weights("keypoint")
    pt = P.backbone_forward(W, g["coord"], g["grid_coord"], g["feat4"], [n], perms=list(g["perms_keypoint"]))
    pred = P.keypoint_head(W, pt.feat, pt.batch)
    ref = torch.from_numpy(g["out_keypoint"])
    assert pred.shape == ref.shape == (1, 6, 3)
    assert ((pred - ref).norm() / ref.norm()).item() < 1e-5
    W = _head_weights("segmentor")
    pt = P.backbone_forward(W, g["coord"], g["grid_coord"], g["feat6"], [n], cfg=dict(P.BASE_CFG, in_channels=6),
                            perms=list(g["perms_segmentor"]))
    logits = P.seg_head(W, pt.feat)
    ref = torch.from_numpy(g["out_segmentor"])
    assert logits.shape == ref.shape == (n, 20)
    assert ((logits - ref).norm() / ref.norm()).item() < 1e-5


def test_training_grads_vs_reference(golden):
    """oracle autograd (train() mode: BatchNorm batch statistics, drop_path 0) against the unmodified
    reference's loss.backward() frozen in tests/golden/train.npz."""
    g = golden("train")
    W = _weights()
    loss, grads = P.offset_keypoint_grads(W, g["coord"], g["grid_coord"], g["feat"], [len(g["coord"])], g["target"],
                                          perms=list(g["perms"]))
    assert abs(loss.item() - float(g["loss"])) < 1e-5 * abs(float(g["loss"]))
    names = [str(n) for n in g["grad_names"]]
    assert set(names) == set(grads)
    from make_golden_names import TRAIN_GRADS
    for i, k in enumerate(TRAIN_GRADS):
        ref = torch.from_numpy(g[f"grad_{i}"])
        assert ((grads[k] - ref).norm() / ref.norm()).item() < 1e-4, k
    got = torch.tensor([grads[n].norm().item() for n in names], dtype=torch.float64)
    ref = torch.from_numpy(g["grad_norms"])
    assert ((got - ref).abs() / ref.clamp_min(1e-12)).max().item() < 1e-3


def test_dense_branch_two_scene_batch_vs_reference(golden):
    """enable_flash=False (both fork configs): the reference uses ONE patch size K = min(smallest scene, patch) for
    the whole batch (point_transformer_v3m1_base.py:173-176).  Oracle with cfg enable_flash=False against the unmodified
    reference on a two-scene batch of unequal sizes: eval prediction, training loss and gradients; and the flash
    window semantics really differ on this batch."""
    from make_golden_names import TRAIN_GRADS
    g = golden("dense_batch")
    W = _weights()
    cfg = dict(P.BASE_CFG, enable_flash=False)
    pt = P.backbone_forward(W, g["coord"], g["grid_coord"], g["feat"], g["offset"], cfg=cfg, perms=list(g["perms"]))
    pred = P.offset_head(W, pt.feat)
    ref = torch.from_numpy(g["pred"])
    assert ((pred - ref).norm() / ref.norm()).item() < 1e-5
    pt = P.backbone_forward(W, g["coord"], g["grid_coord"], g["feat"], g["offset"], perms=list(g["perms"]))
    assert ((P.offset_head(W, pt.feat) - ref).norm() / ref.norm()).item() > 1e-3      # flash windows: another function
    loss, grads = P.offset_keypoint_grads(W, g["coord"], g["grid_coord"], g["feat"], g["offset"], g["target"],
                                          perms=list(g["perms"]), cfg=cfg)
    assert abs(loss.item() - float(g["loss"])) < 1e-5 * abs(float(g["loss"]))
    for i, k in enumerate(TRAIN_GRADS):
        r = torch.from_numpy(g[f"grad_{i}"])
        assert ((grads[k] - r).norm() / r.norm()).item() < 1e-4, k


def test_voxelize_first_vs_reference(golden):
    """A0' (inference voxelisers): oracle restatement against the reference's own `voxelize_points` output, on a scan
    with points on / one ulp off the cell borders (float32 quotient)."""
    from oracle import gridsample as G
    g = golden("voxelize")
    idx, grid = G.voxelize_first(g["coord"], 0.02, keep_order=True)
    assert np.array_equal(idx, g["idx_sorted"]) and np.array_equal(grid, g["grid"])
    idx, _ = G.voxelize_first(g["coord"], 0.02, keep_order=False)
    assert np.array_equal(idx, g["idx_lex"])


def test_spunet_forward_vs_reference(golden):
    """oracle/spunet.py (functional) against the reference SpUNetBase class run over the restated
    spconv ops (tests/golden/make_golden.py::gold_spunet): module wiring, state_dict names, skip order."""
    import json, os
    from conftest import GOLDEN
    from oracle import spunet as U
    g = golden("spunet")
    with open(os.path.join(GOLDEN, "param_shapes_spunet.json")) as f:
        shapes = json.load(f)
    W = {k: torch.from_numpy(v) for k, v in synth.make_weights(shapes, seed=0).items()}
    out = U.spunet_forward(W, g["grid_coord"], g["feat"], g["offset"])
    ref = torch.from_numpy(g["logits"])
    assert out.shape == ref.shape and ((out - ref).norm() / ref.norm()).item() < 1e-5
    # strided conv map: every input site belongs to exactly one output site, offsets in 0..7, inverse conv of a
    # one-hot weight returns the coarse feature
    off = torch.from_numpy(g["offset"])
    batch = torch.repeat_interleave(torch.arange(len(off)), torch.diff(off, prepend=torch.zeros(1, dtype=off.dtype)))
    idx = torch.cat([batch[:, None], torch.from_numpy(g["grid_coord"])], 1)
    out_idx, cluster, koff = U.down_map(idx)
    assert torch.equal(out_idx[cluster].long()[:, 1:], idx[:, 1:] >> 1) and int(koff.min()) >= 0 and int(koff.max()) <= 7
    assert torch.unique(torch.stack([cluster, koff], 1), dim=0).shape[0] == idx.shape[0]


def test_offset_keypoint_eval_vs_reference(golden):
    """oracle.ptv3.offset_keypoint_eval pinned to the unmodified OffsetKeypointEvaluator.eval() (tests/golden/eval.npz,
    make_golden.py::gold_eval): the logged mean distance and the per-keypoint means over two validation batches."""
    from make_golden_names import eval_batches
    g = golden("eval")
    acc = None
    for b in eval_batches():
        dist, valid, acc = P.offset_keypoint_eval(b["coord"], b["pred"], b["target"], b["offset"], b.get("scale"), acc)
    mean_dist, per_kp = P.offset_keypoint_eval_summary(acc, 6)
    assert mean_dist == float(g["mean_dist"])                         # same arithmetic in the same order: exact
    assert np.array_equal(np.array(per_kp), g["per_kp"])
    assert acc[1] == 4 and acc[2 + 6 + 2] == 3                        # scene 1 of batch 0 not a sample; keypoint 2 skipped once


@pytest.mark.parametrize("tag", ["dec", "enc"])
def test_sonata_oracle_vs_reference(tag):
    """oracle/sonata.py (PT-v3m2: GridPooling, LayerScale, Linear stem, LayerNorm, enc_mode feature up-cast) pinned on the
    UNMODIFIED reference's DefaultSegmentorV2(PT-v3m2) run on a two-scene batch (tests/golden/sonata.npz)."""
    import sonata_common as SC
    from oracle import sonata as SO
    G = SC.golden()
    cfg = dict(SC.CFGS[tag]["backbone"], enable_flash=False)
    W = SC.weights(tag)
    c, g, f, off = SC.inputs(cfg["in_channels"])
    trace = {}
    pt = SO.backbone_forward(W, c, g, f, off, cfg=cfg, perms=list(G[f"perms_{tag}"]), trace=trace)
    feat = SO.upcast_features(pt) if cfg["enc_mode"] else pt.feat
    assert feat.shape[1] == SC.CFGS[tag]["out_channels"]
    logits = SO.seg_head(W, feat)
    ref = torch.from_numpy(G[f"logits_{tag}"])
    assert logits.shape == ref.shape and ((logits - ref).norm() / ref.norm()).item() < 1e-5
    for k in ("stem", "enc0", "pool1", "enc4", "dec3"):
        if f"{k}_{tag}" in G.files:
            r = torch.from_numpy(G[f"{k}_{tag}"])
            assert ((trace[k] - r).norm() / r.norm()).item() < 1e-5, k


def test_lovasz_softmax_vs_reference():
    """ptv3_b200.losses.LovaszLoss (one batched sort over all classes) against the reference's class-by-class
    LovaszLoss(mode="multiclass", ignore_index=-1): value and gradient (tests/golden/sonata.npz)."""
    import sonata_common as SC
    from ptv3_b200 import losses
    G = SC.golden()
    lg = torch.from_numpy(G["lovasz_logits"]).requires_grad_(True)
    crit = losses.Criteria([dict(type="LovaszLoss", mode="multiclass", loss_weight=1.0, ignore_index=-1)])
    v = crit(lg, torch.from_numpy(G["lovasz_labels"]))
    v.backward()
    assert abs(v.item() - float(G["lovasz_value"])) < 1e-6
    assert (lg.grad - torch.from_numpy(G["lovasz_grad"])).abs().max().item() < 1e-6
    # all labels ignored: zero loss with a graph (lovasz.py:127-129)
    z = losses.lovasz_softmax(lg, torch.full((lg.shape[0],), -1), ignore_index=-1)
    assert z.item() == 0.0 and z.requires_grad


def test_cross_entropy_gather_form_vs_torch():
    """losses.CrossEntropyLoss (log_softmax + gather + device-wide mean instead of torch's one-block nll_loss reduction)
    against torch.nn.functional.cross_entropy: value and gradient, with ignored rows; other options fall back to torch."""
    from ptv3_b200 import losses
    g = torch.Generator().manual_seed(3)
    x = (torch.randn(5000, 22, generator=g) * 3).requires_grad_(True)
    t = torch.randint(-1, 22, (5000,), generator=g)
    ce = losses.CrossEntropyLoss(loss_weight=0.7, ignore_index=-1)
    assert ce.plain
    a = ce(x, t)
    ga, = torch.autograd.grad(a, x)
    b = torch.nn.functional.cross_entropy(x, t, ignore_index=-1) * 0.7
    gb, = torch.autograd.grad(b, x)
    assert abs(a.item() - b.item()) < 1e-6 and (ga - gb).abs().max().item() < 1e-9
    sm = losses.CrossEntropyLoss(label_smoothing=0.1, ignore_index=-1)
    assert not sm.plain
    assert abs(sm(x, t).item() - torch.nn.functional.cross_entropy(x, t, ignore_index=-1, label_smoothing=0.1).item()) < 1e-6


def test_neighbor_table_many_scenes_brute_force():
    """oracle.ops.subm_neighbor_table against a dictionary lookup on a batch of 5 scenes (the 64-bit key used to put the
    batch id at bit 63, so batch ids >= 2 wrapped around; every earlier fixture had at most two scenes)."""
    g = torch.Generator().manual_seed(0)
    idx = torch.unique(torch.cat([torch.randint(0, 5, (600, 1), generator=g), torch.randint(0, 7, (600, 3), generator=g)], 1), dim=0)
    for k in (3, 5):
        nbr = O.subm_neighbor_table(idx, k)
        where = {tuple(r): i for i, r in enumerate(idx.tolist())}
        r = k // 2
        for i, (b, x, y, z) in enumerate(idx.tolist()):
            col = 0
            for dx in range(-r, r + 1):
                for dy in range(-r, r + 1):
                    for dz in range(-r, r + 1):
                        assert int(nbr[i, col]) == where.get((b, x + dx, y + dy, z + dz), -1)
                        col += 1
