"""GPU parity of the backward path: every custom autograd op against torch.autograd through
the CPU oracle on the same seeded inputs, then the whole OffsetKeypointPTv3 training step against the UNMODIFIED
reference's loss.backward() frozen in tests/golden/train.npz.
Tolerances: fp32 operands rel-L2 <= 1e-4 per op (<= 1e-3 for gradients that went through the whole network),
bf16 operands <= 1e-2 per op."""
import json
import os

import numpy as np
import pytest
import torch

from oracle import gridsample as G
from oracle import ops as O
from oracle import ptv3 as P

pytestmark = pytest.mark.gpu

TOL = {torch.float32: 1e-4, torch.bfloat16: 1e-2}


def rel_l2(a, b):
    a, b = a.detach().float().cpu(), b.detach().float().cpu()
    return ((a - b).norm() / b.norm().clamp_min(1e-30)).item()


def _sites(seed, n_scenes=2, r=2500, scale=0.4):
    from ptv3_b200 import synth
    idx = []
    for b in range(n_scenes):
        coord, _ = synth.pig_scan(r, scale=scale, seed=seed + b)
        gs = G.grid_sample(coord, 0.02)
        g = gs["grid_coord"][gs["idx_unique"]]
        idx.append(np.concatenate([np.full((len(g), 1), b), g], 1))
    return torch.from_numpy(np.concatenate(idx)).int()


@pytest.mark.parametrize("dtype,simt", [(torch.float32, False), (torch.float32, True), (torch.bfloat16, False),
                                        (torch.bfloat16, True)])
@pytest.mark.parametrize("k,cin,cout", [(3, 32, 32), (3, 64, 64), (5, 4, 32), (5, 6, 32), (3, 48, 80), (3, 128, 64),
                                        (3, 64, 128), (3, 32, 64)])
def test_subm_conv_grads(dtype, simt, k, cin, cout):
    """simt = the generic CUDA-core weight-gradient kernel; otherwise the tensor-core (bf16, channels % 32 == 0) or
    the stem (fp32, C_out 32, C_in <= 8) fast path when the shape has one."""
    from ptv3_b200 import autograd as A, ops
    idx = _sites(3)
    n = idx.shape[0]
    g = torch.Generator().manual_seed(k * 1000 + cin)
    x = torch.randn(n, cin, generator=g)
    w = torch.randn(cout, k, k, k, cin, generator=g) / (k ** 3 * cin) ** 0.5
    b = torch.randn(cout, generator=g) * 0.1
    dy = torch.randn(n, cout, generator=g)
    if dtype == torch.bfloat16:      # the oracle sees the same rounded operands
        x, w, dy = x.bfloat16().float(), w.bfloat16().float(), dy.bfloat16().float()
    xr, wr, br = x.clone().requires_grad_(), w.clone().requires_grad_(), b.clone().requires_grad_()
    nbr_ref = O.subm_neighbor_table(idx, k)
    O.subm_conv(xr, nbr_ref, wr, br).backward(dy)
    xg, wg, bg = x.cuda().requires_grad_(), w.cuda().requires_grad_(), b.cuda().requires_grad_()
    nbr = ops.subm_build_map(idx.cuda(), k)
    y = A.subm_conv(xg, wg, bg, nbr, dtype)
    from ptv3_b200 import _lib
    _lib.load().ptv3_debug_set(5, int(simt))
    try:
        y.backward(dy.cuda().to(y.dtype))
    finally:
        _lib.load().ptv3_debug_set(5, 0)
    tol = TOL[dtype]
    assert rel_l2(xg.grad, xr.grad) < tol, rel_l2(xg.grad, xr.grad)
    assert rel_l2(wg.grad, wr.grad) < tol, rel_l2(wg.grad, wr.grad)
    assert rel_l2(bg.grad, br.grad) < tol, rel_l2(bg.grad, br.grad)


@pytest.mark.parametrize("dtype,simt", [(torch.float32, True), (torch.bfloat16, False), (torch.bfloat16, True)])
@pytest.mark.parametrize("sizes,K", [([700, 2300, 1024, 1], 1024), ([300, 130], 128), ([5000], 1024), ([1500, 40], 1000)])
def test_attention_grads(dtype, simt, sizes, K):
    """Ragged scenes: a short scene (one short window), a scene whose last window re-uses the previous window's tail
    (keys shared by two windows -> accumulated dK / dV), an exact multiple of K, and a one-row scene."""
    from ptv3_b200 import autograd as A, ops
    n, H, d = sum(sizes), 4, 16
    g = torch.Generator().manual_seed(n)
    qkv = torch.randn(n, 3 * H * d, generator=g)
    dout = torch.randn(n, H * d, generator=g)
    offset = torch.tensor(np.cumsum(sizes))
    order = torch.cat([torch.randperm(s, generator=g) + (int(offset[i]) - s) for i, s in enumerate(sizes)])
    inverse = torch.empty_like(order)
    inverse[order] = torch.arange(n)
    if dtype == torch.bfloat16:
        qkv, dout = qkv.bfloat16().float(), dout.bfloat16().float()
    qr = qkv.clone().requires_grad_()
    ref = O.window_attention(qr, order, inverse, offset.numpy(), H, K)
    ref.backward(dout)
    windows, nwin = ops.build_windows(offset.cuda(), K, n)
    qg = qkv.cuda().to(dtype).requires_grad_()
    out = A.attn_window(qg, order.cuda(), windows, nwin, H, K)
    from ptv3_b200 import _lib
    _lib.load().ptv3_debug_set(4, int(simt))       # bf16: tensor-core kernel (default) or the CUDA-core one
    try:
        out.backward(dout.cuda().to(dtype))
    finally:
        _lib.load().ptv3_debug_set(4, 0)
    assert rel_l2(out, ref) < TOL[dtype]
    grad, gref = qg.grad.float().cpu().view(n, 3, -1), qr.grad.view(n, 3, -1)
    for part, name in enumerate("qkv"):
        err = rel_l2(grad[:, part], gref[:, part])
        assert err < TOL[dtype], (name, err)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("reduce", ["max", "mean", "sum", "min"])
def test_segment_csr_grads(dtype, reduce):
    from ptv3_b200 import autograd as A
    g = torch.Generator().manual_seed(11)
    n, c = 4000, 48
    lens = torch.randint(0, 9, (1200,), generator=g)
    lens = lens[torch.cumsum(lens, 0) <= n]
    ptr = torch.cat([torch.zeros(1, dtype=torch.long), torch.cumsum(lens, 0)])
    m = ptr.numel() - 1
    perm = torch.randperm(n, generator=g)
    src = torch.randn(n, c, generator=g)
    dout = torch.randn(m, c, generator=g)
    if dtype == torch.bfloat16:
        src = src.bfloat16().float()
        if reduce in ("max", "min"):     # make ties certain in some segments: the FIRST row in CSR order wins
            src[perm[1]] = src[perm[0]]
    sr = src.clone().requires_grad_()
    gathered = sr[perm][: int(ptr[-1])]
    if reduce in ("max", "min"):
        # torch_scatter's arg semantics restated: first row of the segment (CSR order) that attains the extremum
        ref_out = O.segment_csr(gathered.detach(), ptr, reduce)
        seg = torch.repeat_interleave(torch.arange(m), lens)
        hit = gathered.detach() == ref_out[seg]
        pos = torch.arange(gathered.shape[0])[:, None].expand_as(hit)
        big = torch.full((m, c), gathered.shape[0], dtype=torch.long)
        firstpos = big.scatter_reduce(0, seg[:, None].expand_as(hit), torch.where(hit, pos, big[seg]), "amin")
        first = pos == firstpos[seg]
        gref_g = torch.where(first, dout[seg], torch.zeros(()))
        gref = torch.zeros(n, c)
        gref[perm[: int(ptr[-1])]] = gref_g
    else:
        O.segment_csr(gathered, ptr, reduce).backward(dout)
        gref = sr.grad
    sg = src.cuda().to(dtype).requires_grad_()
    out = A.segment_csr(sg, ptr.cuda(), reduce, perm.cuda())
    out.backward(dout.cuda().to(dtype))
    assert rel_l2(sg.grad, gref) < (1e-6 if dtype == torch.float32 else 1e-2)


def test_gather_rows_grad():
    from ptv3_b200 import autograd as A, ops
    g = torch.Generator().manual_seed(2)
    m, n, c = 700, 3000, 64
    inverse = torch.randint(0, m, (n,), generator=g)
    inverse[:m] = torch.arange(m)                        # every cluster has a member, like pooling
    indices = torch.sort(inverse, stable=True)[1]
    idx_ptr = torch.cat([torch.zeros(1, dtype=torch.long), torch.cumsum(torch.bincount(inverse, minlength=m), 0)])
    x = torch.randn(m, c, generator=g)
    dy = torch.randn(n, c, generator=g)
    xr = x.clone().requires_grad_()
    xr[inverse].backward(dy)
    xg = x.cuda().requires_grad_()
    A.gather_rows(xg, inverse.cuda(), indices.cuda(), idx_ptr.cuda()).backward(dy.cuda())
    assert rel_l2(xg.grad, xr.grad) < 1e-6


@pytest.mark.parametrize("c", [32, 64, 128, 256, 512, 48])
@pytest.mark.parametrize("xdt,odt", [(torch.float32, torch.float32), (torch.float32, torch.bfloat16),
                                     (torch.bfloat16, torch.float32)])
def test_layer_norm_grads(c, xdt, odt):
    """LayerNorm forward (ln_residual kernel) + ptv3_layer_norm_bwd against torch autograd in fp32 on the CPU; c = 48
    takes the torch fallback.  n is not a multiple of the rows per warp / CTA."""
    from ptv3_b200 import autograd as A
    g = torch.Generator().manual_seed(c)
    n = 3001
    x = torch.randn(n, c, generator=g) * 2.0 + 0.5
    dy = torch.randn(n, c, generator=g)
    if xdt == torch.bfloat16:
        x = x.bfloat16().float()
    if odt == torch.bfloat16:
        dy = dy.bfloat16().float()
    ln = torch.nn.LayerNorm(c)
    with torch.no_grad():
        ln.weight.copy_(torch.rand(c, generator=g) + 0.5)
        ln.bias.copy_(torch.randn(c, generator=g) * 0.1)
    xr = x.clone().requires_grad_()
    ref = ln(xr)
    ref.backward(dy)
    gw, gb = ln.weight.grad.clone(), ln.bias.grad.clone()
    lng = torch.nn.LayerNorm(c).cuda()
    lng.load_state_dict(ln.state_dict())
    xg = x.cuda().to(xdt).requires_grad_()
    out = A.layer_norm(xg, lng, odt)
    assert out.dtype == odt
    out.backward(dy.cuda().to(odt))
    tol = 1e-5 if odt == torch.float32 else 4e-3
    assert rel_l2(out, ref) < tol
    assert rel_l2(xg.grad, xr.grad) < (1e-5 if xdt == torch.float32 else 4e-3)
    assert rel_l2(lng.weight.grad, gw) < 1e-4 and rel_l2(lng.bias.grad, gb) < 1e-4


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("cin,cout", [(64, 192), (32, 96), (512, 2048), (64, 24), (20, 36)])
def test_linear_grads(dtype, cin, cout):
    """autograd.LinearFn (cuBLAS GEMMs + ptv3_column_sum for the bias gradient; (20, 36) takes torch's sum)."""
    from ptv3_b200 import autograd as A
    g = torch.Generator().manual_seed(cin + cout)
    n = 4099
    x = torch.randn(n, cin, generator=g)
    dy = torch.randn(n, cout, generator=g)
    lin = torch.nn.Linear(cin, cout)
    if dtype == torch.bfloat16:
        x, dy = x.bfloat16().float(), dy.bfloat16().float()
        with torch.no_grad():
            lin.weight.copy_(lin.weight.bfloat16().float())
            lin.bias.copy_(lin.bias.bfloat16().float())
    xr = x.clone().requires_grad_()
    ref = lin(xr)
    ref.backward(dy)
    ling = torch.nn.Linear(cin, cout).cuda()
    ling.load_state_dict(lin.state_dict())
    xg = x.cuda().requires_grad_()
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        out = A.linear(xg, ling, dtype)
        out.backward(dy.cuda().to(dtype))
    finally:
        torch.backends.cuda.matmul.allow_tf32 = prev
    tol = TOL[dtype]
    assert rel_l2(out, ref) < tol and rel_l2(xg.grad, xr.grad) < tol
    assert rel_l2(ling.weight.grad, lin.weight.grad) < tol
    assert rel_l2(ling.bias.grad, lin.bias.grad) < 1e-5


def _weights():
    from conftest import GOLDEN
    from ptv3_b200 import synth
    with open(os.path.join(GOLDEN, "param_shapes.json")) as f:
        shapes = json.load(f)
    return {k: torch.from_numpy(v) for k, v in synth.make_weights(shapes, seed=0).items()}


def _train_model(dtype):
    from ptv3_b200 import model as M
    m = M.OffsetKeypointPTv3(backbone_conf=dict(P.BASE_CFG, drop_path=0.0, compute_dtype=dtype))
    missing, unexpected = m.load_state_dict(_weights(), strict=False)
    assert not unexpected and all(k.endswith("num_batches_tracked") for k in missing)
    return m.cuda().train()


# bf16: the per-op gradient tests above hold 1e-2; through the whole network the 0.6 % forward noise of bf16 operands
# flips a fraction f of the non-smooth selections (sign() of the L1 loss, ReLU masks, segment-max winners), which moves
# gradient mass by ~sqrt(2 f) -- measured 1.4 % at the head, 9 % at the last block, 18-22 % in the encoder with these
# random weights (tools/grad_probe.py: keeping any one module kind in fp32 does not remove it, all-fp32 gives 1e-4).
# The whole-step bf16 check is therefore direction (rel-L2 < 0.3 <=> cosine > 0.95) + loss, not 1e-2.
@pytest.mark.parametrize("dtype,tol_loss,tol_grad", [(torch.float32, 1e-5, 1e-3), (torch.bfloat16, 1e-2, 0.3)])
def test_training_step_golden(golden, dtype, tol_loss, tol_grad):
    """Whole training step (train() mode: BatchNorm batch statistics, 5 randperm draws, loss of
    offset_keypoint_ptv3.py:50-86, backward through every kernel) against the unmodified reference."""
    from make_golden_names import TRAIN_GRADS
    g = golden("train")
    m = _train_model(dtype)
    n = len(g["coord"])
    torch.manual_seed(400)
    out = m(dict(coord=torch.from_numpy(g["coord"]).cuda(), grid_coord=torch.from_numpy(g["grid_coord"]).cuda(),
                 feat=torch.from_numpy(g["feat"]).cuda(), offset=torch.tensor([n]).cuda(), grid_size=0.02,
                 target=torch.from_numpy(g["target"]).cuda()))
    out["loss"].backward()
    ref_loss = float(g["loss"])
    assert abs(out["loss"].item() - ref_loss) < tol_loss * abs(ref_loss), (out["loss"].item(), ref_loss)
    params = dict(m.named_parameters())
    errs = {k: rel_l2(params[k].grad, torch.from_numpy(g[f"grad_{i}"])) for i, k in enumerate(TRAIN_GRADS)}
    print({k: f"{v:.2e}" for k, v in errs.items()})
    worst = max(errs.values())
    assert worst < tol_grad, errs
    names = [str(s) for s in g["grad_names"]]
    assert all(params[k].grad is not None for k in names)
    got = torch.tensor([params[k].grad.norm().item() for k in names], dtype=torch.float64)
    ref = torch.from_numpy(g["grad_norms"])
    # parameters whose gradient is analytically zero (a bias in front of a batch-statistics BatchNorm) hold rounding
    # noise on both sides: measure the error relative to max(own norm, 1e-5 x the largest norm)
    floor = (1e-5 if dtype == torch.float32 else 1e-3) * ref.max().item()
    rel = (got - ref).abs() / ref.clamp_min(floor)
    print(f"{dtype}: loss {out['loss'].item():.6f} vs {ref_loss:.6f}; worst full-tensor grad rel-L2 {worst:.2e}; "
          f"grad-norm rel err max {rel.max().item():.2e} median {rel.median().item():.2e}")
    assert rel.max().item() < (1e-2 if dtype == torch.float32 else 0.5)
    # BatchNorm running statistics were updated like nn.BatchNorm1d(momentum=0.01) does
    assert rel_l2(m.backbone.embedding.stem.norm.running_mean, torch.from_numpy(g["running_mean_stem"])) < 1e-3


@pytest.mark.parametrize("dtype,tol,tol_grad", [(torch.float32, 1e-4, 1e-3), (torch.bfloat16, 1e-2, 0.3)])
def test_dense_branch_two_scene_batch_golden(golden, dtype, tol, tol_grad):
    """enable_flash=False with a batch (the setting of both fork configs, configs/my_dataset/offset_keypoint_ptv3.py:36,
    configs/pigseg/semseg-ptv3-v1m1-0-base.py:61): one patch size K = min(smallest scene, patch) for every scene
    (point_transformer_v3m1_base.py:173-176).  Eval prediction and one training step of the CUDA model against the
    unmodified reference on a two-scene batch of unequal sizes (tests/golden/dense_batch.npz)."""
    from make_golden_names import TRAIN_GRADS
    from ptv3_b200 import model as M
    g = golden("dense_batch")
    W = _weights()
    m = M.OffsetKeypointPTv3(backbone_conf=dict(P.BASE_CFG, drop_path=0.0, enable_flash=False, compute_dtype=dtype))
    m.load_state_dict(W, strict=False)
    m = m.cuda().eval()
    data = lambda: dict(coord=torch.from_numpy(g["coord"]).cuda(), grid_coord=torch.from_numpy(g["grid_coord"]).cuda(),
                        feat=torch.from_numpy(g["feat"]).cuda(), offset=torch.from_numpy(g["offset"]).cuda(), grid_size=0.02)
    torch.manual_seed(500)
    pred = m(data())["pred"]
    e = rel_l2(pred, torch.from_numpy(g["pred"]))
    assert e < tol, e
    flash = M.OffsetKeypointPTv3(backbone_conf=dict(P.BASE_CFG, drop_path=0.0, compute_dtype=torch.float32))
    flash.load_state_dict(W, strict=False)
    torch.manual_seed(500)
    assert rel_l2(flash.cuda().eval()(data())["pred"], torch.from_numpy(g["pred"])) > 1e-3     # really another function
    m.train()
    d = data()
    d["target"] = torch.from_numpy(g["target"]).cuda()
    torch.manual_seed(500)
    out = m(d)
    out["loss"].backward()
    ref_loss = float(g["loss"])
    assert abs(out["loss"].item() - ref_loss) < (1e-5 if dtype == torch.float32 else 1e-2) * abs(ref_loss)
    params = dict(m.named_parameters())
    errs = {k: rel_l2(params[k].grad, torch.from_numpy(g[f"grad_{i}"])) for i, k in enumerate(TRAIN_GRADS)}
    print(f"dense-branch batch {dtype}: pred rel-L2 {e:.2e}, loss {out['loss'].item():.6f} vs {ref_loss:.6f}, "
          f"worst grad rel-L2 {max(errs.values()):.2e}")
    assert max(errs.values()) < tol_grad, errs


def test_training_loss_decreases():
    """Five AdamW steps on a fixed two-scene batch (bf16 operands, drop_path on): the loss goes down and stays
    finite -- optimizer + autograd plumbing end to end."""
    from make_golden_names import make_target
    from ptv3_b200 import model as M, synth
    torch.manual_seed(0)
    m = M.OffsetKeypointPTv3(backbone_conf=dict(P.BASE_CFG, drop_path=0.1)).cuda().train()
    opt = torch.optim.AdamW(m.parameters(), lr=2e-3)
    scenes = []
    for s in (71, 72):
        coord, feat = synth.pig_scan(3000, scale=0.5, seed=s)
        gs = G.grid_sample(coord, 0.02)
        iu = gs["idx_unique"]
        scenes.append((coord[iu], feat[iu], gs["grid_coord"][iu]))
    cat = lambda j: torch.from_numpy(np.concatenate([s[j] for s in scenes])).cuda()
    offset = torch.tensor(np.cumsum([len(s[0]) for s in scenes])).cuda()
    target = torch.from_numpy(make_target(int(offset[-1]), 5)).cuda()
    losses = []
    for _ in range(6):
        opt.zero_grad(set_to_none=True)
        out = m(dict(coord=cat(0), feat=cat(1), grid_coord=cat(2), offset=offset, target=target))
        out["loss"].backward()
        opt.step()
        losses.append(out["loss"].item())
    assert all(np.isfinite(losses)) and losses[-1] < losses[0], losses


def test_training_other_heads_vs_oracle(golden):
    """KeypointPTv3 (MSE on per-scene keypoints, keypoint_ptv3.py:52-63) and DefaultSegmentorV2 (cross entropy on
    per-point logits, default.py:86-88) in train() mode, fp32: loss and gradients against torch autograd through
    the oracle (BatchNorm batch statistics, dropout off, drop_path 0)."""
    import torch.nn.functional as F
    from ptv3_b200 import model as M, synth
    g = golden("heads")
    n = len(g["coord"])
    from conftest import GOLDEN
    with open(os.path.join(GOLDEN, "param_shapes_heads.json")) as f:
        shapes = json.load(f)
    torch.manual_seed(300)
    perms = [torch.randperm(4).numpy() for _ in range(5)]
    batch = lambda feat: dict(coord=torch.from_numpy(g["coord"]).cuda(), grid_coord=torch.from_numpy(g["grid_coord"]).cuda(),
                              feat=torch.from_numpy(feat).cuda(), offset=torch.tensor([n]).cuda(), grid_size=0.02)

    def oracle(W, feat, cfg, loss_fn):
        Wg = {k: (v.clone().requires_grad_(True) if "running_" not in k else v) for k, v in W.items()}
        P.BN_TRAIN = True
        try:
            with torch.enable_grad():
                pt = P.backbone_forward.__wrapped__(Wg, g["coord"], g["grid_coord"], feat, [n], cfg=cfg, perms=list(perms))
                loss = loss_fn(Wg, pt)
                loss.backward()
        finally:
            P.BN_TRAIN = False
        return loss.detach(), {k: v.grad for k, v in Wg.items() if torch.is_tensor(v) and v.requires_grad}

    # ---- KeypointPTv3 ----
    W = {k: torch.from_numpy(v) for k, v in synth.make_weights(shapes["keypoint"], seed=0).items()}
    # two scenes (the head's BatchNorm1d needs more than one row in train()): the golden scene and a copy with
    # negated features
    target = torch.randn(2, 6, 3, generator=torch.Generator().manual_seed(1)) * 0.3
    m = M.KeypointPTv3(backbone_conf=dict(P.BASE_CFG, drop_path=0.0, compute_dtype=torch.float32))
    m.load_state_dict(W, strict=False)
    m = m.cuda().train()
    m.reg_head[3].p = 0.0                                               # nn.Dropout(0.3) is random: off for parity
    c2, g2 = np.concatenate([g["coord"]] * 2), np.concatenate([g["grid_coord"]] * 2)
    f2 = np.concatenate([g["feat4"], -g["feat4"]])
    torch.manual_seed(300)
    out = m(dict(coord=torch.from_numpy(c2).cuda(), grid_coord=torch.from_numpy(g2).cuda(),
                 feat=torch.from_numpy(f2).cuda(), offset=torch.tensor([n, 2 * n]).cuda(), grid_size=0.02,
                 target=target.cuda()))
    out["loss"].backward()

    def kp_oracle():
        Wg = {k: (v.clone().requires_grad_(True) if "running_" not in k else v) for k, v in W.items()}
        P.BN_TRAIN = True
        try:
            with torch.enable_grad():
                pt = P.backbone_forward.__wrapped__(Wg, c2, g2, f2, [n, 2 * n], perms=list(perms))
                loss = F.mse_loss(P.keypoint_head.__wrapped__(Wg, pt.feat, pt.batch), target)
                loss.backward()
        finally:
            P.BN_TRAIN = False
        return loss.detach(), {k: v.grad for k, v in Wg.items() if torch.is_tensor(v) and v.requires_grad}
    ref_loss, ref = kp_oracle()
    # BatchNorm over the TWO pooled rows of this batch turns 1e-6 feature differences into 1e-5 ones: fp32 bound 1e-4
    assert abs(out["loss"].item() - ref_loss.item()) < 1e-4 * abs(ref_loss.item())
    params = dict(m.named_parameters())
    for k in ("reg_head.6.weight", "reg_head.0.weight", "backbone.dec.dec0.block1.attn.qkv.weight",
              "backbone.enc.enc0.block0.cpe.0.weight"):
        assert rel_l2(params[k].grad, ref[k]) < 1e-3, k
    # ---- DefaultSegmentorV2 (C_in = 6, 20 classes) ----
    W = {k: torch.from_numpy(v) for k, v in synth.make_weights(shapes["segmentor"], seed=0).items()}
    labels = torch.randint(0, 20, (n,), generator=torch.Generator().manual_seed(2))
    cfg6 = dict(P.BASE_CFG, in_channels=6)
    m = M.DefaultSegmentorV2(num_classes=20, backbone_out_channels=64,
                             backbone=dict(cfg6, drop_path=0.0, compute_dtype=torch.float32))
    m.load_state_dict(W, strict=False)
    m = m.cuda().train()
    torch.manual_seed(300)
    out = m(dict(batch(g["feat6"]), segment=labels.cuda()))
    out["loss"].backward()
    ref_loss, ref = oracle(W, g["feat6"], cfg6, lambda Wg, pt: F.cross_entropy(P.seg_head.__wrapped__(Wg, pt.feat),
                                                                               labels, ignore_index=-1))
    assert abs(out["loss"].item() - ref_loss.item()) < 1e-5 * abs(ref_loss.item())
    params = dict(m.named_parameters())
    for k in ("seg_head.weight", "backbone.embedding.stem.conv.weight", "backbone.enc.enc3.block2.mlp.0.fc1.weight"):
        assert rel_l2(params[k].grad, ref[k]) < 1e-3, k


def test_eval_returns_loss_when_target_is_present(golden):
    """Validation batches carry `target`: the reference returns the loss next to the sigmoid prediction in eval() mode
    (offset_keypoint_ptv3.py:50-105).  Checked against the oracle's loss on the oracle's features."""
    from make_golden_names import make_target
    g = golden("forward")
    c, f, gc, perms = g["coord_0"], g["feat_0"], g["grid_coord_0"], g["perms_0"]
    W = _weights()
    target = torch.from_numpy(make_target(len(c), 11))
    pt = P.backbone_forward(W, c, gc, f, [len(c)], perms=list(perms))
    ref = P.offset_keypoint_loss(W, pt.feat, target)
    m = _train_model(torch.float32).eval()
    torch.manual_seed(100)
    out = m(dict(coord=torch.from_numpy(c).cuda(), grid_coord=torch.from_numpy(gc).cuda(), feat=torch.from_numpy(f).cuda(),
                 offset=torch.tensor([len(c)]).cuda(), target=target.cuda()))
    assert abs(out["loss"].item() - ref.item()) <= 1e-5 * max(1.0, abs(ref.item()))
    assert rel_l2(out["pred"], torch.from_numpy(g["pred_0"])) < 1e-4
    out2 = m(dict(coord=torch.from_numpy(c).cuda(), grid_coord=torch.from_numpy(gc).cuda(), feat=torch.from_numpy(f).cuda(),
                  offset=torch.tensor([len(c)]).cuda()))
    assert "loss" not in out2
