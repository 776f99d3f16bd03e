"""GPU parity, floating point kernels and the whole forward.
Tolerances (BASELINE.json north_star): rel-L2 <= 1e-4 for fp32, <= 1e-2 for bf16 against the fp32 oracle /
frozen reference outputs."""
import json
import os

import numpy as np
import pytest
import torch

from oracle import gridsample as G
from oracle import ops as O
from oracle import ptv3 as P
from oracle import serialization as S

pytestmark = pytest.mark.gpu

TOL_F32, TOL_BF16 = 1e-4, 1e-2


def rel_l2(a, b):
    a, b = a.detach().float().cpu(), b.detach().float().cpu()
    return ((a - b).norm() / b.norm().clamp_min(1e-30)).item()


def _scene(seed, r=6000, scale=0.6):
    from ptv3_b200 import synth
    coord, feat = synth.pig_scan(r, scale=scale, seed=seed)
    gs = G.grid_sample(coord, 0.02)
    iu = gs["idx_unique"]
    return coord[iu], feat[iu], gs["grid_coord"][iu]


def test_segment_csr_and_scatter_mean():
    from ptv3_b200 import ops, shims
    torch.manual_seed(0)
    n, c = 5000, 64
    src = torch.randn(n, c)
    cs = torch.cumsum(torch.randint(0, 9, (1500,)), 0).clamp(max=n)
    ptr = torch.cat([torch.zeros(1, dtype=torch.long), cs, torch.tensor([n])])
    perm = torch.randperm(n)
    for red in ("max", "mean", "sum", "min"):
        ref = O.segment_csr(src[perm], ptr, red)
        got = ops.segment_csr(src.cuda(), ptr.cuda(), red, gather=perm.cuda())
        assert rel_l2(got, ref) < 1e-6, red
        got = shims.segment_csr(src[perm].cuda(), ptr.cuda(), reduce=red)
        assert rel_l2(got, ref) < 1e-6, red
    got = ops.segment_csr(src.bfloat16().cuda(), ptr.cuda(), "max", gather=perm.cuda())
    assert torch.equal(got.float().cpu(), O.segment_csr(src.bfloat16().float()[perm], ptr, "max"))
    # coordinates: c == 3 goes through the generic kernel
    xyz = torch.randn(n, 3)
    assert rel_l2(ops.segment_csr(xyz.cuda(), ptr.cuda(), "mean", gather=perm.cuda()),
                  O.segment_csr(xyz[perm], ptr, "mean")) < 1e-6
    # long segments (per-scene mean pool, keypoint_ptv3.py:44)
    batch = torch.sort(torch.randint(0, 3, (n,)))[0]
    assert rel_l2(shims.scatter_mean(src.cuda(), batch.cuda(), dim=0), O.scatter_mean(src, batch)) < 1e-5


def test_unpool_add_and_ln_residual():
    from ptv3_b200 import ops
    import torch.nn.functional as F
    torch.manual_seed(1)
    n, m, c = 4000, 900, 64
    skip, up = torch.randn(n, c), torch.randn(m, c)
    inv = torch.randint(0, m, (n,))
    out, act = ops.unpool_add(skip.cuda(), up.cuda(), inv.cuda(), gelu=True)
    assert rel_l2(act, F.gelu(skip)) < 1e-6
    assert rel_l2(out, F.gelu(skip) + F.gelu(up)[inv]) < 1e-6
    out, act = ops.unpool_add(skip.bfloat16().cuda(), up.bfloat16().cuda(), inv.cuda(), gelu=True)
    assert out.dtype == torch.float32 and act.dtype == torch.bfloat16
    assert rel_l2(out, F.gelu(skip.bfloat16().float()) + F.gelu(up.bfloat16().float())[inv]) < 1e-5
    for c in (32, 64, 128, 256, 512, 768, 96):
        x, res = torch.randn(777, c) * 3 + 1, torch.randn(777, c)
        g1, b1, g2, b2 = (torch.randn(c) for _ in range(4))
        r_ref = res + F.layer_norm(x, (c,), g1, b1, 1e-5)
        o_ref = F.layer_norm(r_ref, (c,), g2, b2, 1e-5)
        r, o = ops.ln_residual(x.cuda(), g1.cuda(), b1.cuda(), 1e-5, residual=res.cuda(), gamma2=g2.cuda(),
                               beta2=b2.cuda(), out2_dtype=torch.float32)
        assert rel_l2(r, r_ref) < 1e-5 and rel_l2(o, o_ref) < 1e-5, c
        r, o = ops.ln_residual(x.cuda(), None, None, 1e-5, residual=res.cuda(), out2_dtype=torch.bfloat16)
        assert rel_l2(r, res + x) < 1e-6 and rel_l2(o, (res + x).bfloat16()) < 1e-6, c
        r, o = ops.ln_residual(x.bfloat16().cuda(), g1.cuda(), b1.cuda(), 1e-5, want_res=False, gamma2=None,
                               out2_dtype=torch.bfloat16)
        assert r is None and rel_l2(o, F.layer_norm(x.bfloat16().float(), (c,), g1, b1, 1e-5)) < 5e-3, c


@pytest.mark.parametrize("k,cin,cout", [(3, 32, 32), (3, 64, 64), (3, 48, 80), (5, 4, 32), (5, 6, 48), (3, 128, 128)])
def test_subm_conv_vs_oracle(k, cin, cout):
    from ptv3_b200 import ops
    torch.manual_seed(k * 100 + cin)
    _, _, grid = _scene(21, r=4000, scale=0.5)
    idx = torch.from_numpy(np.concatenate([np.zeros((len(grid), 1), np.int64), grid], 1).astype(np.int32))
    n = idx.shape[0]
    feat = torch.randn(n, cin)
    w = torch.randn(cout, k, k, k, cin) / (cin * k ** 3) ** 0.5
    bias = torch.randn(cout)
    nbr_ref = O.subm_neighbor_table(idx, k)
    ref = O.subm_conv(feat, nbr_ref, w, bias)
    nbr = ops.subm_build_map(idx.cuda(), k)
    wk = w.reshape(cout, k ** 3, cin).permute(1, 0, 2).contiguous()
    got = ops.subm_conv(feat.cuda(), wk.cuda(), bias.cuda(), nbr, algo=1)
    assert rel_l2(got, ref) < TOL_F32
    got = ops.subm_conv(feat.bfloat16().cuda(), wk.bfloat16().cuda(), bias.cuda(), nbr, algo=1)
    assert rel_l2(got, ref) < TOL_BF16
    got = ops.subm_conv(feat.bfloat16().cuda(), wk.bfloat16().cuda(), bias.cuda(), nbr, algo=0)
    assert rel_l2(got, ref) < TOL_BF16


@pytest.mark.parametrize("cin,cout", [(32, 32), (64, 64), (128, 128), (256, 256), (512, 512), (32, 64), (64, 32),
                                      (128, 256), (256, 128), (768, 768), (192, 1024), (768, 256)])
def test_subm_conv_tcgen05_vs_oracle(cin, cout):
    """bf16 tcgen05 implicit-GEMM kernel (algo 2 and the packed-weight entry) against the fp32 oracle evaluated on
    the bf16-rounded operands: only the final bf16 rounding of the output differs (rel-L2 <= 4e-3)."""
    from ptv3_b200 import ops
    torch.manual_seed(cin * 7 + cout)
    _, _, grid = _scene(22, r=5000, scale=0.55)
    idx = torch.from_numpy(np.concatenate([np.zeros((len(grid), 1), np.int64), grid], 1).astype(np.int32))
    n = idx.shape[0]
    assert n % 128 != 0          # ragged last tile
    feat = torch.randn(n, cin).bfloat16()
    w = (torch.randn(cout, 3, 3, 3, cin) / (cin * 14) ** 0.5).bfloat16()
    bias = torch.randn(cout)
    nbr_ref = O.subm_neighbor_table(idx, 3)
    ref = O.subm_conv(feat.float(), nbr_ref, w.float(), bias)
    nbr = ops.subm_build_map(idx.cuda(), 3)
    wk = w.reshape(cout, 27, cin).permute(1, 0, 2).contiguous().cuda()
    assert ops.subm_conv_tc_supported(cin, cout, 27)
    got = ops.subm_conv(feat.cuda(), wk, bias.cuda(), nbr, algo=2)
    assert rel_l2(got, ref) < 4e-3, rel_l2(got, ref)
    packed = ops.subm_pack_weight(wk)
    got2 = ops.subm_conv_packed(feat.cuda(), packed, bias.cuda(), nbr, cout)
    assert torch.equal(got, got2)
    got3 = ops.subm_conv_packed(feat.cuda(), packed, None, nbr, cout)
    assert rel_l2(got3, ref - bias) < 4e-3


@pytest.mark.parametrize("cin,cout", [(32, 32), (64, 64), (256, 256), (512, 512)])
def test_subm_conv_tcgen05_many_tiles(cin, cout):
    """Persistent path: several 128-row tiles per CTA (ring / accumulator phase wrap-around), ragged last tile, two
    scenes; checked against the fp32 oracle on bf16-rounded operands."""
    from ptv3_b200 import ops
    torch.manual_seed(cin + 3 * cout)
    g = torch.Generator().manual_seed(5)
    n_target = 148 * 128 * 3 + 77 if cin <= 64 else 148 * 128 + 391
    cells = torch.randint(0, 48, (n_target * 2, 3), generator=g)
    b = torch.randint(0, 2, (n_target * 2, 1), generator=g)
    idx = torch.unique(torch.cat([b, cells], 1), dim=0)[:n_target].int()
    idx = idx[torch.randperm(idx.shape[0], generator=g)].contiguous()
    n = idx.shape[0]
    assert n > 148 * 128 and n % 128 != 0
    feat = torch.randn(n, cin).bfloat16()
    w = (torch.randn(cout, 3, 3, 3, cin) / (cin * 14) ** 0.5).bfloat16()
    bias = torch.randn(cout)
    ref = O.subm_conv(feat.float(), O.subm_neighbor_table(idx, 3), w.float(), bias)
    nbr = ops.subm_build_map(idx.cuda(), 3)
    wk = w.reshape(cout, 27, cin).permute(1, 0, 2).contiguous().cuda()
    got = ops.subm_conv_packed(feat.cuda(), ops.subm_pack_weight(wk), bias.cuda(), nbr, cout)
    assert rel_l2(got, ref) < 4e-3, rel_l2(got, ref)
    assert torch.isfinite(got.float()).all()


@pytest.mark.parametrize("c,n", [(32, 1000), (64, 777), (32, 148 * 128 * 2 + 5), (64, 148 * 128 * 3 + 129), (128, 1),
                                 (128, 1000), (128, 148 * 128 * 3 + 129), (128, 148 * 128 + 1), (256, 1), (256, 3000),
                                 (256, 148 * 128 * 3 + 129)])
def test_mlp_fused_tcgen05(c, n):
    """fc1 -> GELU(erf) -> fc2 (+ residual) in one tcgen05 kernel against the torch fp32 reference on the same
    bf16-rounded operands (hidden activation rounded to bf16 like the reference under autocast): rel-L2 <= 4e-3."""
    import torch.nn.functional as F
    from ptv3_b200 import ops
    torch.manual_seed(c + n)
    h = 4 * c
    x = torch.randn(n, c).bfloat16()
    w1 = (torch.randn(h, c) / c ** 0.5).bfloat16()
    w2 = (torch.randn(c, h) / h ** 0.5).bfloat16()
    b1, b2 = torch.randn(h) * 0.5, torch.randn(c) * 0.5
    res = torch.randn(n, c)
    hid = F.gelu(x.float() @ w1.float().t() + b1).bfloat16().float()
    ref = hid @ w2.float().t() + b2
    assert ops.mlp_fused_supported(c, h) and not ops.mlp_fused_supported(512, 2048)
    w1p, w2p = ops.pack_kmajor(w1.cuda(), ops.mlp_fc1_tile_rows(c)), ops.pack_kmajor(w2.cuda(), c)
    r, r16 = ops.mlp_fused(x.cuda(), w1p, b1.cuda(), w2p, b2.cuda(), h, residual=res.cuda())
    assert rel_l2(r - res.cuda(), ref) < 4e-3, rel_l2(r - res.cuda(), ref)
    assert torch.equal(r16, r.bfloat16())
    r2, r216 = ops.mlp_fused(x.cuda(), w1p, b1.cuda(), w2p, b2.cuda(), h, residual=None, want_res=False)
    assert r2 is None and rel_l2(r216, ref) < 6e-3


@pytest.mark.parametrize("c,n", [(32, 1), (64, 127), (32, 5000), (64, 40001), (32, 148 * 128 * 4 + 3),
                                 (64, 148 * 128 * 5 + 77)])
@pytest.mark.parametrize("drop", [False, True])
def test_proj_mlp_fused_tcgen05(c, n, drop):
    """attn.proj -> (+ DropPath) + shortcut -> LayerNorm -> fc1 -> GELU(erf) -> fc2 -> (+ DropPath) + shortcut in one
    tcgen05 kernel against the torch fp32 reference on the same bf16-rounded operands (LayerNorm output and hidden
    activation rounded to bf16 like the reference under autocast): rel-L2 <= 4e-3 on the two branches."""
    import torch.nn.functional as F
    from ptv3_b200 import ops
    torch.manual_seed(c * 3 + n)
    h = 4 * c
    a = torch.randn(n, c).bfloat16()
    wp = (torch.randn(c, c) / c ** 0.5).bfloat16()
    w1 = (torch.randn(h, c) / c ** 0.5).bfloat16()
    w2 = (torch.randn(c, h) / h ** 0.5).bfloat16()
    bp, b1, b2 = torch.randn(c) * 0.5, torch.randn(h) * 0.5, torch.randn(c) * 0.5
    g, be = torch.rand(c) + 0.5, torch.randn(c) * 0.2
    x1 = torch.randn(n, c) * 2.0 + 0.7
    ra = (torch.rand(n) > 0.3).float() / 0.7 if drop else None
    rm = (torch.rand(n) > 0.2).float() / 0.8 if drop else None
    br1 = a.float() @ wp.float().t() + bp
    x2 = x1 + (br1 * ra[:, None] if drop else br1)
    hn = F.layer_norm(x2, (c,), g, be, 1e-5).bfloat16().float()
    hid = F.gelu(hn @ w1.float().t() + b1).bfloat16().float()
    br2 = hid @ w2.float().t() + b2
    ref = x2 + (br2 * rm[:, None] if drop else br2)
    cu = lambda t: t.cuda() if t is not None else None
    out, out16 = ops.proj_mlp_fused(a.cuda(), ops.pack_kmajor(wp.cuda(), c), bp.cuda(), x1.cuda(), g.cuda(), be.cuda(),
                                    1e-5, ops.pack_kmajor(w1.cuda(), 128), b1.cuda(), ops.pack_kmajor(w2.cuda(), c),
                                    b2.cuda(), h, rs_attn=cu(ra), rs_mlp=cu(rm))
    assert rel_l2(out - x1.cuda(), ref - x1) < 4e-3, rel_l2(out - x1.cuda(), ref - x1)
    assert torch.equal(out16, out.bfloat16())
    # and against the separate launches it replaces (proj output rounded to bf16 there)
    p16 = F.linear(a.cuda(), wp.cuda(), bp.bfloat16().cuda())
    x2k, hk = ops.ln_residual(p16, None, None, 1e-5, residual=x1.cuda(), gamma2=g.cuda(), beta2=be.cuda(),
                              out2_dtype=torch.bfloat16, row_scale=cu(ra))
    chain, _ = ops.mlp_fused(hk, ops.pack_kmajor(w1.cuda(), 128), b1.cuda(), ops.pack_kmajor(w2.cuda(), c), b2.cuda(),
                             h, residual=x2k, row_scale=cu(rm))
    assert rel_l2(out - x1.cuda(), chain - x1.cuda()) < 8e-3


def test_attention_golden(golden):
    """Reference SerializedAttention.forward (dense branch, one scene) frozen in tests/golden/attention.npz."""
    from ptv3_b200 import ops
    g = golden("attention")
    qkv = torch.from_numpy(g["qkv"]).cuda()
    order = torch.from_numpy(g["order"]).cuda()
    off = torch.from_numpy(g["offset"]).cuda()
    H, K = int(g["H"]), int(g["K"])
    w, nw = ops.build_windows(off, K, qkv.shape[0])
    ref = torch.from_numpy(g["out"])
    pw, pb = torch.from_numpy(g["proj_w"]).cuda(), torch.from_numpy(g["proj_b"]).cuda()
    a = ops.attn_window(qkv, order, w, nw, H, K, algo=1)
    assert rel_l2(a @ pw.t() + pb, ref) < TOL_F32
    a = ops.attn_window(qkv.bfloat16(), order, w, nw, H, K, algo=1)
    assert rel_l2(a.float() @ pw.t() + pb, ref) < TOL_BF16
    for algo in (0, 2, 3, 4, 5):
        a = ops.attn_window(qkv.bfloat16(), order, w, nw, H, K, algo=algo)
        assert rel_l2(a.float() @ pw.t() + pb, ref) < TOL_BF16, algo


@pytest.mark.parametrize("counts,K,H", [([5], 16, 1), ([1024], 1024, 2), ([1025], 1024, 2), ([700, 2300, 64, 1100], 1024, 4),
                                        ([300, 17, 129], 128, 32), ([3000], 1024, 3)])
def test_attention_ragged_vs_oracle(counts, K, H):
    """Ragged scenes: short windows, the overlapping last window, several heads; all three kernels."""
    from ptv3_b200 import ops, shims
    torch.manual_seed(sum(counts) + K)
    n, C = sum(counts), H * 16
    qkv = torch.randn(n, 3 * C) * 1.5
    offset = np.cumsum(counts)
    order = torch.cat([torch.randperm(c) + s for c, s in zip(counts, np.concatenate([[0], offset[:-1]]))])
    inverse = torch.empty_like(order)
    inverse[order] = torch.arange(n)
    ref = O.window_attention(qkv, order, inverse, offset, H, K)
    w, nw = ops.build_windows(torch.from_numpy(offset).cuda(), K, n)
    got = ops.attn_window(qkv.cuda(), order.cuda(), w, nw, H, K, algo=1)
    assert rel_l2(got, ref) < TOL_F32
    qb = qkv.bfloat16()
    refb = O.window_attention(qb.float(), order, inverse, offset, H, K)
    for algo in (1, 2, 3, 4, 5) if K <= 1024 else (1, 2, 4):
        got = ops.attn_window(qb.cuda(), order.cuda(), w, nw, H, K, algo=algo)
        assert rel_l2(got, refb) < 6e-3, (algo, rel_l2(got, refb))       # bf16 P and bf16 output rounding
    for v in (111, 231, 212, 232, 113, 133):                                        # default kernel: ping-pong off / on, 0 or 2/8 of
        ops.debug_set(7, v)                                               # the exponentials on the polynomial, both chunkings
        try:
            got = ops.attn_window(qb.cuda(), order.cuda(), w, nw, H, K, algo=0)
        finally:
            ops.debug_set(7, 0)
        e4 = rel_l2(got, refb)
        assert e4 < 4e-3, (v, e4)                                         # f16 P: only the bf16 output rounding is left
    assert ops.device_error() == 0
    # flash_attn call surface (:208-215): padded layout + cu_seqlens, identity order
    pad, unpad, cu = O.padding_and_inverse(offset, K)
    g = qb[order[torch.from_numpy(pad)]].reshape(-1, 3, H, 16).cuda()
    out = shims.flash_attn_varlen_qkvpacked_func(g, torch.from_numpy(cu).cuda(), max_seqlen=K, dropout_p=0.0,
                                                 softmax_scale=16 ** -0.5)
    out = out.reshape(-1, C)[torch.from_numpy(unpad).cuda()[inverse.cuda()]]
    assert rel_l2(out, refb) < 6e-3


@pytest.mark.parametrize("gain", [4.0, 12.0])
def test_attention_moving_maximum(gain):
    """Lazy rescaling of the tensor-core kernel: keys whose scores keep growing along the window (the reference maximum
    moves in jumps both below and far above the 2^8 threshold) and peaked softmaxes, against the fp32 oracle."""
    from ptv3_b200 import ops
    torch.manual_seed(3)
    counts, K, H = [2048, 900], 1024, 2
    n, C = sum(counts), H * 16
    qkv = torch.randn(n, 3, H, 16)
    ramp = torch.linspace(0.0, 1.0, n)[:, None, None]
    qkv[:, 0] = qkv[:, 0].abs() * gain                     # positive queries ...
    qkv[:, 1] = qkv[:, 1].abs() * (0.2 + ramp * gain)      # ... and keys that grow with the row index
    qkv = qkv.reshape(n, 3 * C)
    offset = np.cumsum(counts)
    order = torch.arange(n)
    qb = qkv.bfloat16()
    refb = O.window_attention(qb.float(), order, order.clone(), offset, H, K)
    w, nw = ops.build_windows(torch.from_numpy(offset).cuda(), K, n)
    for algo in (2, 4, 5):     # 5: the tcgen05 kernel keeps the first chunk's maximum as the reference; rows whose
        got = ops.attn_window(qb.cuda(), order.cuda(), w, nw, H, K, algo=algo)   # scores run away from it by more
        assert torch.isfinite(got).all()                                          # than ~100 octaves take the fallback
        assert rel_l2(got, refb) < 6e-3, (algo, rel_l2(got, refb))


def test_attention_large_scores_and_tiny_values():
    """The default kernel stages Q (pre-scaled), K and V as f16: scores far outside the f16 range of a single product
    (|q.k| ~ 1e4), very negative rows (every score << the first chunk's maximum) and tiny V stay correct."""
    from ptv3_b200 import ops
    torch.manual_seed(5)
    counts, K, H = [1500], 1024, 2
    n, C = sum(counts), H * 16
    qkv = torch.randn(n, 3, H, 16)
    qkv[:, 0] *= 30.0
    qkv[:, 1] *= 30.0                                       # scores ~ N(0, 3600 * 4): softmax is one-hot-ish
    qkv[: n // 2, 2] *= 1e-3                                # small values next to O(1) values
    qkv = qkv.reshape(n, 3 * C)
    offset = np.cumsum(counts)
    order = torch.randperm(n)
    inverse = torch.empty_like(order)
    inverse[order] = torch.arange(n)
    qb = qkv.bfloat16()
    refb = O.window_attention(qb.float(), order, inverse, offset, H, K)
    w, nw = ops.build_windows(torch.from_numpy(offset).cuda(), K, n)
    got5 = ops.attn_window(qb.cuda(), order.cuda(), w, nw, H, K, algo=5)
    assert torch.isfinite(got5).all() and rel_l2(got5, refb) < 2e-2, rel_l2(got5, refb)
    got = ops.attn_window(qb.cuda(), order.cuda(), w, nw, H, K, algo=4)
    assert torch.isfinite(got).all()
    # one-hot softmaxes amplify the f16 rounding of Q * scale * log2(e) (2^-11 relative on scores of ~1e2 log2 units)
    assert rel_l2(got, refb) < 2e-2, rel_l2(got, refb)


def test_attention_window_contract_flag():
    """a window whose keys exceed max_len (or whose queries leave the key range) must not leave
    garbage silently: the CTA returns and ptv3_device_error() reports it."""
    from ptv3_b200 import ops
    n, H = 2048, 2
    qkv = torch.randn(n, 3 * H * 16).bfloat16().cuda()
    order = torch.arange(n).cuda()
    good = torch.tensor([[0, 1024, 0, 1024], [1024, 1024, 1024, 1024]], dtype=torch.int64).cuda()
    nw = torch.tensor([2], dtype=torch.int64).cuda()
    ops.attn_window(qkv, order, good, nw, H, 1024)
    assert ops.device_error() == 0
    bad = torch.tensor([[0, 1024, 0, 2048], [1024, 1024, 0, 512]], dtype=torch.int64).cuda()
    ops.attn_window(qkv, order, bad, nw, H, 1024)          # k_len > max_len ; queries outside the keys
    assert ops.device_error() == 1
    assert ops.device_error() == 0                          # cleared by the read


def test_block_native_call_equals_op_by_op():
    """ptv3_block_fwd (one C-ABI call per Block: conv -> cuBLASLt Linear -> LN ... -> MLP) against the op-by-op path over
    the same kernels, for a fused-MLP width (C = 32 / 64) and a cuBLAS-MLP width (C = 128), on a two-scene batch."""
    from ptv3_b200 import ops
    m = _model(torch.bfloat16)
    scenes = [_scene(71, 6000, 0.6), _scene(72, 2500, 0.35)]
    cat = lambda j: torch.from_numpy(np.concatenate([s[j] for s in scenes])).cuda()
    offset = torch.tensor(np.cumsum([len(s[0]) for s in scenes])).cuda()
    from ptv3_b200.model import Block
    feats, fuse_default = {}, Block.fuse_proj
    for native, fuse in ((True, False), (True, True), (False, False)):
        ops.NATIVE_BLOCK, Block.fuse_proj = native, fuse
        try:
            hooks = []
            for name, mod in (("enc0", m.backbone.enc.enc0), ("enc2", m.backbone.enc.enc2)):
                hooks.append(mod.register_forward_hook(
                    lambda _m, _i, o, name=name, key=(native, fuse): feats.__setitem__((name, key), o.feat.clone())))
            torch.manual_seed(9)
            pred = m(dict(coord=cat(0), feat=cat(1), grid_coord=cat(2), offset=offset))["pred"]
            feats[("pred", (native, fuse))] = pred.clone()
            for h in hooks:
                h.remove()
        finally:
            ops.NATIVE_BLOCK, Block.fuse_proj = True, fuse_default
    for k in ("enc0", "enc2", "pred"):
        e = rel_l2(feats[(k, (True, False))], feats[(k, (False, False))])
        assert e < 2e-3, (k, e)       # same kernels; the GEMMs may pick another cuBLASLt algorithm (bf16 rounding)
        # proj + LayerNorm + MLP in one kernel (C = 32): attn.proj's output is no longer rounded to bf16 on the way, so
        # these are two differently rounded bf16 paths (each is held to the oracle by the golden / full-size tests)
        e = rel_l2(feats[(k, (True, True))], feats[(k, (False, False))])
        assert e < 1e-2, (k, e)


def test_subm_map_site_range_flag():
    """the 64-bit site key holds 10 bits of batch id and 18-bit coordinates; sites outside that range
    used to alias silently.  They are now left out of the map and reported through ptv3_device_error()."""
    from ptv3_b200 import ops
    idx = torch.tensor([[0, 1, 2, 3], [0, 1, 2, 4], [3, 5, 5, 5]], dtype=torch.int32).cuda()
    nbr = ops.subm_build_map(idx, 3)
    assert ops.device_error() == 0 and int(nbr[0, 14]) == 1 and int(nbr[1, 12]) == 0
    bad = torch.tensor([[0, 1, 2, 3], [1024, 1, 2, 3], [0, 1, 2, 1 << 18]], dtype=torch.int32).cuda()
    nbr = ops.subm_build_map(bad, 3)
    assert ops.device_error() == 2
    assert int(nbr[0, 13]) == 0 and (nbr[0] >= 0).sum().item() == 1     # the valid site sees only itself
    assert ops.device_error() == 0


def _weights():
    from conftest import GOLDEN
    from ptv3_b200 import synth
    with open(os.path.join(GOLDEN, "param_shapes.json")) as f:
        shapes = json.load(f)
    return {k: torch.from_numpy(v) for k, v in synth.make_weights(shapes, seed=0).items()}


def _model(dtype):
    from ptv3_b200 import model as M
    m = M.OffsetKeypointPTv3(backbone_conf=dict(P.BASE_CFG, compute_dtype=dtype)).eval()
    missing, unexpected = m.load_state_dict(_weights(), strict=False)
    assert not unexpected and all(k.endswith("num_batches_tracked") for k in missing)
    return m.cuda()


def _hooks(m, feats):
    bb = m.backbone
    for name, mod in (("stem", bb.embedding), ("enc0", bb.enc.enc0), ("pool1", bb.enc.enc1.down),
                      ("enc4", bb.enc.enc4), ("dec3", bb.dec.dec3)):
        mod.register_forward_hook(lambda _m, _i, o, name=name: feats.__setitem__(name, o.feat.detach().clone()))


@pytest.mark.parametrize("dtype,tol", [(torch.float32, TOL_F32), (torch.bfloat16, TOL_BF16)])
def test_forward_golden(golden, dtype, tol):
    """Whole OffsetKeypointPTv3 forward against the UNMODIFIED reference's frozen outputs (tests/golden/forward.npz),
    same weights, same torch.randperm draws (5 per forward, CPU generator)."""
    g = golden("forward")
    m = _model(dtype)
    feats = {}
    _hooks(m, feats)
    for i in range(int(g["n_scenes"])):
        n = len(g[f"coord_{i}"])
        torch.manual_seed(100 + i)
        out = m(dict(coord=torch.from_numpy(g[f"coord_{i}"]).cuda(), grid_coord=torch.from_numpy(g[f"grid_coord_{i}"]).cuda(),
                     feat=torch.from_numpy(g[f"feat_{i}"]).cuda(), offset=torch.tensor([n]).cuda(), grid_size=0.02))
        for k in ("stem", "enc0", "pool1", "enc4", "dec3"):
            assert rel_l2(feats[k], torch.from_numpy(g[f"{k}_{i}"])) < tol, (i, k)
        ref = torch.from_numpy(g[f"pred_{i}"])
        assert rel_l2(out["pred"], ref) < tol, i
        # keypoint positions (tools/infer_offset.py:366-399), 1 mm at the reference scale of 874 mm / unit
        coord = torch.from_numpy(g[f"coord_{i}"])
        kp_ref = P.aggregate_keypoints(coord, ref, mode="weighted")
        pred = out["pred"].float().cpu()
        if dtype == torch.float32:
            kp = P.aggregate_keypoints(coord, pred, mode="weighted")
            assert (kp - kp_ref).norm(dim=1).max().item() * 873.69 < 1.0
        # offsets voted through the SAME point selection as the reference (with random weights the mask sits around
        # the 0.5 threshold, so re-selecting with bf16 masks measures the synthetic weights, not the kernels)
        mixed = torch.cat([pred[..., :3], ref[..., 3:]], dim=-1)
        kp = P.aggregate_keypoints(coord, mixed, mode="weighted")
        err = (kp - kp_ref).norm(dim=1).max().item()
        mag = ref[..., :3].abs().max().item()     # synthetic weights: offsets are O(mag) units, not pig-sized (<0.3)
        print(f"scene {i} {dtype}: pred rel-L2 {rel_l2(pred, ref):.2e}, keypoint error {err * 873.69:.3f} mm "
              f"at offset magnitude {mag:.2f} units ({err / mag:.2e} relative)")
        if dtype == torch.float32:
            assert err * 873.69 < 1.0
        else:
            assert err / mag < 1e-2      # = 1 mm for a real pig whose offsets stay below ~0.1 units (87 mm)


def test_forward_batched_equals_per_scene():
    """Scenes are independent units: a batch of 3 equals 3 single-scene forwards."""
    m = _model(torch.float32)
    m.backbone.shuffle_orders = False
    for mod in m.backbone.modules():
        if hasattr(mod, "shuffle_orders"):
            mod.shuffle_orders = False
    scenes = [_scene(31, 7000, 0.7), _scene(32, 1500, 0.3), _scene(33, 4000, 0.5)]
    singles = []
    for c, f, gcoord in scenes:
        singles.append(m(dict(coord=torch.from_numpy(c).cuda(), grid_coord=torch.from_numpy(gcoord).cuda(),
                              feat=torch.from_numpy(f).cuda(), offset=torch.tensor([len(c)]).cuda()))["pred"])
    cat = lambda j: torch.from_numpy(np.concatenate([s[j] for s in scenes])).cuda()
    offset = torch.tensor(np.cumsum([len(s[0]) for s in scenes])).cuda()
    batched = m(dict(coord=cat(0), feat=cat(1), grid_coord=cat(2), offset=offset))["pred"]
    assert rel_l2(batched, torch.cat(singles)) < 1e-5


def test_forward_vs_oracle_fresh_scene():
    """A scene that is NOT in the golden set, against the oracle restatement run here on the CPU."""
    m = _model(torch.float32)
    W = _weights()
    c, f, gcoord = _scene(41, 5000, 0.55)
    torch.manual_seed(7)
    perms = [torch.randperm(4).numpy() for _ in range(5)]
    pt = P.backbone_forward(W, c, gcoord, f, [len(c)], perms=list(perms))
    ref = P.offset_head(W, pt.feat)
    torch.manual_seed(7)
    out = m(dict(coord=torch.from_numpy(c).cuda(), grid_coord=torch.from_numpy(gcoord).cuda(),
                 feat=torch.from_numpy(f).cuda(), offset=torch.tensor([len(c)]).cuda()))["pred"]
    assert rel_l2(out, ref) < TOL_F32


@pytest.mark.parametrize("mode", ["argmax", "weighted"])
def test_keypoint_aggregate_matches_oracle(mode):
    """A9: ptv3_keypoint_aggregate against oracle.ptv3.aggregate_keypoints (tools/infer_offset.py:366-399) per scene:
    ragged scenes (1 row, no confident row, exact ties -> first index wins)."""
    from ptv3_b200 import ops
    g = torch.Generator().manual_seed(5)
    sizes = [1, 777, 5000, 33, 12345]
    n, K = sum(sizes), 6
    coord = torch.randn(n, 3, generator=g)
    pred = torch.randn(n, K, 4, generator=g) * 0.2
    pred[..., 3] = torch.sigmoid(torch.randn(n, K, generator=g) * 2)
    s3 = sizes[0] + sizes[1] + sizes[2]
    pred[s3:s3 + sizes[3], :, 3] = 0.1                       # scene 3: nothing above the threshold, all tied
    pred[5:9, 2, 3] = 0.999                                   # exact ties inside scene 1: first index wins
    offset = torch.tensor(sizes).cumsum(0)
    out = ops.keypoint_aggregate(coord.cuda(), pred.cuda(), offset.cuda(), mode=mode).cpu()
    assert out.shape == (len(sizes), K, 3)
    start = 0
    for b, sz in enumerate(sizes):
        ref = P.aggregate_keypoints(coord[start:start + sz], pred[start:start + sz], mode=mode)
        assert torch.allclose(out[b], ref, rtol=1e-5, atol=1e-5), (b, (out[b] - ref).abs().max())
        start += sz


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("act", ["none", "gelu", "relu"])
def test_affine_act(dtype, act):
    """ptv3_affine_act = act(x * s + t) (eval BatchNorm + GELU after pooling, :433-437) against torch fp32."""
    from ptv3_b200 import ops
    g = torch.Generator().manual_seed(11)
    x = torch.randn(1003, 64, generator=g).to(dtype)
    s, t = torch.randn(64, generator=g), torch.randn(64, generator=g)
    ref = x.float() * s + t
    ref = {"none": ref, "gelu": torch.nn.functional.gelu(ref), "relu": torch.relu(ref)}[act]
    o32, o16 = ops.affine_act(x.cuda(), s.cuda(), t.cuda(), act, want_f32=True, want_bf16=True)
    assert torch.allclose(o32.cpu(), ref, rtol=1e-5, atol=1e-6)
    assert torch.equal(o16.cpu(), o32.cpu().bfloat16())


@pytest.mark.parametrize("n,c", [(1, 8), (777, 512), (1003, 2048), (4097, 3072)])
def test_gelu_bf16_plain(n, c):
    """The stand-alone bf16 GELU between the library GEMMs of the wide MLPs (MLP.forward :237-247): exact-erf GELU of the
    bf16 input in fp32, rounded once to bf16 (two vectors per thread + tail; in place as the model calls it)."""
    from ptv3_b200 import ops
    g = torch.Generator().manual_seed(n + c)
    x = (torch.randn(n, c, generator=g) * 2.5).bfloat16()
    ref = torch.nn.functional.gelu(x.float()).bfloat16()
    _, o16 = ops.affine_act(x.cuda(), None, None, "gelu", want_f32=False, want_bf16=True)
    o = o16.cpu()
    # the 15-instruction erf is within 1.5e-7 ABSOLUTE of erf: at most one bf16 ulp everywhere, rounding flips only on a
    # handful of ties where the value is not tiny (deep in the negative tail, |gelu| < 1e-3, 4e-7 is a visible fraction)
    assert torch.allclose(o.float(), ref.float(), rtol=8e-3, atol=1e-6)
    big = ref.float().abs() > 1e-3
    assert (o != ref)[big].float().mean() < 1e-3


def _head_weights(tag):
    from conftest import GOLDEN
    from ptv3_b200 import synth
    with open(os.path.join(GOLDEN, "param_shapes_heads.json")) as f:
        shapes = json.load(f)[tag]
    return {k: torch.from_numpy(v) for k, v in synth.make_weights(shapes, seed=0).items()}


def _load(m, W):
    missing, unexpected = m.load_state_dict(W, strict=False)
    assert not unexpected and all(k.endswith("num_batches_tracked") for k in missing), (missing, unexpected)
    return m.eval().cuda()


@pytest.mark.parametrize("dtype,tol", [(torch.float32, TOL_F32), (torch.bfloat16, TOL_BF16)])
def test_keypoint_head_golden(golden, dtype, tol):
    """A9: KeypointPTv3 (per-scene mean pool -> regression MLP) against the unmodified reference's frozen output, and
    a 3-scene batch against per-scene oracle runs (ragged scenes through scatter_mean)."""
    from ptv3_b200 import model as M
    g = golden("heads")
    n = len(g["coord"])
    W = _head_weights("keypoint")
    m = _load(M.KeypointPTv3(backbone_conf=dict(P.BASE_CFG, compute_dtype=dtype)), W)
    torch.manual_seed(300)
    out = m(dict(coord=torch.from_numpy(g["coord"]).cuda(), grid_coord=torch.from_numpy(g["grid_coord"]).cuda(),
                 feat=torch.from_numpy(g["feat4"]).cuda(), offset=torch.tensor([n]).cuda(), grid_size=0.02))["pred"]
    ref = torch.from_numpy(g["out_keypoint"])
    assert out.shape == (1, 6, 3) and rel_l2(out, ref) < tol, rel_l2(out, ref)
    if dtype == torch.float32:
        for mod in m.backbone.modules():
            if hasattr(mod, "shuffle_orders"):
                mod.shuffle_orders = False
        scenes = [_scene(51, 3000, 0.45), _scene(52, 900, 0.25), _scene(53, 2000, 0.4)]
        refs = []
        for c, f, gc in scenes:
            pt = P.backbone_forward(W, c, gc, f, [len(c)], perms=None)
            refs.append(P.keypoint_head(W, pt.feat, pt.batch))
        cat = lambda j: torch.from_numpy(np.concatenate([s[j] for s in scenes])).cuda()
        offset = torch.tensor(np.cumsum([len(s[0]) for s in scenes])).cuda()
        out = m(dict(coord=cat(0), feat=cat(1), grid_coord=cat(2), offset=offset))["pred"]
        assert out.shape == (3, 6, 3) and rel_l2(out, torch.cat(refs)) < tol


@pytest.mark.parametrize("dtype,tol", [(torch.float32, TOL_F32), (torch.bfloat16, TOL_BF16)])
def test_segmentor_head_golden(golden, dtype, tol):
    """A9 / BASELINE configs[2] shape: DefaultSegmentorV2 over PT-v3m1 with C_in = 6 and 20 classes against the
    unmodified reference's frozen logits; predicted classes agree wherever the reference's margin exceeds the
    tolerance."""
    from ptv3_b200 import model as M
    g = golden("heads")
    n = len(g["coord"])
    W = _head_weights("segmentor")
    m = _load(M.DefaultSegmentorV2(num_classes=20, backbone_out_channels=64,
                                   backbone=dict(P.BASE_CFG, in_channels=6, compute_dtype=dtype)), W)
    torch.manual_seed(300)
    out = m(dict(coord=torch.from_numpy(g["coord"]).cuda(), grid_coord=torch.from_numpy(g["grid_coord"]).cuda(),
                 feat=torch.from_numpy(g["feat6"]).cuda(), offset=torch.tensor([n]).cuda(), grid_size=0.02))
    logits = out["seg_logits"].float().cpu()
    ref = torch.from_numpy(g["out_segmentor"])
    assert logits.shape == (n, 20) and rel_l2(logits, ref) < tol, rel_l2(logits, ref)
    top2 = ref.topk(2, dim=1).values
    clear = (top2[:, 0] - top2[:, 1]) > 4 * tol * ref.abs().max()
    assert clear.any() and torch.equal(logits.argmax(1)[clear], ref.argmax(1)[clear])


def test_serving_paths_on_device(golden):
    """collate, the Qt server's detect path (normalise -> first point per voxel -> model -> vote ->
    de-normalise) and test-mode fragment voting, each against the same steps done with the oracle pieces."""
    from ptv3_b200 import infer, model as M, synth, ops
    m = _model(torch.float32)
    for mod in m.modules():
        if hasattr(mod, "shuffle_orders"):
            mod.shuffle_orders = False
    coord, feat = synth.pig_scan(20000, scale=350.0, seed=77)          # a scan in millimetres
    coord = coord + np.array([1200.0, -300.0, 80.0], dtype=np.float32)
    kp = infer.detect_keypoints(m, torch.from_numpy(coord).cuda(), torch.from_numpy(feat).cuda(), 0.02, mode="argmax")
    # the same steps with the oracle's voxeliser / aggregation around OUR model's prediction.  The normalised
    # coordinates are taken from the device function (a 1e-7 difference in the fp32 centroid would move points across
    # voxel borders and, with random weights, the arg-max vote to another point)
    from ptv3_b200 import transform
    cn_t, cen_t, scale_t = transform.normalize_unit_sphere(torch.from_numpy(coord).cuda())
    cn = cn_t.cpu().numpy()
    idx, grid = G.voxelize_first(cn, 0.02, keep_order=True)
    pred = m(dict(coord=torch.from_numpy(cn[idx]).cuda(), grid_coord=torch.from_numpy(grid).cuda(),
                  feat=torch.from_numpy(feat[idx]).cuda(), offset=torch.tensor([len(idx)]).cuda()))["pred"].cpu()
    ref = P.aggregate_keypoints(torch.from_numpy(cn[idx]), pred, mode="argmax") * float(scale_t) + cen_t.cpu()
    assert abs(float(scale_t) - np.linalg.norm(coord - coord.mean(0), axis=1).max()) < 0.05
    assert kp.shape == (6, 3) and (kp.cpu() - ref).abs().max().item() < 0.05      # millimetres
    # collate: two scenes -> offsets, concatenation
    a = dict(coord=torch.randn(5, 3).cuda(), feat=torch.randn(5, 4).cuda(), grid_size=0.02)
    b = dict(coord=torch.randn(7, 3).cuda(), feat=torch.randn(7, 4).cuda(), grid_size=0.02)
    c = infer.collate([a, b])
    assert c["offset"].tolist() == [5, 12] and c["coord"].shape == (12, 3) and c["grid_size"] == 0.02
    # fragment voting with a segmentor: every raw point gets exactly one vote per fragment it belongs to
    seg = M.DefaultSegmentorV2(num_classes=5, backbone_out_channels=64,
                               backbone=dict(P.BASE_CFG, in_channels=4, compute_dtype=torch.float32,
                                             shuffle_orders=False)).eval().cuda()
    c2, f2 = synth.pig_scan(12000, scale=0.5, seed=78)
    data = dict(coord=torch.from_numpy(c2).cuda(), feat=torch.from_numpy(f2).cuda())
    pred, prob = infer.segment_with_fragments(seg, data, 0.02, 5)
    assert pred.shape == (12000,) and prob.shape == (12000, 5)
    gs = ops.gridsample(data["coord"], 0.02, "fnv")
    votes = prob.sum(1)                                           # = number of fragments that contained the point
    cnt = gs["count"][gs["inverse"]]                              # population of the point's voxel
    kmax = int(gs["count"].max())
    # fragment i takes point (i % count) of every voxel: a point at position p of a voxel with c points is picked
    # by the fragments i with i % c == p -> floor((kmax - 1 - p) / c) + 1 votes; summed over a voxel that is kmax
    assert torch.allclose(votes.sum(), torch.tensor(float(kmax * gs["n_voxels"]), device=votes.device), rtol=1e-5)
    assert float(votes.min()) >= 1.0 - 1e-4 and bool((votes <= torch.ceil(kmax / cnt.float()) + 1e-4).all())


def test_fused_stem_equals_table_path():
    """Embedding with the neighbour search fused into the stem conv (no (N,125) table) -- by block-occupancy search over
    the z-order serialization (8 probes per row) and by 125 hash probes per row -- against the table path: same present
    offsets in the same order, so the fp32 outputs are bit-identical; plus the oracle."""
    from ptv3_b200 import ops
    m = _model(torch.bfloat16)
    emb = m.backbone.embedding
    from ptv3_b200.structure import Point
    scenes = [_scene(61, 6000, 0.6), _scene(62, 2500, 0.35)]
    cat = lambda j: torch.from_numpy(np.concatenate([s[j] for s in scenes])).cuda()
    offset = torch.tensor(np.cumsum([len(s[0]) for s in scenes])).cuda()
    outs = []
    torch.manual_seed(5)
    for fuse, block, order, shuffle in ((True, True, ["hilbert", "z-trans", "z", "hilbert-trans"], True),
                                        (True, True, "z", False), (True, False, "z", False), (False, False, "z", False)):
        emb.fuse_lookup, emb.block_search = fuse, block
        p = Point(coord=cat(0), feat=cat(1), grid_coord=cat(2), offset=offset)
        p.serialization(order=order, shuffle_orders=shuffle)
        assert p["_z_row"] is not None
        p.sparsify()
        p = emb(p)
        outs.append((p.feat.clone(), p["_feat_cd"].clone()))
    emb.fuse_lookup = emb.block_search = True
    assert ops.device_error() == 0
    for o in outs[1:]:
        assert torch.equal(outs[0][0], o[0]) and torch.equal(outs[0][1], o[1])
    assert outs[0][1].dtype == torch.bfloat16 and outs[0][0].dtype == torch.float32
    W = {k[len("backbone."):]: v for k, v in _weights().items() if k.startswith("backbone.")}
    idx = torch.cat([torch.repeat_interleave(torch.arange(2), torch.tensor([len(s[0]) for s in scenes]))[:, None],
                     cat(2).cpu()], 1).int()
    ref = torch.nn.functional.gelu(P._bn(O.subm_conv(cat(1).cpu(), O.subm_neighbor_table(idx, 5),
                                                     W["embedding.stem.conv.weight"], None), W, "embedding.stem.norm"))
    assert rel_l2(outs[0][0], ref) < 1e-5


def test_keypoint_eval_matches_oracle_and_reference(golden):
    """A10: ptv3_keypoint_eval (per-(scene, keypoint) distances + device-resident running sums) against
    oracle.ptv3.offset_keypoint_eval and the unmodified evaluator's logged means (tests/golden/eval.npz): skipped
    keypoints, a scene without valid keypoints, a tied predicted mask, batches with and without `scale`."""
    from make_golden_names import eval_batches
    from ptv3_b200 import ops
    from ptv3_b200.evaluator import summarize
    g = golden("eval")
    acc = torch.zeros(2 + 2 * 6, dtype=torch.float64, device="cuda")
    ref_acc = None
    for b in eval_batches():
        t = {k: torch.from_numpy(v).cuda() for k, v in b.items()}
        dist, valid = ops.keypoint_eval(t["coord"], t["pred"], t["target"], t["offset"], t.get("scale"), acc=acc)
        rd, rv, ref_acc = P.offset_keypoint_eval(b["coord"], b["pred"], b["target"], b["offset"], b.get("scale"), ref_acc)
        assert torch.equal(valid.cpu().bool(), rv)
        assert torch.allclose(dist.cpu(), rd, rtol=1e-5, atol=1e-5), (dist.cpu() - rd).abs().max()   # fp32 sum order
    mean_dist, per_kp, counts = summarize(acc, 6)
    assert abs(mean_dist - float(g["mean_dist"])) <= 1e-5 * float(g["mean_dist"])
    assert np.allclose(per_kp, g["per_kp"], rtol=1e-5)
    assert counts == [int(ref_acc[2 + 6 + k]) for k in range(6)] and acc[1].item() == ref_acc[1]


@pytest.mark.parametrize("split", [False, True])
@pytest.mark.parametrize("K", [6, 5])
def test_keypoint_eval_ragged(split, K):
    """A10 on ragged scenes large enough for the split form (several CTAs per scene, whole-row reads): an empty scene in
    the middle, a one-row scene, a sparse mask, tied maxima across split borders (the first row wins), no accumulator;
    K = 5 makes the block size (510 / 5 = 102 threads per keypoint) and the tree reduction uneven."""
    from ptv3_b200 import _lib, ops
    sizes = [1, 0, 40001, 3000]
    gen = torch.Generator().manual_seed(3)
    n = sum(sizes)
    coord, pred = torch.randn(n, 3, generator=gen), torch.randn(n, K, 4, generator=gen)
    target = torch.randn(n, K, 4, generator=gen) * 0.1
    target[..., 3] = (torch.rand(n, K, generator=gen) < 0.02).float()
    target[1:, 1, 3] = 0.0                                            # keypoint 1 masked nowhere in the big scenes
    pred[[500, 20000, 39000], 2, 3] = 9.0                             # ties in different splits of scene 2
    offset = torch.tensor(sizes).cumsum(0)
    assert (_lib.load().ptv3_keypoint_eval_workspace(n, len(sizes), K) > 0)
    dist, valid = ops.keypoint_eval(coord.cuda(), pred.cuda(), target.cuda(), offset.cuda(), split=split)
    rd, rv, _ = P.offset_keypoint_eval(coord, pred, target, offset)
    assert torch.equal(valid.cpu().bool(), rv) and not valid[1].any() and not valid[2:, 1].any()
    assert torch.allclose(dist.cpu(), rd, rtol=2e-5, atol=2e-5), (dist.cpu() - rd).abs().max()
    # the tied maximum resolves to row 500 of scene 2 in both forms: identical prediction vote -> same distance
    other = ops.keypoint_eval(coord.cuda(), pred.cuda(), target.cuda(), offset.cuda(), split=not split)[0]
    assert torch.allclose(dist, other, rtol=2e-5, atol=2e-5)


def test_offset_keypoint_evaluator_hook(golden):
    """The drop-in hook (ptv3_b200.evaluator.OffsetKeypointEvaluator) driven like the reference's by a trainer object:
    same log lines, writer scalars and comm_info metric as hooks/offset_keypoint_evaluator.py produced for eval.npz."""
    import types
    from make_golden_names import eval_batches
    from ptv3_b200.evaluator import OffsetKeypointEvaluator
    g = golden("eval")
    logs, scalars = [], {}

    class Model(torch.nn.Module):
        def forward(self, data_dict):
            return dict(pred=data_dict["pred"])

    loader = [{k: torch.from_numpy(v) for k, v in b.items()} for b in eval_batches()]        # host tensors, as a DataLoader yields
    trainer = types.SimpleNamespace(
        model=Model(), val_loader=loader, epoch=0, comm_info={}, logger=types.SimpleNamespace(info=logs.append),
        writer=types.SimpleNamespace(add_scalar=lambda name, v, step: scalars.__setitem__(name, (float(v), int(step)))))
    hook = OffsetKeypointEvaluator(num_keypoints=6)
    hook.trainer = trainer
    hook.after_epoch()
    # same log lines as the reference's; the 4-decimal means may differ in the last digit (fp32 summation order of the
    # ground-truth mean), so numbers are compared at 1e-5 relative and the text around them exactly
    import re
    num = re.compile(r"-?\d+\.\d+")
    want = [str(s) for s in g["logs"]]
    assert len(logs) == len(want) == 2 + 6
    for a, b in zip(logs, want):
        assert num.sub("#", a) == num.sub("#", b), (a, b)
        for x, y in zip(num.findall(a), num.findall(b)):
            assert abs(float(x) - float(y)) <= 1e-5 * abs(float(y)) + 1e-4, (a, b)
    assert trainer.comm_info["current_metric_name"] == "mean_dist"
    assert abs(-trainer.comm_info["current_metric_value"] - float(g["mean_dist"])) <= 1e-5 * float(g["mean_dist"])
    assert scalars["val/MeanDist"][1] == 1 and len(scalars) == 7
