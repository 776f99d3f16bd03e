"""Torch-tensor wrappers over the C ABI (include/ptv3_b200.h).

torch is plumbing here: it owns device memory and the current stream; every
computation below is one of our sm_100a kernels.  All functions require CUDA
tensors and raise RuntimeError otherwise (no CPU fallback).
"""
import ctypes

import torch

from . import _lib

ORDER_IDS = {"z": 0, "z-trans": 1, "hilbert": 2, "hilbert-trans": 3}
REDUCE_IDS = {"sum": 0, "add": 0, "mean": 1, "max": 2, "min": 3}
F32, BF16 = 0, 1


def _p(t):
    """Raw device pointer for a c_void_p argument (ctypes converts the int; None is NULL)."""
    return t.data_ptr() if t is not None else None


_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)


def _stream():
    """cudaStream_t of torch's current stream (the raw getter is ~20x cheaper than building a torch.cuda.Stream)."""
    if _raw_stream is not None:
        return _raw_stream(torch.cuda.current_device())
    return torch.cuda.current_stream().cuda_stream


def _need_cuda(*ts):
    for t in ts:
        if t is not None and not t.is_cuda:
            raise RuntimeError("ptv3_b200 ops need CUDA tensors (no CPU fallback)")


def _dt(t):
    if t.dtype == torch.float32:
        return F32
    if t.dtype == torch.bfloat16:
        return BF16
    raise RuntimeError(f"unsupported dtype {t.dtype} (float32 / bfloat16 only)")


# ---- optional per-op CUDA-event timing (bench.py's roofline leg); off by default -------------------------
_PROF = None


def profile_start():
    global _PROF
    _PROF = []


def profile_stop():
    """Returns [(name, ms, meta)] for every op launched since profile_start(); synchronises."""
    global _PROF
    rec, _PROF = _PROF or [], None
    torch.cuda.synchronize()
    return [(name, a.elapsed_time(b), meta) for name, a, b, meta in rec]


class _timed:
    def __init__(self, name, **meta):
        self.name, self.meta = name, meta

    def __enter__(self):
        if _PROF is not None:
            self.a = torch.cuda.Event(enable_timing=True)
            self.a.record()
        return self

    def __exit__(self, *exc):
        if _PROF is not None:
            b = torch.cuda.Event(enable_timing=True)
            b.record()
            _PROF.append((self.name, self.a, b, self.meta))
        return False


def _ws(nbytes, device):
    return torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=device)


# ------------------------------------------------------------------ A2 serialization
def encode_orders(grid_coord, batch, depth, orders):
    """(len(orders), N) int64 codes.  serialization/default.py:9-24 for several orders in one pass."""
    _need_cuda(grid_coord, batch)
    lib = _lib.load()
    if grid_coord.dtype not in (torch.int32, torch.int64):
        grid_coord = grid_coord.long()
    g = grid_coord.contiguous()
    n = g.shape[0]
    b = batch.contiguous().long() if batch is not None else None
    ids = (ctypes.c_int * len(orders))(*[ORDER_IDS[o] for o in orders])
    code = torch.empty((len(orders), n), dtype=torch.int64, device=g.device)
    _lib.check(lib.ptv3_encode(_p(g), int(g.dtype == torch.int64), _p(b), n, int(depth), ids, len(orders), _p(code),
                               _stream()), "ptv3_encode")
    return code


def encode(grid_coord, batch=None, depth=16, order="z"):
    """Drop-in for pointcept.models.utils.serialization.encode (default.py:9-24)."""
    if order not in ORDER_IDS:
        raise AssertionError(order)
    return encode_orders(grid_coord, batch, depth, [order])[0]


def grid_max(grid_coord):
    _need_cuda(grid_coord)
    lib = _lib.load()
    g = grid_coord.contiguous()
    if g.dtype not in (torch.int32, torch.int64):
        g = g.long()
    out = torch.empty(3, dtype=torch.int64, device=g.device)
    _lib.check(lib.ptv3_grid_max(_p(g), int(g.dtype == torch.int64), g.shape[0], _p(out), _stream()), "ptv3_grid_max")
    return out


def argsort_rows(code, key_bits=63, want_inverse=True):
    """order, inverse of (rows, N) int64 codes (structure.py:92-99)."""
    _need_cuda(code)
    lib = _lib.load()
    code = code.contiguous()
    rows, n = code.shape
    order = torch.empty_like(code)
    inverse = torch.empty_like(code) if want_inverse else None
    nbytes = lib.ptv3_argsort_rows_workspace(rows, n)
    ws = _ws(nbytes, code.device)
    _lib.check(lib.ptv3_argsort_rows(_p(code), rows, n, int(key_bits), _p(order), _p(inverse), _p(ws), ws.numel(),
                                     _stream()), "ptv3_argsort_rows")
    return order, inverse


# ------------------------------------------------------------------ A0 GridSample
HASH_IDS = {("fnv", "f64"): 0, ("ravel", "f64"): 1, ("ravel", "f32"): 2, ("fnv", "f32"): 3}


def gridsample(coord, grid_size, hash_type="fnv", rand=None, quotient="f32"):
    """Returns dict of device tensors (see include/ptv3_b200.h); one host sync to read the voxel count.
    quotient: precision of `coord / grid_size` before the floor (transform.py:850): "f32" = the reference's pinned
    numpy 1.26.4 (float32 array / 0-d float64 array stays float32), "f64" = numpy >= 2 (NEP 50 promotes to float64)."""
    if hash_type == "ravel_f32":
        hash_type, quotient = "ravel", "f32"
    _need_cuda(coord, rand)
    lib = _lib.load()
    coord = coord.contiguous().float()
    r = coord.shape[0]
    dev = coord.device
    i64 = dict(dtype=torch.int64, device=dev)
    grid = torch.empty((r, 3), **i64)
    key = torch.empty(r, **i64)
    idx_sort, inverse, count, idx_unique = (torch.empty(r, **i64) for _ in range(4))
    min_grid = torch.empty(3, **i64)
    n_vox = torch.zeros(1, **i64)
    if rand is not None:
        rand = rand.contiguous().long()
        if rand.numel() < r:
            rand = torch.cat([rand, rand.new_zeros(r - rand.numel())])
    ws = _ws(lib.ptv3_gridsample_workspace(r), dev)
    _lib.check(lib.ptv3_gridsample(_p(coord), r, float(grid_size), HASH_IDS[(hash_type, quotient)], _p(rand), _p(grid),
                                   _p(key), _p(idx_sort), _p(inverse), _p(count), _p(idx_unique), _p(min_grid),
                                   _p(n_vox), _p(ws), ws.numel(), _stream()), "ptv3_gridsample")
    m = int(n_vox.item())
    return dict(grid_coord=grid, key=key, idx_sort=idx_sort, inverse=inverse, count=count[:m],
                idx_unique=idx_unique[:m], min_grid=min_grid, n_voxels=m)


def gridsample_batch(coord, offset, grid_size, hash_type="fnv", rand=None, quotient="f32"):
    """GridSample of every scene of a batch in one launch chain and ONE host read (coord (R,3) rows of all scenes,
    offset (B) cumulative row counts).  The per-scene results of gridsample() concatenated: idx_unique / idx_sort hold
    global row indices ordered by (scene, ascending key); voxel_offset (B) = cumulative voxel counts."""
    _need_cuda(coord, offset, rand)
    lib = _lib.load()
    coord = coord.contiguous().float()
    offset = offset.contiguous().long()
    r, b = coord.shape[0], offset.numel()
    dev = coord.device
    i64 = dict(dtype=torch.int64, device=dev)
    grid = torch.empty((r, 3), **i64)
    key = torch.empty(r, **i64)
    idx_sort, inverse, count, idx_unique = (torch.empty(r, **i64) for _ in range(4))
    min_grid = torch.empty((b, 3), **i64)
    tail = torch.zeros(b + 1, **i64)              # voxel_offset (b) followed by n_voxels: one D2H read for both
    if rand is not None:
        rand = rand.contiguous().long()
        if rand.numel() < r:
            rand = torch.cat([rand, rand.new_zeros(r - rand.numel())])
    ws = _ws(lib.ptv3_gridsample_batch_workspace(r, b), dev)
    _lib.check(lib.ptv3_gridsample_batch(_p(coord), r, _p(offset), b, float(grid_size), HASH_IDS[(hash_type, quotient)],
                                         _p(rand), _p(grid), _p(key), _p(idx_sort), _p(inverse), _p(count),
                                         _p(idx_unique), _p(min_grid), _p(tail), tail.data_ptr() + 8 * b, _p(ws),
                                         ws.numel(), _stream()), "ptv3_gridsample_batch")
    host = tail.cpu()
    m = int(host[b])
    return dict(grid_coord=grid, key=key, idx_sort=idx_sort, inverse=inverse, count=count[:m], idx_unique=idx_unique[:m],
                min_grid=min_grid, voxel_offset=tail[:b], voxel_offset_host=host[:b], n_voxels=m)


# ------------------------------------------------------------------ A6 pooling front end
def grid_pool(grid_coord, batch, stride, bits, n_scenes):
    """Integer front end of PT-v3m2's GridPooling (point_transformer_v3m2_sonata.py:393-417) in one launch chain and ONE
    host read (pooled row count + the pooled grid's per-axis maximum).  Returns dict(cluster, indices, idx_ptr,
    grid_coord (M,3) int64, batch (M), indices4 (M,4) int32, n_pooled, grid_max (3 ints))."""
    _need_cuda(grid_coord, batch)
    lib = _lib.load()
    g = grid_coord.contiguous()
    if g.dtype not in (torch.int32, torch.int64):
        g = g.long()
    b = batch.contiguous().long()
    n, dev = g.shape[0], g.device
    scene_bits = max(1, (int(n_scenes) - 1).bit_length())
    cluster = torch.empty(n, dtype=torch.int64, device=dev)
    indices = torch.empty(n, dtype=torch.int64, device=dev)
    idx_ptr = torch.empty(n + 1, dtype=torch.int64, device=dev)
    pg = torch.empty((n, 3), dtype=torch.int64, device=dev)
    pb = torch.empty(n, dtype=torch.int64, device=dev)
    i4 = torch.empty((n, 4), dtype=torch.int32, device=dev)
    meta = torch.empty(4, dtype=torch.int64, device=dev)
    nbytes = lib.ptv3_grid_pool_workspace(n)
    ws = _ws(nbytes, dev)
    _lib.check(lib.ptv3_grid_pool(_p(g), int(g.dtype == torch.int64), _p(b), n, int(stride), int(bits), scene_bits,
                                  _p(cluster), _p(indices), _p(idx_ptr), _p(pg), _p(pb), _p(i4), _p(meta), _p(ws), nbytes,
                                  _stream()), "ptv3_grid_pool")
    m, mx, my, mz = (int(v) for v in meta.tolist())      # the level's one host read
    return dict(cluster=cluster, indices=indices, idx_ptr=idx_ptr[:m + 1], grid_coord=pg[:m], batch=pb[:m],
                indices4=i4[:m], n_pooled=m, grid_max=(mx, my, mz))


def pool_front(code, order, pool_depth):
    """SerializedPooling integer front end (:385-412).  One host sync (pooled count)."""
    _need_cuda(code, order)
    lib = _lib.load()
    code, order = code.contiguous(), order.contiguous()
    rows, n = code.shape
    dev = code.device
    i64 = dict(dtype=torch.int64, device=dev)
    cluster, indices, head = (torch.empty(n, **i64) for _ in range(3))
    idx_ptr = torch.empty(n + 1, **i64)
    new_code, new_order, new_inverse = (torch.empty((rows, n), **i64) for _ in range(3))
    n_pooled = torch.zeros(1, **i64)
    ws = _ws(lib.ptv3_pool_front_workspace(rows, n), dev)
    _lib.check(lib.ptv3_pool_front(_p(code), _p(order), rows, n, 3 * int(pool_depth), _p(cluster), _p(indices),
                                   _p(idx_ptr), _p(head), _p(new_code), _p(new_order), _p(new_inverse), _p(n_pooled),
                                   _p(ws), ws.numel(), _stream()), "ptv3_pool_front")
    m = int(n_pooled.item())
    return dict(cluster=cluster, indices=indices, idx_ptr=idx_ptr[:m + 1], head=head[:m],
                code=new_code[:, :m].contiguous(), order=new_order[:, :m].contiguous(),
                inverse=new_inverse[:, :m].contiguous(), n_pooled=m)


def pool_gather_ints(grid, batch, head, pool_depth, m=None, want_grid=True, want_batch=True, want_indices=True):
    _need_cuda(grid, batch, head)
    lib = _lib.load()
    grid, batch = grid.contiguous().long(), batch.contiguous().long()
    m = int(head.shape[0]) if head is not None else int(grid.shape[0] if m is None else m)
    dev = grid.device
    ng = torch.empty((m, 3), dtype=torch.int64, device=dev) if want_grid else None
    nb = torch.empty(m, dtype=torch.int64, device=dev) if want_batch else None
    ni = torch.empty((m, 4), dtype=torch.int32, device=dev) if want_indices else None
    _lib.check(lib.ptv3_pool_gather_ints(_p(grid), _p(batch), _p(head.contiguous() if head is not None else None), m,
                                         int(pool_depth), _p(ng), _p(nb), _p(ni), _stream()), "ptv3_pool_gather_ints")
    return ng, nb, ni


def offset_to_batch(offset, n):
    """offset2batch (utils/misc.py:13-22) for n rows, no host sync."""
    _need_cuda(offset)
    lib = _lib.load()
    offset = offset.contiguous().long()
    batch = torch.empty(int(n), dtype=torch.int64, device=offset.device)
    _lib.check(lib.ptv3_offset_to_batch(_p(offset), offset.numel(), int(n), _p(batch), _stream()),
               "ptv3_offset_to_batch")
    return batch


def batch_to_offset(batch, n_scenes):
    _need_cuda(batch)
    lib = _lib.load()
    batch = batch.contiguous().long()
    off = torch.empty(int(n_scenes), dtype=torch.int64, device=batch.device)
    _lib.check(lib.ptv3_batch_to_offset(_p(batch), batch.shape[0], int(n_scenes), _p(off), _stream()),
               "ptv3_batch_to_offset")
    return off


# ------------------------------------------------------------------ segment reductions
def segment_csr(src, indptr, reduce="sum", gather=None, long_segments=False):
    """torch_scatter.segment_csr along dim 0 (optionally reading src rows through `gather`)."""
    _need_cuda(src, indptr, gather)
    lib = _lib.load()
    shape = src.shape
    src2 = src.contiguous().view(shape[0], -1) if src.dim() != 2 else src.contiguous()
    c = src2.shape[1]
    indptr = indptr.contiguous().long()
    m = indptr.numel() - 1
    out = torch.empty((m, c), dtype=src.dtype, device=src.device)
    g = gather.contiguous().long() if gather is not None else None
    _lib.check(lib.ptv3_segment_csr(_p(src2), _dt(src2), _p(g), _p(indptr), -m if long_segments else m, c,
                                    REDUCE_IDS[reduce], _p(out), _stream()), "ptv3_segment_csr")
    return out.view((m,) + tuple(shape[1:]))


def unpool_add(skip, up, inverse, gelu=True, want_skip_act=True, second="skip"):
    """out = act(skip) + act(up)[inverse] (fp32); the second output, in the input dtype, is act(skip) (PT-v3m1's
    SerializedUnpooling :471-482, where the sparse tensor keeps the skip branch) or, with second="sum", the sum again
    (PT-v3m2's GridUnpooling :498-499: the compute-dtype copy for the next conv)."""
    _need_cuda(skip, up, inverse)
    lib = _lib.load()
    skip, up = skip.contiguous(), up.contiguous()
    assert skip.dtype == up.dtype and skip.shape[1] == up.shape[1]
    n, c = skip.shape
    out = torch.empty((n, c), dtype=torch.float32, device=skip.device)
    act = torch.empty_like(skip) if want_skip_act else None
    _lib.check(lib.ptv3_unpool_add(_p(skip), _p(up), _dt(skip), _p(inverse.contiguous().long()), n, c,
                                   int(bool(gelu)) | (2 if second == "sum" else 0), _p(out), _p(act), _stream()),
               "ptv3_unpool_add")
    return out, act


def ln_residual(x, gamma=None, beta=None, eps=1e-5, residual=None, want_res=True, gamma2=None, beta2=None,
                out2_dtype=None, row_scale=None):
    """t = LN(x) if gamma is given else x; t *= row_scale[:, None] if given (DropPath: fp32 (n), 0 or 1 / keep);
    r = residual + t (or t).  Returns (r fp32 or None, out2 or None) with out2 = LN2(r) if gamma2 is given, else r cast
    to out2_dtype (when out2_dtype is given)."""
    _need_cuda(x)
    lib = _lib.load()
    x = x.contiguous()
    n, c = x.shape
    res_out = torch.empty((n, c), dtype=torch.float32, device=x.device) if want_res else None
    if gamma2 is not None and out2_dtype is None:
        out2_dtype = torch.float32
    out2 = torch.empty((n, c), dtype=out2_dtype, device=x.device) if out2_dtype is not None else None
    if residual is not None:
        residual = residual.contiguous()
        assert residual.dtype == torch.float32
    if row_scale is not None:
        row_scale = row_scale.contiguous().view(-1)
        assert row_scale.dtype == torch.float32 and row_scale.numel() == n
    _lib.check(lib.ptv3_ln_residual_rs(_p(x), _dt(x), _p(gamma), _p(beta), float(eps), _p(row_scale), _p(residual),
                                       _p(res_out), _p(gamma2), _p(beta2), _p(out2),
                                       _dt(out2) if out2 is not None else 0, n, c, _stream()), "ptv3_ln_residual_rs")
    return res_out, out2


# ------------------------------------------------------------------ A8 fused MLP (tcgen05)
def pack_kmajor(w, tile_rows):
    """(rows, k) bf16 row-major weight -> K-blocked, 128-byte-swizzled shared-memory image (tiles of tile_rows rows)."""
    _need_cuda(w)
    lib = _lib.load()
    w = w.contiguous()
    assert w.dtype == torch.bfloat16 and w.dim() == 2
    rows, k = w.shape
    packed = torch.empty(lib.ptv3_pack_kmajor_bytes(rows, k) // 2, dtype=torch.bfloat16, device=w.device)
    _lib.check(lib.ptv3_pack_kmajor(_p(w), rows, k, int(tile_rows), _p(packed), _stream()), "ptv3_pack_kmajor")
    return packed


def mlp_fc1_tile_rows(c):
    """tile_rows for pack_kmajor of fc1's weight (the hidden-block width of the fused-MLP kernel for this width)."""
    return int(_lib.load().ptv3_mlp_fc1_tile_rows(int(c)))


def mlp_fused_supported(c, hidden):
    return bool(_lib.load().ptv3_mlp_fused_supported(int(c), int(hidden)))


def mlp_fused(x, w1_packed, b1, w2_packed, b2, hidden, residual=None, want_res=True, want_bf16=True, row_scale=None):
    """r = (residual +) row_scale[:, None] * fc2(GELU(fc1(x))) in one kernel (row_scale: DropPath, fp32 (N) or None).
    x (N,C) bf16.  Returns (r fp32 or None, r bf16 or None)."""
    _need_cuda(x, w1_packed, w2_packed)
    lib = _lib.load()
    x = x.contiguous()
    assert x.dtype == torch.bfloat16
    n, c = x.shape
    res_out = torch.empty((n, c), dtype=torch.float32, device=x.device) if want_res else None
    out16 = torch.empty((n, c), dtype=torch.bfloat16, device=x.device) if want_bf16 else None
    if residual is not None:
        residual = residual.contiguous()
        assert residual.dtype == torch.float32 and residual.shape == x.shape
    with _timed("mlp_fused", n=n, c=c, hidden=int(hidden)):
        if row_scale is not None:
            row_scale = row_scale.contiguous().view(-1)
            assert row_scale.dtype == torch.float32 and row_scale.numel() == n
        _lib.check(lib.ptv3_mlp_fused_rs(_p(x), _p(w1_packed), _p(b1), _p(w2_packed), _p(b2), _p(residual),
                                         _p(row_scale), n, c, int(hidden), _p(res_out), _p(out16), _stream()),
                   "ptv3_mlp_fused_rs")
    return res_out, out16


def proj_mlp_fused(a, wproj_packed, bproj, x1, ln_gamma, ln_beta, eps, w1_packed, b1, w2_packed, b2, hidden,
                   rs_attn=None, rs_mlp=None, want_bf16=True):
    """Second half of Block.forward in one kernel (C = 32 / 64): x2 = x1 + rs_attn * proj(a); out = x2 + rs_mlp *
    MLP(LayerNorm(x2)).  a (N,C) bf16, x1 (N,C) fp32.  Returns (out fp32, out bf16 or None)."""
    _need_cuda(a, x1, wproj_packed, w1_packed, w2_packed)
    lib = _lib.load()
    a, x1 = a.contiguous(), x1.contiguous()
    assert a.dtype == torch.bfloat16 and x1.dtype == torch.float32 and a.shape == x1.shape
    n, c = a.shape
    out = torch.empty((n, c), dtype=torch.float32, device=a.device)
    out16 = torch.empty((n, c), dtype=torch.bfloat16, device=a.device) if want_bf16 else None
    for r in (rs_attn, rs_mlp):
        assert r is None or (r.dtype == torch.float32 and r.numel() == n and r.is_contiguous())
    with _timed("proj_mlp_fused", n=n, c=c, hidden=int(hidden)):
        _lib.check(lib.ptv3_proj_mlp_fused(_p(a), _p(wproj_packed), _p(bproj), _p(x1), _p(rs_attn), _p(ln_gamma),
                                           _p(ln_beta), float(eps), _p(w1_packed), _p(b1), _p(w2_packed), _p(b2),
                                           _p(rs_mlp), n, c, int(hidden), _p(out), _p(out16), _stream()),
                   "ptv3_proj_mlp_fused")
    return out, out16


# ------------------------------------------------------------------ A8 one native call per Block
NATIVE_BLOCK = True      # bench.py's per-op timing leg and A/B tests switch this off (op-by-op path, same kernels)


def block_params(**kw):
    """ptv3_block_params from keyword arguments: ints / floats by value, tensors by device pointer.  Returns (struct,
    keepalive list) -- the caller caches both next to the prepared weights."""
    bp = _lib.BlockParams()
    keep = []
    for k, v in kw.items():
        if torch.is_tensor(v):
            _need_cuda(v)
            assert v.is_contiguous()
            keep.append(v)
            setattr(bp, k, v.data_ptr())
        else:
            setattr(bp, k, v)
    return bp, keep


_BLOCK_WS = {}


def block_fwd(bp, x, conv_in, nbr, order, windows, n_windows):
    """Block.forward (:318-338) in ONE library call.  x (n,c) fp32, conv_in (n,c) bf16 -> (x_out fp32, x_out bf16).
    The workspace is a per-device buffer that only grows (stream-ordered reuse: every Block of a forward runs on the
    same stream)."""
    _need_cuda(x, conv_in, nbr, order, windows, n_windows)
    lib = _lib.load()
    assert x.dtype == torch.float32 and conv_in.dtype == torch.bfloat16 and x.is_contiguous() and conv_in.is_contiguous()
    n, c = x.shape
    x_out = torch.empty((n, c), dtype=torch.float32, device=x.device)
    x_cd = torch.empty((n, c), dtype=torch.bfloat16, device=x.device)
    if n == 0:
        return x_out, x_cd
    need = lib.ptv3_block_workspace(ctypes.byref(bp), n, windows.shape[0])
    key = (x.device.index, _stream())
    ws = _BLOCK_WS.get(key)
    if ws is None or ws.numel() < need:
        ws = _BLOCK_WS[key] = torch.empty(int(need * 1.25), dtype=torch.uint8, device=x.device)
    with _timed("block_fwd", n=n, c=c):
        _lib.check(lib.ptv3_block_fwd(ctypes.byref(bp), _p(x), _p(conv_in), _p(nbr), _p(order), _p(windows), _p(n_windows),
                                      windows.shape[0], n, _p(x_out), _p(x_cd), _p(ws), ws.numel(), _stream()),
                   "ptv3_block_fwd")
    return x_out, x_cd


# ------------------------------------------------------------------ A5 submanifold conv
def subm_build_map(indices, ksize, z=None):
    """indices (N,4) int32 [b,x,y,z] -> nbr (N, ksize^3) int32 (-1 = absent).  z = (code_z, order_z, depth) from the
    level's serialization selects the block-occupancy search (same table, 8 probes per row instead of ksize^3)."""
    _need_cuda(indices)
    lib = _lib.load()
    indices = indices.contiguous().int()
    n = indices.shape[0]
    nbr = torch.empty((n, ksize ** 3), dtype=torch.int32, device=indices.device)
    if z is not None and ksize in (3, 5) and int(z[2]) >= 2 and z[0].numel() == n:
        nbytes = lib.ptv3_subm_stem_z_workspace(n)
        ws = _ws(nbytes, indices.device)
        _lib.check(lib.ptv3_subm_build_map_z(_p(indices), n, int(ksize), _p(z[0].contiguous()), _p(z[1].contiguous()),
                                             int(z[2]), _p(nbr), _p(ws), nbytes, _stream()), "ptv3_subm_build_map_z")
        return nbr
    ws = _ws(lib.ptv3_subm_map_workspace(n), indices.device)
    _lib.check(lib.ptv3_subm_build_map(_p(indices), n, int(ksize), _p(nbr), _p(ws), ws.numel(), _stream()),
               "ptv3_subm_build_map")
    return nbr


def subm_stem_fused(indices, feat, weight_koc, bias, ksize, gelu=True, want_f32=True, want_bf16=True, z=None):
    """Stem conv (C_in 4 -> 32, fp32) with the neighbour search inside + bias + GELU: returns (fp32, bf16) outputs.
    z = (code_z (n) int64, order_z (n) int64, depth) from Point.serialization selects the block-occupancy search
    (ptv3_subm_stem_fused_z: 8 probes per row instead of 125; bit-identical result)."""
    _need_cuda(indices, feat, weight_koc)
    lib = _lib.load()
    indices, feat, w = indices.contiguous().int(), feat.contiguous().float(), weight_koc.contiguous().float()
    n = feat.shape[0]
    assert feat.shape[1] == 4 and w.shape == (ksize ** 3, 32, 4)
    o32 = torch.empty((n, 32), dtype=torch.float32, device=feat.device) if want_f32 else None
    o16 = torch.empty((n, 32), dtype=torch.bfloat16, device=feat.device) if want_bf16 else None
    b = bias.contiguous().float() if bias is not None else None
    if z is not None and int(z[2]) >= 2 and ksize <= 5:
        code_z, order_z, depth = z
        assert code_z.dtype == torch.int64 and order_z.dtype == torch.int64 and code_z.numel() == n == order_z.numel()
        nbytes = lib.ptv3_subm_stem_z_workspace(n)
        ws = _ws(nbytes, feat.device)
        _lib.check(lib.ptv3_subm_stem_fused_z(_p(indices), n, int(ksize), _p(code_z.contiguous()),
                                              _p(order_z.contiguous()), int(depth), _p(feat), _p(w), _p(b),
                                              1 if gelu else 0, _p(o32), _p(o16), _p(ws), nbytes, _stream()),
                   "ptv3_subm_stem_fused_z")
        return o32, o16
    ws = _ws(lib.ptv3_subm_map_workspace(n), feat.device)
    _lib.check(lib.ptv3_subm_stem_fused(_p(indices), n, int(ksize), _p(feat), _p(w), _p(b), 1 if gelu else 0, _p(o32),
                                        _p(o16), _p(ws), ws.numel(), _stream()), "ptv3_subm_stem_fused")
    return o32, o16


def subm_conv(feat, weight_koc, bias, nbr, algo=0):
    """feat (N,Cin), weight (k^3, Cout, Cin) same dtype, bias fp32 (Cout) or None -> (N,Cout)."""
    _need_cuda(feat, weight_koc, nbr)
    lib = _lib.load()
    feat, weight_koc = feat.contiguous(), weight_koc.contiguous()
    assert feat.dtype == weight_koc.dtype
    cin = feat.shape[1]
    k3, cout, cin2 = weight_koc.shape
    n = nbr.shape[0]          # output rows; SubM convs have n == feat.shape[0], strided / inverse convs do not
    assert cin2 == cin and nbr.shape == (n, k3) and nbr.dtype == torch.int32
    out = torch.empty((n, cout), dtype=feat.dtype, device=feat.device)
    dt = _dt(feat)
    ws = _ws(lib.ptv3_subm_conv_workspace(n, cin, cout, k3, dt, algo), feat.device)
    b = bias.contiguous().float() if bias is not None else None
    with _timed("subm_conv", n=n, cin=cin, cout=cout, k3=k3, dtype=dt, nbr=nbr):
        _lib.check(lib.ptv3_subm_conv(_p(feat), _p(weight_koc), _p(b), _p(nbr.contiguous()), n, cin, cout, k3, dt,
                                      int(algo), _p(out), _p(ws), ws.numel(), _stream()), "ptv3_subm_conv")
    return out


def subm_conv_tc_supported(cin, cout, k3):
    return bool(_lib.load().ptv3_subm_conv_tc_supported(int(cin), int(cout), int(k3)))


def subm_pack_weight(weight_koc):
    """(k^3, Cout, Cin) bf16 -> the K-blocked, swizzled shared-memory image the tcgen05 conv kernel bulk-copies."""
    _need_cuda(weight_koc)
    lib = _lib.load()
    w = weight_koc.contiguous()
    assert w.dtype == torch.bfloat16
    k3, cout, cin = w.shape
    nbytes = lib.ptv3_subm_pack_weight_bytes(cin, cout, k3)
    if nbytes == 0:
        raise RuntimeError(f"subm_pack_weight: unsupported shape cin={cin} cout={cout} k3={k3}")
    packed = torch.empty(nbytes // 2, dtype=torch.bfloat16, device=w.device)
    _lib.check(lib.ptv3_subm_pack_weight(_p(w), cin, cout, k3, _p(packed), _stream()), "ptv3_subm_pack_weight")
    return packed


def subm_conv_packed(feat, packed, bias, nbr, cout):
    """bf16 tcgen05 path with a pre-packed weight (subm_pack_weight)."""
    _need_cuda(feat, packed, nbr)
    lib = _lib.load()
    feat = feat.contiguous()
    assert feat.dtype == torch.bfloat16 and nbr.dtype == torch.int32
    cin = feat.shape[1]
    n, k3 = nbr.shape         # output rows (== feat rows for SubM convs)
    out = torch.empty((n, cout), dtype=feat.dtype, device=feat.device)
    b = bias.contiguous().float() if bias is not None else None
    with _timed("subm_conv", n=n, cin=cin, cout=cout, k3=k3, dtype=BF16, nbr=nbr):
        _lib.check(lib.ptv3_subm_conv_packed(_p(feat), _p(packed), _p(b), _p(nbr.contiguous()), n, cin, int(cout), k3,
                                             _p(out), _stream()), "ptv3_subm_conv_packed")
    return out


# ------------------------------------------------------------------ A3/A4 attention
def build_windows(offset, patch, n_total):
    """Window records (W_max,4) int64 + device count.  W_max = N // patch + n_scenes (upper bound, no sync)."""
    _need_cuda(offset)
    lib = _lib.load()
    offset = offset.contiguous().long()
    b = offset.numel()
    wmax = int(n_total) // int(patch) + b
    windows = torch.empty((wmax, 4), dtype=torch.int64, device=offset.device)
    nwin = torch.empty(1, dtype=torch.int64, device=offset.device)
    _lib.check(lib.ptv3_build_windows(_p(offset), b, int(patch), _p(windows), wmax, _p(nwin), _stream()),
               "ptv3_build_windows")
    return windows, nwin


def attn_window(qkv, order, windows, n_windows, num_heads, max_len, scale=None, algo=0):
    """qkv (N, 3*H*16) rows in storage order -> (N, H*16) in storage order."""
    _need_cuda(qkv, order, windows, n_windows)
    lib = _lib.load()
    qkv = qkv.contiguous()
    n = qkv.shape[0]
    c = qkv.numel() // (3 * n) if n > 0 else 0
    h = int(num_heads)
    d = c // h if h else 0
    out = torch.empty((n, c), dtype=qkv.dtype, device=qkv.device)
    if n == 0:
        return out
    scale = float(scale) if scale is not None else d ** -0.5
    o = order.contiguous().long() if order is not None else None
    ws, ws_bytes = None, 0
    if int(algo) == 5:
        ws_bytes = lib.ptv3_attn_window_workspace(windows.shape[0], h)
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=qkv.device)
    with _timed("attn_window", n=n, h=h, d=d, windows=windows, n_windows=n_windows, dtype=_dt(qkv)):
        _lib.check(lib.ptv3_attn_window_fwd(_p(qkv), _dt(qkv), _p(o), _p(windows), _p(n_windows), windows.shape[0],
                                            int(max_len), n, h, d, scale, int(algo), _p(out), _p(ws), ws_bytes,
                                            _stream()),
                   "ptv3_attn_window_fwd")
    return out


def device_error():
    """Reads and clears the library's sticky device-side error word (synchronises), a bit mask: 1 = an attention window
    record violated the build_windows contract (k_len > max_len or queries outside the key range); 2 = a sparse-conv
    site outside the packed key range (batch id >= 1024 or a coordinate outside [0, 262128))."""
    return int(_lib.load().ptv3_device_error())


def debug_set(key, value):
    _lib.check(_lib.load().ptv3_debug_set(int(key), int(value)), "ptv3_debug_set")


# ------------------------------------------------------------------ A9 keypoint aggregation (offset head)
def keypoint_aggregate(coord, pred, offset, mode="argmax", mask_thresh=0.5):
    """coord (N,3) fp32, pred (N,K,4) fp32 [dx,dy,dz,conf], offset int64 (B) -> (B,K,3) fp32 keypoints in model
    units (tools/infer_offset.py:366-399 per scene; no host sync)."""
    _need_cuda(coord, pred, offset)
    lib = _lib.load()
    coord, pred = coord.contiguous().float(), pred.contiguous().float()
    offset = offset.contiguous().long()
    n, k = pred.shape[0], pred.shape[1]
    assert coord.shape == (n, 3) and pred.shape == (n, k, 4)
    out = torch.empty((offset.numel(), k, 3), dtype=torch.float32, device=pred.device)
    _lib.check(lib.ptv3_keypoint_aggregate(_p(coord), _p(pred), _p(offset), offset.numel(), int(k),
                                           {"argmax": 0, "weighted": 1}[mode], float(mask_thresh), _p(out), _stream()),
               "ptv3_keypoint_aggregate")
    return out


# ------------------------------------------------------------------ A10 offset-keypoint validation metric
def keypoint_eval(coord, pred, target, offset, scale=None, acc=None, split=True):
    """hooks/offset_keypoint_evaluator.py:46-91 for one batch: coord (N,3), pred / target (N,K,4) fp32, offset int64 (B),
    scale fp32 (B) or None -> dist (B,K) fp32, valid (B,K) int32.  `acc`: float64 (2+2K) device tensor of the running sums
    [total_dist, total_samples, dist_per_kp, samples_per_kp], updated in place (no host sync).  split=False keeps the
    one-CTA-per-(scene, keypoint) kernel (no workspace)."""
    _need_cuda(coord, pred, target, offset)
    lib = _lib.load()
    coord, pred, target = coord.contiguous().float(), pred.contiguous().float(), target.contiguous().float()
    offset = offset.contiguous().long()
    n, k = pred.shape[0], pred.shape[1]
    assert coord.shape == (n, 3) and pred.shape == (n, k, 4) and target.shape == (n, k, 4)
    b = offset.numel()
    if scale is not None:
        scale = scale.contiguous().float()
        assert scale.numel() == b and scale.is_cuda
    if acc is not None:
        assert acc.is_cuda and acc.dtype == torch.float64 and acc.is_contiguous() and acc.numel() == 2 + 2 * k
    dist = torch.empty((b, k), dtype=torch.float32, device=pred.device)
    valid = torch.empty((b, k), dtype=torch.int32, device=pred.device)
    ws_bytes = lib.ptv3_keypoint_eval_workspace(n, b, int(k)) if split else 0
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=pred.device) if ws_bytes else None
    _lib.check(lib.ptv3_keypoint_eval(_p(coord), _p(pred), _p(target), _p(offset), _p(scale), n, b, int(k), _p(dist),
                                      _p(valid), _p(acc), _p(ws), ws_bytes, _stream()),
               "ptv3_keypoint_eval")
    return dist, valid


# ------------------------------------------------------------------ A6 glue: per-channel affine + activation
def affine_act(x, scale=None, shift=None, act="none", want_f32=True, want_bf16=False):
    """act(x * scale + shift) per channel in one pass -> (fp32 or None, bf16 or None).  act: none | gelu | relu."""
    _need_cuda(x)
    lib = _lib.load()
    x = x.contiguous()
    n, c = x.shape
    o32 = torch.empty((n, c), dtype=torch.float32, device=x.device) if want_f32 else None
    o16 = torch.empty((n, c), dtype=torch.bfloat16, device=x.device) if want_bf16 else None
    sc = scale.contiguous().float() if scale is not None else None
    sh = shift.contiguous().float() if shift is not None else None
    _lib.check(lib.ptv3_affine_act(_p(x), _dt(x), _p(sc), _p(sh), {"none": 0, "gelu": 1, "relu": 2}[act], n, c,
                                   _p(o32), _p(o16), _stream()), "ptv3_affine_act")
    return o32, o16


# ------------------------------------------------------------------ backward kernels
def subm_wgrad(feat, dout, nbr, k3):
    """dW (k3, Cout, Cin) fp32 = sum_i dout[i] (x) feat[nbr[i,k]]."""
    _need_cuda(feat, dout, nbr)
    lib = _lib.load()
    feat, dout = feat.contiguous(), dout.contiguous()
    assert feat.dtype == dout.dtype and nbr.dtype == torch.int32 and nbr.shape == (feat.shape[0], k3)
    n, cin = feat.shape
    cout = dout.shape[1]
    dw = torch.zeros((k3, cout, cin), dtype=torch.float32, device=feat.device)
    _lib.check(lib.ptv3_subm_wgrad(_p(feat), _p(dout), _p(nbr.contiguous()), n, cin, cout, int(k3), _dt(feat), _p(dw),
                                   _stream()), "ptv3_subm_wgrad")
    return dw


def attn_window_lse(qkv, order, windows, n_windows, num_heads, max_len, scale=None):
    """bf16 attn_window that also returns the rows' log-sum-exp (N, H) fp32 (log2 units) for attn_window_bwd."""
    _need_cuda(qkv, order, windows, n_windows)
    lib = _lib.load()
    qkv = qkv.contiguous()
    assert qkv.dtype == torch.bfloat16
    n = qkv.shape[0]
    c = qkv.numel() // (3 * n) if n > 0 else 0
    h = int(num_heads)
    d = c // h if h else 0
    out = torch.empty((n, c), dtype=qkv.dtype, device=qkv.device)
    lse = torch.empty((n, h), dtype=torch.float32, device=qkv.device)
    if n == 0:
        return out, lse
    scale = float(scale) if scale is not None else d ** -0.5
    o = order.contiguous().long() if order is not None else None
    _lib.check(lib.ptv3_attn_window_fwd_lse(_p(qkv), _dt(qkv), _p(o), _p(windows), _p(n_windows), windows.shape[0],
                                            int(max_len), n, h, d, scale, _p(out), _p(lse), _stream()),
               "ptv3_attn_window_fwd_lse")
    return out, lse


def attn_window_bwd(qkv, out, dout, order, windows, n_windows, num_heads, max_len, scale=None, lse=None):
    """Gradient of attn_window w.r.t. qkv: (N, 3*H*16) fp32."""
    _need_cuda(qkv, out, dout, order, windows, n_windows)
    lib = _lib.load()
    qkv, out = qkv.contiguous(), out.contiguous()
    dout = dout.contiguous().to(qkv.dtype)
    n = qkv.shape[0]
    c = qkv.numel() // (3 * n) if n > 0 else 0
    h = int(num_heads)
    d = c // h if h else 0
    dqkv = torch.zeros((n, 3 * c), dtype=torch.float32, device=qkv.device)
    if n == 0:
        return dqkv
    scale = float(scale) if scale is not None else d ** -0.5
    o = order.contiguous().long() if order is not None else None
    ws = _ws(lib.ptv3_attn_window_bwd_workspace(n, h), qkv.device)
    _lib.check(lib.ptv3_attn_window_bwd(_p(qkv), _p(out), _p(dout), _dt(qkv), _p(o), _p(windows), _p(n_windows),
                                        windows.shape[0], int(max_len), n, h, d, scale, _p(lse), _p(dqkv), _p(ws),
                                        ws.numel(), _stream()), "ptv3_attn_window_bwd")
    return dqkv


def segment_csr_bwd(dout, src, out, indptr, reduce="sum", gather=None):
    """Gradient of segment_csr w.r.t. src: (n_src, c) fp32 (rows outside every segment get 0)."""
    _need_cuda(dout, src, indptr, gather)
    lib = _lib.load()
    src2 = src.contiguous().view(src.shape[0], -1)
    c = src2.shape[1]
    indptr = indptr.contiguous().long()
    m = indptr.numel() - 1
    dout = dout.contiguous().float().view(m, c)
    out2 = out.contiguous().view(m, c) if out is not None else None
    g = gather.contiguous().long() if gather is not None else None
    dsrc = torch.zeros(src2.shape, dtype=torch.float32, device=src.device)
    _lib.check(lib.ptv3_segment_csr_bwd(_p(dout), _p(src2), _p(out2), _dt(src2), _p(g), _p(indptr), m, c,
                                        REDUCE_IDS[reduce], _p(dsrc), _stream()), "ptv3_segment_csr_bwd")
    return dsrc.view(src.shape)


LN_BWD_CHANNELS = (32, 64, 128, 256, 512)


def layer_norm_bwd(x, dy, gamma, eps):
    """(dx fp32 (n,c), dgamma fp32 (c), dbeta fp32 (c)) of y = LayerNorm(x; gamma, beta, eps)."""
    _need_cuda(x, dy, gamma)
    lib = _lib.load()
    x, dy = x.contiguous(), dy.contiguous()
    n, c = x.shape
    dx = torch.empty((n, c), dtype=torch.float32, device=x.device)
    dgb = torch.zeros((2, c), dtype=torch.float32, device=x.device)
    _lib.check(lib.ptv3_layer_norm_bwd(_p(x), _dt(x), _p(dy), _dt(dy), _p(gamma.contiguous().float()), float(eps), n, c,
                                       _p(dx), _p(dgb[0]), _p(dgb[1]), _stream()), "ptv3_layer_norm_bwd")
    return dx, dgb[0], dgb[1]


def column_sum(x):
    """sum over dim 0 of an (n, c) fp32 / bf16 matrix -> fp32 (c); torch for shapes the kernel does not take."""
    _need_cuda(x)
    n, c = x.shape
    e = 8 if x.dtype == torch.bfloat16 else 4
    if x.dtype not in (torch.bfloat16, torch.float32) or c % e != 0 or c // e > 256:
        return x.float().sum(0)
    x = x.contiguous()
    out = torch.empty(c, dtype=torch.float32, device=x.device)
    _lib.check(_lib.load().ptv3_column_sum(_p(x), _dt(x), n, c, _p(out), _stream()), "ptv3_column_sum")
    return out


def sparse_pair_maps(grid_coord, cluster, n_coarse):
    """(nbr_fwd (M,8), nbr_inv (N,8)) int32 gather tables of SparseConv3d(k2,s2) / its SparseInverseConv3d."""
    _need_cuda(grid_coord, cluster)
    g, c = grid_coord.contiguous().long(), cluster.contiguous().long()
    n = g.shape[0]
    fwd = torch.empty((int(n_coarse), 8), dtype=torch.int32, device=g.device)
    inv = torch.empty((n, 8), dtype=torch.int32, device=g.device)
    _lib.check(_lib.load().ptv3_sparse_pair_maps(_p(g), _p(c), n, int(n_coarse), _p(fwd), _p(inv), _stream()),
               "ptv3_sparse_pair_maps")
    return fwd, inv
