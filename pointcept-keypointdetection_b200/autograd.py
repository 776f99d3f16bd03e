"""torch.autograd.Function wrappers of the custom kernels that carry gradients in the training step
(the reference gets these from spconv / flash_attn / torch_scatter autograd functions behind
engines/train.py:184-245).  Forward = the inference kernels; backward = csrc/backward.cu plus re-use of the forward
kernels where the adjoint is the same operator:

  SubMConv3d   dX = conv(dY; W'[k] = W[k3-1-k]^T) over the same (symmetric) neighbour map, dW = ptv3_subm_wgrad
  attention    ptv3_attn_window_bwd (log-sum-exp recomputed per row, no stored probabilities)
  segment_csr  ptv3_segment_csr_bwd (max: first row equal to the maximum, like torch_scatter's arg output)
  x[inverse]   segment sum over the pooling CSR (deterministic; no atomics)

No CPU fallback: every op raises on non-CUDA tensors (ops._need_cuda).
"""
import torch

from . import ops


def _conv(x, w_koc, bias, nbr):
    """x (N,Cin), w (k3,Cout,Cin) in x.dtype -> (N,Cout); tcgen05 kernel when the shape allows it."""
    k3, cout, cin = w_koc.shape
    if x.dtype == torch.bfloat16 and ops.subm_conv_tc_supported(cin, cout, k3):
        return ops.subm_conv_packed(x, ops.subm_pack_weight(w_koc), bias, nbr, cout)
    return ops.subm_conv(x, w_koc, bias, nbr)


class SubMConvFn(torch.autograd.Function):
    """feat (N,Cin) any float dtype, weight (Cout,k,k,k,Cin) fp32 parameter, bias (Cout) or None, nbr (N,k^3) int32.
    Operands are cast to `cd` (what torch.autocast does for the reference's spconv call); output is `cd`."""

    @staticmethod
    def forward(ctx, feat, weight, bias, nbr, cd):
        cout, cin = weight.shape[0], weight.shape[-1]
        k3 = nbr.shape[1]
        x = feat.detach().to(cd).contiguous()
        w = weight.detach().reshape(cout, k3, cin).permute(1, 0, 2).contiguous().to(cd)
        y = _conv(x, w, bias.detach().float() if bias is not None else None, nbr)
        ctx.save_for_backward(x, w, nbr)
        ctx.meta = (feat.dtype, weight.shape, bias is not None)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w, nbr = ctx.saved_tensors
        feat_dtype, wshape, has_bias = ctx.meta
        k3, cout, cin = w.shape
        dy = dy.contiguous().to(x.dtype)
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            wd = w.flip(0).transpose(1, 2).contiguous()          # (k3, Cin, Cout): W'[k] = W[k3-1-k]^T
            dx = _conv(dy, wd, None, nbr).to(feat_dtype)
        if ctx.needs_input_grad[1]:
            dw = ops.subm_wgrad(x, dy, nbr, k3).permute(1, 0, 2).reshape(wshape)
        if has_bias and ctx.needs_input_grad[2]:
            db = ops.column_sum(dy)
        return dx, dw, db, None, None


class WindowAttnFn(torch.autograd.Function):
    """qkv (N, 3*H*16) rows in storage order -> (N, H*16); see ops.attn_window."""

    @staticmethod
    def forward(ctx, qkv, order, windows, n_windows, num_heads, max_len, scale, algo):
        q = qkv.detach().contiguous()
        lse = None
        if q.dtype == torch.bfloat16 and algo in (0, 2) and max_len <= 1024:
            out, lse = ops.attn_window_lse(q, order, windows, n_windows, num_heads, max_len, scale)
        else:
            out = ops.attn_window(q, order, windows, n_windows, num_heads, max_len, scale, algo)
        ctx.save_for_backward(q, out, order, windows, n_windows, lse)
        ctx.meta = (num_heads, max_len, scale)
        return out

    @staticmethod
    def backward(ctx, dout):
        q, out, order, windows, n_windows, lse = ctx.saved_tensors
        num_heads, max_len, scale = ctx.meta
        dqkv = ops.attn_window_bwd(q, out, dout, order, windows, n_windows, num_heads, max_len, scale, lse=lse)
        return dqkv.to(q.dtype), None, None, None, None, None, None, None


class SegmentCSRFn(torch.autograd.Function):
    """torch_scatter.segment_csr(src[gather], indptr, reduce) along dim 0."""

    @staticmethod
    def forward(ctx, src, indptr, reduce, gather):
        s = src.detach().contiguous()
        out = ops.segment_csr(s, indptr, reduce, gather=gather)
        ctx.save_for_backward(s, out, indptr, gather)
        ctx.reduce = reduce
        return out

    @staticmethod
    def backward(ctx, dout):
        s, out, indptr, gather = ctx.saved_tensors
        return ops.segment_csr_bwd(dout, s, out, indptr, ctx.reduce, gather).to(s.dtype), None, None, None


class GatherRowsFn(torch.autograd.Function):
    """y = x[inverse] where (indices, idx_ptr) is the CSR of `inverse` (pooling cluster -> member rows):
    dx[c] = sum of dy over the rows of cluster c, read through the CSR."""

    @staticmethod
    def forward(ctx, x, inverse, indices, idx_ptr):
        ctx.save_for_backward(indices, idx_ptr)
        return x.detach().index_select(0, inverse)

    @staticmethod
    def backward(ctx, dy):
        indices, idx_ptr = ctx.saved_tensors
        return ops.segment_csr(dy.contiguous(), idx_ptr, "sum", gather=indices), None, None, None


class LayerNormFn(torch.autograd.Function):
    """y = LayerNorm(x) over the last dim of an (N, C) tensor, written directly in `out_dtype` (the dtype of the GEMM
    that consumes it): forward = ops.ln_residual, backward = ptv3_layer_norm_bwd (mean / rstd recomputed)."""

    @staticmethod
    def forward(ctx, x, weight, bias, eps, out_dtype):
        xd = x.detach().contiguous()
        _, y = ops.ln_residual(xd, weight.detach().float(), bias.detach().float(), eps, want_res=False,
                               out2_dtype=out_dtype)
        ctx.save_for_backward(xd, weight)
        ctx.eps = eps
        return y

    @staticmethod
    def backward(ctx, dy):
        x, weight = ctx.saved_tensors
        if dy.dtype not in (torch.float32, torch.bfloat16):
            dy = dy.float()
        dx, dg, db = ops.layer_norm_bwd(x, dy, weight.detach(), ctx.eps)
        return dx.to(x.dtype), dg.to(weight.dtype), db.to(weight.dtype), None, None


class ParamCast:
    """Compute-dtype copies of a module tree's Linear parameters, refreshed by ONE multi-tensor copy per optimizer step
    (torch.autocast re-casts every weight at every use: 2 small kernels per Linear per step here before this cache).
    Copies are keyed by parameter identity and re-made when a parameter's version counter, storage or device moved."""

    def __init__(self):
        self.params, self.copies, self.stamp, self.index, self.cd = [], [], None, {}, None

    def refresh(self, module, cd):
        params = [p for m in module.modules() if isinstance(m, torch.nn.Linear)
                  for p in (m.weight, m.bias) if p is not None and p.dtype != cd]
        stamp = [(id(p), p._version, p.data_ptr()) for p in params]
        if stamp == self.stamp and cd == self.cd:
            return self
        if ([s[0] for s in stamp] != [s[0] for s in (self.stamp or [])] or cd != self.cd
                or any(c.device != p.device for c, p in zip(self.copies, params))):
            self.copies = [torch.empty_like(p, dtype=cd) for p in params]
            self.index = {id(p): c for p, c in zip(params, self.copies)}
        with torch.no_grad():
            if params:
                torch._foreach_copy_(self.copies, [p.detach() for p in params])
        self.params, self.stamp, self.cd = params, stamp, cd
        return self

    def get(self, p, cd):
        c = self.index.get(id(p))
        return c if c is not None and c.dtype == cd else None


_param_cast = None            # the ParamCast of the forward in flight (set by train.backbone_forward)


def set_param_cast(cache):
    global _param_cast
    _param_cast = cache


class LinearFn(torch.autograd.Function):
    """y = x W^T + b with operands cast to `cd` (torch.autocast's Linear): cuBLAS GEMMs both ways (the weight
    gradient accumulated and written in fp32 by the GEMM itself), the bias gradient by ptv3_column_sum instead of
    torch's generic dim-0 reduction.  `wc` / `bc`: compute-dtype copies from ParamCast, or None."""

    @staticmethod
    def forward(ctx, x, weight, bias, cd, wc, bc):
        xc = x.detach().to(cd)
        if wc is None:
            wc = weight.detach().to(cd)
        if bias is not None and bc is None:
            bc = bias.detach().to(cd)
        y = torch.nn.functional.linear(xc, wc, bc)
        ctx.save_for_backward(xc, wc)
        ctx.meta = (x.dtype, weight.dtype, bias is not None)
        return y

    @staticmethod
    def backward(ctx, dy):
        xc, wc = ctx.saved_tensors
        xdt, wdt, has_bias = ctx.meta
        dy = dy.contiguous().to(xc.dtype)
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            dx = (dy @ wc).to(xdt)
        if ctx.needs_input_grad[1]:
            if wdt == torch.float32 and xc.dtype != torch.float32:
                dw = torch.mm(dy.t(), xc, out_dtype=torch.float32)
            else:
                dw = (dy.t() @ xc).to(wdt)
        if has_bias and ctx.needs_input_grad[2]:
            db = ops.column_sum(dy).to(wdt)
        return dx, dw, db, None, None, None


def linear(x, lin, cd):
    c = _param_cast
    wc = c.get(lin.weight, cd) if c is not None else None
    bc = c.get(lin.bias, cd) if c is not None and lin.bias is not None else None
    return LinearFn.apply(x, lin.weight, lin.bias, cd, wc, bc)


def layer_norm(x, ln, out_dtype=torch.float32):
    """nn.LayerNorm `ln` applied to (N, C) rows; custom kernels for the PTv3 channel counts, torch otherwise."""
    c = x.shape[-1]
    if (x.dim() == 2 and c in ops.LN_BWD_CHANNELS and ln.elementwise_affine and ln.bias is not None
            and x.dtype in (torch.float32, torch.bfloat16)):
        return LayerNormFn.apply(x, ln.weight, ln.bias, ln.eps, out_dtype)
    return torch.nn.functional.layer_norm(x.float(), ln.normalized_shape, ln.weight, ln.bias, ln.eps).to(out_dtype)


def subm_conv(feat, weight, bias, nbr, cd):
    return SubMConvFn.apply(feat, weight, bias, nbr, cd)


def attn_window(qkv, order, windows, n_windows, num_heads, max_len, scale=None, algo=0):
    return WindowAttnFn.apply(qkv, order, windows, n_windows, num_heads, max_len, scale, algo)


def segment_csr(src, indptr, reduce="sum", gather=None):
    return SegmentCSRFn.apply(src, indptr, reduce, gather)


def gather_rows(x, inverse, indices, idx_ptr):
    return GatherRowsFn.apply(x, inverse, indices, idx_ptr)
