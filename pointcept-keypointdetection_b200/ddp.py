"""Gradient synchronisation for the training config (BASELINE configs[4]: "NCCL grad all-reduce"; the reference
wraps the model in torch DistributedDataParallel, engines/train.py / defaults.py:22-43, launch.py:74).

One process per GPU, scenes sharded per rank (shard.py), replicas of the parameters; the only collective of a step is
the gradient all-reduce.  `GradSync` keeps every parameter's .grad as a VIEW into a few flat fp32 buckets (filled in
reverse registration order ~ the order backward produces them) and launches one asynchronous all-reduce per bucket
from a post-accumulate-grad hook as soon as the bucket's last gradient is written, so the NVLink/NVSwitch transfer of
bucket i overlaps the backward kernels of the layers below it.  NCCL (AVG) on GPUs, gloo (SUM, then divide) in the
CPU tests.  Buckets are sized for launch latency, not link count (NVSwitch gives full bandwidth to every peer).
"""
import torch
import torch.distributed as dist


class GradSync:
    def __init__(self, params, bucket_bytes=32 << 20, group=None):
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.params = [p for p in params if p.requires_grad]
        self.buckets, self._where, self._hooks = [], {}, []
        cur, cur_bytes = [], 0
        for p in reversed(self.params):
            cur.append(p)
            cur_bytes += p.numel() * 4
            if cur_bytes >= bucket_bytes:
                self._make_bucket(cur)
                cur, cur_bytes = [], 0
        if cur:
            self._make_bucket(cur)
        for p in self.params:
            self._hooks.append(p.register_post_accumulate_grad_hook(self._on_grad))
        self._works = []
        self.zero_grad()

    def _make_bucket(self, ps):
        dev = ps[0].device
        flat = torch.zeros(sum(p.numel() for p in ps), dtype=torch.float32, device=dev)
        views, o = [], 0
        for p in ps:
            assert p.dtype == torch.float32, "master parameters are fp32 (compute_dtype only casts operands)"
            views.append(flat[o:o + p.numel()].view_as(p))
            self._where[id(p)] = (len(self.buckets), views[-1])
            o += p.numel()
        self.buckets.append(dict(flat=flat, params=list(ps), views=views, pending=len(ps), launched=False))

    def zero_grad(self):
        """Replaces optimizer.zero_grad(): zero the buckets and (re)attach the views as .grad."""
        for b in self.buckets:
            b["flat"].zero_()
            b["pending"], b["launched"] = len(b["params"]), False
            for p, v in zip(b["params"], b["views"]):
                p.grad = v
        self._works = []

    def _launch(self, b):
        b["launched"] = True
        if self.world == 1:
            return
        if b["flat"].is_cuda:
            self._works.append((dist.all_reduce(b["flat"], op=dist.ReduceOp.AVG, group=self.group, async_op=True), None))
        else:
            self._works.append((dist.all_reduce(b["flat"], op=dist.ReduceOp.SUM, group=self.group, async_op=True), b))

    def _on_grad(self, p):
        bi, v = self._where[id(p)]
        b = self.buckets[bi]
        if p.grad is not None and p.grad.data_ptr() != v.data_ptr():
            # autograd replaced .grad (first accumulation into a None grad): copy into the bucket and re-attach
            v.copy_(p.grad)
            p.grad = v
        b["pending"] -= 1
        if b["pending"] == 0 and not b["launched"]:
            self._launch(b)

    def finish(self):
        """Call after backward(): reduces the buckets whose parameters got no gradient this step, then waits."""
        for b in self.buckets:
            if not b["launched"]:
                self._launch(b)
        for work, b in self._works:
            work.wait()
            if b is not None:
                b["flat"].div_(self.world)
        self._works = []

    def remove(self):
        for h in self._hooks:
            h.remove()


# ---------------------------------------------------------------------------------------------------------------
# SyncBatchNorm (engines/train.py:256-257: `if cfg.sync_bn: model = nn.SyncBatchNorm.convert_sync_batchnorm(model)`).
# PT-v3m1 has BatchNorm1d at the stem, the four poolings and the unpoolings; with scenes sharded per rank their batch
# statistics must span every rank's rows to equal the single-process step.  The modules stay nn.BatchNorm1d (same
# state_dict names, same eval-mode folding into the fused kernels); `convert_sync_batchnorm` only marks them, and the
# training forward (train._bn) then routes through `sync_batch_norm`:
#   forward : per rank (count, mean, M2) -> all_gather -> pooled mean / biased variance (Chan's parallel merge, the
#             cancellation-free form torch's batch_norm_gather_stats_with_counts uses), running stats updated with the
#             UNBIASED pooled variance like nn.BatchNorm1d;
#   backward: all_reduce of (sum dy, sum dy * xhat); dx uses the pooled count; d weight / d bias stay per-rank sums and
#             are averaged with every other parameter gradient by GradSync (what DDP does for nn.SyncBatchNorm).
# Two small collectives per BatchNorm and direction ((2 C + 1) and 2 C floats), torch ops only: runs on NCCL and gloo.
# ---------------------------------------------------------------------------------------------------------------
def convert_sync_batchnorm(model, process_group=None):
    """Marks every BatchNorm of `model` for cross-rank statistics in training; returns the model (same classes)."""
    for m in model.modules():
        if isinstance(m, torch.nn.modules.batchnorm._BatchNorm):
            m.ptv3_sync = True
            m.ptv3_sync_group = process_group
    return model


def _sync_active(bn):
    return (bool(getattr(bn, "ptv3_sync", False)) and bn.training and dist.is_available() and dist.is_initialized()
            and dist.get_world_size(getattr(bn, "ptv3_sync_group", None)) > 1)


class _SyncBNFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, eps, group):
        n, c = x.shape
        cnt = x.new_tensor([float(n)])
        if n > 0:
            mean_l = x.mean(0)
            m2_l = (x - mean_l).square().sum(0)
        else:                                  # a rank without rows (ragged shard) still takes part in the collectives
            mean_l, m2_l = x.new_zeros(c), x.new_zeros(c)
        mine = torch.cat([cnt, mean_l, m2_l])
        world = dist.get_world_size(group)
        parts = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(parts, mine, group=group)
        allp = torch.stack(parts)                                         # (world, 1 + 2 C)
        counts, means, m2s = allp[:, :1], allp[:, 1:1 + c], allp[:, 1 + c:]
        total = counts.sum()
        mean = (counts * means).sum(0) / total
        m2 = m2s.sum(0) + (counts * (means - mean).square()).sum(0)
        var = m2 / total                                                  # biased: what normalises the batch
        invstd = torch.rsqrt(var + eps)
        xhat = (x - mean) * invstd
        ctx.save_for_backward(xhat, weight, invstd)
        ctx.total, ctx.group = total, group
        ctx.mark_non_differentiable(mean, m2, total)
        y = xhat * weight + bias if weight is not None else xhat
        return y, mean, m2, total

    @staticmethod
    def backward(ctx, dy, _m, _v, _t):
        xhat, weight, invstd = ctx.saved_tensors
        c = xhat.shape[1]
        dy = dy.float()
        sum_dy, sum_dy_xhat = dy.sum(0), (dy * xhat).sum(0)
        red = torch.cat([sum_dy, sum_dy_xhat])
        dist.all_reduce(red, group=ctx.group)
        g_dy, g_dyx = red[:c] / ctx.total, red[c:] / ctx.total
        scale = invstd * weight if weight is not None else invstd
        dx = (dy - g_dy - xhat * g_dyx) * scale
        dw = sum_dy_xhat if weight is not None else None
        db = sum_dy if weight is not None else None
        return dx, dw, db, None, None


def sync_batch_norm(x, bn):
    """Training-mode BatchNorm1d over the rows of EVERY rank (x: (n_local, C) fp32).  Updates bn.running_mean /
    running_var / num_batches_tracked exactly as nn.BatchNorm1d would on the concatenated batch."""
    group = getattr(bn, "ptv3_sync_group", None)
    w, b = (bn.weight, bn.bias) if bn.affine else (None, None)
    y, mean, m2, total = _SyncBNFn.apply(x, w, b, bn.eps, group)
    if bn.track_running_stats:
        with torch.no_grad():
            bn.num_batches_tracked += 1
            mom = bn.momentum if bn.momentum is not None else 1.0 / float(bn.num_batches_tracked)
            unbiased = m2 / torch.clamp(total - 1.0, min=1.0)
            bn.running_mean.mul_(1.0 - mom).add_(mean.to(bn.running_mean.dtype), alpha=mom)
            bn.running_var.mul_(1.0 - mom).add_(unbiased.to(bn.running_var.dtype), alpha=mom)
    return y
