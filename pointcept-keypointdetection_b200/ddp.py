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
