"""CPU tests of the host side: the C-ABI library loads and exports every symbol include/ptv3_b200.h declares (no
compute calls), argument validation that needs no device, and the scene sharding over world_size-2 gloo."""
import ctypes
import os
import re
import socket

import pytest
import torch

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    src = open(os.path.join(REPO, "include", "ptv3_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(ptv3_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from ptv3_b200 import _lib
    if not os.path.exists(_lib.LIB_PATH):
        import __graft_entry__ as g
        g.build()
    lib = ctypes.CDLL(_lib.LIB_PATH)
    names = _declared_symbols()
    assert len(names) >= 25
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/ptv3_b200.h but not exported"
    assert set(_lib.SIGNATURES) == set(names), set(_lib.SIGNATURES) ^ set(names)


def test_host_side_queries_need_no_device():
    from ptv3_b200 import _lib
    lib = _lib.load()
    assert lib.ptv3_abi_version() == 1
    assert lib.ptv3_error_string(-2).decode()
    assert lib.ptv3_argsort_rows_workspace(4, 1000) > 0
    assert lib.ptv3_subm_map_workspace(1000) >= 2048 * 12
    assert lib.ptv3_subm_conv_tc_supported(64, 64, 27) == 1
    assert lib.ptv3_subm_conv_tc_supported(4, 32, 125) == 0
    assert lib.ptv3_subm_pack_weight_bytes(32, 32, 27) == 14 * 32 * 128
    assert lib.ptv3_subm_pack_weight_bytes(512, 512, 27) == 27 * 8 * 512 * 128
    # argument validation happens before any launch
    assert lib.ptv3_encode(None, 1, None, 10, 17, (ctypes.c_int * 1)(0), 1, None, None) != 0
    assert lib.ptv3_subm_build_map(None, 10, 4, None, None, 0, None) != 0


def test_ops_refuse_cpu_tensors():
    from ptv3_b200 import ops
    with pytest.raises(RuntimeError):
        ops.encode(torch.zeros(4, 3, dtype=torch.int64), None, 4, "z")
    with pytest.raises(RuntimeError):
        ops.segment_csr(torch.zeros(4, 8), torch.tensor([0, 4]))


def test_assign_scenes_balanced_and_complete():
    from ptv3_b200.shard import assign_scenes
    sizes = [72934, 50000, 73000, 10, 64000, 72000, 100, 30000]
    for w in (1, 2, 3, 4, 8):
        parts = assign_scenes(sizes, w)
        assert sorted(i for p in parts for i in p) == list(range(len(sizes)))
        loads = [sum(sizes[i] for i in p) for p in parts]
        assert max(loads) - min(loads) <= max(sizes)
    assert assign_scenes([5, 5, 5, 5], 2) == [[0, 2], [1, 3]]
    assert assign_scenes([], 2) == [[], []]


def _worker(rank, world, port, q):
    import torch.distributed as dist
    from ptv3_b200.shard import assign_scenes, gather_rows, take_scenes
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = torch.Generator().manual_seed(0)
        sizes = [7, 3, 9, 1, 4]
        n = sum(sizes)
        batch = dict(coord=torch.randn(n, 3, generator=g), feat=torch.randn(n, 4, generator=g),
                     grid_coord=torch.randint(0, 50, (n, 3), generator=g),
                     offset=torch.cumsum(torch.tensor(sizes), 0), grid_size=0.02)
        parts = assign_scenes(sizes, world)
        mine = take_scenes(batch, parts[rank])
        assert mine["offset"].tolist() == torch.cumsum(torch.tensor([sizes[i] for i in parts[rank]]), 0).tolist()
        assert mine["grid_size"] == 0.02
        # stand-in for the per-point forward: a row-wise function, so sharding must not change the result
        local = torch.cat([mine["coord"] * 2.0, mine["feat"].sum(1, keepdim=True)], 1)
        full = gather_rows(local, parts, sizes)
        want = torch.cat([batch["coord"] * 2.0, batch["feat"].sum(1, keepdim=True)], 1)
        t = torch.tensor([float(local.shape[0])])
        dist.all_reduce(t)                                # units processed by all ranks = whole batch
        q.put((rank, bool(torch.equal(full, want)), int(t.item()) == n))
    finally:
        dist.destroy_process_group()


def test_scene_sharding_world_size_2_gloo():
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert sorted(res) == [(0, True, True), (1, True, True)]


def _grad_worker(rank, world, port, q):
    import torch.distributed as dist
    from ptv3_b200.ddp import GradSync
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.manual_seed(0)                       # replicas start from identical parameters
        net = torch.nn.Sequential(torch.nn.Linear(8, 16), torch.nn.GELU(), torch.nn.Linear(16, 16), torch.nn.GELU(),
                                  torch.nn.Linear(16, 4))
        unused = torch.nn.Parameter(torch.ones(3))             # a parameter that gets no gradient
        sync = GradSync(list(net.parameters()) + [unused], bucket_bytes=300)      # three small buckets
        assert len(sync.buckets) >= 3
        g = torch.Generator().manual_seed(1)
        data = torch.randn(world, 5, 8, generator=g)
        out = []
        for step in range(2):                      # second step checks zero_grad() re-attaches the bucket views
            sync.zero_grad()
            net(data[rank] * (step + 1)).square().mean().backward()
            sync.finish()
            out.append([p.grad.numpy().copy() for p in net.parameters()])   # numpy pickles by value (a tensor would
        q.put((rank, out, unused.grad.numpy().copy()))                      # travel as an fd of a process that exits)
    finally:
        dist.destroy_process_group()


def test_grad_sync_world_size_2_gloo():
    """Training config: replicas + one bucketed gradient all-reduce per step; the averaged gradients must
    equal the mean of the per-rank gradients computed in one process."""
    import torch.multiprocessing as mp
    world, port = 2, 29500 + (os.getpid() % 2000) + 7
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_grad_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=120) for _ in procs], key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    torch.manual_seed(0)
    net = torch.nn.Sequential(torch.nn.Linear(8, 16), torch.nn.GELU(), torch.nn.Linear(16, 16), torch.nn.GELU(),
                              torch.nn.Linear(16, 4))
    g = torch.Generator().manual_seed(1)
    data = torch.randn(world, 5, 8, generator=g)
    for step in range(2):
        want = None
        for r in range(world):
            net.zero_grad()
            net(data[r] * (step + 1)).square().mean().backward()
            gr = [p.grad.clone() / world for p in net.parameters()]
            want = gr if want is None else [a + b for a, b in zip(want, gr)]
        for r in range(world):
            for a, b in zip(res[r][1][step], want):
                assert torch.allclose(torch.from_numpy(a), b, atol=1e-6)
    assert all((r[2] == 0).all() for r in res)


_SBN_ROWS = [[5, 9], [7, 0], [1, 1]]          # rows per rank for three consecutive steps (one rank empty, two single rows)


def _sbn_data(step, rank, c=6):
    g = torch.Generator().manual_seed(100 * step + rank)
    n = _SBN_ROWS[step][rank]
    return torch.randn(n, c, generator=g) * (1.0 + rank) + 3.0 * step, torch.randn(n, c, generator=g)


def _syncbn_worker(rank, world, port, q):
    import torch.distributed as dist
    from ptv3_b200 import ddp, train
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.manual_seed(0)
        bn = torch.nn.BatchNorm1d(6, eps=1e-3, momentum=0.01)
        with torch.no_grad():
            bn.weight.copy_(torch.linspace(0.5, 1.5, 6))
            bn.bias.copy_(torch.linspace(-1, 1, 6))
        holder = torch.nn.Sequential(bn)
        assert ddp.convert_sync_batchnorm(holder) is holder and type(bn) is torch.nn.BatchNorm1d and bn.ptv3_sync
        out = []
        for step in range(len(_SBN_ROWS)):
            x, gy = _sbn_data(step, rank)
            x.requires_grad_(True)
            bn.zero_grad()
            y = train._bn(x, bn)                       # the call site of the training forward
            (y * gy).sum().backward()
            out.append([t.detach().numpy().copy() for t in (y, x.grad, bn.weight.grad, bn.bias.grad)])
        stats = [bn.running_mean.numpy().copy(), bn.running_var.numpy().copy(), int(bn.num_batches_tracked)]
        bn.eval()                                      # eval(): plain running statistics, no collective
        x, _ = _sbn_data(0, rank)
        ev = train._bn(x, bn).detach().numpy().copy()
        q.put((rank, out, stats, ev))
    finally:
        dist.destroy_process_group()


def test_sync_batchnorm_world_size_2_gloo():
    """cfg.sync_bn (engines/train.py:256-257): BatchNorm statistics over the rows of every rank.  Two gloo ranks with
    unequal (and empty) shards must reproduce nn.BatchNorm1d on the concatenated batch: outputs, input gradients,
    parameter gradients (per-rank sums; GradSync averages them like DDP) and the running statistics."""
    import torch.multiprocessing as mp
    world, port = 2, 29500 + (os.getpid() % 2000) + 23
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_syncbn_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=120) for _ in procs], key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    bn = torch.nn.BatchNorm1d(6, eps=1e-3, momentum=0.01)
    with torch.no_grad():
        bn.weight.copy_(torch.linspace(0.5, 1.5, 6))
        bn.bias.copy_(torch.linspace(-1, 1, 6))
    for step in range(len(_SBN_ROWS)):
        xs, gs = zip(*[_sbn_data(step, r) for r in range(world)])
        x = torch.cat(xs).requires_grad_(True)
        bn.zero_grad()
        y = bn(x)
        (y * torch.cat(gs)).sum().backward()
        o = 0
        for r in range(world):
            n = _SBN_ROWS[step][r]
            yr, dxr, _, _ = res[r][1][step]
            assert torch.allclose(torch.from_numpy(yr), y[o:o + n].detach(), atol=2e-5), (step, r)
            assert torch.allclose(torch.from_numpy(dxr), x.grad[o:o + n], atol=2e-5), (step, r)
            o += n
        dw = sum(torch.from_numpy(res[r][1][step][2]) for r in range(world))
        db = sum(torch.from_numpy(res[r][1][step][3]) for r in range(world))
        assert torch.allclose(dw, bn.weight.grad, atol=2e-5) and torch.allclose(db, bn.bias.grad, atol=2e-5)
    for r in range(world):
        rm, rv, nb = res[r][2]
        assert torch.allclose(torch.from_numpy(rm), bn.running_mean, atol=1e-6)
        assert torch.allclose(torch.from_numpy(rv), bn.running_var, atol=1e-6) and nb == len(_SBN_ROWS)
        bn.eval()
        assert torch.allclose(torch.from_numpy(res[r][3]), bn(_sbn_data(0, r)[0]).detach(), atol=1e-6)
        bn.train()


def _eval_reduce_worker(rank, world, port, q):
    import torch.distributed as dist
    from ptv3_b200.evaluator import reduce_sums, summarize
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        k = 3          # per-rank running sums [total_dist, total_samples, dist_per_kp[k], samples_per_kp[k]]
        acc = torch.tensor([[4.0, 2, 1.5, 2.5, 0.0, 2, 2, 0], [9.0, 3, 3.0, 3.0, 3.0, 3, 3, 1]][rank], dtype=torch.float64)
        q.put((rank,) + summarize(reduce_sums(acc), k))
    finally:
        dist.destroy_process_group()


def test_evaluator_sums_world_size_2_gloo():
    """Validation metric across ranks (offset_keypoint_evaluator.py:94-103): the (2 + 2K) sums are all-reduced, then every
    rank derives the same means."""
    import torch.multiprocessing as mp
    world, port = 2, 29500 + (os.getpid() % 2000) + 11
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_eval_reduce_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=120) for _ in procs], key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, mean_dist, per_kp, counts in res:
        assert mean_dist == 13.0 / (5 + 1e-6)
        assert per_kp == [4.5 / (5 + 1e-6), 5.5 / (5 + 1e-6), 3.0 / (1 + 1e-6)] and counts == [5, 5, 1]


def test_reference_builds_over_shims():
    """VERDICT r01 weak #4 / north star "the reference runs unchanged": in a fresh interpreter ptv3_b200.shims.install(
    force=True) registers OUR spconv / torch_scatter / flash_attn modules, the UNMODIFIED reference files
    point_transformer_v3m1_base.py + offset_keypoint_ptv3.py are imported over them, the model is built and its state_dict
    names / shapes equal the frozen reference ones (tests/golden/param_shapes.json).  Needs /root/reference: build
    container only; the GPU-side half is tests/test_gpu_shims.py (replay of the frozen call trace)."""
    import json
    import subprocess
    import sys
    sys.path.insert(0, os.path.join(REPO, "tests", "golden"))
    import ref_loader
    if not ref_loader.available():
        pytest.skip("reference tree not present (GPU box)")
    out = subprocess.run([sys.executable, os.path.join(REPO, "tests", "golden", "make_shim_trace.py"), "structure"],
                         capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    got = json.loads(out.stdout.strip().splitlines()[-1])
    with open(os.path.join(REPO, "tests", "golden", "param_shapes.json")) as f:
        ref = json.load(f)
    assert got == ref


def test_shim_trace_covers_every_third_party_call_site():
    """The frozen call sequence of one reference forward (tests/golden/shim_trace.json): 4 encodes, 23 SubMConv3d
    forwards (stem + 22 Blocks), 22 flash-attention calls, 8 segment_csr calls (max + mean per pooling) -- exactly the
    call sites SURVEY.md 8b lists; shims.py exposes an entry point for each."""
    import json
    from ptv3_b200 import shims
    with open(os.path.join(REPO, "tests", "golden", "shim_trace.json")) as f:
        tr = json.load(f)
    kinds = {}
    for c in tr["calls"]:
        kinds[c["op"]] = kinds.get(c["op"], 0) + 1
    assert kinds == {"encode": 4, "SubMConv3d.forward": 23, "flash_attn_varlen_qkvpacked_func": 22, "segment_csr": 8}
    assert callable(shims.segment_csr) and callable(shims.flash_attn_varlen_qkvpacked_func)
    assert hasattr(shims.SubMConv3d, "forward")


def test_sonata_state_dict_matches_reference():
    """PT-v3m2 modules carry the reference's state_dict names and shapes (tests/golden/param_shapes_sonata.json, written
    from the unmodified reference by make_golden.py sonata), so its checkpoints load unchanged."""
    import json
    import sonata_common as SC
    from ptv3_b200 import model as M
    with open(os.path.join(REPO, "tests", "golden", "param_shapes_sonata.json")) as f:
        shapes = json.load(f)
    for tag, cfg in SC.CFGS.items():
        m = M.DefaultSegmentorV2(cfg["num_classes"], cfg["out_channels"], backbone=dict(type="PT-v3m2", **cfg["backbone"]))
        got = {k: list(v.shape) for k, v in m.state_dict().items() if not k.endswith("num_batches_tracked")}
        assert got == shapes[tag], (tag, set(got) ^ set(shapes[tag]))


class _Toy(torch.nn.Module):
    def __init__(self):
        super().__init__()
        torch.manual_seed(0)
        self.net = torch.nn.Sequential(torch.nn.Linear(4, 8), torch.nn.GELU(), torch.nn.Linear(8, 1))

    def forward(self, d):
        return dict(loss=(self.net(d["x"]) * d["w"]).square().mean())


def _ref_run_step(model, opt, sched, scaler, batches, accum, clip):
    """The reference sequence, engines/train.py:196-239, written out literally."""
    count, stepped = 0, []
    for b in batches:
        if count == 0:
            opt.zero_grad()
        loss = model(b)["loss"] / accum
        if scaler is not None:
            scaler.scale(loss).backward()
        else:
            loss.backward()
        count += 1
        if count >= accum:
            if scaler is not None:
                scaler.unscale_(opt)
                if clip is not None:
                    torch.nn.utils.clip_grad_norm_(model.parameters(), clip)
                scaler.step(opt)
                scale = scaler.get_scale()
                scaler.update()
                if scale <= scaler.get_scale():
                    sched.step()
                    stepped.append(True)
                else:
                    stepped.append(False)
            else:
                if clip is not None:
                    torch.nn.utils.clip_grad_norm_(model.parameters(), clip)
                opt.step()
                sched.step()
                stepped.append(True)
            count = 0
    return stepped


@pytest.mark.parametrize("amp", [False, True])
def test_train_step_follows_reference_run_step(amp):
    """train.TrainStep against the reference's Trainer.run_step order of operations (engines/train.py:184-245) on a toy
    model: gradient accumulation, clipping after unscale_, GradScaler skip (a non-finite batch) holding the scheduler
    back.  CPU: torch.amp.GradScaler('cpu'); the model-side mixed precision (compute_dtype) is covered by the GPU tests."""
    from ptv3_b200 import train
    g = torch.Generator().manual_seed(5)
    batches = [dict(x=torch.randn(6, 4, generator=g), w=torch.ones(6, 1)) for _ in range(8)]
    if amp:
        batches[4]["w"] = torch.full((6, 1), float("inf"))       # the third optimizer step overflows -> skipped
    out = []
    for mine in (False, True):
        m = _Toy()
        opt = torch.optim.AdamW(m.parameters(), lr=1e-2, weight_decay=0.05)
        sched = torch.optim.lr_scheduler.StepLR(opt, step_size=1, gamma=0.5)
        scaler = torch.amp.GradScaler(device="cpu", init_scale=1024.0) if amp else None
        if mine:
            ts = train.TrainStep(m, opt, sched, enable_amp=amp, clip_grad=0.5, gradient_accumulation_steps=2, scaler=scaler)
            stepped = []
            for i, b in enumerate(batches):
                ts(b)
                if i % 2 == 1:
                    stepped.append(ts.stepped)
        else:
            stepped = _ref_run_step(m, opt, sched, scaler, batches, 2, 0.5)
        out.append((stepped, [p.detach().clone() for p in m.parameters()], sched.get_last_lr(),
                    scaler.get_scale() if amp else None))
    assert out[0][0] == out[1][0] == ([True, True, False, True] if amp else [True] * 4)
    assert all(torch.equal(a, b) for a, b in zip(out[0][1], out[1][1]))
    assert out[0][2] == out[1][2] and out[0][3] == out[1][3]
    with pytest.raises(NotImplementedError):
        train.TrainStep(_Toy(), None, enable_amp=True, amp_dtype="float16")
