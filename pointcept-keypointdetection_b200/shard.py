"""Scene-level sharding of a batch across ranks.

Scenes are independent in the PTv3 forward (batch id sits in the top bits of every serialized code, in column 0 of
the conv indices, and BatchNorm is in eval mode), so the path shards scene-by-scene with no data-path collective:
rank r runs the forward on its own scenes; results are gathered only if the caller wants them.
Reference launcher this mirrors: pointcept/engines/launch.py:74 (one process per GPU)."""
import torch


def assign_scenes(sizes, world_size):
    """Largest-first greedy balance by voxel count.  Returns a list (per rank) of scene indices, each ascending.
    Deterministic: ties go to the lowest rank."""
    loads = [0] * world_size
    out = [[] for _ in range(world_size)]
    for i in sorted(range(len(sizes)), key=lambda i: (-int(sizes[i]), i)):
        r = min(range(world_size), key=lambda r: (loads[r], r))
        out[r].append(i)
        loads[r] += int(sizes[i])
    return [sorted(x) for x in out]


def take_scenes(batch, scene_ids):
    """Sub-batch dict (coord, grid_coord, feat, offset, ...) holding only `scene_ids` (in that order)."""
    offset = batch["offset"]
    starts = torch.nn.functional.pad(offset, (1, 0))[:-1].tolist()
    ends = offset.tolist()
    rows = [torch.arange(starts[i], ends[i], device=offset.device) for i in scene_ids]
    idx = torch.cat(rows) if rows else torch.zeros(0, dtype=torch.long, device=offset.device)
    n = int(offset[-1]) if offset.numel() else 0
    out = {}
    for k, v in batch.items():
        if k == "offset":
            out[k] = torch.cumsum(torch.tensor([ends[i] - starts[i] for i in scene_ids], dtype=offset.dtype,
                                               device=offset.device), 0)
        elif torch.is_tensor(v) and v.dim() >= 1 and v.shape[0] == n:
            out[k] = v[idx]
        else:
            out[k] = v
    return out


def gather_rows(local_rows, scene_ids_per_rank, sizes, group=None):
    """All-gather per-point results (rows of this rank's scenes, scene after scene) and restore the original scene
    order.  Host-side convenience (NCCL or gloo); not part of the timed path."""
    import torch.distributed as dist
    world = dist.get_world_size(group)
    counts = [sum(int(sizes[i]) for i in ids) for ids in scene_ids_per_rank]
    width = local_rows.shape[1:]
    bufs = [torch.empty((c,) + tuple(width), dtype=local_rows.dtype, device=local_rows.device) for c in counts]
    pad = max(counts)
    send = torch.zeros((pad,) + tuple(width), dtype=local_rows.dtype, device=local_rows.device)
    send[: local_rows.shape[0]] = local_rows
    recv = [torch.empty_like(send) for _ in range(world)]
    dist.all_gather(recv, send, group=group)
    pieces = {}
    for r, ids in enumerate(scene_ids_per_rank):
        bufs[r] = recv[r][: counts[r]]
        o = 0
        for i in ids:
            pieces[i] = bufs[r][o:o + int(sizes[i])]
            o += int(sizes[i])
    return torch.cat([pieces[i] for i in range(len(sizes))])
