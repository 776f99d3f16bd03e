"""Device-resident serving path: everything between the raw scan and the answer stays on the
GPU -- no numpy, no DataLoader workers.

  collate                    datasets/utils.py:16-69 `collate_fn` for the keys of the hot path (concatenate, `offset`
                             = running row count)
  detect_keypoints           tools/OffsetKeyPointPrediction_Qt.py:403-434 `predict`: normalise to the unit sphere ->
                             first point per voxel (scan order) -> OffsetKeypointPTv3 -> argmax / weighted vote ->
                             back to the scan's own coordinates (`(vote) * scale + centroid`)
  segment_with_fragments     engines/test.py:185-204 + transform.py:897-925: GridSample in test mode cuts the scan into
                             count.max() fragments that together cover every raw point; each fragment is segmented and
                             its softmax is accumulated on the raw points it came from
"""
import torch
import torch.nn.functional as F

from . import ops
from .transform import GridSample, normalize_unit_sphere, voxelize_first


def collate(scenes):
    """list of per-scene dicts of device tensors -> one batch dict; every (n_i, ...) tensor is concatenated along
    dim 0, `offset` (or the row count when a scene has none) becomes the running total, scalars are kept from scene 0."""
    out, counts = {}, []
    for s in scenes:
        rows = next(s[k] for k in ("coord", "grid_coord", "feat") if k in s)
        counts.append(int(rows.shape[0]))
    for k in scenes[0]:
        v0 = scenes[0][k]
        if k == "offset":
            continue
        if torch.is_tensor(v0) and v0.dim() >= 1 and v0.shape[0] == counts[0] and all(
                torch.is_tensor(s[k]) and s[k].shape[0] == c for s, c in zip(scenes, counts)):
            out[k] = torch.cat([s[k] for s in scenes], 0)
        else:
            out[k] = v0
    dev = next(v.device for v in out.values() if torch.is_tensor(v))
    out["offset"] = torch.cumsum(torch.tensor(counts, dtype=torch.int64), 0).to(dev)
    return out


@torch.no_grad()
def detect_keypoints(model, coord_raw, feat_raw, grid_size, mode="argmax", mask_thresh=0.5):
    """coord_raw (R,3), feat_raw (R,C) device tensors of ONE scan in its own units -> keypoints (K,3) in the same
    units.  One host sync (the voxel count inside GridSample)."""
    coord_n, centroid, scale = normalize_unit_sphere(coord_raw)
    idx, grid = voxelize_first(coord_n, grid_size, keep_order=True)
    coord = coord_n[idx]
    batch = dict(coord=coord, grid_coord=grid, feat=feat_raw[idx].float(),
                 offset=torch.tensor([idx.numel()], device=coord.device), grid_size=grid_size)
    pred = model(batch)["pred"]
    kp = ops.keypoint_aggregate(coord, pred.float(), batch["offset"], mode=mode, mask_thresh=mask_thresh)[0]
    return kp * scale + centroid


@torch.no_grad()
def segment_with_fragments(model, data_dict, grid_size, num_classes, hash_type="fnv", feat_keys=("feat",)):
    """data_dict: raw scan (coord (R,3) + per-point features) on the device.  Returns (pred (R,) int64, prob (R,C))."""
    data = dict(data_dict)
    data.setdefault("index_valid_keys", ["coord"] + [k for k in feat_keys if k in data])
    gs = GridSample(grid_size=grid_size, hash_type=hash_type, mode="test", return_grid_coord=True)
    parts = gs(data)
    r = data_dict["coord"].shape[0]
    prob = torch.zeros((r, num_classes), dtype=torch.float32, device=data_dict["coord"].device)
    for part in parts:
        n = part["coord"].shape[0]
        batch = dict(coord=part["coord"], grid_coord=part["grid_coord"],
                     feat=torch.cat([part[k].float() for k in feat_keys], 1),
                     offset=torch.tensor([n], device=prob.device), grid_size=grid_size)
        logits = model(batch)["seg_logits"]
        prob.index_add_(0, part["index"], F.softmax(logits.float(), -1))
    return prob.argmax(1), prob
