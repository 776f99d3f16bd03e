"""GridSample on the device: the reference's transform with tensors instead of numpy arrays.

Mirrors pointcept/datasets/transform.py:825-963 (GridSample) and :24-51
(index_operator): same constructor, same dict contract ("index_valid_keys",
"grid_coord", "inverse", "min_coord", "displacement", "index"; a list of parts in
test mode).  Differences, all stated in SURVEY.md 8c: values are CUDA tensors;
within a voxel points are ordered by ascending original index (the reference's
np.argsort is unstable); the per-voxel random draw comes from `generator` (a
torch CPU generator; the reference uses np.random) or can be passed in as
data_dict["gridsample_rand"]; "sampled_index" (ScanNet data-efficient) is not
supported.  `quotient` selects the precision of `coord / np.array(grid_size)`
(transform.py:850): "f32" (default) is what the reference's pinned numpy 1.26.4
computes for float32 coordinates, "f64" what numpy >= 2 computes.
"""
import copy

import torch

from . import ops

DEFAULT_INDEX_VALID_KEYS = ["coord", "color", "normal", "superpoint", "strength", "segment", "instance"]


def index_operator(data_dict, index, duplicate=False):
    if "index_valid_keys" not in data_dict:
        data_dict["index_valid_keys"] = list(DEFAULT_INDEX_VALID_KEYS)
    if not duplicate:
        for key in data_dict["index_valid_keys"]:
            if key in data_dict:
                data_dict[key] = data_dict[key][index]
        return data_dict
    out = dict()
    for key in data_dict.keys():
        if key in data_dict["index_valid_keys"]:
            out[key] = data_dict[key][index]
        elif key == "index_valid_keys":
            out[key] = copy.copy(data_dict[key])
        else:
            out[key] = data_dict[key]
    return out


class GridSample(object):
    def __init__(self, grid_size=0.05, hash_type="fnv", mode="train", return_inverse=False, return_grid_coord=False,
                 return_min_coord=False, return_displacement=False, project_displacement=False, generator=None,
                 quotient="f32"):
        assert mode in ["train", "test"]
        assert hash_type in ["fnv", "ravel"]
        assert quotient in ["f32", "f64"]
        self.quotient = quotient
        self.grid_size = grid_size
        self.hash_type = hash_type
        self.mode = mode
        self.return_inverse = return_inverse
        self.return_grid_coord = return_grid_coord
        self.return_min_coord = return_min_coord
        self.return_displacement = return_displacement
        self.project_displacement = project_displacement
        self.generator = generator

    def _extras(self, part, data_dict, r, idx):
        if self.return_inverse:
            part["inverse"] = r["inverse"]
        if self.return_grid_coord:
            part["grid_coord"] = r["grid_coord"][idx]
            if "grid_coord" not in part["index_valid_keys"]:
                part["index_valid_keys"].append("grid_coord")
        if self.return_min_coord:
            part["min_coord"] = (r["min_grid"].double() * float(self.grid_size)).reshape(1, 3)
        if self.return_displacement:
            scaled = r["_coord"].double() / float(self.grid_size) - r["min_grid"].double()
            disp = scaled - r["grid_coord"].double() - 0.5
            if self.project_displacement:
                disp = torch.sum(disp * data_dict["normal"].double(), dim=-1, keepdim=True)
            part["displacement"] = disp[idx]
            if "displacement" not in part["index_valid_keys"]:
                part["index_valid_keys"].append("displacement")

    def __call__(self, data_dict):
        assert "coord" in data_dict.keys()
        if "sampled_index" in data_dict:
            raise NotImplementedError("GridSample: sampled_index (ScanNet data-efficient) is not on the hot path")
        coord = data_dict["coord"]
        n = coord.shape[0]
        rand = None
        if self.mode == "train":
            rand = data_dict.pop("gridsample_rand", None)
            if rand is None:
                # the reference draws randint(0, count.max(), count.size) % count; any non-negative draw modulo the
                # voxel population is an equally valid pick
                rand = torch.randint(0, 1 << 30, (n,), generator=self.generator).to(coord.device, non_blocking=True)
        r = ops.gridsample(coord, self.grid_size, self.hash_type, rand=rand, quotient=self.quotient)
        r["_coord"] = coord
        full = dict(data_dict)
        if self.mode == "train":
            idx_unique = r["idx_unique"]
            data_dict = index_operator(data_dict, idx_unique)
            self._extras(data_dict, full, r, idx_unique)
            return data_dict
        # test mode: count.max() partitions, the i-th takes point (i % count) of every voxel (:897-925)
        count, idx_sort = r["count"], r["idx_sort"]
        seg_start = torch.cumsum(count, 0) - count
        parts = []
        for i in range(int(count.max().item())):
            idx_part = idx_sort[seg_start + i % count]
            part = index_operator(full, idx_part, duplicate=True)
            part["index"] = idx_part
            self._extras(part, full, r, idx_part)
            parts.append(part)
        return parts


def normalize_unit_sphere(coord):
    """tools/OffsetKeyPointPrediction_Qt.py:262-273 (and offset_keypoint_dataset.py:166-177): centre on the centroid,
    divide by the largest distance to it.  Returns (normalised coord fp32, centroid (3,), scale ())."""
    c = coord.float()
    centroid = c.mean(0)
    centered = c - centroid
    scale = centered.norm(dim=1).max()
    scale = torch.where(scale < 1e-6, torch.ones_like(scale), scale)
    return centered / scale, centroid, scale


def voxelize_first(coord, grid_size, keep_order=True):
    """The reference's inference-time voxelisers (SURVEY 8a row A0'): `np.unique(floor(coord / grid), axis=0,
    return_index=True)` keeps the FIRST point of every voxel; the Qt server then sorts the indices back to scan order
    (tools/OffsetKeyPointPrediction_Qt.py:288-289, keep_order=True), pig-seg keeps lexicographic voxel order
    (tools/pigseg_shared.py:133-134, keep_order=False).  Returns (indices int64 (M,), grid_coord int64 (M,3) already
    shifted to start at 0).  One kernel chain (float32 quotient like numpy, ravel key, stable radix sort, run heads)."""
    r = ops.gridsample(coord, grid_size, "ravel", rand=None, quotient="f32")
    idx = r["idx_unique"]
    if keep_order:
        idx = torch.sort(idx)[0]
    return idx, r["grid_coord"][idx]
