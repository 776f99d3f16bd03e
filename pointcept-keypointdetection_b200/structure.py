"""`Point` and `SparseConvTensor`: the reference's data surface over our kernels.

Mirrors pointcept/models/utils/structure.py:20-146 (Point: batch/offset
inference, serialization(), sparsify()), pointcept/models/utils/misc.py:13-34
(offset2batch / batch2offset / offset2bincount) and the part of
spconv.pytorch.SparseConvTensor the hot path touches (structure.py:137-146,
modules.py:84-108): features / indices / spatial_shape / batch_size /
indice_dict / replace_feature.
"""
import torch

from . import ops


@torch.no_grad()
def offset2bincount(offset):
    return torch.diff(offset, prepend=torch.zeros(1, device=offset.device, dtype=torch.long))


@torch.no_grad()
def offset2batch(offset):
    bincount = offset2bincount(offset)
    return torch.arange(len(bincount), device=offset.device, dtype=torch.long).repeat_interleave(bincount)


@torch.no_grad()
def batch2offset(batch):
    return torch.cumsum(batch.bincount(), dim=0).long()


_PERM_CACHE = {}


def device_perm(perm, device):
    """Row permutation drawn on the host (torch.randperm, CPU generator, like the reference) as a device index tensor.
    There are only k! of them: cached, so that a forward does not issue a synchronous pageable H2D copy per draw."""
    key = (device, tuple(perm.tolist()))
    t = _PERM_CACHE.get(key)
    if t is None:
        k = perm.numel()
        if k <= 5 and (device, tuple(range(k))) not in _PERM_CACHE:      # first draw on this device: upload all k!
            import itertools
            allp = torch.tensor(list(itertools.permutations(range(k))), dtype=perm.dtype).to(device)
            for i, pp in enumerate(itertools.permutations(range(k))):
                _PERM_CACHE[(device, pp)] = allp[i]
            t = _PERM_CACHE[key]
        else:
            t = _PERM_CACHE[key] = perm.to(device)
    return t


def z_info(point):
    """(code_z, order_z, depth) of a serialized Point -- the row of plain Morton codes and its argsort -- or None."""
    zr = point.get("_z_row")
    if zr is None or "serialized_code" not in point or int(point["serialized_depth"]) < 2:
        return None
    return point["serialized_code"][zr], point["serialized_order"][zr], int(point["serialized_depth"])


class SparseConvTensor:
    """features (N,C), indices (N,4) int32 [b,x,y,z]; `indice_dict` is the spconv-style cache shared by every
    tensor derived through replace_feature (holds the neighbour maps keyed by (indice_key, ksize))."""

    def __init__(self, features, indices, spatial_shape, batch_size, indice_dict=None):
        self.features = features
        self.indices = indices
        self.spatial_shape = spatial_shape
        self.batch_size = batch_size
        self.indice_dict = indice_dict if indice_dict is not None else {}

    def replace_feature(self, feature):
        return SparseConvTensor(feature, self.indices, self.spatial_shape, self.batch_size, self.indice_dict)

    @property
    def spatial_size(self):
        return self.spatial_shape


class Point(dict):
    """Dict with attribute access (the reference uses addict.Dict).  Keys with a defined meaning are the
    reference's: coord, grid_coord, feat, offset, batch, grid_size, serialized_depth / code / order / inverse,
    sparse_shape, sparse_conv_feat, pooling_inverse, pooling_parent, order."""

    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        if "batch" not in self and "offset" in self:
            rows = next((self[k] for k in ("feat", "coord", "grid_coord") if k in self), None)
            if rows is not None and self["offset"].is_cuda:
                self["batch"] = ops.offset_to_batch(self["offset"], rows.shape[0])   # row count known: no host sync
            else:
                self["batch"] = offset2batch(self["offset"])
        elif "offset" not in self and "batch" in self:
            self["offset"] = batch2offset(self["batch"])

    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError:
            raise AttributeError(k)

    def __setattr__(self, k, v):
        self[k] = v

    def __delattr__(self, k):
        del self[k]

    def serialization(self, order="z", depth=None, shuffle_orders=False):
        """structure.py:52-109.  Codes for all orders in one kernel, one batched radix argsort with the inverse
        fused into its last pass.  `torch.randperm` is drawn from the CPU generator exactly like the reference."""
        order = [order] if isinstance(order, str) else list(order)
        self["order"] = order
        assert "batch" in self
        if "grid_coord" not in self:
            assert {"grid_size", "coord"}.issubset(self.keys())
            self["grid_coord"] = torch.div(self.coord - self.coord.min(0)[0], self.grid_size,
                                           rounding_mode="trunc").int()
        if depth is None:
            gmax = ops.grid_max(self.grid_coord)           # device (3,)
            self["_grid_max"] = gmax
            depth = int(int(gmax.max().item()) + 1).bit_length()   # the one host sync the reference also has (:73)
        self["serialized_depth"] = depth
        assert depth * 3 + len(self.offset).bit_length() <= 63
        assert depth <= 16
        code = ops.encode_orders(self.grid_coord, self.batch, depth, order)
        key_bits = depth * 3 + max(1, (len(self.offset) - 1).bit_length())
        order_, inverse = ops.argsort_rows(code, key_bits=key_bits)
        z_row = order.index("z") if "z" in order else None       # row of the plain Morton codes (stem block search)
        if shuffle_orders:
            perm = torch.randperm(code.shape[0])
            if z_row is not None:
                z_row = perm.tolist().index(z_row)
            if not torch.equal(perm, torch.arange(code.shape[0])):
                perm = device_perm(perm, code.device)
                code, order_, inverse = code[perm], order_[perm], inverse[perm]
        self["_z_row"] = z_row
        self["serialized_code"] = code
        self["serialized_order"] = order_
        self["serialized_inverse"] = inverse

    def sparsify(self, pad=96):
        """structure.py:111-146.  sparse_shape: the reference syncs for max(grid_coord)+pad; we use the bound
        2**serialized_depth + pad unless the exact per-axis maximum is already on the host (no extra sync) --
        our kernels do not consume spatial_shape."""
        assert {"feat", "batch"}.issubset(self.keys())
        if "grid_coord" not in self:
            assert {"grid_size", "coord"}.issubset(self.keys())
            self["grid_coord"] = torch.div(self.coord - self.coord.min(0)[0], self.grid_size,
                                           rounding_mode="trunc").int()
        if "sparse_shape" in self:
            sparse_shape = self.sparse_shape
        elif "serialized_depth" in self:
            sparse_shape = [(1 << int(self.serialized_depth)) + pad] * 3
        else:
            sparse_shape = torch.add(torch.max(self.grid_coord, dim=0).values, pad).tolist()
        _, _, indices = ops.pool_gather_ints(self.grid_coord, self.batch, None, 0, want_grid=False, want_batch=False)
        self["sparse_shape"] = sparse_shape
        self["sparse_conv_feat"] = SparseConvTensor(self.feat, indices, sparse_shape, len(self.offset))
        z = z_info(self)
        if z is not None:       # lets the neighbour search of this level use 4^3-block occupancy masks (ops.subm_build_map)
            self["sparse_conv_feat"].indice_dict["_z"] = z
