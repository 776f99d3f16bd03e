"""Oracle: GridSample voxel hashing / unique / representative selection (numpy).

Restates pointcept/datasets/transform.py:848-927 (GridSample.__call__),
:930-946 (ravel_hash_vec), :949-963 (fnv_hash_vec) with the two points the
reference leaves implementation-defined *defined* (SURVEY.md 8c):
  * within a voxel, points are ordered by ascending original index
    (= np.argsort(kind="stable"); the reference uses the unstable default);
  * the per-voxel random draw is an input array `rand` (the reference draws
    np.random.randint(0, count.max(), count.size)).
Test infrastructure only (see oracle/__init__.py).
"""
import numpy as np

FNV_OFFSET = np.uint64(14695981039346656037)
FNV_PRIME = np.uint64(1099511628211)


def fnv_hash_vec(arr):
    """transform.py:949-963: multiply-then-xor over whole uint64 words."""
    arr = np.asarray(arr).astype(np.int64).astype(np.uint64)
    h = np.full(arr.shape[0], FNV_OFFSET, dtype=np.uint64)
    for j in range(arr.shape[1]):
        h = h * FNV_PRIME
        h = h ^ arr[:, j]
    return h


def ravel_hash_vec(arr):
    """transform.py:930-946."""
    arr = np.asarray(arr).astype(np.int64)
    arr = (arr - arr.min(0)).astype(np.uint64)
    amax = arr.max(0).astype(np.uint64) + np.uint64(1)
    keys = np.zeros(arr.shape[0], dtype=np.uint64)
    for j in range(arr.shape[1] - 1):
        keys = keys + arr[:, j]
        keys = keys * amax[j + 1]
    keys = keys + arr[:, -1]
    return keys


def voxelize(coord, grid_size, quotient="f32"):
    """transform.py:850-855: `coord / np.array(grid_size)`.  For float32 coordinates the reference's pinned numpy
    1.26.4 (uv_requirements.txt:84) keeps the division in float32 (value-based casting treats the 0-d float64 array
    like a scalar): quotient="f32", the default.  numpy >= 2 (NEP 50) promotes to float64: quotient="f64".  Both are
    frozen from the unmodified reference class in tests/golden/gridsample.npz (keys *_q32 / plain)."""
    if quotient == "f32":
        scaled = np.asarray(coord).astype(np.float32) / np.float32(grid_size)
    else:
        scaled = np.asarray(coord).astype(np.float64) / np.float64(grid_size)
    grid = np.floor(scaled).astype(np.int64)
    gmin = grid.min(0)
    return grid - gmin, gmin


def grid_sample(coord, grid_size, hash_type="fnv", rand=None, quotient="f32"):
    """Returns dict(grid_coord (R,3) shifted, min_grid, key, idx_sort, count,
    inverse (R,), idx_unique (M,), seg_start (M,)).  rand=None picks the first
    point of every voxel (rand == 0)."""
    grid, gmin = voxelize(coord, grid_size, quotient)
    key = fnv_hash_vec(grid) if hash_type == "fnv" else ravel_hash_vec(grid)
    idx_sort = np.argsort(key, kind="stable")
    key_sort = key[idx_sort]
    uniq, inv_sorted, count = np.unique(key_sort, return_inverse=True, return_counts=True)
    seg_start = np.cumsum(np.insert(count, 0, 0)[:-1])
    if rand is None:
        rand = np.zeros(count.size, dtype=np.int64)
    idx_select = seg_start + np.asarray(rand).astype(np.int64) % count
    idx_unique = idx_sort[idx_select]
    inverse = np.zeros(key.shape[0], dtype=np.int64)
    inverse[idx_sort] = inv_sorted
    return dict(grid_coord=grid, min_grid=gmin, key=key, idx_sort=idx_sort, key_unique=uniq,
                count=count, seg_start=seg_start, inverse=inverse, idx_unique=idx_unique)


def grid_sample_test_parts(count, idx_sort):
    """mode='test' (transform.py:897-925): count.max() partitions, the i-th takes
    point (i % count) of every voxel."""
    seg_start = np.cumsum(np.insert(count, 0, 0)[:-1])
    return [idx_sort[seg_start + i % count] for i in range(int(count.max()))]


def voxelize_first(coord, grid_size, keep_order=True):
    """Inference-time voxelisers (SURVEY 8a row A0'): tools/OffsetKeyPointPrediction_Qt.py:276-296 (keep_order=True:
    first point per voxel, indices sorted back to scan order) and tools/pigseg_shared.py:132-146 (keep_order=False:
    lexicographic voxel order).  `coord / grid_size` with a Python float stays float32 in numpy.
    Returns (indices, grid_coord shifted to start at 0)."""
    coord = np.asarray(coord, dtype=np.float32)
    discrete = np.floor(coord / grid_size).astype(np.int32)
    discrete -= discrete.min(axis=0, keepdims=True)
    _, idx = np.unique(discrete, axis=0, return_index=True)
    if keep_order:
        idx = np.sort(idx)
    return idx, discrete[idx].astype(np.int64)
