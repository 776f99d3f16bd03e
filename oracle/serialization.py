"""Oracle: space-filling-curve codes and serialized orders (numpy, int64).

Restates, as plain integer arithmetic, what the reference computes with torch
bit-matrix code:
  * Morton:  pointcept/models/utils/serialization/z_order.py:66-101 (xyz2key)
  * Hilbert: pointcept/models/utils/serialization/hilbert.py:91-198 (encode)
  * order dispatch + batch bits: serialization/default.py:9-24 (encode)
  * argsort / inverse: pointcept/models/utils/structure.py:52-109
Test infrastructure only (see oracle/__init__.py).
"""
import numpy as np

ORDERS = ("z", "z-trans", "hilbert", "hilbert-trans")


def _mask_to_depth(v, depth):
    # z_order.py:90-95 keeps 8 low bits and (depth-8) further bits; hilbert.py:143
    # truncates the bit matrix to the last num_bits columns.  Both == low `depth` bits.
    return v.astype(np.int64) & np.int64((1 << depth) - 1)


def z_order_encode(grid_coord, depth):
    """x -> bit 3i+2, y -> bit 3i+1, z -> bit 3i  (z_order.py:40-52)."""
    g = np.asarray(grid_coord)
    x = _mask_to_depth(g[:, 0], depth)
    y = _mask_to_depth(g[:, 1], depth)
    z = _mask_to_depth(g[:, 2], depth)
    key = np.zeros(g.shape[0], dtype=np.int64)
    for i in range(depth):
        bit = np.int64(1) << np.int64(i)
        key |= ((x & bit) << np.int64(2 * i + 2)) | ((y & bit) << np.int64(2 * i + 1)) | ((z & bit) << np.int64(2 * i))
    return key


def hilbert_encode(grid_coord, depth):
    """Skilling's transpose algorithm followed by Gray decoding of the
    interleaved word (hilbert.py:156-198): walk bits MSB->LSB, dims 0..2; a set
    bit inverts the lower bits of dim 0, a clear bit exchanges the lower bits of
    dim 0 and this dim; interleave with dim 0 most significant; then
    binary = prefix-xor of the gray word."""
    g = np.asarray(grid_coord)
    X = [_mask_to_depth(g[:, d], depth).copy() for d in range(3)]
    for bit in range(depth - 1, -1, -1):
        q = np.int64(1) << np.int64(bit)
        p = q - np.int64(1)
        for d in range(3):
            on = (X[d] & q) != 0
            # bit set: invert low bits of dim 0
            X[0] = np.where(on, X[0] ^ p, X[0])
            # bit clear: swap differing low bits between dim 0 and dim d
            t = np.where(on, np.int64(0), (X[0] ^ X[d]) & p)
            X[0] = X[0] ^ t
            X[d] = X[d] ^ t
    gray = np.zeros(g.shape[0], dtype=np.int64)
    for i in range(depth):
        bit = np.int64(1) << np.int64(i)
        gray |= ((X[0] & bit) << np.int64(2 * i + 2)) | ((X[1] & bit) << np.int64(2 * i + 1)) | ((X[2] & bit) << np.int64(2 * i))
    # gray2binary (hilbert.py:70-88): b = g ^ g>>1 ^ g>>2 ^ ...
    h = gray
    s = 1
    while s < 64:
        h = h ^ (h >> np.int64(s))
        s *= 2
    return h


def encode(grid_coord, batch, depth, order):
    """serialization/default.py:9-24."""
    assert order in ORDERS
    g = np.asarray(grid_coord)
    if order.endswith("-trans"):
        g = g[:, [1, 0, 2]]
    code = z_order_encode(g, depth) if order.startswith("z") else hilbert_encode(g, depth)
    if batch is not None:
        code = (np.asarray(batch).astype(np.int64) << np.int64(3 * depth)) | code
    return code


def serialized_depth(grid_coord):
    """structure.py:73."""
    return int(int(np.asarray(grid_coord).max()) + 1).bit_length()


def argsort_rows(code):
    """structure.py:92-99.  Stable so that the result is defined for duplicate
    codes too (the reference's torch.argsort is not; they coincide whenever the
    codes of a row are unique, which holds after GridSample / pooling)."""
    code = np.asarray(code)
    order = np.argsort(code, axis=1, kind="stable").astype(np.int64)
    inverse = np.empty_like(order)
    k, n = code.shape
    for r in range(k):
        inverse[r, order[r]] = np.arange(n, dtype=np.int64)
    return order, inverse


def serialization(grid_coord, batch, orders=ORDERS, depth=None, perm=None):
    """Point.serialization (structure.py:52-109).  `perm` is the row permutation
    the reference draws with torch.randperm when shuffle_orders=True."""
    if depth is None:
        depth = serialized_depth(grid_coord)
    nb = int(np.asarray(batch).max()) + 1 if len(batch) else 1
    assert depth * 3 + nb.bit_length() <= 63 and depth <= 16
    code = np.stack([encode(grid_coord, batch, depth, o) for o in orders])
    order, inverse = argsort_rows(code)
    if perm is not None:
        perm = np.asarray(perm)
        code, order, inverse = code[perm], order[perm], inverse[perm]
    return depth, code, order, inverse
