"""The allocation-free entry point ptv3_argsort_rows_static (scratch in a file-scope `static __device__` array inside the
library; no counterpart in the reference) against the existing
ptv3_argsort_rows, through the C ABI, on two calls in a row.  Integer work: the results must be bit-identical.
Named test_zz_* so that pytest collects it last."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


def _p(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _codes(rows, n, key_bits, seed):
    g = torch.Generator().manual_seed(seed)
    hi = 1 << min(key_bits, 62)
    code = torch.randint(0, hi, (rows, n), generator=g, dtype=torch.int64)
    if n > 8:
        code[:, n // 2:n // 2 + n // 4] = code[:, :n // 4]          # duplicates: the stable order matters
    return code.cuda()


def _existing(lib, code, key_bits):
    rows, n = code.shape
    order = torch.empty_like(code)
    inverse = torch.empty_like(code)
    nbytes = lib.ptv3_argsort_rows_workspace(rows, n)
    ws = torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=code.device)
    rc = lib.ptv3_argsort_rows(_p(code), rows, n, key_bits, _p(order), _p(inverse), _p(ws), nbytes, _stream())
    assert rc == 0, lib.ptv3_error_string(rc)
    return order, inverse


def _static(lib, code, key_bits, want_inverse=True):
    rows, n = code.shape
    order = torch.full_like(code, -1)
    inverse = torch.full_like(code, -1) if want_inverse else None
    rc = lib.ptv3_argsort_rows_static(_p(code), rows, n, key_bits, _p(order), _p(inverse), _stream())
    assert rc == 0, lib.ptv3_error_string(rc)
    return order, inverse


def test_argsort_rows_static_matches_argsort_rows_on_two_calls_in_a_row():
    from ptv3_b200 import _lib
    lib = _lib.load()
    assert lib.ptv3_argsort_rows_static_max_elems() == 262144
    # first call: the 4 orders of a 60 000-voxel scene (5 radix passes); second call: a smaller, differently shaped one
    # that finds the first call's keys / histograms still in the shared scratch array
    calls = [(4, 60000, 34, 1), (3, 7001, 17, 2), (1, 262144, 63, 3), (4, 60000, 34, 1)]
    for rows, n, key_bits, seed in calls:
        code = _codes(rows, n, key_bits, seed)
        ref_order, ref_inverse = _existing(lib, code, key_bits)
        order, inverse = _static(lib, code, key_bits)
        torch.cuda.synchronize()
        assert torch.equal(order, ref_order), (rows, n, key_bits)
        assert torch.equal(inverse, ref_inverse), (rows, n, key_bits)
        # and against the definition: a stable argsort
        assert torch.equal(order.cpu(), torch.argsort(code.cpu(), dim=1, stable=True))
    order, inverse = _static(lib, code, 34, want_inverse=False)              # inverse may be NULL
    torch.cuda.synchronize()
    assert inverse is None and torch.equal(order, ref_order)


def test_argsort_rows_static_rejects_sizes_beyond_the_maximum():
    from ptv3_b200 import _lib
    lib = _lib.load()
    code = torch.zeros(8, dtype=torch.int64, device="cuda")
    order = torch.full((8,), -7, dtype=torch.int64, device="cuda")
    too_large = -5                                                            # PTV3_ETOOLARGE
    # sizes are checked before anything is read or enqueued, so the small buffers are never touched
    assert lib.ptv3_argsort_rows_static(_p(code), 1, 262145, 30, _p(order), None, _stream()) == too_large
    assert lib.ptv3_argsort_rows_static(_p(code), 4, 65537, 30, _p(order), None, _stream()) == too_large
    assert lib.ptv3_argsort_rows_static(_p(code), 262144, 1, 30, _p(order), None, _stream()) == too_large   # histograms
    assert b"capacity" in lib.ptv3_error_string(too_large)
    torch.cuda.synchronize()
    assert bool((order == -7).all())
    assert lib.ptv3_argsort_rows_static(_p(code), 0, 0, 30, None, None, _stream()) == 0
