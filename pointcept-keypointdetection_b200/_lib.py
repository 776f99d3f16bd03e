"""ctypes binding of libptv3_b200.so (the flat C ABI declared in include/ptv3_b200.h).

There is no CPU fallback: if the shared library is missing or a call returns a
non-zero code a RuntimeError is raised (the reference's convention is Python
asserts/exceptions, pointcept/models/utils/structure.py:59,76,81).
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libptv3_b200.so")

c_i64, c_int, c_vp, c_sz = ctypes.c_int64, ctypes.c_int, ctypes.c_void_p, ctypes.c_size_t
c_f32, c_f64 = ctypes.c_float, ctypes.c_double

class BlockParams(ctypes.Structure):
    """ptv3_block_params (include/ptv3_b200.h)."""
    _fields_ = [(k, ctypes.c_int) for k in ("c", "heads", "hidden", "k3", "patch", "mlp_fused", "attn_algo")] + \
               [(k, ctypes.c_float) for k in ("scale", "eps_cpe", "eps_norm2")] + \
               [(k, ctypes.c_void_p) for k in ("conv_w_packed", "conv_bias", "cpe_lin_w", "cpe_lin_b", "cpe_ln_g",
                                               "cpe_ln_b", "norm1_g", "norm1_b", "norm2_g", "norm2_b", "qkv_w", "qkv_b",
                                               "proj_w", "proj_b", "fc1_w", "fc1_b", "fc2_w", "fc2_b", "drop_attn",
                                               "drop_mlp", "proj_w_packed", "proj_b_f32")]


# name -> (restype, argtypes); order mirrors include/ptv3_b200.h
SIGNATURES = {
    "ptv3_abi_version": (c_int, []),
    "ptv3_error_string": (ctypes.c_char_p, [c_int]),
    "ptv3_launch_count": (c_i64, []),
    "ptv3_encode": (c_int, [c_vp, c_int, c_vp, c_i64, c_int, ctypes.POINTER(c_int), c_int, c_vp, c_vp]),
    "ptv3_grid_max": (c_int, [c_vp, c_int, c_i64, c_vp, c_vp]),
    "ptv3_argsort_rows_workspace": (c_sz, [c_i64, c_i64]),
    "ptv3_argsort_rows": (c_int, [c_vp, c_i64, c_i64, c_int, c_vp, c_vp, c_vp, c_sz, c_vp]),
    "ptv3_argsort_rows_static_max_elems": (c_i64, []),
    "ptv3_argsort_rows_static": (c_int, [c_vp, c_i64, c_i64, c_int, c_vp, c_vp, c_vp]),
    "ptv3_gridsample_workspace": (c_sz, [c_i64]),
    "ptv3_gridsample": (c_int, [c_vp, c_i64, c_f64, c_int, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp,
                                c_vp, c_sz, c_vp]),
    "ptv3_gridsample_batch_workspace": (c_sz, [c_i64, c_i64]),
    "ptv3_gridsample_batch": (c_int, [c_vp, c_i64, c_vp, c_i64, c_f64, c_int, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp,
                                      c_vp, c_vp, c_vp, c_vp, c_sz, c_vp]),
    "ptv3_grid_pool_workspace": (c_sz, [c_i64]),
    "ptv3_grid_pool": (c_int, [c_vp, c_int, c_vp, c_i64, c_int, c_int, c_int, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp,
                               c_vp, c_sz, c_vp]),
    "ptv3_pool_front_workspace": (c_sz, [c_i64, c_i64]),
    "ptv3_pool_front": (c_int, [c_vp, c_vp, c_i64, c_i64, c_int, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp,
                                c_vp, c_sz, c_vp]),
    "ptv3_pool_gather_ints": (c_int, [c_vp, c_vp, c_vp, c_i64, c_int, c_vp, c_vp, c_vp, c_vp]),
    "ptv3_offset_to_batch": (c_int, [c_vp, c_i64, c_i64, c_vp, c_vp]),
    "ptv3_batch_to_offset": (c_int, [c_vp, c_i64, c_i64, c_vp, c_vp]),
    "ptv3_segment_csr": (c_int, [c_vp, c_int, c_vp, c_vp, c_i64, c_i64, c_int, c_vp, c_vp]),
    "ptv3_unpool_add": (c_int, [c_vp, c_vp, c_int, c_vp, c_i64, c_i64, c_int, c_vp, c_vp, c_vp]),
    "ptv3_subm_map_workspace": (c_sz, [c_i64]),
    "ptv3_subm_build_map": (c_int, [c_vp, c_i64, c_int, c_vp, c_vp, c_sz, c_vp]),
    "ptv3_subm_stem_fused": (c_int, [c_vp, c_i64, c_int, c_vp, c_vp, c_vp, c_int, c_vp, c_vp, c_vp, c_sz, c_vp]),
    "ptv3_subm_stem_z_workspace": (c_sz, [c_i64]),
    "ptv3_subm_build_map_z": (c_int, [c_vp, c_i64, c_int, c_vp, c_vp, c_int, c_vp, c_vp, c_sz, c_vp]),
    "ptv3_subm_stem_fused_z": (c_int, [c_vp, c_i64, c_int, c_vp, c_vp, c_int, c_vp, c_vp, c_vp, c_int, c_vp, c_vp, c_vp,
                                       c_sz, c_vp]),
    "ptv3_subm_conv_workspace": (c_sz, [c_i64, c_int, c_int, c_int, c_int, c_int]),
    "ptv3_subm_conv": (c_int, [c_vp, c_vp, c_vp, c_vp, c_i64, c_int, c_int, c_int, c_int, c_int, c_vp, c_vp, c_sz,
                               c_vp]),
    "ptv3_subm_conv_tc_supported": (c_int, [c_int, c_int, c_int]),
    "ptv3_subm_pack_weight_bytes": (c_sz, [c_int, c_int, c_int]),
    "ptv3_subm_pack_weight": (c_int, [c_vp, c_int, c_int, c_int, c_vp, c_vp]),
    "ptv3_subm_conv_packed": (c_int, [c_vp, c_vp, c_vp, c_vp, c_i64, c_int, c_int, c_int, c_vp, c_vp]),
    "ptv3_build_windows": (c_int, [c_vp, c_i64, c_i64, c_vp, c_i64, c_vp, c_vp]),
    "ptv3_attn_window_workspace": (c_sz, [c_i64, c_int]),
    "ptv3_attn_window_fwd": (c_int, [c_vp, c_int, c_vp, c_vp, c_vp, c_i64, c_i64, c_i64, c_int, c_int, c_f32, c_int,
                                     c_vp, c_vp, c_sz, c_vp]),
    "ptv3_pack_kmajor_bytes": (c_sz, [c_int, c_int]),
    "ptv3_pack_kmajor": (c_int, [c_vp, c_int, c_int, c_int, c_vp, c_vp]),
    "ptv3_mlp_fused_supported": (c_int, [c_int, c_int]),
    "ptv3_mlp_fc1_tile_rows": (c_int, [c_int]),
    "ptv3_mlp_fused": (c_int, [c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_i64, c_int, c_int, c_vp, c_vp, c_vp]),
    "ptv3_mlp_fused_rs": (c_int, [c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_i64, c_int, c_int, c_vp, c_vp, c_vp]),
    "ptv3_proj_mlp_fused": (c_int, [c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_f32, c_vp, c_vp, c_vp, c_vp, c_vp, c_i64,
                                    c_int, c_int, c_vp, c_vp, c_vp]),
    "ptv3_block_workspace": (c_sz, [ctypes.POINTER(BlockParams), c_i64, c_i64]),
    "ptv3_block_fwd": (c_int, [ctypes.POINTER(BlockParams), c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_i64, c_i64, c_vp, c_vp,
                               c_vp, c_sz, c_vp]),
    "ptv3_debug_set": (c_int, [c_int, c_int]),
    "ptv3_device_error": (c_int, []),
    "ptv3_ln_residual": (c_int, [c_vp, c_int, c_vp, c_vp, c_f32, c_vp, c_vp, c_vp, c_vp, c_vp, c_int, c_i64, c_int,
                                 c_vp]),
    "ptv3_ln_residual_rs": (c_int, [c_vp, c_int, c_vp, c_vp, c_f32, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_int, c_i64,
                                    c_int, c_vp]),
    "ptv3_affine_act": (c_int, [c_vp, c_int, c_vp, c_vp, c_int, c_i64, c_int, c_vp, c_vp, c_vp]),
    "ptv3_subm_wgrad": (c_int, [c_vp, c_vp, c_vp, c_i64, c_int, c_int, c_int, c_int, c_vp, c_vp]),
    "ptv3_attn_window_bwd_workspace": (c_sz, [c_i64, c_int]),
    "ptv3_attn_window_fwd_lse": (c_int, [c_vp, c_int, c_vp, c_vp, c_vp, c_i64, c_i64, c_i64, c_int, c_int, c_f32, c_vp,
                                         c_vp, c_vp]),
    "ptv3_attn_window_bwd": (c_int, [c_vp, c_vp, c_vp, c_int, c_vp, c_vp, c_vp, c_i64, c_i64, c_i64, c_int, c_int, c_f32,
                                     c_vp, c_vp, c_vp, c_sz, c_vp]),
    "ptv3_segment_csr_bwd": (c_int, [c_vp, c_vp, c_vp, c_int, c_vp, c_vp, c_i64, c_i64, c_int, c_vp, c_vp]),
    "ptv3_layer_norm_bwd": (c_int, [c_vp, c_int, c_vp, c_int, c_vp, c_f32, c_i64, c_int, c_vp, c_vp, c_vp, c_vp]),
    "ptv3_sparse_pair_maps": (c_int, [c_vp, c_vp, c_i64, c_i64, c_vp, c_vp, c_vp]),
    "ptv3_column_sum": (c_int, [c_vp, c_int, c_i64, c_int, c_vp, c_vp]),
    "ptv3_keypoint_aggregate": (c_int, [c_vp, c_vp, c_vp, c_i64, c_int, c_int, c_f32, c_vp, c_vp]),
    "ptv3_keypoint_eval_workspace": (c_sz, [c_i64, c_i64, c_int]),
    "ptv3_keypoint_eval": (c_int, [c_vp, c_vp, c_vp, c_vp, c_vp, c_i64, c_i64, c_int, c_vp, c_vp, c_vp, c_vp, c_sz, c_vp]),
}

_lib = None


def load():
    """Load (once) and return the ctypes handle.  Raises RuntimeError if the library has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc -gencode arch=compute_100a,code=sm_100a).  ptv3_b200 has no CPU or PyTorch fallback.")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError here = header / library mismatch
        fn.restype, fn.argtypes = res, args
    _lib = lib
    return lib


def check(rc, what):
    if rc != 0:
        msg = load().ptv3_error_string(int(rc)).decode()
        raise RuntimeError(f"{what} failed: {msg} (code {rc})")


def launch_count():
    return int(load().ptv3_launch_count())
