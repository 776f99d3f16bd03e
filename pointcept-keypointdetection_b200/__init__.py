"""ptv3_b200: B200-native (sm_100a) PTv3 backbone hot path behind the reference's operator surface.

    ops         torch-tensor wrappers over the C ABI (include/ptv3_b200.h, libptv3_b200.so)
    structure   Point / SparseConvTensor (reference: pointcept/models/utils/structure.py)
    transform   GridSample on the device (reference: pointcept/datasets/transform.py:825-963)
    model       PointTransformerV3 + OffsetKeypointPTv3 / KeypointPTv3 / DefaultSegmentorV2
    shims       drop-in `spconv.pytorch`, `torch_scatter`, `flash_attn` names for the unmodified reference
    synth       synthetic scans / deterministic weights
There is no CPU fallback: ops raise if libptv3_b200.so is missing or a tensor is not on a CUDA device.
"""
__version__ = "0.1.0"
