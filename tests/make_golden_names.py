"""Names shared by tests/golden/make_golden.py::gold_train and the tests that read tests/golden/train.npz."""
TRAIN_GRADS = ["backbone.embedding.stem.conv.weight", "backbone.enc.enc0.block0.cpe.0.weight",
               "backbone.enc.enc1.down.proj.weight", "backbone.enc.enc2.block0.attn.qkv.weight",
               "backbone.dec.dec0.block1.mlp.0.fc2.weight", "head.3.weight"]


def make_target(n, seed, k=6):
    """Synthetic (N, K, 4) offset-keypoint target: small offsets, ~30 % of the points inside each keypoint's mask."""
    import numpy as np
    rng = np.random.default_rng(seed)
    t = rng.normal(0.0, 0.1, size=(n, k, 4)).astype(np.float32)
    t[..., 3] = (rng.random((n, k)) < 0.3).astype(np.float32)
    return t
