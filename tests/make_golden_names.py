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


def eval_batches():
    """Two synthetic validation batches of the offset head: the inputs of tests/golden/eval.npz."""
    import numpy as np
    rng = np.random.default_rng(91)
    K = 6
    batches = []
    for sizes, with_scale in (([1500, 700, 2300], True), ([900, 1200], False)):
        n = sum(sizes)
        coord = rng.normal(0.0, 0.4, size=(n, 3)).astype(np.float32)
        pred = rng.normal(0.0, 0.2, size=(n, K, 4)).astype(np.float32)
        target = rng.normal(0.0, 0.1, size=(n, K, 4)).astype(np.float32)
        target[..., 3] = (rng.random((n, K)) < 0.05).astype(np.float32)
        d = dict(coord=coord, pred=pred, target=target, offset=np.cumsum(sizes).astype(np.int64))
        if with_scale:
            d["scale"] = np.array([873.69, 912.5, 640.25], dtype=np.float32)
            target[:1500, 2, 3] = 0.0                     # scene 0: keypoint 2 has no masked row -> skipped
            target[1500:2200, :, 3] = 0.0                 # scene 1: no valid keypoint -> not counted as a sample
            pred[2300, 4, 3] = pred[2900, 4, 3] = 7.5     # scene 2: tied maximum, the first row wins
        batches.append(d)
    return batches
