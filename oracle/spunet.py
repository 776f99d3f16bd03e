"""Oracle: SpUNet-v1m1 (BASELINE configs[3]) -- functional torch-CPU fp32 restatement.

Test infrastructure only (see oracle/__init__.py).  Follows
  pointcept/models/sparse_unet/spconv_unet_v1m1_base.py:23-86  BasicBlock
  :89-281                                                        SpUNetBase (conv_input, down / enc / up / dec, final)
over a flat weight dict with the reference's state_dict names.  The strided / inverse convolutions live in the
un-vendored spconv 2.3.6 (uv_requirements.txt:152); their published semantics are restated here ("parity unpinned"
against real spconv, like SubMConv3d):

  SparseConv3d(k=2, s=2, padding=0)   out site o = floor(i / 2) for every active input site i of the same batch
                                      element; out[o] = sum_{i in cell o} W[:, i - 2 o, :] in[i]; output sites are the
                                      distinct o (order: ascending (b, x, y, z) -- spconv's order is unspecified and
                                      does not matter, the decoder returns to the input sites)
  SparseInverseConv3d(k=2, indice_key) re-uses the pairs of the forward conv with the same kernel offset:
                                      up[i] = W[:, i - 2 o(i), :] coarse[o(i)]
"""
import torch
import torch.nn.functional as F

from . import ops as O


def down_map(indices):
    """indices int (N,4) [b,x,y,z] -> (out_indices (M,4) int32, cluster (N,) int64 = o(i), koff (N,) int64 in 0..7 with
    koff = ((x&1)*2 + (y&1))*2 + (z&1), matching weight[:, kx, ky, kz, :])."""
    idx = indices.to(torch.int64)
    coarse = torch.cat([idx[:, :1], idx[:, 1:] >> 1], 1)
    key = O._pack_key(coarse[:, 0], coarse[:, 1:])
    uk, cluster = torch.unique(key, sorted=True, return_inverse=True)
    first = torch.full((uk.numel(),), idx.shape[0], dtype=torch.int64).scatter_reduce(
        0, cluster, torch.arange(idx.shape[0]), "amin")
    koff = ((idx[:, 1] & 1) * 2 + (idx[:, 2] & 1)) * 2 + (idx[:, 3] & 1)
    return coarse[first].to(torch.int32), cluster, koff


def sparse_conv_k2s2(feat, cluster, koff, n_out, weight):
    """weight (Cout, 2,2,2, Cin).  out[o] = sum_i W[:, koff(i), :] feat[i]."""
    cout, cin = weight.shape[0], weight.shape[-1]
    w = weight.reshape(cout, 8, cin)
    out = torch.zeros(n_out, cout, dtype=feat.dtype)
    for k in range(8):
        rows = torch.nonzero(koff == k).squeeze(1)
        if rows.numel():
            out.index_add_(0, cluster[rows], feat[rows] @ w[:, k, :].t())
    return out


def sparse_inverse_conv_k2(coarse, cluster, koff, weight):
    """weight (Cout, 2,2,2, Cin).  up[i] = W[:, koff(i), :] coarse[cluster(i)]."""
    cout, cin = weight.shape[0], weight.shape[-1]
    w = weight.reshape(cout, 8, cin)
    out = torch.zeros(cluster.shape[0], cout, dtype=coarse.dtype)
    for k in range(8):
        rows = torch.nonzero(koff == k).squeeze(1)
        if rows.numel():
            out[rows] = coarse[cluster[rows]] @ w[:, k, :].t()
    return out


def _bn(x, W, p, eps=1e-3):
    return (x - W[p + ".running_mean"]) / torch.sqrt(W[p + ".running_var"] + eps) * W[p + ".weight"] + W[p + ".bias"]


def _block(x, nbr, W, p):
    """BasicBlock (:23-86)."""
    if (p + ".proj.0.weight") in W:
        res = _bn(x @ W[p + ".proj.0.weight"].reshape(W[p + ".proj.0.weight"].shape[0], -1).t(), W, p + ".proj.1")
    else:
        res = x
    y = F.relu(_bn(O.subm_conv(x, nbr, W[p + ".conv1.weight"]), W, p + ".bn1"))
    y = _bn(O.subm_conv(y, nbr, W[p + ".conv2.weight"]), W, p + ".bn2")
    return F.relu(y + res)


@torch.no_grad()
def spunet_forward(W, grid_coord, feat, offset, channels=(32, 64, 128, 256, 256, 128, 96, 96),
                   layers=(2, 3, 4, 6, 2, 2, 2, 2)):
    """SpUNetBase.forward (:240-281), eval mode, decoder on.  Returns logits (N, num_classes)."""
    offset = torch.as_tensor(offset, dtype=torch.int64)
    counts = torch.diff(offset, prepend=torch.zeros(1, dtype=torch.int64))
    batch = torch.repeat_interleave(torch.arange(len(offset)), counts)
    idx = torch.cat([batch[:, None], torch.as_tensor(grid_coord).to(torch.int64)], 1).to(torch.int32)
    ns = len(layers) // 2
    x = torch.as_tensor(feat, dtype=torch.float32)
    x = F.relu(_bn(O.subm_conv(x, O.subm_neighbor_table(idx, 5), W["conv_input.0.weight"]), W, "conv_input.1"))
    levels = [dict(idx=idx, nbr=O.subm_neighbor_table(idx, 3))]
    skips = [x]
    for s in range(ns):
        out_idx, cluster, koff = down_map(levels[-1]["idx"])
        x = F.relu(_bn(sparse_conv_k2s2(x, cluster, koff, out_idx.shape[0], W[f"down.{s}.0.weight"]), W, f"down.{s}.1"))
        levels.append(dict(idx=out_idx, nbr=O.subm_neighbor_table(out_idx, 3), cluster=cluster, koff=koff))
        for i in range(layers[s]):
            x = _block(x, levels[-1]["nbr"], W, f"enc.{s}.block{i}")
        skips.append(x)
    x = skips.pop(-1)
    for s in reversed(range(ns)):
        lv = levels[s + 1]
        x = F.relu(_bn(sparse_inverse_conv_k2(x, lv["cluster"], lv["koff"], W[f"up.{s}.0.weight"]), W, f"up.{s}.1"))
        x = torch.cat([x, skips.pop(-1)], 1)
        for i in range(layers[len(channels) - s - 1]):
            x = _block(x, levels[s]["nbr"], W, f"dec.{s}.block{i}")
    wf = W["final.weight"]
    return x @ wf.reshape(wf.shape[0], -1).t() + W["final.bias"]
