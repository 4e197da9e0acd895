"""Multi-GPU plumbing: one process per GPU, units of work are independent
(image, layer, feature-map) triples, so the only communication is ONE broadcast of the packed
weights at start-up (58.9 MB fp32 for the 13 VGG16 convs) over NCCL/NVLink.  No per-step collective.
"""
import numpy as np


def partition(n_items, world, rank):
    """Contiguous, balanced slice [lo, hi) of `n_items` units for `rank` (first n%world ranks get one more)."""
    base, extra = divmod(int(n_items), int(world))
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard_work(n_images, n_features, world, rank):
    """Static (image, feature) partition: by image when there are at least as many images as ranks
    (each rank runs the forward of its own images once and all their projections locally); otherwise
    every rank replicates the cheap forward of every image and takes a contiguous feature range.
    Returns a list of (image_index, feat_lo, feat_hi)."""
    if n_images >= world:
        lo, hi = partition(n_images, world, rank)
        return [(i, 0, n_features) for i in range(lo, hi)]
    flo, fhi = partition(n_features, world, rank)
    return [(i, flo, fhi) for i in range(n_images)] if fhi > flo else []


def _weighted_layers(model, include_dense=False):
    """The layer set is a function of the model's topology and `include_dense` only — never of which lazy dense
    matrices happen to be materialised on this rank — so every rank packs the same vector."""
    return [l for l in model.layers if l.kind == "conv" or (include_dense and l.kind == "dense")]


def pack_weights(model, include_dense=False):
    """All conv (and, on request, dense) weights of `model` as one flat float32 vector (layer order: W then b).
    Dense layers are materialised when they are included (fc1 alone is 411 MB)."""
    parts = []
    for l in _weighted_layers(model, include_dense):
        for w in l.get_weights():
            parts.append(np.asarray(w, np.float32).reshape(-1))
    return np.concatenate(parts) if parts else np.zeros((0,), np.float32)


def unpack_weights(model, flat, include_dense=False):
    off = 0
    for l in _weighted_layers(model, include_dense):
        new = []
        for w in l.get_weights():
            n = int(np.prod(w.shape))
            new.append(np.asarray(flat[off:off + n], np.float32).reshape(w.shape).copy())
            off += n
        l.set_weights(new)
    assert off == flat.size, "weight vector length mismatch"


def broadcast_model_weights(model, dist, src=0, device=None, include_dense=False):
    """One data collective: rank `src`'s weights replace everybody's.  `dist` is torch.distributed (NCCL on
    GPUs, gloo in the CPU tests); `device` is the tensor device to stage through.  The ranks first agree on the
    vector length (a MAX/MIN all-reduce of one integer): a mismatch would hang or corrupt the broadcast."""
    import torch
    flat = pack_weights(model, include_dense)
    n = torch.tensor([flat.size, -flat.size], dtype=torch.int64)
    if device is not None:
        n = n.to(device)
    dist.all_reduce(n, op=dist.ReduceOp.MAX)
    if int(n[0]) != flat.size or int(-n[1]) != flat.size:
        raise RuntimeError("broadcast_model_weights: ranks disagree on the weight vector length (%d here, %d..%d overall)"
                           % (flat.size, int(-n[1]), int(n[0])))
    t = torch.from_numpy(flat)
    if device is not None:
        t = t.to(device)
    dist.broadcast(t, src=src)
    if dist.get_rank() != src:
        unpack_weights(model, t.cpu().numpy(), include_dense)
    return flat.size * 4


def gather_outputs(local, dist, dst=0):
    """Optional per-request gather of each rank's projections (numpy, first axis = units) to `dst`."""
    objs = [None] * dist.get_world_size() if dist.get_rank() == dst else None
    dist.gather_object(local, objs, dst=dst)
    if dist.get_rank() != dst:
        return None
    return np.concatenate([o for o in objs if o is not None and len(o)], axis=0)
