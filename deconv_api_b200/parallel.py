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


def gather_outputs(local, dist, dst=0, device=None):
    """Per-request gather of each rank's results (numpy array, first axis = units, any length per rank) to `dst`:
    two tensor collectives (the lengths, then the zero-padded payload) — no pickling, so it works for gigabytes and
    for NCCL (`device` = the rank's CUDA device) as well as gloo.  Returns the concatenation on `dst`, None elsewhere."""
    import torch
    local = np.ascontiguousarray(local)
    world, rank = dist.get_world_size(), dist.get_rank()
    tail = tuple(local.shape[1:])
    row = int(np.prod(tail)) if tail else 1
    nbytes = local.dtype.itemsize * row
    n = torch.tensor([local.shape[0]], dtype=torch.int64)
    if device is not None:
        n = n.to(device)
    counts = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(counts, n)
    counts = [int(c.item()) for c in counts]
    cap = max(max(counts), 1)
    pad = torch.zeros((cap * nbytes,), dtype=torch.uint8)
    pad[:local.shape[0] * nbytes] = torch.from_numpy(local.reshape(-1).view(np.uint8))
    if device is not None:
        pad = pad.to(device)
    bufs = [torch.empty_like(pad) for _ in range(world)] if rank == dst else None
    dist.gather(pad, bufs, dst=dst)
    if rank != dst:
        return None
    parts = [b.cpu().numpy()[:c * nbytes].view(local.dtype).reshape((c,) + tail) for b, c in zip(bufs, counts) if c]
    return np.concatenate(parts, axis=0) if parts else np.zeros((0,) + tail, local.dtype)


# ------------------------------------------------------------------------------------------------------------
# Sharded drivers (BASELINE.json configs[3] and [4]).  The unit of work is one (image, layer, feature map) backward
# pass of visualize_all_layers (reference app/deepdream.py:441-474), resp. one image of a DeepDream batch.
class _NoDist(object):
    """Stand-in for torch.distributed in a single process."""

    @staticmethod
    def get_rank():
        return 0

    @staticmethod
    def get_world_size():
        return 1


def _engine_runner(model, layer_name, precision, out_format, visualize_mode, device):
    """runner(images (B,H,W,3) f32, feats) -> (n_feat, B, H, W, 3): one forward tape, one batched projection."""
    from . import engine as _engine
    eng = model.bound_engine(device, upto=layer_name)
    li = eng.index[layer_name]
    dtype = _engine.OUT_FORMATS[out_format][1]

    def run(images, feats, out=None, pinned=False):
        eng.forward(images, li)
        if out is None and pinned:          # a buffer the caller will reuse: page-locked, so the copies run at PCIe rate
            shape = (len(feats),) + tuple(images.shape)
            cache = eng.__dict__.setdefault("_sink_buffers", {})      # kept across calls: cudaHostAlloc of 600 MB takes ~0.3 s
            out = cache.get((shape, dtype))
            if out is None:
                cache.clear()
                out = cache[(shape, dtype)] = eng.pinned_empty(shape, dtype)
        return eng.project(li, feats, mode=visualize_mode, precision=precision, out=out, out_format=out_format)
    run.supports_pinned = True
    return run


def visualize_batch(model, images, layer_name, features=None, dist=None, precision="bf16", out_format="u8",
                    visualize_mode="all", images_per_tape=8, gather=False, device=None, runner=None, sink=None,
                    torch_device=None):
    """Project feature maps `features` (default: every map) of `layer_name` for every image of `images`
    ((N,H,W,3) float32, preprocessed; every rank passes the same array), sharded over the ranks of `dist`
    (torch.distributed, or None for one process) by `shard_work`: by image when N >= world, otherwise every rank
    replicates the cheap forward and takes a contiguous slice of `features`.  No collective touches the data path.

    Returns a list of blocks ``(image_indices, feature_indices, array)`` with array of shape
    (len(feature_indices), len(image_indices), H, W, 3) in `out_format` ('u8' = every projection de-processed on the
    GPU, 'bf16' bit patterns, or 'f32') — this rank's share, left where it is.  `gather=True` additionally collects
    every rank's blocks on rank 0 (tensor collectives) and returns there ONE array (N, n_features, H, W, 3), None on
    the other ranks.  `sink(image_indices, feature_indices, array)`, when given, consumes each block as soon as it is
    complete and the block list is not kept: the array is ONE page-locked buffer reused for every block (full-rate
    device-to-host copies), valid only inside the call."""
    from . import deepdream as _dd
    m = _dd.adopt(model)
    d = dist if dist is not None else _NoDist
    world, rank = d.get_world_size(), d.get_rank()
    images = np.asarray(images, np.float32)
    assert images.ndim == 4, "images must be (N,H,W,3)"
    N = images.shape[0]
    if features is None:
        features = np.arange(m.get_layer(layer_name).output_shape[-1], dtype=np.int32)
    features = np.ascontiguousarray(features, np.int32)
    units = shard_work(N, int(features.size), world, rank)
    run = runner if runner is not None else _engine_runner(m, layer_name, precision, out_format, visualize_mode, device)
    blocks = []
    if units:
        flo, fhi = units[0][1], units[0][2]
        feats = features[flo:fhi]
        mine = [i for i, _, _ in units]
        reuse = None
        for k in range(0, len(mine), images_per_tape):
            idx = mine[k:k + images_per_tape]
            batch = np.ascontiguousarray(images[idx])
            if sink is not None and reuse is not None and reuse.shape[1] == len(idx):
                arr = run(batch, feats, reuse)
            elif sink is not None and getattr(run, "supports_pinned", False):
                arr = run(batch, feats, None, True)
            else:
                arr = run(batch, feats)
            if sink is not None:
                sink(idx, feats, arr)
                reuse = arr
            else:
                blocks.append((idx, feats, arr))
    if not gather:
        return blocks
    # rows = (image, feature) pairs in this rank's order, plus their coordinates
    rows, coords = [], []
    for idx, feats, arr in blocks:
        a = np.moveaxis(arr, 0, 1)                       # (images, features, H, W, 3)
        rows.append(a.reshape((-1,) + a.shape[2:]))
        coords += [(i, int(f)) for i in idx for f in feats]
    tail = images.shape[1:]
    dtype = {"u8": np.uint8, "bf16": np.uint16, "f32": np.float32}[out_format]
    local = np.concatenate(rows, axis=0) if rows else np.zeros((0,) + tail, dtype)
    cl = np.asarray(coords, np.int64).reshape(-1, 2)
    data = gather_outputs(local, d, 0, torch_device) if world > 1 else local
    coord = gather_outputs(cl, d, 0, torch_device) if world > 1 else cl
    if rank != 0:
        return None
    out = np.zeros((N, int(features.size)) + tail, dtype)
    pos = {int(f): j for j, f in enumerate(features)}
    for r, (i, f) in enumerate(coord):
        out[int(i), pos[int(f)]] = data[r]
    return out


def dream_batch(model, images, dist=None, gather=False, device=None, runner=None, torch_device=None, **dream_kwargs):
    """DeepDream on a batch of images ((N,H,W,3) float32), the batch sharded over the ranks (config 5: one or more
    2048x2048 images per GPU).  A single image is never split across GPUs (it would need halo exchange): replicas
    only.  Returns [(image_index, dreamed image)] for this rank, or with gather=True the (N,H,W,3) array on rank 0."""
    from . import dream as _dm
    d = dist if dist is not None else _NoDist
    world, rank = d.get_world_size(), d.get_rank()
    images = np.asarray(images, np.float32)
    lo, hi = partition(images.shape[0], world, rank)
    run = runner if runner is not None else (lambda img: _dm.dream(model, img, device=device, **dream_kwargs))
    local = [(i, np.asarray(run(images[i]), np.float32)) for i in range(lo, hi)]
    if not gather:
        return local
    arr = np.stack([a for _, a in local]) if local else np.zeros((0,) + images.shape[1:], np.float32)
    data = gather_outputs(arr, d, 0, torch_device) if world > 1 else arr      # partition() is contiguous and ordered by rank
    return data if rank == 0 else None
