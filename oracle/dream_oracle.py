"""TEST INFRASTRUCTURE ONLY — CPU oracle of the DeepDream loop: torch autograd on the spec of SURVEY.md
Appendix A.  There is NO reference implementation of DeepDream (reference app/deepdream.py contains only
deconvnet code), so this oracle is ours and parity against the reference is UNPINNED for this row.
Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import it."""
import numpy as np
import torch
import torch.nn.functional as F


def _convs(model, upto):
    ops = []
    for layer in model.layers:
        n = type(layer).__name__
        kind = getattr(layer, "kind", None) or {"Conv2D": "conv", "MaxPooling2D": "pool", "InputLayer": "input"}.get(n, n)
        if kind == "conv":
            W, b = layer.get_weights()
            ops.append(("conv", layer.name, torch.from_numpy(np.ascontiguousarray(W)).permute(3, 2, 0, 1).contiguous(),
                        torch.from_numpy(np.ascontiguousarray(b))))
        elif kind == "pool":
            ops.append(("pool", layer.name, None, None))
        if layer.name == upto:
            break
    return ops


def _pool_with_idx(t, idx_nhwc):
    """2x2/2 max pool that takes, for every window, the element a given switch index (N,h,w,C; r*2+c) names — the
    same value as F.max_pool2d wherever the index is the window's arg-max, and differentiable the same way."""
    n, c, h, w = t.shape
    ph, pw = h // 2, w // 2
    win = t[:, :, :2 * ph, :2 * pw].reshape(n, c, ph, 2, pw, 2).permute(0, 1, 2, 4, 3, 5).reshape(n, c, ph, pw, 4)
    idx = torch.from_numpy(np.ascontiguousarray(idx_nhwc)).permute(0, 3, 1, 2).long()
    return win.gather(-1, idx[..., None])[..., 0]


def loss_and_grad(ops, x_nhwc, layers, pool_idx=None):
    """loss = sum_l coeff_l * mean(act_l^2); returns (loss, dloss/dx NHWC).  pool_idx (optional): {pool layer name:
    switch indices (N,h,w,C)} to route a pool by (tests that follow an implementation's choice at near-ties)."""
    x = torch.from_numpy(np.ascontiguousarray(x_nhwc, np.float32)).permute(0, 3, 1, 2).clone().requires_grad_(True)
    t, loss = x, 0.0
    for kind, name, W, b in ops:
        if kind == "conv":
            t = F.relu(F.conv2d(t, W, b, padding=1))
        elif pool_idx is not None and name in pool_idx:
            t = _pool_with_idx(t, pool_idx[name])
        else:
            t = F.max_pool2d(t, 2, 2)
        if name in layers:
            loss = loss + layers[name] * (t * t).mean()
    loss.backward()
    return float(loss.detach()), x.grad.permute(0, 2, 3, 1).contiguous().numpy()


def resize(x_nhwc, hw):
    t = torch.from_numpy(np.ascontiguousarray(x_nhwc, np.float32)).permute(0, 3, 1, 2)
    return F.interpolate(t, size=tuple(hw), mode="bilinear", align_corners=False).permute(0, 2, 3, 1).contiguous().numpy()


def dream(model, image, layers, shapes, iterations, step, jitter):
    names = [l.name for l in model.layers]
    deepest = max(layers, key=lambda n: names.index(n))
    ops = _convs(model, deepest)
    orig = np.asarray(image, np.float32)
    img = orig.copy()
    shrunk = resize(orig, shapes[0])
    losses, k = [], 0
    for shape in shapes:
        if tuple(img.shape[1:3]) != tuple(shape):
            img = resize(img, shape)
        for _ in range(iterations):
            oy, ox = int(jitter[k][0]), int(jitter[k][1]); k += 1
            x = np.roll(img, (oy, ox), axis=(1, 2))
            loss, g = loss_and_grad(ops, x, layers)
            losses.append(loss)
            for b in range(g.shape[0]):
                rms = np.sqrt(np.mean(g[b].astype(np.float64) ** 2))
                x[b] = x[b] + np.float32(step / max(rms, 1e-7)) * g[b]
            img = np.roll(x, (-oy, -ox), axis=(1, 2))
        same = resize(orig, shape)
        img = img + (same - resize(shrunk, shape))
        shrunk = same
    return img, np.array(losses)
