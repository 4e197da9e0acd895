"""DeepDream on the B200 path.  The reference has no DeepDream code despite its file name
(SURVEY.md §0.2); this is the derived spec of SURVEY.md Appendix A, every constant ours:

    dream(model, image, layers={'block3_conv3': 0.5, 'block4_conv3': 2.0, 'block5_conv3': 1.5},
          num_octave=3, octave_scale=1.4, iterations=20, step=1.0, jitter=32, seed=0) -> image

`image` is (H,W,3) or (B,H,W,3) float32 in the model's input space (caffe-preprocessed pixels, so
`step` is in pixel units); the result has the same shape.  The whole octave loop is one C-ABI call.
"""
import numpy as np

from . import deepdream as _dd

DEFAULT_LAYERS = {"block3_conv3": 0.5, "block4_conv3": 2.0, "block5_conv3": 1.5}


def octave_shapes(h, w, num_octave=3, octave_scale=1.4):
    """[(h_k, w_k)] smallest first: int(h / scale**k) for k = num_octave-1 .. 0."""
    shapes = [(int(h / (octave_scale ** k)), int(w / (octave_scale ** k))) for k in range(num_octave)]
    return shapes[::-1]


def jitter_offsets(n, jitter=32, seed=0):
    """n (oy, ox) pairs from numpy's default_rng(seed): integers in [-jitter, jitter]."""
    rng = np.random.default_rng(seed)
    return rng.integers(-jitter, jitter + 1, (n, 2)).astype(np.int32)


def dream(model, image, layers=None, num_octave=3, octave_scale=1.4, iterations=20, step=1.0, jitter=32, seed=0,
          device=None, return_losses=False, precision="fp32"):
    layers = dict(DEFAULT_LAYERS if layers is None else layers)
    m = _dd.adopt(model)
    names = [l.name for l in m.layers]
    for n in layers:
        if n not in names:
            raise KeyError(n)
    deepest = max(layers, key=lambda n: names.index(n))
    eng = m.bound_engine(device, upto=deepest)
    img = np.asarray(image, np.float32)
    squeeze = img.ndim == 3
    if squeeze:
        img = img[None]
    B, H, W, _ = img.shape
    shapes = octave_shapes(H, W, num_octave, octave_scale)
    jt = jitter_offsets(num_octave * iterations, jitter, seed)
    idx = [eng.index[n] for n in layers]
    coeffs = [float(layers[n]) for n in layers]
    res = eng.dream(img, idx, coeffs, shapes, iterations, step, jt, want_losses=return_losses, precision=precision)
    out, losses = res if return_losses else (res, None)
    out = out[0] if squeeze else out
    return (out, losses) if return_losses else out
