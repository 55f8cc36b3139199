"""``tf.image`` helpers the reference's input pipelines call inside ``Dataset.map``:
``tf.image.resize(image, size)`` (core/tests/examples/dogs_classification.ipynb,
call_run_within_script_with_keras_fit.py:60-64).  Host side, numpy in / numpy out."""
from __future__ import annotations

import numpy as np


def resize(images, size, method="bilinear"):
    """Resizes ``[H, W, C]`` or ``[N, H, W, C]`` images to ``size = (height, width)``; float32 out,
    half-pixel-centre sampling like ``tf.image.resize`` (v2)."""
    a = np.asarray(images)
    single = a.ndim == 3
    if single:
        a = a[None]
    if a.ndim != 4:
        raise ValueError("'images' must have either 3 or 4 dimensions.")
    h, w = int(size[0]), int(size[1])
    n, ih, iw, _c = a.shape
    a = a.astype(np.float32)
    if (ih, iw) == (h, w):
        out = a
    elif method == "nearest":
        ys = np.minimum((np.arange(h) + 0.5) * ih / h, ih - 1).astype(np.int64)
        xs = np.minimum((np.arange(w) + 0.5) * iw / w, iw - 1).astype(np.int64)
        out = a[:, ys][:, :, xs]
    elif method == "bilinear":
        fy = np.clip((np.arange(h) + 0.5) * ih / h - 0.5, 0, ih - 1)
        fx = np.clip((np.arange(w) + 0.5) * iw / w - 0.5, 0, iw - 1)
        y0, x0 = np.floor(fy).astype(np.int64), np.floor(fx).astype(np.int64)
        y1, x1 = np.minimum(y0 + 1, ih - 1), np.minimum(x0 + 1, iw - 1)
        wy = (fy - y0)[None, :, None, None].astype(np.float32)
        wx = (fx - x0)[None, None, :, None].astype(np.float32)
        top = a[:, y0][:, :, x0] * (1 - wx) + a[:, y0][:, :, x1] * wx
        bottom = a[:, y1][:, :, x0] * (1 - wx) + a[:, y1][:, :, x1] * wx
        out = top * (1 - wy) + bottom * wy
    else:
        raise ValueError("Unsupported resize method: {}".format(method))
    return out[0] if single else out


def convert_image_dtype(image, dtype="float32"):
    a = np.asarray(image)
    if a.dtype == np.uint8 and np.dtype(dtype).kind == "f":
        return a.astype(dtype) / 255.0
    return a.astype(dtype)


def flip_left_right(image):
    return np.asarray(image)[..., ::-1, :]
