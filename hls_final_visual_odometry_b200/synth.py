"""Synthetic KITTI-shaped grayscale frames.

There is no dataset on the build or GPU box (the reference's own sequence directory is git-ignored
upstream, /root/reference/.gitignore:6), so every parity test, golden vector and benchmark uses
frames from this generator.  It is integer-only on purpose: the committed golden hashes must be
reproduced bit-for-bit on any host, independent of numpy's SIMD dispatch for float reductions.

Recipe: uniform u8 noise canvas -> `blur_passes` passes of a 3x3 box sum (edge replicated) ->
normalise to mean 128 / std `contrast` -> clip to u8.  Frame f of a sequence is the crop at offset
(x0 + 3 f, y0 + f) (wrapping toroidally), i.e. a true optical flow of (-3, -1) px per frame, well
inside the matcher's +-180 x +-135 px search window (/root/reference/src/size.h:14-15).
"""
from __future__ import annotations

import math

import numpy as np

_MARGIN_X = 256
_MARGIN_Y = 64


def make_canvas(width: int, height: int, seed: int = 0, blur_passes: int = 10,
                contrast: int = 45) -> np.ndarray:
    """Blurred-noise texture canvas of shape (height + 64, width + 256), dtype uint8."""
    rng = np.random.default_rng(seed)
    ch, cw = height + _MARGIN_Y, width + _MARGIN_X
    x = rng.integers(0, 256, size=(ch, cw), dtype=np.int64)
    for _ in range(blur_passes):
        p = np.pad(x, 1, mode="edge")
        rows = p[:, :-2] + p[:, 1:-1] + p[:, 2:]
        x = rows[:-2, :] + rows[1:-1, :] + rows[2:, :]
    # keep ~20 significant bits so that the variance sum stays inside int64
    shift = max(0, int(x.max()).bit_length() - 20)
    x = x >> shift
    n = x.size
    mean = int(x.sum()) // n
    d = x - mean
    var = int((d * d).sum()) // n
    std = max(1, math.isqrt(var))
    y = 128 + (d * contrast) // std
    return np.clip(y, 0, 255).astype(np.uint8)


def frame_from_canvas(canvas: np.ndarray, width: int, height: int, f: int,
                      x0: int = 16, y0: int = 8, flow=(3, 1)) -> np.ndarray:
    """Frame f = crop of the canvas at (x0 + flow[0]*f, y0 + flow[1]*f), wrapping toroidally."""
    ch, cw = canvas.shape
    xs = (x0 + flow[0] * f + np.arange(width)) % cw
    ys = (y0 + flow[1] * f + np.arange(height)) % ch
    return np.ascontiguousarray(canvas[np.ix_(ys, xs)])


def make_sequence(width: int, height: int, n_frames: int, seed: int = 0, blur_passes: int = 10,
                  contrast: int = 45) -> np.ndarray:
    """(n_frames, height, width) uint8 frames of one synthetic sequence."""
    canvas = make_canvas(width, height, seed, blur_passes, contrast)
    return np.stack([frame_from_canvas(canvas, width, height, f) for f in range(n_frames)])
