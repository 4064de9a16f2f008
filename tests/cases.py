"""Adversarial / edge-case frame pairs for the parity tests.

Every case returns (prev, curr) uint8 images of the requested size, deterministic and integer-only
(see hls_final_visual_odometry_b200/synth.py for why).
"""
from __future__ import annotations

import numpy as np

from hls_final_visual_odometry_b200.synth import frame_from_canvas, make_canvas


def _pair(canvas, w, h, shift=(3, 1)):
    return (frame_from_canvas(canvas, w, h, 0), frame_from_canvas(canvas, w, h, 1, flow=shift))


def textured(w, h, seed=0):
    return _pair(make_canvas(w, h, seed), w, h)


def flat(w, h, seed=0):
    """no gradient at all -> zero features, zero matches"""
    img = np.full((h, w), 128, np.uint8)
    return img, img.copy()


def low_contrast(w, h, seed=0):
    """responses below nms_tau almost everywhere"""
    return _pair(make_canvas(w, h, seed, contrast=4), w, h)


def dense(w, h, seed=0):
    """2 blur passes: ~2.4x the feature density of the standard texture (long lists, big windows)"""
    return _pair(make_canvas(w, h, seed, blur_passes=2), w, h)


def holes(w, h, seed=0):
    """flat rectangles punched into the texture: empty (bin,class) lists above populated ones
    exercise the lagging v_bin of the reference's column scan (src/myMatch.cpp:153-158)"""
    a, b = textured(w, h, seed)
    out = []
    for img in (a, b):
        img = img.copy()
        img[: h // 2, 100:300] = 128          # empty top v-bins in u-bins 1..4
        img[h // 3: 2 * h // 3, 420:600] = 128  # hole in the middle of a column
        img[:, 700:760] = 128                  # an entirely empty u-bin column
        out.append(img)
    return tuple(out)


def dead_zone_only(w, h, seed=0):
    """texture only in the last u-bin column and the last v-bin row: features exist, but the
    reference never matches them (src/myMatch.cpp:141-146, :92-99)"""
    a, b = textured(w, h, seed)
    u_last = ((9 + 3 * ((w - 19) // 3)) // 60) * 60
    v_last = (h // 45) * 45
    out = []
    for img in (a, b):
        z = np.full_like(img, 128)
        z[:, u_last + 6:] = img[:, u_last + 6:]
        z[v_last + 6:, :] = img[v_last + 6:, :]
        out.append(z)
    return tuple(out)


def tiny_patch(w, h, seed=0):
    """one 48x40 textured patch: a handful of features, fewer than 5 matches -> the '<5 -> 0' rule
    (src/myMatch.cpp:198-203)"""
    a, b = textured(w, h, seed)
    out = []
    for img in (a, b):
        z = np.full_like(img, 128)
        z[60:100, 200:248] = img[60:100, 200:248]
        out.append(z)
    return tuple(out)


def periodic(w, h, seed=0):
    """a 60x45 tile repeated over the frame, zero motion: many exactly equal descriptors, so the
    SAD argmin is decided by the first-found-wins order (src/myMatch.cpp:107)"""
    tile = make_canvas(60, 45, seed)[8:53, 16:76]
    reps = (h // 45 + 1, w // 60 + 1)
    img = np.tile(tile, reps)[:h, :w].copy()
    return img, img.copy()


def big_motion(w, h, seed=0):
    """flow of (-150, -100) px: matches sit at the edge of the +-180 x +-135 window"""
    return _pair(make_canvas(w, h, seed), w, h, shift=(150, 100))


def out_of_window(w, h, seed=0):
    """flow of (-200, 0): the true match is outside the search window"""
    return _pair(make_canvas(w, h, seed), w, h, shift=(200, 0))


CASES = {
    "textured": textured, "flat": flat, "low_contrast": low_contrast, "dense": dense, "holes": holes,
    "dead_zone_only": dead_zone_only, "tiny_patch": tiny_patch, "periodic": periodic,
    "big_motion": big_motion, "out_of_window": out_of_window,
}


# ---------------------------------------------------------------------------------------------
# match lists for the removeOutliers (Delaunay support filter) parity tests, SURVEY.md row A17.
# Integer-lattice current-frame points (as the matcher emits), unique per list, with a smooth flow
# field plus a fraction of gross outliers.  Pure integer numpy, deterministic.
# ---------------------------------------------------------------------------------------------
def _match_list(dtype, uc, vc, du, dv):
    m = np.zeros(len(uc), dtype)
    for k in m.dtype.names:
        m[k] = -1
    m["u1c"], m["v1c"] = uc, vc
    m["u1p"], m["v1p"] = uc - du, vc - dv
    return m


def _lattice_points(rng, n, w, h, step=1):
    cells = rng.permutation((w // step) * (h // step))[:n]
    cells.sort()  # the matcher emits in a spatial order; keep lists ordered but irregular
    return (cells % (w // step)) * step + 9, (cells // (w // step)) * step + 9


def outlier_lists(dtype, seed=0):
    """name -> p_match array.  Sizes stay below the reference's POINT_L = 14002."""
    rng = np.random.default_rng(1000 + seed)
    out = {}

    def add(name, uc, vc, frac=0.1, spread=12):
        n = len(uc)
        du = -3 + (uc // 200) + rng.integers(-1, 2, n)      # smooth flow + 1 px jitter
        dv = -1 + (vc // 150) + rng.integers(-1, 2, n)
        bad = rng.permutation(n)[: int(n * frac)]
        du[bad] += rng.integers(-spread, spread + 1, len(bad))
        dv[bad] += rng.integers(-spread, spread + 1, len(bad))
        out[name] = _match_list(dtype, uc, vc, du, dv)

    add("three", *_lattice_points(rng, 3, 100, 80))                       # <= 3: returned unchanged
    add("four", *_lattice_points(rng, 4, 100, 80))
    add("small", *_lattice_points(rng, 60, 100, 80))
    add("mid", *_lattice_points(rng, 2500, 500, 150))
    add("kitti_10k", *_lattice_points(rng, 10000, 1220, 356))
    add("near_cap", *_lattice_points(rng, 13500, 1220, 356), frac=0.2)  # 3*(2n-5) stays below TRIANGLE_L = 83598
    add("coarse_lattice", *_lattice_points(rng, 5000, 1220, 356, step=3))  # 3-px lattice: many co-circular quadruples
    gu, gv = np.meshgrid(np.arange(60) * 3 + 9, np.arange(40) * 3 + 9)     # full regular grid: every quad is co-circular
    add("regular_grid", gu.ravel(), gv.ravel(), frac=0.15)
    cu = np.concatenate([rng.integers(0, 80, 700) + c for c in (50, 500, 1000)])
    cv = np.concatenate([rng.integers(0, 60, 700) + c for c in (30, 200, 120)])
    key = np.unique(cv * 4096 + cu)                                         # unique pixels, raster order
    add("clustered", key % 4096, key // 4096, frac=0.3, spread=30)
    add("all_outliers", *_lattice_points(rng, 800, 400, 300), frac=1.0, spread=40)
    return out
