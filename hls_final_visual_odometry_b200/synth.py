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


def make_stereo_sequence(width: int, height: int, n_frames: int, seed: int = 0, disparity=12, blur_passes: int = 10,
                         contrast: int = 45):
    """(left, right): two (n_frames, height, width) uint8 stacks of a rectified stereo rig looking at the translating
    texture: right frame f is the same canvas cropped `disparity` pixels further right, so a feature at u in the left
    image sits at u - disparity in the right one (disparity may be a per-frame sequence).  BASELINE.json config 4."""
    canvas = make_canvas(width, height, seed, blur_passes, contrast)
    d = [int(disparity)] * n_frames if np.isscalar(disparity) else [int(x) for x in disparity]
    left = np.stack([frame_from_canvas(canvas, width, height, f) for f in range(n_frames)])
    right = np.stack([frame_from_canvas(canvas, width, height, f, x0=16 + d[f]) for f in range(n_frames)])
    return left, right


# ------------------------------------------------------------------------------------------------
# Match lists from a 3-D scene: the blurred-noise sequence above is a pure
# image translation (a planar, degenerate configuration for the 8-point algorithm: the reference's
# mono egomotion returns false on it), so the egomotion stage is exercised with match lists projected
# from a static, non-planar point cloud seen by a forward-moving KITTI-like camera.
# ------------------------------------------------------------------------------------------------
KITTI_CALIB = dict(f=718.856, cu=607.1928, cv=185.2157)   # KITTI odometry sequence 00 colour-left intrinsics
KITTI_HEIGHT, KITTI_PITCH = 1.6, -0.08                      # the reference demo's values (src/demo.cpp:63-64)


def _rot(rx, ry, rz):
    sx, cx, sy, cy, sz, cz = math.sin(rx), math.cos(rx), math.sin(ry), math.cos(ry), math.sin(rz), math.cos(rz)
    return np.array([[cy * cz, -cy * sz, sy],
                     [sx * sy * cz + cx * sz, -sx * sy * sz + cx * cz, -sx * cy],
                     [-cx * sy * cz + sx * sz, cx * sy * sz + sx * cz, cx * cy]])


def make_scene_matches(n_lists: int, seed: int = 0, width: int = 1241, height: int = 376, n_points: int = 320,
                       step: float = 0.9, outlier_frac: float = 0.05, calib=None, cam_height: float = KITTI_HEIGHT,
                       base: float = 0.0):
    """n_lists match lists (structured arrays with fields u1p,v1p,u1c,v1c,... like Matcher::p_match) of a camera
    driving forward `step` metres per frame with a small yaw/pitch wobble through a static scene: a ground
    plane cam_height below the camera plus random structure on both sides.  Coordinates are rounded to integer
    pixels like the matcher's output; outlier_frac of the matches get a random wrong partner.
    base > 0 adds the right camera of a rectified rig (u2 = u1 - f*base/Z, v2 = v1): quad match lists for the stereo
    egomotion.  Returns (lists, true_motions[n_lists, 6]) with motion = (rx,ry,rz,tx,ty,tz) of Tr_delta (prev -> curr)."""
    from .api import P_MATCH_DTYPE
    calib = calib or KITTI_CALIB
    f, cu, cv = calib["f"], calib["cu"], calib["cv"]
    rng = np.random.default_rng(seed)
    lists, motions = [], []
    for _ in range(n_lists):
        # scene in the PREVIOUS camera frame (x right, y down, z forward)
        n_ground = n_points // 2
        gz = rng.uniform(5.0, 40.0, n_ground)
        gx = rng.uniform(-8.0, 8.0, n_ground)
        ground = np.stack([gx, np.full(n_ground, cam_height), gz], 1)
        n_struct = 3 * n_points
        sz_ = rng.uniform(6.0, 60.0, n_struct)
        side = rng.choice([-1.0, 1.0], n_struct)
        sx_ = side * rng.uniform(4.0, 15.0, n_struct)
        sy_ = rng.uniform(-6.0, cam_height, n_struct)
        pts = np.concatenate([ground, np.stack([sx_, sy_, sz_], 1)])
        rx, ry, rz = rng.normal(0, 0.002), rng.normal(0, 0.01), rng.normal(0, 0.002)
        R = _rot(rx, ry, rz)
        t = np.array([rng.normal(0, 0.02), rng.normal(0, 0.01), -step * rng.uniform(0.8, 1.2)])  # p_curr = R p_prev + t
        pc = pts @ R.T + t
        up, vp = f * pts[:, 0] / pts[:, 2] + cu, f * pts[:, 1] / pts[:, 2] + cv
        uc, vc = f * pc[:, 0] / pc[:, 2] + cu, f * pc[:, 1] / pc[:, 2] + cv
        ok = (pc[:, 2] > 1.0) & (up >= 9) & (up <= width - 9) & (vp >= 9) & (vp <= height - 9) & \
             (uc >= 9) & (uc <= width - 9) & (vc >= 9) & (vc <= height - 9)
        idx = np.flatnonzero(ok)[:n_points]
        m = np.zeros(len(idx), dtype=P_MATCH_DTYPE)
        m["u1p"], m["v1p"] = np.rint(up[idx]), np.rint(vp[idx])
        m["u1c"], m["v1c"] = np.rint(uc[idx]), np.rint(vc[idx])
        n_out = int(outlier_frac * len(idx))
        if n_out:
            bad = rng.choice(len(idx), n_out, replace=False)
            m["u1c"][bad] = np.rint(rng.uniform(9, width - 9, n_out))
            m["v1c"][bad] = np.rint(rng.uniform(9, height - 9, n_out))
        for fld in ("u2p", "v2p", "u2c", "v2c"):
            m[fld] = -1.0
        if base > 0:
            m["u2p"], m["v2p"] = np.rint(up[idx] - f * base / pts[idx, 2]), m["v1p"]
            m["u2c"], m["v2c"] = np.rint(uc[idx] - f * base / pc[idx, 2]), m["v1c"]
        for fld in ("i1p", "i2p", "i1c", "i2c"):
            m[fld] = -1
        lists.append(m)
        motions.append([rx, ry, rz, t[0], t[1], t[2]])
    return lists, np.array(motions)


# ------------------------------------------------------------------------------------------------
# Rendered 3-D sequence (BASELINE.json config 3, "full process() loop"): a textured corridor -- ground
# plane, two side walls, an end wall -- seen by a forward-moving camera.  Three orthogonal planes make the
# epipolar geometry well conditioned, so the reference's mono egomotion succeeds on these frames (it fails
# on the pure-translation crops above).  Integer-only texture, float ray casting with fixed evaluation
# order: frames are reproducible across hosts to the last bit in practice, but the committed goldens of
# the hot path do NOT depend on this generator.
# ------------------------------------------------------------------------------------------------
def _noise_texture(size: int, seed: int, blur_passes: int = 10, contrast: int = 45) -> np.ndarray:
    rng = np.random.default_rng(seed)
    x = rng.integers(0, 256, size=(size, size), dtype=np.int64)
    for _ in range(blur_passes):
        rows = np.roll(x, 1, 1) + x + np.roll(x, -1, 1)
        x = np.roll(rows, 1, 0) + rows + np.roll(rows, -1, 0)
    shift = max(0, int(x.max()).bit_length() - 20)
    x = x >> shift
    mean = int(x.sum()) // x.size
    d = x - mean
    std = max(1, math.isqrt(int((d * d).sum()) // x.size))
    return np.clip(128 + (d * contrast) // std, 0, 255).astype(np.uint8)


def make_corridor_sequence(width: int = 1241, height: int = 376, n_frames: int = 8, seed: int = 0, step: float = 0.6,
                           calib=None, cam_height: float = KITTI_HEIGHT, half_width: float = 6.0, yaw_amp: float = 0.01,
                           tex_size: int = 1024, levels: int = 6, texel_px: float = 2.0, cam_offset_x: float = 0.0):
    """-> (frames uint8 (n, height, width), poses (n, 4, 4) camera-to-world).  The camera drives along +z, `step`
    metres per frame, with a slow sinusoidal yaw; y points down (ground at y = +cam_height).  cam_offset_x shifts the
    camera centre along its own x axis (the right camera of a rectified rig: cam_offset_x = baseline)."""
    calib = calib or KITTI_CALIB
    f, cu, cv = calib["f"], calib["cu"], calib["cv"]
    tex = np.stack([_noise_texture(tex_size, seed * 1000 + 10 * lv + s) for lv in range(levels) for s in range(4)]).astype(np.float64)
    uu, vv = np.meshgrid(np.arange(width, dtype=np.float64), np.arange(height, dtype=np.float64))
    rays_c = np.stack([(uu - cu) / f, (vv - cv) / f, np.ones_like(uu)], -1)      # camera frame
    base_texel = texel_px * 10.0 / f                                                 # one texel ~ texel_px pixels at 10 m
    wall_h_top = -8.0                                                                # walls reach 8 m above the camera
    frames, poses = [], []
    for k in range(n_frames):
        yaw = yaw_amp * math.sin(0.35 * k)
        R = _rot(0.0, yaw, 0.0)                                                      # camera-to-world rotation
        c = np.array([0.3 * math.sin(0.2 * k), 0.0, step * k]) + R[:, 0] * cam_offset_x  # camera centre in the world
        d = rays_c @ R.T
        big = 1e9
        with np.errstate(divide="ignore", invalid="ignore"):
            tg = np.where(d[..., 1] > 1e-9, (cam_height - c[1]) / d[..., 1], big)                   # ground y = cam_height
            tl = np.where(d[..., 0] < -1e-9, (-half_width - c[0]) / d[..., 0], big)                 # left wall
            tr = np.where(d[..., 0] > 1e-9, (half_width - c[0]) / d[..., 0], big)                   # right wall
            zend = step * n_frames + 80.0
            te = np.where(d[..., 2] > 1e-9, (zend - c[2]) / d[..., 2], big)                         # end wall
        t = np.minimum(np.minimum(tg, tl), np.minimum(tr, te))
        which = np.where(t == tg, 0, np.where(t == tl, 1, np.where(t == tr, 2, 3)))
        P = c + d * t[..., None]
        depth = t * (d @ R[:, 2])                                                    # distance along the optical axis
        lv = np.clip(np.floor(np.log2(np.maximum(depth, 1e-3) / 10.0) + 0.5).astype(np.int64), 0, levels - 1)
        texel = base_texel * (2.0 ** lv)
        a = np.where(which == 0, P[..., 0], np.where(which == 3, P[..., 0], P[..., 1]))              # texture axes per surface
        b = np.where(which == 3, P[..., 1], P[..., 2])
        fa, fb = a / texel, b / texel                                                # bilinear texture lookup
        a0, b0 = np.floor(fa), np.floor(fb)
        wa, wb = fa - a0, fb - b0
        ia, ib = a0.astype(np.int64) % tex_size, b0.astype(np.int64) % tex_size
        ia1, ib1 = (ia + 1) % tex_size, (ib + 1) % tex_size
        ls = lv * 4 + which
        val = ((1 - wa) * (1 - wb) * tex[ls, ib, ia] + wa * (1 - wb) * tex[ls, ib, ia1] +
               (1 - wa) * wb * tex[ls, ib1, ia] + wa * wb * tex[ls, ib1, ia1])
        img = np.clip(np.floor(val + 0.5), 0, 255).astype(np.uint8)
        sky = (which != 0) & (P[..., 1] < wall_h_top)
        img[sky] = 128
        frames.append(img)
        T = np.eye(4)
        T[:3, :3], T[:3, 3] = R, c
        poses.append(T)
    return np.stack(frames), np.stack(poses)
