"""CPU: properties of the oracle's EXTENSION section (quad matching, sub-pixel refinement).
These stages are absent from the reference (parity unpinned), so what can be checked on the CPU is internal
consistency: the quad matcher against the pinned flow matcher it is built from, and the fit against its definition."""
import numpy as np

from hls_final_visual_odometry_b200.synth import make_stereo_sequence
from oracle.oracle_lib import Oracle

W, H = 500, 150


def _setup(seed=3, disparity=12, n=2):
    left, right = make_stereo_sequence(W, H, n, seed=seed, disparity=disparity)
    orc = Oracle(W, H)
    fl = [orc.compute_features(f)[:2] for f in left]
    fr = [orc.compute_features(f)[:2] for f in right]
    return left, right, orc, fl, fr


def test_quad_with_identical_cameras_is_a_subset_of_flow():
    left, _, orc, fl, _ = _setup()
    quad = orc.quad_matching(fl[0], fl[0], fl[1], fl[1], 2)
    flow, _ = orc.matching(*fl[0], *fl[1])
    key = lambda m: set(zip(m["u1p"].tolist(), m["v1p"].tolist(), m["u1c"].tolist(), m["v1c"].tolist()))
    assert len(quad) > 0.9 * len(flow) and key(quad) <= key(flow)
    assert (quad["u2p"] == quad["u1p"]).all() and (quad["v2c"] == quad["v1c"]).all()


def test_quad_recovers_rig_geometry_and_rejects_negative_disparity():
    left, right, orc, fl, fr = _setup(disparity=[9, 17])
    m = orc.quad_matching(fl[0], fr[0], fl[1], fr[1], 2)
    assert len(m) > 1000
    ok = (m["u1p"] - m["u2p"] == 9) & (m["u1c"] - m["u2c"] == 17) & (m["v1p"] == m["v2p"]) & (m["u1c"] - m["u1p"] == -3)
    assert ok.mean() > 0.99
    assert len(orc.quad_matching(fr[0], fl[0], fr[1], fl[1], 2)) == 0      # cameras swapped
    assert (m["i1p"] == -1).all() and (m["i2c"] == -1).all()


def test_disp_tolerance_is_the_row_window():
    left, right, orc, fl, fr = _setup()
    frs = [orc.compute_features(np.roll(f, 3, axis=0))[:2] for f in right]   # right camera 3 rows low
    assert len(orc.quad_matching(fl[0], frs[0], fl[1], frs[1], 2)) < 0.3 * len(orc.quad_matching(fl[0], frs[0], fl[1], frs[1], 3))


def test_refine_keeps_order_moves_less_than_a_pixel_and_is_idempotent_on_rejects():
    left, right, orc, fl, fr = _setup()
    m = orc.quad_matching(fl[0], fr[0], fl[1], fr[1], 2)
    r = orc.refine(m, left[0], left[1], right[0], right[1])
    assert 0.5 * len(m) < len(r) <= len(m)
    # anchors are untouched and still in the original order
    pos = {(a, b): i for i, (a, b) in enumerate(zip(m["u1c"].tolist(), m["v1c"].tolist()))}
    idx = [pos[(a, b)] for a, b in zip(r["u1c"].tolist(), r["v1c"].tolist())]
    assert idx == sorted(idx)
    for f in ("u1p", "v1p", "u2p", "v2p", "u2c", "v2c"):
        assert np.abs(r[f] - m[f][idx]).max() < 2.0                           # integer shift <= 1 plus |dd| < 1
    # flow lists (u2c = -1) only touch 1p
    flow, _ = orc.matching(*fl[0], *fl[1])
    rf = orc.refine(flow, left[0], left[1])
    assert (rf["u2c"] == -1).all() and (rf["u1p"] != np.rint(rf["u1p"])).any()
    assert len(orc.refine(flow[:0], left[0], left[1])) == 0
