"""Stereo egomotion (the half of SURVEY.md row F3 the reference DOES implement): vo_estimate_motion_stereo against the
compiled reference's VisualOdometryStereo::estimateMotion (/root/reference/src/viso_stereo.cpp:54-157) on quad match
lists of a synthetic rectified rig.

Floating point: |delta| <= 1e-6 on each of the six motion parameters (measured: 0), same success flag, same inlier
count.  tests/golden/egomotion_stereo.json (made by tests/golden/make_golden_egomotion_stereo.py from oracle/_ref) pins
it on any box; where oracle/_ref is present the live reference is compared as well.  No GPU involved: CPU suite."""
import json
import os

import numpy as np
import pytest

from hls_final_visual_odometry_b200 import EgomotionStereo, P_MATCH_DTYPE
from hls_final_visual_odometry_b200.synth import KITTI_CALIB, make_scene_matches

TOL = 1e-6
GOLD = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "egomotion_stereo.json")))
BASE = GOLD["base"]


@pytest.mark.parametrize("gi", range(len(GOLD["cases"])))
def test_stereo_motion_matches_reference_golden(gi):
    g = GOLD["cases"][gi]
    kw = {k: v for k, v in g["case"].items() if k not in ("seed", "n_lists")}
    lists = make_scene_matches(g["case"]["n_lists"], seed=g["case"]["seed"], base=BASE, **kw)[0]
    assert [len(m) for m in lists] == g["n_matches"], "synthetic scene generator changed"
    eg = EgomotionStereo(base=BASE, **KITTI_CALIB)   # one handle per sequence: the sample stream carries over
    for i, m in enumerate(lists):
        tr = eg.estimate(m)
        assert (tr is not None) == bool(g["ok"][i])
        assert len(eg.inliers()) == g["n_inliers"][i]
        if tr is not None:
            assert np.abs(tr - np.array([float.fromhex(x) for x in g["tr"][i]])).max() <= TOL


def test_stereo_motion_matches_live_reference_and_truth():
    from oracle.ref_lib import RefLib, have_ref, stereo_motion_seq
    if not have_ref("native"):
        pytest.skip("oracle/_ref not built")
    lists, truth = make_scene_matches(4, seed=91, base=BASE, outlier_frac=0.1)
    tr_ref, ok_ref, ninl_ref = stereo_motion_seq(RefLib("native"), lists, base=BASE, **KITTI_CALIB)
    eg = EgomotionStereo(base=BASE, **KITTI_CALIB)
    for i, m in enumerate(lists):
        tr = eg.estimate(m)
        assert (tr is not None) == bool(ok_ref[i]) and len(eg.inliers()) == ninl_ref[i]
        assert np.abs(tr - tr_ref[i]).max() <= TOL
        # metric scale comes from the baseline: the full 6-vector is the scene's motion (pixel-rounded matches)
        assert np.abs(tr[:3] - truth[i, :3]).max() < 5e-3 and np.abs(tr[3:] - truth[i, 3:]).max() < 0.05


def test_stereo_failure_paths_and_sample_stream():
    eg = EgomotionStereo(base=BASE, **KITTI_CALIB)
    first = eg.rand()
    eg2 = EgomotionStereo(base=BASE, **KITTI_CALIB)
    assert eg2.estimate(np.zeros(0, dtype=P_MATCH_DTYPE)) is None          # N < 6: src/viso_stereo.cpp:69-70 ...
    lists, _ = make_scene_matches(1, seed=3, base=BASE)
    assert eg2.estimate(lists[0][:5]) is None
    assert eg2.rand() == first                                              # ... before any rand() is consumed
    assert eg2.estimate(lists[0]) is not None
    flow_only = lists[0].copy()                                             # a flow list (u2 = -1) is not a stereo observation
    for f in ("u2p", "v2p", "u2c", "v2c"):
        flow_only[f] = -1.0
    tr = eg2.estimate(flow_only)
    assert tr is None or not np.allclose(tr, eg.estimate(lists[0]), atol=1e-3)
