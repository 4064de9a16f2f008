"""Row F2: host egomotion, vo_estimate_motion_mono, against the compiled reference's
VisualOdometryMono::estimateMotion (/root/reference/src/viso_mono.cpp:41-160).

Floating point, so the bar is a tolerance, stated here: |delta| <= 1e-6 on each of the six motion parameters
(BASELINE.md section 3), with the same success flag and the same RANSAC inlier count.  The committed golden
(tests/golden/egomotion.json, made by tests/golden/make_golden_egomotion.py from oracle/_ref) pins it on any
box; where oracle/_ref is present the live reference is compared as well.  No GPU involved: CPU suite."""
import ctypes
import json
import os

import numpy as np
import pytest

from hls_final_visual_odometry_b200 import Egomotion, P_MATCH_DTYPE, transformation_vector_to_matrix
from hls_final_visual_odometry_b200.synth import KITTI_CALIB, KITTI_HEIGHT, KITTI_PITCH, make_scene_matches

TOL = 1e-6
GOLD = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "egomotion.json")))


def _lists(case):
    kw = {k: v for k, v in case.items() if k not in ("seed", "n_lists")}
    return make_scene_matches(case["n_lists"], seed=case["seed"], **kw)[0]


def _new():
    return Egomotion(height=KITTI_HEIGHT, pitch=KITTI_PITCH, **KITTI_CALIB)


def test_sample_stream_is_glibc_rand():
    """getRandomSample draws from rand() after srand(0) (src/viso.cpp:34,98); ours is a private copy of it."""
    libc = ctypes.CDLL("libc.so.6")
    for seed in (0, 1, 12345):
        libc.srand(seed)
        eg = Egomotion(seed=seed)
        assert [eg.rand() for _ in range(2000)] == [libc.rand() for _ in range(2000)]


@pytest.mark.parametrize("gi", range(len(GOLD["cases"])))
def test_motion_matches_reference_golden(gi):
    g = GOLD["cases"][gi]
    lists = _lists(g["case"])
    assert [len(m) for m in lists] == g["n_matches"], "synthetic scene generator changed"
    eg = _new()   # one handle per sequence: the sample stream carries over from list to list like the reference's
    for i, m in enumerate(lists):
        tr = eg.estimate(m)
        assert (tr is not None) == bool(g["ok"][i])
        assert len(eg.inliers()) == g["n_inliers"][i]
        if tr is not None:
            want = np.array([float.fromhex(x) for x in g["tr"][i]])
            assert np.abs(tr - want).max() <= TOL


def test_motion_matches_live_reference():
    from oracle.ref_lib import RefLib, have_ref, mono_motion_seq
    if not have_ref("native"):
        pytest.skip("oracle/_ref not built")
    lists, truth = make_scene_matches(4, seed=77)
    tr_ref, ok_ref, ninl_ref = mono_motion_seq(RefLib("native"), lists, height=KITTI_HEIGHT, pitch=KITTI_PITCH, **KITTI_CALIB)
    eg = _new()
    for i, m in enumerate(lists):
        tr = eg.estimate(m)
        assert (tr is not None) == bool(ok_ref[i]) and len(eg.inliers()) == ninl_ref[i]
        assert np.abs(tr - tr_ref[i]).max() <= TOL
        # and the estimate is the scene's motion: rotation to 5 mrad (pixel-rounded matches), translation direction to 5 degrees
        assert np.abs(tr[:3] - truth[i, :3]).max() < 5e-3
        cosang = np.dot(tr[3:], truth[i, 3:]) / (np.linalg.norm(tr[3:]) * np.linalg.norm(truth[i, 3:]))
        assert cosang > np.cos(np.radians(5.0))


def test_result_independent_of_thread_count():
    lists, _ = make_scene_matches(2, seed=5)
    outs = []
    for threads in (1, 3, 8):
        eg = Egomotion(height=KITTI_HEIGHT, pitch=KITTI_PITCH, threads=threads, **KITTI_CALIB)
        outs.append([eg.estimate(m) for m in lists] + [eg.inliers()])
    for o in outs[1:]:
        for a, b in zip(outs[0], o):
            assert np.array_equal(a, b)


def test_failure_paths_return_none():
    eg = _new()
    assert eg.estimate(np.zeros(0, dtype=P_MATCH_DTYPE)) is None            # N < 10: src/viso_mono.cpp:45-46
    lists, _ = make_scene_matches(1, seed=3)
    assert eg.estimate(lists[0][:9]) is None
    same = lists[0].copy()                                                    # all points identical: normalisation fails (:213-214)
    for f in ("u1p", "v1p", "u1c", "v1c"):
        same[f] = 100.0
    assert eg.estimate(same) is None


def test_transformation_vector_to_matrix():
    rx, ry, rz, t = 0.01, -0.02, 0.03, (0.1, -0.2, 0.9)
    M = transformation_vector_to_matrix([rx, ry, rz, *t])
    cx, sx, cy, sy, cz, sz = np.cos(rx), np.sin(rx), np.cos(ry), np.sin(ry), np.cos(rz), np.sin(rz)
    Rx = np.array([[1, 0, 0], [0, cx, -sx], [0, sx, cx]])
    Ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
    Rz = np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]])
    assert np.allclose(M[:3, :3], Rx @ Ry @ Rz, atol=1e-15)   # src/viso.cpp:76-79 is Rx*Ry*Rz written out
    assert np.allclose(M[:3, 3], t) and np.allclose(M[3], [0, 0, 0, 1])


def test_sample_sets_equal_the_references_erase_loop_and_batch_equals_sequential():
    """getRandomSample (src/viso.cpp:86-106) erases each drawn element from a vector of all N indices; ours ranks the draw among
    the indices not taken yet.  Checked against a literal restatement of the reference loop on libc's rand(), and the batched
    drawing (vo_egomotion_draw_samples_batch) against list-by-list drawing, including lists that never reach RANSAC."""
    lists = _lists(GOLD["cases"][0]["case"])[:3]
    lists = [lists[0], lists[1][:9], lists[2], lists[0][:0], lists[1]]      # 9 matches / empty: no random numbers consumed
    a, b = Egomotion(ransac_iters=300), Egomotion(ransac_iters=300)
    seq = [a.draw_samples(m) for m in lists]
    bat = b.draw_samples_batch(lists, threads=3)
    libc = ctypes.CDLL("libc.so.6")
    libc.srand(0)
    for (s1, n1), (s2, n2), m in zip(seq, bat, lists):
        assert n1 == n2 == (300 if len(m) >= 10 else 0)
        assert np.array_equal(s1[:n1], s2[:n2])
        for k in range(n1):
            pool = list(range(len(m)))
            want = []
            for _ in range(8):
                j = libc.rand() % len(pool)
                want.append(pool.pop(j))
            assert list(s1[k]) == want
    assert a.rand() == b.rand() == libc.rand()                                # all three streams stand at the same place
