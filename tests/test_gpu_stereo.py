"""BASELINE.json config 4: quad (left/right x prev/curr) circular matching and sub-pixel
refinement on the GPU, through the C ABI, against oracle/vo_oracle.c's EXTENSION section.

PARITY UNPINNED: the reference names these stages in its API but does not implement them (its matcher is flow-only:
/root/reference/src/matcher.cpp:51-111 ignores I2 and `method`), so the oracle here is our own restatement (the
reference's find_match on every edge + libviso2's published quad rule and parabolic fit).  Integer coordinates:
bit-exact vs that oracle.  Refined coordinates: bit-exact floats as well (integer costs, closed-form fit, one IEEE
division per offset), asserted with tolerance 0."""
import numpy as np
import pytest

from hls_final_visual_odometry_b200 import Engine
from hls_final_visual_odometry_b200.api import VoError, VO_ERR_STATE
from hls_final_visual_odometry_b200.synth import make_sequence, make_stereo_sequence
from oracle.oracle_lib import Oracle, matches_digest

pytestmark = pytest.mark.gpu


def _feats(orc, img):
    return orc.compute_features(img)[:2]


def _quad_case(w, h, seed, disparity, n=3, **kw):
    left, right = make_stereo_sequence(w, h, n, seed=seed, disparity=disparity, **kw)
    orc = Oracle(w, h)
    fl = [_feats(orc, f) for f in left]
    fr = [_feats(orc, f) for f in right]
    return left, right, orc, fl, fr


@pytest.mark.parametrize("w,h,seed,disp", [(500, 150, 3, 12), (500, 150, 4, [0, 7, 30]), (1241, 376, 0, [14, 15, 13])])
def test_quad_matching_bit_exact_vs_oracle(w, h, seed, disp):
    left, right, orc, fl, fr = _quad_case(w, h, seed, disp)
    n = len(left)
    eng = Engine(w, h, max_frames=2 * n, max_pairs=n - 1, stereo=True)
    try:
        eng.batch_extract_host(0, left)
        eng.batch_extract_host(n, right)
        k = np.arange(1, n)
        eng.batch_match_quad(k - 1, n + k - 1, k, n + k, disp_tolerance=2, max_features=2, bw=50.0, bh=50.0)
        raw_n, bk_n = eng.batch_counts(n - 1)
        for i in range(n - 1):
            want = orc.quad_matching(fl[i], fr[i], fl[i + 1], fr[i + 1], 2)
            got = eng.batch_get_matches(i, 0)
            assert raw_n[i] == len(want) > 100
            assert matches_digest(got) == matches_digest(want)
            # every match is a true quad of the synthetic rig: flow (-3,-1), non-negative disparity
            assert ((got["u1c"] - got["u1p"] == -3) & (got["v1c"] - got["v1p"] == -1)).mean() > 0.99
            assert (got["u1p"] >= got["u2p"]).all() and (got["u1c"] >= got["u2c"]).all()
            wb = orc.bucket_features(want, 2, 50.0, 50.0)
            assert bk_n[i] == len(wb) and matches_digest(eng.batch_get_matches(i, 1)) == matches_digest(wb)
    finally:
        eng.close()


def test_quad_exhaustive_equals_pruned_and_tolerance_matters():
    w, h = 500, 150
    left, right, orc, fl, fr = _quad_case(w, h, 8, 9, n=2)
    # a one-pixel vertical misalignment of the right camera: tolerance 0 loses the true quads (a few hundred self-consistent false ones remain), tolerance 2 keeps it
    right_shift = np.roll(right, 1, axis=1)
    frs = [_feats(orc, f) for f in right_shift]
    out = {}
    for exhaustive in (False, True):
        eng = Engine(w, h, max_frames=4, max_pairs=1, stereo=True, match_exhaustive=exhaustive)
        try:
            eng.batch_extract_host(0, left)
            eng.batch_extract_host(2, right_shift)
            for tol in (0, 2):
                eng.batch_match_quad([0], [2], [1], [3], disp_tolerance=tol)
                out[(exhaustive, tol)] = eng.batch_get_matches(0, 0)
        finally:
            eng.close()
    for tol in (0, 2):
        want = orc.quad_matching(fl[0], frs[0], fl[1], frs[1], tol)
        assert matches_digest(out[(False, tol)]) == matches_digest(out[(True, tol)]) == matches_digest(want)
    assert len(out[(False, 0)]) < 0.3 * len(out[(False, 2)])


def test_quad_degenerate_inputs():
    w, h = 500, 150
    left, right, orc, fl, fr = _quad_case(w, h, 5, 10, n=2)
    flat = np.full((2, h, w), 90, np.uint8)
    eng = Engine(w, h, max_frames=4, max_pairs=1, stereo=True)
    try:
        eng.batch_extract_host(0, left)
        eng.batch_extract_host(2, flat)                       # featureless right camera: no cycle can close
        eng.batch_match_quad([0], [2], [1], [3])
        assert eng.batch_counts(1)[0][0] == 0
        eng.batch_extract_host(2, left)                       # right == left: disparity 0 everywhere
        eng.batch_match_quad([0], [2], [1], [3])
        got = eng.batch_get_matches(0, 0)
        want = orc.quad_matching(fl[0], fl[0], fl[1], fl[1], 2)
        assert matches_digest(got) == matches_digest(want) and (got["u1c"] == got["u2c"]).all()
        # negative disparity (cameras swapped): rejected by the disparity rule
        eng.batch_extract_host(0, right)
        eng.batch_extract_host(2, left)
        eng.batch_match_quad([0], [2], [1], [3])
        assert eng.batch_counts(1)[0][0] == len(orc.quad_matching(fr[0], fl[0], fr[1], fl[1], 2)) == 0
    finally:
        eng.close()


def test_mono_engine_keeps_reference_semantics_for_I2_and_method():
    """the reference ignores I2 and `method` (src/matcher.cpp:51-111): a default engine must do the same"""
    w, h = 500, 150
    left, right = make_stereo_sequence(w, h, 2, seed=6, disparity=11)
    orc = Oracle(w, h)
    eng = Engine(w, h)
    try:
        for k in range(2):
            eng.push_back(left[k], img2=right[k])
        eng.match_features(2)
        want, _ = orc.matching(*_feats(orc, left[0]), *_feats(orc, left[1]))
        got = eng.get_matches()
        assert matches_digest(got) == matches_digest(want) and (got["u2c"] == -1).all()
        with pytest.raises(VoError):
            eng.batch_match_quad([0], [0], [1], [1])
    finally:
        eng.close()


def test_matcher_level_quad_ring():
    w, h = 500, 150
    left, right, orc, fl, fr = _quad_case(w, h, 7, [10, 12, 9], n=3)
    eng = Engine(w, h, stereo=True)
    try:
        eng.push_back(left[0], img2=right[0])
        eng.match_features(2)
        assert eng.num_matches() == 0                          # frame 0
        for k in (1, 2):
            eng.push_back(left[k], img2=right[k])
            eng.match_features(2)
            want = orc.quad_matching(fl[k - 1], fr[k - 1], fl[k], fr[k], 2)
            assert matches_digest(eng.get_matches()) == matches_digest(want)
            eng.match_features(0)                              # the flow matcher on the same ring
            wf, _ = orc.matching(*fl[k - 1], *fl[k])
            assert matches_digest(eng.get_matches()) == matches_digest(wf)
    finally:
        eng.close()


# (306 x 186: the response planes K1 leaves for K5 need one more tile column and one more tile row than the feature lists do)
@pytest.mark.parametrize("w,h", [(500, 150), (1241, 376), (306, 186)])
def test_refinement_flow_and_quad_vs_oracle(w, h):
    left, right, orc, fl, fr = _quad_case(w, h, 2, 13, n=2, blur_passes=6)
    eng = Engine(w, h, max_frames=4, max_pairs=1, stereo=True)
    try:
        eng.batch_extract_host(0, left)
        eng.batch_extract_host(2, right)
        # flow list
        eng.batch_match([0], [1])
        raw = eng.batch_get_matches(0, 0)
        eng.batch_refine(1)
        got = eng.batch_get_matches(0, 0)
        want = orc.refine(raw, left[0], left[1])
        assert 0.5 * len(raw) < len(want) <= len(raw)
        assert len(got) == len(want) and got.tobytes() == want.tobytes()
        assert np.abs(got["u1p"] - got["u1c"] - 3).max() < 1.0 and (got["u1p"] != np.rint(got["u1p"])).any()
        # quad list, refinement fused into the match call, then bucketing
        eng.batch_match_quad([0], [2], [1], [3], refinement=2, max_features=2)
        rawq = orc.quad_matching(fl[0], fr[0], fl[1], fr[1], 2)
        wantq = orc.refine(rawq, left[0], left[1], right[0], right[1])
        gotq = eng.batch_get_matches(0, 0)
        assert len(gotq) == len(wantq) > 100 and gotq.tobytes() == wantq.tobytes()
        wb = orc.bucket_features(wantq, 2, 50.0, 50.0)
        assert eng.batch_get_matches(0, 1).tobytes() == wb.tobytes()
        # sub-pixel disparities stay close to the rig's 13 px
        assert np.abs((gotq["u1c"] - gotq["u2c"]) - 13).max() < 1.0
    finally:
        eng.close()


def test_refinement_matcher_level_and_state_errors():
    w, h = 500, 150
    seq = make_sequence(w, h, 2, seed=9)
    orc = Oracle(w, h)
    eng = Engine(w, h)
    try:
        eng.push_back(seq[0]); eng.push_back(seq[1])
        eng.match_features(0)
        raw = eng.get_matches()
        eng.refine_matches()
        assert eng.get_matches().tobytes() == orc.refine(raw, seq[0], seq[1]).tobytes()
        eng.bucket_features(2, 50.0, 50.0)
        with pytest.raises(VoError) as ei:
            eng.refine_matches()                               # refinement comes before bucketing
        assert ei.value.status == VO_ERR_STATE
    finally:
        eng.close()


def test_stereo_process_loop_on_a_rendered_rig(tmp_path):
    """VisualOdometryStereo::process (/root/reference/src/viso_stereo.cpp:33-52) over a rendered 3-D corridor seen by a
    rectified rig: bucketed quad lists == oracle quad matching -> removeOutliers -> bucketing (unpinned extension),
    motion == the compiled reference's estimateMotion on those lists within 1e-6 (pinned half), and it is the rendered
    motion in metres (the baseline fixes the scale).  The C++ shim (vo_demo --stereo) must report the same counts."""
    import re
    import subprocess
    from hls_final_visual_odometry_b200 import VisualOdometryStereo, build
    from hls_final_visual_odometry_b200.api import remove_outliers_host
    from hls_final_visual_odometry_b200.synth import KITTI_CALIB, make_corridor_sequence
    from oracle.ref_lib import RefLib, have_ref, stereo_motion_seq
    w, h, n, base = 1241, 376, 4, 0.54
    left, poses = make_corridor_sequence(w, h, n_frames=n, seed=1)
    right, _ = make_corridor_sequence(w, h, n_frames=n, seed=1, cam_offset_x=base)
    p = VisualOdometryStereo.parameters()
    p.f, p.cu, p.cv, p.base = KITTI_CALIB["f"], KITTI_CALIB["cu"], KITTI_CALIB["cv"], base
    vo = VisualOdometryStereo(w, h, p)
    orc = Oracle(w, h)
    fl = [_feats(orc, f) for f in left]
    fr = [_feats(orc, f) for f in right]
    lists, got_tr, counts = [], [], []
    for k in range(n):
        ok = vo.process(left[k], right[k])
        counts.append(vo.getNumberOfMatches())
        if k == 0:
            assert not ok and vo.getNumberOfMatches() == 0
            continue
        want = orc.bucket_features(remove_outliers_host(orc.quad_matching(fl[k - 1], fr[k - 1], fl[k], fr[k], 2)), 2, 50.0, 50.0)
        assert matches_digest(vo.p_matched) == matches_digest(want) and len(want) > 100
        assert ok
        lists.append(want)
        got_tr.append(np.array(vo.getMotion()))
        rel = np.linalg.inv(poses[k]) @ poses[k - 1]                    # rendered motion prev -> curr: 0.6 m forward
        assert np.abs(vo.getMotion()[:3, 3] - rel[:3, 3]).max() < 0.05 and np.abs(vo.getMotion()[:3, :3] - rel[:3, :3]).max() < 5e-3
    if have_ref("native"):
        # the reference object sees list 1 twice (bootstrap + regular pass of the first valid frame)
        from hls_final_visual_odometry_b200 import transformation_vector_to_matrix
        tr_ref, ok_ref, _ = stereo_motion_seq(RefLib("native"), [lists[0]] + lists, base=base, **KITTI_CALIB)
        for k in range(1, n):
            assert ok_ref[k] == 1
            assert np.abs(got_tr[k - 1] - transformation_vector_to_matrix(tr_ref[k])).max() <= 1e-6
    # C++ drop-in
    build.build_all()
    for k in range(n):
        left[k].tofile(str(tmp_path / f"I1_{k:06d}.dat"))
        right[k].tofile(str(tmp_path / f"I2_{k:06d}.dat"))
    out = subprocess.run([build.DEMO, str(tmp_path), str(w), str(h), str(n), str(p.f), str(p.cu), str(p.cv), "--stereo", str(base)],
                         capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr
    assert [int(re.search(r"Matches: (\d+)", l).group(1)) for l in out.stdout.splitlines() if "Matches:" in l] == counts


@pytest.mark.parametrize("name", ["holes", "periodic", "dense", "dead_zone_only", "tiny_patch", "low_contrast", "big_motion", "flat"])
@pytest.mark.parametrize("size", [(1024, 284), (500, 150)])
def test_quad_adversarial_vs_oracle(name, size):
    """the mono path's adversarial frame pairs (tests/cases.py: lagging v_bin, first-wins ties, dead column/row, '<5 -> 0',
    chunked windows) as the left camera of a rig whose right camera sees the same scene 11 px to the left, one row down"""
    from cases import CASES
    w, h = size
    prev, curr = CASES[name](w, h, seed=2)
    shift = lambda img: np.ascontiguousarray(np.roll(np.roll(img, -11, axis=1), 1, axis=0))
    frames = [prev, shift(prev), curr, shift(curr)]                       # slots 1p, 2p, 1c, 2c
    orc = Oracle(w, h)
    f = [_feats(orc, x) for x in frames]
    want = orc.quad_matching(f[0], f[1], f[2], f[3], 2)
    eng = Engine(w, h, max_frames=4, max_pairs=1, stereo=True)
    try:
        eng.batch_extract_host(0, np.stack(frames))
        eng.batch_match_quad([0], [1], [2], [3], disp_tolerance=2, refinement=0, max_features=2)
        got = eng.batch_get_matches(0, 0)
        assert len(got) == len(want) and matches_digest(got) == matches_digest(want)
        assert matches_digest(eng.batch_get_matches(0, 1)) == matches_digest(orc.bucket_features(want, 2, 50.0, 50.0))
        eng.batch_refine(1)
        wr = orc.refine(want, frames[0], frames[2], frames[1], frames[3])
        assert eng.batch_get_matches(0, 0).tobytes() == wr.tobytes()
    finally:
        eng.close()


def test_quad_tall_frames_row_groups():
    """2048x1024 (V = 23: origin rows are split into groups of 8 per CTA, windows staged in chunks) on a dense texture"""
    w, h = 2048, 1024
    left, right, orc, fl, fr = _quad_case(w, h, 1, 21, n=2, blur_passes=4)
    eng = Engine(w, h, max_frames=4, max_pairs=1, stereo=True)
    try:
        eng.batch_extract_host(0, left)
        eng.batch_extract_host(2, right)
        eng.batch_match_quad([0], [2], [1], [3], disp_tolerance=2, refinement=2)
        want = orc.refine(orc.quad_matching(fl[0], fr[0], fl[1], fr[1], 2, cap=1 << 19), left[0], left[1], right[0], right[1])
        got = eng.batch_get_matches(0, 0)
        assert len(want) > 50000 and got.tobytes() == want.tobytes()
    finally:
        eng.close()


@pytest.mark.parametrize("amp", [2, 10])
def test_quad_matching_under_sensor_noise_vs_oracle(amp):
    """Independent noise on all four images: K2's two-word filtering on flow AND left<->right edges; refinement on top."""
    w, h, n = 800, 240, 3
    left, right = make_stereo_sequence(w, h, n, seed=8, disparity=[9, 10, 11])
    rng = np.random.default_rng(amp)
    left = np.clip(left.astype(np.int16) + rng.integers(-amp, amp + 1, left.shape, dtype=np.int16), 0, 255).astype(np.uint8)
    right = np.clip(right.astype(np.int16) + rng.integers(-amp, amp + 1, right.shape, dtype=np.int16), 0, 255).astype(np.uint8)
    orc = Oracle(w, h)
    fl = [_feats(orc, f) for f in left]
    fr = [_feats(orc, f) for f in right]
    eng = Engine(w, h, max_frames=2 * n, max_pairs=n - 1, stereo=True)
    try:
        eng.batch_extract_host(0, left)
        eng.batch_extract_host(n, right)
        k = np.arange(1, n)
        eng.batch_match_quad(k - 1, n + k - 1, k, n + k, disp_tolerance=2, refinement=2, max_features=0)
        for i in range(n - 1):
            want = orc.quad_matching(fl[i], fr[i], fl[i + 1], fr[i + 1], 2)
            want = orc.refine(want, left[i], left[i + 1], right[i], right[i + 1])
            got = eng.batch_get_matches(i, 0)
            assert len(want) > 100
            assert got.tobytes() == want.tobytes(), f"noise {amp} item {i}"
    finally:
        eng.close()
