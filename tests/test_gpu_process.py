"""BASELINE.json config 3: the full VisualOdometryMono::process loop (/root/reference/src/viso_mono.cpp:33-39) --
GPU extraction + circular matching, host removeOutliers, GPU bucketing, host egomotion -- through the sequence-level
C ABI (vo_pipeline_*), against the compiled reference run stage by stage on the same frames.

Integer stages (the bucketed match list that feeds egomotion): bit-exact.  Egomotion: |delta| <= 1e-6 per motion
parameter, same success flag, same inlier count.  Frames: a rendered 3-D corridor (synth.make_corridor_sequence),
because the reference's mono egomotion fails by construction on pure image translations."""
import numpy as np
import pytest

from hls_final_visual_odometry_b200 import Pipeline, VisualOdometryMono
from hls_final_visual_odometry_b200.synth import KITTI_CALIB, KITTI_HEIGHT, make_corridor_sequence, make_sequence
from oracle.oracle_lib import Oracle, matches_digest
from oracle.ref_lib import RefLib, have_ref, mono_motion_seq

pytestmark = pytest.mark.gpu
W, H = 1241, 376
TOL = 1e-6


@pytest.fixture(scope="module")
def corridor():
    frames, poses = make_corridor_sequence(W, H, n_frames=6, seed=0)
    return frames, poses


def _reference_chain(frames):
    """matchFeatures (incl. removeOutliers) from the compiled reference; bucketing through the oracle port (the
    reference's bucket array overflows at this size, src/matcher.cpp:154-155); estimateMotion from the reference."""
    ref, orc = RefLib(f"{W}x{H}"), Oracle(W, H)
    m = ref.matcher()
    lists = []
    for k in range(len(frames)):
        m.push_back(frames[k])
        if k == 0:
            continue
        m.match_features()
        lists.append(orc.bucket_features(m.get_matches(), 2, 50.0, 50.0))
    m.close()
    tr, ok, ninl = mono_motion_seq(ref, lists, height=KITTI_HEIGHT, pitch=0.0, **KITTI_CALIB)
    return lists, tr, ok, ninl


def test_pipeline_matches_reference_process_loop(corridor):
    if not have_ref(f"{W}x{H}"):
        pytest.skip("oracle/_ref not built")
    frames, poses = corridor
    lists, tr_ref, ok_ref, ninl_ref = _reference_chain(frames)
    pipe = Pipeline(W, H, batch=4, cam_height=KITTI_HEIGHT, pitch=0.0, **KITTI_CALIB)
    got = [pipe.process(frames[:4], want_matches=True), pipe.process(frames[4:], want_matches=True)]  # second call pairs with frame 3
    pipe.close()
    ok = np.concatenate([g[0] for g in got]); tr = np.concatenate([g[1] for g in got])
    nm = np.concatenate([g[2] for g in got]); ni = np.concatenate([g[3] for g in got])
    ml = got[0][4] + got[1][4]
    assert ok[0] == 0 and nm[0] == 0                                   # frame 0: no previous frame
    for k in range(1, len(frames)):
        assert matches_digest(ml[k]) == matches_digest(lists[k - 1]), f"frame {k}: bucketed list differs"
        assert ok[k] == ok_ref[k - 1] == 1 and ni[k] == ninl_ref[k - 1]
        assert np.abs(tr[k] - tr_ref[k - 1]).max() <= TOL
        rel = np.linalg.inv(poses[k]) @ poses[k - 1]                   # and it is the rendered motion (0.6 m forward)
        assert abs(tr[k][5] - rel[2, 3]) < 0.05 and np.abs(tr[k][:3]).max() < 0.01


def test_pipeline_equals_frame_by_frame_process(corridor):
    frames, _ = corridor
    p = VisualOdometryMono.parameters()
    p.f, p.cu, p.cv, p.height, p.pitch = KITTI_CALIB["f"], KITTI_CALIB["cu"], KITTI_CALIB["cv"], KITTI_HEIGHT, 0.0
    vo = VisualOdometryMono(W, H, p)
    seq = []
    for k in range(len(frames)):
        ok = vo.process(frames[k])
        seq.append((ok, vo.getNumberOfMatches(), vo.getNumberOfInliers() if ok else 0, vo.getMotion().copy()))
    pipe = Pipeline(W, H, batch=len(frames), cam_height=KITTI_HEIGHT, pitch=0.0, **KITTI_CALIB)
    ok, tr, nm, ni = pipe.process(frames)
    pipe.close()
    from hls_final_visual_odometry_b200 import transformation_vector_to_matrix
    for k in range(len(frames)):
        assert bool(ok[k]) == seq[k][0] and nm[k] == seq[k][1]
        if ok[k]:
            assert ni[k] == seq[k][2]
            assert np.array_equal(transformation_vector_to_matrix(tr[k]), seq[k][3])   # same code, same samples: bit-exact


def test_a_large_batch_equals_small_batches():
    """The batch size is a schedule, not a parameter of the result: one 40-frame call of vo_pipeline_process must give exactly
    what eight 5-frame calls give -- same bucketed lists, same sample sets in frame order, hence the same motions bit for bit."""
    frames, _ = make_corridor_sequence(W, H, n_frames=40, seed=1)
    big = Pipeline(W, H, batch=40, cam_height=KITTI_HEIGHT, pitch=0.0, **KITTI_CALIB)
    ok_b, tr_b, nm_b, ni_b, ml_b = big.process(frames, want_matches=True)
    big.close()
    small = Pipeline(W, H, batch=5, cam_height=KITTI_HEIGHT, pitch=0.0, **KITTI_CALIB)
    parts = [small.process(frames[i:i + 5], want_matches=True) for i in range(0, 40, 5)]
    small.close()
    ok_s = np.concatenate([q[0] for q in parts]); tr_s = np.concatenate([q[1] for q in parts])
    nm_s = np.concatenate([q[2] for q in parts]); ni_s = np.concatenate([q[3] for q in parts])
    ml_s = [m for q in parts for m in q[4]]
    assert ok_b[1:].all() and (nm_b[1:] > 100).all()
    assert np.array_equal(ok_b, ok_s) and np.array_equal(nm_b, nm_s) and np.array_equal(ni_b, ni_s)
    assert np.array_equal(tr_b, tr_s)
    for k in range(40):
        assert matches_digest(ml_b[k]) == matches_digest(ml_s[k]), f"frame {k}"


def test_pipeline_on_pure_translation():
    """The blurred-noise crops are a pure image translation: a planar, degenerate configuration on which the reference's
    mono egomotion returns false.  (Not replayed through the compiled reference here: its
    removeOutliers holds at most POINT_L = 14002 matches, /root/reference/src/delaunator.hpp:10, and this texture
    yields ~17k.)  The integer stages must still equal oracle matching -> removeOutliers -> bucketing."""
    from hls_final_visual_odometry_b200.api import remove_outliers_host
    frames = make_sequence(W, H, 3, seed=2)
    pipe = Pipeline(W, H, batch=3, **KITTI_CALIB)
    ok, tr, nm, ni, ml = pipe.process(frames, want_matches=True)
    pipe.close()
    orc = Oracle(W, H)
    feats = [orc.compute_features(f)[:2] for f in frames]
    assert nm[0] == 0 and not ok.any()
    for k in (1, 2):
        raw, _ = orc.matching(feats[k - 1][0], feats[k - 1][1], feats[k][0], feats[k][1])
        want = orc.bucket_features(remove_outliers_host(raw), 2, 50.0, 50.0)
        assert matches_digest(ml[k]) == matches_digest(want)


def test_gpu_ransac_scoring_equals_the_host_stage(corridor):
    """vo_ransac_gpu_score runs the host stage's arithmetic (csrc/vo_ransac_core.h) one thread per hypothesis: the winner, its
    inlier count, and therefore motion and inlier set are the host's, bit for bit (tolerance 0)."""
    from hls_final_visual_odometry_b200 import Egomotion, RansacGpu
    frames, _ = corridor
    pipe = Pipeline(W, H, batch=len(frames), cam_height=KITTI_HEIGHT, pitch=0.0, **KITTI_CALIB)
    lists = pipe.process(frames, want_matches=True)[4][1:]                # the bucketed lists of the five pairs
    pipe.close()
    rng = np.random.default_rng(3)
    lists.append(lists[0][rng.permutation(len(lists[0]))[:40]])           # a short list
    lists.append(lists[1][:9])                                            # never reaches RANSAC (< 10 matches)
    junk = lists[2].copy()                                                # no geometry at all: low, tied counts -> first-wins rule
    junk["u1c"] = rng.uniform(0, W, len(junk)).astype(np.float32); junk["v1c"] = rng.uniform(0, H, len(junk)).astype(np.float32)
    lists.append(junk)
    ego_a = Egomotion(height=KITTI_HEIGHT, pitch=0.0, ransac_iters=500, **KITTI_CALIB)
    ego_b = Egomotion(height=KITTI_HEIGHT, pitch=0.0, ransac_iters=500, **KITTI_CALIB)
    gpu = RansacGpu(max_lists=len(lists), max_points=max(len(l) for l in lists), max_iters=500)
    try:
        drawn = [ego_a.draw_samples(l) for l in lists]
        for l in lists:
            ego_b.draw_samples(l)                                          # both handles consumed the same stream
        best_k, best_n = gpu.score(lists, ego_a.prm, [d[0] for d in drawn], [d[1] for d in drawn])
        assert best_k[-2] == -1 and drawn[-2][1] == 0
        for i, l in enumerate(lists):
            tr_h, inl_h = ego_a.estimate_with_samples(l, drawn[i][0], drawn[i][1])
            tr_g, inl_g = ego_a.estimate_with_samples(l, drawn[i][0], drawn[i][1], best_k=int(best_k[i]))
            assert (tr_h is None) == (tr_g is None)
            assert np.array_equal(inl_h, inl_g), f"list {i}: inlier sets differ"
            if tr_h is not None:
                assert np.array_equal(tr_h, tr_g), f"list {i}: motion differs"
            if drawn[i][1] and len(inl_h):
                assert best_n[i] == len(inl_h)
    finally:
        gpu.close(); ego_a.close(); ego_b.close()
