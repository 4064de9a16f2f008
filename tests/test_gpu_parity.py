"""GPU parity tests proper: the CUDA path, called through the C ABI, against the CPU oracle
(oracle/vo_oracle.c, itself pinned to the compiled reference) on the same seeded inputs.
Integer work: bit-exact or fail."""
import numpy as np
import pytest

from hls_final_visual_odometry_b200 import Engine
from hls_final_visual_odometry_b200.synth import make_sequence
from oracle.oracle_lib import Oracle, compact_features

pytestmark = pytest.mark.gpu

SIZES = [(500, 150), (1024, 284), (1241, 376)]


def assert_features_equal(got, want, tag=""):
    (gm, gn), (wm, wn) = got, want
    assert gn.shape == wn.shape
    if not (gn == wn).all():
        bad = np.argwhere(gn != wn)
        raise AssertionError(f"{tag}: {len(bad)} (bin,class) counts differ, first {bad[:5].tolist()} "
                             f"got {gn[tuple(bad[0])]} want {wn[tuple(bad[0])]}; totals {gn.sum()} vs {wn.sum()}")
    for i, (g, w) in enumerate(zip(compact_features(gm, gn), compact_features(wm, wn))):
        if not (g == w).all():
            k = int(np.argwhere((g != w).any(axis=1))[0][0])
            raise AssertionError(f"{tag}: list {i} record {k} differs\n got {g[k]}\nwant {w[k]}")


def assert_matches_equal(got, want, tag=""):
    assert len(got) == len(want), f"{tag}: {len(got)} matches, want {len(want)}"
    if got.tobytes() != want.tobytes():
        for i in range(len(got)):
            if got[i] != want[i]:
                raise AssertionError(f"{tag}: match {i} differs\n got {got[i]}\nwant {want[i]}")


CASES = [(w, h, ex) for (w, h) in SIZES for ex in (False, True)]


@pytest.fixture(scope="module", params=CASES, ids=lambda s: f"{s[0]}x{s[1]}-{'exhaustive' if s[2] else 'pruned'}")
def setup(request):
    w, h, exhaustive = request.param
    seq = make_sequence(w, h, 3, seed=0)
    orc = Oracle(w, h)
    eng = Engine(w, h, max_frames=4, max_pairs=2, match_exhaustive=exhaustive)
    yield w, h, seq, orc, eng
    eng.close()


def test_compute_features_bit_exact(setup):
    w, h, seq, orc, eng = setup
    for f in range(2):
        wm, wn, dropped = orc.compute_features(seq[f])
        assert dropped == 0
        gm, gn = eng.compute_features(seq[f])
        assert_features_equal((gm, gn), (wm, wn), f"{w}x{h} frame {f}")


def test_matching_operator_bit_exact(setup):
    w, h, seq, orc, eng = setup
    m0, n0, _ = orc.compute_features(seq[0])
    m1, n1, _ = orc.compute_features(seq[1])
    want, stats = orc.matching(m0, n0, m1, n1)
    got = eng.matching(m0, n0, m1, n1)
    assert len(want) > 100
    assert_matches_equal(got, want, f"{w}x{h}")


def test_matcher_flow_bit_exact(setup):
    w, h, seq, orc, eng = setup
    feats = [orc.compute_features(seq[f])[:2] for f in range(3)]
    eng.push_back(seq[0])
    eng.match_features(0)
    assert eng.num_matches() == 0  # frame-0 contract
    for f in (1, 2):
        eng.push_back(seq[f])
        assert_features_equal(eng.dump_features(1), feats[f], f"curr {f}")
        assert_features_equal(eng.dump_features(0), feats[f - 1], f"prev {f}")
        eng.match_features(0)
        want, _ = orc.matching(feats[f - 1][0], feats[f - 1][1], feats[f][0], feats[f][1])
        got = eng.get_matches()
        assert_matches_equal(got, want, f"{w}x{h} pair {f}")
        eng.bucket_features(2, 50.0, 50.0)
        wb = orc.bucket_features(want, 2, 50.0, 50.0)
        assert_matches_equal(eng.get_matches(), wb, f"{w}x{h} bucketed {f}")


def test_bucket_operator_bit_exact(setup):
    w, h, seq, orc, eng = setup
    m0, n0, _ = orc.compute_features(seq[0])
    m1, n1, _ = orc.compute_features(seq[1])
    raw, _ = orc.matching(m0, n0, m1, n1)
    for mf, bw, bh in [(2, 50.0, 50.0), (1, 50.0, 50.0), (3, 37.5, 61.0), (5, 100.0, 20.0)]:
        want = orc.bucket_features(raw, mf, bw, bh)
        got = eng.bucket_matches(raw, mf, bw, bh)
        assert_matches_equal(got, want, f"bucket mf={mf} {bw}x{bh}")
    # short / empty lists
    for n in (0, 1, 4, 7):
        want = orc.bucket_features(raw[:n], 2, 50.0, 50.0)
        got = eng.bucket_matches(raw[:n], 2, 50.0, 50.0)
        assert_matches_equal(got, want, f"bucket n={n}")


def test_batch_equals_oracle(setup):
    w, h, seq, orc, eng = setup
    feats = [orc.compute_features(seq[f])[:2] for f in range(3)]
    eng.batch_extract_host(0, seq)
    for f in range(3):
        assert_features_equal(eng.slot_dump_features(f), feats[f], f"slot {f}")
        assert eng.slot_feature_count(f) == int(feats[f][1].sum())
    eng.batch_match([0, 1], [1, 2], max_features=2, bw=50.0, bh=50.0)
    raw_n, bk_n = eng.batch_counts(2)
    for p in range(2):
        want, _ = orc.matching(feats[p][0], feats[p][1], feats[p + 1][0], feats[p + 1][1])
        got = eng.batch_get_matches(p, 0)
        assert raw_n[p] == len(want)
        assert_matches_equal(got, want, f"batch pair {p}")
        wb = orc.bucket_features(want, 2, 50.0, 50.0)
        assert bk_n[p] == len(wb)
        assert_matches_equal(eng.batch_get_matches(p, 1), wb, f"batch bucketed {p}")


def test_fpga_operator_signatures(setup):
    """the reference's FPGA-side operator signatures, 1:1: myMatching(..., iter) (hls_opencl/kernel/myMatch.hpp:51-57, iter == 0 ->
    zero matches) and the fused removeOutliers = Delaunay filter then bucketing 2 per 50x50 (remove_outliers.hpp:9-11)"""
    from hls_final_visual_odometry_b200.api import remove_outliers_host
    w, h, seq, orc, eng = setup
    m0, n0, _ = orc.compute_features(seq[0])
    m1, n1, _ = orc.compute_features(seq[1])
    want, _ = orc.matching(m0, n0, m1, n1)
    assert len(eng.matching_iter(m0, n0, m1, n1, 0)) == 0
    assert_matches_equal(eng.matching_iter(m0, n0, m1, n1, 3), want, f"{w}x{h} iter 3")
    fused = eng.remove_outliers_bucket(want)
    assert_matches_equal(fused, orc.bucket_features(remove_outliers_host(want), 2, 50.0, 50.0), f"{w}x{h} fused removeOutliers")
    assert eng.remove_outliers_bucket(want[:3]) is None    # <= 3 matches: outputs untouched (remove_outliers.cpp:20-21)


def test_bad_slots_leave_the_pair_tables_intact(setup):
    """an out-of-range slot is refused before the table generation advances (round-1 advisor finding): the previous
    batch's tables stay valid and the next good call works"""
    from hls_final_visual_odometry_b200 import VoError
    from hls_final_visual_odometry_b200 import api
    w, h, seq, orc, eng = setup
    eng.batch_extract_host(0, seq[:2])
    eng.batch_match([0], [1], max_features=2, bw=50.0, bh=50.0)
    good = eng.batch_get_matches(0, 0)
    for prev, curr in (([0], [99]), ([-1], [1]), ([0, 0, 0], [1, 1, 1])):   # bad slot, bad slot, more pairs than buffers
        with pytest.raises(VoError) as ei:
            eng.batch_match(prev, curr, max_features=2, bw=50.0, bh=50.0)
        assert ei.value.status == api.VO_ERR_INVALID
    eng.batch_match([0], [1], max_features=2, bw=50.0, bh=50.0)
    assert_matches_equal(eng.batch_get_matches(0, 0), good, "after refused calls")
