"""CPU: the oracle restatement against the reference's own sources compiled in place (oracle/_ref).
Runs wherever oracle/_ref/*.so exists (built from /root/reference by oracle/build_ref.sh)."""
import numpy as np
import pytest

import cases
from hls_final_visual_odometry_b200.synth import make_sequence
from oracle.oracle_lib import Oracle, features_digest, matches_digest
from oracle.ref_lib import RefLib, have_ref

needs_ref = pytest.mark.skipif(not have_ref("native"), reason="oracle/_ref not built (no /root/reference here)")


@needs_ref
@pytest.mark.parametrize("tag", ["native", "1024x284", "500x150", "1241x376"])
def test_operators_bit_exact(tag):
    ref = RefLib(tag)
    orc = Oracle(ref.W, ref.H)
    assert (orc.U, orc.V) == (ref.U, ref.V)
    seq = make_sequence(ref.W, ref.H, 2, seed=7)
    s = (slice(2, -2), slice(2, -2))
    du, dv = ref.sobel(seq[0]); odu, odv = orc.sobel(seq[0])
    assert (du[s] == odu[s]).all() and (dv[s] == odv[s]).all()
    assert (ref.blob(seq[0])[s] == orc.blob(seq[0])[s]).all()
    assert (ref.checkerboard(seq[0])[s] == orc.checkerboard(seq[0])[s]).all()
    feats_r = [ref.compute_features(seq[f]) for f in range(2)]
    feats_o = [orc.compute_features(seq[f])[:2] for f in range(2)]
    for (mr, nr), (mo, no) in zip(feats_r, feats_o):
        assert (nr == no).all()
        assert features_digest(mr, nr) == features_digest(mo, no)
    raw_r = ref.matching(feats_r[0][0], feats_r[0][1], feats_r[1][0], feats_r[1][1])
    raw_o, stats = orc.matching(feats_o[0][0], feats_o[0][1], feats_o[1][0], feats_o[1][1])
    assert stats[1] == 0  # UB fence: no origin without candidates => reference output is defined
    assert len(raw_r) == len(raw_o) and matches_digest(raw_r) == matches_digest(raw_o)


@pytest.mark.parametrize("tag", ["1241x376_b45x45", "1024x284_b60x30", "1241x376_b90x30"])
def test_bin_sizes_bit_exact(tag):
    """live bin size: the reference's own sources with other BIN_W / BIN_H numbers (src/size.h:10-11) against
    the oracle generalised the same way"""
    if not have_ref(tag):
        pytest.skip("oracle/_ref not built for this bin size")
    ref = RefLib(tag)
    orc = Oracle(ref.W, ref.H, ref.BIN_W, ref.BIN_H)
    assert (orc.U, orc.V, orc.CAP) == (ref.U, ref.V, ref.CAP)
    seq = make_sequence(ref.W, ref.H, 3, seed=5)
    feats_r = [ref.compute_features(seq[f]) for f in range(3)]
    feats_o = [orc.compute_features(seq[f])[:2] for f in range(3)]
    for (mr, nr), (mo, no) in zip(feats_r, feats_o):
        assert features_digest(mr, nr) == features_digest(mo, no)
    for a, b in ((0, 1), (1, 2)):
        raw_r = ref.matching(feats_r[a][0], feats_r[a][1], feats_r[b][0], feats_r[b][1])
        raw_o, stats = orc.matching(feats_o[a][0], feats_o[a][1], feats_o[b][0], feats_o[b][1])
        assert stats[1] == 0
        assert len(raw_r) == len(raw_o) and matches_digest(raw_r) == matches_digest(raw_o)


@needs_ref
def test_native_and_override_builds_agree():
    """the geometry-override build (only numbers replaced) reproduces the unmodified reference"""
    a, b = RefLib("native"), RefLib("1024x284")
    seq = make_sequence(1024, 284, 2, seed=11)
    fa = [a.compute_features(seq[f]) for f in range(2)]
    fb = [b.compute_features(seq[f]) for f in range(2)]
    for (ma, na), (mb, nb) in zip(fa, fb):
        assert features_digest(ma, na) == features_digest(mb, nb)
    ra = a.matching(fa[0][0], fa[0][1], fa[1][0], fa[1][1])
    rb = b.matching(fb[0][0], fb[0][1], fb[1][0], fb[1][1])
    assert matches_digest(ra) == matches_digest(rb)


@needs_ref
@pytest.mark.parametrize("name", ["holes", "periodic", "big_motion", "dead_zone_only", "tiny_patch"])
def test_adversarial_cases(name):
    ref = RefLib("1024x284")
    orc = Oracle(ref.W, ref.H)
    a, b = cases.CASES[name](ref.W, ref.H, 0)
    ma, na = ref.compute_features(a); mb, nb = ref.compute_features(b)
    oa, ona, _ = orc.compute_features(a); ob, onb, _ = orc.compute_features(b)
    assert features_digest(ma, na) == features_digest(oa, ona)
    assert features_digest(mb, nb) == features_digest(ob, onb)
    rr = ref.matching(ma, na, mb, nb, cap=1 << 18)
    ro, _ = orc.matching(oa, ona, ob, onb, cap=1 << 18)
    assert len(rr) == len(ro) and matches_digest(rr) == matches_digest(ro)


@needs_ref
def test_bucketing_and_lfsr():
    ref = RefLib("native")
    orc = Oracle(ref.W, ref.H)
    x = 5
    for _ in range(20000):
        assert ref.rand_number(x) == orc.rand_number(x)
        x = ref.rand_number(x)
    for v in (0, 1, 2, 0x7FFFFFFF, 0x80000000, 0xFFFFFFFF, 0x80000001):
        assert ref.rand_number(v) == orc.rand_number(v)
    seq = make_sequence(ref.W, ref.H, 2, seed=3)
    f0 = ref.compute_features(seq[0]); f1 = ref.compute_features(seq[1])
    raw = ref.matching(f0[0], f0[1], f1[0], f1[1])
    assert len(raw) <= ref.POINT_L
    for mf in (1, 2, 3):
        assert matches_digest(ref.bucket_features(raw, mf, 50.0, 50.0)) == matches_digest(orc.bucket_features(raw, mf, 50.0, 50.0))
    for n in (0, 1, 3, 9, 100):
        assert matches_digest(ref.bucket_features(raw[:n], 2, 50.0, 50.0)) == matches_digest(orc.bucket_features(raw[:n], 2, 50.0, 50.0))
