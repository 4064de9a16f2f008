"""CPU: the oracle restatement (oracle/vo_oracle.c) against the committed golden digests, which were
produced by the reference's own sources compiled in place (tests/golden/make_golden.py)."""
import hashlib
import json
import os

import numpy as np
import pytest

import cases
from hls_final_visual_odometry_b200.synth import make_sequence
from oracle.oracle_lib import Oracle, features_digest, matches_digest

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = json.load(open(os.path.join(HERE, "golden", "golden.json")))


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def _inputs(geom, name):
    W, H = geom["W"], geom["H"]
    if name.startswith("textured_seed"):
        seed = int(name[len("textured_seed"):].split("_")[0])
        f = name.split("_f")[1]
        seq = make_sequence(W, H, 3, seed=seed)
        return seq[int(f[0])], seq[int(f[1])]
    return cases.CASES[name](W, H, 0)


CASE_IDS = [(tag, name) for tag, g in GOLD["geometries"].items() for name in g["cases"]]


@pytest.mark.parametrize("tag,name", CASE_IDS, ids=[f"{t}-{n}" for t, n in CASE_IDS])
def test_oracle_matches_reference_golden(tag, name):
    geom = GOLD["geometries"][tag]
    rec = geom["cases"][name]
    orc = Oracle(geom["W"], geom["H"], geom.get("bin_w", 60), geom.get("bin_h", 45), geom.get("rad_u", 0), geom.get("rad_v", 0))
    assert (orc.U, orc.V) == (geom["U"], geom["V"])
    prev, curr = _inputs(geom, name)
    mp, npv, d0 = orc.compute_features(prev)
    mc, ncv, d1 = orc.compute_features(curr)
    assert d0 == 0 and d1 == 0
    assert int(npv.sum()) == rec["n_prev"] and int(ncv.sum()) == rec["n_curr"]
    assert features_digest(mp, npv) == rec["features_prev"]
    assert features_digest(mc, ncv) == rec["features_curr"]
    raw, stats = orc.matching(mp, npv, mc, ncv, cap=1 << 18)
    assert len(raw) == rec["n_raw"]
    assert matches_digest(raw) == rec["raw_matches"]
    if "bucketed_raw_2_50_50" in rec and rec.get("n_after_outliers") == rec["n_raw"]:
        # removeOutliers kept everything here, so bucketing the raw list is what the reference did
        b = orc.bucket_features(raw, 2, 50.0, 50.0)
        assert len(b) == rec["n_bucketed"]
        assert matches_digest(b) == rec["bucketed_2_50_50"]


def test_lfsr_known_answers():
    orc = Oracle(1024, 284)
    x, seq = 5, []
    for i in range(100000):
        if i < 16:
            seq.append(x)
        x = orc.rand_number(x)
    assert seq == GOLD["lfsr"]["first16"]
    assert x == GOLD["lfsr"]["after_100000"]
    # SURVEY.md row A13 known-answer vector
    assert seq[:8] == [5, 2147483650, 3221225473, 1610612736, 2952790016, 3623878656, 1811939328, 3053453312]


def test_filters_match_reference_golden():
    f = GOLD["filters"]
    orc = Oracle(1024, 284)
    img = make_sequence(1024, 284, 1, seed=f["seed"])[0]
    s = (slice(2, -2), slice(2, -2))
    du, dv = orc.sobel(img)
    assert sha(du[s]) == f["du"] and sha(dv[s]) == f["dv"]
    assert sha(orc.blob(img)[s]) == f["blob"]
    assert sha(orc.checkerboard(img)[s]) == f["checkerboard"]


def test_small_full_vectors():
    z = np.load(os.path.join(HERE, "golden", "small_500x150.npz"))
    orc = Oracle(500, 150)
    seq = make_sequence(500, 150, 2, seed=0)
    mp, npv, _ = orc.compute_features(seq[0])
    mc, ncv, _ = orc.compute_features(seq[1])
    assert (npv == z["num_prev"]).all() and (ncv == z["num_curr"]).all()
    assert (mc[0, 0, :, :4, :] == z["first_records_curr"]).all()
    raw, _ = orc.matching(mp, npv, mc, ncv)
    xy = np.stack([raw["u1p"], raw["v1p"], raw["u1c"], raw["v1c"]], 1).astype(np.int16)
    assert xy.shape == z["raw_xy"].shape and (xy == z["raw_xy"]).all()
    # the other p_match fields are all -1 (src/myMatch.cpp:189)
    for k in ("i1p", "i2p", "i1c", "i2c"):
        assert (raw[k] == -1).all()
    for k in ("u2p", "v2p", "u2c", "v2c"):
        assert (raw[k] == -1.0).all()


def test_oracle_properties():
    """size-independent properties of the matcher output"""
    orc = Oracle(1024, 284)
    seq = make_sequence(1024, 284, 2, seed=5)
    mp, npv, _ = orc.compute_features(seq[0])
    mc, ncv, _ = orc.compute_features(seq[1])
    raw, stats = orc.matching(mp, npv, mc, ncv)
    # every current pixel is used at most once (the M[][] dedupe)
    cur = raw["u1c"].astype(np.int64) * 65536 + raw["v1c"].astype(np.int64)
    assert len(np.unique(cur)) == len(cur)
    # matches stay inside the search window
    assert (np.abs(raw["u1c"] - raw["u1p"]) <= 180).all() and (np.abs(raw["v1c"] - raw["v1p"]) <= 135).all()
    # nothing from the dead last u-bin column
    assert (raw["u1c"] < 60 * (orc.U - 1)).all() and (raw["u1p"] < 60 * (orc.U - 1)).all()
    # synthetic flow is (-3,-1)
    flow = np.stack([raw["u1c"] - raw["u1p"], raw["v1c"] - raw["v1p"]], 1)
    assert (flow == [-3, -1]).all(axis=1).mean() > 0.99
    # bucketing: idempotent at max_features=1 per bucket count and bounded by buckets * k
    b2 = orc.bucket_features(raw, 2, 50.0, 50.0)
    cols = int(raw["u1c"].max() // 50) + 1
    rows = int(raw["v1c"].max() // 50) + 1
    assert len(b2) <= 2 * cols * rows
    ids = (b2["v1c"] // 50).astype(int) * cols + (b2["u1c"] // 50).astype(int)
    assert (np.diff(ids) >= 0).all() and np.bincount(ids).max() <= 2
    # a bucketed list is a sub-multiset of the input
    key = lambda m: set(zip(m["u1p"].tolist(), m["v1p"].tolist(), m["u1c"].tolist(), m["v1c"].tolist()))
    assert key(b2) <= key(raw)
