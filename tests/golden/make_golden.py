"""Generate tests/golden/golden.json and small_500x150.npz from the REFERENCE ITSELF.

Runs only where /root/reference is mounted: oracle/build_ref.sh compiles the reference's own
sources in place into oracle/_ref/*.so, and this script records digests (sha256) of what those
binaries produce on seeded synthetic inputs.  The committed files let every other box (no
reference checkout) pin oracle/vo_oracle.c -- and through it the CUDA path -- to the reference.

    python tests/golden/make_golden.py

Digest definitions: oracle/oracle_lib.py features_digest / matches_digest.
"""
from __future__ import annotations

import hashlib
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np  # noqa: E402

import cases  # noqa: E402
from hls_final_visual_odometry_b200.synth import make_sequence  # noqa: E402
from oracle.oracle_lib import features_digest, matches_digest  # noqa: E402
from oracle.ref_lib import RefLib  # noqa: E402

# cases whose reference output is well defined (no origin without candidates outside the dead
# column -- SURVEY.md section 8c fence (2)); low_contrast@1024x284 is excluded for that reason.
ADVERSARIAL = ["flat", "dense", "holes", "dead_zone_only", "tiny_patch", "periodic", "big_motion", "out_of_window"]


def sha(a: np.ndarray) -> str:
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def pair_record(ref: RefLib, prev, curr, with_post=False):
    mp, npv = ref.compute_features(prev)
    mc, ncv = ref.compute_features(curr)
    raw = ref.matching(mp, npv, mc, ncv, cap=1 << 18)
    rec = {
        "features_prev": features_digest(mp, npv), "features_curr": features_digest(mc, ncv),
        "n_prev": int(npv.sum()), "n_curr": int(ncv.sum()), "max_list": int(max(npv.max(), ncv.max())),
        "raw_matches": matches_digest(raw), "n_raw": int(len(raw)),
    }
    if with_post and 3 < len(raw) <= ref.POINT_L:
        post = ref.remove_outliers(raw)
        rec["after_outliers"] = matches_digest(post)
        rec["n_after_outliers"] = int(len(post))
        u = post["u1c"].max() if len(post) else 0
        v = post["v1c"].max() if len(post) else 0
        if (int(u // 50) + 1) * (int(v // 50) + 1) <= 126:  # the reference's fixed bucket array
            b = ref.bucket_features(post, 2, 50.0, 50.0)
            rec["bucketed_2_50_50"] = matches_digest(b)
            rec["n_bucketed"] = int(len(b))
            b1 = ref.bucket_features(raw if len(raw) <= ref.POINT_L else post, 2, 50.0, 50.0)
            rec["bucketed_raw_2_50_50"] = matches_digest(b1)
    return rec


BIN_TAGS = ["1241x376_b45x45", "1024x284_b60x30", "1241x376_b90x30"]  # live bin size: the reference with other BIN_W / BIN_H


# live search radius: the reference with other SEARCH_RAD_U / SEARCH_RAD_V (96x64: inside the three-bin range;
# 40x150: the v window reaches past the bin range, which then limits it)
RAD_TAGS = ["1241x376_r96x64", "1024x284_r40x150"]


def bin_geometries(tags=None):
    """geometries of the reference compiled with other bin sizes / search radii (oracle/build_ref.sh <W>x<H>[_b<BW>x<BH>][_r<RU>x<RV>])"""
    out = {}
    for tag in (tags or BIN_TAGS):
        ref = RefLib(tag)
        g = {"W": ref.W, "H": ref.H, "U": ref.U, "V": ref.V, "bin_w": ref.BIN_W, "bin_h": ref.BIN_H, "cases": {}}
        rad = (0, 0)
        if "_r" in tag:
            rad = tuple(int(x) for x in tag.split("_r")[1].split("_")[0].split("x"))
            g["rad_u"], g["rad_v"] = rad
        seq = make_sequence(ref.W, ref.H, 3, seed=0)
        g["cases"]["textured_seed0_f01"] = pair_record(ref, seq[0], seq[1], with_post=True)
        g["cases"]["textured_seed0_f12"] = pair_record(ref, seq[1], seq[2], with_post=True)
        from oracle.oracle_lib import Oracle
        orc = Oracle(ref.W, ref.H, ref.BIN_W, ref.BIN_H, rad[0], rad[1])
        for name in ADVERSARIAL:
            a, b = cases.CASES[name](ref.W, ref.H, 0)
            fa, fb = orc.compute_features(a), orc.compute_features(b)
            if orc.matching(fa[0], fa[1], fb[0], fb[1], cap=1 << 18)[1][1] != 0:
                continue  # an origin without candidates outside the dead column: the reference reads an uninitialised
                          # min_cand there (SURVEY.md section 8c fence (2)), its output is not defined
            g["cases"][name] = pair_record(ref, a, b, with_post=(name != "dense"))
        out[tag] = g
        print(tag, "done", sorted(g["cases"]), flush=True)
    return out


def main():
    if "--bins-only" in sys.argv:  # add the bin-size geometries to the committed file, leave the rest as it is
        path = os.path.join(HERE, "golden.json")
        out = json.load(open(path))
        for tag in BIN_TAGS:
            out["geometries"].pop(tag, None)
        out["geometries"].update(bin_geometries())
        json.dump(out, open(path, "w"), indent=1, sort_keys=True)
        return
    if "--radius-only" in sys.argv:  # same for the search-radius geometries
        path = os.path.join(HERE, "golden.json")
        out = json.load(open(path))
        for tag in RAD_TAGS:
            out["geometries"].pop(tag, None)
        out["geometries"].update(bin_geometries(RAD_TAGS))
        json.dump(out, open(path, "w"), indent=1, sort_keys=True)
        return
    out = {"about": "sha256 digests of outputs of the reference sources compiled in place (oracle/_ref)",
           "geometries": {}, "lfsr": {}, "filters": {}, "matcher_flow": {}, "outliers": {}}
    for tag in ["native", "1024x284", "500x150", "1241x376", "2048x1024"]:
        ref = RefLib(tag)
        g = {"W": ref.W, "H": ref.H, "U": ref.U, "V": ref.V, "cases": {}}
        seeds = [0, 1, 2] if tag in ("native", "1241x376", "500x150") else [0]
        for seed in seeds:
            seq = make_sequence(ref.W, ref.H, 3, seed=seed)
            g["cases"][f"textured_seed{seed}_f01"] = pair_record(ref, seq[0], seq[1], with_post=True)
            g["cases"][f"textured_seed{seed}_f12"] = pair_record(ref, seq[1], seq[2], with_post=True)
        if tag != "native":  # every override geometry, incl. the BASELINE sizes (V = 9: one row group; V = 23: row groups, chunked windows)
            for name in ADVERSARIAL:
                a, b = cases.CASES[name](ref.W, ref.H, 0)
                g["cases"][name] = pair_record(ref, a, b, with_post=(name != "dense"))
        out["geometries"][tag] = g
        print(tag, "done", flush=True)
    out["geometries"].update(bin_geometries())
    out["geometries"].update(bin_geometries(RAD_TAGS))

    ref = RefLib("native")
    x, seq_vals = 5, []
    for i in range(100000):
        if i < 16:
            seq_vals.append(x)
        x = ref.rand_number(x)
    out["lfsr"] = {"first16": seq_vals, "after_100000": x}
    img = make_sequence(ref.W, ref.H, 1, seed=0)[0]
    du, dv = ref.sobel(img)
    s = (slice(2, -2), slice(2, -2))
    out["filters"] = {"geometry": "native", "seed": 0, "du": sha(du[s]), "dv": sha(dv[s]),
                      "blob": sha(ref.blob(img)[s]), "checkerboard": sha(ref.checkerboard(img)[s])}

    # the Matcher facade end to end (pushBack / matchFeatures incl. removeOutliers / bucketFeatures /
    # getMatches), native geometry, 4 frames
    seq = make_sequence(ref.W, ref.H, 4, seed=0)
    m = ref.matcher()
    flow = []
    for f in range(4):
        m.push_back(seq[f])
        m.match_features(0)
        after = m.get_matches()
        m.bucket_features(2, 50.0, 50.0)
        b = m.get_matches()
        flow.append({"n_after_outliers": int(len(after)), "after_outliers": matches_digest(after),
                     "n_bucketed": int(len(b)), "bucketed": matches_digest(b)})
    m.close()
    out["matcher_flow"] = {"geometry": "native", "seed": 0, "frames": flow}

    # removeOutliers on explicit match lists with real outliers (tests/cases.py outlier_lists)
    from oracle.ref_lib import P_MATCH_DTYPE
    out["outliers"] = {}
    for seed in (0, 1):
        for name, lst in cases.outlier_lists(P_MATCH_DTYPE, seed).items():
            post = ref.remove_outliers(lst)
            out["outliers"][f"{name}_seed{seed}"] = {"n_in": int(len(lst)), "n_out": int(len(post)),
                                                     "in": matches_digest(lst), "out": matches_digest(post)}
    print("outliers done", flush=True)

    with open(os.path.join(HERE, "golden.json"), "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)

    # one small case with the full vectors (not just digests)
    ref = RefLib("500x150")
    seq = make_sequence(ref.W, ref.H, 2, seed=0)
    mp, npv = ref.compute_features(seq[0])
    mc, ncv = ref.compute_features(seq[1])
    raw = ref.matching(mp, npv, mc, ncv)
    np.savez_compressed(
        os.path.join(HERE, "small_500x150.npz"),
        num_prev=npv.astype(np.int16), num_curr=ncv.astype(np.int16),
        raw_xy=np.stack([raw["u1p"], raw["v1p"], raw["u1c"], raw["v1c"]], 1).astype(np.int16),
        first_records_curr=mc[0, 0, :, :4, :].copy())
    print("wrote golden.json, small_500x150.npz")


if __name__ == "__main__":
    main()
