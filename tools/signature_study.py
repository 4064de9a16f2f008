#!/usr/bin/env python
"""Offline study behind the grouping of K2's lower-bound signature (csrc/vo_common.cuh, desc_q).

Any partition of the 32 descriptor bytes into groups gives a valid bound  SAD(a,b) >= sum_g |S_g(a) - S_g(b)|;  what it is worth
depends on how little the byte differences inside a group cancel.  On the oracle's features of two synthetic frames with sensor
noise this script measures, with every origin's own best SAD as the bound, the share of its window candidates a partition lets
through (separately for origins with a partner, best SAD <= 96, and without), and can search the partitions of the 16 du and the
16 dv samples into four groups each by simulated annealing.

CPU only (numpy + the oracle); nothing here is on the product path.

    python tools/signature_study.py [--noise 2] [--seed 1] [--origins 600] [--anneal 1500]
"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from filter_study import flat  # noqa: E402
from hls_final_visual_odometry_b200 import synth  # noqa: E402
from oracle.oracle_lib import Oracle  # noqa: E402

W, H = 1241, 376

# group index of sample k (du, dv); sample offsets: vo_extract.cu sdx/sdy
PARTITIONS = {
    "r1e (by sample index)": ([0] * 4 + [1] * 4 + [2] * 4 + [3] * 4,) * 2,
    "r2t (du left/right/top/bottom, dv rows)": ([2, 2, 3, 0, 1, 1, 3, 0, 3, 0, 3, 0, 2, 2, 1, 1],
                                                [0, 1, 0, 1, 0, 1, 0, 1, 2, 3, 2, 3, 2, 3, 2, 3]),
    "du as r2t, dv as r1e": ([2, 2, 3, 0, 1, 1, 3, 0, 3, 0, 3, 0, 2, 2, 1, 1], [0] * 4 + [1] * 4 + [2] * 4 + [3] * 4),
    "outer ring split top/bottom": ([0, 0, 0, 0, 1, 1, 1, 1, 2, 3, 2, 3, 2, 3, 2, 3],) * 2,
}


def pairs(noise, seed, n_origins):
    """signed byte differences (du and dv parts) of sampled origins against all their window candidates + the origin's best SAD"""
    o = Oracle(W, H)
    canvas = synth.make_canvas(W, H, 1)
    rng = np.random.default_rng(500)
    fr = []
    for f in range(2):
        x = synth.frame_from_canvas(canvas, W, H, f).astype(np.int16)
        if noise:
            x = np.clip(x + rng.integers(-noise, noise + 1, x.shape, dtype=np.int16), 0, 255)
        fr.append(x.astype(np.uint8))
    up, vp, vbp, ubp, cp, dp = flat(*o.compute_features(fr[0])[:2])
    uc, vc, vbc, ubc, cc, dc = flat(*o.compute_features(fr[1])[:2])
    rs = np.random.default_rng(seed)
    ds, ts = [], []
    for i in rs.choice(len(uc), n_origins, replace=False):
        m = (cp == cc[i]) & (np.abs(up - uc[i]) <= 180) & (np.abs(vp - vc[i]) <= 135) & (ubp >= ubc[i] - 3) & (ubp < ubc[i] + 3) & \
            (vbp >= vbc[i] - 2) & (vbp < vbc[i] + 4)
        idx = np.nonzero(m)[0]
        if len(idx) == 0:
            continue
        d = dp[idx].astype(np.int32) - dc[i].astype(np.int32)
        ds.append(d)
        ts.append(np.full(len(idx), np.abs(d).sum(1).min()))
    d = np.concatenate(ds)
    return d[:, 0::2].copy(), d[:, 1::2].copy(), np.concatenate(ts)


def q_of(dx, assign):
    q = 0
    for g in range(4):
        cols = np.nonzero(np.asarray(assign) == g)[0]
        if len(cols):
            q = q + np.abs(dx[:, cols].sum(1))
    return q


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--noise", type=int, default=2)
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--origins", type=int, default=600)
    ap.add_argument("--anneal", type=int, default=0, help="annealing steps per restart (3 restarts)")
    a = ap.parse_args()
    ddu, ddv, t = pairs(a.noise, a.seed, a.origins)
    poor = t > 96
    print(f"noise +-{a.noise}: {len(t)} (origin, candidate) pairs, {poor.mean():.2f} of them of origins without a partner (best SAD > 96)")

    def score(adu, adv):
        h = (q_of(ddu, adu) + q_of(ddv, adv)) <= t
        return h[poor].mean(), h[~poor].mean()

    for name, (adu, adv) in PARTITIONS.items():
        s = score(adu, adv)
        print(f"  {name:44s} pass: no partner {s[0]:.4f}   partner {s[1]:.5f}")
    if not a.anneal:
        return
    rs = np.random.default_rng(a.seed + 100)
    for restart in range(3):
        adu, adv = (np.array(p) for p in PARTITIONS["outer ring split top/bottom"])
        if restart:
            adu, adv = rs.integers(0, 4, 16), rs.integers(0, 4, 16)
        qdu, qdv = q_of(ddu, adu), q_of(ddv, adv)
        cur = ((qdu + qdv) <= t)[poor].mean()
        temp = 0.004
        for _ in range(a.anneal):
            which = rs.integers(2)
            p = (adu if which == 0 else adv).copy()
            if rs.random() < 0.5:
                p[rs.integers(16)] = rs.integers(4)
            else:
                i, j = rs.integers(16, size=2)
                p[i], p[j] = p[j], p[i]
            if len(np.unique(p)) < 4:
                continue
            qn = q_of(ddu if which == 0 else ddv, p)
            s = ((qn + (qdv if which == 0 else qdu)) <= t)[poor].mean()
            if s < cur or rs.random() < np.exp((cur - s) / temp):
                cur = s
                if which == 0:
                    adu, qdu = p, qn
                else:
                    adv, qdv = p, qn
            temp *= 0.995
        print(f"  restart {restart}: {cur:.4f}  du {adu.tolist()}  dv {adv.tolist()}", flush=True)


if __name__ == "__main__":
    main()
