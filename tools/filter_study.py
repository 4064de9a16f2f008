#!/usr/bin/env python
"""Offline study behind K2's candidate index: on the oracle's features of two synthetic frames with
sensor noise, how tight is the best SAD of an origin, and what fraction of its window candidates survives

  * the two-word L1 bound        Q0 + Q1 <= R                         (r1 filter, 2 VABSDIFF4 per candidate)
  * the per-group interval test  |f(S_g(a)) - f(S_g(b))| <= R for all 8 groups, on cells of `w` codes
                                 (what a bitmap index answers for 32 candidates with 2 loads + 1 LOP3 per group)

CPU only (numpy + the oracle); nothing here is on the product path.

    python tools/filter_study.py [noise ...]
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from hls_final_visual_odometry_b200 import synth  # noqa: E402
from oracle.oracle_lib import Oracle  # noqa: E402

W, H = 1241, 376


def sum_code(s):
    s = s.astype(np.int64)
    lo = np.minimum(s, 410) >> 4
    mid = np.clip(s, 410, 610) - 410
    hi = (np.maximum(s, 610) - 610) >> 4
    return lo + mid + hi


def signature(desc):
    """desc (n, 32) u8 -> (n, 8) codes, the groups of csrc/vo_common.cuh desc_q (sample k: du = byte 2k, dv = byte 2k+1)"""
    d = desc.astype(np.int64)
    du, dv = d[:, 0::2], d[:, 1::2]
    g = [du[:, [0, 1, 12, 13]], du[:, [4, 5, 14, 15]], dv[:, [0, 2, 4, 6]], dv[:, [9, 11, 13, 15]],
         du[:, [2, 6, 8, 10]], du[:, [3, 7, 9, 11]], dv[:, [1, 3, 5, 7]], dv[:, [8, 10, 12, 14]]]
    return sum_code(np.stack([x.sum(axis=1) for x in g], axis=1))


def flat(max2, num):
    V, U, _ = num.shape
    rec = []
    for v in range(V):
        for u in range(U):
            for c in range(4):
                r = max2[v, u, c, : num[v, u, c]]
                if len(r):
                    rec.append(np.concatenate([r[:, :2], np.full((len(r), 1), v), np.full((len(r), 1), u), np.full((len(r), 1), c),
                                               r[:, 4:]], axis=1))
    a = np.concatenate(rec)
    desc = np.ascontiguousarray(a[:, 5:].astype("<i4")).view(np.uint8).reshape(-1, 32)
    return a[:, 0], a[:, 1], a[:, 2], a[:, 3], a[:, 4], desc


def main():
    amps = [int(x) for x in sys.argv[1:]] or [0, 2, 8]
    o = Oracle(W, H)
    canvas = synth.make_canvas(W, H, 1)
    for amp in amps:
        rng = np.random.default_rng(500)
        fr = []
        for f in range(2):
            x = synth.frame_from_canvas(canvas, W, H, f).astype(np.int16)
            if amp:
                x = np.clip(x + rng.integers(-amp, amp + 1, x.shape, dtype=np.int16), 0, 255)
            fr.append(x.astype(np.uint8))
        fp = flat(*o.compute_features(fr[0])[:2])
        fc = flat(*o.compute_features(fr[1])[:2])
        up, vp, vbp, ubp, cp, dp = fp
        uc, vc, vbc, ubc, cc, dc = fc
        sp, sc = signature(dp), signature(dc)
        rs = np.random.default_rng(1)
        pick = rs.choice(len(uc), 1500, replace=False)
        best, own_best, n_win = [], [], []
        stats = {}
        for i in pick:
            m = (cp == cc[i]) & (np.abs(up - uc[i]) <= 180) & (np.abs(vp - vc[i]) <= 135) & (ubp >= ubc[i] - 3) & (ubp < ubc[i] + 3) & \
                (vbp >= vbc[i] - 2) & (vbp < vbc[i] + 4)
            idx = np.nonzero(m)[0]
            if len(idx) == 0:
                continue
            sad = np.abs(dp[idx].astype(np.int32) - dc[i].astype(np.int32)).sum(axis=1)
            dq = np.abs(sp[idx] - sc[i])
            q = dq.sum(axis=1)
            linf = dq.max(axis=1)
            b = sad.min()
            own = (ubp[idx] == ubc[i]) & (vbp[idx] == vbc[i])
            ob = sad[own].min() if own.any() else 9999
            best.append(b); own_best.append(ob); n_win.append(len(idx))
            for R in (8, 16, 24, 32, 48, 64):
                st = stats.setdefault(R, dict(q=0, linf=0, cells8=0, cells16=0, n=0, d4=0))
                st["n"] += len(idx)
                st["q"] += int((q <= R).sum())
                st["linf"] += int((linf <= R).sum())
                st["d4"] += int((dq[:, :4].max(axis=1) <= R).sum())
                for w in (8, 16):
                    lo = np.maximum(sc[i] - R, 0) // w
                    hi = (sc[i] + R) // w
                    cell = sp[idx] // w
                    st[f"cells{w}"] += int(((cell >= lo) & (cell <= hi)).all(axis=1).sum())
        best, own_best = np.array(best), np.array(own_best)
        print(f"noise +-{amp}: features {len(up)}/{len(uc)}, window candidates per origin {np.mean(n_win):.0f}")
        print("  best SAD percentiles 50/75/90/95/99:", np.percentile(best, [50, 75, 90, 95, 99]).astype(int),
              " own-bin seed == best:", f"{(own_best == best).mean():.3f}", " own-bin seed <= 64:", f"{(own_best <= 64).mean():.3f}")
        for R, st in stats.items():
            n = st["n"]
            print(f"  R={R:3d}: pass  Q0+Q1<=R {st['q'] / n:8.5f}   Linf(8 groups) {st['linf'] / n:8.5f}   cells w=8 {st['cells8'] / n:8.5f}"
                  f"   cells w=16 {st['cells16'] / n:8.5f}   Linf(4 groups) {st['d4'] / n:8.5f}")


if __name__ == "__main__":
    main()
