#!/usr/bin/env python
"""Developer A/B timing of several builds of libvo_b200.so (tools/ab_build.py) inside ONE process, interleaved, so that clock and
tenant drift of the box hits every build alike: per build one engine on the same noisy frames, REPS rounds of ITERS matched
batches each in rotating order, median of the per-round K2 / K1 times.  Parity (pruned search == oracle on the first pair) is
checked for every build when VO_PROBE_PARITY is set.

    python tools/ab_probe.py base split2 ... [--amps 2,8] [--reps 7] [--iters 6]
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

from hls_final_visual_odometry_b200 import api  # noqa: E402
from hls_final_visual_odometry_b200.synth import make_sequence  # noqa: E402

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def engine_for(tag, w, h, b, **kw):
    api._lib = None
    os.environ["VO_B200_LIB"] = os.path.join(ROOT, "build", "ab", f"libvo_{tag}.so")
    return api.Engine(w, h, max_frames=b + 1, max_pairs=b, **kw)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("tags", nargs="+")
    ap.add_argument("--amps", default="2,8")
    ap.add_argument("--reps", type=int, default=7)
    ap.add_argument("--iters", type=int, default=6)
    ap.add_argument("--size", default="1241x376")
    ap.add_argument("--pairs", type=int, default=64)
    ap.add_argument("--seed-good", type=int, default=0)
    a = ap.parse_args()
    w, h = (int(x) for x in a.size.split("x"))
    b = a.pairs
    base = make_sequence(w, h, b + 1, seed=1).astype(np.int16)
    prev = np.arange(b, dtype=np.int32)
    for amp in (int(x) for x in a.amps.split(",")):
        rng = np.random.default_rng(100 + amp)
        seq = np.clip(base + (rng.integers(-amp, amp + 1, base.shape, dtype=np.int16) if amp else 0), 0, 255).astype(np.uint8)
        engs = {}
        want = None
        for tag in a.tags:
            e = engine_for(tag, w, h, b, match_seed_good=a.seed_good)
            e.batch_extract_host(0, seq)
            e.batch_match(prev, prev + 1, 2, 50.0, 50.0)
            e.synchronize()
            if os.environ.get("VO_PROBE_PARITY"):
                from oracle.oracle_lib import Oracle, matches_digest
                if want is None:
                    orc = Oracle(w, h)
                    f0, f1 = orc.compute_features(seq[0])[:2], orc.compute_features(seq[1])[:2]
                    want = matches_digest(orc.matching(*f0, *f1)[0])
                assert matches_digest(e.batch_get_matches(0, 0)) == want, f"{tag}: parity lost at +-{amp}"
            engs[tag] = e
        times = {t: {"match": [], "extract": []} for t in a.tags}
        for rep in range(a.reps):
            order = a.tags[rep % len(a.tags):] + a.tags[:rep % len(a.tags)]
            for tag in order:
                e = engs[tag]
                e.profile_enable(True)
                for _ in range(a.iters):
                    e.batch_extract_host(0, seq)
                    e.batch_match(prev, prev + 1, 2, 50.0, 50.0)
                e.synchronize()
                prof = e.profile_read()
                e.profile_enable(False)
                for k in times[tag]:
                    times[tag][k].append(prof[k][0] / max(prof[k][1], 1))
        ref = np.median(times[a.tags[0]]["match"])
        for tag in a.tags:
            m = np.array(times[tag]["match"])
            print(f"+-{amp:2d} {tag:12s} K2 median {np.median(m):.4f} ms (min {m.min():.4f}, max {m.max():.4f}) per {b} pairs  "
                  f"{np.median(m) / ref:6.3f} x {a.tags[0]}   K1 {np.median(times[tag]['extract']):.4f}", flush=True)
        for e in engs.values():
            e.close()


if __name__ == "__main__":
    main()
