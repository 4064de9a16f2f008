#!/usr/bin/env python
"""The two small runs the sanitizer logs under profiles/ come from:
`__graft_entry__.smoke()` plus one 1241x376 frame pair with +-2 grey levels of sensor noise through extract -> match
(pruned, hinted second direction) -> bucket, compared with the oracle.

    compute-sanitizer --tool memcheck  python tools/sanitize_case.py
    compute-sanitizer --tool racecheck python tools/sanitize_case.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402
from hls_final_visual_odometry_b200 import Engine  # noqa: E402
from hls_final_visual_odometry_b200.synth import make_sequence  # noqa: E402
from oracle.oracle_lib import Oracle, matches_digest  # noqa: E402


def main():
    ge.smoke()
    w, h = 1241, 376
    rng = np.random.default_rng(500)
    seq = make_sequence(w, h, 2, seed=1).astype(np.int16)
    seq = np.clip(seq + rng.integers(-2, 3, seq.shape, dtype=np.int16), 0, 255).astype(np.uint8)
    orc = Oracle(w, h)
    eng = Engine(w, h, max_frames=2, max_pairs=1)
    eng.batch_extract_host(0, seq)
    eng.batch_match([0], [1], max_features=2, bw=50.0, bh=50.0)
    fp = orc.compute_features(seq[0])[:2]
    fc = orc.compute_features(seq[1])[:2]
    want, _ = orc.matching(fp[0], fp[1], fc[0], fc[1], cap=1 << 18)
    got = eng.batch_get_matches(0, 0)
    assert matches_digest(got) == matches_digest(want), "noisy 1241x376 pair differs from the oracle"
    gb = eng.batch_get_matches(0, 1)
    assert matches_digest(gb) == matches_digest(orc.bucket_features(want, 2, 50.0, 50.0))
    eng.close()
    print(f"sanitize_case ok: noisy 1241x376 pair, {len(want)} raw matches, bit-exact vs oracle")


if __name__ == "__main__":
    main()
