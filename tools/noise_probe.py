"""Developer probe: how does the matcher's exact lower-bound pruning hold up when consecutive frames are NOT bit-identical
translations?  Adds independent sensor noise (uniform +-amp grey levels) to every frame of the synthetic sequence and times
extract + match + bucket; parity against the oracle is checked on the first pair.

    python tools/noise_probe.py [amp ...]
"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

from hls_final_visual_odometry_b200 import Engine  # noqa: E402
from hls_final_visual_odometry_b200.synth import make_sequence  # noqa: E402

W, H, B = 1241, 376, 64
amps = [int(a) for a in sys.argv[1:]] or [0, 1, 2, 4, 8]
budget = int(os.environ.get("VO_SEED_GOOD", "0"))
base = make_sequence(W, H, B + 1, seed=1).astype(np.int16)
for amp in amps:
    rng = np.random.default_rng(100 + amp)
    seq = np.clip(base + (rng.integers(-amp, amp + 1, base.shape, dtype=np.int16) if amp else 0), 0, 255).astype(np.uint8)
    eng = Engine(W, H, max_frames=B + 1, max_pairs=B, match_seed_good=budget,
                 match_threads=int(os.environ.get("VO_MATCH_THREADS", "0")))
    prev = np.arange(B, dtype=np.int32)
    eng.batch_extract_host(0, seq)
    eng.batch_match(prev, prev + 1, 2, 50.0, 50.0)
    raw, bk = eng.batch_counts(B)
    if os.environ.get("VO_PROBE_PARITY"):
        from oracle.oracle_lib import Oracle, matches_digest
        orc = Oracle(W, H)
        f0, f1 = orc.compute_features(seq[0])[:2], orc.compute_features(seq[1])[:2]
        want, _ = orc.matching(*f0, *f1)
        assert matches_digest(eng.batch_get_matches(0, 0)) == matches_digest(want), "parity lost"
    eng.profile_enable(True)
    for _ in range(int(os.environ.get("VO_PROBE_ITERS", "5"))):
        eng.batch_extract_host(0, seq)
        eng.batch_match(prev, prev + 1, 2, 50.0, 50.0)
    eng.synchronize()
    prof = eng.profile_read()
    ms = {k: v[0] / max(v[1], 1) for k, v in prof.items()}
    print(f"noise +-{amp}: features {eng.slot_feature_count(1)}, raw matches {raw.mean():.0f}; per {B} pairs: "
          f"K1 {ms['extract']:.3f} K2 {ms['match']:.3f} K3 {ms['emit']:.3f} K4 {ms['bucket']:.3f} ms "
          f"-> {B / (sum(ms.values()) * 1e-3):.0f} pairs/s kernel rate", flush=True)
    eng.close()
