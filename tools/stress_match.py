"""Developer stress test for K2's pruned / hinted search: random frame sizes, flows (up to the limits of the matching
window, where the bin ranges of the two directions stop being symmetric), sensor-noise levels and texture densities;
the pruned two-launch search must give exactly the matches of the search that evaluates every SAD (match_exhaustive).

    python tools/stress_match.py [cases] [seed]
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

from hls_final_visual_odometry_b200 import Engine  # noqa: E402
from hls_final_visual_odometry_b200.synth import frame_from_canvas, make_canvas  # noqa: E402


def run_case(rng, idx):
    w = int(rng.integers(200, 1400))
    h = int(rng.integers(100, 520))
    n = 5
    blur = int(rng.choice([2, 4, 6, 10, 14]))
    contrast = int(rng.choice([10, 25, 45, 80]))
    canvas = make_canvas(w, h, seed=int(rng.integers(1 << 30)), blur_passes=blur, contrast=contrast)
    big = rng.random() < 0.4
    frames = []
    x, y = 16, 8
    for f in range(n):
        frames.append(frame_from_canvas(canvas, w, h, 1, x0=x, y0=y, flow=(0, 0)))
        lim_u, lim_v = (200, 150) if big else (12, 6)
        x += int(rng.integers(-lim_u, lim_u + 1))
        y += int(rng.integers(-lim_v, lim_v + 1))
    seq = np.stack(frames)
    amps = rng.choice([0, 0, 1, 2, 4, 10, 30], size=n)
    for f, a in enumerate(amps):
        if a:
            seq[f] = np.clip(seq[f].astype(np.int16) + rng.integers(-a, a + 1, seq[f].shape, dtype=np.int16), 0, 255)
    prev = np.arange(n - 1, dtype=np.int32)
    res = []
    for exhaustive in (False, True):
        eng = Engine(w, h, max_frames=n, max_pairs=n - 1, match_exhaustive=exhaustive)
        try:
            eng.batch_extract_host(0, seq)
            eng.batch_match(prev, prev + 1, max_features=0)
            res.append([eng.batch_get_matches(i, 0).tobytes() for i in range(n - 1)])
            cnt = [len(eng.batch_get_matches(i, 0)) for i in range(n - 1)]
        finally:
            eng.close()
    ok = res[0] == res[1]
    print(f"case {idx}: {w}x{h} blur {blur} contrast {contrast} big_motion {big} noise {list(map(int, amps))} matches {cnt} "
          f"{'ok' if ok else 'MISMATCH'}", flush=True)
    return ok


def main():
    cases = int(sys.argv[1]) if len(sys.argv) > 1 else 40
    seed = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    rng = np.random.default_rng(seed)
    bad = sum(not run_case(rng, i) for i in range(cases))
    print(f"{cases - bad}/{cases} cases identical")
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
