#!/usr/bin/env python
"""BASELINE.json config 4: stereo (left/right + prev/curr quad) circular matching, 1241x376, batched on one B200.
EXTENSION of the reference (its matcher is flow-only) -- reported separately from the pinned mono
headline of bench.py.  One JSON line.

A step = B consecutive stereo frames of one synthetic rig sequence: 2B extractions (left + right), B quad matchings
(four searches each + cycle/emit), optional sub-pixel refinement, B bucketings.  Frames go host -> device every step
(pinned memory), counts and bucketed lists come back; kernel times from CUDA events on the engine stream.
CPU leg: the oracle port of the same stages (kind "port": the reference has no stereo matcher to time), one thread.

    python tools/bench_stereo.py [--batch 64] [--steps 20] [--warmup 3] [--refinement 2] [--cpu-pairs 2]
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--refinement", type=int, default=2)
    ap.add_argument("--disparity", type=int, default=14)
    ap.add_argument("--cpu-pairs", type=int, default=2)
    args = ap.parse_args()
    from hls_final_visual_odometry_b200 import Engine, pinned_empty
    from hls_final_visual_odometry_b200.synth import make_stereo_sequence
    W, H, B = 1241, 376, args.batch
    left, right = make_stereo_sequence(W, H, B + 1, seed=1, disparity=args.disparity)
    pitch = (W + 15) & ~15  # the engine's slot pitch: one linear H2D copy per stack instead of a strided one
    hl, hr = pinned_empty((B + 1, H, pitch), np.uint8), pinned_empty((B + 1, H, pitch), np.uint8)
    hl[:, :, :W], hr[:, :, :W] = left, right
    eng = Engine(W, H, max_frames=2 * (B + 1), max_pairs=B, stereo=True)
    k = np.arange(1, B + 1, dtype=np.int32)
    s1p, s2p, s1c, s2c = k - 1, B + 1 + k - 1, k, B + 1 + k
    cap = 4096
    counts = pinned_empty((2 * B,), np.int32)
    lists = pinned_empty((B, cap), eng.get_matches().dtype)

    def step():
        eng.batch_extract_host(0, hl.ctypes.data, B + 1, H * pitch, pitch)
        eng.batch_extract_host(B + 1, hr.ctypes.data, B + 1, H * pitch, pitch)
        eng.batch_match_quad(s1p, s2p, s1c, s2c, 2, args.refinement, 2, 50.0, 50.0)
        eng.batch_download_async(B, 1, counts, lists, cap, cap)

    for _ in range(max(args.warmup, 3)):
        step()
    eng.synchronize()
    raw_n, bk_n = eng.batch_counts(B)
    eng.profile_enable(True)
    launches0 = eng.launch_count()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    eng.synchronize()
    wall = time.perf_counter() - t0
    prof = eng.profile_read_ext()
    launches = eng.launch_count() - launches0
    feats = np.mean([eng.slot_feature_count(s) for s in (0, 1, B + 1, B + 2)])
    out = {"workload": f"configs[3]: stereo quad matching {W}x{H}, batch of {B} stereo frame pairs per step, 1 B200",
           "extension": "quad matching + sub-pixel refinement are absent from the reference (parity unpinned; oracle = oracle/vo_oracle.c EXTENSION)",
           "quads_per_s": args.steps * B / wall, "ms_per_step": 1e3 * wall / args.steps, "batch": B, "steps": args.steps,
           "refinement": args.refinement, "features_per_frame": float(feats), "raw_quads_per_pair": float(raw_n.mean()),
           "bucketed_per_pair": float(bk_n.mean()), "gpu_launches": int(launches),
           "h2d_bytes_per_step": int(hl.nbytes + hr.nbytes), "d2h_bytes_per_step": int(counts.nbytes + lists.nbytes),
           "kernel_ms_per_step": {k2: ms / args.steps for k2, (ms, calls) in prof.items()}}
    eng.close()
    # ---- CPU leg: the oracle port of the same stages, single thread ------------------------------------------------
    if args.cpu_pairs > 0:
        from oracle.oracle_lib import Oracle
        orc = Oracle(W, H)
        n = args.cpu_pairs
        t0 = time.perf_counter()
        fl = [orc.compute_features(left[i])[:2] for i in range(n + 1)]
        fr = [orc.compute_features(right[i])[:2] for i in range(n + 1)]
        for i in range(1, n + 1):
            m = orc.quad_matching(fl[i - 1], fr[i - 1], fl[i], fr[i], 2)
            if args.refinement == 2:
                m = orc.refine(m, left[i - 1], left[i], right[i - 1], right[i])
            orc.bucket_features(m, 2, 50.0, 50.0)
        dt = time.perf_counter() - t0
        out["cpu_baseline"] = {"value": n / dt, "unit": "quads/s", "cores": 1, "kind": "port",
                               "sample": f"{n} stereo frame pairs ({2 * (n + 1)} extractions), oracle port, one thread"}
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
