#!/usr/bin/env python
"""BASELINE.json config 4: stereo (left/right + prev/curr quad) circular matching, 1241x376, batched on one B200.
EXTENSION of the reference (its matcher is flow-only) -- reported separately from the pinned mono
headline of bench.py.  One JSON line.

A step = B new consecutive stereo frames of one synthetic rig sequence: 2B extractions (left + right), B quad matchings
(four searches each + cycle/emit), optional sub-pixel refinement, B bucketings.  Frames go host -> device every step
(pinned memory), counts and bucketed lists come back, software-pipelined to depth 2 like bench.py's e2e leg; kernel
times from CUDA events on the engine stream.
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
    ap.add_argument("--no-profile", action="store_true", help="no per-kernel CUDA events in the timed loop")
    args = ap.parse_args()
    from hls_final_visual_odometry_b200 import Engine, pinned_empty
    from hls_final_visual_odometry_b200.synth import make_stereo_sequence
    W, H, B = 1241, 376, args.batch
    pool = 4 * B                                            # frames of the rig sequence cycled through (122 MB per camera at B = 64: > L2)
    left, right = make_stereo_sequence(W, H, pool, seed=1, disparity=args.disparity)
    pitch = (W + 15) & ~15  # the engine's slot pitch: one linear H2D copy per stack instead of a strided one
    # host frames in step-sized blocks [B left frames, B right frames], so one call uploads and extracts a whole step
    hs = pinned_empty((pool // B, 2 * B, H, pitch), np.uint8)
    hs[:, :B, :, :W] = left.reshape(pool // B, B, H, W)
    hs[:, B:, :, :W] = right.reshape(pool // B, B, H, W)
    # slots: step s fills bank (s % 3) = slots [b*2B, b*2B + 2B): B left frames then B right frames.  Three banks, because
    # step s still reads bank (s-1) % 3 (previous frame of its first item; the refinement reads its image): the H2D copy
    # of step s+1 (upload stream) then touches nothing step s reads and overlaps all of its kernels; the D2H of step s
    # (download stream) overlaps the kernels of step s+1
    eng = Engine(W, H, max_frames=6 * B, max_pairs=B, stereo=True)
    cap = 512
    counts = [pinned_empty((2 * B,), np.int32) for _ in range(2)]
    lists = [pinned_empty((B, cap), eng.get_matches().dtype) for _ in range(2)]
    kept = [0]

    def consume(buf):
        eng.fence_wait(buf)                                 # that step's counts + bucketed lists are on the host
        kept[0] += int(counts[buf][B:].sum())

    def step(sidx, first):
        bank, buf = sidx % 3, sidx % 2
        lc = bank * 2 * B + np.arange(B)                     # left current slots; the right ones are B further
        lp = lc - 1
        lp[0] = ((sidx - 1) % 3) * 2 * B + B - 1             # previous of the first frame: last left frame of the previous bank
        blk = (sidx * B % pool) // B
        eng.batch_extract_host(bank * 2 * B, hs.ctypes.data + blk * 2 * B * H * pitch, 2 * B, H * pitch, pitch)
        eng.batch_match_quad(lp, lp + B, lc, lc + B, 2, args.refinement, 2, 50.0, 50.0)
        eng.batch_download_async(B, 1, counts[buf], lists[buf], cap, cap)
        eng.fence_record(buf)
        if sidx > first:
            consume(1 - buf)

    # fill the bank before the first one, so the very first pair has a resident previous frame
    eng.batch_extract_host(2 * 2 * B, hs.ctypes.data + (pool // B - 1) * 2 * B * H * pitch, 2 * B, H * pitch, pitch)
    nw = max(args.warmup, 3)
    for i in range(nw):
        step(i, 0)
    consume((nw - 1) % 2)
    eng.synchronize()
    raw_n, bk_n = eng.batch_counts(B)
    eng.profile_enable(not args.no_profile)
    launches0 = eng.launch_count()
    t0 = time.perf_counter()
    for i in range(nw, nw + args.steps):
        step(i, nw)
    consume((nw + args.steps - 1) % 2)
    eng.synchronize()
    wall = time.perf_counter() - t0
    prof = eng.profile_read_ext()
    launches = eng.launch_count() - launches0
    feats = np.mean([eng.slot_feature_count(s) for s in (0, 1, B, B + 1)])
    out = {"workload": f"configs[3]: stereo quad matching {W}x{H}, batch of {B} consecutive stereo frames per step, 1 B200",
           "extension": "quad matching + sub-pixel refinement are absent from the reference (parity unpinned; oracle = oracle/vo_oracle.c EXTENSION)",
           "quads_per_s": args.steps * B / wall, "ms_per_step": 1e3 * wall / args.steps, "batch": B, "steps": args.steps,
           "refinement": args.refinement, "features_per_frame": float(feats), "raw_quads_per_pair": float(raw_n.mean()),
           "bucketed_per_pair": float(bk_n.mean()), "bucketed_matches_read_back": kept[0], "gpu_launches": int(launches),
           "pipeline": "depth 2: H2D(s+1) | kernels(s) | D2H(s-1); every step's 2B frames go host->device and its counts + bucketed lists come back inside the timed region",
           "l2": f"pool of {pool} stereo frames ({2 * pool * H * pitch / 1e6:.0f} MB) cycled: inputs larger than L2",
           "h2d_bytes_per_step": int(2 * B * H * pitch), "d2h_bytes_per_step": int(counts[0].nbytes + lists[0].nbytes),
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
