#!/usr/bin/env python
"""BASELINE.json config 3: a 1000-frame synthetic KITTI-shaped mono sequence through the FULL process() loop
(GPU extraction + circular matching + bucketing, host removeOutliers + egomotion) on one B200, next to the
reference's own process() chain on the host cores.  Not the headline bench (bench.py); writes one JSON line.

    python tools/bench_process.py [--frames 1000] [--batch 128] [--threads 0] [--ref-frames 6]
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
    ap.add_argument("--frames", type=int, default=1000)
    ap.add_argument("--batch", type=int, default=128)
    ap.add_argument("--threads", type=int, default=0)
    ap.add_argument("--pool", type=int, default=24, help="distinct rendered frames, played forward and backward")
    ap.add_argument("--ref-frames", type=int, default=6)
    args = ap.parse_args()
    from hls_final_visual_odometry_b200 import Pipeline, pinned_empty
    from hls_final_visual_odometry_b200.synth import KITTI_CALIB, KITTI_HEIGHT, make_corridor_sequence
    W, H = 1241, 376
    frames, _ = make_corridor_sequence(W, H, n_frames=args.pool, seed=0)
    # ping-pong playback: every consecutive pair is a real camera motion (forward or backward)
    order = list(range(args.pool)) + list(range(args.pool - 2, 0, -1))
    idx = [order[i % len(order)] for i in range(args.frames)]
    host = pinned_empty((args.frames, H, W), np.uint8)
    for i, j in enumerate(idx):
        host[i] = frames[j]
    pipe = Pipeline(W, H, batch=args.batch, cam_height=KITTI_HEIGHT, pitch=0.0, host_threads=args.threads, **KITTI_CALIB)
    pipe.process(host[: min(args.batch, 8)])                         # warm-up (CUDA context, pinned buffers)
    pipe.close()
    pipe = Pipeline(W, H, batch=args.batch, cam_height=KITTI_HEIGHT, pitch=0.0, host_threads=args.threads, **KITTI_CALIB)
    oks, nms, nis = [], [], []
    t0 = time.perf_counter()
    for s in range(0, args.frames, args.batch):
        ok, tr, nm, ni = pipe.process(host[s:s + args.batch])
        oks.append(ok); nms.append(nm); nis.append(ni)
    wall = time.perf_counter() - t0
    stages = pipe.stage_seconds()
    pipe.close()
    ok, nm, ni = np.concatenate(oks), np.concatenate(nms), np.concatenate(nis)
    out = {"workload": f"configs[2]: {args.frames}-frame rendered corridor sequence {W}x{H}, full process() loop",
           "frames": args.frames, "seconds": wall, "frames_per_s": args.frames / wall,
           "process_ok_frames": int(ok.sum()), "mean_bucketed_matches": float(nm[1:].mean()),
           "mean_inliers": float(ni[ok == 1].mean()) if ok.any() else 0.0,
           "host_threads": args.threads or os.cpu_count(), "batch": args.batch,
           "stage_seconds": stages, "stage_ms_per_frame": {k: 1e3 * v / args.frames for k, v in stages.items()}}
    # ---- the reference's own chain, one thread (it has no threads), on a few frames -----------------------------
    try:
        from oracle.oracle_lib import Oracle
        from oracle.ref_lib import RefLib, have_ref, mono_motion_seq
        if have_ref(f"{W}x{H}") and args.ref_frames > 1:
            ref, orc = RefLib(f"{W}x{H}"), Oracle(W, H)
            m = ref.matcher()
            m.push_back(frames[0])
            t0 = time.perf_counter()
            lists = []
            for k in range(1, args.ref_frames):
                m.push_back(frames[k]); m.match_features()
                lists.append(orc.bucket_features(m.get_matches(), 2, 50.0, 50.0))
            mono_motion_seq(ref, lists, height=KITTI_HEIGHT, pitch=0.0, **KITTI_CALIB)
            dt = time.perf_counter() - t0
            out["reference_frames_per_s_per_core"] = (args.ref_frames - 1) / dt
            out["reference_note"] = ("compiled reference (oracle/_ref): pushBack + matchFeatures (incl. removeOutliers) + estimateMotion, "
                                     "bucketing via the oracle port; single thread, the reference has none")
            out["host_cores"] = os.cpu_count()
    except Exception as e:  # the reference leg is a report, not a gate
        out["reference_error"] = repr(e)
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
