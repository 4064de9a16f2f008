#!/usr/bin/env python
"""ONE sequence over several GPUs by contiguous frame ranges (second half; the pipeline shape of the
reference's FPGA host, /root/reference/hls_opencl/host/main2.cpp:395-519, spread over devices instead of over queue slots).

Rank r extracts the frames [f0, f1) of sharding.split_frame_ranges and matches its pairs f0+1 .. f1-1; the boundary frame is
extracted by two ranks, nothing is exchanged on the data path.  The per-pair results (counts + sha256 of the raw and the
bucketed lists) are gathered on rank 0, which -- with --verify -- also runs the whole sequence alone and compares: the split
must reproduce it bit for bit.

    python tools/split_sequence.py --frames 257 --verify                                   # one GPU, 1 range
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29511 \
        tools/split_sequence.py --frames 1025 --verify
"""
import argparse
import hashlib
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

from hls_final_visual_odometry_b200 import Engine, sharding  # noqa: E402
from hls_final_visual_odometry_b200.synth import make_sequence  # noqa: E402


def digest(m: np.ndarray) -> str:
    return hashlib.sha256(np.ascontiguousarray(m).tobytes()).hexdigest()[:16]


def run_range(seq: np.ndarray, f0: int, f1: int, device: int = 0, batch: int = 64, want_lists: bool = True):
    """extract frames [f0, f1) and match the pairs f0+1 .. f1-1 in batches; returns {pair: (n_raw, n_bucketed, raw digest,
    bucketed digest)} and the seconds spent between the first upload and the last result"""
    out = {}
    if f1 - f0 < 2:
        return out, 0.0
    h, w = seq.shape[1:]
    eng = Engine(w, h, max_frames=batch + 1, max_pairs=batch, device=device)
    try:
        t0 = time.perf_counter()
        f = f0
        while f < f1 - 1:  # frames f .. f+n: the first one is the previous frame of the batch's first pair
            n = min(batch, f1 - 1 - f)
            eng.batch_extract_host(0, seq[f:f + n + 1])
            prev = np.arange(n, dtype=np.int32)
            eng.batch_match(prev, prev + 1, 2, 50.0, 50.0)
            raw_n, bk_n = eng.batch_counts(n)
            for i in range(n):
                if want_lists:
                    out[f + 1 + i] = (int(raw_n[i]), int(bk_n[i]), digest(eng.batch_get_matches(i, 0)), digest(eng.batch_get_matches(i, 1)))
                else:
                    out[f + 1 + i] = (int(raw_n[i]), int(bk_n[i]), "", "")
            f += n
        eng.synchronize()
        return out, time.perf_counter() - t0
    finally:
        eng.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=257)
    ap.add_argument("--width", type=int, default=1241)
    ap.add_argument("--height", type=int, default=376)
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--verify", action="store_true", help="rank 0 also runs the whole sequence alone and compares every pair")
    a = ap.parse_args()
    world, rank, local = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl")
    seq = make_sequence(a.width, a.height, a.frames, seed=1)
    f0, f1 = sharding.split_frame_ranges(a.frames, world, rank)
    if dist is not None:
        dist.barrier()
    mine, secs = run_range(seq, f0, f1, device=local, batch=a.batch, want_lists=a.verify)
    secs = sharding.reduce_max([secs], dist, device=f"cuda:{local}" if dist is not None else None)[0]
    parts = [mine]
    if dist is not None:
        parts = [None] * world
        dist.all_gather_object(parts, mine)
    if rank == 0:
        merged = {}
        for p in parts:
            assert not (set(p) & set(merged)), "a pair was matched by two ranks"
            merged.update(p)
        assert sorted(merged) == list(range(1, a.frames)), "pairs missing"
        line = {"workload": f"one {a.width}x{a.height} sequence of {a.frames} frames split into {world} contiguous frame ranges",
                "n_gpus": world, "pairs": a.frames - 1, "seconds_max_over_ranks": secs, "pairs_per_s": (a.frames - 1) / secs if secs else None,
                "frames_extracted_twice": world - 1}
        if a.verify:
            whole, _ = run_range(seq, 0, a.frames, device=local, batch=a.batch)
            bad = [p for p in whole if whole[p] != merged[p]]
            line["verified_bit_identical_to_one_range"] = not bad
            assert not bad, f"pairs {bad[:5]} differ from the unsplit run"
        print(json.dumps(line))
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
