#!/usr/bin/env python
"""Bare host->device ceiling of this box for bench.py's end-to-end leg: every rank streams its own pinned
buffer of the bench's step size (60 MB = 128 frames of 1248 x 376) to its GPU back to back, no kernels, and rank 0 prints the
aggregate GB/s (bytes of all ranks / max over ranks of the CUDA-event time).  bench.py's e2e at N GPUs needs
N x h2d_bytes_per_step / ms_per_step of it.

    python tools/h2d_ceiling.py                                   (one GPU)
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29700 tools/h2d_ceiling.py
"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    nbytes = int(os.environ.get("H2D_BYTES", 128 * 1248 * 376))
    reps = int(os.environ.get("H2D_REPS", 200))
    out = {}
    for kind in ("pinned", "pinned_2streams"):
        host = [torch.empty(nbytes, dtype=torch.uint8).pin_memory() for _ in range(3)]
        dev = [torch.empty(nbytes, dtype=torch.uint8, device="cuda") for _ in range(3)]
        streams = [torch.cuda.Stream() for _ in range(2 if kind.endswith("2streams") else 1)]
        for i in range(6):
            with torch.cuda.stream(streams[i % len(streams)]):
                dev[i % 3].copy_(host[i % 3], non_blocking=True)
        torch.cuda.synchronize()
        if dist:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for s in streams:
            s.wait_event(e0)
        for i in range(reps):
            with torch.cuda.stream(streams[i % len(streams)]):
                dev[i % 3].copy_(host[i % 3], non_blocking=True)
        for s in streams:
            torch.cuda.current_stream().wait_stream(s)
        e1.record()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1)], device="cuda")
        if dist:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        out[kind] = {"ms_max_over_ranks": float(ms.item()), "aggregate_gb_s": world * reps * nbytes / (float(ms.item()) * 1e-3) / 1e9,
                     "per_rank_gb_s": reps * nbytes / (float(ms.item()) * 1e-3) / 1e9}
    if rank == 0:
        print(json.dumps({"what": "bare pinned host->device copies, one process per GPU, no kernels", "n_gpus": world, "bytes_per_copy": nbytes,
                          "copies_per_rank": reps, "host_cpus": len(os.sched_getaffinity(0)), **out}))
    if dist:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
