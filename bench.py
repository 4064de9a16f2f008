#!/usr/bin/env python
"""bench.py -- frame-pairs/s of the VO front-end hot path (extract -> circular match -> bucket).

Contract (see the task statement):  python bench.py --gpus N --steps K --warmup W  prints ONE JSON
line from rank 0.  For N > 1 it is launched by torchrun, one rank per GPU; ranks shard INDEPENDENT
synthetic sequences (no data-path collective, "scaling": "weak").

A step = one batch of B consecutive new frames of a synthetic 1241x376 sequence: B extractions
(one new frame per pair, the previous frame's features are reused from HBM -- the steady state
of Matcher::pushBack, /root/reference/src/matcher.cpp:51-91), B circular matchings and B
bucketings.  `value` times that with the frames already resident in HBM; `e2e` times the same step
through the C ABI from pinned HOST frames (H2D inside) with the bucketed match lists and counts
copied back (D2H inside).

--impl reference times the reference's own CPU code for the same path (oracle/_ref, compiled from
/root/reference/src in place) on the host cores.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

W_DEFAULT, H_DEFAULT = 1241, 376
UNIT = "frame-pairs/s"
NOISE_DEFAULT = 2   # grey levels of uniform sensor noise added independently to every frame of the headline workload


def metric_name(width, height):
    return f"frame_pairs_per_s_{width}x{height}"


def workload_name(args):
    """One string for both arms (the driver compares them)."""
    cfg = "configs[4]: 8 independent sequences" if args.config == 5 else "configs[1]"
    return (f"{cfg}: {args.width}x{args.height} frame pairs, extract+match+bucket, synthetic sequence with uniform "
            f"+-{args.noise} grey levels of sensor noise per frame")


def noisy_frame(frame, amp, rng):
    """frame + independent uniform noise in [-amp, amp], clipped (both arms and every leg use this)."""
    if amp <= 0:
        return frame
    return np.clip(frame.astype(np.int16) + rng.integers(-amp, amp + 1, frame.shape, dtype=np.int16), 0, 255).astype(np.uint8)


def source_hash():
    """sha256 over the DEVICE sources (csrc/*.cu, *.cuh, *.h) + ABI header: ties profiles/traffic.json to the kernels it was
    captured from.  The host stages (csrc/*.cpp: Delaunay filter, egomotion, pipeline) do not move a byte of device traffic and
    are left out, so that work on them does not orphan a capture of unchanged kernels."""
    import hashlib
    h = hashlib.sha256()
    base = os.path.join(ROOT, "hls_final_visual_odometry_b200", "csrc")
    files = sorted(os.path.join(base, f) for f in os.listdir(base) if f.endswith((".cu", ".cuh", ".h")))
    for f in files + [os.path.join(ROOT, "include", "vo_b200.h")]:
        h.update(os.path.basename(f).encode())
        h.update(open(f, "rb").read())
    return h.hexdigest()[:16]


# --------------------------------------------------------------------------------------------
# CPU legs (the only place outside tests/ and smoke() that touches oracle/)
# --------------------------------------------------------------------------------------------
def _cpu_worker(args):
    """Process n consecutive pairs of one synthetic sequence with the reference code; returns
    (pairs, seconds, kind)."""
    width, height, seed, n_pairs, noise = args
    from hls_final_visual_odometry_b200.synth import make_sequence
    from oracle.oracle_lib import Oracle
    from oracle.ref_lib import RefLib, have_ref

    seq = make_sequence(width, height, n_pairs + 1, seed=seed)
    rng = np.random.default_rng(9000 + seed)
    seq = np.stack([noisy_frame(f, noise, rng) for f in seq])
    tag = f"{width}x{height}"
    orc = Oracle(width, height)
    if have_ref(tag):
        ref = RefLib(tag)
        kind = "reference"
        extract = lambda img: ref.compute_features(img)  # noqa: E731
        match = lambda a, b: ref.matching(a[0], a[1], b[0], b[1])  # noqa: E731
    else:
        kind = "port"
        extract = lambda img: orc.compute_features(img)[:2]  # noqa: E731
        match = lambda a, b: orc.matching(a[0], a[1], b[0], b[1])[0]  # noqa: E731
    prev = extract(seq[0])  # untimed: steady state has the previous frame's features already
    t0 = time.perf_counter()
    for i in range(1, n_pairs + 1):
        cur = extract(seq[i])
        raw = match(prev, cur)
        # the reference's bucketFeatures keeps its buckets in a fixed [126][256] stack array
        # (src/matcher.cpp:154-155) that overflows at this image size, so the (sub-millisecond)
        # bucketing stage of the CPU leg runs through the oracle port
        orc.bucket_features(raw, 2, 50.0, 50.0)
        prev = cur
    dt = time.perf_counter() - t0
    # untimed: how many 32-byte SADs the reference's exhaustive search evaluates per pair on these frames (the unit of the
    # integer-pipe roofline); the oracle port counts them
    fa, fb = orc.compute_features(seq[0])[:2], orc.compute_features(seq[1])[:2]
    _, st = orc.matching(fa[0], fa[1], fb[0], fb[1], cap=1 << 19)
    return n_pairs, dt, kind, int(st[0])


def cpu_throughput(width, height, pairs_per_worker, workers, noise):
    """Aggregate pairs/s of `workers` independent single-threaded reference processes."""
    import multiprocessing as mp
    ctx = mp.get_context("fork")
    jobs = [(width, height, 100 + i, pairs_per_worker, noise) for i in range(workers)]
    t0 = time.perf_counter()
    with ctx.Pool(workers) as pool:
        res = pool.map(_cpu_worker, jobs)
    wall = time.perf_counter() - t0
    slowest = max(r[1] for r in res)
    total = sum(r[0] for r in res)
    return total / slowest, res[0][2], wall, total, float(np.mean([r[3] for r in res]))


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    cores = os.cpu_count() or 1
    workers = max(1, min(cores, 64))
    pairs_per_worker = 2
    vals = []
    kind = "reference"
    for s in range(args.warmup + args.steps):
        v, kind, _, total, _ = cpu_throughput(args.width, args.height, pairs_per_worker, workers, args.noise)
        if s >= args.warmup:
            vals.append(v)
    value = float(np.mean(vals)) if vals else 0.0
    sample = (f"{workers} single-threaded processes x {pairs_per_worker} consecutive pairs per step of the "
              f"{args.width}x{args.height} synthetic sequence (+-{args.noise} grey levels of sensor noise): myComputeFeatures + myMatching "
              f"({'oracle/_ref, reference sources compiled in place' if kind == 'reference' else 'oracle port'})"
              f" + bucketing (oracle port)")
    line = {
        "impl": "reference", "metric": metric_name(args.width, args.height), "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * workers * pairs_per_worker / value if value else None,
        "higher_is_better": True, "scaling": "strong" if args.config == 5 else "weak", "vs_baseline": None, "dtype": "u8",
        "data": "synthetic",
        "config": {"workload": workload_name(args), "pairs_per_step": workers * pairs_per_worker},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": workers, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)
    return 0


# --------------------------------------------------------------------------------------------
# clocks
# --------------------------------------------------------------------------------------------
class ClockSampler:
    """SM clock + throttle reasons sampled DURING the timed region: NVML polled from a thread every
    few ms (nvidia-smi's own loop is too coarse for a sub-second region), nvidia-smi as fallback."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int, period_s: float = 0.004):
        self.index, self.period = index, period_s
        self.sm, self.reasons, self.mx = [], set(), None
        self.stop_flag = threading.Event()
        self.thread = None
        self.how = None

    def _nvml_loop(self, nv, h):
        bits = {}
        for nm, attr in (("hw_slowdown", "nvmlClocksEventReasonHwSlowdown"), ("hw_thermal_slowdown", "nvmlClocksEventReasonHwThermalSlowdown"),
                         ("sw_thermal_slowdown", "nvmlClocksEventReasonSwThermalSlowdown"), ("sw_power_cap", "nvmlClocksEventReasonSwPowerCap")):
            alt = attr.replace("ClocksEventReason", "ClocksThrottleReason")
            bits[nm] = getattr(nv, attr, getattr(nv, alt, 0))
        get_reasons = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or nv.nvmlDeviceGetCurrentClocksThrottleReasons
        while not self.stop_flag.is_set():
            try:
                self.sm.append(float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)))
                r = get_reasons(h)
                for nm, b in bits.items():
                    if b and (r & b):
                        self.reasons.add(nm)
            except Exception:
                pass
            time.sleep(self.period)

    def _smi_loop(self):
        proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "20", "-i",
                                 str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        import select
        while not self.stop_flag.is_set():
            if not select.select([proc.stdout], [], [], 0.05)[0]:
                continue
            ln = proc.stdout.readline()
            parts = [p.strip() for p in ln.split(",")]
            if len(parts) < 6:
                continue
            try:
                self.sm.append(float(parts[0])); self.mx = max(self.mx or 0.0, float(parts[1]))
            except ValueError:
                continue
            for nm, val in zip(names, parts[2:6]):
                if val.lower().startswith("active"):
                    self.reasons.add(nm)
        proc.terminate()

    def start(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            idx = self.index
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            if vis:
                try:
                    idx = int(vis.split(",")[self.index])
                except (ValueError, IndexError):
                    pass
            h = nv.nvmlDeviceGetHandleByIndex(idx)
            self.mx = float(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM))
            self.how = "nvml"
            self.thread = threading.Thread(target=self._nvml_loop, args=(nv, h), daemon=True)
        except Exception:
            self.how = "nvidia-smi"
            self.thread = threading.Thread(target=self._smi_loop, daemon=True)
        self.thread.start()

    def stop(self):
        self.stop_flag.set()
        if self.thread:
            self.thread.join(timeout=5)
        if not self.sm:
            return {"sm_mhz": None, "sm_max_mhz": self.mx, "reasons": ["no clock samples"], "samples": 0, "how": self.how}
        return {"sm_mhz": float(np.median(self.sm)), "sm_max_mhz": self.mx, "reasons": sorted(self.reasons),
                "samples": len(self.sm), "how": self.how}


# --------------------------------------------------------------------------------------------
# our arm
# --------------------------------------------------------------------------------------------
def _bind_host_to_gpu(torch, local_rank):
    """Multi-rank runs: pin this process to the CPUs NVML names as local to its GPU BEFORE the pinned frame pool is
    allocated and first touched, so that the pool lands on the GPU's NUMA node (the e2e leg streams 60 MB per step and rank
    from it).  Returns a short description for the JSON line, or None if NVML cannot tell."""
    try:
        import pynvml as nv
        nv.nvmlInit()
        try:
            uuid = str(torch.cuda.get_device_properties(local_rank).uuid)
            h = nv.nvmlDeviceGetHandleByUUID(("GPU-" + uuid) if not uuid.startswith("GPU-") else uuid)
        except Exception:
            h = nv.nvmlDeviceGetHandleByIndex(local_rank)
        nv.nvmlDeviceSetCpuAffinity(h)
        cpus = sorted(os.sched_getaffinity(0))
        return f"{len(cpus)} cpus local to the GPU ({cpus[0]}..{cpus[-1]}; nvmlDeviceSetCpuAffinity)"
    except Exception as exc:  # not fatal: the run is valid without it
        return f"unbound ({type(exc).__name__})"


def run_ours(args):
    import torch

    from hls_final_visual_odometry_b200 import Engine, pinned_empty
    from hls_final_visual_odometry_b200.api import P_MATCH_DTYPE
    from hls_final_visual_odometry_b200.sharding import assign_sequences, sequence_seed
    from hls_final_visual_odometry_b200.synth import frame_from_canvas, make_canvas

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    host_affinity = _bind_host_to_gpu(torch, local_rank) if world > 1 else None
    dist = None
    if world > 1:
        if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"             # keep stdout to the one JSON line
        import torch.distributed as dist_mod
        dist_mod.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        dist = dist_mod

    W, H = args.width, args.height
    pitch = (W + 15) & ~15
    # ---- which sequences this rank owns, and how many new frames of each make one step ------------------------------
    if args.config == 5:
        # BASELINE configs[4]: a FIXED job of 8 independent sequences (seeds 1..8), sharded 8/4/2/1 per GPU over 1/2/4/8
        # GPUs -- strong scaling; one GPU interleaves all eight
        n_total = args.sequences
        if n_total % world:
            raise SystemExit(f"bench.py --config 5: {n_total} sequences do not split evenly over {world} ranks")
        mine = assign_sequences(n_total, world, rank)
        bs = args.batch // n_total                        # frames per sequence and step (global step = args.batch pairs)
        pairs_per_step = bs * n_total
        scaling = "strong"
    else:
        mine = [rank]                                     # independent sequence per rank, seed 1 + rank: weak scaling
        bs = args.batch
        pairs_per_step = bs * world
        scaling = "weak"
    nq = len(mine)
    B = bs * nq                                           # pairs per step on THIS rank
    pool_frames = max(args.pool // nq, 2 * bs)
    pool_frames = (pool_frames + bs - 1) // bs * bs       # whole steps
    canvases = [make_canvas(W, H, seed=sequence_seed(q)) for q in mine]
    host_pools = [pinned_empty((pool_frames, H, pitch), np.uint8) for _ in mine]

    def fill_pools(amp):
        for qi, q in enumerate(mine):
            host_pools[qi][:] = 0
            rng = np.random.default_rng(500 + q)
            for f in range(pool_frames):
                host_pools[qi][f, :, :W] = noisy_frame(frame_from_canvas(canvases[qi], W, H, f), amp, rng)

    fill_pools(args.noise)
    dev_pools = [torch.from_numpy(hp).cuda() for hp in host_pools]   # HBM-resident copies, TMA-friendly pitch
    torch.cuda.synchronize()
    # the engine launches on THIS torch stream, so torch.cuda.Event timing brackets our kernels
    # (torch's default stream has handle 0, which the C ABI reads as "create your own stream")
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    eng = Engine(W, H, max_frames=2 * B, max_pairs=B, device=local_rank, stream=stream.cuda_stream,
                 match_rows_per_cta=args.rows)
    max_keep = 512
    # double-buffered host results: the download of step s overlaps the kernels of step s+1
    counts = [pinned_empty((2 * B,), np.int32) for _ in range(2)]
    out_lists = [pinned_empty((B, max_keep), P_MATCH_DTYPE) for _ in range(2)]
    consumed = [0]

    def slots_of(step):
        """per sequence: first slot of this step's frames; all pairs of the rank: prev/curr slot arrays.  Sequence qi owns
        the ring of slots [qi * 2 bs, (qi + 1) * 2 bs)."""
        bases, prev, curr = [], [], []
        for qi in range(nq):
            ring0 = qi * 2 * bs
            base = ring0 + (step % 2) * bs
            c = ring0 + (base - ring0 + np.arange(bs)) % (2 * bs)
            bases.append(base); curr.append(c); prev.append(ring0 + (c - ring0 - 1) % (2 * bs))
        return bases, np.concatenate(prev).astype(np.int32), np.concatenate(curr).astype(np.int32)

    def step_device(step):
        bases, prev, curr = slots_of(step)
        f0 = (step * bs) % pool_frames
        for qi in range(nq):
            eng.batch_extract_device(bases[qi], dev_pools[qi].data_ptr() + f0 * H * pitch, bs, H * pitch, pitch)
        eng.batch_match(prev, curr, 2, 50.0, 50.0)

    def consume(buf):
        eng.fence_wait(buf)                                # step's counts + bucketed lists are on the host
        consumed[0] += int(counts[buf][B:].sum())

    def step_e2e(step, first):
        # software pipeline, depth 2 (the reference's FPGA host does the same, hls_opencl/host/main2.cpp:449-518):
        # H2D of this step runs on the upload stream while the previous step's kernels run; its D2H runs on
        # the download stream; the host blocks only on the PREVIOUS step's result.
        bases, prev, curr = slots_of(step)
        f0 = (step * bs) % pool_frames
        buf = step % 2
        for qi in range(nq):
            eng.batch_extract_host(bases[qi], host_pools[qi][f0:f0 + bs])
        eng.batch_match(prev, curr, 2, 50.0, 50.0)
        eng.batch_download_async(B, 1, counts[buf], out_lists[buf], max_keep, max_keep)
        eng.fence_record(buf)
        if step > first:
            consume(1 - buf)

    def barrier():
        if dist is not None:
            dist.barrier()
        eng.synchronize()
        torch.cuda.synchronize()

    def timed(fn, first_step, drain=None, n_steps=None):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record(stream)
        for s in range(n_steps or args.steps):
            fn(first_step + s)
        if drain:
            drain()
        e1.record(stream)
        barrier()
        wall = time.perf_counter() - t0
        ms = e0.elapsed_time(e1)
        if dist is not None:
            t = torch.tensor([ms, wall * 1e3], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms, wall = float(t[0]), float(t[1]) / 1e3
        return ms, wall

    # ---- warm-up (fills the rings so every timed pair has a real previous frame) ------------
    for s in range(max(args.warmup, 3)):
        step_device(s)
    step0 = max(args.warmup, 3)
    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()
    launches0 = eng.launch_count()
    eng.profile_enable(True)
    eng.profile_read(reset=True)
    dev_ms, _ = timed(step_device, step0)
    prof = eng.profile_read(reset=True)
    eng.profile_enable(False)
    launches = eng.launch_count() - launches0
    raw_n, bk_n = eng.batch_counts(B)
    n_feat = [eng.slot_feature_count(s) for s in range(0, 2 * B, max(1, (2 * B) // 8))]
    e2e_first = step0 + args.steps
    for s in range(3):                                     # e2e warm-up (pinned pages, copy engines)
        step_e2e(e2e_first + s, e2e_first)
    consume((e2e_first + 2) % 2)
    e2e_first += 3
    last = e2e_first + args.steps - 1
    e2e_ms, e2e_wall = timed(lambda st: step_e2e(st, e2e_first), e2e_first, drain=lambda: consume(last % 2))
    e2e_ms = max(e2e_ms, e2e_wall * 1e3)                  # host-visible completion is what counts
    clocks = sampler.stop() if sampler else None

    # ---- the same device-resident step at other noise levels (N = 1 only; reported beside the headline, never in it).
    #      Noise-free frames are exact integer translations of one canvas: K2's exact pruning collapses on them (92 % of the
    #      origins have a SAD-0 partner), so that figure is a BEST CASE, not the headline. ---------------------------------
    extra = {}
    if world == 1 and not args.no_extra_legs:
        nfirst = last + 1
        for name, amp in (("best_case_noise_free", 0), ("noise_8", 8)):
            if amp == args.noise:
                continue
            n_leg = min(args.steps, 50)
            fill_pools(amp)
            for qi in range(nq):
                dev_pools[qi].copy_(torch.from_numpy(host_pools[qi]))
            for s in range(3):
                step_device(nfirst + s)
            leg_ms, _ = timed(step_device, nfirst + 3, n_steps=n_leg)
            nraw, _ = eng.batch_counts(B)
            extra[name] = {"value": B * n_leg / (leg_ms / 1e3), "unit": UNIT, "steps": n_leg, "ms_per_step": leg_ms / n_leg,
                           "noise": f"uniform +-{amp} grey levels, independent per frame" if amp else
                                    "none: every frame is an exact integer translation of one canvas",
                           "raw_matches_per_pair": float(np.mean(nraw)),
                           "what": "device-resident step as `value`, same sequence at another sensor-noise level (results stay "
                                   "bit-exact at every level: tests/test_gpu_cases.py::test_matching_under_sensor_noise_vs_oracle)"}
            nfirst += 3 + n_leg

    value = pairs_per_step * args.steps / (dev_ms / 1e3)
    e2e_value = pairs_per_step * args.steps / (e2e_ms / 1e3)
    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return 0

    # ---- CPU baseline on this box's host cores (bounded sample, the SAME noisy sequence type) ----------------------------
    cpu, sads_per_pair = None, None
    if world == 1 and not args.no_cpu:
        cores = max(1, min(os.cpu_count() or 1, 64))
        ppw = 12 if W * H < 1_000_000 else 3
        v, kind, wall, total, sads_per_pair = cpu_throughput(W, H, ppw, cores, args.noise)
        cpu = {"value": v, "unit": UNIT, "cores": cores, "kind": kind,
               "sample": f"{cores} single-threaded processes x {ppw} consecutive pairs ({total} pairs, {wall:.1f} s wall) of "
                         f"the same {W}x{H} synthetic sequence type with +-{args.noise} grey levels of sensor noise: "
                         f"myComputeFeatures+myMatching from oracle/_ref, bucketing via the oracle port"}

    # ---- rooflines of the dominant kernel (k2_match) ----------------------------------------
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    if os.path.exists(peaks_path):
        try:
            peak = float(json.load(open(peaks_path))["hbm_gbs"]); peak_src = "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    n_c = float(np.mean(n_feat)); n_m = float(np.mean(raw_n)); n_b = float(np.mean(bk_n))
    # algorithmic bytes per pair
    b_pair = W * H + 48 * n_c + 48 * (2 * n_c) + 48 * n_m + 48 * n_m + 48 * n_b
    b_match = 48 * (2 * n_c) + 4 * (2 * n_c)              # k2: both feature lists in, one best-id per origin out
    k2_ms, k2_calls = prof["match"]
    k2_avg_s = k2_ms / max(k2_calls, 1) / 1e3
    achieved = b_match * B / k2_avg_s / 1e9 if k2_avg_s > 0 else 0.0
    step_ms = dev_ms / args.steps
    # measured DRAM bytes and ALU-pipe utilisation of K2 come from an `ncu --set full` capture of THIS build
    # (tools/ncu_traffic.py writes profiles/traffic.json with the hash of the kernel sources); a capture of another
    # build is refused instead of silently reported
    traffic, alu_pct, traffic_note = None, None, None
    src_hash = source_hash()
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        if tj.get("source_sha256") == src_hash:
            k2 = tj["k2_match"]
            traffic = (k2["dram_bytes_read"] + k2["dram_bytes_write"]) / k2["pairs_per_launch"] * B
            alu_pct = k2.get("alu_pipe_pct_of_peak")
            traffic_note = f"{k2.get('capture')} (sources {src_hash})"
        else:
            traffic_note = (f"profiles/traffic.json was captured from sources {tj.get('source_sha256')}, this build is {src_hash}: "
                            "not reported (re-run tools/profile_round.sh)")
            print("bench.py: " + traffic_note, file=sys.stderr)
    except Exception as exc:
        traffic_note = f"profiles/traffic.json unreadable ({type(exc).__name__})"
    # integer-pipe roofline: a 32-byte SAD is 8 VABSDIFF4 on the ALU pipe; tools/intpipe_bench.cu measured how many the
    # machine issues (profiles/r2_intpipe.json).  The reference's search evaluates `sads_per_pair` of them per frame pair
    # (counted by the oracle on the CPU leg's frames); K2's pruning returns the same argmin with fewer, so the figure is
    # "reference-equivalent SAD32/s" and may exceed the exhaustive peak.
    int_pipe = None
    try:
        ip = json.load(open(os.path.join(ROOT, "profiles", "r2_intpipe.json")))
        rate = next(r["warp_inst_per_clk_per_sm"] for r in ip["results"] if r["op"] == "vabsdiff4_acc")
        ghz = (clocks or {}).get("sm_mhz") or ip["sm_clock_max_ghz"] * 1e3
        peak_sad = rate * 32 / 8 * ip["sms"] * ghz * 1e6
        int_pipe = {"measured_peak_sad32_per_s": peak_sad, "vabsdiff4_warp_inst_per_clk_per_sm": rate, "sm_mhz": ghz,
                    "source": "profiles/r2_intpipe.json (tools/intpipe_bench.cu on a B200 of this pool)",
                    "alu_pipe_pct_of_peak": alu_pct,
                    "alu_pipe_note": "sm__inst_executed_pipe_alu of the K2 launches in the ncu capture of this build" if alu_pct is not None else traffic_note}
        if sads_per_pair:
            eq = sads_per_pair * (B / k2_avg_s) if k2_avg_s > 0 else 0.0
            int_pipe.update({"reference_sad32_per_pair": sads_per_pair, "sad32_equiv_per_s": eq, "frac": eq / peak_sad,
                             "what": "SAD32 evaluations of the reference's exhaustive search per second of K2 time / measured VABSDIFF4 peak"})
    except Exception as exc:
        int_pipe = {"unavailable": f"profiles/r2_intpipe.json: {type(exc).__name__}"}
    fam = {k: (v[0] / max(v[1], 1)) for k, v in prof.items()}
    roofline = {
        "kernel": "k2_match", "launches": "per step the matching of a batch is k2_match twice (prev->curr in full, curr->prev from "
                  "hints) + k2_hint_scatter; algorithmic bytes, traffic and avg_launch_ms are for that group",
        # BASELINE.json's metric asks for the HBM fraction, so achieved/peak/frac are the HBM figures; the kernel is bound
        # by the SM integer (ALU) pipe -- `int_pipe` holds that roofline
        "bound": "hbm", "binding_resource": "sm_integer_alu_pipe", "achieved": achieved, "peak": peak, "unit": "GB/s",
        "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_note, "peak_source": peak_src,
        "algorithmic_bytes_per_launch": b_match * B, "avg_launch_ms": k2_avg_s * 1e3,
        "share_of_step": (k2_avg_s * 1e3) / step_ms if step_ms else None,
        "kernel_ms_per_step": fam,
        "path": {"bytes_per_pair": b_pair, "achieved": b_pair * value / world / 1e9, "frac": b_pair * value / world / 1e9 / peak},
        "int_pipe": int_pipe,
        "note": "SAD matching is integer-pipe bound (about 100 int ops per algorithmic byte, SURVEY.md row D3); "
                "the HBM fraction is reported because BASELINE.json's metric asks for it",
    }

    line = {
        "metric": metric_name(W, H), "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": step_ms, "higher_is_better": True, "scaling": scaling, "vs_baseline": None, "dtype": "u8",
        "data": "synthetic",
        "config": {"workload": workload_name(args),
                   "pairs_per_step": pairs_per_step, "batch_per_gpu": B, "sequences_per_gpu": nq, "bucket": "2 per 50x50",
                   "l2": f"inputs larger than L2: pool of {pool_frames * nq} frames ({pool_frames * nq * H * pitch / 1e6:.0f} MB) cycled",
                   "features_per_frame": n_c, "raw_matches_per_pair": n_m, "bucketed_per_pair": n_b,
                   "sharding": ("8 sequences (seeds 1..8) round-robin over the ranks, no collective" if args.config == 5 else
                                "independent sequences per GPU, no collective"), "host_affinity": host_affinity,
                   "sensor_noise": f"uniform +-{args.noise} grey levels, independent per frame (both arms)"},
        "clocks": clocks,
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(B * H * pitch) * world,
                "d2h_bytes_per_step": int(counts[0].nbytes + B * max_keep * 48) * world,
                "ms_per_step": e2e_ms / args.steps, "pipeline": "depth 2: H2D(s+1) | kernels(s) | D2H(s-1) on three streams; "
                "every step's frames go host->device and its counts+bucketed lists come back inside the timed region",
                "bucketed_matches_read_back": consumed[0]},
        "gpu_launches": int(launches) * world,
        "roofline": roofline,
        "cpu_baseline": cpu,
        "other_noise_levels": extra or None,
    }
    print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", type=int, default=1, choices=[1, 5],
                    help="BASELINE.json config: 1 = configs[1], 1241x376 frame pairs (the headline); 5 = configs[4], 8 sequences of 2048x1024")
    ap.add_argument("--width", type=int, default=None)
    ap.add_argument("--height", type=int, default=None)
    ap.add_argument("--batch", type=int, default=None, help="frame pairs per step: per GPU (config 1), of the whole job (config 5)")
    ap.add_argument("--sequences", type=int, default=8, help="config 5: independent sequences of the job")
    ap.add_argument("--pool", type=int, default=None, help="distinct frames cycled per GPU (keeps inputs larger than L2)")
    ap.add_argument("--rows", type=int, default=0, help="match_rows_per_cta tuning knob (0 = library default)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the CPU baseline leg")
    ap.add_argument("--no-extra-legs", "--no-noise-leg", dest="no_extra_legs", action="store_true",
                    help="skip the extra device-resident legs at other noise levels (N=1)")
    ap.add_argument("--noise", type=int, default=NOISE_DEFAULT,
                    help="uniform +-N grey levels of independent sensor noise on every frame, both arms (0 = best case)")
    args = ap.parse_args()
    if args.config == 5:
        args.width, args.height = args.width or 2048, args.height or 1024
        args.batch = args.batch or 256   # 32 new frames per sequence and step: at 8 GPUs a rank still has 32 pairs per launch
        args.pool = args.pool or 512
        if args.steps == 200:
            args.steps = 40
    else:
        args.width, args.height = args.width or W_DEFAULT, args.height or H_DEFAULT
        args.batch = args.batch or 128
        args.pool = args.pool or 384
    if args.impl == "reference":
        return run_reference(args)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.gpus != world and world == 1 and args.gpus > 1:
        # convenience: re-launch under torchrun when called directly with --gpus N
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", "29511", os.path.abspath(__file__)] + sys.argv[1:]
        return subprocess.call(cmd)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
