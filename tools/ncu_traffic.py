#!/usr/bin/env python
"""Write profiles/traffic.json from `ncu --set full` captures of the CURRENT build: measured DRAM bytes per launch and the
ALU-pipe utilisation of the dominant kernels, stamped with the hash of the kernel sources (bench.py refuses a stamp that
does not match the build it benchmarks).

    python tools/ncu_traffic.py gpurun_out/<tag>_k2_match.ncu-rep gpurun_out/<tag>_k1_extract.ncu-rep <pairs per launch>
"""
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import source_hash  # noqa: E402


def rows(path):
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    r = list(csv.reader(raw.splitlines()))
    return [dict(zip(r[0], v)) for v in r[2:]]


def num(d, k):
    return float(d[k].replace(",", ""))


def summarise(path):
    rs = rows(path)
    t = sum(num(d, "gpu__time_duration.sum") for d in rs)
    return {
        "capture": os.path.basename(path) + f" ({len(rs)} launch{'es' if len(rs) > 1 else ''})",
        "dram_bytes_read": sum(num(d, "dram__bytes_read.sum") for d in rs),
        "dram_bytes_write": sum(num(d, "dram__bytes_write.sum") for d in rs),
        "gpu_time_ns": t,
        "alu_pipe_pct_of_peak": sum(num(d, "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active") * num(d, "gpu__time_duration.sum")
                                    for d in rs) / t,
        "issue_active_pct": sum(num(d, "smsp__issue_active.avg.pct_of_peak_sustained_active") * num(d, "gpu__time_duration.sum") for d in rs) / t,
        "warp_instructions": sum(num(d, "smsp__inst_executed.sum") for d in rs),
    }


def main():
    k2, k1, n = sys.argv[1], sys.argv[2], int(sys.argv[3])
    out = {"_comment": "dram__bytes_read.sum + dram__bytes_write.sum per step of the dominant kernels and their ALU-pipe utilisation, "
                       "from `ncu --set full` captures of the build whose kernel sources hash to source_sha256 (bench.py source_hash); "
                       "k2_match = its two launches of a step; ncu flushes caches before every profiled launch, so these are cold-cache upper bounds",
           "source_sha256": source_hash(),
           "k2_match": dict(summarise(k2), pairs_per_launch=n), "k1_extract": dict(summarise(k1), frames_per_launch=n)}
    with open(os.path.join(ROOT, "profiles", "traffic.json"), "w") as f:
        json.dump(out, f, indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
