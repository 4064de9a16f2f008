#!/usr/bin/env python
"""Executed warp instructions and stall samples of a k2_match ncu capture, aggregated by code region of csrc/vo_match.cu
(regions are found by anchor strings in the CURRENT source, so the capture must come from the current build).

    python tools/k2_regions.py gpurun_out/<capture>.ncu-rep
"""
import collections
import csv
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "hls_final_visual_odometry_b200", "csrc", "vo_match.cu")
ANCHORS = [  # (first line containing the string starts the region)
    ("helpers: sad4 / cand_key / origin_vb", "__device__ __forceinline__ uint32_t sad4("),
    ("tables", "const int32_t* cnt_to = fa.cnt"),
    ("chunk setup", "bool first_chunk = true;"),
    ("staging", "// ---- stage: warp per (row, column) bin list"),
    ("load_origin", "auto load_origin = [&]"),
    ("pass 1: setup / hint", "// ================= pass 1"),
    ("pass 1: seed", "if (PRUNE && __any_sync(FULL, need_seed && rows_here)) {"),
    ("pass 1: class + ordered lists", "// ---- class: Z (bound 0)"),
    ("pass 2: item claim", "// ================= pass 2"),
    ("good task: setup", "// ---------- filter task"),
    ("good task: flush (cooperative evaluation)", "auto flush = [&]() {"),
    ("good task: row loop", "// rows in the order own row (of the first lane), then the rest, each claimed"),
    ("good task: filter", "// 32 candidates (8 groups) at a time"),
    ("good task: hits (queue / own loop)", "bool own = true;  // every lane evaluates its own hits"),
    ("P item", "// ---------- P item"),
    ("write-out", "// every searched origin's key, from the shared bounds"),
    ("kernel entry / rest", "first_chunk = false;"),
]


def main():
    path = sys.argv[1]
    src = open(SRC).read().splitlines()
    starts = []
    for name, needle in ANCHORS:
        ln = next((i + 1 for i, l in enumerate(src) if needle in l), None)
        if ln is None:
            raise SystemExit(f"anchor not found: {needle}")
        starts.append((ln, name))
    starts.sort()
    raw = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdrs = [i for i, r in enumerate(rows) if r and r[0] == "Line No"]
    files = [i for i, r in enumerate(rows) if r and r[0] == "File Path"]
    # sections: each "Line No" header belongs to the closest preceding "Function Name"/"File Path" rows
    launches = collections.OrderedDict()
    cur_fn, cur_file = None, None
    ii = si = None
    for i, r in enumerate(rows):
        if not r:
            continue
        if r[0] == "Function Name":
            cur_fn = r[1]
            if cur_fn not in launches:
                launches[cur_fn + f"#{len(launches)}"] = None
            key = list(launches.keys())[-1]
            launches[key] = [collections.Counter(), collections.Counter()]
        elif r[0] == "File Path":
            cur_file = r[1]
        elif r[0] == "Line No":
            ii, si = r.index("Instructions Executed"), r.index("# Samples")
        elif r[0].isdigit() and ii is not None:
            try:
                ln, n, smp = int(r[0]), int(r[ii]), int(r[si])
            except (ValueError, IndexError):
                continue
            agg, samp = launches[list(launches.keys())[-1]]
            if cur_file and cur_file.endswith("vo_match.cu"):
                name = "before"
                for s0, nm in starts:
                    if ln >= s0:
                        name = nm
            else:
                name = "inlined headers (" + os.path.basename(cur_file or "?") + ")"
            agg[name] += n
            samp[name] += smp
    # merge sections of the same launch: ncu prints one Function section per file; launches alternate
    for key, (agg, samp) in launches.items():
        tot, ts = sum(agg.values()), sum(samp.values())
        if tot < 1e6:
            continue
        print(f"== {key[:70]}: {tot / 1e6:.1f} M warp instructions")
        for nm, n in sorted(agg.items(), key=lambda kv: -kv[1]):
            print(f"   {nm:46s} {n / 1e6:9.1f}M {n / tot:6.1%}   stall samples {samp[nm] / max(ts, 1):6.1%}")


if __name__ == "__main__":
    main()
