#!/usr/bin/env python
"""Top CUDA source lines of csrc/vo_match.cu by warp stall samples in one launch of an ncu capture (--import-source on).

    python tools/k2_stalls.py gpurun_out/<capture>.ncu-rep [launch index, default 0] [lines, default 30]
"""
import collections
import csv
import subprocess
import sys


def main():
    path = sys.argv[1]
    launch = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 30
    raw = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    starts = [i for i, r in enumerate(rows) if r and r[0] == "File Path" and r[1].endswith("vo_match.cu")]
    a = starts[launch]
    b = next((i for i in range(a + 1, len(rows)) if rows[i] and rows[i][0] == "File Path"), len(rows))
    hdr = next(r for r in rows[a:b] if r and r[0] == "Line No")
    ii, si = hdr.index("Instructions Executed"), hdr.index("# Samples")
    stall_cols = [(i, h) for i, h in enumerate(hdr) if h.startswith("stall_")]
    samp, inst, text, why = collections.Counter(), collections.Counter(), {}, collections.defaultdict(collections.Counter)
    for r in rows[a:b]:
        if r and r[0].isdigit():
            try:
                ln = int(r[0])
                samp[ln] += int(r[si]); inst[ln] += int(r[ii]); text[ln] = r[1].strip()[:86]
                for i, h in stall_cols:
                    if r[i] and r[i] != "0":
                        why[ln][h[6:]] += int(r[i])
            except (ValueError, IndexError):
                pass
    tot = sum(samp.values()) or 1
    print(f"# {path} launch {launch}: {tot} samples with line info in vo_match.cu")
    for ln, n in samp.most_common(top):
        w = ", ".join(f"{k} {v / n:.0%}" for k, v in why[ln].most_common(2)) if n else ""
        print(f"L{ln:4d} samp {n / tot:6.1%} inst {inst[ln] / 1e6:7.1f}M  [{w}]  {text[ln]}")


if __name__ == "__main__":
    main()
