#!/usr/bin/env python
"""Turn gpurun_out/ ncu artefacts into the small text summaries committed under profiles/.

  python tools/ncu_summary.py launches gpurun_out/launches.csv > profiles/rN_launches.txt
  python tools/ncu_summary.py kernel   gpurun_out/k2_full.ncu-rep > profiles/rN_k2_full.txt
  python tools/ncu_summary.py sass     gpurun_out/k2_full.ncu-rep [min_share] > hot SASS lines
  python tools/ncu_summary.py lines    gpurun_out/k2_full.ncu-rep [min_share] > executed instructions per CUDA source line
"""
import collections
import csv
import subprocess
import sys

KEEP = (
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed.sum", "sm__inst_executed.avg.per_cycle_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_uniform.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_adu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_cbu.avg.pct_of_peak_sustained_active",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
    "launch__shared_mem_per_block_dynamic", "launch__shared_mem_per_block_static", "launch__occupancy_limit_registers",
    "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_warps", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "smsp__thread_inst_executed_per_inst_executed.ratio", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
    "lts__t_bytes.sum", "sm__cycles_active.avg",
)


def launches(path):
    rows = [r for r in csv.reader(open(path, errors="replace")) if len(r) > 10]
    hdr = rows[0]
    ki, vi, gi, bi = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Grid Size"), hdr.index("Block Size")
    agg = collections.OrderedDict()
    for r in rows[1:]:
        try:
            v = float(r[vi].replace(",", ""))
        except ValueError:
            continue
        name = r[ki].split("(")[0].replace("void ", "")
        a = agg.setdefault(name, [0, 0.0, r[gi], r[bi]])
        a[0] += 1
        a[1] += v
    tot = sum(a[1] for a in agg.values())
    print(f"# ncu --metrics gpu__time_duration.sum --clock-control none (serialised, cold-cache launch times); source: {path}")
    print(f"{'kernel':28s} {'grid':>16s} {'block':>14s} {'launches':>8s} {'avg_us':>10s} {'total_us':>10s} {'share':>7s}")
    for k, a in agg.items():
        print(f"{k:28s} {a[2]:>16s} {a[3]:>14s} {a[0]:8d} {a[1] / a[0] / 1e3:10.1f} {a[1] / 1e3:10.1f} {a[1] / tot:7.3f}")


def kernel(path):
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    h, u = rows[0], rows[1]
    for v in rows[2:]:
        d = dict(zip(h, zip(u, v)))
        print(f"# {path}: {d.get('Kernel Name', ('', '?'))[1]}  grid {d.get('Grid Size', ('', '?'))[1]} block {d.get('Block Size', ('', '?'))[1]}")
        for k in KEEP:
            if k in d:
                print(f"{k:90s} {d[k][1]:>18s} {d[k][0]}")
    det = subprocess.run(["ncu", "-i", path, "--page", "details"], capture_output=True, text=True).stdout
    print("\n# ---- ncu --page details (trimmed) ----")
    keep = False
    for ln in det.splitlines():
        s = ln.strip()
        if s.startswith("Section:"):
            keep = any(x in s for x in ("Speed Of Light Throughput", "Compute Workload", "Memory Workload Analysis", "Scheduler", "Warp State",
                                        "Instruction Statistics", "Launch Statistics", "Occupancy"))
            keep = keep and "Chart" not in s and "Tables" not in s
        if keep and s and not s.startswith(("OPT", "INF")) and len(s) < 110:
            print(ln.rstrip())


def sass(path, min_share=0.002):
    raw = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr = rows[1]
    ii = hdr.index("Instructions Executed")
    out = []
    for r in rows[2:]:
        try:
            out.append((r[0][-5:], r[1].strip(), int(r[ii]), int(r[4]), r[ii + 3]))
        except (ValueError, IndexError):
            pass
    tot = sum(o[2] for o in out)
    print(f"# {rows[0][1]}: {tot} warp instructions executed; lines with >= {min_share:.3%} shown (addr, Minst, stall samples, avg threads)")
    for o in out:
        if o[2] >= tot * min_share:
            print(f"{o[0]} {o[2] / 1e6:8.2f}M s={o[3]:5d} thr={o[4]:>3s} {o[1]}")


def lines(path, min_share=0.004):
    """Executed warp instructions and stall samples aggregated per CUDA source line (needs -lineinfo + --import-source on)."""
    raw = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr = next(r for r in rows if r and r[0] == "Line No")
    ii, si = hdr.index("Instructions Executed"), hdr.index("# Samples")
    out = []
    for r in rows:
        if r and r[0].isdigit():
            try:
                out.append((int(r[0]), int(r[ii]), int(r[si]), r[1].strip()))
            except (ValueError, IndexError):
                pass
    tot, ts = sum(o[1] for o in out) or 1, sum(o[2] for o in out) or 1
    print(f"# {path}: {tot} warp instructions with line info; lines with >= {min_share:.2%} of instructions or samples (line, Minst, share, stall-sample share)")
    for ln, n, smp, src in sorted(out):
        if n >= tot * min_share or smp >= ts * min_share:
            print(f"L{ln:4d} {n / 1e6:8.1f}M {n / tot:6.1%} samp {smp / ts:6.1%}  {src[:110]}")


if __name__ == "__main__":
    fn = {"launches": launches, "kernel": kernel, "sass": sass, "lines": lines}[sys.argv[1]]
    if sys.argv[1] in ("sass", "lines") and len(sys.argv) > 3:
        fn(sys.argv[2], float(sys.argv[3]))
    else:
        fn(sys.argv[2])
