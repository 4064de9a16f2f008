"""Developer check: bench.py prints ONE JSON line with the keys the driver reads (small batch, no CPU leg).  Needs a GPU.

    python tools/check_bench_contract.py
"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "3", "--warmup", "3", "--batch", "8",
                          "--pool", "16", "--no-cpu"], capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [ln for ln in out.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1, lines
    d = json.loads(lines[0])
    assert d["metric"] == "frame_pairs_per_s_1241x376" and d["unit"] == "frame-pairs/s" and d["n_gpus"] == 1
    assert d["steps"] == 3 and d["warmup"] >= 3 and d["higher_is_better"] is True and d["scaling"] == "weak"
    assert d["dtype"] == "u8" and d["data"] == "synthetic" and d["vs_baseline"] is None
    assert d["value"] > 0 and d["ms_per_step"] > 0 and "workload" in d["config"]
    e = d["e2e"]
    assert e["value"] > 0 and e["h2d_bytes_per_step"] == 8 * 376 * 1248 and e["d2h_bytes_per_step"] > 0
    assert d["gpu_launches"] == 3 * 7, d["gpu_launches"]  # K1, K2 x 2, hint scatter, K3 flags + emit, K4 per step
    r = d["roofline"]
    assert r["bound"] == "hbm" and r["unit"] == "GB/s" and r["peak"] > 0 and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-9
    assert set(d["clocks"]) >= {"sm_mhz", "sm_max_mhz", "reasons"}
    assert d["with_sensor_noise"]["value"] > 0
    print("bench contract ok")


if __name__ == "__main__":
    main()
