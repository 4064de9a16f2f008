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
    assert "int_pipe" in r and r["binding_resource"] == "sm_integer_alu_pipe"
    assert "+-2 grey levels" in d["config"]["workload"]      # the headline runs on frames with sensor noise
    legs = d["other_noise_levels"]
    assert legs["best_case_noise_free"]["value"] > 0 and legs["noise_8"]["value"] > 0
    # the reference arm names the same workload and metric (the driver compares the two lines)
    ref = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0"],
                         capture_output=True, text=True, timeout=900, cwd=ROOT)
    assert ref.returncode == 0, ref.stderr[-2000:]
    rd = json.loads([ln for ln in ref.stdout.splitlines() if ln.strip()][-1])
    assert rd["impl"] == "reference" and rd["metric"] == d["metric"] and rd["config"]["workload"] == d["config"]["workload"]
    assert rd["cpu_baseline"]["value"] == rd["value"] and rd["e2e"]["h2d_bytes_per_step"] == 0
    # config 5 (BASELINE configs[4]) is a runnable, correctly labelled configuration
    c5 = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--config", "5", "--steps", "2", "--warmup", "3", "--batch", "16",
                         "--pool", "32", "--no-cpu", "--no-extra-legs"], capture_output=True, text=True, timeout=900, cwd=ROOT)
    assert c5.returncode == 0, c5.stderr[-2000:]
    cd = json.loads([ln for ln in c5.stdout.splitlines() if ln.strip()][-1])
    assert cd["metric"] == "frame_pairs_per_s_2048x1024" and cd["scaling"] == "strong" and cd["config"]["sequences_per_gpu"] == 8
    assert cd["config"]["pairs_per_step"] == 16 and cd["value"] > 0
    print("bench contract ok")


if __name__ == "__main__":
    main()
