set -u
out=gpurun_out; tag=r2g
timeout 600 python -m pytest tests/test_gpu_cases.py tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -4
for sg in 0 32 64 128 160; do echo "== seed_good $sg"; VO_SEED_GOOD=$sg VO_PROBE_PARITY=1 timeout 300 python tools/noise_probe.py 0 2 8 2>&1 | grep -E "noise|parity|Error" | sed 's/features.*per 64 pairs://'; done > $out/${tag}_seed_sweep.txt 2>&1
cat $out/${tag}_seed_sweep.txt
timeout 300 python tools/stress_match.py 30 11 2>&1 | tail -1
