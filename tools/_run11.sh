set -u
out=gpurun_out; tag=r2m
timeout 600 python -m pytest tests/test_gpu_cases.py tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -3
VO_PROBE_PARITY=1 timeout 300 python tools/noise_probe.py 0 1 2 4 8 16 2>&1 | grep -E "noise|parity|Error" | sed 's/features.*per 64 pairs://' > $out/${tag}_noise_probe.txt 2>&1
cat $out/${tag}_noise_probe.txt
for v in f8 f16 f24; do echo "== $v"; VO_B200_LIB=build/ab/libvo_$v.so VO_PROBE_PARITY=1 python tools/noise_probe.py 0 2 8 2>&1 | grep -E "noise|parity|Error" | sed 's/features.*per 64 pairs://'; done
timeout 300 python tools/stress_match.py 30 29 2>&1 | tail -1
