set -u
out=gpurun_out; tag=r2r
timeout 900 python -m pytest tests/test_gpu_cases.py tests/test_gpu_parity.py tests/test_gpu_stereo.py -m gpu -x -q 2>&1 | tail -3
VO_PROBE_PARITY=1 timeout 300 python tools/noise_probe.py 0 1 2 4 8 16 2>&1 | grep -E "noise|parity|Error" | sed 's/features.*per 64 pairs://' > $out/${tag}_noise_probe.txt 2>&1
cat $out/${tag}_noise_probe.txt
timeout 300 python tools/stress_match.py 40 31 2>&1 | tail -1
python tools/bench_stereo.py --cpu-pairs 0 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('stereo', d['quads_per_s'], d['kernel_ms_per_step'])"
