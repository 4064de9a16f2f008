set -u
out=gpurun_out; tag=r2o
timeout 600 python -m pytest tests/test_gpu_cases.py tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -4
VO_PROBE_PARITY=1 timeout 300 python tools/noise_probe.py 0 1 2 4 8 16 2>&1 | grep -E "noise|parity|Error" | sed 's/features.*per 64 pairs://' > $out/${tag}_noise_probe.txt 2>&1
cat $out/${tag}_noise_probe.txt
true
timeout 300 python tools/stress_match.py 30 19 2>&1 | tail -1
ncu --set full --clock-control none --import-source on -k regex:k2_match -s 8 -c 2 -o $out/${tag}_noise2_k2_match python bench.py --noise 2 --steps 2 --warmup 3 --no-cpu --no-extra-legs > $out/${tag}_ncu_noise2_k2.log 2>&1
