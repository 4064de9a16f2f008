set -u
out=gpurun_out; tag=r2h
timeout 600 python -m pytest tests/test_gpu_cases.py tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -4
for sg in 0 48 160; do echo "== seed_good $sg"; VO_SEED_GOOD=$sg VO_PROBE_PARITY=1 timeout 300 python tools/noise_probe.py 0 2 8 2>&1 | grep -E "noise|parity|Error" | sed 's/features.*per 64 pairs://'; done > $out/${tag}_seed_sweep.txt 2>&1
cat $out/${tag}_seed_sweep.txt
timeout 300 python tools/stress_match.py 30 13 2>&1 | tail -1
ncu --set full --clock-control none --import-source on -k regex:k2_match -s 8 -c 2 -o $out/${tag}_noise2_k2_match python bench.py --noise 2 --steps 2 --warmup 3 --no-cpu --no-extra-legs > $out/${tag}_ncu_noise2_k2.log 2>&1
