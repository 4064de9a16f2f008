set -u
out=gpurun_out; tag=r2e
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > $out/${tag}_pytest.txt
python bench.py > $out/${tag}_bench.json 2> $out/${tag}_bench.err
VO_PROBE_PARITY=1 python tools/noise_probe.py 0 1 2 4 8 16 > $out/${tag}_noise_probe.txt 2>&1
ncu --set full --clock-control none --import-source on -k regex:k2_match -s 8 -c 2 -o $out/${tag}_noise2_k2_match python bench.py --noise 2 --steps 2 --warmup 3 --no-cpu --no-extra-legs > $out/${tag}_ncu_noise2_k2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k1_extract -s 4 -c 1 -o $out/${tag}_k1_extract python bench.py --steps 2 --warmup 3 --no-cpu --no-extra-legs > $out/${tag}_ncu_k1.log 2>&1
ls -la $out
