set -u
out=gpurun_out; tag=r2n
python tools/bench_process.py > $out/${tag}_process_loop.json 2> $out/${tag}_process.err; cat $out/${tag}_process_loop.json
python tools/bench_stereo.py --cpu-pairs 0 > $out/${tag}_stereo_bench.json 2> $out/${tag}_stereo.err; cut -c1-400 $out/${tag}_stereo_bench.json
for r in 4 5 6 8; do echo "rows $r"; python bench.py --config 5 --no-cpu --no-extra-legs --steps 20 --rows $r 2>/dev/null | cut -c1-130; done
timeout 300 python -m pytest tests/test_gpu_process.py tests/test_gpu_stereo.py -m gpu -x -q 2>&1 | tail -2
