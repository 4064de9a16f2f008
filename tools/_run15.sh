set -u
out=gpurun_out; tag=r2q
timeout 600 python -m pytest tests/test_gpu_process.py -m gpu -x -q 2>&1 | tail -3
python tools/bench_process.py > $out/${tag}_process_loop.json 2> $out/${tag}_process.err; python -c "
import json; d=json.load(open('$out/${tag}_process_loop.json')); print(d['frames_per_s'], d['process_ok_frames'], d['mean_inliers'], d['stage_ms_per_frame'])"
python tools/bench_process.py --threads 8 --ref-frames 0 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('8 threads', d['frames_per_s'], d['stage_ms_per_frame'])"
