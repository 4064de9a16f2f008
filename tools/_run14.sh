set -u
out=gpurun_out; tag=r2p
timeout 900 python -m pytest tests/test_gpu_process.py -m gpu -x -q 2>&1 | tail -6
