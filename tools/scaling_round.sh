#!/usr/bin/env bash
# 1/2/4/8-GPU scaling of bench.py on one box (run through `gpurun --gpus 8`): bash tools/scaling_round.sh r1d
set -u
tag="${1:-rX}"; out=gpurun_out; mkdir -p "$out"
python bench.py --steps 100 --no-cpu > "$out/${tag}_scale_1.json" 2> "$out/${tag}_scale.err"
for n in 2 4 8; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) \
      bench.py --gpus $n --steps 100 --no-cpu > "$out/${tag}_scale_$n.json" 2>> "$out/${tag}_scale.err"
done
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29600 \
    bench.py --gpus 8 --steps 30 --no-cpu --width 2048 --height 1024 --batch 32 > "$out/${tag}_scale_8_2048x1024.json" 2>> "$out/${tag}_scale.err"
for f in "$out/${tag}"_scale_*.json; do echo "$f"; cut -c1-160 "$f"; done
tail -3 "$out/${tag}_scale.err"
