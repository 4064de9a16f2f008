#!/usr/bin/env bash
# Multi-GPU legs of the scaling record on one box (run through `gpurun --gpus 8`; the N=1 legs come from a 1-GPU call and
# the driver's own SCALE run covers the headline workload at 1/2/4/8):
#   bash tools/scaling_round.sh r2
# BASELINE configs[4] (8 sequences of 2048x1024, strong scaling) at 2/4/8 GPUs, the headline workload at 8, and the box's
# bare host->device ceiling at 2/4/8 ranks next to the end-to-end figures.
set -u
tag="${1:-rX}"; out=gpurun_out; mkdir -p "$out"
run() { n=$1; shift; python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n + ${PORT_OFF:-0})) "$@"; }
# SCALE_QUICK=1: the 8-GPU headline run and configs[4] at 8 and 4 only (an 8-GPU box is charged 8x)
h2d_ns="8 4 2"; c5_ns="8 4 2"
if [ -n "${SCALE_QUICK:-}" ]; then h2d_ns=""; c5_ns="8 4"; fi
for n in $h2d_ns; do
  PORT_OFF=40 run $n tools/h2d_ceiling.py > "$out/${tag}_h2d_$n.json" 2>> "$out/${tag}_scale.err"
done
run 8 bench.py --gpus 8 --steps 60 --no-cpu --no-extra-legs > "$out/${tag}_scale_8.json" 2>> "$out/${tag}_scale.err"
for n in $c5_ns; do
  PORT_OFF=20 run $n bench.py --gpus $n --config 5 --steps 30 --no-cpu --no-extra-legs > "$out/${tag}_scale_config5_$n.json" 2>> "$out/${tag}_scale.err"
done
for f in "$out/${tag}"_scale_*.json "$out/${tag}"_h2d_*.json; do echo "$f"; cut -c1-200 "$f"; done
tail -3 "$out/${tag}_scale.err"
