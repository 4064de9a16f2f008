set -u
out=gpurun_out
for n in 8 4; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29560 + n)) bench.py --gpus $n --config 5 --batch 256 --steps 20 --no-cpu --no-extra-legs > $out/r2s_config5_b256_$n.json 2>> $out/r2s_config5_b256.err
cut -c1-140 $out/r2s_config5_b256_$n.json | tail -1
done
