#!/usr/bin/env bash
# Collects the evidence committed under profiles/ for one capture tag (run on the GPU box through gpurun):
#   bash tools/profile_round.sh r1d
# Timed runs first (never under a profiler), then the ncu launch lists, then one full capture per kernel.
set -u
tag="${1:-rX}"
out=gpurun_out
mkdir -p "$out"
python bench.py > "$out/${tag}_bench.json" 2> "$out/${tag}_bench.err" || echo "bench failed"
python bench.py --impl reference --steps 3 --warmup 1 > "$out/${tag}_bench_reference.json" 2>> "$out/${tag}_bench.err" || echo "reference arm failed"
python tools/bench_stereo.py > "$out/${tag}_stereo_bench.json" 2> "$out/${tag}_stereo.err" || echo "stereo bench failed"
python tools/bench_stereo.py --refinement 0 --cpu-pairs 0 > "$out/${tag}_stereo_bench_norefine.json" 2>> "$out/${tag}_stereo.err"
python bench.py --noise 2 --steps 100 --no-cpu > "$out/${tag}_bench_noise2.json" 2>> "$out/${tag}_bench.err" || echo "noise bench failed"
python bench.py --noise 8 --steps 50 --no-cpu > "$out/${tag}_bench_noise8.json" 2>> "$out/${tag}_bench.err"
VO_PROBE_PARITY=1 python tools/noise_probe.py 0 1 2 4 8 16 > "$out/${tag}_noise_probe.txt" 2>&1 || echo "noise probe failed"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file "$out/${tag}_launches.csv" \
    python bench.py --steps 2 --warmup 3 --no-cpu --no-extra-legs > "$out/${tag}_ncu_launches.log" 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file "$out/${tag}_stereo_launches.csv" \
    python tools/bench_stereo.py --steps 2 --warmup 3 --cpu-pairs 0 > "$out/${tag}_ncu_stereo_launches.log" 2>&1
for k in k2_match k1_extract; do
  ncu --set full --clock-control none --import-source on -k regex:$k -s $([ $k = k2_match ] && echo 8 || echo 4) -c $([ $k = k2_match ] && echo 2 || echo 1) -o "$out/${tag}_${k}" \
      python bench.py --steps 2 --warmup 3 --no-cpu --no-extra-legs > "$out/${tag}_ncu_${k}.log" 2>&1
done
ncu --set full --clock-control none --import-source on -k regex:k2_match -s 8 -c 2 -o "$out/${tag}_noise2_k2_match" \
    python bench.py --noise 2 --steps 2 --warmup 3 --no-cpu > "$out/${tag}_ncu_noise2_k2.log" 2>&1
for k in k5_refine k3q_flags k2_match; do
  ncu --set full --clock-control none --import-source on -k regex:$k -s 3 -c 1 -o "$out/${tag}_stereo_${k}" \
      python tools/bench_stereo.py --steps 1 --warmup 3 --cpu-pairs 0 > "$out/${tag}_ncu_stereo_${k}.log" 2>&1
done
ls -la "$out" | grep "$tag"
