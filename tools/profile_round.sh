#!/usr/bin/env bash
# Collects the evidence committed under profiles/ for one capture tag (run on the GPU box through gpurun, one GPU):
#   bash tools/profile_round.sh r2s
# First the ncu captures of the dominant kernels (each after its own command has run without ncu inside this script's
# earlier calls), so that traffic.json is stamped with this build; then the timed runs (never under a profiler), which
# read it; then the launch lists and the remaining captures.
set -u
tag="${1:-rX}"
out=gpurun_out
mkdir -p "$out"
python bench.py --steps 5 --warmup 3 --no-cpu --no-extra-legs > "$out/${tag}_bench_short.json" 2> "$out/${tag}_bench.err" || echo "bench failed"
for k in k2_match k1_extract; do
  ncu --set full --clock-control none --import-source on -k regex:$k -s $([ $k = k2_match ] && echo 8 || echo 4) -c $([ $k = k2_match ] && echo 2 || echo 1) -o "$out/${tag}_${k}" \
      python bench.py --steps 2 --warmup 3 --no-cpu --no-extra-legs > "$out/${tag}_ncu_${k}.log" 2>&1
done
python tools/ncu_traffic.py "$out/${tag}_k2_match.ncu-rep" "$out/${tag}_k1_extract.ncu-rep" 128 > "$out/${tag}_traffic.json" 2>> "$out/${tag}_bench.err" && cp profiles/traffic.json "$out/${tag}_traffic_written.json"
python bench.py > "$out/${tag}_bench.json" 2>> "$out/${tag}_bench.err" || echo "bench failed"
python bench.py --impl reference --steps 3 --warmup 1 > "$out/${tag}_bench_reference.json" 2>> "$out/${tag}_bench.err" || echo "reference arm failed"
python bench.py --config 5 --no-cpu > "$out/${tag}_bench_config5.json" 2>> "$out/${tag}_bench.err" || echo "config 5 failed"
python tools/bench_stereo.py > "$out/${tag}_stereo_bench.json" 2> "$out/${tag}_stereo.err" || echo "stereo bench failed"
python tools/bench_stereo.py --refinement 0 --cpu-pairs 0 > "$out/${tag}_stereo_bench_norefine.json" 2>> "$out/${tag}_stereo.err"
python tools/bench_process.py > "$out/${tag}_process_loop.json" 2> "$out/${tag}_process.err" || echo "process bench failed"
VO_PROBE_PARITY=1 python tools/noise_probe.py 0 1 2 4 8 16 > "$out/${tag}_noise_probe.txt" 2>&1 || echo "noise probe failed"
python tools/h2d_ceiling.py > "$out/${tag}_h2d_1.json" 2>> "$out/${tag}_bench.err"
tools/intpipe_bench > "$out/${tag}_intpipe.json" 2>> "$out/${tag}_bench.err"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file "$out/${tag}_launches.csv" \
    python bench.py --steps 2 --warmup 3 --no-cpu --no-extra-legs > "$out/${tag}_ncu_launches.log" 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file "$out/${tag}_stereo_launches.csv" \
    python tools/bench_stereo.py --steps 2 --warmup 3 --cpu-pairs 0 > "$out/${tag}_ncu_stereo_launches.log" 2>&1
ncu --set full --clock-control none --import-source on -k regex:k2_match -s 8 -c 2 -o "$out/${tag}_clean_k2_match" \
    python bench.py --noise 0 --steps 2 --warmup 3 --no-cpu --no-extra-legs > "$out/${tag}_ncu_clean_k2.log" 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_ransac_score -c 1 -o "$out/${tag}_k_ransac_score" \
    python tools/bench_process.py --frames 256 --ref-frames 0 > "$out/${tag}_ncu_ransac.log" 2>&1
for k in k5_refine k2_match; do
  ncu --set full --clock-control none --import-source on -k regex:$k -s 3 -c 1 -o "$out/${tag}_stereo_${k}" \
      python tools/bench_stereo.py --steps 1 --warmup 3 --cpu-pairs 0 > "$out/${tag}_ncu_stereo_${k}.log" 2>&1
done
ls -la "$out" | grep "$tag"
