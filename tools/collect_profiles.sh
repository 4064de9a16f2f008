#!/usr/bin/env bash
# Turns the raw output of tools/profile_round.sh <tag> (gpurun_out/<tag>_*) into the committed summaries profiles/<tag>_*:
#   bash tools/collect_profiles.sh r2t
set -u
tag="${1:?capture tag}"
in=gpurun_out; out=profiles
for f in bench bench_reference bench_config5 stereo_bench stereo_bench_norefine process_loop h2d_1 intpipe; do
  [ -s "$in/${tag}_$f.json" ] && cp "$in/${tag}_$f.json" "$out/${tag}_$f.json"
done
[ -s "$in/${tag}_noise_probe.txt" ] && sed 's/features \([0-9]*\), raw matches \([0-9]*\); //' "$in/${tag}_noise_probe.txt" > "$out/${tag}_noise_probe.txt"
[ -s "$in/${tag}_launches.csv" ] && python tools/ncu_summary.py launches "$in/${tag}_launches.csv" > "$out/${tag}_launches.txt"
[ -s "$in/${tag}_stereo_launches.csv" ] && python tools/ncu_summary.py launches "$in/${tag}_stereo_launches.csv" > "$out/${tag}_stereo_launches.txt"
for k in k2_match k1_extract clean_k2_match k_ransac_score stereo_k2_match stereo_k5_refine; do
  [ -s "$in/${tag}_$k.ncu-rep" ] && python tools/ncu_summary.py kernel "$in/${tag}_$k.ncu-rep" > "$out/${tag}_${k}_full.txt"
done
for k in k2_match k1_extract stereo_k5_refine; do
  [ -s "$in/${tag}_$k.ncu-rep" ] && python tools/ncu_summary.py lines "$in/${tag}_$k.ncu-rep" > "$out/${tag}_${k}_lines.txt"
done
if [ -s "$in/${tag}_k2_match.ncu-rep" ]; then
  { python tools/k2_regions.py "$in/${tag}_k2_match.ncu-rep"; echo; echo "# clean (noise-free) frames"; python tools/k2_regions.py "$in/${tag}_clean_k2_match.ncu-rep"; } > "$out/${tag}_k2_match_regions.txt"
  python tools/k2_stalls.py "$in/${tag}_k2_match.ncu-rep" > "$out/${tag}_k2_match_stalls.txt"
fi
ls -la "$out" | grep "${tag}_"
