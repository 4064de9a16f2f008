set -u
out=gpurun_out; tag=r2l
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
VO_PROBE_PARITY=1 timeout 300 python tools/noise_probe.py 0 1 2 4 8 16 2>&1 | grep -E "noise|parity|Error" | sed 's/features.*per 64 pairs://' > $out/${tag}_noise_probe.txt 2>&1
cat $out/${tag}_noise_probe.txt
for v in rc2 sg64 sg128; do echo "== $v"; VO_B200_LIB=build/ab/libvo_$v.so VO_PROBE_PARITY=1 python tools/noise_probe.py 0 2 8 2>&1 | grep -E "noise|parity|Error" | sed 's/features.*per 64 pairs://'; done
timeout 300 python tools/stress_match.py 40 23 2>&1 | tail -1
python bench.py --no-cpu > $out/${tag}_bench.json 2> $out/${tag}_bench.err; cut -c1-200 $out/${tag}_bench.json
python bench.py --config 5 --no-cpu > $out/${tag}_bench_config5.json 2>> $out/${tag}_bench.err; cut -c1-200 $out/${tag}_bench_config5.json
