set -u
out=gpurun_out; tag=r2l
for a in 0 2 8; do VO_B200_LIB=build/ab/libvo_stats.so timeout 120 python tools/k2_stats.py $a; done 2>&1 | grep -v "^origins"
ncu --set full --clock-control none --import-source on -k regex:k2_match -s 8 -c 2 -o $out/${tag}_noise2_k2_match python bench.py --noise 2 --steps 2 --warmup 3 --no-cpu --no-extra-legs > $out/${tag}_ncu_noise2_k2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k2_match -s 8 -c 2 -o $out/${tag}_noise0_k2_match python bench.py --noise 0 --steps 2 --warmup 3 --no-cpu --no-extra-legs > $out/${tag}_ncu_noise0_k2.log 2>&1
