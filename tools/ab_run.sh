for v in base flood8 flood20 append2 seedx; do echo "== $v"; VO_B200_LIB=build/ab/libvo_$v.so VO_PROBE_PARITY=1 python tools/noise_probe.py 0 2 8 2>&1 | grep noise | sed 's/features.*per 64 pairs://'; done
echo "== base threads 256"; VO_MATCH_THREADS=256 VO_B200_LIB=build/ab/libvo_base.so python tools/noise_probe.py 0 2 8 2>&1 | grep noise | sed 's/features.*per 64 pairs://'
echo "== base threads 384"; VO_MATCH_THREADS=384 VO_B200_LIB=build/ab/libvo_base.so python tools/noise_probe.py 0 2 8 2>&1 | grep noise | sed 's/features.*per 64 pairs://'
