#!/usr/bin/env bash
# Poor man's memcheck (compute-sanitizer is closed on this pool): the parity suites, the noisy 1241x376 pair and the
# pruned-vs-exhaustive stress test against a build whose shared-memory indices are all checked (-DVO_DEBUG_BOUNDS traps),
# once with the library's own choice of K2 staging layout and once with the lean layout forced on every size.
#   python tools/ab_build.py bounds "-DVO_DEBUG_BOUNDS"      (on the build box)
#   bash tools/bounds_check.sh > gpurun_out/r2u_bounds_check.log
set -u
export VO_B200_LIB=build/ab/libvo_bounds.so
for lean in "" 1; do
  echo "== VO_MATCH_LEAN='${lean}' (empty: the library's choice)"
  if [ -n "$lean" ]; then export VO_MATCH_LEAN=$lean; fi
  python tools/sanitize_case.py 2>&1 | tail -2
  python tools/stress_match.py 40 21 2>&1 | tail -2
  python -m pytest tests/test_gpu_cases.py tests/test_gpu_parity.py tests/test_gpu_stereo.py tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -3
done
