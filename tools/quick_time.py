"""Developer timing probe (not the benchmark): batched extract + match at 1241x376."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from hls_final_visual_odometry_b200 import Engine
from hls_final_visual_odometry_b200.synth import make_sequence

W, H = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (1241, 376)
B = int(sys.argv[3]) if len(sys.argv) > 3 else 32
rows = int(sys.argv[4]) if len(sys.argv) > 4 else 0
exh = bool(int(sys.argv[5])) if len(sys.argv) > 5 else False
threads = int(sys.argv[6]) if len(sys.argv) > 6 else 0
seq = make_sequence(W, H, B + 1, seed=1)
eng = Engine(W, H, max_frames=B + 1, max_pairs=B, match_rows_per_cta=rows, match_exhaustive=exh, match_threads=threads)
eng.batch_extract_host(0, seq)
prev = np.arange(B, dtype=np.int32); curr = prev + 1
eng.batch_match(prev, curr, 2, 50.0, 50.0)
raw, bk = eng.batch_counts(B)
print("raw", raw[:4], "bucketed", bk[:4], "features", eng.slot_feature_count(0))
eng.profile_enable(True)
for it in range(5):
    eng.batch_extract_host(0, seq)
    eng.batch_match(prev, curr, 2, 50.0, 50.0)
eng.synchronize()
t0 = time.perf_counter()
for it in range(10):
    eng.batch_extract_host(0, seq)
    eng.batch_match(prev, curr, 2, 50.0, 50.0)
eng.synchronize()
dt = time.perf_counter() - t0
print(f"{W}x{H} B={B} rows={rows} exhaustive={exh} threads={threads}: {10*B/dt:.1f} pairs/s wall (incl H2D of {B+1} frames)")
prof = eng.profile_read()
for k, (ms, calls) in prof.items():
    print(f"  {k}: {ms/max(calls,1)*1e3:.1f} us/launch over {calls} launches -> {ms/max(calls,1)*1e3/B:.2f} us/pair")
