"""Developer probe: K2 filter statistics (needs libvo_b200.so built with -DVO_K2_STATS)."""
import sys, os, ctypes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from hls_final_visual_odometry_b200 import Engine
from hls_final_visual_odometry_b200.build import LIB
from hls_final_visual_odometry_b200.synth import make_sequence
W, H, B = 1241, 376, 8
seq = make_sequence(W, H, B + 1, seed=1)
eng = Engine(W, H, max_frames=B + 1, max_pairs=B)
lib = ctypes.CDLL(LIB)
out = (ctypes.c_ulonglong * 8)()
eng.batch_extract_host(0, seq)
lib.vo_debug_k2_stats(out)
prev = np.arange(B, dtype=np.int32); curr = prev + 1
eng.batch_match(prev, curr, 2, 50.0, 50.0)
eng.synchronize()
lib.vo_debug_k2_stats(out)
names = ["origins", "with rows", "seed none", "seed T==0", "filter hits (groups)", "budget handovers", "-", "-"]
for n, v in zip(names, out):
    print(f"{n:24s} {v / B:12.1f} per pair")
