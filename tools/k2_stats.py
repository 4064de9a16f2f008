"""Developer probe: K2 filter statistics (needs libvo_b200.so built with -DVO_K2_STATS)."""
import sys, os, ctypes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from hls_final_visual_odometry_b200 import Engine
from hls_final_visual_odometry_b200.build import LIB
from hls_final_visual_odometry_b200.synth import make_sequence
W, H, B = 1241, 376, 8
amp = int(sys.argv[1]) if len(sys.argv) > 1 else 0   # independent sensor noise of +-amp grey levels per frame
seq = make_sequence(W, H, B + 1, seed=1)
if amp:
    rng = np.random.default_rng(100 + amp)
    seq = np.clip(seq.astype(np.int16) + rng.integers(-amp, amp + 1, seq.shape, dtype=np.int16), 0, 255).astype(np.uint8)
eng = Engine(W, H, max_frames=B + 1, max_pairs=B)
lib = ctypes.CDLL(os.environ.get("VO_B200_LIB", LIB))
out = (ctypes.c_ulonglong * 16)()
eng.batch_extract_host(0, seq)
lib.vo_debug_k2_stats(out)
prev = np.arange(B, dtype=np.int32); curr = prev + 1
eng.batch_match(prev, curr, 2, 50.0, 50.0)
eng.synchronize()
lib.vo_debug_k2_stats(out)
names = ["origins (per chunk)", "Z: bound 0", "N: bound <= seed_good", "P: exhaustive", "M tasks: 32-candidate words", "M tasks: filter hits", "M tasks: sum over words of the busiest lane's hits", "M tasks: sum over words of lanes with a hit",
         "Z/N tasks: 32-candidate words", "Z/N tasks: filter hits", "Z/N tasks: sum over words of the busiest lane's hits",
         "Z/N tasks: words with a hit", "Z/N tasks: two-word words", "task joins (warp, task)", "-", "-"]
print(f"noise +-{amp}")
for n, v in zip(names, out):
    print(f"{n:24s} {v / B:12.1f} per pair")
