#!/usr/bin/env python
"""Developer A/B builds of libvo_b200.so with extra -D flags, kept under build/ab/ (git-ignored, but shipped to the GPU box):

    python tools/ab_build.py base "" flood8 "-DVO_K2_FLOOD=8" append2 "-DVO_K2_APPEND=2"
    VO_B200_LIB=build/ab/libvo_flood8.so python tools/noise_probe.py 0 2 8
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from hls_final_visual_odometry_b200 import build as b  # noqa: E402


def one(tag, flags):
    out = os.path.join(ROOT, "build", "ab", f"libvo_{tag}.so")
    os.makedirs(os.path.dirname(out), exist_ok=True)
    srcs = b._sources(b.CSRC, (".cu", ".cpp"))
    subprocess.check_call([b._nvcc(), *b.NVCC_FLAGS, *flags.split(), "-shared", "-o", out, *srcs])
    return out


if __name__ == "__main__":
    pairs = list(zip(sys.argv[1::2], sys.argv[2::2]))
    with ThreadPoolExecutor(max_workers=4) as ex:
        for o in ex.map(lambda p: one(*p), pairs):
            print(o)
