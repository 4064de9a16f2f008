"""In-tree build of libvo_b200.so (hand-written CUDA for sm_100a + the C ABI of include/vo_b200.h)
and of the C++ host shim (host/libvo_host.so: drop-in Matcher / VisualOdometryMono classes).

nvcc cross-compiles without a GPU, so this runs on the CPU build box; the resulting .so files are
git-ignored but travel to the GPU box with the repository snapshot.
"""
from __future__ import annotations

import os
import shutil
import subprocess

_PKG = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_PKG)
CSRC = os.path.join(_PKG, "csrc")
HOST = os.path.join(_PKG, "host")
LIB = os.path.join(_PKG, "libvo_b200.so")
HOST_LIB = os.path.join(_PKG, "libvo_host.so")
DEMO = os.path.join(_PKG, "vo_demo")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
    # no FMA contraction in device code either: vo_ransac.cu scores RANSAC hypotheses with the host stage's own arithmetic
    # (vo_ransac_core.h) and must land on the same inlier counts; the other kernels are integer code
    "-fmad=false",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-Wall",
    # the host Delaunay stage (vo_outliers.cpp) must round exactly like the reference's x86 build
    "-Xcompiler", "-ffp-contract=off",
]


def _nvcc() -> str:
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def _sources(d: str, exts) -> list[str]:
    return sorted(os.path.join(d, f) for f in os.listdir(d) if f.endswith(exts))


def _digest(deps: list[str], flags: list[str]) -> str:
    import hashlib
    h = hashlib.sha256(" ".join(flags).encode())
    for d in sorted(deps):
        if os.path.exists(d):
            h.update(os.path.basename(d).encode())
            h.update(open(d, "rb").read())
    return h.hexdigest()


def _stale(target: str, deps: list[str], flags: list[str] | None = None) -> bool:
    """A target is current when the digest of its sources (written beside it when it was built) matches: file times do not
    survive the snapshot that carries the tree to a GPU box, contents do.  Without a digest file: modification times."""
    if not os.path.exists(target):
        return True
    side = target + ".srchash"
    if flags is not None and os.path.exists(side):
        return open(side).read().strip() != _digest(deps, flags)
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def _stamp(target: str, deps: list[str], flags: list[str]) -> None:
    tmp = f"{target}.srchash.{os.getpid()}.tmp"
    with open(tmp, "w") as f:
        f.write(_digest(deps, flags))
    os.replace(tmp, target + ".srchash")


def build_library(force: bool = False, verbose: bool = False) -> str:
    srcs = _sources(CSRC, (".cu", ".cpp"))
    deps = srcs + _sources(CSRC, (".cuh", ".h")) + [os.path.join(_ROOT, "include", "vo_b200.h")]
    if force or _stale(LIB, deps, NVCC_FLAGS):
        # several ranks of one job may find the library stale at the same time (a fresh snapshot can carry sources newer than
        # the .so): every builder writes its own file and renames it into place, so a reader never maps a half-written library
        tmp = f"{LIB}.{os.getpid()}.tmp"
        cmd = [_nvcc(), *NVCC_FLAGS, "-shared", "-o", tmp, *srcs]
        if verbose:
            cmd.insert(1, "-Xptxas")
            cmd.insert(2, "-v")
        try:
            subprocess.check_call(cmd)
            os.replace(tmp, LIB)
            _stamp(LIB, deps, NVCC_FLAGS)
        finally:
            if os.path.exists(tmp):
                os.remove(tmp)
    return LIB


def build_host(force: bool = False) -> str | None:
    if not os.path.isdir(HOST):
        return None
    srcs = [s for s in _sources(HOST, (".cpp",)) if not s.endswith("demo_main.cpp")]
    if not srcs:
        return None
    deps = srcs + _sources(HOST, (".h",)) + [LIB]
    inc = ["-I", os.path.join(_ROOT, "include"), "-I", HOST]
    if force or _stale(HOST_LIB, deps):
        tmp = f"{HOST_LIB}.{os.getpid()}.tmp"
        try:
            subprocess.check_call([
                "g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-Wall", *inc,
                "-o", tmp, *srcs, "-L", _PKG, "-lvo_b200", "-lz", "-Wl,-rpath,$ORIGIN",
            ])
            os.replace(tmp, HOST_LIB)
        finally:
            if os.path.exists(tmp):
                os.remove(tmp)
    demo_src = os.path.join(HOST, "demo_main.cpp")
    if os.path.exists(demo_src) and (force or _stale(DEMO, deps + [demo_src, HOST_LIB])):
        subprocess.check_call([
            "g++", "-O2", "-std=c++17", "-Wall", *inc, "-o", DEMO, demo_src, "-L", _PKG, "-lvo_host", "-lvo_b200", "-lz",
            "-Wl,-rpath,$ORIGIN",
        ])
    return HOST_LIB


def build_all(force: bool = False) -> None:
    build_library(force)
    build_host(force)


if __name__ == "__main__":
    import sys
    build_library(force="--force" in sys.argv, verbose="-v" in sys.argv)
    build_host(force="--force" in sys.argv)
    print(LIB)
