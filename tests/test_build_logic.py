"""Build plumbing that a GPU box depends on (no GPU needed): a library is rebuilt when the CONTENT of its sources changes, not
when file times do (a snapshot to a fresh box scrambles them: eight ranks once rebuilt libvo_b200.so at the same moment and one
mapped a half-written file), builders write to a private file and rename it into place, and the hash that ties
profiles/traffic.json to the kernels covers the device sources only."""
import os
import time

import bench
from hls_final_visual_odometry_b200 import build


def _deps():
    srcs = build._sources(build.CSRC, (".cu", ".cpp"))
    return srcs + build._sources(build.CSRC, (".cuh", ".h")) + [os.path.join(build._ROOT, "include", "vo_b200.h")]


def test_staleness_follows_content_not_file_times():
    lib = build.build_library()
    assert os.path.exists(lib + ".srchash"), "a built library carries the digest of its sources"
    deps = _deps()
    assert not build._stale(lib, deps, build.NVCC_FLAGS)
    src = os.path.join(build.CSRC, "vo_common.cuh")
    st = os.stat(src)
    try:
        os.utime(src, (time.time() + 3600, time.time() + 3600))  # newer than the library, same bytes
        assert not build._stale(lib, deps, build.NVCC_FLAGS)
    finally:
        os.utime(src, (st.st_atime, st.st_mtime))
    assert build._stale(lib, deps, build.NVCC_FLAGS + ["-DSOMETHING_ELSE"]), "other flags are another build"
    assert build._digest(deps, build.NVCC_FLAGS) != build._digest(deps[:-1], build.NVCC_FLAGS)


def test_no_partial_files_are_left_behind():
    build.build_library()
    build.build_host()
    left = [f for f in os.listdir(build._PKG) if f.endswith(".tmp")]
    assert not left, left


def test_traffic_hash_covers_device_sources_only():
    import hashlib
    base = build.CSRC
    files = sorted(os.path.join(base, f) for f in os.listdir(base) if f.endswith((".cu", ".cuh", ".h")))
    h = hashlib.sha256()
    for f in files + [os.path.join(build._ROOT, "include", "vo_b200.h")]:
        h.update(os.path.basename(f).encode())
        h.update(open(f, "rb").read())
    assert bench.source_hash() == h.hexdigest()[:16]
    assert not any(f.endswith(".cpp") for f in files)
