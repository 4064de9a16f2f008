"""CPU: the C-ABI library loads and exports every symbol include/vo_b200.h declares; without a GPU
the product path fails loudly instead of falling back to anything."""
import ctypes as C
import os
import re

import pytest

from hls_final_visual_odometry_b200 import api, build

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_functions():
    text = open(os.path.join(ROOT, "include", "vo_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(vo_[a-z0-9_]+)\s*\(", text)))


def test_header_and_binding_agree():
    assert header_functions() == sorted(api.EXPORTS)


def test_library_exports_every_declared_symbol():
    path = build.build_library()
    lib = C.CDLL(path)
    for name in header_functions():
        assert hasattr(lib, name), f"{name} declared in include/vo_b200.h but not exported"


def test_struct_sizes_match_header():
    assert C.sizeof(api.VoConfig) == 16 * 4
    assert C.sizeof(api.VoGeometry) == 16 * 4
    assert api.P_MATCH_DTYPE.itemsize == 48


def test_no_gpu_means_loud_failure():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(api.VoError) as ei:
        api.Engine(1241, 376)
    assert ei.value.status == api.VO_ERR_CUDA
    assert "no CPU fallback" in str(ei.value)


def test_argument_validation_happens_before_cuda():
    lib = api.load_library()
    h = C.c_void_p()
    assert lib.vo_engine_create(None, 0, None, C.byref(h)) == api.VO_ERR_INVALID
    cfg = api.VoConfig(width=32, height=32)
    assert lib.vo_engine_create(C.byref(cfg), 0, None, C.byref(h)) == api.VO_ERR_INVALID
    cfg = api.VoConfig(width=1241, height=376, bin_w=50, bin_h=50)
    assert lib.vo_engine_create(C.byref(cfg), 0, None, C.byref(h)) == api.VO_ERR_UNSUPPORTED
    assert b"60x45" in lib.vo_last_error()


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "hls_final_visual_odometry_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                src = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "oracle_lib" not in src and "liboracle" not in src and "ref_lib" not in src, f
