"""TEST INFRASTRUCTURE ONLY -- ctypes loader for oracle/liboracle.so (vo_oracle.c, the plain-C
restatement of the reference hot path).  Allowed importers: tests/, __graft_entry__.smoke(),
bench.py's cpu_baseline / `--impl reference` legs.  Never imported by the product package.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

from .ref_lib import P_MATCH_DTYPE

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liboracle.so")


class Geom(C.Structure):
    _fields_ = [(n, C.c_int32) for n in
                ("width", "height", "pitch", "bin_w", "bin_h", "u_bins", "v_bins", "cap", "tau", "rad_u", "rad_v")]


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "vo_oracle.c")
    if force or not os.path.exists(_SO) or (
            os.path.exists(src) and os.path.getmtime(src) > os.path.getmtime(_SO)):
        # (written to a private file and renamed into place: the CPU legs of bench.py load this from many processes at once)
        tmp = f"{_SO}.{os.getpid()}.tmp"
        try:
            subprocess.check_call(["gcc", "-O2", "-fPIC", "-std=c99", "-shared", "-o", tmp, src, "-lm"])
            os.replace(tmp, _SO)
        finally:
            if os.path.exists(tmp):
                os.remove(tmp)
    return _SO


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


class Oracle:
    def __init__(self, width: int, height: int, bin_w: int = 60, bin_h: int = 45, rad_u: int = 0, rad_v: int = 0):
        self.lib = C.CDLL(build())
        self.g = Geom()
        self.lib.vo_oracle_geom_init_bins(C.byref(self.g), width, height, width, bin_w, bin_h)
        if rad_u > 0:  # SEARCH_RAD_U / SEARCH_RAD_V other than BIN_W*3 / BIN_H*3 (src/size.h:14-15): the live search radius
            self.g.rad_u = rad_u
        if rad_v > 0:
            self.g.rad_v = rad_v
        self.W, self.H = width, height
        self.U, self.V, self.CAP = self.g.u_bins, self.g.v_bins, self.g.cap
        self.lib.vo_oracle_rand_number.restype = C.c_uint32
        self.lib.vo_oracle_rand_number.argtypes = [C.c_uint32]
        self.lib.vo_oracle_compute_features.restype = C.c_int32
        self.lib.vo_oracle_matching.restype = C.c_int32

    def _img(self, img):
        img = np.ascontiguousarray(img, dtype=np.uint8)
        assert img.shape == (self.H, self.W)
        return img

    def sobel(self, img):
        img = self._img(img)
        du = np.zeros_like(img); dv = np.zeros_like(img)
        self.lib.vo_oracle_sobel5x5(_p(img), self.W, self.H, self.W, _p(du), _p(dv))
        return du, dv

    def blob(self, img):
        img = self._img(img)
        out = np.zeros(img.shape, np.int16)
        self.lib.vo_oracle_blob5x5(_p(img), self.W, self.H, self.W, _p(out))
        return out

    def checkerboard(self, img):
        img = self._img(img)
        out = np.zeros(img.shape, np.int16)
        self.lib.vo_oracle_checkerboard5x5(_p(img), self.W, self.H, self.W, _p(out))
        return out

    def compute_features(self, img):
        """-> (max2 int32[V,U,4,CAP,12], num int32[V,U,4], dropped)"""
        img = self._img(img)
        max2 = np.zeros((self.V, self.U, 4, self.CAP, 12), np.int32)
        num = np.zeros((self.V, self.U, 4), np.int32)
        dropped = self.lib.vo_oracle_compute_features(_p(img), C.byref(self.g), _p(max2), _p(num))
        return max2, num, int(dropped)

    def matching(self, m1p, n1p, m1c, n1c, cap: int = 1 << 17):
        """-> (matches, stats[3] = {sad evaluations, zero-candidate origins, range tests})"""
        out = np.zeros(cap, P_MATCH_DTYPE)
        n = C.c_int32(0)
        stats = np.zeros(3, np.int64)
        ovf = self.lib.vo_oracle_matching(C.byref(self.g), _p(m1p), _p(n1p), _p(m1c), _p(n1c), _p(out),
                                          C.c_int32(cap), C.byref(n), _p(stats))
        assert ovf == 0
        return out[: n.value].copy(), stats

    def bucket_features(self, matches, max_features=2, bw=50.0, bh=50.0):
        buf = np.array(matches, dtype=P_MATCH_DTYPE, copy=True)
        if len(buf) == 0:
            buf = np.zeros(1, P_MATCH_DTYPE)
        n = C.c_int32(len(matches))
        self.lib.vo_oracle_bucket_features(_p(buf), C.byref(n), C.c_int32(max_features), C.c_float(bw),
                                           C.c_float(bh))
        return buf[: n.value].copy()

    # ---- EXTENSION, parity unpinned (vo_oracle.c "EXTENSION" section) ----
    def quad_matching(self, f1p, f2p, f1c, f2c, disp_tolerance: int = 2, cap: int = 1 << 17):
        """matchFeatures(method=2) on four feature sets (max2, num) -> matches with all 8 coordinates"""
        out = np.zeros(cap, P_MATCH_DTYPE)
        n = C.c_int32(0)
        self.lib.vo_oracle_quad_matching.restype = C.c_int32
        ovf = self.lib.vo_oracle_quad_matching(C.byref(self.g), _p(f1p[0]), _p(f1p[1]), _p(f2p[0]), _p(f2p[1]),
                                               _p(f1c[0]), _p(f1c[1]), _p(f2c[0]), _p(f2c[1]),
                                               C.c_int32(disp_tolerance), _p(out), C.c_int32(cap), C.byref(n))
        assert ovf == 0
        return out[: n.value].copy()

    def refine(self, matches, img_1p, img_1c, img_2p=None, img_2c=None):
        """parabolic sub-pixel refinement (param.refinement == 2) of a flow or quad list; drops rejected matches"""
        buf = np.array(matches, dtype=P_MATCH_DTYPE, copy=True)
        if len(buf) == 0:
            return buf
        n = C.c_int32(len(buf))
        imgs = [None if i is None else self._img(i) for i in (img_1p, img_2p, img_1c, img_2c)]
        self.lib.vo_oracle_refine(C.byref(self.g), *[C.c_void_p(None) if i is None else _p(i) for i in imgs],
                                  _p(buf), C.byref(n))
        return buf[: n.value].copy()

    def rand_number(self, x: int) -> int:
        return int(self.lib.vo_oracle_rand_number(C.c_uint32(x)))


def compact_features(max2: np.ndarray, num: np.ndarray):
    """Layout-independent view of a reference feature array: list over (v,u,c) of (n,12) records."""
    V, U, _ = num.shape
    out = []
    for v in range(V):
        for u in range(U):
            for c in range(4):
                out.append(np.ascontiguousarray(max2[v, u, c, : num[v, u, c]]))
    return out


def features_digest(max2: np.ndarray, num: np.ndarray) -> str:
    """sha256 over counts + payload records in (v_bin,u_bin,class,slot) order (stride independent)."""
    import hashlib
    h = hashlib.sha256()
    h.update(np.ascontiguousarray(num, dtype="<i4").tobytes())
    for rec in compact_features(max2, num):
        h.update(np.ascontiguousarray(rec, dtype="<i4").tobytes())
    return h.hexdigest()


def matches_digest(m: np.ndarray) -> str:
    import hashlib
    return hashlib.sha256(np.ascontiguousarray(m).tobytes()).hexdigest()
