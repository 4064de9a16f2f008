"""TEST INFRASTRUCTURE ONLY -- ctypes loader for oracle/_ref/libvoref_*.so (the UNMODIFIED
reference sources compiled by oracle/build_ref.sh).  Never imported by the product package.

Layouts (reference):
  max2      int32[V_BIN_NUM][U_BIN_NUM][4][MAX_FP_IN_BIN][12]   /root/reference/src/size.h:17-20
            record = {u, v, val(=0), class, 8 x int32 descriptor words}
            /root/reference/src/myComputeFeature.cpp:262-270
  max2_num  int32[V_BIN_NUM][U_BIN_NUM][4]
  p_match   12 x 4 bytes {f u1p,v1p; i i1p; f u2p,v2p; i i2p; f u1c,v1c; i i1c; f u2c,v2c; i i2c}
            /root/reference/src/matcher.h:89-104
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))

P_MATCH_DTYPE = np.dtype([
    ("u1p", "<f4"), ("v1p", "<f4"), ("i1p", "<i4"),
    ("u2p", "<f4"), ("v2p", "<f4"), ("i2p", "<i4"),
    ("u1c", "<f4"), ("v1c", "<f4"), ("i1c", "<i4"),
    ("u2c", "<f4"), ("v2c", "<f4"), ("i2c", "<i4"),
])
assert P_MATCH_DTYPE.itemsize == 48


def ref_path(tag: str) -> str:
    return os.path.join(_HERE, "_ref", f"libvoref_{tag}.so")


def have_ref(tag: str) -> bool:
    return os.path.exists(ref_path(tag))


def _p(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


class RefLib:
    """One compiled geometry of the reference (tag = 'native' or '<W>x<H>')."""

    def __init__(self, tag: str):
        self.tag = tag
        self.lib = C.CDLL(ref_path(tag))
        info = (C.c_int32 * 10)()
        self.lib.ref_info(info)
        (self.W, self.H, self.BIN_W, self.BIN_H, self.U, self.V, self.CAP, self.COL_MAX,
         self.POINT_L, psz) = list(info)
        assert psz == 48
        self.BIN_NUM = self.U * self.V * 4
        self.lib.ref_rand_number.restype = C.c_uint32
        self.lib.ref_rand_number.argtypes = [C.c_uint32]
        self.lib.ref_matcher_new.restype = C.c_void_p
        self.lib.ref_matcher_num_matches.restype = C.c_int32
        self.lib.ref_matcher_num_matches.argtypes = [C.c_void_p]

    # ---- filters -------------------------------------------------------------------------
    def sobel(self, img: np.ndarray):
        img = np.ascontiguousarray(img, dtype=np.uint8)
        assert img.shape == (self.H, self.W)
        du = np.zeros_like(img)
        dv = np.zeros_like(img)
        self.lib.ref_sobel5x5(_p(img), _p(du), _p(dv))
        return du, dv

    def blob(self, img: np.ndarray):
        img = np.ascontiguousarray(img, dtype=np.uint8)
        out = np.zeros(img.shape, dtype=np.int16)
        self.lib.ref_blob5x5(_p(img), _p(out))
        return out

    def checkerboard(self, img: np.ndarray):
        img = np.ascontiguousarray(img, dtype=np.uint8)
        out = np.zeros(img.shape, dtype=np.int16)
        self.lib.ref_checkerboard5x5(_p(img), _p(out))
        return out

    # ---- the three operators -------------------------------------------------------------
    def new_feature_arrays(self):
        max2 = np.zeros((self.V, self.U, 4, self.CAP, 12), dtype=np.int32)
        num = np.zeros((self.V, self.U, 4), dtype=np.int32)
        return max2, num

    def compute_features(self, img: np.ndarray):
        img = np.ascontiguousarray(img, dtype=np.uint8)
        assert img.shape == (self.H, self.W), (img.shape, self.H, self.W)
        max2, num = self.new_feature_arrays()
        self.lib.ref_compute_features(_p(img), _p(max2), _p(num))
        return max2, num

    def matching(self, m1p, n1p, m1c, n1c, cap: int = 1 << 17):
        out = np.zeros(cap, dtype=P_MATCH_DTYPE)
        n = C.c_int32(0)
        self.lib.ref_matching(_p(m1p), _p(m1c), _p(n1p), _p(n1c), _p(out), C.byref(n))
        return out[: n.value].copy()

    def remove_outliers(self, matches: np.ndarray):
        assert len(matches) <= self.POINT_L
        buf = np.zeros(self.POINT_L, dtype=P_MATCH_DTYPE)
        buf[: len(matches)] = matches
        n = C.c_int32(len(matches))
        self.lib.ref_remove_outliers(_p(buf), C.byref(n))
        return buf[: n.value].copy()

    def bucket_features(self, matches: np.ndarray, max_features=2, bw=50.0, bh=50.0):
        assert len(matches) <= self.POINT_L
        buf = np.zeros(self.POINT_L, dtype=P_MATCH_DTYPE)
        buf[: len(matches)] = matches
        n = C.c_int32(len(matches))
        self.lib.ref_bucket_features(_p(buf), C.byref(n), C.c_int32(max_features),
                                     C.c_float(bw), C.c_float(bh))
        return buf[: n.value].copy()

    def rand_number(self, x: int) -> int:
        return int(self.lib.ref_rand_number(C.c_uint32(x)))

    # ---- Matcher facade ------------------------------------------------------------------
    def matcher(self):
        return RefMatcher(self)


class RefMatcher:
    def __init__(self, ref: RefLib):
        self.ref = ref
        self.h = C.c_void_p(ref.lib.ref_matcher_new())

    def close(self):
        if self.h:
            self.ref.lib.ref_matcher_delete(self.h)
            self.h = None

    def push_back(self, img: np.ndarray, replace=False):
        img = np.ascontiguousarray(img, dtype=np.uint8)
        dims = (C.c_int32 * 3)(img.shape[1], img.shape[0], img.shape[1])
        self._keep = img
        self.ref.lib.ref_matcher_push_back(self.h, _p(img), dims, C.c_int(int(replace)))

    def match_features(self, method=0):
        self.ref.lib.ref_matcher_match_features(self.h, C.c_int(method))

    def bucket_features(self, max_features=2, bw=50.0, bh=50.0):
        self.ref.lib.ref_matcher_bucket_features(self.h, C.c_int32(max_features), C.c_float(bw),
                                                 C.c_float(bh))

    def get_matches(self):
        n = self.ref.lib.ref_matcher_num_matches(self.h)
        out = np.zeros(max(n, 1), dtype=P_MATCH_DTYPE)
        self.ref.lib.ref_matcher_get_matches(self.h, _p(out))
        return out[:n].copy()

    def dump_features(self, which: int):
        max2, num = self.ref.new_feature_arrays()
        self.ref.lib.ref_matcher_dump_features(self.h, C.c_int(which), _p(max2), _p(num))
        return max2, num


def mono_motion_seq(ref: "RefLib", lists, f, cu, cv, height=1.0, pitch=0.0, ransac_iters=2000, inlier_threshold=1e-5,
                    motion_threshold=100.0):
    """VisualOdometryMono::estimateMotion (/root/reference/src/viso_mono.cpp:41) on a sequence of match lists
    processed by ONE reference object (srand(0) once, rand() stream carried over).  Returns (tr[n,6], ok[n], n_inliers[n])."""
    lists = [np.ascontiguousarray(m, dtype=P_MATCH_DTYPE) for m in lists]
    counts = np.array([len(m) for m in lists], dtype=np.int32)
    flat = np.concatenate(lists) if lists else np.zeros(0, dtype=P_MATCH_DTYPE)
    prm = np.array([f, cu, cv, height, pitch, ransac_iters, inlier_threshold, motion_threshold], dtype=np.float64)
    tr = np.zeros((len(lists), 6), dtype=np.float64)
    ok = np.zeros(len(lists), dtype=np.int32)
    ninl = np.zeros(len(lists), dtype=np.int32)
    ref.lib.ref_mono_motion_seq(_p(flat), _p(counts), C.c_int32(len(lists)), _p(prm), _p(tr), _p(ok), _p(ninl))
    return tr, ok, ninl


def stereo_motion_seq(ref: "RefLib", lists, f, cu, cv, base=1.0, ransac_iters=200, inlier_threshold=2.0, reweighting=True):
    """VisualOdometryStereo::estimateMotion (/root/reference/src/viso_stereo.cpp:54) on a sequence of quad match lists
    processed by ONE reference object.  Returns (tr[n,6], ok[n], n_inliers[n])."""
    lists = [np.ascontiguousarray(m, dtype=P_MATCH_DTYPE) for m in lists]
    counts = np.array([len(m) for m in lists], dtype=np.int32)
    flat = np.concatenate(lists) if lists else np.zeros(0, dtype=P_MATCH_DTYPE)
    prm = np.array([f, cu, cv, base, ransac_iters, inlier_threshold, 1.0 if reweighting else 0.0], dtype=np.float64)
    tr = np.zeros((len(lists), 6), dtype=np.float64)
    ok = np.zeros(len(lists), dtype=np.int32)
    ninl = np.zeros(len(lists), dtype=np.int32)
    ref.lib.ref_stereo_motion_seq(_p(flat), _p(counts), C.c_int32(len(lists)), _p(prm), _p(tr), _p(ok), _p(ninl))
    return tr, ok, ninl
