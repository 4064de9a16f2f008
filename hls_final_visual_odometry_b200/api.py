"""ctypes binding of libvo_b200.so (include/vo_b200.h) + a Python mirror of the reference's Matcher.

This module is host-side plumbing only: every operator call goes through the C ABI into
hand-written sm_100a kernels.  There is no Python/numpy/torch implementation of any operator here
and no CPU fallback: if the library is missing or no B200 is present, calls raise.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import build as _build

P_MATCH_DTYPE = np.dtype([
    ("u1p", "<f4"), ("v1p", "<f4"), ("i1p", "<i4"),
    ("u2p", "<f4"), ("v2p", "<f4"), ("i2p", "<i4"),
    ("u1c", "<f4"), ("v1c", "<f4"), ("i1c", "<i4"),
    ("u2c", "<f4"), ("v2c", "<f4"), ("i2c", "<i4"),
])  # == Matcher::p_match, /root/reference/src/matcher.h:89-104

VO_OK, VO_ERR_INVALID, VO_ERR_UNSUPPORTED, VO_ERR_CUDA, VO_ERR_CAPACITY, VO_ERR_STATE = range(6)


class VoError(RuntimeError):
    def __init__(self, status: int, msg: str):
        super().__init__(f"vo_b200 status {status}: {msg}")
        self.status = status


class VoConfig(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("bin_w", C.c_int32), ("bin_h", C.c_int32),
                ("nms_tau", C.c_int32), ("max_frames", C.c_int32), ("max_pairs", C.c_int32),
                ("max_matches", C.c_int32), ("match_rows_per_cta", C.c_int32), ("match_exhaustive", C.c_int32),
                ("match_threads", C.c_int32), ("stereo", C.c_int32), ("match_disp_tolerance", C.c_int32),
                ("match_seed_good", C.c_int32), ("match_radius_u", C.c_int32), ("match_radius_v", C.c_int32)]


class VoGeometry(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("pitch", C.c_int32), ("bin_w", C.c_int32),
                ("bin_h", C.c_int32), ("u_bins", C.c_int32), ("v_bins", C.c_int32), ("bin_cap", C.c_int32),
                ("col_cap", C.c_int32), ("bin_num", C.c_int32), ("max_frames", C.c_int32),
                ("max_pairs", C.c_int32), ("max_matches", C.c_int32), ("reserved", C.c_int32 * 3)]


# every symbol include/vo_b200.h declares (tests/test_abi.py checks the header against this list)
EXPORTS = [
    "vo_engine_create", "vo_engine_destroy", "vo_engine_geometry", "vo_engine_synchronize", "vo_remove_outliers_host",
    "vo_remove_outliers", "vo_fence_record",
    "vo_fence_wait", "vo_last_error", "vo_version", "vo_compute_features", "vo_matching", "vo_matching_iter",
    "vo_remove_outliers_bucket", "vo_bucket_matches", "vo_push_back", "vo_match_features",
    "vo_bucket_features", "vo_get_matches", "vo_num_matches", "vo_set_matches", "vo_dump_features",
    "vo_batch_extract_host", "vo_batch_extract_device", "vo_batch_match", "vo_batch_counts",
    "vo_batch_get_matches", "vo_batch_download_async", "vo_slot_dump_features", "vo_slot_feature_count",
    "vo_launch_count", "vo_profile_enable", "vo_profile_read", "vo_host_alloc", "vo_host_free",
    "vo_egomotion_create", "vo_egomotion_destroy", "vo_estimate_motion_mono", "vo_egomotion_inliers",
    "vo_egomotion_rand", "vo_transformation_vector_to_matrix", "vo_egomotion_draw_samples",
    "vo_estimate_motion_mono_samples", "vo_batch_bucket_host", "vo_pipeline_create", "vo_pipeline_destroy",
    "vo_pipeline_process", "vo_pipeline_stage_seconds", "vo_engine_device", "vo_egomotion_draw_samples_batch", "vo_estimate_motion_mono_best",
    "vo_ransac_gpu_create", "vo_ransac_gpu_destroy", "vo_ransac_gpu_score", "vo_estimate_motion_stereo", "vo_batch_match_quad", "vo_batch_refine", "vo_refine_matches", "vo_profile_read_ext", "vo_descriptor_signature",
    "vo_bucket_matches_noalloc",
]

class VoStereoParams(C.Structure):  # vo_stereo_params: /root/reference/src/viso_stereo.h:31-43
    _fields_ = [("f", C.c_double), ("cu", C.c_double), ("cv", C.c_double), ("base", C.c_double),
                ("ransac_iters", C.c_int32), ("reweighting", C.c_int32), ("inlier_threshold", C.c_double)]


class VoMonoParams(C.Structure):  # vo_mono_params: /root/reference/src/viso_mono.h:32-46, src/viso.h:56-64
    _fields_ = [("f", C.c_double), ("cu", C.c_double), ("cv", C.c_double), ("height", C.c_double), ("pitch", C.c_double),
                ("ransac_iters", C.c_int32), ("reserved", C.c_int32), ("inlier_threshold", C.c_double),
                ("motion_threshold", C.c_double)]


_lib = None


def load_library() -> C.CDLL:
    """Load libvo_b200.so (building it in-tree first if nvcc is available and sources are newer)."""
    global _lib
    if _lib is not None:
        return _lib
    path = _build.LIB
    alt = os.environ.get("VO_B200_LIB")  # developer A/B switch: another build of the same sources (tools/ab_build.py)
    if alt:
        path = alt
    else:
        try:
            path = _build.build_library()
        except Exception:
            if not os.path.exists(path):
                raise
    lib = C.CDLL(path)
    lib.vo_last_error.restype = C.c_char_p
    lib.vo_version.restype = C.c_char_p
    lib.vo_engine_destroy.restype = None
    vp, i32, i64, f32 = C.c_void_p, C.c_int32, C.c_int64, C.c_float
    sigs = {
        "vo_engine_create": [C.POINTER(VoConfig), i32, vp, C.POINTER(vp)],
        "vo_engine_destroy": [vp],
        "vo_descriptor_signature": [vp, vp],
        "vo_engine_geometry": [vp, C.POINTER(VoGeometry)],
        "vo_engine_synchronize": [vp],
        "vo_compute_features": [vp, vp, i32, vp, vp, i32],
        "vo_matching": [vp, vp, vp, vp, vp, i32, vp, i32, C.POINTER(i32)],
        "vo_matching_iter": [vp, vp, vp, vp, vp, i32, vp, i32, C.POINTER(i32), i32],
        "vo_remove_outliers_bucket": [vp, vp, C.POINTER(i32), vp, i32, C.POINTER(i32)],
        "vo_bucket_matches": [vp, vp, i32, i32, f32, f32, vp, i32, C.POINTER(i32)],
        "vo_bucket_matches_noalloc": [vp, vp, i32, i32, f32, f32, vp, i32, vp],
        "vo_push_back": [vp, vp, vp, C.POINTER(i32), i32],
        "vo_match_features": [vp, i32],
        "vo_bucket_features": [vp, i32, f32, f32],
        "vo_get_matches": [vp, vp, i32, C.POINTER(i32)],
        "vo_num_matches": [vp, C.POINTER(i32)],
        "vo_set_matches": [vp, vp, i32],
        "vo_dump_features": [vp, i32, vp, vp, i32],
        "vo_batch_extract_host": [vp, i32, i32, vp, i64, i32],
        "vo_batch_extract_device": [vp, i32, i32, vp, i64, i32],
        "vo_batch_match": [vp, i32, vp, vp, i32, f32, f32],
        "vo_batch_counts": [vp, i32, vp, vp],
        "vo_batch_get_matches": [vp, i32, i32, vp, i32, C.POINTER(i32)],
        "vo_batch_download_async": [vp, i32, i32, vp, vp, i64, i32],
        "vo_slot_dump_features": [vp, i32, vp, vp, i32],
        "vo_slot_feature_count": [vp, i32, C.POINTER(i32)],
        "vo_launch_count": [vp, C.POINTER(i64)],
        "vo_profile_enable": [vp, i32],
        "vo_profile_read": [vp, C.POINTER(C.c_double), C.POINTER(i64), i32],
        "vo_remove_outliers_host": [vp, C.POINTER(i32)],
        "vo_remove_outliers": [vp],
        "vo_fence_record": [vp, i32],
        "vo_fence_wait": [vp, i32],
        "vo_host_alloc": [C.POINTER(vp), i64],
        "vo_host_free": [vp],
        "vo_egomotion_create": [C.c_uint32, i32, C.POINTER(vp)],
        "vo_egomotion_destroy": [vp],
        "vo_estimate_motion_mono": [vp, vp, i32, C.POINTER(VoMonoParams), C.POINTER(C.c_double), C.POINTER(i32)],
        "vo_egomotion_inliers": [vp, vp, i32, C.POINTER(i32)],
        "vo_egomotion_rand": [vp],
        "vo_transformation_vector_to_matrix": [C.POINTER(C.c_double), C.POINTER(C.c_double)],
        "vo_egomotion_draw_samples": [vp, vp, i32, C.POINTER(VoMonoParams), vp, C.POINTER(i32)],
        "vo_estimate_motion_mono_samples": [vp, i32, C.POINTER(VoMonoParams), vp, i32, i32, C.POINTER(C.c_double), C.POINTER(i32),
                                            vp, i32, C.POINTER(i32)],
        "vo_estimate_motion_mono_best": [vp, i32, C.POINTER(VoMonoParams), vp, i32, i32, C.POINTER(C.c_double), C.POINTER(i32),
                                         vp, i32, C.POINTER(i32)],
        "vo_egomotion_draw_samples_batch": [vp, i32, vp, i64, vp, C.POINTER(VoMonoParams), vp, i64, vp, i32],
        "vo_ransac_gpu_create": [i32, i32, i32, i32, C.POINTER(vp)],
        "vo_ransac_gpu_destroy": [vp],
        "vo_ransac_gpu_score": [vp, i32, vp, i64, vp, C.POINTER(VoMonoParams), vp, i64, vp, vp, vp],
        "vo_engine_device": [vp, C.POINTER(i32)],
        "vo_batch_bucket_host": [vp, i32, vp, i64, vp, i32, f32, f32],
        "vo_pipeline_create": [vp, C.POINTER(VoMonoParams), i32, f32, f32, i32, C.POINTER(vp)],
        "vo_pipeline_destroy": [vp],
        "vo_pipeline_process": [vp, i32, vp, i64, i32, vp, vp, vp, vp, vp, i64],
        "vo_pipeline_stage_seconds": [vp, vp],
        "vo_estimate_motion_stereo": [vp, vp, i32, C.POINTER(VoStereoParams), C.POINTER(C.c_double), C.POINTER(i32)],
        "vo_batch_match_quad": [vp, i32, vp, vp, vp, vp, i32, i32, i32, f32, f32],
        "vo_batch_refine": [vp, i32],
        "vo_refine_matches": [vp],
        "vo_profile_read_ext": [vp, C.POINTER(C.c_double), C.POINTER(i64), i32],
    }
    for name, args in sigs.items():
        fn = getattr(lib, name)
        fn.argtypes = args
        if name in ("vo_engine_destroy", "vo_egomotion_destroy", "vo_transformation_vector_to_matrix", "vo_pipeline_destroy",
                    "vo_ransac_gpu_destroy"):
            fn.restype = None
        elif name == "vo_egomotion_rand":
            fn.restype = C.c_uint32
        else:
            fn.restype = C.c_int
    _lib = lib
    return lib


def _ptr(a):
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        return a.ctypes.data_as(C.c_void_p)
    return C.c_void_p(int(a))


class Engine:
    """One vo_engine: frame slots + pair buffers on one GPU, one CUDA stream."""

    def __init__(self, width: int, height: int, max_frames: int = 2, max_pairs: int = 1, device: int = 0,
                 stream: int | None = None, max_matches: int = 0, match_rows_per_cta: int = 0, nms_tau: int = 0,
                 bin_w: int = 0, bin_h: int = 0, match_exhaustive: bool = False, match_threads: int = 0,
                 stereo: bool = False, match_disp_tolerance: int = 0, match_seed_good: int = 0,
                 match_radius_u: int = 0, match_radius_v: int = 0):
        self.lib = load_library()
        if stereo:
            max_frames = max(max_frames, 4)  # left and right prev/curr ring
        cfg = VoConfig(width=width, height=height, bin_w=bin_w, bin_h=bin_h, nms_tau=nms_tau,
                       max_frames=max_frames, max_pairs=max_pairs, max_matches=max_matches,
                       match_rows_per_cta=match_rows_per_cta, match_exhaustive=int(match_exhaustive),
                       match_threads=match_threads, stereo=int(stereo), match_disp_tolerance=match_disp_tolerance,
                       match_seed_good=match_seed_good, match_radius_u=match_radius_u, match_radius_v=match_radius_v)
        h = C.c_void_p()
        self.h = None
        self._check(self.lib.vo_engine_create(C.byref(cfg), device, C.c_void_p(stream or 0), C.byref(h)))
        self.h = h
        g = VoGeometry()
        self._check(self.lib.vo_engine_geometry(self.h, C.byref(g)))
        self.geom = g
        self.W, self.H, self.U, self.V = g.width, g.height, g.u_bins, g.v_bins
        self.CAP, self.BIN_NUM, self.max_matches = g.bin_cap, g.bin_num, g.max_matches

    def _check(self, st: int):
        if st != VO_OK:
            raise VoError(st, self.lib.vo_last_error().decode())

    def close(self):
        if self.h:
            self.lib.vo_engine_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def synchronize(self):
        self._check(self.lib.vo_engine_synchronize(self.h))

    def fence_record(self, idx: int):
        """Mark 'everything enqueued so far, downloads included' (pipeline fence idx in [0,8))."""
        self._check(self.lib.vo_fence_record(self.h, idx))

    def fence_wait(self, idx: int):
        self._check(self.lib.vo_fence_wait(self.h, idx))

    # ---- operator level ------------------------------------------------------------------
    def new_feature_arrays(self, cap_stride: int | None = None):
        cap = cap_stride or self.CAP
        return (np.zeros((self.V, self.U, 4, cap, 12), np.int32), np.zeros((self.V, self.U, 4), np.int32))

    def compute_features(self, img: np.ndarray, cap_stride: int | None = None):
        img = np.ascontiguousarray(img, dtype=np.uint8)
        max2, num = self.new_feature_arrays(cap_stride)
        self._check(self.lib.vo_compute_features(self.h, _ptr(img), img.strides[0], _ptr(max2), _ptr(num),
                                                 max2.shape[3]))
        return max2, num

    def matching(self, m1p, n1p, m1c, n1c, cap: int | None = None):
        cap = cap or self.max_matches
        out = np.zeros(cap, P_MATCH_DTYPE)
        n = C.c_int32(0)
        assert m1p.shape == m1c.shape
        self._check(self.lib.vo_matching(self.h, _ptr(m1p), _ptr(n1p), _ptr(m1c), _ptr(n1c), m1p.shape[3],
                                         _ptr(out), cap, C.byref(n)))
        return out[: n.value].copy()

    def matching_iter(self, m1p, n1p, m1c, n1c, iter_: int, cap: int | None = None):
        """myMatching(..., iter) of the FPGA build (hls_opencl/kernel/myMatch.hpp:51-57): iter == 0 -> zero matches"""
        cap = cap or self.max_matches
        out = np.zeros(cap, P_MATCH_DTYPE)
        n = C.c_int32(-1)
        self._check(self.lib.vo_matching_iter(self.h, _ptr(m1p), _ptr(n1p), _ptr(m1c), _ptr(n1c), m1p.shape[3], _ptr(out), cap,
                                              C.byref(n), iter_))
        return out[: n.value].copy()

    def remove_outliers_bucket(self, matches: np.ndarray):
        """removeOutliers(in, in_cnt, out, out_cnt) of the FPGA build (hls_opencl/kernel/remove_outliers.hpp:9-11): Delaunay
        filter then bucketing 2 per 50x50.  Returns None when the operator left its outputs untouched (<= 3 matches)."""
        buf = np.array(matches, dtype=P_MATCH_DTYPE, copy=True)
        out = np.zeros(max(len(buf), 1), P_MATCH_DTYPE)
        n_in, n_out = C.c_int32(len(buf)), C.c_int32(-1)
        self._check(self.lib.vo_remove_outliers_bucket(self.h, _ptr(buf), C.byref(n_in), _ptr(out), len(out), C.byref(n_out)))
        return None if n_out.value < 0 else out[: n_out.value].copy()

    def bucket_matches(self, matches: np.ndarray, max_features=2, bw=50.0, bh=50.0):
        matches = np.ascontiguousarray(matches, dtype=P_MATCH_DTYPE)
        out = np.zeros(max(len(matches), 1), P_MATCH_DTYPE)
        n = C.c_int32(0)
        self._check(self.lib.vo_bucket_matches(self.h, _ptr(matches) if len(matches) else None, len(matches),
                                               max_features, bw, bh, _ptr(out), len(out), C.byref(n)))
        return out[: n.value].copy()

    def bucket_matches_noalloc(self, d_in: int, n_in: int, max_features, bw, bh, d_out: int, out_cap: int, d_result: int):
        """vo_bucket_matches_noalloc: DEVICE addresses in, nothing allocated, asynchronous on the engine's stream."""
        self._check(self.lib.vo_bucket_matches_noalloc(self.h, C.c_void_p(d_in), n_in, max_features, bw, bh,
                                                       C.c_void_p(d_out), out_cap, C.c_void_p(d_result)))

    # ---- Matcher level -------------------------------------------------------------------
    def push_back(self, img: np.ndarray, replace: bool = False, img2=None):
        img = np.ascontiguousarray(img, dtype=np.uint8)
        if img2 is not None:
            img2 = np.ascontiguousarray(img2, dtype=np.uint8)
            assert img2.shape == img.shape and img2.strides == img.strides
        dims = (C.c_int32 * 3)(img.shape[1], img.shape[0], img.strides[0])
        self._check(self.lib.vo_push_back(self.h, _ptr(img), _ptr(img2), dims, int(replace)))

    def match_features(self, method: int = 0):
        self._check(self.lib.vo_match_features(self.h, method))

    def remove_outliers(self):
        """The removeOutliers call of Matcher::matchFeatures (src/matcher.cpp:108) on the current raw list."""
        self._check(self.lib.vo_remove_outliers(self.h))

    def refine_matches(self):
        """EXTENSION (row F3): sub-pixel refinement of the current list (parameters::refinement == 2)."""
        self._check(self.lib.vo_refine_matches(self.h))

    def bucket_features(self, max_features=2, bw=50.0, bh=50.0):
        self._check(self.lib.vo_bucket_features(self.h, max_features, bw, bh))

    def num_matches(self) -> int:
        n = C.c_int32(0)
        self._check(self.lib.vo_num_matches(self.h, C.byref(n)))
        return n.value

    def get_matches(self) -> np.ndarray:
        out = np.zeros(self.max_matches, P_MATCH_DTYPE)
        n = C.c_int32(0)
        self._check(self.lib.vo_get_matches(self.h, _ptr(out), len(out), C.byref(n)))
        return out[: n.value].copy()

    def set_matches(self, matches: np.ndarray):
        matches = np.ascontiguousarray(matches, dtype=P_MATCH_DTYPE)
        self._check(self.lib.vo_set_matches(self.h, _ptr(matches) if len(matches) else None, len(matches)))

    def dump_features(self, which: int, cap_stride: int | None = None):
        max2, num = self.new_feature_arrays(cap_stride)
        self._check(self.lib.vo_dump_features(self.h, which, _ptr(max2), _ptr(num), max2.shape[3]))
        return max2, num

    # ---- batch level ---------------------------------------------------------------------
    def batch_extract_host(self, first_slot: int, frames, n: int | None = None, frame_stride: int | None = None,
                           pitch: int | None = None):
        """frames: (n,H,W) uint8 ndarray, or a raw host address with n/frame_stride/pitch given."""
        if isinstance(frames, np.ndarray):
            assert frames.dtype == np.uint8 and frames.ndim == 3 and frames.strides[2] == 1
            n, frame_stride, pitch = frames.shape[0], frames.strides[0], frames.strides[1]
        self._check(self.lib.vo_batch_extract_host(self.h, first_slot, n, _ptr(frames), frame_stride, pitch))

    def batch_extract_device(self, first_slot: int, d_ptr: int, n: int, frame_stride: int, pitch: int):
        self._check(self.lib.vo_batch_extract_device(self.h, first_slot, n, C.c_void_p(d_ptr), frame_stride, pitch))

    def batch_match(self, prev_slots, curr_slots, max_features=0, bw=50.0, bh=50.0):
        pv = np.ascontiguousarray(prev_slots, dtype=np.int32)
        cu = np.ascontiguousarray(curr_slots, dtype=np.int32)
        self._check(self.lib.vo_batch_match(self.h, len(pv), _ptr(pv), _ptr(cu), max_features, bw, bh))

    def batch_match_quad(self, s1p, s2p, s1c, s2c, disp_tolerance=2, refinement=0, max_features=0, bw=50.0, bh=50.0):
        """EXTENSION (row F3): quad matching of stereo frame pairs; slots of left/right x prev/curr per item."""
        a = [np.ascontiguousarray(x, dtype=np.int32) for x in (s1p, s2p, s1c, s2c)]
        self._check(self.lib.vo_batch_match_quad(self.h, len(a[0]), _ptr(a[0]), _ptr(a[1]), _ptr(a[2]), _ptr(a[3]),
                                                 disp_tolerance, refinement, max_features, bw, bh))

    def batch_refine(self, n_items: int):
        self._check(self.lib.vo_batch_refine(self.h, n_items))

    def batch_counts(self, n_pairs: int):
        raw = np.zeros(n_pairs, np.int32)
        bk = np.zeros(n_pairs, np.int32)
        self._check(self.lib.vo_batch_counts(self.h, n_pairs, _ptr(raw), _ptr(bk)))
        return raw, bk

    def batch_get_matches(self, pair: int, which: int = 0):
        out = np.zeros(self.max_matches, P_MATCH_DTYPE)
        n = C.c_int32(0)
        self._check(self.lib.vo_batch_get_matches(self.h, pair, which, _ptr(out), len(out), C.byref(n)))
        return out[: n.value].copy()

    def batch_download_async(self, n_pairs: int, which: int, counts, out, stride_matches: int, max_per_pair: int):
        self._check(self.lib.vo_batch_download_async(self.h, n_pairs, which, _ptr(counts), _ptr(out), stride_matches,
                                                     max_per_pair))

    def slot_dump_features(self, slot: int, cap_stride: int | None = None):
        max2, num = self.new_feature_arrays(cap_stride)
        self._check(self.lib.vo_slot_dump_features(self.h, slot, _ptr(max2), _ptr(num), max2.shape[3]))
        return max2, num

    def slot_feature_count(self, slot: int) -> int:
        n = C.c_int32(0)
        self._check(self.lib.vo_slot_feature_count(self.h, slot, C.byref(n)))
        return n.value

    # ---- instrumentation -----------------------------------------------------------------
    def launch_count(self) -> int:
        n = C.c_int64(0)
        self._check(self.lib.vo_launch_count(self.h, C.byref(n)))
        return n.value

    def profile_enable(self, on: bool = True):
        self._check(self.lib.vo_profile_enable(self.h, int(on)))

    def profile_read(self, reset: bool = True):
        ms = (C.c_double * 4)()
        calls = (C.c_int64 * 4)()
        self._check(self.lib.vo_profile_read(self.h, ms, calls, int(reset)))
        names = ("extract", "match", "emit", "bucket")
        return {k: (ms[i], calls[i]) for i, k in enumerate(names)}

    def profile_read_ext(self, reset: bool = True):
        ms = (C.c_double * 8)()
        calls = (C.c_int64 * 8)()
        self._check(self.lib.vo_profile_read_ext(self.h, ms, calls, int(reset)))
        names = ("extract", "match", "emit", "bucket", "refine")
        return {k: (ms[i], calls[i]) for i, k in enumerate(names)}


def remove_outliers_host(matches: np.ndarray) -> np.ndarray:
    """removeOutliers (src/remove_outliers.cpp:4) on a host list; needs no GPU (host stage of the path)."""
    lib = load_library()
    buf = np.array(matches, dtype=P_MATCH_DTYPE, copy=True)
    n = C.c_int32(len(buf))
    st = lib.vo_remove_outliers_host(_ptr(buf) if len(buf) else None, C.byref(n))
    if st != VO_OK:
        raise VoError(st, "vo_remove_outliers_host")
    return buf[: n.value].copy()


def pinned_empty(shape, dtype) -> np.ndarray:
    """numpy array backed by page-locked host memory from vo_host_alloc (never freed: bench lifetime)."""
    lib = load_library()
    dtype = np.dtype(dtype)
    n = int(np.prod(shape)) * dtype.itemsize
    p = C.c_void_p()
    st = lib.vo_host_alloc(C.byref(p), max(n, 1))
    if st != VO_OK:
        raise VoError(st, lib.vo_last_error().decode())
    buf = (C.c_uint8 * max(n, 1)).from_address(p.value)
    return np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)


class Matcher:
    """Python mirror of the reference's Matcher facade (/root/reference/src/matcher.h:40-143): same
    method names, argument meaning and error behaviour; every call is a C-ABI call."""

    class parameters:  # /root/reference/src/matcher.h:45-72 (dead on the reference's executed path)
        def __init__(self):
            self.nms_n = 2
            self.nms_tau = 50
            self.match_binsize = 50
            self.match_radius = 200
            self.match_disp_tolerance = 2
            self.outlier_disp_tolerance = 5
            self.outlier_flow_tolerance = 5
            self.multi_stage = 0
            self.half_resolution = 0
            self.refinement = 0
            self.f = self.cu = self.cv = self.base = 0.0

    def __init__(self, width: int, height: int, param: "Matcher.parameters | None" = None, device: int = 0,
                 stereo: bool = False):
        """stereo=True (EXTENSION, absent from the reference): pushBack keeps I2, matchFeatures(2) is the quad matcher;
        param.refinement == 2 adds the sub-pixel stage.  Defaults = the reference's behaviour (I2/method ignored)."""
        self.param = param or Matcher.parameters()
        self.engine = Engine(width, height, device=device, stereo=stereo,
                             match_disp_tolerance=self.param.match_disp_tolerance if stereo else 0)

    def setIntrinsics(self, f, cu, cv, base):
        self.param.f, self.param.cu, self.param.cv, self.param.base = f, cu, cv, base

    def pushBack(self, I1, dims=None, replace=False, I2=None):
        I1 = np.ascontiguousarray(I1, dtype=np.uint8)
        if dims is not None and (dims[0] <= 0 or dims[1] <= 0 or dims[2] < dims[0]):
            import sys
            print("ERROR: Image dimension mismatch!", file=sys.stderr)  # src/matcher.cpp:59-62: message, return
            return
        self.engine.push_back(I1, replace, img2=I2)

    def matchFeatures(self, method=0, Tr_delta=None):
        # src/matcher.cpp:93-111: myMatching, then the host-side removeOutliers
        self.engine.match_features(method)
        self.engine.remove_outliers()
        if self.param.refinement == 2:
            self.engine.refine_matches()

    def bucketFeatures(self, max_features, bucket_width, bucket_height):
        self.engine.bucket_features(max_features, bucket_width, bucket_height)

    def getMatches(self):
        return self.engine.get_matches()


class Egomotion:
    """Host stage F2 (vo_estimate_motion_mono): VisualOdometryMono::estimateMotion,
    /root/reference/src/viso_mono.cpp:41-160, with the reference's rand() sample stream (seed 0)."""

    def __init__(self, f=1.0, cu=0.0, cv=0.0, height=1.0, pitch=0.0, ransac_iters=2000, inlier_threshold=1e-5,
                 motion_threshold=100.0, seed: int = 0, threads: int = 0):
        self.lib = load_library()
        self.prm = VoMonoParams(f=f, cu=cu, cv=cv, height=height, pitch=pitch, ransac_iters=ransac_iters,
                                inlier_threshold=inlier_threshold, motion_threshold=motion_threshold)
        h = C.c_void_p()
        st = self.lib.vo_egomotion_create(seed, threads, C.byref(h))
        if st != VO_OK:
            raise VoError(st, "vo_egomotion_create")
        self.h = h

    def close(self):
        if getattr(self, "h", None):
            self.lib.vo_egomotion_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def estimate(self, matches: np.ndarray):
        """-> 6-vector (rx,ry,rz,tx,ty,tz) or None where the reference returns an empty vector."""
        m = np.ascontiguousarray(matches, dtype=P_MATCH_DTYPE)
        tr = (C.c_double * 6)()
        ok = C.c_int32(0)
        st = self.lib.vo_estimate_motion_mono(self.h, _ptr(m) if len(m) else None, len(m), C.byref(self.prm), tr, C.byref(ok))
        if st != VO_OK:
            raise VoError(st, "vo_estimate_motion_mono")
        return np.array(tr[:], dtype=np.float64) if ok.value else None

    # ---- the two-step form: sample sets drawn in frame order, estimation stateless (and scorable on the GPU) ----
    def draw_samples(self, matches: np.ndarray):
        """-> (samples int32[iters, 8], iters drawn): the sets the reference would draw for this list (advances the stream)."""
        m = np.ascontiguousarray(matches, dtype=P_MATCH_DTYPE)
        smp = np.zeros((max(self.prm.ransac_iters, 1), 8), np.int32)
        n = C.c_int32(0)
        st = self.lib.vo_egomotion_draw_samples(self.h, _ptr(m) if len(m) else None, len(m), C.byref(self.prm), _ptr(smp), C.byref(n))
        if st != VO_OK:
            raise VoError(st, "vo_egomotion_draw_samples")
        return smp, n.value

    def draw_samples_batch(self, lists, threads: int = 4):
        """the same for consecutive lists in one call (vo_egomotion_draw_samples_batch) -> [(samples, iters), ...]"""
        n = len(lists)
        cap = max([len(l) for l in lists] + [1])
        buf = np.zeros((n, cap), dtype=P_MATCH_DTYPE)
        cnt = np.zeros(n, np.int32)
        for i, l in enumerate(lists):
            buf[i, : len(l)] = l
            cnt[i] = len(l)
        iters = max(self.prm.ransac_iters, 1)
        smp = np.zeros((n, iters, 8), np.int32)
        drawn = np.zeros(n, np.int32)
        st = self.lib.vo_egomotion_draw_samples_batch(self.h, n, _ptr(buf), cap, _ptr(cnt), C.byref(self.prm), _ptr(smp), iters,
                                                      _ptr(drawn), threads)
        if st != VO_OK:
            raise VoError(st, "vo_egomotion_draw_samples_batch")
        return [(smp[i], int(drawn[i])) for i in range(n)]

    def estimate_with_samples(self, matches: np.ndarray, samples: np.ndarray, iters: int, best_k: int | None = None):
        """-> (tr or None, inlier indices).  best_k: the RANSAC winner if already known (RansacGpu.score), else scored here."""
        m = np.ascontiguousarray(matches, dtype=P_MATCH_DTYPE)
        tr = (C.c_double * 6)()
        ok, ni = C.c_int32(0), C.c_int32(0)
        inl = np.zeros(max(len(m), 1), np.int32)
        if best_k is None:
            st = self.lib.vo_estimate_motion_mono_samples(_ptr(m) if len(m) else None, len(m), C.byref(self.prm), _ptr(samples), iters, 1,
                                                          tr, C.byref(ok), _ptr(inl), len(inl), C.byref(ni))
        else:
            st = self.lib.vo_estimate_motion_mono_best(_ptr(m) if len(m) else None, len(m), C.byref(self.prm), _ptr(samples), iters,
                                                       int(best_k), tr, C.byref(ok), _ptr(inl), len(inl), C.byref(ni))
        if st != VO_OK:
            raise VoError(st, "vo_estimate_motion_mono_samples/_best")
        return (np.array(tr[:], dtype=np.float64) if ok.value else None), inl[: ni.value].copy()

    def inliers(self) -> np.ndarray:
        n = C.c_int32(0)
        self.lib.vo_egomotion_inliers(self.h, None, 0, C.byref(n))
        out = np.zeros(max(n.value, 1), dtype=np.int32)
        self.lib.vo_egomotion_inliers(self.h, _ptr(out), len(out), C.byref(n))
        return out[: n.value]

    def rand(self) -> int:
        return int(self.lib.vo_egomotion_rand(self.h))


class RansacGpu:
    """vo_ransac_gpu_*: the RANSAC hypotheses of VisualOdometryMono::estimateMotion (/root/reference/src/viso_mono.cpp:63-77)
    scored on the GPU, one thread per hypothesis, with the host stage's own arithmetic."""

    def __init__(self, max_lists: int, max_points: int, max_iters: int = 2000, device: int = 0):
        self.lib = load_library()
        h = C.c_void_p()
        st = self.lib.vo_ransac_gpu_create(device, max_lists, max_points, max_iters, C.byref(h))
        if st != VO_OK:
            raise VoError(st, "vo_ransac_gpu_create")
        self.h, self.max_points, self.max_iters = h, max_points, max_iters

    def close(self):
        if getattr(self, "h", None):
            self.lib.vo_ransac_gpu_destroy(self.h)
            self.h = None

    def score(self, lists, prm: "VoMonoParams", samples, drawn):
        """lists: sequence of match arrays; samples: sequence of int32[iters, 8]; drawn: iterations per list.
        -> (best_k[n], best_count[n])"""
        n = len(lists)
        buf = np.zeros((n, self.max_points), dtype=P_MATCH_DTYPE)
        cnt = np.zeros(n, np.int32)
        smp = np.zeros((n, self.max_iters, 8), np.int32)
        for i, (l, s) in enumerate(zip(lists, samples)):
            buf[i, : len(l)] = l
            cnt[i] = len(l)
            smp[i, : len(s)] = s
        dr = np.ascontiguousarray(drawn, dtype=np.int32)
        bk, bc = np.zeros(n, np.int32), np.zeros(n, np.int32)
        st = self.lib.vo_ransac_gpu_score(self.h, n, _ptr(buf), self.max_points, _ptr(cnt), C.byref(prm), _ptr(smp), self.max_iters,
                                          _ptr(dr), _ptr(bk), _ptr(bc))
        if st != VO_OK:
            raise VoError(st, "vo_ransac_gpu_score")
        return bk, bc


class EgomotionStereo(Egomotion):
    """vo_estimate_motion_stereo: VisualOdometryStereo::estimateMotion, /root/reference/src/viso_stereo.cpp:54-157,
    on quad match lists; same sample stream handle as the monocular stage."""

    def __init__(self, f=1.0, cu=0.0, cv=0.0, base=1.0, ransac_iters=200, inlier_threshold=2.0, reweighting=True,
                 seed: int = 0):
        super().__init__(seed=seed)
        self.sprm = VoStereoParams(f=f, cu=cu, cv=cv, base=base, ransac_iters=ransac_iters, reweighting=int(reweighting),
                                   inlier_threshold=inlier_threshold)

    def estimate(self, matches: np.ndarray):
        m = np.ascontiguousarray(matches, dtype=P_MATCH_DTYPE)
        tr = (C.c_double * 6)()
        ok = C.c_int32(0)
        st = self.lib.vo_estimate_motion_stereo(self.h, _ptr(m) if len(m) else None, len(m), C.byref(self.sprm), tr, C.byref(ok))
        if st != VO_OK:
            raise VoError(st, "vo_estimate_motion_stereo")
        return np.array(tr[:], dtype=np.float64) if ok.value else None


def transformation_vector_to_matrix(tr) -> np.ndarray:
    """VisualOdometry::transformationVectorToMatrix, /root/reference/src/viso.cpp:59-84."""
    lib = load_library()
    a = (C.c_double * 6)(*[float(x) for x in tr])
    out = (C.c_double * 16)()
    lib.vo_transformation_vector_to_matrix(a, out)
    return np.array(out[:], dtype=np.float64).reshape(4, 4)


class VisualOdometryMono:
    """Python mirror of /root/reference/src/viso_mono.h:27-84 + src/viso.h:31-113: process() = pushBack ->
    matchFeatures (GPU circular matching + host removeOutliers) -> bucketFeatures -> getMatches ->
    updateMotion (host egomotion)."""

    class parameters:
        def __init__(self):
            self.match = Matcher.parameters()
            self.max_features, self.bucket_width, self.bucket_height = 2, 50.0, 50.0   # src/viso.h:43-53
            self.f, self.cu, self.cv = 1.0, 0.0, 0.0                                   # src/viso.h:56-64
            self.height, self.pitch = 1.0, 0.0                                         # src/viso_mono.h:32-46
            self.ransac_iters, self.inlier_threshold, self.motion_threshold = 2000, 1e-5, 100.0

    def __init__(self, width: int, height: int, param: "VisualOdometryMono.parameters | None" = None, device: int = 0):
        self.param = param or VisualOdometryMono.parameters()
        p = self.param
        self.matcher = Matcher(width, height, p.match, device=device)
        self.ego = Egomotion(p.f, p.cu, p.cv, p.height, p.pitch, p.ransac_iters, p.inlier_threshold, p.motion_threshold)
        self.Tr_delta = np.eye(4)
        self.Tr_valid = False
        self.p_matched = np.zeros(0, dtype=P_MATCH_DTYPE)

    def process(self, I, dims=None, replace=False) -> bool:
        if isinstance(I, np.ndarray) and I.dtype == P_MATCH_DTYPE:      # process(p_matched): src/viso.h:75-78
            self.p_matched = I
            return self.updateMotion()
        self.matcher.pushBack(I, dims, replace)
        self.matcher.matchFeatures(0)
        self.matcher.bucketFeatures(self.param.max_features, self.param.bucket_width, self.param.bucket_height)
        self.p_matched = self.matcher.getMatches()
        return self.updateMotion()

    def updateMotion(self) -> bool:  # src/viso.cpp:42-57
        tr = self.ego.estimate(self.p_matched)
        if tr is None:
            return False
        self.Tr_delta = transformation_vector_to_matrix(tr)
        self.Tr_valid = True
        return True

    def getMotion(self):
        return self.Tr_delta

    def getMatches(self):
        return self.matcher.getMatches()

    def getNumberOfMatches(self):
        return len(self.p_matched)

    def getNumberOfInliers(self):
        return len(self.ego.inliers())

    def getInlierIndices(self):
        return self.ego.inliers()


class VisualOdometryStereo:
    """Python mirror of /root/reference/src/viso_stereo.h:27-96: process(I1, I2) = pushBack -> [bootstrap while no valid
    motion] -> matchFeatures(2) -> bucketFeatures -> getMatches -> updateMotion (src/viso_stereo.cpp:33-52).  The
    matching half is the quad-matching EXTENSION (the reference's matcher is flow-only, so its own class cannot work);
    the egomotion half is the reference's, bit for bit (vo_estimate_motion_stereo)."""

    class parameters:
        def __init__(self):
            self.match = Matcher.parameters()
            self.max_features, self.bucket_width, self.bucket_height = 2, 50.0, 50.0   # src/viso.h:43-53
            self.f, self.cu, self.cv = 1.0, 0.0, 0.0                                   # src/viso.h:56-64
            self.base, self.ransac_iters, self.inlier_threshold, self.reweighting = 1.0, 200, 2.0, True

    def __init__(self, width: int, height: int, param: "VisualOdometryStereo.parameters | None" = None, device: int = 0):
        self.param = param or VisualOdometryStereo.parameters()
        p = self.param
        self.matcher = Matcher(width, height, p.match, device=device, stereo=True)
        self.matcher.setIntrinsics(p.f, p.cu, p.cv, p.base)
        self.ego = EgomotionStereo(p.f, p.cu, p.cv, p.base, p.ransac_iters, p.inlier_threshold, p.reweighting)
        self.Tr_delta = np.eye(4)
        self.Tr_valid = False
        self.p_matched = np.zeros(0, dtype=P_MATCH_DTYPE)

    def _match(self):
        self.matcher.matchFeatures(2, self.Tr_delta if self.Tr_valid else None)
        self.matcher.bucketFeatures(self.param.max_features, self.param.bucket_width, self.param.bucket_height)
        self.p_matched = self.matcher.getMatches()

    def process(self, I1, I2=None, dims=None, replace=False) -> bool:
        if isinstance(I1, np.ndarray) and I1.dtype == P_MATCH_DTYPE:    # process(p_matched): src/viso.h:75-78
            self.p_matched = I1
            return self.updateMotion()
        self.matcher.pushBack(I1, dims, replace, I2=I2)
        if not self.Tr_valid:                                           # bootstrap: src/viso_stereo.cpp:39-44
            self._match()
            self.updateMotion()
        self._match()
        return self.updateMotion()

    def updateMotion(self) -> bool:
        tr = self.ego.estimate(self.p_matched)
        if tr is None:
            return False
        self.Tr_delta = transformation_vector_to_matrix(tr)
        self.Tr_valid = True
        return True

    def getMotion(self):
        return self.Tr_delta

    def getMatches(self):
        return self.matcher.getMatches()

    def getNumberOfMatches(self):
        return len(self.p_matched)

    def getNumberOfInliers(self):
        return len(self.ego.inliers())

    def getInlierIndices(self):
        return self.ego.inliers()


class Pipeline:
    """Sequence level (vo_pipeline_*): VisualOdometryMono::process for batches of consecutive frames of one
    sequence -- GPU extraction/matching/bucketing of the whole batch, host Delaunay filter and egomotion of all
    pairs on a thread pool; per-frame results equal frame-by-frame process() (/root/reference/src/viso_mono.cpp:33-39)."""

    def __init__(self, width: int, height: int, batch: int, f=1.0, cu=0.0, cv=0.0, cam_height=1.0, pitch=0.0, ransac_iters=2000,
                 inlier_threshold=1e-5, motion_threshold=100.0, max_features=2, bucket_w=50.0, bucket_h=50.0,
                 host_threads: int = 0, device: int = 0):
        self.engine = Engine(width, height, max_frames=batch + 1, max_pairs=batch, device=device)
        self.lib = self.engine.lib
        self.batch = batch
        self.prm = VoMonoParams(f=f, cu=cu, cv=cv, height=cam_height, pitch=pitch, ransac_iters=ransac_iters,
                                inlier_threshold=inlier_threshold, motion_threshold=motion_threshold)
        h = C.c_void_p()
        st = self.lib.vo_pipeline_create(self.engine.h, C.byref(self.prm), max_features, bucket_w, bucket_h, host_threads, C.byref(h))
        if st != VO_OK:
            raise VoError(st, "vo_pipeline_create")
        self.h = h
        self.cap = (int(width / bucket_w) + 2) * (int(height / bucket_h) + 2) * max_features

    def close(self):
        if getattr(self, "h", None):
            self.lib.vo_pipeline_destroy(self.h)
            self.h = None
            self.engine.close()

    STAGES = ("extract_match_gpu", "raw_lists_d2h", "remove_outliers_host", "bucket_h2d_gpu_d2h", "ransac_samples_host",
              "ransac_score_gpu", "egomotion_host")

    def stage_seconds(self) -> dict:
        """Wall seconds process() has spent per stage since creation (vo_pipeline_stage_seconds)."""
        out = np.zeros(7, np.float64)
        st = self.lib.vo_pipeline_stage_seconds(self.h, _ptr(out))
        if st != VO_OK:
            raise VoError(st, "vo_pipeline_stage_seconds")
        return dict(zip(self.STAGES, out.tolist()))

    def process(self, frames: np.ndarray, want_matches: bool = False):
        """frames: (n, H, pitch>=W) uint8, n <= batch.  -> ok[n], tr[n,6], n_matches[n], n_inliers[n] (, lists)."""
        frames = np.ascontiguousarray(frames, dtype=np.uint8)
        n = frames.shape[0]
        ok = np.zeros(n, np.int32); tr = np.zeros((n, 6), np.float64)
        nm = np.zeros(n, np.int32); ni = np.zeros(n, np.int32)
        lists = np.zeros((n, self.cap), dtype=P_MATCH_DTYPE) if want_matches else None
        st = self.lib.vo_pipeline_process(self.h, n, _ptr(frames), frames.strides[0], frames.strides[1], _ptr(ok), _ptr(tr),
                                          _ptr(nm), _ptr(ni), _ptr(lists), self.cap)
        if st != VO_OK:
            raise VoError(st, self.lib.vo_last_error().decode() or "vo_pipeline_process")
        if want_matches:
            return ok, tr, nm, ni, [lists[i, : nm[i]].copy() for i in range(n)]
        return ok, tr, nm, ni
