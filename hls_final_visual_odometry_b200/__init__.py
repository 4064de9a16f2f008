"""B200-native visual-odometry front end: feature extraction -> circular matching -> bucketing.

Hand-written CUDA for sm_100a behind a C ABI (include/vo_b200.h); this package is the host-side
mirror of the reference's Matcher interface plus the synthetic-frame generator used by tests and
bench.py.  No CPU implementation of the hot path lives here.
"""
from .api import (Engine, Matcher, VoError, P_MATCH_DTYPE, load_library, pinned_empty, Egomotion, RansacGpu,  # noqa: F401
                  VisualOdometryMono, transformation_vector_to_matrix, Pipeline, EgomotionStereo, VisualOdometryStereo)
from . import synth  # noqa: F401

__all__ = ["Engine", "Matcher", "VoError", "P_MATCH_DTYPE", "load_library", "pinned_empty", "synth", "Egomotion", "RansacGpu",
           "VisualOdometryMono", "transformation_vector_to_matrix", "Pipeline", "EgomotionStereo", "VisualOdometryStereo"]
