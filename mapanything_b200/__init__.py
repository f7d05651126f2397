"""mapanything_b200: B200-native (sm_100a) implementation of the MapAnything inference hot path.

`MapAnything` mirrors the reference class's inference API; all arithmetic runs in the hand-written
CUDA kernels of libmab200.so (C ABI in include/mab200.h).
"""
from .model import MapAnything, validate_input_views_for_inference  # noqa: F401

__all__ = ["MapAnything", "validate_input_views_for_inference"]
