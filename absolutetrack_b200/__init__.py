"""absolutetrack_b200: B200-native (sm_100a) implementation of UmeTrack's per-frame multi-view
inference hot path behind the reference's own lib.tracker / lib.models API.

`absolutetrack_b200.lib` mirrors the reference's `lib` package for this path; put this directory
on sys.path and run_eval_*_skeleton.py / run_inference_torch_data.py import
it unchanged.  All arithmetic of the hot path runs in libabsolutetrack_b200.so (CUDA, C-ABI in
include/absolutetrack_b200.h); there is no CPU fallback.
"""
from . import _cabi  # noqa: F401

__all__ = ["_cabi"]
