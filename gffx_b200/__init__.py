"""gffx_b200 — B200-native interval-overlap engine for GFFx's hot path (intersect / depth / coverage).

The compute path is hand-written sm_100a CUDA behind a C ABI (include/gffx_b200.h,
gffx_b200/lib/libgffx_b200.so).  This package is the thin host-side mirror of the reference's
public interface over that ABI; it has no CPU fallback and never touches oracle/.
"""
from .api import (CountRows, CoverageArgs, DepthArgs, GffxError, IntersectArgs, OverlapMode, Result, TreeIndexData, run_coverage,
                  run_depth, run_intersect, synth_gff, build_index)
from ._lib import F_DEVICE_INPUT, F_NO_HOST_COPY, F_PAIRS_ON_DEVICE, F_WANT_PAIRS, LIB_PATH

__all__ = ["CountRows", "CoverageArgs", "DepthArgs", "GffxError", "IntersectArgs", "OverlapMode", "Result", "TreeIndexData",
           "run_coverage", "run_depth", "run_intersect", "synth_gff", "build_index", "F_DEVICE_INPUT", "F_NO_HOST_COPY", "F_PAIRS_ON_DEVICE",
           "F_WANT_PAIRS", "LIB_PATH"]
