"""duckdb-spatial_b200: B200-native (sm_100a) implementation of duckdb-spatial's data-parallel hot path
(bbox / Hilbert -> R-tree bulk load -> batched probe -> exact refinement) behind a C-ABI
(include/spatial_b200.h, libspatial_b200.so). This package is the thin host mirror of the reference's
function / index / operator surface for that path. There is no CPU fallback."""
from . import build as _build
from . import capi
from .capi import SpatialB200Error
from .functions import ENTRY_DTYPE, Context, FlatRTree, RTreeIndex, SpatialJoinOp, spatial_join_points_boxes_multi


def build_native(force=False, verbose=False):
    return _build.build(force=force, verbose=verbose)


__all__ = ["capi", "Context", "RTreeIndex", "FlatRTree", "SpatialJoinOp", "spatial_join_points_boxes_multi", "ENTRY_DTYPE", "SpatialB200Error", "build_native"]
