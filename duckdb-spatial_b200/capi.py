"""ctypes binding of libspatial_b200.so (include/spatial_b200.h). Every signature is declared here; loading fails
loudly when the library is missing -- there is no CPU fallback anywhere in this package."""
import ctypes as C
import os

from . import build as _build

HOST, DEVICE = 0, 1
PRED_INTERSECTS, PRED_WITHIN = 0, 1
OK, ERR_CUDA, ERR_ARG, ERR_CAPACITY, ERR_GEOMETRY, ERR_EMPTY = 0, -1, -2, -3, -4, -5

vp = C.c_void_p
u32, u64, i32 = C.c_uint32, C.c_uint64, C.c_int
f64 = C.c_double
i64 = C.c_int64

# name -> (restype, argtypes). Data pointers are void* so numpy / torch / raw addresses all pass as ints.
SIGNATURES = {
    "sb_last_error": (C.c_char_p, []),
    "sb_version": (C.c_char_p, []),
    "sb_context_create": (i32, [i32, C.POINTER(vp)]),
    "sb_context_destroy": (None, [vp]),
    "sb_context_set_stream": (i32, [vp, vp]),
    "sb_context_sync": (i32, [vp]),
    "sb_context_set_option": (i32, [vp, C.c_char_p, C.c_int64]),
    "sb_context_launches": (u64, [vp]),
    "sb_context_profile": (i32, [vp, i32]),
    "sb_context_profile_read": (i32, [vp, C.c_char_p, C.POINTER(f64), C.POINTER(u64)]),
    "sb_bbox_from_points": (i32, [vp, i32, vp, vp, u64, vp]),
    "sb_bbox_from_box2d": (i32, [vp, i32, vp, vp, vp, vp, u64, vp]),
    "sb_bbox_from_blobs": (i32, [vp, i32, vp, vp, u64, vp, vp, vp]),
    "sb_extent_from_blobs": (i32, [vp, i32, vp, vp, u64, vp, vp, vp]),
    "sb_join_op_create": (i32, [vp, i32, i32, vp, vp, vp, vp, u32, u64, vp]),
    "sb_join_op_push": (i32, [vp, vp, vp, u64, vp]),
    "sb_join_op_flush": (i32, [vp, vp, vp, u64, vp, vp, u64, vp]),
    "sb_join_op_finish": (i32, [vp, vp, u64, vp]),
    "sb_join_op_rows_pushed": (u64, [vp]),
    "sb_join_op_destroy": (None, [vp]),
    "sb_extent_from_wkb": (i32, [vp, i32, vp, vp, u64, vp, vp, vp]),
    "sb_points_from_blobs": (i32, [vp, i32, vp, vp, u64, vp, vp, vp]),
    "sb_polygons_from_blobs": (i32, [vp, i32, vp, vp, u64, vp, vp, vp, vp, u64, u64, vp, vp, vp]),
    "sb_hilbert_xy": (i32, [vp, i32, vp, vp, u64, vp, vp]),
    "sb_hilbert_box2d": (i32, [vp, i32, vp, u64, vp, vp]),
    "sb_hilbert_box2df": (i32, [vp, i32, vp, u64, vp, vp]),
    "sb_hilbert_geom": (i32, [vp, i32, vp, vp, u64, vp, vp]),
    "sb_hilbert_geom_bounds": (i32, [vp, i32, vp, vp, u64, vp, vp, vp]),
    "sb_sort_pairs_u32": (i32, [vp, i32, vp, vp, u64]),
    "sb_sort_pairs_u64": (i32, [vp, i32, vp, vp, u64]),
    "sb_rtree_build_str": (i32, [vp, i32, vp, vp, u64, u32, C.POINTER(vp)]),
    "sb_rtree_destroy": (None, [vp]),
    "sb_rtree_num_levels": (i32, [vp]),
    "sb_rtree_level_count": (u64, [vp, i32]),
    "sb_rtree_slice_size": (u64, [vp]),
    "sb_rtree_export_level": (i32, [vp, vp, i32, i32, vp]),
    "sb_rtree_pages_layout": (i32, [vp, u64, vp]),
    "sb_rtree_export_pages": (i32, [vp, vp, u64, i32, vp, vp, vp]),
    "sb_rtree_scan": (i32, [vp, vp, i32, vp, u32, vp, vp, u64, C.POINTER(u64)]),
    "sb_flat_build": (i32, [vp, i32, vp, u32, u32, C.POINTER(vp)]),
    "sb_flat_destroy": (None, [vp]),
    "sb_flat_item_count": (u32, [vp]),
    "sb_flat_node_count": (u32, [vp]),
    "sb_flat_num_layers": (u32, [vp]),
    "sb_flat_layer_bounds": (C.POINTER(u32), [vp]),
    "sb_flat_device_buffer": (i32, [vp, C.POINTER(vp), C.POINTER(u64)]),
    "sb_flat_attach": (i32, [vp, vp, u64, C.POINTER(vp)]),
    "sb_flat_export": (i32, [vp, vp, i32, vp, vp]),
    "sb_flat_probe": (i32, [vp, vp, i32, vp, u64, vp, vp, u64, C.POINTER(u64)]),
    "sb_flat_probe_points": (i32, [vp, vp, i32, i32, vp, vp, u64, u32, vp, vp, vp, u64, C.POINTER(u64)]),
    "sb_refine_point_box": (i32, [vp, i32, i32, vp, vp, vp, vp, vp, vp, vp, vp, C.POINTER(u64)]),
    "sb_refine_point_polygon": (i32, [vp, i32, vp, vp, vp, vp, vp, vp, u64, u64, u64, u64, vp, vp, C.POINTER(u64)]),
    "sb_refine_point_polygon_exact": (i32, [vp, i32, i32, vp, vp, vp, vp, vp, vp, u64, u64, u64, u64, vp, vp, C.POINTER(u64)]),
    "sb_refine_dwithin": (i32, [vp, i32, vp, vp, vp, vp, f64, vp, vp, vp, C.POINTER(u64)]),
    "sb_distance_points": (i32, [vp, i32, vp, vp, vp, vp, u64, vp]),
    "sb_intersects_box2d": (i32, [vp, i32] + [vp] * 8 + [u32, vp, vp, u64, vp]),
    "sb_distance_point_linestring": (i32, [vp, i32, vp, vp, vp, u64, vp, vp, u64, vp]),
    "sb_join_points_boxes": (i32, [vp, i32, i32, vp, vp, u64, vp, vp, vp, vp, u32, vp, vp, u64, C.POINTER(u64)]),
    "sb_join_points_boxes_multi": (i32, [vp, i32, i32, vp, vp, u64, vp, vp, vp, vp, u32, vp, vp, u64, C.POINTER(u64)]),
    "sb_join_mark_matched": (i32, [vp, i32, vp, vp, u64, u32, vp, u64, vp, u64]),
    "sb_join_unmatched_rows": (i32, [vp, i32, vp, u64, u32, vp, u64, C.POINTER(u64)]),
    "sb_rtree_build_from_blobs": (i32, [vp, i32, vp, vp, u64, vp, u32, C.POINTER(vp), C.POINTER(u64), vp]),
    "sb_rtree_slice_for": (u64, [u64, u32]),
    "sb_rtree_build_from_points": (i32, [vp, i32, vp, vp, vp, u64, u32, C.POINTER(vp)]),
    "sb_rtree_build_leaf_shard": (i32, [vp, i32, vp, vp, u64, u32, u64, u64, i64, i64, vp, vp]),
    "sb_rtree_from_leaf_level": (i32, [vp, i32, vp, u64, vp, u32, C.POINTER(vp)]),
    "sb_rtree_dump": (i32, [vp, vp, i32, vp, vp, vp, u64, C.POINTER(u64)]),
    "sb_context_trim": (i32, [vp]),
    "sb_pairs_checksum": (i32, [vp, i32, vp, vp, u64, u32, u32, C.POINTER(u64), C.POINTER(u64)]),
    "sb_pairs_checksum_noalloc": (i32, [i32, vp, vp, vp, u64, u32, u32, vp]),
    "sb_join_points_polygons": (i32, [vp, i32, vp, vp, u64, vp, vp, vp, vp, u64, u64, u64, vp, vp, u64, C.POINTER(u64)]),
    "sb_join_points_polygons_exact": (i32, [vp, i32, i32, vp, vp, u64, vp, vp, vp, vp, u64, u64, u64, vp, vp, u64, C.POINTER(u64)]),
    "sb_generate_points": (i32, [vp, i32, u64, u64, u64, vp, vp, vp]),
    "sb_join_dwithin": (i32, [vp, i32, vp, vp, u64, vp, vp, u32, f64, vp, vp, vp, u64, C.POINTER(u64)]),
}

_lib = None


class SpatialB200Error(RuntimeError):
    def __init__(self, status, message):
        super().__init__("spatial_b200 status %d: %s" % (status, message))
        self.status = status


def library_path():
    # SPATIAL_B200_LIB: another build of the same library (A/B measurements of kernel variants, scripts/build_variant.py)
    return os.environ.get("SPATIAL_B200_LIB") or _build.LIB


def lib(build_if_missing=True):
    """Load libspatial_b200.so; raises (never falls back) when it cannot be loaded."""
    global _lib
    if _lib is None:
        path = library_path()
        if not os.path.exists(path):
            if not build_if_missing:
                raise SpatialB200Error(ERR_CUDA, "libspatial_b200.so is not built (run __graft_entry__.build())")
            _build.build()
        L = C.CDLL(path)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)  # AttributeError if the symbol is missing: the header and the .so must agree
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(status):
    if status != OK:
        raise SpatialB200Error(status, lib().sb_last_error().decode())
    return status
