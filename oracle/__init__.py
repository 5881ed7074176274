"""CPU oracle for the duckdb-spatial hot path -- TEST INFRASTRUCTURE ONLY.

ctypes/numpy front-end of ``oracle/oracle.c`` (a plain-C restatement of the reference algorithms, each
function citing the reference file:line) and, when present, of ``oracle/_ref/libref.so`` (the reference's
own headers + FlatRTree compiled from /root/reference by ``oracle/Makefile``).

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py`` (cpu_baseline / ``--impl reference``) may
import this package; the product (``duckdb-spatial_b200/``) never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None
_REF = None

u8p = C.POINTER(C.c_uint8)
u32p = C.POINTER(C.c_uint32)
u64p = C.POINTER(C.c_uint64)
i32p = C.POINTER(C.c_int32)
i64p = C.POINTER(C.c_int64)
f32p = C.POINTER(C.c_float)
f64p = C.POINTER(C.c_double)


def build(force=False):
    """Compile liboracle.so (and _ref/libref.so when /root/reference exists)."""
    so = os.path.join(_HERE, "_build", "liboracle.so")
    src = os.path.join(_HERE, "oracle.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", _HERE, "_build/liboracle.so"])
    ref = os.path.join(_HERE, "_ref", "libref.so")
    rsrc = os.path.join(_HERE, "ref_harness.cpp")
    if os.path.isdir("/root/reference/src/spatial") and (
        force or not os.path.exists(ref) or os.path.getmtime(ref) < os.path.getmtime(rsrc)
    ):
        subprocess.check_call(["make", "-s", "-C", _HERE, "_ref/libref.so"])
    ref2 = os.path.join(_HERE, "_ref", "libref_scalar.so")
    rsrc2 = os.path.join(_HERE, "ref_scalar.cpp")
    if os.path.isdir("/root/reference/src/spatial") and (
        force or not os.path.exists(ref2) or os.path.getmtime(ref2) < os.path.getmtime(rsrc2)
    ):
        subprocess.check_call(["make", "-s", "-C", _HERE, "_ref/libref_scalar.so"])
    # the DuckDB-linked reference (CREATE INDEX / rtree_index_dump / ST_Within / SPATIAL_JOIN through SQL): DuckDB core
    # (~6 min the first time) + the unmodified reference translation units, see the `ref_duckdb` rules of the Makefile
    if os.path.isdir("/root/reference/duckdb/src") and not os.environ.get("ORACLE_SKIP_DUCKDB"):
        subprocess.check_call(["make", "-s", "-j8", "-C", _HERE, "ref_duckdb"])


def _p(a, t):
    return a.ctypes.data_as(t) if a is not None else None


def lib():
    global _LIB
    if _LIB is None:
        build()
        L = C.CDLL(os.path.join(_HERE, "_build", "liboracle.so"))
        L.orc_pcg32_nth.restype = C.c_uint32
        L.orc_pcg32_nth.argtypes = [C.c_uint64, C.c_uint64]
        L.orc_double_to_float_down.restype = C.c_float
        L.orc_double_to_float_down.argtypes = [C.c_double]
        L.orc_double_to_float_up.restype = C.c_float
        L.orc_double_to_float_up.argtypes = [C.c_double]
        L.orc_hilbert_encode.restype = C.c_uint32
        L.orc_hilbert_encode.argtypes = [C.c_uint32] * 3
        L.orc_hilbert_f32_to_u32.restype = C.c_uint32
        L.orc_hilbert_f32_to_u32.argtypes = [C.c_float]
        L.orc_write_polygon_blobs.restype = C.c_uint64
        L.orc_cached_bounds.restype = C.c_int
        L.orc_extent_exact.restype = C.c_int
        L.orc_str_slice_size.restype = C.c_uint64
        L.orc_str_slice_size.argtypes = [C.c_uint64, C.c_uint32]
        L.orc_rtree_build_str.restype = C.c_void_p
        L.orc_rtree_free.argtypes = [C.c_void_p]
        L.orc_rtree_nlevels.argtypes = [C.c_void_p]
        L.orc_rtree_level_count.restype = C.c_uint64
        L.orc_rtree_level_count.argtypes = [C.c_void_p, C.c_int]
        L.orc_rtree_level_entries.restype = C.c_void_p
        L.orc_rtree_level_entries.argtypes = [C.c_void_p, C.c_int]
        L.orc_rtree_slice.restype = C.c_uint64
        L.orc_rtree_slice.argtypes = [C.c_void_p]
        L.orc_rtree_dump.restype = C.c_uint64
        L.orc_rtree_scan.restype = C.c_uint64
        L.orc_flat_build.restype = C.c_void_p
        L.orc_flat_free.argtypes = [C.c_void_p]
        L.orc_flat_nodes.restype = C.c_uint32
        L.orc_flat_nodes.argtypes = [C.c_void_p]
        L.orc_flat_nlayers.restype = C.c_uint32
        L.orc_flat_nlayers.argtypes = [C.c_void_p]
        L.orc_flat_layer_bounds.restype = u32p
        L.orc_flat_layer_bounds.argtypes = [C.c_void_p]
        L.orc_flat_boxes.restype = f32p
        L.orc_flat_boxes.argtypes = [C.c_void_p]
        L.orc_flat_idx.restype = u32p
        L.orc_flat_idx.argtypes = [C.c_void_p]
        L.orc_flat_tree_box.restype = f32p
        L.orc_flat_tree_box.argtypes = [C.c_void_p]
        L.orc_flat_lookup.restype = C.c_uint64
        L.orc_point_intersects_box.argtypes = [C.c_double] * 6
        L.orc_point_within_box.argtypes = [C.c_double] * 6
        L.orc_pip_native.restype = C.c_int
        L.orc_pip_native.argtypes = [C.c_double, C.c_double, f64p, f64p, u64p, C.c_uint64, C.c_uint64]
        L.orc_orient2d_sign.restype = C.c_int
        L.orc_orient2d_sign.argtypes = [C.c_double] * 6
        L.orc_point_polygon_locate.restype = C.c_int
        L.orc_point_polygon_locate.argtypes = [C.c_double, C.c_double, f64p, f64p, u64p, C.c_uint64, C.c_uint64]
        L.orc_distance_pp.restype = C.c_double
        L.orc_distance_pp.argtypes = [C.c_double] * 4
        L.orc_join_points_boxes.restype = C.c_uint64
        L.orc_dwithin_brute.restype = C.c_uint64
        L.orc_rtree_build_from_points.restype = C.c_int
        _LIB = L
    return _LIB


def ref():
    """The compiled reference (oracle/_ref/libref.so) or None when it was never built."""
    global _REF
    if _REF is None:
        path = os.path.join(_HERE, "_ref", "libref.so")
        if not os.path.exists(path):
            try:
                build()
            except Exception:
                pass
        if not os.path.exists(path):
            return None
        R = C.CDLL(path, mode=os.RTLD_LAZY)
        R.ref_double_to_float_down.restype = C.c_float
        R.ref_double_to_float_down.argtypes = [C.c_double]
        R.ref_double_to_float_up.restype = C.c_float
        R.ref_double_to_float_up.argtypes = [C.c_double]
        R.ref_hilbert_encode.restype = C.c_uint32
        R.ref_hilbert_encode.argtypes = [C.c_uint32] * 3
        R.ref_hilbert_f32_to_u32.restype = C.c_uint32
        R.ref_hilbert_f32_to_u32.argtypes = [C.c_float]
        R.ref_cast_f32_u32.restype = C.c_uint32
        R.ref_cast_f32_u32.argtypes = [C.c_float]
        R.ref_cast_f64_u32.restype = C.c_uint32
        R.ref_cast_f64_u32.argtypes = [C.c_double]
        R.ref_flat_build.restype = C.c_void_p
        R.ref_flat_free.argtypes = [C.c_void_p]
        R.ref_flat_nodes.restype = C.c_uint32
        R.ref_flat_nodes.argtypes = [C.c_void_p]
        R.ref_flat_nlayers.restype = C.c_uint32
        R.ref_flat_nlayers.argtypes = [C.c_void_p]
        R.ref_flat_export.argtypes = [C.c_void_p, f32p, u32p, u32p]
        R.ref_flat_lookup.restype = C.c_uint64
        R.ref_flat_lookup.argtypes = [C.c_void_p, f32p, u32p, C.c_uint64]
        R.ref_flat_join_count.restype = C.c_uint64
        R.ref_flat_join_hash.restype = C.c_uint64
        R.ref_cached_bounds.restype = C.c_int
        _REF = R
    return _REF


_REF_SCALAR = None


def ref_scalar():
    """The reference's scalar-function TU (oracle/_ref/libref_scalar.so, see ref_scalar.cpp) or None."""
    global _REF_SCALAR
    if _REF_SCALAR is None:
        path = os.path.join(_HERE, "_ref", "libref_scalar.so")
        if not os.path.exists(path):
            try:
                build()
            except Exception:
                pass
        if not os.path.exists(path):
            return None
        R = C.CDLL(path, mode=os.RTLD_LAZY)
        R.ref_dist_to_segment_squared.restype = C.c_double
        R.ref_dist_to_segment_squared.argtypes = [C.c_double] * 6
        _REF_SCALAR = R
    return _REF_SCALAR


# ---------------------------------------------------------------------------------------------------
# generators / serialisation
# ---------------------------------------------------------------------------------------------------
def generate_points(seed, n, bounds=(0.0, 0.0, 10000.0, 10000.0), skip=0):
    """ST_GeneratePoints(bounds, n, seed) -> (x, y) float64 arrays."""
    x = np.empty(n, np.float64)
    y = np.empty(n, np.float64)
    lib().orc_generate_points(
        C.c_uint64(seed), C.c_uint64(skip), C.c_uint64(n), *[C.c_double(b) for b in bounds], _p(x, f64p), _p(y, f64p)
    )
    return x, y


def pcg32_stream(seed, n, skip=0):
    out = np.empty(n, np.uint32)
    lib().orc_pcg32_fill(C.c_uint64(seed), C.c_uint64(skip), C.c_uint64(n), _p(out, u32p))
    return out


def uniform01(seed, n, skip=0):
    """ldexp(pcg32, -32) stream (RandomEngine::NextRandom32)."""
    return pcg32_stream(seed, n, skip).astype(np.float64) * (2.0 ** -32)


def point_blobs(x, y):
    n = len(x)
    out = np.empty(32 * n, np.uint8)
    lib().orc_write_point_blobs(_p(np.ascontiguousarray(x), f64p), _p(np.ascontiguousarray(y), f64p), C.c_uint64(n), _p(out, u8p))
    off = np.arange(n + 1, dtype=np.uint64) * 32
    return out, off


def envelope_blobs(minx, miny, maxx, maxy):
    n = len(minx)
    out = np.empty(120 * n, np.uint8)
    a = [np.ascontiguousarray(v, np.float64) for v in (minx, miny, maxx, maxy)]
    lib().orc_write_envelope_blobs(*[_p(v, f64p) for v in a], C.c_uint64(n), _p(out, u8p))
    off = np.arange(n + 1, dtype=np.uint64) * 120
    return out, off


def polygon_blobs(vx, vy, voff):
    """single-ring polygons; ring i = vertices [voff[i], voff[i+1])"""
    vx = np.ascontiguousarray(vx, np.float64)
    vy = np.ascontiguousarray(vy, np.float64)
    voff = np.ascontiguousarray(voff, np.uint64)
    n = len(voff) - 1
    total = lib().orc_write_polygon_blobs(_p(vx, f64p), _p(vy, f64p), _p(voff, u64p), C.c_uint64(n), None, None)
    out = np.empty(total, np.uint8)
    off = np.empty(n + 1, np.uint64)
    lib().orc_write_polygon_blobs(_p(vx, f64p), _p(vy, f64p), _p(voff, u64p), C.c_uint64(n), _p(out, u8p), _p(off, u64p))
    return out, off


def concat_blobs(blobs):
    """list of bytes -> (data uint8, offsets uint64)"""
    off = np.zeros(len(blobs) + 1, np.uint64)
    if blobs:
        off[1:] = np.cumsum([len(b) for b in blobs])
    data = np.frombuffer(b"".join(blobs), np.uint8).copy() if blobs else np.zeros(0, np.uint8)
    return data, off


# ---------------------------------------------------------------------------------------------------
# bbox / extent / hilbert
# ---------------------------------------------------------------------------------------------------
def bbox_from_points(x, y):
    n = len(x)
    out = np.empty((n, 4), np.float32)
    lib().orc_bbox_from_points(_p(np.ascontiguousarray(x), f64p), _p(np.ascontiguousarray(y), f64p), C.c_uint64(n), _p(out, f32p))
    return out


def bbox_from_box2d(minx, miny, maxx, maxy):
    n = len(minx)
    out = np.empty((n, 4), np.float32)
    a = [np.ascontiguousarray(v, np.float64) for v in (minx, miny, maxx, maxy)]
    lib().orc_bbox_from_box2d(*[_p(v, f64p) for v in a], C.c_uint64(n), _p(out, f32p))
    return out


def bbox_from_blobs(data, off):
    n = len(off) - 1
    box = np.zeros((n, 4), np.float32)
    valid = np.zeros(n, np.uint8)
    status = np.zeros(n, np.int32)
    lib().orc_bbox_from_blobs(_p(data, u8p), _p(off, u64p), C.c_uint64(n), _p(box, f32p), _p(valid, u8p), _p(status, i32p))
    return box, valid, status


def extent_from_blobs(data, off):
    n = len(off) - 1
    box = np.zeros((n, 4), np.float64)
    valid = np.zeros(n, np.uint8)
    status = np.zeros(n, np.int32)
    lib().orc_extent_from_blobs(_p(data, u8p), _p(off, u64p), C.c_uint64(n), _p(box, f64p), _p(valid, u8p), _p(status, i32p))
    return box, valid, status


def extent_from_wkb(data, off):
    """ST_Extent(WKB_BLOB): (box [n,4] f64, valid, status)."""
    n = len(off) - 1
    box = np.zeros((n, 4), np.float64)
    valid = np.zeros(n, np.uint8)
    status = np.zeros(n, np.int32)
    lib().orc_extent_from_wkb(_p(data, u8p), _p(off, u64p), C.c_uint64(n), _p(box, f64p), _p(valid, u8p), _p(status, i32p))
    return box, valid, status


def ref_wkb_extent(blob):
    """The reference's own wkb_reader_try_parse_stats on one blob (libref.so): (rc, box4, vertex_count)."""
    R = ref()
    n = len(blob)
    b = np.frombuffer(bytes(blob) + b"\0", np.uint8)  # never a NULL buffer (the reader asserts on it)
    box = np.zeros(4, np.float64)
    vc = C.c_uint64(0)
    R.ref_wkb_extent.restype = C.c_int
    rc = R.ref_wkb_extent(_p(b, u8p), C.c_uint64(n), _p(box, f64p), C.byref(vc))
    return rc, box, vc.value


def points_from_blobs(data, off):
    """GEOMETRY -> POINT_2D cast: (x, y, status)."""
    n = len(off) - 1
    x = np.zeros(n, np.float64)
    y = np.zeros(n, np.float64)
    status = np.zeros(n, np.int32)
    lib().orc_points_from_blobs(_p(data, u8p), _p(off, u64p), C.c_uint64(n), _p(x, f64p), _p(y, f64p), _p(status, i32p))
    return x, y, status


def polygons_from_blobs(data, off):
    """GEOMETRY -> POLYGON_2D cast in list layout: (poly_ring_off, ring_vert_off, vx, vy, status)."""
    n = len(off) - 1
    poly_off = np.zeros(n + 1, np.uint64)
    status = np.zeros(n, np.int32)
    nr, nv = C.c_uint64(0), C.c_uint64(0)
    lib().orc_polygons_from_blobs(_p(data, u8p), _p(off, u64p), C.c_uint64(n), _p(poly_off, u64p), None, None, None,
                                  C.byref(nr), C.byref(nv), _p(status, i32p))
    ring_off = np.zeros(nr.value + 1, np.uint64)
    vx = np.zeros(nv.value, np.float64)
    vy = np.zeros(nv.value, np.float64)
    lib().orc_polygons_from_blobs(_p(data, u8p), _p(off, u64p), C.c_uint64(n), _p(poly_off, u64p), _p(ring_off, u64p),
                                  _p(vx, f64p), _p(vy, f64p), C.byref(nr), C.byref(nv), _p(status, i32p))
    return poly_off, ring_off, vx, vy, status


def hilbert_xy(x, y, bounds):
    n = len(x)
    out = np.empty(n, np.uint32)
    b = np.asarray(bounds, np.float64)
    lib().orc_hilbert_xy(_p(np.ascontiguousarray(x), f64p), _p(np.ascontiguousarray(y), f64p), C.c_uint64(n), _p(b, f64p), _p(out, u32p))
    return out


def hilbert_box2d(box4, bounds):
    box4 = np.ascontiguousarray(box4, np.float64)
    out = np.empty(len(box4), np.uint32)
    b = np.asarray(bounds, np.float64)
    lib().orc_hilbert_box2d(_p(box4, f64p), C.c_uint64(len(box4)), _p(b, f64p), _p(out, u32p))
    return out


def hilbert_box2df(box4, bounds):
    box4 = np.ascontiguousarray(box4, np.float32)
    out = np.empty(len(box4), np.uint32)
    b = np.asarray(bounds, np.float32)
    lib().orc_hilbert_box2df(_p(box4, f32p), C.c_uint64(len(box4)), _p(b, f32p), _p(out, u32p))
    return out


def hilbert_geom(data, off):
    n = len(off) - 1
    out = np.empty(n, np.uint32)
    valid = np.empty(n, np.uint8)
    lib().orc_hilbert_geom(_p(data, u8p), _p(off, u64p), C.c_uint64(n), _p(out, u32p), _p(valid, u8p))
    return out, valid


def hilbert_geom_bounds(data, off, bounds):
    n = len(off) - 1
    out = np.empty(n, np.uint32)
    status = np.empty(n, np.int32)
    b = np.asarray(bounds, np.float64)
    lib().orc_hilbert_geom_bounds(_p(data, u8p), _p(off, u64p), C.c_uint64(n), _p(b, f64p), _p(out, u32p), _p(status, i32p))
    return out, status


# ---------------------------------------------------------------------------------------------------
# STR R-tree
# ---------------------------------------------------------------------------------------------------
ENTRY_DTYPE = np.dtype([("ptr", np.uint64), ("box", np.float32, 4)])


class StrTree:
    def __init__(self, box4, rowid, cap=128):
        box4 = np.ascontiguousarray(box4, np.float32)
        rowid = np.ascontiguousarray(rowid, np.int64)
        self.cap = cap
        self.h = lib().orc_rtree_build_str(_p(box4, f32p), _p(rowid, i64p), C.c_uint64(len(rowid)), C.c_uint32(cap))

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_rtree_free(self.h)
            self.h = None

    @property
    def nlevels(self):
        return lib().orc_rtree_nlevels(self.h)

    @property
    def slice(self):
        return lib().orc_rtree_slice(self.h)

    def level_counts(self):
        return [lib().orc_rtree_level_count(self.h, l) for l in range(self.nlevels)]

    def level(self, l):
        n = lib().orc_rtree_level_count(self.h, l)
        ptr = lib().orc_rtree_level_entries(self.h, l)
        buf = (C.c_uint8 * (24 * n)).from_address(ptr)
        return np.frombuffer(buf, ENTRY_DTYPE).copy()

    def dump(self):
        n = lib().orc_rtree_dump(C.c_void_p(self.h), None, None, None)
        lvl = np.empty(n, np.int32)
        box = np.empty((n, 4), np.float32)
        rid = np.empty(n, np.int64)
        lib().orc_rtree_dump(C.c_void_p(self.h), _p(lvl, i32p), _p(box, f32p), _p(rid, i64p))
        return lvl, box, rid

    def pages(self, block_size=262136):
        """FixedSizeAllocator block images of the tree -> (counts dict, leaf images, branch images, root entry)."""
        L = lib()
        L.orc_rtree_export_pages.restype = C.c_int
        L.orc_fsa_layout.restype = C.c_int
        lay = (C.c_uint64 * 3)()
        if L.orc_fsa_layout(C.c_uint64(24 + 24 * self.cap), C.c_uint64(block_size), lay):
            raise ValueError("segment does not fit the block")
        cnt = (C.c_uint64 * 4)()
        root = np.zeros(1, ENTRY_DTYPE)
        L.orc_rtree_export_pages(C.c_void_p(self.h), C.c_uint64(block_size), None, None, _p(root, C.c_void_p), cnt)
        leaf = np.empty(cnt[1] * block_size, np.uint8)
        branch = np.empty(cnt[3] * block_size, np.uint8)
        L.orc_rtree_export_pages(C.c_void_p(self.h), C.c_uint64(block_size), _p(leaf, C.c_void_p),
                                 _p(branch, C.c_void_p), _p(root, C.c_void_p), cnt)
        info = dict(segment_size=24 + 24 * self.cap, segments_per_buffer=lay[0], bitmask_count=lay[1],
                    bitmask_offset=lay[2], leaf_pages=cnt[0], leaf_buffers=cnt[1], branch_pages=cnt[2],
                    branch_buffers=cnt[3])
        return info, leaf, branch, root[0]

    def scan(self, qbox):
        q = np.asarray(qbox, np.float32)
        n = lib().orc_rtree_scan(C.c_void_p(self.h), _p(q, f32p), None)
        out = np.empty(n, np.int64)
        lib().orc_rtree_scan(C.c_void_p(self.h), _p(q, f32p), _p(out, i64p))
        return out


# ---------------------------------------------------------------------------------------------------
# FlatRTree
# ---------------------------------------------------------------------------------------------------
class FlatTree:
    """sort_mode 0 = reference quicksort order, 1 = stable (curve, index) order (CUDA order)."""

    def __init__(self, box4, node_size=32, sort_mode=1):
        box4 = np.ascontiguousarray(box4, np.float32)
        self.n = len(box4)
        self.h = lib().orc_flat_build(_p(box4, f32p), C.c_uint32(self.n), C.c_uint32(node_size), C.c_int(sort_mode))

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_flat_free(self.h)
            self.h = None

    @property
    def nodes(self):
        return lib().orc_flat_nodes(self.h)

    @property
    def layer_bounds(self):
        nl = lib().orc_flat_nlayers(self.h)
        return np.ctypeslib.as_array(lib().orc_flat_layer_bounds(self.h), (nl,)).copy()

    @property
    def boxes(self):
        return np.ctypeslib.as_array(lib().orc_flat_boxes(self.h), (self.nodes, 4)).copy()

    @property
    def idx(self):
        return np.ctypeslib.as_array(lib().orc_flat_idx(self.h), (self.nodes,)).copy()

    @property
    def tree_box(self):
        return np.ctypeslib.as_array(lib().orc_flat_tree_box(self.h), (4,)).copy()

    def lookup(self, qbox):
        q = np.asarray(qbox, np.float32)
        out = np.empty(max(self.n, 1), np.uint32)
        n = lib().orc_flat_lookup(C.c_void_p(self.h), _p(q, f32p), _p(out, u32p), C.c_uint64(len(out)))
        return out[:n].copy()


def flat_curve(box4, tree_box):
    box4 = np.ascontiguousarray(box4, np.float32)
    tb = np.asarray(tree_box, np.float32)
    out = np.empty(len(box4), np.uint32)
    lib().orc_flat_curve(_p(box4, f32p), C.c_uint32(len(box4)), _p(tb, f32p), _p(out, u32p))
    return out


# ---------------------------------------------------------------------------------------------------
# refinement / joins
# ---------------------------------------------------------------------------------------------------
def join_points_boxes(px, py, bminx, bminy, bmaxx, bmaxy, pred=0, threads=0, want_pairs=True):
    """(pairs[n,2] uint32 sorted probe-major, n_candidates)"""
    px, py = np.ascontiguousarray(px, np.float64), np.ascontiguousarray(py, np.float64)
    b = [np.ascontiguousarray(v, np.float64) for v in (bminx, bminy, bmaxx, bmaxy)]
    cand = C.c_uint64(0)
    args = [_p(px, f64p), _p(py, f64p), C.c_uint64(len(px))] + [_p(v, f64p) for v in b] + [C.c_uint32(len(b[0])), C.c_int(pred), C.c_int(threads)]
    n = lib().orc_join_points_boxes(*args, None, None, C.c_uint64(0), C.byref(cand))
    if not want_pairs:
        return n, cand.value
    op = np.empty(n, np.uint32)
    ob = np.empty(n, np.uint32)
    lib().orc_join_points_boxes(*args, _p(op, u32p), _p(ob, u32p), C.c_uint64(n), C.byref(cand))
    return np.stack([op, ob], 1), cand.value


def pair_checksum(pairs_probe, pairs_build):
    """(count, hash): the order-independent checksum of include/spatial_b200.h's sb_pairs_checksum, in numpy"""
    p = np.asarray(pairs_probe).astype(np.uint64)
    b = np.asarray(pairs_build).astype(np.uint64)
    with np.errstate(over="ignore"):
        k = p * np.uint64(2654435761) + b * np.uint64(2246822519)
        k ^= k >> np.uint64(29)
        k *= np.uint64(0xBF58476D1CE4E5B9)
        k ^= k >> np.uint64(32)
        return len(p), int(k.sum(dtype=np.uint64))


def pip_native(x, y, vx, vy, ring_off, r0, r1):
    return lib().orc_pip_native(x, y, _p(vx, f64p), _p(vy, f64p), _p(ring_off, u64p), C.c_uint64(r0), C.c_uint64(r1))


def orient2d_sign(ax, ay, bx, by, cx, cy):
    """exact sign of (b-a) x (c-a): +1 left / counter-clockwise, -1 right, 0 collinear"""
    return lib().orc_orient2d_sign(ax, ay, bx, by, cx, cy)


def point_polygon_locate(x, y, vx, vy, ring_off, r0, r1):
    """GEOS point-in-area location (exact orientation): 0 exterior, 1 boundary, 2 interior"""
    return lib().orc_point_polygon_locate(x, y, _p(vx, f64p), _p(vy, f64p), _p(ring_off, u64p), C.c_uint64(r0), C.c_uint64(r1))


def point_polygon_exact(px, py, pi, qi, vx, vy, ring_off, poly_off, pred=0):
    """pairs (pi[k], qi[k]) -> uint8 mask; pred 0 = ST_Intersects, 1 = ST_Within(point, polygon) (GEOS semantics)"""
    px, py, vx, vy = [np.ascontiguousarray(v, np.float64) for v in (px, py, vx, vy)]
    ring_off = np.ascontiguousarray(ring_off, np.uint64)
    poly_off = np.ascontiguousarray(poly_off, np.uint64)
    pi = None if pi is None else np.ascontiguousarray(pi, np.uint32)
    qi = None if qi is None else np.ascontiguousarray(qi, np.uint32)
    n = len(pi) if pi is not None else (len(qi) if qi is not None else len(px))
    out = np.empty(n, np.uint8)
    lib().orc_point_polygon_exact(_p(px, f64p), _p(py, f64p), _p(pi, u32p), _p(qi, u32p), C.c_uint64(n), _p(vx, f64p),
                                  _p(vy, f64p), _p(ring_off, u64p), _p(poly_off, u64p), C.c_int(pred), _p(out, u8p))
    return out


def dwithin_brute(ax, ay, bx, by, r):
    ax, ay, bx, by = [np.ascontiguousarray(v, np.float64) for v in (ax, ay, bx, by)]
    args = [_p(ax, f64p), _p(ay, f64p), C.c_uint64(len(ax)), _p(bx, f64p), _p(by, f64p), C.c_uint64(len(bx)), C.c_double(r)]
    n = lib().orc_dwithin_brute(*args, None, None, None, C.c_uint64(0))
    oa = np.empty(n, np.uint32)
    ob = np.empty(n, np.uint32)
    od = np.empty(n, np.float64)
    lib().orc_dwithin_brute(*args, _p(oa, u32p), _p(ob, u32p), _p(od, f64p), C.c_uint64(n))
    return np.stack([oa, ob], 1), od


def box2d_intersects(a_box4, b_box4, valid_a=None, valid_b=None):
    a = np.ascontiguousarray(a_box4, np.float64).reshape(-1, 4)
    b = np.ascontiguousarray(b_box4, np.float64).reshape(-1, 4)
    va = None if valid_a is None else np.ascontiguousarray(valid_a, np.uint8)
    vb = None if valid_b is None else np.ascontiguousarray(valid_b, np.uint8)
    out = np.empty(len(a), np.uint8)
    u8p = C.POINTER(C.c_uint8)
    lib().orc_box2d_intersects(_p(a, f64p), _p(b, f64p), None if va is None else _p(va, u8p),
                               None if vb is None else _p(vb, u8p), C.c_uint64(len(a)), _p(out, u8p))
    return out


def distance_point_linestring(px, py, lists, vx, vy):
    px, py, vx, vy = [np.ascontiguousarray(v, np.float64) for v in (px, py, vx, vy)]
    lists = np.ascontiguousarray(lists, np.uint64).reshape(-1, 2)
    out = np.empty(len(px), np.float64)
    lib().orc_distance_point_linestring(_p(px, f64p), _p(py, f64p), _p(lists, C.POINTER(C.c_uint64)), C.c_uint64(len(px)),
                                        _p(vx, f64p), _p(vy, f64p), _p(out, f64p))
    return out
