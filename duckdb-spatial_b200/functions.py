"""Host-side mirror of the reference's function / index / operator surface for the hot path, over the C-ABI.

Names follow the reference (ST_Extent_Approx, ST_Extent, ST_Hilbert, RTreeIndex, FlatRTree, SpatialJoin).
Inputs may be numpy arrays (host memory: the C-ABI stages them through HBM, SB_HOST) or torch CUDA tensors
(device memory, SB_DEVICE, zero-copy). Outputs come back in the same kind of container.
"""
import ctypes as C

import numpy as np

from . import capi

ENTRY_DTYPE = np.dtype([("ptr", np.uint64), ("box", np.float32, 4)])  # RTreeEntry, rtree_node.hpp:66-75


def _is_torch(a):
    return a is not None and type(a).__module__.split(".")[0] == "torch"


def _ptr(a):
    if a is None:
        return None
    if _is_torch(a):
        assert a.is_contiguous()
        return a.data_ptr()
    assert a.flags["C_CONTIGUOUS"]
    return a.ctypes.data


def _pred_code(name):
    return {"intersects": capi.PRED_INTERSECTS, "within": capi.PRED_WITHIN, "contains": capi.PRED_WITHIN}[name]


class Context:
    """One per GPU / calling thread (sb_context): owns a stream and a scratch arena."""

    def __init__(self, device=0):
        self.L = capi.lib()
        h = C.c_void_p()
        capi.check(self.L.sb_context_create(device, C.byref(h)))
        self.h = h
        self.device = device

    def close(self):
        if getattr(self, "h", None):
            self.L.sb_context_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- plumbing ------------------------------------------------------------------------------------------
    def use_torch_stream(self):
        import torch

        # torch's default stream has handle 0, which sb_context_set_stream reads as "back to the private stream":
        # name the legacy default stream explicitly (cudaStreamLegacy = 0x1) so the work is ordered with torch's
        h = torch.cuda.current_stream().cuda_stream
        capi.check(self.L.sb_context_set_stream(self.h, h if h else 1))

    def use_private_stream(self):
        capi.check(self.L.sb_context_set_stream(self.h, None))

    def set_option(self, key, value):
        capi.check(self.L.sb_context_set_option(self.h, key.encode(), int(value)))

    def sync(self):
        capi.check(self.L.sb_context_sync(self.h))

    @property
    def launches(self):
        return self.L.sb_context_launches(self.h)

    def profile(self, enable=True):
        capi.check(self.L.sb_context_profile(self.h, 1 if enable else 0))

    def profile_read(self, tag):
        ms, n = C.c_double(0), C.c_uint64(0)
        capi.check(self.L.sb_context_profile_read(self.h, tag.encode(), C.byref(ms), C.byref(n)))
        return ms.value, n.value

    def _mem(self, *arrays):
        kinds = {_is_torch(a) for a in arrays if a is not None}
        assert len(kinds) == 1, "all arrays of one call must be numpy (host) or torch CUDA (device)"
        return capi.DEVICE if kinds.pop() else capi.HOST

    def _empty(self, like, shape, dtype):
        if _is_torch(like):
            import torch

            tdt = {np.float32: torch.float32, np.float64: torch.float64, np.uint32: torch.int32, np.uint8: torch.uint8,
                   np.int32: torch.int32, np.int64: torch.int64, np.uint64: torch.int64}[dtype]
            return torch.empty(shape, dtype=tdt, device=like.device)
        return np.empty(shape, dtype)

    @staticmethod
    def _f64(a):
        return a if _is_torch(a) else np.ascontiguousarray(a, np.float64)

    # -- harness: ST_GeneratePoints ------------------------------------------------------------------------
    def st_generate_points(self, seed, n, bounds=(0.0, 0.0, 10000.0, 10000.0), skip=0, device=False):
        """points [skip, skip+n) of ST_GeneratePoints(bounds, *, seed): numpy arrays, or torch CUDA tensors."""
        b = np.ascontiguousarray(bounds, np.float64)
        if device:
            import torch

            x = torch.empty(n, dtype=torch.float64, device="cuda:%d" % self.device)
            y = torch.empty_like(x)
        else:
            x, y = np.empty(n, np.float64), np.empty(n, np.float64)
        capi.check(self.L.sb_generate_points(self.h, self._mem(x, y), seed, skip, n, _ptr(b), _ptr(x), _ptr(y)))
        return x, y

    # -- K1 ------------------------------------------------------------------------------------------------
    def bbox_from_points(self, x, y):
        """POINT_2D -> BOX_2DF (TryGetCachedBounds point branch)."""
        x, y = self._f64(x), self._f64(y)
        out = self._empty(x, (len(x), 4), np.float32)
        capi.check(self.L.sb_bbox_from_points(self.h, self._mem(x, y), _ptr(x), _ptr(y), len(x), _ptr(out)))
        return out

    def bbox_from_box2d(self, minx, miny, maxx, maxy):
        a = [self._f64(v) for v in (minx, miny, maxx, maxy)]
        out = self._empty(a[0], (len(a[0]), 4), np.float32)
        capi.check(self.L.sb_bbox_from_box2d(self.h, self._mem(*a), *[_ptr(v) for v in a], len(a[0]), _ptr(out)))
        return out

    def st_extent_approx(self, data, offsets, want_status=False):
        """ST_Extent_Approx(GEOMETRY) -> (BOX_2DF [n,4], valid [n])"""
        n = len(offsets) - 1
        box = self._empty(data, (n, 4), np.float32)
        valid = self._empty(data, (n,), np.uint8)
        status = self._empty(data, (n,), np.int32)
        rc = self.L.sb_bbox_from_blobs(self.h, self._mem(data, offsets), _ptr(data), _ptr(offsets), n, _ptr(box),
                                       _ptr(valid), _ptr(status))
        if want_status:
            return box, valid, status, rc
        capi.check(rc)
        return box, valid

    def st_extent(self, data, offsets, want_status=False):
        """ST_Extent(GEOMETRY) -> (BOX_2D [n,4], valid [n])"""
        n = len(offsets) - 1
        box = self._empty(data, (n, 4), np.float64)
        valid = self._empty(data, (n,), np.uint8)
        status = self._empty(data, (n,), np.int32)
        rc = self.L.sb_extent_from_blobs(self.h, self._mem(data, offsets), _ptr(data), _ptr(offsets), n, _ptr(box),
                                         _ptr(valid), _ptr(status))
        if want_status:
            return box, valid, status, rc
        capi.check(rc)
        return box, valid

    def st_extent_wkb(self, data, offsets, want_status=False):
        """ST_Extent(WKB_BLOB) -> (BOX_2D [n,4], valid [n])"""
        n = len(offsets) - 1
        box = self._empty(data, (n, 4), np.float64)
        valid = self._empty(data, (n,), np.uint8)
        status = self._empty(data, (n,), np.int32)
        rc = self.L.sb_extent_from_wkb(self.h, self._mem(data, offsets), _ptr(data), _ptr(offsets), n, _ptr(box),
                                       _ptr(valid), _ptr(status))
        if want_status:
            return box, valid, status, rc
        capi.check(rc)
        return box, valid

    # -- GEOMETRY -> coordinate columns --------------------------------------------------------------------
    def points_from_blobs(self, data, offsets, want_status=False):
        """GEOMETRY -> POINT_2D cast (PointCasts::FromGeometryCast): (x, y); raises where the reference throws."""
        n = len(offsets) - 1
        x = self._empty(data, (n,), np.float64)
        y = self._empty(data, (n,), np.float64)
        status = self._empty(data, (n,), np.int32)
        rc = self.L.sb_points_from_blobs(self.h, self._mem(data, offsets), _ptr(data), _ptr(offsets), n, _ptr(x), _ptr(y),
                                         _ptr(status))
        if want_status:
            return x, y, status, rc
        capi.check(rc)
        return x, y

    def polygons_from_blobs(self, data, offsets, want_status=False):
        """GEOMETRY -> POLYGON_2D cast in list layout: (poly_ring_off [n+1], ring_vert_off [rings+1], vx, vy).
        Two calls through the C-ABI: the first sizes the outputs (SB_ERR_CAPACITY with the counts)."""
        n = len(offsets) - 1
        mem = self._mem(data, offsets)
        poly_off = self._empty(data, (n + 1,), np.uint64)
        status = self._empty(data, (n,), np.int32)
        nr, nv = C.c_uint64(0), C.c_uint64(0)
        rc = self.L.sb_polygons_from_blobs(self.h, mem, _ptr(data), _ptr(offsets), n, _ptr(poly_off), None, None, None,
                                           0, 0, C.byref(nr), C.byref(nv), _ptr(status))
        if rc not in (capi.OK, capi.ERR_CAPACITY, capi.ERR_GEOMETRY):
            capi.check(rc)
        ring_off = self._empty(data, (nr.value + 1,), np.uint64)
        vx = self._empty(data, (nv.value,), np.float64)
        vy = self._empty(data, (nv.value,), np.float64)
        rc = self.L.sb_polygons_from_blobs(self.h, mem, _ptr(data), _ptr(offsets), n, _ptr(poly_off), _ptr(ring_off),
                                           _ptr(vx), _ptr(vy), nr.value, nv.value, C.byref(nr), C.byref(nv), _ptr(status))
        if want_status:
            return poly_off, ring_off, vx, vy, status, rc
        capi.check(rc)
        return poly_off, ring_off, vx, vy

    # -- K2 ------------------------------------------------------------------------------------------------
    def st_hilbert_xy(self, x, y, bounds):
        x, y = self._f64(x), self._f64(y)
        b = np.ascontiguousarray(bounds, np.float64)
        out = self._empty(x, (len(x),), np.uint32)
        capi.check(self.L.sb_hilbert_xy(self.h, self._mem(x, y), _ptr(x), _ptr(y), len(x), _ptr(b), _ptr(out)))
        return out

    def st_hilbert_box2d(self, box4, bounds):
        box4 = box4 if _is_torch(box4) else np.ascontiguousarray(box4, np.float64)
        b = np.ascontiguousarray(bounds, np.float64)
        out = self._empty(box4, (len(box4),), np.uint32)
        capi.check(self.L.sb_hilbert_box2d(self.h, self._mem(box4), _ptr(box4), len(box4), _ptr(b), _ptr(out)))
        return out

    def st_hilbert_box2df(self, box4, bounds):
        box4 = box4 if _is_torch(box4) else np.ascontiguousarray(box4, np.float32)
        b = np.ascontiguousarray(bounds, np.float32)
        out = self._empty(box4, (len(box4),), np.uint32)
        capi.check(self.L.sb_hilbert_box2df(self.h, self._mem(box4), _ptr(box4), len(box4), _ptr(b), _ptr(out)))
        return out

    def st_hilbert_geom(self, data, offsets):
        n = len(offsets) - 1
        out = self._empty(data, (n,), np.uint32)
        valid = self._empty(data, (n,), np.uint8)
        capi.check(self.L.sb_hilbert_geom(self.h, self._mem(data, offsets), _ptr(data), _ptr(offsets), n, _ptr(out),
                                          _ptr(valid)))
        return out, valid

    def st_hilbert_geom_bounds(self, data, offsets, bounds, want_status=False):
        n = len(offsets) - 1
        b = np.ascontiguousarray(bounds, np.float64)
        out = self._empty(data, (n,), np.uint32)
        status = self._empty(data, (n,), np.int32)
        rc = self.L.sb_hilbert_geom_bounds(self.h, self._mem(data, offsets), _ptr(data), _ptr(offsets), n, _ptr(b),
                                           _ptr(out), _ptr(status))
        if want_status:
            return out, status, rc
        capi.check(rc)
        return out

    # -- K3 ------------------------------------------------------------------------------------------------
    def sort_pairs(self, keys, values):
        """in-place stable sort of (keys, values); keys uint32 or uint64 (torch: int32 / int64 bit patterns)."""
        itemsize = keys.element_size() if _is_torch(keys) else keys.dtype.itemsize
        fn = self.L.sb_sort_pairs_u32 if itemsize == 4 else self.L.sb_sort_pairs_u64
        capi.check(fn(self.h, self._mem(keys, values), _ptr(keys), _ptr(values), len(keys)))
        return keys, values

    # -- K6 ------------------------------------------------------------------------------------------------
    def st_distance_points(self, ax, ay, bx, by):
        a = [self._f64(v) for v in (ax, ay, bx, by)]
        out = self._empty(a[0], (len(a[0]),), np.float64)
        capi.check(self.L.sb_distance_points(self.h, self._mem(*a), *[_ptr(v) for v in a], len(a[0]), _ptr(out)))
        return out

    def st_distance_point_linestring(self, px, py, list_entries, vx, vy):
        """ST_Distance(POINT_2D, LINESTRING_2D); list_entries = [n,2] uint64 (offset, length) rows into vx / vy."""
        px, py, vx, vy = [self._f64(v) for v in (px, py, vx, vy)]
        le = list_entries if _is_torch(list_entries) else np.ascontiguousarray(list_entries, np.uint64).reshape(-1, 2)
        out = self._empty(px, (len(px),), np.float64)
        capi.check(self.L.sb_distance_point_linestring(self.h, self._mem(px, py, le, vx, vy), _ptr(px), _ptr(py),
                                                       _ptr(le), len(px), _ptr(vx), _ptr(vy), len(vx), _ptr(out)))
        return out

    def st_intersects_box2d(self, a_box4, b_box4, valid_a=None, valid_b=None):
        """ST_Intersects(BOX_2D, BOX_2D) over AoS [n,4] float64 boxes -> uint8[n]; rows whose validity byte is 0 are 0."""
        a = a_box4 if _is_torch(a_box4) else np.ascontiguousarray(a_box4, np.float64).reshape(-1, 4)
        b = b_box4 if _is_torch(b_box4) else np.ascontiguousarray(b_box4, np.float64).reshape(-1, 4)
        out = self._empty(a, (len(a),), np.uint8)
        pa, pb = _ptr(a), _ptr(b)
        cols = [pa + 8 * c for c in range(4)] + [pb + 8 * c for c in range(4)]
        capi.check(self.L.sb_intersects_box2d(self.h, self._mem(a, b), *cols, 4, _ptr(valid_a), _ptr(valid_b),
                                              len(a), _ptr(out)))
        return out

    def st_intersects_box2d_columns(self, a_cols, b_cols):
        """same over DuckDB's struct child vectors: two lists of four float64 columns (min_x, min_y, max_x, max_y)."""
        cols = [self._f64(v) for v in list(a_cols) + list(b_cols)]
        out = self._empty(cols[0], (len(cols[0]),), np.uint8)
        capi.check(self.L.sb_intersects_box2d(self.h, self._mem(*cols), *[_ptr(v) for v in cols], 1, None, None,
                                              len(cols[0]), _ptr(out)))
        return out

    def st_intersects_extent(self, data_a, off_a, data_b, off_b):
        """ST_Intersects_Extent(GEOMETRY, GEOMETRY), spatial_functions_scalar.cpp:5456-5481: the exact extents of both
        sides (sb_extent_from_blobs) and the box test; rows where either side has no extent are false."""
        ea, va = self.st_extent(data_a, off_a)
        eb, vb = self.st_extent(data_b, off_b)
        return self.st_intersects_box2d(ea, eb, va, vb)

    def st_within_point_polygon(self, px, py, vx, vy, ring_off, poly_off, pair_point, pair_poly, geos=None):
        """ST_Within(POINT_2D, POLYGON_2D) over candidate pairs -> surviving (point, polygon) pairs.
        geos="intersects" | "within": the GEOMETRY-argument (GEOS route) predicate with exact orientation instead."""
        px, py, vx, vy = [self._f64(v) for v in (px, py, vx, vy)]
        cnt = C.c_uint64(len(pair_point))
        pp = pair_point.clone() if _is_torch(pair_point) else np.ascontiguousarray(pair_point, np.uint32).copy()
        pb = pair_poly.clone() if _is_torch(pair_poly) else np.ascontiguousarray(pair_poly, np.uint32).copy()
        mem = self._mem(px, py, vx, vy, ring_off, poly_off, pp, pb)
        tail = (_ptr(px), _ptr(py), _ptr(vx), _ptr(vy), _ptr(ring_off), _ptr(poly_off), len(poly_off) - 1, len(ring_off) - 1,
                len(vx), len(px), _ptr(pp), _ptr(pb), C.byref(cnt))
        if geos is None:
            capi.check(self.L.sb_refine_point_polygon(self.h, mem, *tail))
        else:
            capi.check(self.L.sb_refine_point_polygon_exact(self.h, mem, _pred_code(geos), *tail))
        return pp[: cnt.value], pb[: cnt.value]

    # -- whole joins ---------------------------------------------------------------------------------------
    def spatial_join_points_boxes(self, px, py, bminx, bminy, bmaxx, bmaxy, predicate="intersects", capacity=None):
        """PhysicalSpatialJoin for POINT probe side x envelope-polygon build side. Returns pairs [k,2] uint32
        (probe row, build row), unordered."""
        px, py = self._f64(px), self._f64(py)
        b = [self._f64(v) for v in (bminx, bminy, bmaxx, bmaxy)]
        pred = capi.PRED_INTERSECTS if predicate == "intersects" else capi.PRED_WITHIN
        cap = int(capacity if capacity is not None else max(1024, 2 * len(px)))
        while True:
            op = self._empty(px, (cap,), np.uint32)
            ob = self._empty(px, (cap,), np.uint32)
            cnt = C.c_uint64(0)
            rc = self.L.sb_join_points_boxes(self.h, self._mem(px, py, *b), pred, _ptr(px), _ptr(py), len(px),
                                             *[_ptr(v) for v in b], len(b[0]), _ptr(op), _ptr(ob), cap, C.byref(cnt))
            if rc == capi.ERR_CAPACITY:
                cap = int(cnt.value)
                continue
            capi.check(rc)
            return op[: cnt.value], ob[: cnt.value]

    def join_mark_matched(self, pair_probe, pair_build, n_probe, n_build, probe_index_base=0, probe_matched=None,
                          build_matched=None):
        """LEFT / RIGHT / FULL bookkeeping: flag arrays (uint8) with 1 for every row that appears in the pair list.
        Existing flag arrays are OR-ed into (the build side accumulates over probe batches)."""
        like = pair_probe
        if probe_matched is None:
            probe_matched = self._zeros_u8(like, n_probe)
        if build_matched is None:
            build_matched = self._zeros_u8(like, n_build)
        capi.check(self.L.sb_join_mark_matched(self.h, self._mem(pair_probe, pair_build, probe_matched, build_matched),
                                               _ptr(pair_probe), _ptr(pair_build), len(pair_probe), probe_index_base,
                                               _ptr(probe_matched), n_probe, _ptr(build_matched), n_build))
        return probe_matched, build_matched

    def pairs_checksum(self, pair_probe, pair_build, probe_lo=0, probe_hi=0xFFFFFFFF, count=None):
        """(count, hash) of the pairs whose probe id lies in [probe_lo, probe_hi): order-independent 64-bit checksum
        (see include/spatial_b200.h); compares a join result with a reference pair set without sorting it."""
        n = len(pair_probe) if count is None else int(count)
        c, h = C.c_uint64(0), C.c_uint64(0)
        capi.check(self.L.sb_pairs_checksum(self.h, self._mem(pair_probe, pair_build), _ptr(pair_probe), _ptr(pair_build),
                                            n, probe_lo, probe_hi, C.byref(c), C.byref(h)))
        return c.value, h.value

    def join_unmatched_rows(self, matched, row_base=0):
        """ascending row ids (row_base + position) whose flag is 0."""
        n = len(matched)
        out = self._empty(matched, (max(n, 1),), np.uint32)
        cnt = C.c_uint64(0)
        capi.check(self.L.sb_join_unmatched_rows(self.h, self._mem(matched), _ptr(matched), n, row_base, _ptr(out), n,
                                                 C.byref(cnt)))
        return out[: cnt.value]

    def _zeros_u8(self, like, n):
        if _is_torch(like):
            import torch

            return torch.zeros(n, dtype=torch.uint8, device=like.device)
        return np.zeros(n, np.uint8)

    def spatial_join(self, px, py, bminx, bminy, bmaxx, bmaxy, predicate="intersects", join_type="inner", capacity=None):
        """PhysicalSpatialJoin with join type INNER / LEFT / RIGHT / FULL (spatial_join_physical.cpp:748-938,
        1022-1068). Returns (pairs_probe, pairs_build, unmatched_probe_rows, unmatched_build_rows): the unmatched
        lists are what the reference emits NULL-padded; they are empty for join types that do not propagate
        that side."""
        jt = join_type.lower()
        assert jt in ("inner", "left", "right", "full", "outer")
        p, b = self.spatial_join_points_boxes(px, py, bminx, bminy, bmaxx, bmaxy, predicate, capacity)
        empty = p[:0]
        if jt == "inner":
            return p, b, empty, empty
        pm, bm = self.join_mark_matched(p, b, len(px), len(bminx))
        up = self.join_unmatched_rows(pm) if jt in ("left", "full", "outer") else empty
        ub = self.join_unmatched_rows(bm) if jt in ("right", "full", "outer") else empty
        return p, b, up, ub

    def join_points_polygons(self, px, py, vx, vy, ring_off, poly_off, capacity=None, out=None, geos=None):
        """ST_Within(point, polygon) join over POLYGON_2D list columns -> (point rows, polygon rows), unordered.
        geos="intersects" | "within": GEOMETRY-argument (GEOS route) predicate with exact orientation instead of the
        native winding routine."""
        px, py, vx, vy = [self._f64(v) for v in (px, py, vx, vy)]
        cap = int(capacity if capacity is not None else max(1024, 2 * len(px)))
        while True:
            if out is not None:
                op, ob = out
                cap = len(op)
            else:
                op = self._empty(px, (cap,), np.uint32)
                ob = self._empty(px, (cap,), np.uint32)
            cnt = C.c_uint64(0)
            tail = (_ptr(px), _ptr(py), len(px), _ptr(vx), _ptr(vy), _ptr(ring_off), _ptr(poly_off), len(poly_off) - 1,
                    len(ring_off) - 1, len(vx), _ptr(op), _ptr(ob), cap, C.byref(cnt))
            mem = self._mem(px, py, vx, vy, ring_off, poly_off)
            if geos is None:
                rc = self.L.sb_join_points_polygons(self.h, mem, *tail)
            else:
                rc = self.L.sb_join_points_polygons_exact(self.h, mem, _pred_code(geos), *tail)
            if rc == capi.ERR_CAPACITY and out is None:
                cap = int(cnt.value)
                continue
            capi.check(rc)
            return op[: cnt.value], ob[: cnt.value]

    def join_dwithin(self, ax, ay, bx, by, radius, capacity=None):
        ax, ay, bx, by = [self._f64(v) for v in (ax, ay, bx, by)]
        cap = int(capacity if capacity is not None else max(1024, 4 * len(ax)))
        while True:
            oa = self._empty(ax, (cap,), np.uint32)
            ob = self._empty(ax, (cap,), np.uint32)
            od = self._empty(ax, (cap,), np.float64)
            cnt = C.c_uint64(0)
            rc = self.L.sb_join_dwithin(self.h, self._mem(ax, ay, bx, by), _ptr(ax), _ptr(ay), len(ax), _ptr(bx),
                                        _ptr(by), len(bx), float(radius), _ptr(oa), _ptr(ob), _ptr(od), cap,
                                        C.byref(cnt))
            if rc == capi.ERR_CAPACITY:
                cap = int(cnt.value)
                continue
            capi.check(rc)
            return oa[: cnt.value], ob[: cnt.value], od[: cnt.value]


def spatial_join_points_boxes_multi(ctxs, px, py, bminx, bminy, bmaxx, bmaxy, predicate="intersects", capacity=None, out=None):
    """ONE point x box join split over the devices of `ctxs` (contexts on distinct GPUs of this process): host (numpy or
    pinned torch CPU) columns in, one unordered pair list out -> (probe rows, build rows)."""
    arrs = [np.ascontiguousarray(v, np.float64) if not _is_torch(v) else v for v in (px, py, bminx, bminy, bmaxx, bmaxy)]
    n, m = len(arrs[0]), len(arrs[2])
    cap = int(capacity if capacity is not None else max(1024, 4 * n))
    hs = (C.c_void_p * len(ctxs))(*[c.h for c in ctxs])
    while True:
        if out is not None:
            op, ob = out
            cap = len(op)
        else:
            op, ob = np.empty(cap, np.uint32), np.empty(cap, np.uint32)
        cnt = C.c_uint64(0)
        rc = ctxs[0].L.sb_join_points_boxes_multi(hs, len(ctxs), _pred_code(predicate), *[_ptr(v) for v in arrs[:2]], n,
                                                  *[_ptr(v) for v in arrs[2:]], m, _ptr(op), _ptr(ob), cap, C.byref(cnt))
        if rc == capi.ERR_CAPACITY and out is None:
            cap = int(cnt.value)
            continue
        capi.check(rc)
        return op[: cnt.value], ob[: cnt.value]


class SpatialJoinOp:
    """PhysicalSpatialJoin for a streaming probe side (sb_join_op): the build side is finalized at construction,
    probe chunks (DuckDB: <= 2048 rows) are pushed and probed in device batches. Host (numpy) columns."""

    TYPES = {"inner": 0, "left": 1, "right": 2, "full": 3, "outer": 3}

    def __init__(self, ctx, bminx, bminy, bmaxx, bmaxy, predicate="intersects", join_type="inner", batch_rows=1 << 20):
        self.ctx = ctx
        self.jt = self.TYPES[join_type.lower()]
        b = [np.ascontiguousarray(v, np.float64) for v in (bminx, bminy, bmaxx, bmaxy)]
        self.n_build = len(b[0])
        self.batch_rows = int(batch_rows)
        pred = capi.PRED_INTERSECTS if predicate == "intersects" else capi.PRED_WITHIN
        h = C.c_void_p()
        capi.check(ctx.L.sb_join_op_create(ctx.h, pred, self.jt, *[_ptr(v) for v in b], self.n_build, self.batch_rows,
                                           C.byref(h)))
        self.h = h
        self.pending = 0

    def push(self, px, py):
        """append one probe chunk; flushes first when the chunk would not fit. Returns the flush result or None."""
        px, py = np.ascontiguousarray(px, np.float64), np.ascontiguousarray(py, np.float64)
        out = None
        if self.pending + len(px) > self.batch_rows:
            out = self.flush()
        pend = C.c_uint64(0)
        capi.check(self.ctx.L.sb_join_op_push(self.h, _ptr(px), _ptr(py), len(px), C.byref(pend)))
        self.pending = pend.value
        return out

    def flush(self, capacity=None):
        """probe the pending rows: (pair_probe, pair_build, unmatched_probe_rows)"""
        cap = int(capacity if capacity is not None else max(1024, 2 * self.pending))
        lhs = np.empty(max(self.pending, 1), np.uint32)
        while True:
            op, ob = np.empty(cap, np.uint32), np.empty(cap, np.uint32)
            n, nl = C.c_uint64(0), C.c_uint64(0)
            rc = self.ctx.L.sb_join_op_flush(self.h, _ptr(op), _ptr(ob), cap, C.byref(n), _ptr(lhs), len(lhs), C.byref(nl))
            if rc == capi.ERR_CAPACITY and n.value > cap:
                cap = int(n.value)
                continue
            capi.check(rc)
            self.pending = 0
            return op[: n.value], ob[: n.value], lhs[: nl.value]

    def finish(self):
        """after the last flush: unmatched build rows (RIGHT / FULL), ascending"""
        out = np.empty(max(self.n_build, 1), np.uint32)
        n = C.c_uint64(0)
        capi.check(self.ctx.L.sb_join_op_finish(self.h, _ptr(out), len(out), C.byref(n)))
        return out[: n.value]

    def close(self):
        if getattr(self, "h", None):
            self.ctx.L.sb_join_op_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class RTreeIndex:
    """CREATE INDEX ... USING RTREE: STR bulk load on the GPU + batched index scans (sb_rtree)."""

    def __init__(self, ctx, box4, row_ids, max_node_capacity=128):
        self.ctx = ctx
        if not _is_torch(box4):
            box4 = np.ascontiguousarray(box4, np.float32)
            row_ids = np.ascontiguousarray(row_ids, np.int64)
        h = C.c_void_p()
        capi.check(ctx.L.sb_rtree_build_str(ctx.h, ctx._mem(box4, row_ids), _ptr(box4), _ptr(row_ids), len(row_ids),
                                            max_node_capacity, C.byref(h)))
        self.h = h
        self.cap = max_node_capacity

    @classmethod
    def from_points(cls, ctx, x, y, row_ids, max_node_capacity=128):
        """CREATE INDEX over a POINT_2D column: the float boxes are computed inside the build (sb_rtree_build_from_points);
        same tree as RTreeIndex(ctx, ctx.bbox_from_points(x, y), row_ids)."""
        self = cls.__new__(cls)
        self.ctx = ctx
        self.cap = max_node_capacity
        h = C.c_void_p()
        capi.check(ctx.L.sb_rtree_build_from_points(ctx.h, ctx._mem(x, y, row_ids), _ptr(x), _ptr(y), _ptr(row_ids), len(row_ids),
                                                    max_node_capacity, C.byref(h)))
        self.h = h
        return self

    @staticmethod
    def leaf_shard(ctx, box4, row_ids, max_node_capacity, slice_size, node_base, x_key_before=-1, x_key_after=-1):
        """multi-GPU build, per rank: leaf entries + node entries (uint8 [., 24]) of a slice-aligned shard of the global x
        order, rows in ascending global input position; x_key_before / x_key_after: the global order's x keys at the
        shard's first rank and just after its last (-1 at the layer's ends) (sb_rtree_build_leaf_shard)."""
        n = len(row_ids)
        nodes = (n + max_node_capacity - 1) // max_node_capacity
        leaf = ctx._empty(box4, (n, 24), np.uint8)
        node = ctx._empty(box4, (nodes, 24), np.uint8)
        capi.check(ctx.L.sb_rtree_build_leaf_shard(ctx.h, ctx._mem(box4, row_ids), _ptr(box4), _ptr(row_ids), n,
                                                   max_node_capacity, slice_size, node_base, int(x_key_before),
                                                   int(x_key_after), _ptr(leaf), _ptr(node)))
        return leaf, node

    @classmethod
    def from_leaf_level(cls, ctx, leaf_entries, n, node_entries, max_node_capacity):
        """multi-GPU build, on the gathering rank: the whole tree from all leaf entries and all leaf-page entries"""
        self = cls.__new__(cls)
        self.ctx = ctx
        self.cap = max_node_capacity
        h = C.c_void_p()
        capi.check(ctx.L.sb_rtree_from_leaf_level(ctx.h, ctx._mem(leaf_entries, node_entries), _ptr(leaf_entries), n,
                                                  _ptr(node_entries), max_node_capacity, C.byref(h)))
        self.h = h
        return self

    @classmethod
    def from_blobs(cls, ctx, data, offsets, row_ids=None, max_node_capacity=128):
        """CREATE INDEX ... USING RTREE over a serialized GEOMETRY column; rows without bounds are not indexed."""
        self = cls.__new__(cls)
        self.ctx = ctx
        self.cap = max_node_capacity
        if not _is_torch(data):
            data = np.ascontiguousarray(data, np.uint8)
            offsets = np.ascontiguousarray(offsets, np.uint64)
            row_ids = None if row_ids is None else np.ascontiguousarray(row_ids, np.int64)
        h, kept = C.c_void_p(), C.c_uint64(0)
        capi.check(ctx.L.sb_rtree_build_from_blobs(ctx.h, ctx._mem(data, offsets), _ptr(data), _ptr(offsets), len(offsets) - 1,
                                                   _ptr(row_ids) if row_ids is not None else None, max_node_capacity,
                                                   C.byref(h), C.byref(kept), None))
        self.h = h
        self.indexed_rows = kept.value
        return self

    def close(self):
        if getattr(self, "h", None):
            self.ctx.L.sb_rtree_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def nlevels(self):
        return self.ctx.L.sb_rtree_num_levels(self.h)

    @property
    def slice(self):
        return self.ctx.L.sb_rtree_slice_size(self.h)

    def level_counts(self):
        return [self.ctx.L.sb_rtree_level_count(self.h, l) for l in range(self.nlevels)]

    def level(self, l):
        """level l as a numpy structured array of 24-byte RTreeEntry records (host copy)."""
        n = self.ctx.L.sb_rtree_level_count(self.h, l)
        out = np.empty(n, ENTRY_DTYPE)
        capi.check(self.ctx.L.sb_rtree_export_level(self.ctx.h, self.h, l, capi.HOST, _ptr(out)))
        return out

    def dump(self, device=False):
        """rtree_index_dump: (level int32[n], bounds float32[n,4], row_id int64[n], -1 for branch entries) in the
        reference pragma's DFS order (rtree_index_pragmas.cpp:137-227)."""
        cnt = C.c_uint64(0)
        capi.check(self.ctx.L.sb_rtree_dump(self.ctx.h, self.h, capi.HOST, None, None, None, 0, C.byref(cnt)))
        n = cnt.value
        if device:
            import torch

            dev = torch.device("cuda", self.ctx.device)
            lv = torch.empty(n, dtype=torch.int32, device=dev)
            box = torch.empty((n, 4), dtype=torch.float32, device=dev)
            rid = torch.empty(n, dtype=torch.int64, device=dev)
            mem = capi.DEVICE
        else:
            lv, box, rid = np.empty(n, np.int32), np.empty((n, 4), np.float32), np.empty(n, np.int64)
            mem = capi.HOST
        capi.check(self.ctx.L.sb_rtree_dump(self.ctx.h, self.h, mem, _ptr(lv), _ptr(box), _ptr(rid), n, C.byref(cnt)))
        return lv, box, rid

    def pages_layout(self, block_size=262136):
        """FixedSizeAllocator geometry + page / buffer counts of the exported tree (sb_rtree_pages_layout)."""
        info = (C.c_uint64 * 8)()
        capi.check(self.ctx.L.sb_rtree_pages_layout(self.h, block_size, info))
        names = ("segment_size", "segments_per_buffer", "bitmask_count", "bitmask_offset", "leaf_pages",
                 "leaf_buffers", "branch_pages", "branch_buffers")
        return dict(zip(names, (int(v) for v in info)))

    def pages(self, block_size=262136, device=False):
        """The tree as the block images of the reference's leaf / branch FixedSizeAllocators, IndexPointers
        included (GetStorageInfo path, rtree_index.cpp:234-259) -> (layout, leaf images, branch images, root entry).
        Images are uint8 arrays (numpy, or CUDA torch tensors with device=True)."""
        info = self.pages_layout(block_size)
        root = np.zeros(1, ENTRY_DTYPE)
        if device:
            import torch

            dev = torch.device("cuda", self.ctx.device)
            leaf = torch.empty(info["leaf_buffers"] * block_size, dtype=torch.uint8, device=dev)
            branch = torch.empty(info["branch_buffers"] * block_size, dtype=torch.uint8, device=dev)
            mem = capi.DEVICE
        else:
            leaf = np.empty(info["leaf_buffers"] * block_size, np.uint8)
            branch = np.empty(info["branch_buffers"] * block_size, np.uint8)
            mem = capi.HOST
        capi.check(self.ctx.L.sb_rtree_export_pages(self.ctx.h, self.h, block_size, mem, _ptr(leaf), _ptr(branch),
                                                    _ptr(root)))
        return info, leaf, branch, root[0]

    def scan(self, query_box4, capacity=None):
        """batched RTreeIndex scan -> (query index [k], row id [k]); per query in the reference's DFS order."""
        q = query_box4 if _is_torch(query_box4) else np.ascontiguousarray(query_box4, np.float32).reshape(-1, 4)
        cap = int(capacity if capacity is not None else 1 << 16)
        while True:
            oq = self.ctx._empty(q, (cap,), np.uint32)
            orow = self.ctx._empty(q, (cap,), np.int64)
            cnt = C.c_uint64(0)
            rc = self.ctx.L.sb_rtree_scan(self.ctx.h, self.h, self.ctx._mem(q), _ptr(q), len(q), _ptr(oq), _ptr(orow),
                                          cap, C.byref(cnt))
            if rc == capi.ERR_CAPACITY:
                cap = int(cnt.value)
                continue
            capi.check(rc)
            return oq[: cnt.value], orow[: cnt.value]


class FlatRTree:
    """The join build side (FlatRTree, node size 32): Hilbert sort + bottom-up pack on the GPU."""

    def __init__(self, ctx, box4=None, node_size=32, _handle=None, _keepalive=None):
        self.ctx = ctx
        self._keepalive = _keepalive
        if _handle is not None:
            self.h = _handle
            return
        if not _is_torch(box4):
            box4 = np.ascontiguousarray(box4, np.float32)
        h = C.c_void_p()
        capi.check(ctx.L.sb_flat_build(ctx.h, ctx._mem(box4), _ptr(box4), len(box4), node_size, C.byref(h)))
        self.h = h

    @classmethod
    def attach(cls, ctx, device_tensor):
        """view over a tree buffer received from another GPU (torch uint8 CUDA tensor)."""
        h = C.c_void_p()
        capi.check(ctx.L.sb_flat_attach(ctx.h, device_tensor.data_ptr(), device_tensor.numel(), C.byref(h)))
        return cls(ctx, _handle=h, _keepalive=device_tensor)

    def close(self):
        if getattr(self, "h", None):
            self.ctx.L.sb_flat_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def item_count(self):
        return self.ctx.L.sb_flat_item_count(self.h)

    @property
    def nodes(self):
        return self.ctx.L.sb_flat_node_count(self.h)

    @property
    def layer_bounds(self):
        n = self.ctx.L.sb_flat_num_layers(self.h)
        p = self.ctx.L.sb_flat_layer_bounds(self.h)
        return np.array([p[i] for i in range(n)], np.uint32)

    def device_buffer(self):
        """(device pointer, bytes) of the packed tree, for NCCL broadcast."""
        p, b = C.c_void_p(), C.c_uint64()
        capi.check(self.ctx.L.sb_flat_device_buffer(self.h, C.byref(p), C.byref(b)))
        return p.value, b.value

    def export(self):
        n = self.nodes
        box = np.empty((n, 4), np.float32)
        idx = np.empty(n, np.uint32)
        capi.check(self.ctx.L.sb_flat_export(self.ctx.h, self.h, capi.HOST, _ptr(box), _ptr(idx)))
        return box, idx

    def probe(self, probe_box4, capacity=None):
        """candidate pairs (probe index, build index) for float probe boxes."""
        q = probe_box4 if _is_torch(probe_box4) else np.ascontiguousarray(probe_box4, np.float32).reshape(-1, 4)
        cap = int(capacity if capacity is not None else max(1024, 2 * len(q)))
        while True:
            op = self.ctx._empty(q, (cap,), np.uint32)
            ob = self.ctx._empty(q, (cap,), np.uint32)
            cnt = C.c_uint64(0)
            rc = self.ctx.L.sb_flat_probe(self.ctx.h, self.h, self.ctx._mem(q), _ptr(q), len(q), _ptr(op), _ptr(ob),
                                          cap, C.byref(cnt))
            if rc == capi.ERR_CAPACITY:
                cap = int(cnt.value)
                continue
            capi.check(rc)
            return op[: cnt.value], ob[: cnt.value]

    def probe_points(self, px, py, build_box2d4=None, predicate="intersects", index_base=0, capacity=None, out=None):
        """probe with points; with build_box2d4 ([n,4] float64 AoS) the exact predicate is fused into the kernel."""
        px, py = self.ctx._f64(px), self.ctx._f64(py)
        pred = capi.PRED_INTERSECTS if predicate == "intersects" else capi.PRED_WITHIN
        cap = int(capacity if capacity is not None else max(1024, 2 * len(px)))
        while True:
            if out is not None:
                op, ob = out
                cap = len(op)
            else:
                op = self.ctx._empty(px, (cap,), np.uint32)
                ob = self.ctx._empty(px, (cap,), np.uint32)
            cnt = C.c_uint64(0)
            arrays = (px, py) if build_box2d4 is None else (px, py, build_box2d4)
            rc = self.ctx.L.sb_flat_probe_points(self.ctx.h, self.h, self.ctx._mem(*arrays), pred, _ptr(px), _ptr(py),
                                                 len(px), index_base, _ptr(build_box2d4), _ptr(op), _ptr(ob), cap,
                                                 C.byref(cnt))
            if rc == capi.ERR_CAPACITY and out is None:
                cap = int(cnt.value)
                continue
            capi.check(rc)
            return op[: cnt.value], ob[: cnt.value]
