"""SQL front-end of oracle/_ref/libref_duckdb.so -- TEST INFRASTRUCTURE ONLY.

The reference's own DuckDB-linked code (CREATE INDEX ... USING RTREE, rtree_index_dump, ST_Within(POINT_2D, POLYGON_2D),
SPATIAL_JOIN, ...) driven through SQL; see oracle/ref_duckdb.cpp and the `ref_duckdb` target of oracle/Makefile.
Only tests/ and bench.py's cpu_baseline / --impl reference legs may import this.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_L = None


def available():
    return os.path.exists(os.path.join(_HERE, "_ref", "libref_duckdb.so"))


def lib():
    global _L
    if _L is None:
        path = os.path.join(_HERE, "_ref", "libref_duckdb.so")
        if not os.path.exists(path):
            return None
        L = C.CDLL(path, mode=os.RTLD_LAZY)
        L.rd_last_error.restype = C.c_char_p
        L.rd_open.restype = C.c_void_p
        L.rd_open.argtypes = [C.c_char_p, C.c_int]
        L.rd_close.argtypes = [C.c_void_p]
        L.rd_exec.argtypes = [C.c_void_p, C.c_char_p]
        L.rd_query.restype = C.c_void_p
        L.rd_query.argtypes = [C.c_void_p, C.c_char_p]
        L.rd_rows.restype = C.c_int64
        L.rd_rows.argtypes = [C.c_void_p]
        L.rd_cols.argtypes = [C.c_void_p]
        L.rd_free_result.argtypes = [C.c_void_p]
        L.rd_col_i64.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int64), C.POINTER(C.c_uint8)]
        L.rd_col_f64.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_uint8)]
        L.rd_create_table.argtypes = [C.c_void_p, C.c_char_p, C.c_int, C.POINTER(C.c_char_p), C.POINTER(C.c_int),
                                      C.POINTER(C.c_void_p), C.c_int64]
        _L = L
    return _L


class RefDB:
    """In-memory (or file-backed) DuckDB with the reference's spatial registrations."""

    def __init__(self, path=None, threads=0):
        self.L = lib()
        if self.L is None:
            raise RuntimeError("oracle/_ref/libref_duckdb.so is not built (make -C oracle ref_duckdb)")
        self.h = self.L.rd_open((path or "").encode(), int(threads))
        if not self.h:
            raise RuntimeError("rd_open: " + self.L.rd_last_error().decode())

    def close(self):
        if self.h:
            self.L.rd_close(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def execute(self, sql):
        if self.L.rd_exec(self.h, sql.encode()):
            raise RuntimeError(self.L.rd_last_error().decode())

    def query(self, sql, kinds=None):
        """-> list of numpy columns; kinds: string of 'i' (int64) / 'f' (float64) per column (default: all 'f')"""
        r = self.L.rd_query(self.h, sql.encode())
        if not r:
            raise RuntimeError(self.L.rd_last_error().decode())
        try:
            n, nc = self.L.rd_rows(r), self.L.rd_cols(r)
            kinds = kinds or "f" * nc
            out = []
            for c in range(nc):
                if kinds[c] == "i":
                    a = np.empty(n, np.int64)
                    rc = self.L.rd_col_i64(r, c, a.ctypes.data_as(C.POINTER(C.c_int64)), None)
                else:
                    a = np.empty(n, np.float64)
                    rc = self.L.rd_col_f64(r, c, a.ctypes.data_as(C.POINTER(C.c_double)), None)
                if rc:
                    raise RuntimeError(self.L.rd_last_error().decode())
                out.append(a)
            return out
        finally:
            self.L.rd_free_result(r)

    def scalar(self, sql, kind="i"):
        return self.query(sql, kind)[0][0].item()

    def create_table(self, name, **cols):
        """columns from numpy arrays: float64 -> DOUBLE, integer -> BIGINT"""
        names = list(cols)
        arrs, types = [], []
        for k in names:
            a = np.asarray(cols[k])
            if a.dtype.kind == "f":
                a = np.ascontiguousarray(a, np.float64)
                types.append(0)
            else:
                a = np.ascontiguousarray(a, np.int64)
                types.append(1)
            arrs.append(a)
        n = len(arrs[0])
        cn = (C.c_char_p * len(names))(*[s.encode() for s in names])
        ct = (C.c_int * len(names))(*types)
        cp = (C.c_void_p * len(names))(*[a.ctypes.data for a in arrs])
        if self.L.rd_create_table(self.h, name.encode(), len(names), cn, ct, cp, n):
            raise RuntimeError(self.L.rd_last_error().decode())

    def rtree_dump(self, index):
        """rtree_index_dump(index) -> (level int64[n], box float32[n,4], row_id int64[n]) in the pragma's DFS order"""
        lv, x0, y0, x1, y1, rid = self.query(
            "SELECT level, ST_XMin(bounds), ST_YMin(bounds), ST_XMax(bounds), ST_YMax(bounds), row_id FROM rtree_index_dump('%s')" % index,
            "iffffi")
        return lv, np.stack([x0, y0, x1, y1], 1).astype(np.float32), rid


class RefRTree:
    """A bulk-loaded tree given as FixedSizeAllocator block images, loaded into the REFERENCE's RTree through the
    reference's own allocators (rd_rtree_load); scan / insert / delete run the reference's rtree.cpp on it."""

    def __init__(self, db, cap, leaf_img, leaf_pages, branch_img, branch_pages, root24, block_size=262136, min_cap=None):
        self.L = db.L
        self.db = db
        leaf_img = np.ascontiguousarray(leaf_img, np.uint8)
        branch_img = np.ascontiguousarray(branch_img, np.uint8)
        root24 = np.ascontiguousarray(root24, np.uint8)
        mm = C.c_int(0)
        self.L.rd_rtree_load.restype = C.c_void_p
        self.L.rd_rtree_dump.restype = C.c_uint64
        self.L.rd_rtree_scan.restype = C.c_uint64
        u8p = C.POINTER(C.c_uint8)
        self.h = self.L.rd_rtree_load(C.c_void_p(db.h), C.c_uint32(cap), C.c_uint32(min_cap or max(cap // 2, 1)),
                                      leaf_img.ctypes.data_as(u8p), C.c_uint64(leaf_pages), branch_img.ctypes.data_as(u8p),
                                      C.c_uint64(branch_pages), root24.ctypes.data_as(u8p), C.c_uint64(block_size), C.byref(mm))
        if not self.h:
            raise RuntimeError("rd_rtree_load: " + self.L.rd_last_error().decode())
        self.mismatch = mm.value

    def close(self):
        if self.h:
            self.L.rd_rtree_free(C.c_void_p(self.h))
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def dump(self):
        n = self.L.rd_rtree_dump(C.c_void_p(self.h), None, None, None, C.c_uint64(0))
        lv, box, rid = np.empty(n, np.int32), np.empty((n, 4), np.float32), np.empty(n, np.int64)
        self.L.rd_rtree_dump(C.c_void_p(self.h), lv.ctypes.data_as(C.POINTER(C.c_int32)), box.ctypes.data_as(C.POINTER(C.c_float)),
                             rid.ctypes.data_as(C.POINTER(C.c_int64)), C.c_uint64(n))
        return lv, box, rid

    def scan(self, qbox):
        q = np.ascontiguousarray(qbox, np.float32)
        n = self.L.rd_rtree_scan(C.c_void_p(self.h), q.ctypes.data_as(C.POINTER(C.c_float)), None, C.c_uint64(0))
        out = np.empty(n, np.int64)
        self.L.rd_rtree_scan(C.c_void_p(self.h), q.ctypes.data_as(C.POINTER(C.c_float)), out.ctypes.data_as(C.POINTER(C.c_int64)), C.c_uint64(n))
        return out

    def _mut(self, fn, box4, rid):
        b = np.ascontiguousarray(box4, np.float32).reshape(-1, 4)
        r = np.ascontiguousarray(rid, np.int64)
        if fn(C.c_void_p(self.h), b.ctypes.data_as(C.POINTER(C.c_float)), r.ctypes.data_as(C.POINTER(C.c_int64)), C.c_uint64(len(r))):
            raise RuntimeError(self.L.rd_last_error().decode())

    def insert(self, box4, rid):
        self._mut(self.L.rd_rtree_insert, box4, rid)

    def delete(self, box4, rid):
        self._mut(self.L.rd_rtree_delete, box4, rid)
