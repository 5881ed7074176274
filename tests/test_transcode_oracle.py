"""CPU: the oracle's restatement of the GEOMETRY -> POINT_2D / POLYGON_2D casts (spatial_functions_cast.cpp:247-270,
473-537) against the reference's own expectations for those casts and against round trips through the blob writers."""
import numpy as np

import geomblob

SQ = [(0, 0), (1, 0), (1, 1), (0, 1), (0, 0)]
HOLE = [(0.5, 0.5), (0.5, 0.75), (0.75, 0.75), (0.75, 0.5), (0.5, 0.5)]


def test_polygon_cast_reference_counts(O):
    # test/sql/geometry/st_npoints.test:19-26: ST_NumPoints(...::POLYGON_2D) = 5, 10, 0
    # test/sql/geometry/st_ninteriorrings.test: interior rings 0, 0 (POLYGON EMPTY), 1
    geoms = [("POLYGON", [SQ]), ("POLYGON", [SQ, HOLE]), ("POLYGON", [])]
    data, off = O.concat_blobs([geomblob.blob(g, O) for g in geoms])
    poly_off, ring_off, vx, vy, status = O.polygons_from_blobs(data, off)
    assert (status == 0).all()
    npoints = [int(ring_off[poly_off[i + 1]] - ring_off[poly_off[i]]) for i in range(3)]
    assert npoints == [5, 10, 0]
    interior = [max(int(poly_off[i + 1] - poly_off[i]) - 1, 0) for i in range(3)]
    assert interior == [0, 1, 0]
    assert list(zip(vx[:5], vy[:5])) == [tuple(map(float, v)) for v in SQ]
    assert list(zip(vx[10:15], vy[10:15])) == [tuple(map(float, v)) for v in HOLE]


def test_point_cast_statuses(O):
    # st_npoints.test:6 POINT(0 0)::POINT_2D is legal; POINT EMPTY and non-points throw (cast.cpp:258-264)
    geoms = [("POINT", [(0, 0)]), ("POINT", []), ("LINESTRING", [(0, 0), (1, 1)]), ("POINT", [(-3.5, 1e300)]),
             ("POLYGON", [SQ]), ("MULTIPOINT", [("POINT", [(1, 2)])])]
    data, off = O.concat_blobs([geomblob.blob(g, O) for g in geoms])
    x, y, status = O.points_from_blobs(data, off)
    assert list(status) == [0, 1, -3, 0, -3, -3]
    assert (x[0], y[0]) == (0.0, 0.0) and (x[3], y[3]) == (-3.5, 1e300)
    assert np.isnan(x[[1, 2, 4, 5]]).all() and np.isnan(y[[1, 2, 4, 5]]).all()
    # version bits and truncation
    d2 = data.copy()
    d2[1] |= 0x80
    _, _, s2 = O.points_from_blobs(d2, off)
    assert s2[0] == -1
    off3 = off.copy()
    off3[1] -= 8
    off3 = np.array([0, off3[1]], np.uint64)
    _, _, s3 = O.points_from_blobs(data, off3)
    assert s3[0] == -2


def test_round_trips(O):
    rng = np.random.default_rng(2)
    n = 500
    x, y = rng.uniform(-1e6, 1e6, n), rng.uniform(-1e6, 1e6, n)
    d, off = O.point_blobs(x, y)
    gx, gy, st = O.points_from_blobs(d, off)
    assert (st == 0).all() and (gx == x).all() and (gy == y).all()
    cnt = rng.integers(4, 40, 200)
    voff = np.zeros(201, np.uint64)
    voff[1:] = np.cumsum(cnt)
    vx, vy = rng.uniform(-10, 10, int(voff[-1])), rng.uniform(-10, 10, int(voff[-1]))
    d, off = O.polygon_blobs(vx, vy, voff)
    poly_off, ring_off, gx, gy, st = O.polygons_from_blobs(d, off)
    assert (st == 0).all() and (poly_off == np.arange(201)).all() and (ring_off == voff).all()
    assert (gx == vx).all() and (gy == vy).all()


def test_zm_vertices_drop_extra_ordinates(O):
    # a POLYGON Z blob (flags bit 0): vertex = 24 bytes, bbox = 24 bytes; the cast keeps x, y (get_vertex_xy)
    import struct
    ring = [(0.0, 0.0, 9.0), (2.0, 0.0, 8.0), (2.0, 2.0, 7.0), (0.0, 0.0, 6.0)]
    b = struct.pack("<BBHI", 2, 0x05, 0, 0) + struct.pack("<ffffff", 0, 0, 6, 2, 2, 9)
    b += struct.pack("<II", 2, 1) + struct.pack("<I", len(ring)) + b"\0\0\0\0"
    for v in ring:
        b += struct.pack("<ddd", *v)
    p = struct.pack("<BBHI", 0, 0x03, 0, 0) + struct.pack("<II", 0, 1) + struct.pack("<dddd", 1.5, 2.5, 3.5, 4.5)
    data, off = O.concat_blobs([b, p])
    poly_off, ring_off, vx, vy, st = O.polygons_from_blobs(data, off)
    assert list(st) == [0, -3] and list(poly_off) == [0, 1, 1] and list(ring_off) == [0, 4]
    assert list(vx) == [0.0, 2.0, 2.0, 0.0] and list(vy) == [0.0, 0.0, 2.0, 0.0]
    x, y, st = O.points_from_blobs(data, off)
    assert list(st) == [-3, 0] and (x[1], y[1]) == (1.5, 2.5)
