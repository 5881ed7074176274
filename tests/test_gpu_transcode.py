"""GPU parity for the GEOMETRY -> POINT_2D / POLYGON_2D transcoders (sb_points_from_blobs, sb_polygons_from_blobs)
through the C-ABI against the oracle's restatement of the reference casts (spatial_functions_cast.cpp:247-270, 473-537):
bit-exact coordinates and offsets, identical row statuses, host and device memory, and the blob -> columns -> join path."""
import struct

import numpy as np
import pytest

import geomblob

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx(sb):
    c = sb.Context(0)
    yield c
    c.close()


def bits(a):
    return np.ascontiguousarray(a).view(np.uint64)


def _rings(rng, k):
    ang = np.sort(rng.uniform(0, 2 * np.pi, k))
    cx, cy, r = rng.uniform(-1e3, 1e3), rng.uniform(-1e3, 1e3), rng.uniform(1, 50)
    ring = [(cx + r * np.cos(t), cy + r * np.sin(t)) for t in ang]
    ring.append(ring[0])
    return ring, cx, cy, r


def _polygon_column(O, rng, n, bad=True):
    blobs = []
    for i in range(n):
        kind = int(rng.integers(0, 10))
        ring, cx, cy, r = _rings(rng, int(rng.integers(3, 90)))
        if kind <= 4:
            g = ("POLYGON", [ring])
        elif kind <= 6:
            holes = []
            for h in range(int(rng.integers(1, 40))):  # up to 40 rings: more than one warp round of ring sizes
                hr, _, _, _ = _rings(rng, int(rng.integers(3, 8)))
                holes.append(hr)
            g = ("POLYGON", [ring] + holes)
        elif kind == 7:
            g = ("POLYGON", [])
        elif kind == 8 and bad:
            g = ("LINESTRING", ring[:4])
        elif kind == 9 and bad:
            g = ("MULTIPOLYGON", [("POLYGON", [ring])])
        else:
            g = ("POLYGON", [ring, ring[::-1]])
        blobs.append(geomblob.blob(g, O))
    return O.concat_blobs(blobs)


def _check_polys(got, exp):
    for a, b in zip(got[:2], exp[:2]):
        assert (np.asarray(a) == np.asarray(b)).all()
    assert (bits(got[2]) == bits(exp[2])).all() and (bits(got[3]) == bits(exp[3])).all()
    assert (np.asarray(got[4]) == np.asarray(exp[4])).all()


@pytest.mark.parametrize("n", [1, 31, 1024, 1025, 5000])
def test_polygons_from_blobs(ctx, O, sb, n):
    rng = np.random.default_rng(n)
    data, off = _polygon_column(O, rng, n)
    exp = O.polygons_from_blobs(data, off)
    got = ctx.polygons_from_blobs(data, off, want_status=True)
    _check_polys(got, exp)
    assert got[5] == (sb.capi.ERR_GEOMETRY if (exp[4] != 0).any() else sb.capi.OK)
    # device memory
    import torch
    dd, do = torch.from_numpy(data).cuda(), torch.from_numpy(off.view(np.int64)).cuda()
    gd = ctx.polygons_from_blobs(dd, do, want_status=True)
    _check_polys([gd[0].cpu().numpy().view(np.uint64), gd[1].cpu().numpy().view(np.uint64), gd[2].cpu().numpy(),
                  gd[3].cpu().numpy(), gd[4].cpu().numpy()], exp)


def test_polygons_capacity_and_empty(ctx, O, sb):
    import ctypes as C
    rng = np.random.default_rng(9)
    data, off = _polygon_column(O, rng, 200, bad=False)
    exp = O.polygons_from_blobs(data, off)
    nr, nv = C.c_uint64(0), C.c_uint64(0)
    poly_off = np.zeros(201, np.uint64)
    ring_off = np.zeros(4, np.uint64)
    vx, vy = np.zeros(8), np.zeros(8)
    rc = ctx.L.sb_polygons_from_blobs(ctx.h, 0, data.ctypes.data, off.ctypes.data, 200, poly_off.ctypes.data,
                                      ring_off.ctypes.data, vx.ctypes.data, vy.ctypes.data, 3, 8, C.byref(nr), C.byref(nv),
                                      None)
    assert rc == sb.capi.ERR_CAPACITY and nr.value == len(exp[1]) - 1 and nv.value == len(exp[2])
    assert (poly_off == exp[0]).all() and (vx == 0).all()
    # zero rows
    p0 = np.full(1, 7, np.uint64)
    rc = ctx.L.sb_polygons_from_blobs(ctx.h, 0, None, np.zeros(1, np.uint64).ctypes.data, 0, p0.ctypes.data, None, None,
                                      None, 0, 0, C.byref(nr), C.byref(nv), None)
    assert rc == 0 and p0[0] == 0 and nr.value == 0 and nv.value == 0
    # a column of empty polygons only: no rings, no vertices, still OK
    d2, o2 = O.concat_blobs([geomblob.blob(("POLYGON", []), O)] * 5)
    g = ctx.polygons_from_blobs(d2, o2)
    assert (g[0] == 0).all() and list(g[1]) == [0] and len(g[2]) == 0


def test_points_from_blobs(ctx, O, sb):
    rng = np.random.default_rng(4)
    n = 100003
    x, y = rng.uniform(-1e6, 1e6, n), rng.uniform(-1e6, 1e6, n)
    x[:5] = [0.0, -0.0, np.inf, 5e-324, 1e308]
    data, off = O.point_blobs(x, y)
    gx, gy = ctx.points_from_blobs(data, off)
    assert (bits(gx) == bits(x)).all() and (bits(gy) == bits(y)).all()
    import torch
    dx, dy = ctx.points_from_blobs(torch.from_numpy(data).cuda(), torch.from_numpy(off.view(np.int64)).cuda())
    assert (bits(dx.cpu().numpy()) == bits(x)).all() and (bits(dy.cpu().numpy()) == bits(y)).all()
    # packed at an odd byte offset
    pad = np.concatenate([np.zeros(5, np.uint8), data[:3200]])
    ux, uy = ctx.points_from_blobs(pad, off[:101] + np.uint64(5))
    assert (bits(ux) == bits(x[:100])).all() and (bits(uy) == bits(y[:100])).all()
    # the rows on which the reference throws
    pz = struct.pack("<BBHI", 0, 0x03, 0, 0) + struct.pack("<II", 0, 1) + struct.pack("<dddd", 1.5, 2.5, 3.5, 4.5)
    geoms = [("POINT", [(3, 4)]), ("POINT", []), ("LINESTRING", [(0, 0), (1, 1)]), ("POLYGON", [geomblob.SQ])]
    blobs = [geomblob.blob(g, O) for g in geoms] + [pz]
    data, off = O.concat_blobs(blobs)
    ex, ey, es = O.points_from_blobs(data, off)
    gx, gy, gs, rc = ctx.points_from_blobs(data, off, want_status=True)
    assert rc == sb.capi.ERR_GEOMETRY and (gs == es).all() and list(es) == [0, 1, -3, -3, 0]
    ok = es == 0
    assert (gx[ok] == ex[ok]).all() and (gy[ok] == ey[ok]).all() and np.isnan(gx[~ok]).all() and np.isnan(gy[~ok]).all()
    # only empty points: the reference's "Cannot cast empty point GEOMETRY"
    data, off = O.concat_blobs(blobs[:2])
    _, _, gs, rc = ctx.points_from_blobs(data, off, want_status=True)
    assert rc == sb.capi.ERR_EMPTY and list(gs) == [0, 1]
    # malformed: truncated vertex, bad version
    d = np.frombuffer(blobs[0], np.uint8).copy()
    _, _, gs, rc = ctx.points_from_blobs(d, np.array([0, 24], np.uint64), want_status=True)
    assert rc == sb.capi.ERR_GEOMETRY and gs[0] == -2
    d[1] |= 0x40
    _, _, gs, rc = ctx.points_from_blobs(d, np.array([0, 32], np.uint64), want_status=True)
    assert rc == sb.capi.ERR_GEOMETRY and gs[0] == -1
    assert len(ctx.points_from_blobs(np.zeros(0, np.uint8), np.zeros(1, np.uint64))[0]) == 0


def test_blob_columns_feed_the_polygon_join(ctx, O):
    # GEOMETRY columns on both sides -> transcoders -> sb_join_points_polygons == the join on the original columns
    rng = np.random.default_rng(17)
    m = 300
    cnt = rng.integers(5, 60, m)
    voff = np.zeros(m + 1, np.uint64)
    voff[1:] = np.cumsum(cnt)
    vx, vy = np.zeros(int(voff[-1])), np.zeros(int(voff[-1]))
    for i in range(m):
        k = int(cnt[i]) - 1
        ang = np.sort(rng.uniform(0, 2 * np.pi, k))
        cx, cy, r = rng.uniform(0, 1000), rng.uniform(0, 1000), rng.uniform(5, 60)
        b = int(voff[i])
        vx[b:b + k], vy[b:b + k] = cx + r * np.cos(ang), cy + r * np.sin(ang)
        vx[b + k], vy[b + k] = vx[b], vy[b]
    n = 20000
    px, py = rng.uniform(0, 1000, n), rng.uniform(0, 1000, n)
    pd, po = O.point_blobs(px, py)
    gd, go = O.polygon_blobs(vx, vy, voff)
    tx, ty = ctx.points_from_blobs(pd, po)
    poly_off, ring_off, tvx, tvy = ctx.polygons_from_blobs(gd, go)
    assert (ring_off == voff).all() and (poly_off == np.arange(m + 1)).all()
    a = ctx.join_points_polygons(tx, ty, tvx, tvy, ring_off, poly_off, capacity=4 * n)
    b = ctx.join_points_polygons(px, py, vx, vy, voff, np.arange(m + 1, dtype=np.uint64), capacity=4 * n)
    sa = np.lexsort((np.asarray(a[1]), np.asarray(a[0])))
    sb_ = np.lexsort((np.asarray(b[1]), np.asarray(b[0])))
    assert len(a[0]) == len(b[0]) > 0
    assert (np.asarray(a[0])[sa] == np.asarray(b[0])[sb_]).all() and (np.asarray(a[1])[sa] == np.asarray(b[1])[sb_]).all()


def test_extent_from_wkb(ctx, O, sb):
    # ST_Extent(WKB_BLOB): every flavour and every rejected input of tests/test_wkb_oracle.py, host and device memory
    from test_wkb_oracle import wkb_cases
    rng = np.random.default_rng(8)
    blobs = wkb_cases(rng, 1500)
    data, off = O.concat_blobs(blobs)
    ebox, evalid, estatus = O.extent_from_wkb(data, off)
    box, valid, status, rc = ctx.st_extent_wkb(data, off, want_status=True)
    assert rc == sb.capi.ERR_GEOMETRY
    # oracle / reference codes are -(SGL_WKB_READER_ERROR): 1 unsupported type, 2 out of bounds, 3 recursion limit;
    # the library reports them as SB_ROW_WRONG_TYPE (-3), SB_ROW_TRUNCATED (-2), SB_ROW_RECURSION (-4)
    to_row = {0: 0, -1: -3, -2: -2, -3: -4}
    assert (valid == evalid).all() and (status == np.array([to_row[int(s)] for s in estatus])).all()
    assert {-2, -3, -4} <= set(status.tolist())
    assert (bits(box)[evalid == 1] == bits(ebox)[evalid == 1]).all()
    good = estatus == 0
    gd, go = O.concat_blobs([b for b, g in zip(blobs, good) if g])
    import torch
    dbox, dvalid = ctx.st_extent_wkb(torch.from_numpy(gd).cuda(), torch.from_numpy(go.view(np.int64)).cuda())
    assert (dvalid.cpu().numpy() == evalid[good]).all()
    v = evalid[good] == 1
    assert (bits(dbox.cpu().numpy())[v] == bits(ebox[good])[v]).all()
    # a long ring: the lanes stride over the vertices
    t = np.linspace(0, 2 * np.pi, 100001)
    ring = list(zip(1e3 * np.cos(t), 1e3 * np.sin(t)))
    for kw in ({}, {"le": False, "z": True}):
        d1, o1 = O.concat_blobs([geomblob.wkb(("POLYGON", [ring]), **kw)])
        b1, v1 = ctx.st_extent_wkb(d1, o1)
        e1, ev1, _ = O.extent_from_wkb(d1, o1)
        assert v1[0] == 1 and (bits(b1) == bits(e1)).all()
    assert len(ctx.st_extent_wkb(np.zeros(0, np.uint8), np.zeros(1, np.uint64))[0]) == 0
