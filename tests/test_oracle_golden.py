"""Pins the CPU oracle on the reference's own known-answer vectors."""
import numpy as np
import pytest

import geomblob

ENV = (450.0, 450.0, 650.0, 650.0)


def _index_scan_count(O, n, dom, cap=128):
    x, y = O.generate_points(1337, n, (0, 0, dom, dom))
    tree = O.StrTree(O.bbox_from_points(x, y), np.arange(n), cap)
    q = O.bbox_from_box2d([ENV[0]], [ENV[1]], [ENV[2]], [ENV[3]])[0]
    cand = tree.scan(q)
    # ST_Within(point, envelope): GEOS semantics = strictly inside (boundary excluded)
    ok = (x[cand] > ENV[0]) & (x[cand] < ENV[2]) & (y[cand] > ENV[1]) & (y[cand] < ENV[3])
    brute = (x > ENV[0]) & (x < ENV[2]) & (y > ENV[1]) & (y < ENV[3])
    assert sorted(cand[ok]) == list(np.nonzero(brute)[0])
    return int(ok.sum()), tree


def test_generate_points_row_50000(O):
    # test/sql/index/rtree_block_reclaim.test_slow:24-27
    x, y = O.generate_points(1337, 100000)
    assert x[49999] == 4732.4060765095055 and y[49999] == 128.3674407750368
    # skip-ahead agrees with the sequential stream
    xs, ys = O.generate_points(1337, 10, skip=49999)
    assert xs[0] == x[49999] and ys[0] == y[49999] and xs[9] == x[50008]


@pytest.mark.parametrize("n,dom,expect", [(10000, 1000, 390), (100000, 10000, 43), (1000000, 10000, 352)])
def test_index_scan_counts(O, n, dom, expect):
    # rtree_basic_points.test:7-26, rtree_persistence_load.test:7-13, rtree_persistence.test:7-13
    got, _ = _index_scan_count(O, n, dom)
    assert got == expect


def test_rtree_dump_levels(O):
    # test/sql/index/rtree_dump.test:4-20 : 100789 rows; per level 7 / 782 / 100000
    _, tree = _index_scan_count(O, 100000, 10000)
    lvl, box, rid = tree.dump()
    assert len(lvl) == 100789
    assert list(np.bincount(lvl)) == [7, 782, 100000]
    assert sorted(rid[lvl == 2]) == list(range(100000))
    assert (rid[lvl < 2] == -1).all()


def test_projection_ids(O):
    # test/sql/index/rtree_projection.test:18-37: ids (1-based row_number) 351, 472, 775
    x, y = O.generate_points(1337, 1000)
    tree = O.StrTree(O.bbox_from_points(x, y), np.arange(1000))
    cand = tree.scan(O.bbox_from_box2d([0.0], [0.0], [500.0], [500.0])[0])
    ok = (x[cand] >= 0) & (x[cand] <= 500) & (y[cand] >= 0) & (y[cand] <= 500)
    ids = sorted(cand[ok] + 1)
    assert ids == [351, 472, 775]
    assert x[350] == 359.812940005213 and y[350] == 406.6655575297773
    assert x[471] == 169.11179292947054 and y[471] == 129.24372218549252
    assert x[774] == 173.61568519845605 and y[774] == 455.52933821454644


def test_single_and_empty(O):
    # rtree_single.test / rtree_empty.test
    t = O.StrTree(O.bbox_from_points(np.array([0.5]), np.array([0.5])), np.arange(1))
    assert t.level_counts() == [1, 1]
    assert list(t.scan(np.array([0, 0, 1, 1], np.float32))) == [0]
    t0 = O.StrTree(np.zeros((0, 4), np.float32), np.zeros(0, np.int64))
    assert t0.nlevels == 0 and len(t0.scan(np.array([0, 0, 1, 1], np.float32))) == 0


def test_join_grid_14641(O):
    # test/sql/join/spatial_join_test.test_slow:8-20
    gx, gy = np.meshgrid(np.arange(501.0), np.arange(501.0), indexing="ij")
    bx, by = np.meshgrid(np.arange(11.0), np.arange(11.0), indexing="ij")
    bx, by = bx.ravel() * 10, by.ravel() * 10
    pairs, cand = O.join_points_boxes(gx.ravel(), gy.ravel(), bx, by, bx + 10, by + 10, pred=0)
    assert len(pairs) == 14641
    # threads must not change the pair set
    pairs1, _ = O.join_points_boxes(gx.ravel(), gy.ravel(), bx, by, bx + 10, by + 10, pred=0, threads=1)
    assert (pairs == pairs1).all()


def test_filter_prune_contains(O):
    # rtree_filter_prune.test:14-20 / rtree_filter_pullup_2.test: ST_Contains(POLYGON((0 0,0 50,50 50,50 0,0 0)), pt)
    # over the 3x3 grid st_point(i,j), i,j in 0..2, j (row_number) 1..9 -> 5,6,8,9 (boundary excluded)
    pts = [(i, j) for i in range(3) for j in range(3)]
    got = [k + 1 for k, (x, y) in enumerate(pts) if O.lib().orc_point_within_box(*map(float, (x, y, 0, 0, 50, 50)))]
    assert got == [5, 6, 8, 9]
    # The native POLYGON_2D winding routine (spatial_functions_scalar.cpp:1938-2051) is NOT GEOS: on an
    # on-edge hit it does `contains = false; break;` which only leaves the coordinate loop (:2011-2016), and
    # the ring is then judged by the winding number accumulated so far (:2025-2033). Hand trace for (1,0):
    # edge (50,50)->(50,0) gives wn=-1, edge (50,0)->(0,0) is the on-edge hit -> in_ring -> TRUE.
    # So ids 4 and 7 (on the y=0 edge, reached after the wn-- edge) are reported inside; ids 1,2,3 (x=0
    # edge, hit first with wn==0) are not. Parity target = this literal behaviour.
    vx = np.array([0.0, 0, 50, 50, 0])
    vy = np.array([0.0, 50, 50, 0, 0])
    ro = np.array([0, 5], np.uint64)
    got2 = [k + 1 for k, (x, y) in enumerate(pts) if O.pip_native(float(x), float(y), vx, vy, ro, 0, 1)]
    assert got2 == [4, 5, 6, 7, 8, 9]


def test_pip_ticket_580(O):
    # test/sql/geos/predicates.test:38-46 (GEOS ticket 580): point strictly inside a 41-vertex polygon
    wkt = """-1.183864 52.951915, -1.183862 52.951903, -1.183890 52.951900, -1.183924 52.951897, -1.183958 52.951894, -1.183954 52.951880, -1.183954 52.951878, -1.183932 52.951841, -1.183904 52.951844, -1.183870 52.951847, -1.183832 52.951852, -1.183824 52.951838, -1.183820 52.951830, -1.183870 52.951819, -1.183886 52.951815, -1.183890 52.951819, -1.183929 52.951810, -1.183909 52.951776, -1.183861 52.951787, -1.183853 52.951788, -1.183842 52.951770, -1.183970 52.951742, -1.183983 52.951763, -1.183963 52.951768, -1.183975 52.951788, -1.183994 52.951785, -1.184009 52.951807, -1.184002 52.951808, -1.184009 52.951835, -1.183990 52.951836, -1.183990 52.951836, -1.183990 52.951838, -1.184001 52.951880, -1.184018 52.951954, -1.184020 52.951956, -1.183998 52.951957, -1.183998 52.951956, -1.183996 52.951948, -1.183970 52.951906, -1.183936 52.951909, -1.183864 52.951915"""
    v = np.array([[float(t) for t in p.split()] for p in wkt.split(",")])
    vx, vy = np.ascontiguousarray(v[:, 0]), np.ascontiguousarray(v[:, 1])
    ro = np.array([0, len(vx)], np.uint64)
    assert O.pip_native(-1.183972, 52.951871, vx, vy, ro, 0, 1) == 1
    assert O.pip_native(-1.19, 52.951871, vx, vy, ro, 0, 1) == 0


def test_extent_15_kinds(O):
    # test/sql/geometry/st_extent.test:6-39
    blobs = [geomblob.blob(g, O) for g, _ in geomblob.EXTENT_CASES]
    data, off = O.concat_blobs(blobs)
    box, valid, status = O.extent_from_blobs(data, off)
    assert (status == 0).all()
    for i, (_, exp) in enumerate(geomblob.EXTENT_CASES):
        if exp is None:
            assert valid[i] == 0
        else:
            assert valid[i] == 1 and tuple(box[i]) == tuple(map(float, exp))
    # ST_Extent_Approx: cached float box for everything non-empty, NULL for empties
    fbox, fvalid, _ = O.bbox_from_blobs(data, off)
    for i, (_, exp) in enumerate(geomblob.EXTENT_CASES):
        assert fvalid[i] == (exp is not None)
        if exp is not None:
            assert tuple(fbox[i]) == tuple(map(float, exp))


def test_hilbert_reference_values(O):
    # SURVEY appendix A.9/A.10 (values produced by the reference code)
    L = O.lib()
    assert L.orc_hilbert_encode(16, 0, 0) == 0
    assert L.orc_hilbert_encode(16, 1, 1) == 2
    assert L.orc_hilbert_encode(16, 65535, 0) == 4294967295
    x, y = O.generate_points(1337, 3)
    data, off = O.point_blobs(x, y)
    h, valid = O.hilbert_geom(data, off)
    assert list(h) == [4244548068, 3198156442, 4156441508] and valid.all()


def test_blob_sizes(O):
    d, off = O.point_blobs(np.array([1.0]), np.array([2.0]))
    assert len(d) == 32
    d, off = O.envelope_blobs([0.0], [0.0], [1.0], [1.0])
    assert len(d) == 120
    box, valid, _ = O.bbox_from_blobs(d, off)
    assert valid[0] and tuple(box[0]) == (0, 0, 1, 1)
    ebox, evalid, _ = O.extent_from_blobs(d, off)
    assert evalid[0] and tuple(ebox[0]) == (0, 0, 1, 1)
    # version bits set -> status -1 (reference throws NotImplementedException)
    d2 = d.copy()
    d2[1] |= 0x40
    _, v, st = O.bbox_from_blobs(d2, off)
    assert v[0] == 0 and st[0] == -1


@pytest.mark.slow
def test_benchmark_10m_3986(O):
    # benchmark/rtree_points_index.benchmark:11-19
    got, tree = _index_scan_count(O, 10_000_000, 10000)
    assert got == 3986
    assert tree.level_counts() == [10_000_000, 78125, 611, 5, 1]
