"""-m gpu (round 2): the pair-set checksum utility, the dense regime in which `auto` really takes the cell-sorted
kernel (>= 40 Mi probes) checked against the reference FlatRTree's checksum, FlatRTree attach validation, and the
winding routine on hole boundaries (pinned on the reference's ST_Within in tests/test_oracle_vs_refdb.py)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx(sb):
    c = sb.Context(0)
    yield c
    c.close()


def test_pairs_checksum_matches_numpy(ctx, O):
    rng = np.random.default_rng(1)
    n = 1_000_003
    p = rng.integers(0, 2**32, n, dtype=np.uint64).astype(np.uint32)
    b = rng.integers(0, 2**32, n, dtype=np.uint64).astype(np.uint32)
    c, h = ctx.pairs_checksum(p, b)
    sel = p < 0xFFFFFFFF
    assert (c, h) == O.pair_checksum(p[sel], b[sel])
    lo, hi = 1 << 30, 3 << 30
    c, h = ctx.pairs_checksum(p, b, lo, hi)
    sel = (p >= lo) & (p < hi)
    assert (c, h) == O.pair_checksum(p[sel], b[sel])
    assert ctx.pairs_checksum(p[:0], b[:0]) == (0, 0)
    # order independence
    perm = rng.permutation(n)
    assert ctx.pairs_checksum(p[perm], b[perm], lo, hi) == (c, h)


def test_auto_path_dense_regime_vs_reference(ctx, O, R, sb):
    # 64 Mi probes: above the threshold at which `auto` switches to the cell-sorted bulk kernel (the path the headline
    # is timed on). Compared with the reference's own FlatRTree through count + checksum of the whole pair set.
    import torch

    n, m = 64 << 20, 500_000
    dev = torch.device("cuda", 0)
    px, py = ctx.st_generate_points(1337, n, (0.0, 0.0, 10000.0, 10000.0), device=True)
    u = O.uniform01(42, 2 * m)
    cx, cy = u[0::2] * 10000.0, u[1::2] * 10000.0
    b = [np.ascontiguousarray(v) for v in (cx - 10, cy - 10, cx + 10, cy + 10)]
    tb = [torch.from_numpy(v).to(dev) for v in b]
    for pred, name in ((0, "intersects"), (1, "within")):
        cap = int(n * m * 400.0 / 1e8 * 1.2) + (1 << 20)
        gp, gb = ctx.spatial_join_points_boxes(px, py, *tb, predicate=name, capacity=cap)
        gc, gh = ctx.pairs_checksum(gp, gb)
        hx, hy = px.cpu().numpy(), py.cpu().numpy()
        f32p, f64p = C.POINTER(C.c_float), C.POINTER(C.c_double)
        fbox = O.bbox_from_box2d(*b)
        tree = R.ref_flat_build(fbox.ctypes.data_as(f32p), C.c_uint32(m), C.c_uint32(32))
        h = C.c_uint64(0)
        rc = R.ref_flat_join_hash(C.c_void_p(tree), hx.ctypes.data_as(f64p), hy.ctypes.data_as(f64p), C.c_uint64(n),
                                  *[v.ctypes.data_as(f64p) for v in b], C.c_int(pred), C.c_int(0 or 16), C.c_uint32(0), C.byref(h))
        R.ref_flat_free(C.c_void_p(tree))
        assert (gc, gh) == (int(rc), int(h.value)), name
        del gp, gb


def test_flat_attach_rejects_foreign_buffers(ctx, sb, O):
    import torch

    dev = torch.device("cuda", 0)
    box = O.bbox_from_points(*O.generate_points(3, 5000))
    t = sb.FlatRTree(ctx, box, 32)
    ptr, nb = t.device_buffer()
    raw = torch.empty(nb, dtype=torch.uint8, device=dev)
    class _R:
        __cuda_array_interface__ = {"shape": (nb,), "typestr": "|u1", "data": (ptr, False), "version": 2}
    raw.copy_(torch.as_tensor(_R(), device=dev))
    ok = sb.FlatRTree.attach(ctx, raw)
    assert ok.item_count == 5000
    ok.close()
    hdr = raw[:256].view(torch.int32)
    for word, val in ((2, 64), (4, 17), (4, 0), (5, 4999), (3, 1 << 28)):  # node_size, nlayers, nlayers, layer_bounds[0], nodes
        bad = raw.clone()
        bad[:256].view(torch.int32)[word] = val
        with pytest.raises(sb.SpatialB200Error):
            sb.FlatRTree.attach(ctx, bad)
    bad = raw.clone()
    bad[:256].view(torch.int32)[0] = 0
    with pytest.raises(sb.SpatialB200Error):
        sb.FlatRTree.attach(ctx, bad)
    t.close()
    del hdr


def test_winding_on_hole_boundary(ctx, O):
    # a point on a hole's boundary: the on-edge branch sets contains = false and a hole that is not "in_ring" leaves it
    # false (spatial_functions_scalar.cpp:2011-2044; the reference's own ST_Within agrees, tests/test_oracle_vs_refdb.py)
    vx = np.array([10.0, 90, 90, 10, 10, 40, 40, 60, 60, 40])
    vy = np.array([10.0, 10, 90, 90, 10, 40, 60, 60, 40, 40])
    ro = np.array([0, 5, 10], np.uint64)
    po = np.array([0, 2], np.uint64)
    g = np.arange(0, 101, 2.5)
    gx, gy = np.meshgrid(g, g)
    px, py = gx.ravel().copy(), gy.ravel().copy()
    n = len(px)
    pp, pq = np.arange(n, dtype=np.uint32), np.zeros(n, np.uint32)
    keep_p, keep_q = ctx.st_within_point_polygon(px, py, vx, vy, ro, po, pp, pq)
    got = np.zeros(n, np.uint8)
    got[keep_p] = 1
    exp = np.array([O.pip_native(float(a), float(b), vx, vy, ro, 0, 2) for a, b in zip(px, py)], np.uint8)
    assert (got == exp).all()
    # hole boundary points are outside
    on_hole = ((px == 40) | (px == 60)) & (py >= 40) & (py <= 60) | ((py == 40) | (py == 60)) & (px >= 40) & (px <= 60)
    assert exp[on_hole].sum() == 0 and on_hole.sum() > 0
    # the whole-join entry point (polygon with a hole takes the literal loop)
    jp, jq = ctx.join_points_polygons(px, py, vx, vy, ro, po)
    got2 = np.zeros(n, np.uint8)
    got2[jp] = 1
    # the join only evaluates the predicate on float-box candidates (spatial_join_physical.cpp:764-849); the native
    # routine itself reports a few points OUTSIDE the box as inside (e.g. (0, 10): the half-open y rules of upward and
    # downward edges are asymmetric at y == ymin), which ST_Within alone returns and the join never sees
    in_box = (px >= 10) & (px <= 90) & (py >= 10) & (py <= 90)
    assert exp[~in_box].sum() > 0
    exp = exp & in_box.astype(np.uint8)
    bad = np.nonzero(got2 != exp)[0]
    assert len(bad) == 0, [(px[i], py[i], int(got2[i]), int(exp[i])) for i in bad[:12]]


@pytest.mark.parametrize("n,cap", [(1, 128), (2, 4), (129, 128), (5000, 16), (100_000, 128), (300_001, 7)])
def test_rtree_dump_matches_oracle_dfs(ctx, sb, O, n, cap):
    # sb_rtree_dump == the oracle's DFS dump, which is pinned row for row on the reference's rtree_index_dump
    # (tests/test_oracle_vs_refdb.py)
    x, y = O.generate_points(77, n, (0, 0, 1000, 1000))
    box = O.bbox_from_points(x, y)
    rows = np.arange(n, dtype=np.int64) * 3 + 5
    idx = sb.RTreeIndex(ctx, box, rows, cap)
    lv, b, rid = idx.dump()
    olv, obox, orid = O.StrTree(box, rows, cap).dump()
    assert np.array_equal(lv, olv)
    assert np.array_equal(b.view(np.uint32), obox.reshape(-1, 4).view(np.uint32))
    assert np.array_equal(rid, orid)
    idx.close()


def test_rtree_build_from_blobs_skips_rows_without_bounds(ctx, sb, O):
    import geomblob

    n = 20_000
    x, y = O.generate_points(5, n, (0, 0, 1000, 1000))
    raw, _ = O.point_blobs(x, y)
    blobs = [raw[i * 32:(i + 1) * 32].tobytes() for i in range(n)]
    empty = geomblob.blob(("POINT", []), O)
    keep = np.ones(n, bool)
    for i in range(0, n, 97):
        blobs[i] = empty
        keep[i] = False
    data, off = O.concat_blobs(blobs)
    idx = sb.RTreeIndex.from_blobs(ctx, data, off, None, 64)
    assert idx.indexed_rows == int(keep.sum())
    rows = np.nonzero(keep)[0].astype(np.int64)
    ref = sb.RTreeIndex(ctx, O.bbox_from_points(x[keep], y[keep]), rows, 64)
    for l in range(idx.nlevels):
        assert np.array_equal(idx.level(l), ref.level(l))
    idx.close()
    ref.close()


def test_sampled_histogram_overflow_falls_back_to_exact(ctx, O, R, sb):
    # inputs >= 32 Mi rows size the bucket regions of the cell partition from every 16th 4096-row tile. Here every sampled
    # tile lies in one corner and all other tiles in the opposite corner, so the sampled regions are far too small: the
    # partition must notice (k_cp_actual) and repeat with the exact histogram. Checked against the reference FlatRTree.
    import torch

    n, m = 40 << 20, 100_000
    rng = np.random.default_rng(8)
    tile = np.arange(n) // 4096
    sampled = (tile % 16) == 0
    px = np.where(sampled, rng.uniform(0, 500, n), rng.uniform(9000, 10000, n))
    py = np.where(sampled, rng.uniform(0, 500, n), rng.uniform(9000, 10000, n))
    cx = np.concatenate([rng.uniform(0, 500, m // 2), rng.uniform(9000, 10000, m // 2)])
    cy = np.concatenate([rng.uniform(0, 500, m // 2), rng.uniform(9000, 10000, m // 2)])
    b = [np.ascontiguousarray(v) for v in (cx - 2, cy - 2, cx + 2, cy + 2)]
    dev = torch.device("cuda", 0)
    tx, ty = torch.from_numpy(px).to(dev), torch.from_numpy(py).to(dev)
    tb = [torch.from_numpy(v).to(dev) for v in b]
    f32p, f64p = C.POINTER(C.c_float), C.POINTER(C.c_double)
    fbox = O.bbox_from_box2d(*b)
    tree = R.ref_flat_build(fbox.ctypes.data_as(f32p), C.c_uint32(m), C.c_uint32(32))
    h = C.c_uint64(0)
    rc = R.ref_flat_join_hash(C.c_void_p(tree), px.ctypes.data_as(f64p), py.ctypes.data_as(f64p), C.c_uint64(n),
                              *[v.ctypes.data_as(f64p) for v in b], C.c_int(0), C.c_int(16), C.c_uint32(0), C.byref(h))
    R.ref_flat_free(C.c_void_p(tree))
    for opt in (1, 0):
        ctx.set_option("hist_sample", opt)
        try:
            gp, gb = ctx.spatial_join_points_boxes(tx, ty, *tb, predicate="intersects", capacity=int(rc) + 1024)
        finally:
            ctx.set_option("hist_sample", 1)
        assert ctx.pairs_checksum(gp, gb) == (int(rc), int(h.value)), opt
        del gp, gb
