"""-m gpu: the bulk point probes (bucket_probe.cu: probe_path=2 bucketed v2, probe_path=3 cell-sorted v3, the default)
against the oracle and against the per-probe tree walk (probe_path=1): identical pair sets on uniform, skewed,
degenerate and boundary-heavy inputs."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx(sb):
    c = sb.Context(0)
    yield c
    c.close()


def sort_pairs(p, b):
    a = np.stack([np.asarray(p, np.int64), np.asarray(b, np.int64)], 1) if len(p) else np.zeros((0, 2), np.int64)
    return a[np.lexsort((a[:, 1], a[:, 0]))]


PATHS = (1, 2, 3)
NAMES = {1: "tree walk", 2: "bucketed probe (v2)", 3: "cell-sorted probe (v3)"}


def all_paths(ctx, fn):
    out = []
    for path in PATHS:
        ctx.set_option("probe_path", path)
        try:
            out.append(fn())
        finally:
            ctx.set_option("probe_path", 0)
    return out


def check_join(ctx, O, px, py, b, pred):
    name = "intersects" if pred == 0 else "within"
    res = all_paths(ctx, lambda: ctx.spatial_join_points_boxes(px, py, *b, predicate=name, capacity=4096))
    exp, _ = O.join_points_boxes(px, py, *b, pred=pred)
    e = sort_pairs(exp[:, 0], exp[:, 1]) if len(exp) else np.zeros((0, 2), np.int64)
    for path, (p, bb) in zip(PATHS, res):
        g = sort_pairs(p, bb)
        assert g.shape == e.shape and (g == e).all(), NAMES[path] + " differs from the oracle"
    return len(e)


@pytest.mark.parametrize("pred", [0, 1])
def test_grid_boundaries(ctx, O, pred):
    # spatial_join_test.test_slow:8-20: every box edge carries probe points (closed vs strict semantics)
    gx, gy = np.meshgrid(np.arange(501.0), np.arange(501.0), indexing="ij")
    bx, by = np.meshgrid(np.arange(11.0), np.arange(11.0), indexing="ij")
    bx, by = bx.ravel() * 10, by.ravel() * 10
    n = check_join(ctx, O, gx.ravel(), gy.ravel(), [bx, by, bx + 10, by + 10], pred)
    assert n == (14641 if pred == 0 else 121 * 81)


@pytest.mark.parametrize("n,m,side", [(1, 1, 10.0), (1000, 3, 10.0), (300000, 30000, 20.0), (2000000, 200000, 20.0),
                                      (500000, 2000, 900.0)])
@pytest.mark.parametrize("pred", [0, 1])
def test_uniform(ctx, O, n, m, side, pred):
    px, py = O.generate_points(1337, n, (0, 0, 10000, 10000))
    u = O.uniform01(42, 2 * m)
    cx, cy = u[0::2] * 10000, u[1::2] * 10000
    h = side / 2
    check_join(ctx, O, px, py, [cx - h, cy - h, cx + h, cy + h], pred)


def test_skewed_cluster_and_outliers(ctx, O):
    # 60k boxes packed into one corner (one bucket holds > BPP_CMAX candidates -> chunked directory), two far
    # outliers stretch the tree box, most points fall into the same bucket, some fall outside the tree box
    rng = np.random.default_rng(5)
    m = 60000
    cx, cy = rng.uniform(0, 30, m), rng.uniform(0, 30, m)
    w = rng.uniform(0.01, 0.5, m)
    cx[:2], cy[:2] = [5000.0, -4000.0], [9000.0, -100.0]
    n = 400000
    px, py = rng.uniform(-5, 35, n), rng.uniform(-5, 35, n)
    px[:1000] = rng.uniform(-6000, 6000, 1000)
    py[:1000] = rng.uniform(-6000, 12000, 1000)
    check_join(ctx, O, px, py, [cx - w, cy - w, cx + w, cy + w], 0)


def test_wide_boxes_and_dense_hits(ctx, O):
    # boxes covering most of the domain (the "wide" list) + 300 identical boxes: every round overflows the
    # shared-memory staging buffer and a point's list is longer than 64 entries
    rng = np.random.default_rng(11)
    m = 400
    cx, cy = rng.uniform(400, 600, m), rng.uniform(400, 600, m)
    w = rng.uniform(300, 500, m)
    cx[:300], cy[:300], w[:300] = 500.0, 500.0, 100.0
    n = 150000
    px, py = rng.uniform(0, 1000, n), rng.uniform(0, 1000, n)
    px[:64] = 400.0  # on the shared edge of the identical boxes
    check_join(ctx, O, px, py, [cx - w, cy - w, cx + w, cy + w], 0)
    check_join(ctx, O, px[:50000], py[:50000], [cx - w, cy - w, cx + w, cy + w], 1)


def test_degenerate_extents(ctx, O):
    rng = np.random.default_rng(3)
    n = 200000
    px, py = rng.uniform(-1, 1, n), rng.uniform(-1, 1, n)
    px[:100], py[:100] = 0.25, 0.5
    # zero-area tree box (a single degenerate box): the quantiser scale is 0 -> every test is a tie
    z = np.array([0.25])
    check_join(ctx, O, px, py, [z, z * 2, z, z * 2], 0)
    # vertical line of boxes: zero width in x
    k = 500
    yy = np.linspace(-1, 1, k)
    check_join(ctx, O, px, py, [np.full(k, 0.25), yy - 0.01, np.full(k, 0.25), yy + 0.01], 0)


def test_quantiser_ties_are_exact(ctx, O):
    # coordinates that differ from box edges by one ulp on either side, at a large offset
    base = 1.0e6
    ex = np.array([base, base + 100.0])
    bx0, bx1 = np.repeat(ex[:1], 50), np.repeat(ex[1:], 50)
    by0 = np.arange(50.0) * 3
    by1 = by0 + 2.0
    pts_x, pts_y = [], []
    for v in (ex[0], ex[1]):
        for d in (np.nextafter(v, -np.inf), v, np.nextafter(v, np.inf)):
            pts_x.append(np.full(50, d))
            pts_y.append(by0 + 1.0)
    for j in range(50):
        for d in (np.nextafter(by0[j], -np.inf), by0[j], np.nextafter(by1[j], np.inf), by1[j]):
            pts_x.append(np.array([base + 50.0]))
            pts_y.append(np.array([d]))
    px = np.concatenate(pts_x + [np.random.default_rng(1).uniform(base - 10, base + 110, 150000)])
    py = np.concatenate(pts_y + [np.random.default_rng(2).uniform(-5, 155, 150000)])
    for pred in (0, 1):
        check_join(ctx, O, px, py, [bx0, by0, bx1, by1], pred)


def test_candidate_mode_matches_tree_lookup(ctx, sb, O):
    # candidate semantics (float boxes): bucketed == tree walk == the reference FlatRTree lookup, incl. NaN probes
    rng = np.random.default_rng(17)
    m = 20000
    cx, cy = rng.uniform(0, 1000, m), rng.uniform(0, 1000, m)
    w = rng.uniform(0, 8, m)
    box = O.bbox_from_box2d(cx - w, cy - w, cx + w, cy + w)
    tree = sb.FlatRTree(ctx, box, 32)
    n = 200000
    px, py = rng.uniform(-20, 1020, n), rng.uniform(-20, 1020, n)
    # points exactly on float box edges and one ulp (of the float) outside
    fb = box[:500]
    px[:500] = fb[:, 0].astype(np.float64)
    py[:500] = (fb[:, 1].astype(np.float64) + fb[:, 3].astype(np.float64)) / 2
    px[500:1000] = np.nextafter(fb[:, 0], np.float32(-np.inf)).astype(np.float64)
    py[500:1000] = py[:500]
    px[1000:1500] = np.nextafter(px[:500], -np.inf)
    py[1000:1500] = py[:500]
    px[1500], py[1501] = np.nan, np.nan
    (p1, b1), (p2, b2), (p3, b3) = all_paths(ctx, lambda: tree.probe_points(px, py))
    g1, g2, g3 = sort_pairs(p1, b1), sort_pairs(p2, b2), sort_pairs(p3, b3)
    assert g1.shape == g2.shape and (g1 == g2).all()
    assert g1.shape == g3.shape and (g1 == g3).all()
    ot = O.FlatTree(box, 32, sort_mode=0)
    k = 3000
    qb = O.bbox_from_points(px[:k], py[:k])
    exp = [(i, int(b)) for i in range(k) if not (np.isnan(px[i]) or np.isnan(py[i])) for b in ot.lookup(qb[i])]
    sel = (g2[:, 0] < k) & ~np.isin(g2[:, 0], [1500, 1501])
    assert (g2[sel] == sort_pairs([e[0] for e in exp], [e[1] for e in exp])).all()


@pytest.mark.parametrize("path", [2, 3])
def test_index_base_and_capacity(ctx, sb, O, path):
    import ctypes as C

    rng = np.random.default_rng(23)
    m, n = 5000, 300000
    cx, cy = rng.uniform(0, 1000, m), rng.uniform(0, 1000, m)
    b = [cx - 5, cy - 5, cx + 5, cy + 5]
    box = O.bbox_from_box2d(*b)
    tree = sb.FlatRTree(ctx, box, 32)
    px, py = rng.uniform(0, 1000, n), rng.uniform(0, 1000, n)
    b2d = np.ascontiguousarray(np.stack(b, 1))
    ctx.set_option("probe_path", path)
    try:
        p, bb = tree.probe_points(px, py, b2d, "intersects", index_base=1000000)
        # too-small capacity: SB_ERR_CAPACITY with the needed count, nothing written past the end
        op = np.full(1000 + 16, 0xFFFFFFFF, np.uint32)
        ob = np.full(1000 + 16, 0xFFFFFFFF, np.uint32)
        cnt = C.c_uint64(0)
        rc = ctx.L.sb_flat_probe_points(ctx.h, tree.h, sb.capi.HOST, 0, px.ctypes.data, py.ctypes.data, n, 0,
                                        b2d.ctypes.data, op.ctypes.data, ob.ctypes.data, 1000, C.byref(cnt))
    finally:
        ctx.set_option("probe_path", 0)
    exp, _ = O.join_points_boxes(px, py, *b, pred=0)
    assert (sort_pairs(p - 1000000, bb) == sort_pairs(exp[:, 0], exp[:, 1])).all()
    assert rc == sb.capi.ERR_CAPACITY and cnt.value == len(exp)
    assert (op[1000:] == 0xFFFFFFFF).all() and (ob[1000:] == 0xFFFFFFFF).all()


def test_host_pipeline_chunks(ctx, O):
    # host-pointer join streamed in chunks (H2D / probe / D2H overlapped); small chunks force many iterations, the
    # dense cluster overflows an output slot (12 pairs per chunk row) and splits the chunk
    rng = np.random.default_rng(31)
    m = 3000
    cx, cy = rng.uniform(0, 1000, m), rng.uniform(0, 1000, m)
    w = rng.uniform(1, 12, m)
    cx[:200], cy[:200], w[:200] = 300.0, 300.0, 40.0
    n = 150001
    px, py = rng.uniform(0, 1000, n), rng.uniform(0, 1000, n)
    px[20000:24000] = rng.uniform(280, 320, 4000)
    py[20000:24000] = rng.uniform(280, 320, 4000)
    b = [cx - w, cy - w, cx + w, cy + w]
    exp, _ = O.join_points_boxes(px, py, *b, pred=0)
    e = sort_pairs(exp[:, 0], exp[:, 1])
    ctx.set_option("pipe_chunk", 4096)
    try:
        for path in PATHS:
            ctx.set_option("probe_path", path)
            p, bb = ctx.spatial_join_points_boxes(px, py, *b, capacity=len(e) + 5)
            g = sort_pairs(p, bb)
            assert g.shape == e.shape and (g == e).all()
        # capacity too small: needed count comes back
        import ctypes as C
        op = np.zeros(1000, np.uint32)
        cnt = C.c_uint64(0)
        args = [v.ctypes.data for v in (px, py)] + [n] + [np.ascontiguousarray(v).ctypes.data for v in b]
        bb4 = [np.ascontiguousarray(v) for v in b]
        rc = ctx.L.sb_join_points_boxes(ctx.h, 0, 0, px.ctypes.data, py.ctypes.data, n, *[v.ctypes.data for v in bb4], m,
                                        op.ctypes.data, op.ctypes.data, 1000, C.byref(cnt))
        assert rc == -3 and cnt.value == len(e)
    finally:
        ctx.set_option("probe_path", 0)
        ctx.set_option("pipe_chunk", 8 << 20)


def test_heavy_sub_cell(ctx, O):
    # 400k probes inside ONE sub-cell of the quantiser grid (70k of them identical): a v3 warp owns a sub-cell, so this
    # is its worst-balanced input; boxes around the cluster are partly definite covers, partly edge cases
    rng = np.random.default_rng(41)
    n = 400000
    px = np.concatenate([rng.uniform(500.0, 500.5, n - 70000), np.full(70000, 500.25)])
    py = np.concatenate([rng.uniform(500.0, 500.5, n - 70000), np.full(70000, 500.25)])
    m = 3000
    cx, cy = rng.uniform(0, 1000, m), rng.uniform(0, 1000, m)
    w = rng.uniform(1, 30, m)
    cx[:40], cy[:40] = rng.uniform(499.5, 501, 40), rng.uniform(499.5, 501, 40)
    w[:40] = rng.uniform(0.05, 3, 40)
    cx[40], cy[40], w[40] = 500.25, 500.25, 0.0  # degenerate box exactly on the identical points
    check_join(ctx, O, px, py, [cx - w, cy - w, cx + w, cy + w], 0)
    check_join(ctx, O, px, py, [cx - w, cy - w, cx + w, cy + w], 1)
