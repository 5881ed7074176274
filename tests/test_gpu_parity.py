"""GPU parity: the CUDA path (through the C-ABI) against the CPU oracle, bit-exact for boxes / keys / trees /
row-id and pair sets, 1e-12 relative for ST_Distance. Run with `-m gpu` on a B200."""
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
    a = np.ascontiguousarray(a)
    return a.view(np.uint32 if a.dtype.itemsize == 4 else np.uint64)


def sort_pairs(p, b):
    a = np.stack([np.asarray(p, np.int64), np.asarray(b, np.int64)], 1)
    return a[np.lexsort((a[:, 1], a[:, 0]))] if len(a) else a.reshape(0, 2)


SPECIAL = np.array([0.0, -0.0, 1.0, 1 + 2**-30, 1 - 2**-30, 3.4028234663852886e38, 3.4028235e38, 3.5e38, -3.5e38,
                    -3.4028234663852886e38, 1e-46, -1e-46, 1e-40, -1e-40, 16777217.0, -16777217.0, 0.1, -0.1, 1e300,
                    -1e300, np.inf, -np.inf, 5e-324, 1.401298464324817e-45])


# ---- K1 ---------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("n", [0, 1, 2, 3, 1000, 100003])
def test_bbox_points(ctx, O, n):
    rng = np.random.default_rng(n)
    x = rng.uniform(-1e4, 1e4, n)
    y = rng.uniform(-1e4, 1e4, n) * 10.0 ** rng.integers(-30, 30, n)
    k = min(n, len(SPECIAL))
    x[:k] = SPECIAL[:k]
    y[:k] = SPECIAL[::-1][:k]
    got = ctx.bbox_from_points(x, y)
    assert (bits(got) == bits(O.bbox_from_points(x, y))).all()
    # unaligned views take the scalar path
    if n > 3:
        got = ctx.bbox_from_points(x[1:].copy(), y[1:].copy())
        assert (bits(got) == bits(O.bbox_from_points(x[1:], y[1:]))).all()


def test_bbox_box2d(ctx, O):
    rng = np.random.default_rng(3)
    a = [rng.uniform(-1e5, 1e5, 5000) for _ in range(4)]
    a[0][: len(SPECIAL)] = SPECIAL
    a[3][: len(SPECIAL)] = SPECIAL
    assert (bits(ctx.bbox_from_box2d(*a)) == bits(O.bbox_from_box2d(*a))).all()


def _mixed_blobs(O, rng, n=300):
    blobs = [geomblob.blob(g, O) for g, _ in geomblob.EXTENT_CASES]
    for _ in range(n):
        k = int(rng.integers(3, 70))
        ang = np.sort(rng.uniform(0, 2 * np.pi, k))
        cx, cy, r = rng.uniform(-1e3, 1e3), rng.uniform(-1e3, 1e3), rng.uniform(1, 50)
        ring = [(cx + r * np.cos(t), cy + r * np.sin(t)) for t in ang]
        ring.append(ring[0])
        hole = [(cx + 0.1 * r * np.cos(t), cy + 0.1 * r * np.sin(t)) for t in ang[::-1]]
        hole.append(hole[0])
        kind = rng.integers(0, 6)
        if kind == 0:
            g = ("POLYGON", [ring])
        elif kind == 1:
            g = ("POLYGON", [ring, hole])
        elif kind == 2:
            g = ("LINESTRING", ring[: max(2, k // 2)])
        elif kind == 3:
            g = ("MULTIPOLYGON", [("POLYGON", [ring]), ("POLYGON", [hole, ring])])
        elif kind == 4:
            g = ("GEOMETRYCOLLECTION", [("POINT", [ring[0]]), ("GEOMETRYCOLLECTION", [("LINESTRING", ring[:3]), ("POLYGON", [])]), ("MULTIPOINT", [("POINT", [ring[1]]), ("POINT", [])])])
        else:
            g = ("POINT", [ring[0]])
        blobs.append(geomblob.blob(g, O))
    return blobs


def test_extent_and_approx_blobs(ctx, O):
    rng = np.random.default_rng(11)
    data, off = O.concat_blobs(_mixed_blobs(O, rng))
    box, valid = ctx.st_extent_approx(data, off)
    obox, ovalid, _ = O.bbox_from_blobs(data, off)
    assert (valid == ovalid).all() and (bits(box)[valid == 1] == bits(obox)[ovalid == 1]).all()
    ebox, evalid = ctx.st_extent(data, off)
    oebox, oevalid, _ = O.extent_from_blobs(data, off)
    assert (evalid == oevalid).all() and (bits(ebox)[evalid == 1] == bits(oebox)[oevalid == 1]).all()
    # the reference's own expectations (test/sql/geometry/st_extent.test)
    for i, (_, exp) in enumerate(geomblob.EXTENT_CASES):
        assert evalid[i] == (exp is not None)
        if exp is not None:
            assert tuple(ebox[i]) == tuple(map(float, exp))


def test_blob_errors_and_unaligned(ctx, O, sb):
    d, off = O.envelope_blobs([0.0, 5.0], [0.0, 5.0], [1.0, 6.0], [1.0, 6.0])
    d2 = d.copy()
    d2[1] |= 0x40  # unsupported version in row 0
    box, valid, status, rc = ctx.st_extent_approx(d2, off, want_status=True)
    assert rc == sb.capi.ERR_GEOMETRY and status[0] == -1 and status[1] == 0 and valid[1] == 1
    # truncated last blob
    off3 = off.copy()
    off3[2] -= 100
    _, v3, s3, rc3 = ctx.st_extent(d, off3, want_status=True)
    assert rc3 == sb.capi.ERR_GEOMETRY and s3[1] == -2 and v3[0] == 1
    # blobs packed at odd byte offsets (host glue may pack without padding)
    pad = np.concatenate([np.zeros(3, np.uint8), d])
    off_u = off + np.uint64(3)
    b1, v1 = ctx.st_extent(pad, off_u)
    b0, v0 = ctx.st_extent(d, off)
    assert (b1 == b0).all() and (v1 == v0).all()
    a1, w1 = ctx.st_extent_approx(pad, off_u)
    a0, w0 = ctx.st_extent_approx(d, off)
    assert (a1 == a0).all()
    # empty input
    e, ev = ctx.st_extent_approx(np.zeros(0, np.uint8), np.zeros(1, np.uint64))
    assert len(e) == 0


# ---- K2 ---------------------------------------------------------------------------------------------------------
def test_hilbert_variants(ctx, O, sb):
    rng = np.random.default_rng(5)
    n = 20000
    x, y = rng.uniform(-180, 180, n), rng.uniform(-90, 90, n)
    bounds = (-180.0, -90.0, 180.0, 90.0)
    assert (ctx.st_hilbert_xy(x, y, bounds) == O.hilbert_xy(x, y, bounds)).all()
    # out-of-bounds inputs: casts must wrap like the x86 reference
    xo = x * 3
    assert (ctx.st_hilbert_xy(xo, y, bounds) == O.hilbert_xy(xo, y, bounds)).all()
    w = rng.uniform(0, 5, n)
    box = np.stack([x - w, y - w, x + w, y + w], 1)
    assert (ctx.st_hilbert_box2d(box, bounds) == O.hilbert_box2d(box, bounds)).all()
    boxf = box.astype(np.float32)
    assert (ctx.st_hilbert_box2df(boxf, bounds) == O.hilbert_box2df(boxf, bounds)).all()
    data, off = O.concat_blobs(_mixed_blobs(O, rng, 200))
    h, v = ctx.st_hilbert_geom(data, off)
    oh, ov = O.hilbert_geom(data, off)
    assert (v == ov).all() and (h[v == 1] == oh[ov == 1]).all()
    hb, st, rc = ctx.st_hilbert_geom_bounds(data, off, (-2000.0, -2000.0, 2000.0, 2000.0), want_status=True)
    ohb, ost = O.hilbert_geom_bounds(data, off, (-2000.0, -2000.0, 2000.0, 2000.0))
    assert rc == sb.capi.ERR_EMPTY  # the column holds empties: the reference throws
    assert (st == ost).all() and (hb[st == 0] == ohb[ost == 0]).all()
    # golden: first three seed-1337 points
    px, py = O.generate_points(1337, 3)
    d, o = O.point_blobs(px, py)
    h3, _ = ctx.st_hilbert_geom(d, o)
    assert list(h3) == [4244548068, 3198156442, 4156441508]


# ---- K3 ---------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("n", [0, 1, 31, 4095, 4096, 4097, 100000, 1234567])
@pytest.mark.parametrize("dtype", [np.uint32, np.uint64])
def test_sort_pairs(ctx, n, dtype):
    rng = np.random.default_rng(n + 7)
    if dtype == np.uint32:
        keys = rng.integers(0, 2**32, n, dtype=np.uint64).astype(np.uint32)
    else:
        keys = rng.integers(0, 2**63, n, dtype=np.uint64) * np.uint64(2) + rng.integers(0, 2, n, dtype=np.uint64)
    if n > 100:
        keys[: n // 2] = keys[n // 2 : 2 * (n // 2)] & dtype(0xFF00FF)  # heavy duplicates / skewed digits
    vals = np.arange(n, dtype=np.uint32)
    order = np.argsort(keys, kind="stable")
    k, v = ctx.sort_pairs(keys.copy(), vals.copy())
    assert (k == keys[order]).all()
    assert (v == vals[order]).all()  # stability


# ---- K4a --------------------------------------------------------------------------------------------------------
def _tree_equal(idx, ot):
    assert idx.level_counts() == ot.level_counts()
    assert idx.slice == ot.slice or idx.nlevels == 0
    for l in range(idx.nlevels):
        g, e = idx.level(l), ot.level(l)
        assert (g["ptr"] == e["ptr"]).all(), "level %d pointers" % l
        assert (bits(g["box"]) == bits(e["box"])).all(), "level %d boxes" % l


@pytest.mark.parametrize("n,cap", [(1, 128), (2, 128), (127, 128), (128, 128), (129, 128), (5000, 4), (10000, 128),
                                    (100000, 128), (100000, 255), (300001, 50)])
def test_str_build_node_for_node(ctx, sb, O, n, cap):
    rng = np.random.default_rng(n * 7 + cap)
    x, y = O.generate_points(1337, n, (0, 0, 1000, 1000))
    if n > 1000:  # duplicate coordinates => x/y key ties, resolved by the documented tie-break
        x[: n // 3] = np.round(x[: n // 3])
        y[: n // 3] = np.round(y[: n // 3])
    w = rng.uniform(0, 3, n)
    box = O.bbox_from_box2d(x - w, y - w, x + w, y + w)
    rows = rng.permutation(n).astype(np.int64) + 5
    idx = sb.RTreeIndex(ctx, box, rows, cap)
    ot = O.StrTree(box, rows, cap)
    _tree_equal(idx, ot)
    # batched scans, row ids in the reference's DFS order
    nq = 64
    qx, qy, qw = rng.uniform(0, 1000, nq), rng.uniform(0, 1000, nq), rng.uniform(0, 120, nq)
    q = O.bbox_from_box2d(qx - qw, qy - qw, qx + qw, qy + qw)
    q[0] = (-1e9, -1e9, 1e9, 1e9)  # everything
    q[1] = (5000, 5000, 6000, 6000)  # nothing
    oq, orow = idx.scan(q)
    pos = 0
    for i in range(nq):
        exp = ot.scan(q[i])
        assert (oq[pos : pos + len(exp)] == i).all()
        assert (orow[pos : pos + len(exp)] == exp).all()
        pos += len(exp)
    assert pos == len(orow)


@pytest.mark.parametrize("n,cap,spread", [(20000, 128, 1 << 33), (20000, 100, 1 << 31), (3000, 255, 1 << 40)])
def test_str_build_wide_row_ids(ctx, sb, O, n, cap, spread):
    # row ids further apart than 2^31 inside a node: the leaf pack's 64-bit ranking path
    rng = np.random.default_rng(n + cap)
    x, y = O.generate_points(99, n, (0, 0, 1000, 1000))
    box = O.bbox_from_box2d(x - 1, y - 1, x + 1, y + 1)
    rows = rng.permutation(n).astype(np.int64) * spread + 7
    _tree_equal(sb.RTreeIndex(ctx, box, rows, cap), O.StrTree(box, rows, cap))


@pytest.mark.parametrize("case", ["same_x", "same_xy", "two_x", "grid", "y_runs", "sorted", "reverse", "clustered_rows", "wide_keys"])
@pytest.mark.parametrize("path", [0, 1, 2, 3])
def test_str_build_degenerate_inputs(ctx, sb, O, case, path):
    # the x order is a radix select of the slice-boundary keys + a stable partition (str_partition.cu): inputs whose x keys
    # repeat across slice boundaries, whose keys sit in one histogram bucket / in all of them, presorted inputs; both
    # forms of the leaf level (str_path 0: record partition + node partition per slice, 1: bin-id sort + gather,
    # 2: record partition + segmented y sort, 3: node partition by separate kernels)
    rng = np.random.default_rng(11)
    n, cap = 70_001, 16
    x, y = O.generate_points(5, n, (0, 0, 1000, 1000))
    rows = rng.permutation(n).astype(np.int64) + 1
    if case == "same_x":
        x[:] = 123.25
    elif case == "same_xy":
        x[:] = 7.0
        y[:] = -3.0
    elif case == "two_x":
        x = np.where(rng.random(n) < 0.37, 10.0, 20.0)
    elif case == "grid":
        x, y = np.floor(x / 50) * 50, np.floor(y / 25) * 25
    elif case == "y_runs":
        y = np.floor(y / 100) * 100  # runs of ~100 equal y keys per slice: every run spans several nodes of 16
    elif case == "sorted":
        x = np.sort(x)
    elif case == "reverse":
        x = np.sort(x)[::-1].copy()
    elif case == "clustered_rows":
        rows = np.arange(n, dtype=np.int64)
        rows[::97] += 1 << 30  # one far row id in many nodes: the bucketed rank crowds and falls back
    elif case == "wide_keys":
        x = np.ldexp(rng.uniform(1, 2, n), rng.integers(-120, 120, n)) * np.where(rng.random(n) < 0.5, -1.0, 1.0)
    box = O.bbox_from_box2d(x, y, x, y)
    ctx.set_option("str_path", path)
    try:
        idx = sb.RTreeIndex(ctx, box, rows, cap)
    finally:
        ctx.set_option("str_path", 0)
    _tree_equal(idx, O.StrTree(box, rows, cap))


@pytest.mark.parametrize("n,cap,path", [(1, 128, 0), (129, 128, 0), (50_001, 16, 0), (200_000, 128, 0), (50_001, 16, 1)])
def test_str_build_from_points_equals_build_from_boxes(ctx, sb, O, n, cap, path):
    # sb_rtree_build_from_points: the float boxes are computed inside the build; host and device memory
    rng = np.random.default_rng(n)
    x, y = O.generate_points(17, n, (-500, -500, 500, 500))
    rows = rng.permutation(n).astype(np.int64)
    ot = O.StrTree(O.bbox_from_points(x, y), rows, cap)
    ctx.set_option("str_path", path)
    try:
        _tree_equal(sb.RTreeIndex.from_points(ctx, x, y, rows, cap), ot)
        import torch

        dev = torch.device("cuda", 0)
        tx, ty, tr = (torch.from_numpy(v).to(dev) for v in (x, y, rows))
        _tree_equal(sb.RTreeIndex.from_points(ctx, tx, ty, tr, cap), ot)
    finally:
        ctx.set_option("str_path", 0)


@pytest.mark.parametrize("dist", ["uniform", "wide_keys"])
def test_str_build_many_slices(ctx, sb, O, dist):
    # ~350 slices of ~350 nodes each: hundreds of cut keys per select (several chunks of level-1 buckets in the per-slice
    # select, the global-atomic form of the x select when the keys spread over more than 96 level-1 buckets)
    rng = np.random.default_rng(5)
    n, cap = 2_000_003, 16
    if dist == "uniform":
        x, y = O.generate_points(23, n, (0, 0, 1000, 1000))
    else:
        x = np.ldexp(rng.uniform(1, 2, n), rng.integers(-100, 100, n)) * np.where(rng.random(n) < 0.5, -1.0, 1.0)
        y = np.ldexp(rng.uniform(1, 2, n), rng.integers(-100, 100, n)) * np.where(rng.random(n) < 0.5, -1.0, 1.0)
    rows = rng.permutation(n).astype(np.int64)
    box = O.bbox_from_box2d(x, y, x, y)
    _tree_equal(sb.RTreeIndex(ctx, box, rows, cap), O.StrTree(box, rows, cap))


def test_str_empty(ctx, sb):
    idx = sb.RTreeIndex(ctx, np.zeros((0, 4), np.float32), np.zeros(0, np.int64))
    assert idx.nlevels == 0
    oq, orow = idx.scan(np.array([[0, 0, 1, 1]], np.float32))
    assert len(orow) == 0


@pytest.mark.parametrize("n,dom,expect", [(10000, 1000, 390), (100000, 10000, 43), (1000000, 10000, 352)])
def test_index_scan_golden_counts(ctx, sb, O, n, dom, expect):
    # rtree_basic_points.test / rtree_persistence_load.test / rtree_persistence.test, from POINT blobs
    x, y = O.generate_points(1337, n, (0, 0, dom, dom))
    data, off = O.point_blobs(x, y)
    box, valid = ctx.st_extent_approx(data, off)
    assert valid.all()
    idx = sb.RTreeIndex(ctx, box, np.arange(n), 128)
    q = ctx.bbox_from_box2d([450.0], [450.0], [650.0], [650.0])
    _, rows = idx.scan(q)
    ok = (x[rows] > 450) & (x[rows] < 650) & (y[rows] > 450) & (y[rows] < 650)  # ST_Within(pt, envelope)
    assert int(ok.sum()) == expect
    if n == 100000:  # rtree_dump.test: 7 / 782 / 100000
        assert idx.level_counts() == [100000, 782, 7, 1]


# ---- K4b / K5 ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("n", [1, 2, 31, 32, 33, 1024, 1025, 50000, 400000])
def test_flat_build_and_probe(ctx, sb, O, n):
    rng = np.random.default_rng(n)
    cx, cy = rng.uniform(0, 1000, n), rng.uniform(0, 1000, n)
    w = rng.uniform(0, 20, n)
    if n > 100:
        cx[: n // 4] = cx[n // 4 : 2 * (n // 4)]
        cy[: n // 4] = cy[n // 4 : 2 * (n // 4)]
    box = O.bbox_from_box2d(cx - w, cy - w, cx + w, cy + w)
    t = sb.FlatRTree(ctx, box, 32)
    ot = O.FlatTree(box, 32, sort_mode=1)
    assert list(t.layer_bounds) == list(ot.layer_bounds) and t.nodes == ot.nodes
    gb, gi = t.export()
    assert (gi == ot.idx).all()
    assert (bits(gb) == bits(ot.boxes)).all()
    m = 3000
    qx, qy, qw = rng.uniform(-50, 1050, m), rng.uniform(-50, 1050, m), rng.uniform(0, 30, m)
    q = O.bbox_from_box2d(qx - qw, qy - qw, qx + qw, qy + qw)
    q[0] = (-1e9, -1e9, 1e9, 1e9)
    gp, gbd = t.probe(q)
    got = sort_pairs(gp, gbd)
    exp = []
    oref = O.FlatTree(box, 32, sort_mode=0)  # the reference's own (quicksort) tree
    for i in range(m):
        for b in oref.lookup(q[i]):
            exp.append((i, int(b)))
    exp = sort_pairs([e[0] for e in exp], [e[1] for e in exp])
    assert got.shape == exp.shape and (got == exp).all()


def test_flat_empty(ctx, sb):
    t = sb.FlatRTree(ctx, np.zeros((0, 4), np.float32), 32)
    p, b = t.probe(np.array([[0, 0, 1, 1]], np.float32))
    assert len(p) == 0


# ---- joins / K6 -------------------------------------------------------------------------------------------------
def test_join_grid_14641(ctx, O):
    # test/sql/join/spatial_join_test.test_slow:8-20 (boundary-touching pairs count: closed ST_Intersects)
    gx, gy = np.meshgrid(np.arange(501.0), np.arange(501.0), indexing="ij")
    bx, by = np.meshgrid(np.arange(11.0), np.arange(11.0), indexing="ij")
    bx, by = bx.ravel() * 10, by.ravel() * 10
    p, b = ctx.spatial_join_points_boxes(gx.ravel(), gy.ravel(), bx, by, bx + 10, by + 10)
    assert len(p) == 14641
    exp, _ = O.join_points_boxes(gx.ravel(), gy.ravel(), bx, by, bx + 10, by + 10, pred=0)
    assert (sort_pairs(p, b) == sort_pairs(exp[:, 0], exp[:, 1])).all()
    # ST_Within: strict
    p2, b2 = ctx.spatial_join_points_boxes(gx.ravel(), gy.ravel(), bx, by, bx + 10, by + 10, predicate="within")
    exp2, _ = O.join_points_boxes(gx.ravel(), gy.ravel(), bx, by, bx + 10, by + 10, pred=1)
    assert len(p2) == len(exp2) == 121 * 81
    assert (sort_pairs(p2, b2) == sort_pairs(exp2[:, 0], exp2[:, 1])).all()


@pytest.mark.parametrize("n,m", [(0, 10), (10, 0), (1, 1), (200000, 20000), (2000000, 200000)])
def test_join_random_vs_oracle(ctx, O, n, m):
    px, py = O.generate_points(1337, n, (0, 0, 10000, 10000))
    u = O.uniform01(42, 2 * m)
    cx, cy = u[0::2] * 10000, u[1::2] * 10000
    b = [cx - 10, cy - 10, cx + 10, cy + 10]
    # small capacity forces the SB_ERR_CAPACITY retry path
    p, bb = ctx.spatial_join_points_boxes(px, py, *b, capacity=1024)
    exp, _ = O.join_points_boxes(px, py, *b, pred=0)
    assert (sort_pairs(p, bb) == sort_pairs(exp[:, 0], exp[:, 1])).all()


def test_join_device_pointers_torch(ctx, sb, O):
    import torch

    n, m = 300000, 30000
    px, py = O.generate_points(7, n, (0, 0, 10000, 10000))
    u = O.uniform01(9, 2 * m)
    cx, cy = u[0::2] * 10000, u[1::2] * 10000
    b = [cx - 15, cy - 15, cx + 15, cy + 15]
    dev = torch.device("cuda:0")
    tp = [torch.from_numpy(v).to(dev) for v in (px, py)]
    tb = [torch.from_numpy(np.ascontiguousarray(v)).to(dev) for v in b]
    ctx.use_torch_stream()
    try:
        p, bb = ctx.spatial_join_points_boxes(tp[0], tp[1], *tb)
        # separate build / broadcast-style attach / fused probe
        fbox = ctx.bbox_from_box2d(*tb)
        tree = sb.FlatRTree(ctx, fbox, 32)
        ptr, nbytes = tree.device_buffer()
        ctx.sync()
        buf = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        class _Raw:
            def __init__(self, p, n):
                self.__cuda_array_interface__ = {"shape": (n,), "typestr": "|u1", "data": (p, False), "version": 2}

        buf.copy_(torch.as_tensor(_Raw(ptr, nbytes), device=dev))
        t2 = sb.FlatRTree.attach(ctx, buf)
        assert t2.nodes == tree.nodes
        b2d = torch.stack(tb, 1).contiguous()
        p3, b3 = t2.probe_points(tp[0], tp[1], b2d, "intersects", index_base=0)
        torch.cuda.synchronize()
    finally:
        ctx.use_private_stream()
    exp, _ = O.join_points_boxes(px, py, *b, pred=0)
    e = sort_pairs(exp[:, 0], exp[:, 1])
    assert (sort_pairs(p.cpu().numpy().view(np.uint32), bb.cpu().numpy().view(np.uint32)) == e).all()
    assert (sort_pairs(p3.cpu().numpy().view(np.uint32), b3.cpu().numpy().view(np.uint32)) == e).all()


def _star_polygons(O, seed, npoly, k, rmin, rmax, dom):
    u = O.uniform01(seed, npoly * (3 + 2 * k))
    u = u.reshape(npoly, 3 + 2 * k)
    cx, cy = u[:, 0] * dom, u[:, 1] * dom
    vx, vy, off = [], [], [0]
    for i in range(npoly):
        ang = np.sort(u[i, 3 : 3 + k]) * 2 * np.pi
        rad = rmin + u[i, 3 + k : 3 + 2 * k] * (rmax - rmin)
        xs = cx[i] + rad * np.cos(ang)
        ys = cy[i] + rad * np.sin(ang)
        vx.append(np.append(xs, xs[0]))
        vy.append(np.append(ys, ys[0]))
        off.append(off[-1] + k + 1)
    return np.concatenate(vx), np.concatenate(vy), np.array(off, np.uint64)


def test_point_in_polygon_join(ctx, sb, O):
    # C4-shaped: points x star polygons; probe on the polygons' cached float boxes, refine with the native
    # winding routine (ST_Within(POINT_2D, POLYGON_2D))
    npoly, k, n = 300, 40, 100000
    vx, vy, voff = _star_polygons(O, 99, npoly, k, 20, 60, 1000)
    # some exact-boundary probes: polygon vertices themselves and edge midpoints on axis-aligned squares
    sq = np.array([100.0, 100, 200, 200])
    vx = np.concatenate([vx, [sq[0], sq[0], sq[2], sq[2], sq[0]]])
    vy = np.concatenate([vy, [sq[1], sq[3], sq[3], sq[1], sq[1]]])
    voff = np.append(voff, voff[-1] + 5).astype(np.uint64)
    npoly += 1
    poly_off = np.arange(npoly + 1, dtype=np.uint64)  # one ring per polygon
    px, py = O.generate_points(1337, n, (0, 0, 1000, 1000))
    px[:8] = [100, 150, 200, 150, 100, 200, 150, 99.999]
    py[:8] = [150, 100, 150, 200, 100, 200, 150, 150]
    data, off = O.polygon_blobs(vx, vy, voff)
    fbox, valid = ctx.st_extent_approx(data, off)
    tree = sb.FlatRTree(ctx, fbox, 32)
    cp, cb = tree.probe_points(px, py)
    gp, gb = ctx.st_within_point_polygon(px, py, vx, vy, voff, poly_off, cp, cb)
    exp = [(int(a), int(b)) for a, b in zip(cp, cb) if O.pip_native(px[a], py[a], vx, vy, voff, int(b), int(b) + 1)]
    assert len(exp) > 1000
    assert (sort_pairs(gp, gb) == sort_pairs([e[0] for e in exp], [e[1] for e in exp])).all()
    # candidates = exactly the float-box hits
    ot = O.FlatTree(fbox, 32, sort_mode=0)
    qb = O.bbox_from_points(px[:2000], py[:2000])
    expc = [(i, int(b)) for i in range(2000) for b in ot.lookup(qb[i])]
    m = cp < 2000
    assert (sort_pairs(cp[m], cb[m]) == sort_pairs([e[0] for e in expc], [e[1] for e in expc])).all()


def test_dwithin_join_and_distance(ctx, O):
    na, nb, r = 3000, 2500, 1.0
    ax, ay = O.generate_points(1337, na, (0, 0, 100, 100))
    bx, by = O.generate_points(4242, nb, (0, 0, 100, 100))
    bx[:10] = ax[:10] + r  # exactly-at-radius candidates
    by[:10] = ay[:10]
    ga, gb, gd = ctx.join_dwithin(ax, ay, bx, by, r, capacity=256)
    exp, ed = O.dwithin_brute(ax, ay, bx, by, r)
    g = sort_pairs(ga, gb)
    assert (g == sort_pairs(exp[:, 0], exp[:, 1])).all()
    # ST_Distance within 1e-12 relative (north_star tolerance); in practice bit-identical
    order_g = np.lexsort((gb, ga))
    order_e = np.lexsort((exp[:, 1], exp[:, 0]))
    rel = np.abs(gd[order_g] - ed[order_e]) / np.maximum(ed[order_e], 1e-300)
    assert (rel <= 1e-12).all()
    d = ctx.st_distance_points(ax[:2000], ay[:2000], bx[:2000], by[:2000])
    od = np.array([O.lib().orc_distance_pp(*map(float, t)) for t in zip(ax[:2000], ay[:2000], bx[:2000], by[:2000])])
    assert (np.abs(d - od) <= 1e-12 * np.maximum(od, 1e-300)).all()


def test_generate_points_bit_exact(ctx, O):
    # golden: row 50000 of the seed-1337 stream (rtree_block_reclaim.test_slow:24-27) + the whole stream vs oracle
    x, y = ctx.st_generate_points(1337, 100003)
    assert x[49999] == 4732.4060765095055 and y[49999] == 128.3674407750368
    ox, oy = O.generate_points(1337, 100003)
    assert (x == ox).all() and (y == oy).all()
    xs, ys = ctx.st_generate_points(42, 1000, (-10.0, 5.0, 300.0, 77.0), skip=123457)
    ox, oy = O.generate_points(42, 1000, (-10.0, 5.0, 300.0, 77.0), skip=123457)
    assert (xs == ox).all() and (ys == oy).all()


@pytest.mark.parametrize("jt", ["inner", "left", "right", "full"])
def test_outer_join_bookkeeping(ctx, O, jt):
    # LEFT / RIGHT / FULL semantics of PhysicalSpatialJoin (spatial_join_physical.cpp:867-938, 1022-1068): unmatched
    # probe rows / unmatched build rows are emitted once, in row order
    n, m = 150000, 40000
    px, py = O.generate_points(1337, n, (0, 0, 1000, 1000))
    u = O.uniform01(42, 2 * m)
    cx, cy = u[0::2] * 1400 - 200, u[1::2] * 1400 - 200  # some boxes lie outside every point
    b = [cx - 0.8, cy - 0.8, cx + 0.8, cy + 0.8]
    p, bb, up, ub = ctx.spatial_join(px, py, *b, predicate="intersects", join_type=jt)
    exp, _ = O.join_points_boxes(px, py, *b, pred=0)
    assert (sort_pairs(p, bb) == sort_pairs(exp[:, 0], exp[:, 1])).all()
    eup = np.setdiff1d(np.arange(n), exp[:, 0]) if jt in ("left", "full") else np.zeros(0, np.int64)
    eub = np.setdiff1d(np.arange(m), exp[:, 1]) if jt in ("right", "full") else np.zeros(0, np.int64)
    assert len(up) == len(eup) and (np.asarray(up, np.int64) == eup).all()
    assert len(ub) == len(eub) and (np.asarray(ub, np.int64) == eub).all()
    if jt == "full":
        assert len(eup) > 1000 and len(eub) > 1000


def test_build_flags_accumulate_over_probe_batches(ctx, O):
    # the build-side match column accumulates over probe chunks (the tuple's bool column, :878-891)
    n, m = 60000, 5000
    px, py = O.generate_points(7, n, (0, 0, 100, 100))
    u = O.uniform01(9, 2 * m)
    cx, cy = u[0::2] * 100, u[1::2] * 100
    b = [cx - 0.05, cy - 0.05, cx + 0.05, cy + 0.05]
    bm = None
    for lo in range(0, n, 20000):
        p, bb = ctx.spatial_join_points_boxes(px[lo:lo + 20000], py[lo:lo + 20000], *b)
        pm, bm = ctx.join_mark_matched(p, bb, 20000, m, build_matched=bm)
        exp, _ = O.join_points_boxes(px[lo:lo + 20000], py[lo:lo + 20000], *b, pred=0)
        assert (ctx.join_unmatched_rows(pm, row_base=lo) == np.setdiff1d(np.arange(20000), exp[:, 0]) + lo).all()
    exp, _ = O.join_points_boxes(px, py, *b, pred=0)
    assert (ctx.join_unmatched_rows(bm) == np.setdiff1d(np.arange(m), exp[:, 1])).all()
    assert len(ctx.join_unmatched_rows(np.zeros(0, np.uint8))) == 0


def _pip_expected(O, px, py, vx, vy, voff, poly_off, fbox):
    """oracle: candidates = float-box hits (FlatRTree lookup semantics), predicate = the native winding routine"""
    qb = O.bbox_from_points(px, py)
    out = []
    for b in range(len(poly_off) - 1):
        fb = fbox[b]
        if not (fb[0] <= fb[2]):
            continue
        m = ~((qb[:, 0] > fb[2]) | (qb[:, 2] < fb[0]) | (qb[:, 1] > fb[3]) | (qb[:, 3] < fb[1]))
        for a in np.nonzero(m)[0]:
            if O.pip_native(px[a], py[a], vx, vy, voff, int(poly_off[b]), int(poly_off[b + 1])):
                out.append((int(a), b))
    return out


def test_join_points_polygons_both_routes(ctx, sb, O):
    # C4-shaped join through both routes of sb_join_points_polygons (polygon-centric kernel with the shared-memory
    # edge index / FlatRTree candidates + refine) against the oracle: jagged stars, an axis-aligned square with probes
    # on its vertices and edges, a polygon with a hole (literal-loop path), a degenerate and an empty polygon
    npoly, k, n = 40, 60, 140000
    vx, vy, voff = _star_polygons(O, 99, npoly, k, 20, 60, 1000)
    rings = [(vx[int(voff[i]):int(voff[i + 1])], vy[int(voff[i]):int(voff[i + 1])]) for i in range(npoly)]
    poly_rings = [[r] for r in rings]
    sq = (np.array([100.0, 100, 200, 200, 100]), np.array([100.0, 200, 200, 100, 100]))
    poly_rings.append([sq])
    outer = (np.array([600.0, 600, 800, 800, 600]), np.array([600.0, 800, 800, 600, 600]))
    hole = (np.array([650.0, 750, 750, 650, 650]), np.array([650.0, 650, 750, 750, 650]))
    poly_rings.append([outer, hole])
    poly_rings.append([(np.array([300.0, 300, 300]), np.array([300.0, 300, 300]))])  # zero-area
    poly_rings.append([])  # no rings at all
    big = 900  # a ring with more vertices than the index path would ever need
    t = np.linspace(0, 2 * np.pi, big, endpoint=False)
    poly_rings.append([(np.append(500 + 90 * np.cos(t), 590.0), np.append(400 + 70 * np.sin(t), 400.0))])
    fx, fy, roff, poff = [], [], [0], [0]
    for pr in poly_rings:
        for rx, ry in pr:
            fx.append(rx)
            fy.append(ry)
            roff.append(roff[-1] + len(rx))
        poff.append(poff[-1] + len(pr))
    vx, vy = np.concatenate(fx), np.concatenate(fy)
    roff, poff = np.array(roff, np.uint64), np.array(poff, np.uint64)
    px, py = O.generate_points(1337, n, (0, 0, 1000, 1000))
    px[:8] = [100, 150, 200, 150, 100, 200, 150, 99.999]
    py[:8] = [150, 100, 150, 200, 100, 200, 150, 150]
    px[8:12], py[8:12] = [650, 700, 600, 300], [700, 650, 700, 300]  # on the hole / the shell / the degenerate polygon
    px[12], py[13] = np.nan, np.nan
    # probes exactly on star vertices and edge midpoints
    px[20:60], py[20:60] = rings[3][0][:40], rings[3][1][:40]
    px[60:100] = (rings[5][0][:40] + rings[5][0][1:41]) / 2
    py[60:100] = (rings[5][1][:40] + rings[5][1][1:41]) / 2
    fbox = np.zeros((len(poly_rings), 4), np.float32)
    for i, pr in enumerate(poly_rings):
        if not pr:
            fbox[i] = [1, 1, 0, 0]
            continue
        ax, ay = np.concatenate([r[0] for r in pr]), np.concatenate([r[1] for r in pr])
        fbox[i] = O.bbox_from_box2d([ax.min()], [ay.min()], [ax.max()], [ay.max()])[0]
    exp = _pip_expected(O, px, py, vx, vy, roff, poff, fbox)
    e = sort_pairs([t[0] for t in exp], [t[1] for t in exp])
    assert len(e) > 20000
    for route in (2, 1):
        ctx.set_option("pip_path", route)
        try:
            gp, gb = ctx.join_points_polygons(px, py, vx, vy, roff, poff, capacity=1024)
        finally:
            ctx.set_option("pip_path", 0)
        g = sort_pairs(gp, gb)
        assert g.shape == e.shape and (g == e).all(), "route %d differs from the oracle" % route


def test_join_points_polygons_routes_agree_large(ctx, O):
    # larger randomised case: the polygon-centric kernel (edge index, W0 prefix counts) must produce exactly the pair
    # set of FlatRTree candidates + the literal winding loop (route 1 is itself checked against the oracle above)
    npoly, k, n = 300, 200, 3000000
    vx, vy, voff = _star_polygons(O, 7, npoly, k, 5, 90, 2000)
    poff = np.arange(npoly + 1, dtype=np.uint64)
    px, py = O.generate_points(11, n, (0, 0, 2000, 2000))
    res = []
    for route in (1, 2):
        ctx.set_option("pip_path", route)
        try:
            res.append(sort_pairs(*ctx.join_points_polygons(px, py, vx, vy, voff, poff, capacity=4 * n)))
        finally:
            ctx.set_option("pip_path", 0)
    assert len(res[0]) > 1000000
    assert res[0].shape == res[1].shape and (res[0] == res[1]).all()
