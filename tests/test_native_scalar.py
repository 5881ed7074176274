"""Native-type scalar functions next to the join path: ST_Intersects(BOX_2D, BOX_2D),
ST_Intersects_Extent(GEOMETRY, GEOMETRY), ST_Distance(POINT_2D, LINESTRING_2D).
CPU: the oracle restatement against hand-checked values and a numpy restatement; GPU: bit-exact against the oracle."""
import numpy as np
import pytest


@pytest.fixture(scope="module")
def ctx(sb):
    c = sb.Context(0)
    yield c
    c.close()


def bits(a):
    return np.ascontiguousarray(a, np.float64).view(np.uint64)


def make_lines(rng, n, max_len=40, degenerate=True):
    lens = rng.integers(0, max_len, n).astype(np.uint64)
    lens[: min(n, 8)] = [0, 1, 2, 2, 3, 5, 1, 2][: min(n, 8)]
    if n > 20:
        lens[10] = 5000  # a long line: many rounds of the warp loop
    off = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    nv = int(off[-1])
    vx = np.cumsum(rng.normal(0, 3, nv)) % 1000
    vy = np.cumsum(rng.normal(0, 3, nv)) % 1000
    if degenerate and nv > 100:
        k = rng.integers(1, nv, nv // 10)
        vx[k] = vx[k - 1] + rng.choice([0.0, 5e-7, -9.9e-7, 1e-6, 2e-6], len(k))  # around the 1e-6 ApproxEqualTo test
        vy[k] = vy[k - 1] + rng.choice([0.0, 5e-7, 1.1e-6], len(k))
        vx[rng.integers(0, nv, 5)] = np.nan
        vy[rng.integers(0, nv, 5)] = np.inf
    lists = np.stack([off[:-1], lens], 1).astype(np.uint64)
    # also exercise offsets that are not cumulative: reverse the row order
    lists = lists[::-1].copy()
    px, py = rng.uniform(0, 1000, n), rng.uniform(0, 1000, n)
    # some points exactly on vertices (distance 0: the reference's early exit)
    for i in range(0, n, 7):
        o, l = int(lists[i, 0]), int(lists[i, 1])
        if l >= 2 and np.isfinite(vx[o + 1]) and np.isfinite(vy[o + 1]):
            px[i], py[i] = vx[o + 1], vy[o + 1]
    return px, py, lists, vx, vy


def test_oracle_point_linestring_known_values(O):
    # 3-4-5 triangle, projection inside / before / after the segment, degenerate segment, single vertex
    vx = np.array([0.0, 10.0, 10.0, 5.0, 5.0 + 5e-7, 7.0])
    vy = np.array([0.0, 0.0, 10.0, 5.0, 5.0, 7.0])
    lists = np.array([[0, 2], [0, 2], [0, 3], [3, 2], [5, 1], [0, 0]], np.uint64)
    px = np.array([5.0, -3.0, 13.0, 8.0, 0.0, 0.0])
    py = np.array([4.0, 4.0, 14.0, 9.0, 0.0, 0.0])
    d = O.distance_point_linestring(px, py, lists, vx, vy)
    assert d[0] == 4.0 and d[1] == 5.0 and d[2] == 5.0 and d[3] == 5.0
    assert d[4] == np.sqrt(np.finfo(np.float64).max) and d[5] == d[4]


def test_oracle_point_linestring_vs_numpy(O):
    rng = np.random.default_rng(3)
    px, py, lists, vx, vy = make_lines(rng, 300, degenerate=False)
    d = O.distance_point_linestring(px, py, lists, vx, vy)
    for i in range(len(px)):
        o, l = int(lists[i, 0]), int(lists[i, 1])
        best = np.finfo(np.float64).max
        for j in range(l - 1):
            x1, y1, x2, y2 = vx[o + j], vy[o + j], vx[o + j + 1], vy[o + j + 1]
            if abs(x1 - x2) < 1e-6 and abs(y1 - y2) < 1e-6:
                cx, cy = x1, y1
            else:
                r = ((px[i] - x1) * (x2 - x1) + (py[i] - y1) * (y2 - y1)) / ((x2 - x1) * (x2 - x1) + (y2 - y1) * (y2 - y1))
                cx, cy = (x1, y1) if r <= 0 else (x2, y2) if r >= 1 else (x1 + r * (x2 - x1), y1 + r * (y2 - y1))
            dd = (px[i] - cx) * (px[i] - cx) + (py[i] - cy) * (py[i] - cy)
            best = min(best, dd)
        assert bits(d[i]) == bits(np.sqrt(best))


def test_oracle_box_intersects(O):
    a = np.array([[0, 0, 1, 1], [0, 0, 1, 1], [0, 0, 1, 1], [0, 0, 1, 1], [np.nan, 0, 1, 1]], np.float64)
    b = np.array([[1, 1, 2, 2], [1.0000001, 0, 2, 1], [-1, -1, 0, 0], [0.2, 0.2, 0.3, 0.3], [5, 5, 6, 6]], np.float64)
    assert list(O.box2d_intersects(a, b)) == [1, 0, 1, 1, 0]  # closed intervals; NaN compares false on min.x only
    assert list(O.box2d_intersects(a, b, valid_a=[1, 1, 0, 1, 1], valid_b=[1, 1, 1, 0, 1])) == [1, 0, 0, 0, 0]


@pytest.mark.gpu
@pytest.mark.parametrize("n", [1, 8, 1000, 200000])
@pytest.mark.parametrize("device", [False, True])
def test_gpu_point_linestring_bit_exact(ctx, O, n, device):
    rng = np.random.default_rng(n)
    px, py, lists, vx, vy = make_lines(rng, n)
    exp = O.distance_point_linestring(px, py, lists, vx, vy)
    if device:
        import torch

        t = [torch.from_numpy(v).cuda() for v in (px, py, lists.view(np.int64), vx, vy)]
        torch.cuda.synchronize()  # the context runs on its own stream: inputs must have landed ...
        got = ctx.st_distance_point_linestring(*t)
        ctx.sync()  # ... and the result must be complete before torch reads it
        got = got.cpu().numpy()
    else:
        got = ctx.st_distance_point_linestring(px, py, lists, vx, vy)
    assert (bits(got) == bits(exp)).all()
    if n >= 1000:
        assert (exp == 0).sum() > 10 and np.isnan(exp).sum() == 0


@pytest.mark.gpu
def test_gpu_point_linestring_rejects_bad_list(ctx, sb):
    with pytest.raises(sb.SpatialB200Error):
        ctx.st_distance_point_linestring([0.0], [0.0], np.array([[1, 5]], np.uint64), [0.0, 1.0], [0.0, 1.0])


@pytest.mark.gpu
def test_gpu_box_intersects_and_extent(ctx, O):
    rng = np.random.default_rng(11)
    n = 100000
    c = rng.uniform(0, 100, (2, n, 2))
    w = rng.uniform(0, 3, (2, n, 2))
    a = np.concatenate([c[0] - w[0], c[0] + w[0]], 1)
    b = np.concatenate([c[1] - w[1], c[1] + w[1]], 1)
    b[:100, :2] = a[:100, 2:]  # touching corners: closed intervals
    b[:100, 2:] = b[:100, :2] + 1
    a[100:110, 0] = np.nan
    exp = O.box2d_intersects(a, b)
    assert 0 < exp.sum() < n and exp[:100].all()
    assert (ctx.st_intersects_box2d(a, b) == exp).all()
    assert (ctx.st_intersects_box2d_columns(a.T.copy(), b.T.copy()) == exp).all()
    ra, rb = (rng.uniform(0, 1, n) < 0.9).astype(np.uint8), (rng.uniform(0, 1, n) < 0.9).astype(np.uint8)
    assert (ctx.st_intersects_box2d(a, b, ra, rb) == O.box2d_intersects(a, b, ra, rb)).all()
    # ST_Intersects_Extent over GEOMETRY blobs: envelope polygons, points, and empty geometries (no extent -> false)
    m = 5000
    da, oa = O.envelope_blobs(a[110:110 + m, 0], a[110:110 + m, 1], a[110:110 + m, 2], a[110:110 + m, 3])
    db, ob = O.point_blobs(b[110:110 + m, 0], b[110:110 + m, 1])
    got = ctx.st_intersects_extent(da, oa, db, ob)
    ea, va, _ = O.extent_from_blobs(da, oa)
    eb, vb, _ = O.extent_from_blobs(db, ob)
    exp2 = O.box2d_intersects(ea, eb, va, vb)
    assert (got == exp2).all() and 0 < exp2.sum() < m
    inside = (b[110:110 + m, 0] >= a[110:110 + m, 0]) & (b[110:110 + m, 0] <= a[110:110 + m, 2]) & \
             (b[110:110 + m, 1] >= a[110:110 + m, 1]) & (b[110:110 + m, 1] <= a[110:110 + m, 3])
    assert (exp2.astype(bool) == inside).all()
