"""-m gpu: the GEOMETRY-argument point x polygon predicates (GEOS route) -- sb_refine_point_polygon_exact
and sb_join_points_polygons_exact (candidate route and polygon-centric route) against the oracle's exact point-in-area
location, on the reference's own vectors (test/sql/geos/predicates.test, index/rtree_filter_prune.test) and on random
/ on-vertex / on-edge probes."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx(sb):
    c = sb.Context(0)
    yield c
    c.close()


def _ring(cx, cy, k, r0, r1, rng):
    ang = np.sort(rng.uniform(0, 2 * np.pi, k))
    rad = rng.uniform(r0, r1, k)
    vx, vy = cx + rad * np.cos(ang), cy + rad * np.sin(ang)
    return np.append(vx, vx[0]), np.append(vy, vy[0])


def _columns(polys):
    """polys: list of list of (vx, vy) rings -> vx, vy, ring_off, poly_off"""
    vx = np.concatenate([r[0] for p in polys for r in p])
    vy = np.concatenate([r[1] for p in polys for r in p])
    ro = np.cumsum([0] + [len(r[0]) for p in polys for r in p]).astype(np.uint64)
    po = np.cumsum([0] + [len(p) for p in polys]).astype(np.uint64)
    return np.ascontiguousarray(vx), np.ascontiguousarray(vy), ro, po


def test_reference_vectors(ctx, O):
    # rtree_filter_prune.test:14-20: ST_Contains(POLYGON((0 0,0 50,50 50,50 0,0 0)), pt) over the 3x3 grid -> 5,6,8,9
    sq = [[(np.array([0.0, 0, 50, 50, 0]), np.array([0.0, 50, 50, 0, 0]))]]
    vx, vy, ro, po = _columns(sq)
    pts = np.array([(i, j) for i in range(3) for j in range(3)], np.float64)
    px, py = np.ascontiguousarray(pts[:, 0]), np.ascontiguousarray(pts[:, 1])
    pp, pq = np.arange(9, dtype=np.uint32), np.zeros(9, np.uint32)
    kp, _ = ctx.st_within_point_polygon(px, py, vx, vy, ro, po, pp, pq, geos="within")
    assert sorted(int(k) + 1 for k in kp) == [5, 6, 8, 9]
    kp, _ = ctx.st_within_point_polygon(px, py, vx, vy, ro, po, pp, pq, geos="intersects")
    assert sorted(int(k) + 1 for k in kp) == list(range(1, 10))  # boundary points intersect (predicates.test:15-36)
    # GEOS ticket 580 (predicates.test:38-46)
    wkt = """-1.183864 52.951915, -1.183862 52.951903, -1.183890 52.951900, -1.183924 52.951897, -1.183958 52.951894, -1.183954 52.951880, -1.183954 52.951878, -1.183932 52.951841, -1.183904 52.951844, -1.183870 52.951847, -1.183832 52.951852, -1.183824 52.951838, -1.183820 52.951830, -1.183870 52.951819, -1.183886 52.951815, -1.183890 52.951819, -1.183929 52.951810, -1.183909 52.951776, -1.183861 52.951787, -1.183853 52.951788, -1.183842 52.951770, -1.183970 52.951742, -1.183983 52.951763, -1.183963 52.951768, -1.183975 52.951788, -1.183994 52.951785, -1.184009 52.951807, -1.184002 52.951808, -1.184009 52.951835, -1.183990 52.951836, -1.183990 52.951836, -1.183990 52.951838, -1.184001 52.951880, -1.184018 52.951954, -1.184020 52.951956, -1.183998 52.951957, -1.183998 52.951956, -1.183996 52.951948, -1.183970 52.951906, -1.183936 52.951909, -1.183864 52.951915"""
    v = np.array([[float(t) for t in p.split()] for p in wkt.split(",")])
    vx, vy, ro, po = _columns([[(np.ascontiguousarray(v[:, 0]), np.ascontiguousarray(v[:, 1]))]])
    px = np.concatenate([[-1.183972, -1.19], v[:, 0]])
    py = np.concatenate([[52.951871, 52.951871], v[:, 1]])
    n = len(px)
    kp, _ = ctx.st_within_point_polygon(px, py, vx, vy, ro, po, np.arange(n, dtype=np.uint32), np.zeros(n, np.uint32), geos="within")
    assert list(kp) == [0]
    kp, _ = ctx.st_within_point_polygon(px, py, vx, vy, ro, po, np.arange(n, dtype=np.uint32), np.zeros(n, np.uint32), geos="intersects")
    assert sorted(kp) == [0] + list(range(2, n))


def _probe_set(polys, rng, n_random, lo, hi):
    px = [rng.uniform(lo, hi, n_random)]
    py = [rng.uniform(lo, hi, n_random)]
    for p in polys:
        for vx, vy in p:
            for t in (0.0, 0.5, 0.25, 1.0 / 3.0):  # vertices, midpoints, other points of the edges (mostly inexact)
                px.append(vx[:-1] + t * (vx[1:] - vx[:-1]))
                py.append(vy[:-1] + t * (vy[1:] - vy[:-1]))
    return np.concatenate(px), np.concatenate(py)


@pytest.mark.parametrize("pred", ["intersects", "within"])
def test_exact_refine_random_and_on_edge(ctx, O, pred):
    rng = np.random.default_rng(21)
    polys = []
    for k in (3, 5, 17, 64, 500):
        polys.append([_ring(50, 50, k, 10, 45, rng)])
    polys.append([_ring(50, 50, 12, 30, 45, rng), _ring(50, 50, 6, 5, 12, rng)])
    # grid-aligned polygons: probes on vertices / edges are exactly representable, collinear triples abound
    polys.append([(np.array([10.0, 90, 90, 10, 10]), np.array([10.0, 10, 90, 90, 10])),
                  (np.array([40.0, 40, 60, 60, 40]), np.array([40.0, 60, 60, 40, 40]))])
    polys.append([(np.array([0.0, 64, 64, 32, 0, 0]), np.array([0.0, 0, 32, 64, 64, 0]))])
    polys.append([(np.array([5.0, 5, 5, 5]), np.array([5.0, 5, 5, 5]))])  # degenerate: a single repeated vertex
    vx, vy, ro, po = _columns(polys)
    px, py = _probe_set(polys, rng, 1_000_000, 0, 100)
    g = np.arange(0, 101, 1.0)
    gx, gy = np.meshgrid(g, g)
    px = np.concatenate([px, gx.ravel(), [np.nan, 5.0]])
    py = np.concatenate([py, gy.ravel(), [50.0, np.nan]])
    n, m = len(px), len(polys)
    for q in range(m):
        pp, pq = np.arange(n, dtype=np.uint32), np.full(n, q, np.uint32)
        kp, kq = ctx.st_within_point_polygon(px, py, vx, vy, ro, po, pp, pq, geos=pred)
        got = np.zeros(n, np.uint8)
        got[kp] = 1
        exp = O.point_polygon_exact(px, py, pp, pq, vx, vy, ro, po, pred=0 if pred == "intersects" else 1)
        bad = np.nonzero(got != exp)[0]
        assert len(bad) == 0, (q, [(px[i], py[i], int(got[i]), int(exp[i])) for i in bad[:8]])
        assert (kq == q).all()


@pytest.mark.parametrize("pred", ["intersects", "within"])
@pytest.mark.parametrize("route", [1, 2])
def test_exact_join_both_routes(ctx, O, pred, route):
    # candidate route (FlatRTree probe + exact refine) and polygon-centric route (shared-memory edge index, exact mode)
    rng = np.random.default_rng(33)
    polys = []
    for i in range(60):
        cx, cy = rng.uniform(100, 900, 2)
        polys.append([_ring(cx, cy, int(rng.integers(3, 400)), 20, 60, rng)])
    polys.append([_ring(500, 500, 30, 150, 200, rng), _ring(500, 500, 8, 20, 60, rng)])  # with a hole: literal loop
    polys.append([(np.array([200.0, 328, 328, 200, 200]), np.array([200.0, 200, 328, 328, 200]))])
    polys.append([(np.array([0.0, 0, 0]), np.array([0.0, 0, 0]))])
    vx, vy, ro, po = _columns(polys)
    px, py = _probe_set(polys, rng, 600_000, 0, 1000)
    g = np.arange(192.0, 337.0, 1.0)
    gx, gy = np.meshgrid(g, g)
    px, py = np.concatenate([px, gx.ravel()]), np.concatenate([py, gy.ravel()])
    ctx.set_option("pip_path", route)
    try:
        jp, jq = ctx.join_points_polygons(px, py, vx, vy, ro, po, geos=pred)
    finally:
        ctx.set_option("pip_path", 0)
    got = np.stack([np.asarray(jp, np.int64), np.asarray(jq, np.int64)], 1)
    got = got[np.lexsort((got[:, 1], got[:, 0]))]
    # expectation: every (point, polygon) whose float boxes intersect and whose exact location satisfies the predicate;
    # a satisfied predicate implies box intersection, so brute force over a bounding-box prefilter is enough
    exp = []
    for q in range(len(polys)):
        a, b = int(ro[int(po[q])]), int(ro[int(po[q + 1])])
        if a == b:
            continue
        x0, x1, y0, y1 = vx[a:b].min(), vx[a:b].max(), vy[a:b].min(), vy[a:b].max()
        cand = np.nonzero((px >= x0) & (px <= x1) & (py >= y0) & (py <= y1))[0].astype(np.uint32)
        keep = O.point_polygon_exact(px, py, cand, np.full(len(cand), q, np.uint32), vx, vy, ro, po, pred=0 if pred == "intersects" else 1)
        exp.append(np.stack([cand[keep == 1].astype(np.int64), np.full(int(keep.sum()), q, np.int64)], 1))
    exp = np.concatenate(exp)
    exp = exp[np.lexsort((exp[:, 1], exp[:, 0]))]
    assert got.shape == exp.shape and (got == exp).all()
