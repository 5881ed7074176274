#!/usr/bin/env python
"""Small invocation of every kernel family for compute-sanitizer (memcheck / racecheck): the look-back of the radix
sort, the named barriers and shared-memory tickets of the bulk probes, the STR build, the polygon joins (native and
exact), transcoders, dumps. Results are checked against the oracle, so a race that corrupts data also fails here.

    compute-sanitizer --tool memcheck  python scripts/sanitize_small.py
    compute-sanitizer --tool racecheck python scripts/sanitize_small.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle as O  # noqa: E402
import sb200  # noqa: E402

sb = sb200.load()
ctx = sb.Context(0)


def sort_pairs(p, b):
    a = np.stack([np.asarray(p, np.int64), np.asarray(b, np.int64)], 1) if len(p) else np.zeros((0, 2), np.int64)
    return a[np.lexsort((a[:, 1], a[:, 0]))]


# joins through the three probe paths
n, m = 150_000, 12_000
px, py = O.generate_points(1337, n, (0, 0, 1000, 1000))
u = O.uniform01(42, 2 * m)
cx, cy = u[0::2] * 1000, u[1::2] * 1000
b = [cx - 4, cy - 4, cx + 4, cy + 4]
exp, _ = O.join_points_boxes(px, py, *b, pred=0)
exp = sort_pairs(exp[:, 0], exp[:, 1])
for path in (1, 2, 3):
    ctx.set_option("probe_path", path)
    gp, gb = ctx.spatial_join_points_boxes(px, py, *b, predicate="intersects")
    assert (sort_pairs(gp, gb) == exp).all(), path
ctx.set_option("probe_path", 0)
print("joins ok", len(exp))

# radix sort (onesweep look-back) + STR build + scan + dump + pages
rng = np.random.default_rng(1)
k = rng.integers(0, 2**32, 300_000, dtype=np.uint64).astype(np.uint32)
v = np.arange(len(k), dtype=np.uint32)
sk, sv = ctx.sort_pairs(k.copy(), v.copy())
o = np.argsort(k, kind="stable")
assert (sk == k[o]).all() and (sv == o).all()
k64 = rng.integers(0, 2**63, 200_000, dtype=np.uint64)
sk, sv = ctx.sort_pairs(k64.copy(), np.arange(len(k64), dtype=np.uint32))
o = np.argsort(k64, kind="stable")
assert (sk == k64[o]).all() and (sv == o).all()
box = O.bbox_from_points(px, py)
rows = np.arange(n, dtype=np.int64)
idx = sb.RTreeIndex(ctx, box, rows, 64)
ot = O.StrTree(box, rows, 64)
for l in range(idx.nlevels):
    assert np.array_equal(idx.level(l), ot.level(l))
lv, bb, rid = idx.dump()
olv, obox, orid = ot.dump()
assert np.array_equal(lv, olv) and np.array_equal(rid, orid)
q = O.bbox_from_box2d([100.0, 500.0], [100.0, 500.0], [300.0, 520.0], [300.0, 530.0])
_, r = idx.scan(q)
assert len(r) == len(ot.scan(q[0])) + len(ot.scan(q[1]))
idx.pages()
idx.close()
print("sort / STR ok")

# polygon joins, native + exact, both routes
import test_gpu_exact_predicate as T  # noqa: E402

polys = [[T._ring(rng.uniform(100, 900), rng.uniform(100, 900), int(rng.integers(3, 200)), 20, 60, rng)] for _ in range(40)]
polys.append([T._ring(500, 500, 30, 150, 200, rng), T._ring(500, 500, 8, 20, 60, rng)])
vx, vy, ro, po = T._columns(polys)
qx, qy = T._probe_set(polys, rng, 140_000, 0, 1000)
for route in (1, 2):
    ctx.set_option("pip_path", route)
    a1 = sort_pairs(*ctx.join_points_polygons(qx, qy, vx, vy, ro, po))
    a2 = sort_pairs(*ctx.join_points_polygons(qx, qy, vx, vy, ro, po, geos="intersects"))
    if route == 1:
        r1, r2 = a1, a2
    else:
        assert a1.shape == r1.shape and (a1 == r1).all() and a2.shape == r2.shape and (a2 == r2).all()
ctx.set_option("pip_path", 0)
print("polygon joins ok", len(r1), len(r2))

# blobs: approx / exact extents (all lane-group widths), transcoders, index from blobs, dwithin
data, off = O.point_blobs(px[:50_000], py[:50_000])
assert np.array_equal(ctx.st_extent_approx(data, off)[0].view(np.uint32), O.bbox_from_blobs(data, off)[0].view(np.uint32))
for kk in (4, 20, 70, 300):
    rings = [T._ring(500, 500, kk, 10, 40, rng) for _ in range(500)]
    pvx = np.concatenate([r[0] for r in rings])
    pvy = np.concatenate([r[1] for r in rings])
    voff = np.cumsum([0] + [len(r[0]) for r in rings]).astype(np.uint64)
    pd, poff = O.polygon_blobs(pvx, pvy, voff)
    e1, v1 = ctx.st_extent(pd, poff)
    e2, v2, _ = O.extent_from_blobs(pd, poff)
    assert np.array_equal(e1, e2) and np.array_equal(v1, v2)
ix = sb.RTreeIndex.from_blobs(ctx, data, off, None, 32)
assert ix.indexed_rows == 50_000
ix.close()
ga, gb2, gd = ctx.join_dwithin(px[:60_000], py[:60_000], cx, cy, 3.0)
ea, ed = O.dwithin_brute(px[:60_000], py[:60_000], cx, cy, 3.0)
assert len(ga) == len(ea)
print("blobs / dwithin ok")
ctx.close()
print("SANITIZE_SMALL_OK")
