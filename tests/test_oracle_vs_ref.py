"""Pins oracle.c against the REFERENCE'S OWN code compiled into oracle/_ref/libref.so
(sgl Hilbert, MathUtil rounding, pcg32, TryGetCachedBounds, FlatRTree build + Lookup)."""
import ctypes as C

import numpy as np
import pytest


def _special_doubles(rng):
    v = rng.uniform(-1e4, 1e4, 2000)
    extra = np.array([0.0, -0.0, 1.0, 1 + 2**-30, 1 - 2**-30, 3.4028234663852886e38, 3.4028235e38, 3.5e38, -3.5e38,
                      -3.4028234663852886e38, 1e-46, -1e-46, 1e-40, 16777217.0, -16777217.0, 0.1, -0.1, 1e300, -1e300])
    return np.concatenate([v, extra, rng.uniform(-1, 1, 500) * 10.0 ** rng.integers(-40, 38, 500)])


def test_rounding_matches_reference(O, R):
    rng = np.random.default_rng(1)
    L = O.lib()
    for d in _special_doubles(rng):
        a, b = L.orc_double_to_float_down(d), R.ref_double_to_float_down(d)
        assert a == b, d
        a, b = L.orc_double_to_float_up(d), R.ref_double_to_float_up(d)
        assert a == b, d


def test_hilbert_matches_reference(O, R):
    rng = np.random.default_rng(2)
    L = O.lib()
    xs = rng.integers(0, 2**16, 3000)
    ys = rng.integers(0, 2**16, 3000)
    for x, y in zip(xs, ys):
        assert L.orc_hilbert_encode(16, int(x), int(y)) == R.ref_hilbert_encode(16, int(x), int(y))
    # 32-bit inputs into the 16-bit routine (ST_Hilbert(geom))
    xs = rng.integers(0, 2**32, 3000)
    ys = rng.integers(0, 2**32, 3000)
    for x, y in zip(xs, ys):
        assert L.orc_hilbert_encode(16, int(x), int(y)) == R.ref_hilbert_encode(16, int(x), int(y))
    for n in (1, 5, 8, 15):
        for x, y in zip(xs[:200], ys[:200]):
            m = (1 << n) - 1
            assert L.orc_hilbert_encode(n, int(x) & m, int(y) & m) == R.ref_hilbert_encode(n, int(x) & m, int(y) & m)
    for f in list(rng.uniform(-1e5, 1e5, 500).astype(np.float32)) + [np.float32(0), np.float32(-0.0), np.float32("nan"), np.float32("inf"), np.float32("-inf")]:
        assert L.orc_hilbert_f32_to_u32(C.c_float(f)) == R.ref_hilbert_f32_to_u32(C.c_float(f))


def test_float_to_u32_casts_match_reference(O, R):
    # oracle's explicit cast rule == what the reference's static_cast<uint32_t> compiles to on x86-64
    bounds = np.array([0.0, 0.0, 1.0, 1.0])
    vals = np.array([0.0, 0.5, 65535.0, 65535.9, 65536.0, 4294967295.0, 4294967296.0, 4294967297.5, 1e12, -1.0, -1.5,
                     -65536.0, 9.3e18, 1e30, -1e30, np.nan, np.inf, -np.inf])
    for v in vals:
        # hilbert_xy with bounds (0,0,1,1): hx = cast((v - 0) * 65535)
        got = O.hilbert_xy(np.array([v / 65535.0]), np.array([0.0]), bounds)[0]
        scaled = (v / 65535.0 - 0.0) * 65535.0
        exp = R.ref_hilbert_encode(16, R.ref_cast_f64_u32(scaled), 0)
        assert got == exp, v
    fb = np.array([0, 0, 1, 1], np.float32)
    for v in vals.astype(np.float32):
        box = np.array([[v / np.float32(65535), 0, v / np.float32(65535), 0]], np.float32)
        got = O.hilbert_box2df(box, fb)[0]
        x = box[0, 0] + (box[0, 2] - box[0, 0]) / np.float32(2)
        scaled = np.float32((x - np.float32(0)) * (np.float32(65535) / np.float32(1)))
        exp = R.ref_hilbert_encode(16, R.ref_cast_f32_u32(C.c_float(scaled)), 0)
        assert got == exp, v


def test_generate_points_matches_reference(O, R):
    n = 5000
    x, y = O.generate_points(42, n, (-10, 5, 300, 77))
    rx = np.empty(n)
    ry = np.empty(n)
    f64p = C.POINTER(C.c_double)
    R.ref_generate_points(C.c_uint64(42), C.c_uint64(n), C.c_double(-10), C.c_double(5), C.c_double(300), C.c_double(77),
                          rx.ctypes.data_as(f64p), ry.ctypes.data_as(f64p))
    assert (x == rx).all() and (y == ry).all()


def test_cached_bounds_matches_reference(O, R):
    import geomblob

    rng = np.random.default_rng(5)
    x, y = rng.uniform(-1e3, 1e3, 50), rng.uniform(-1e3, 1e3, 50)
    blobs = [geomblob.blob(g, O) for g, _ in geomblob.EXTENT_CASES]
    blobs += [geomblob.blob(("POINT", [(a, b)]), O) for a, b in zip(x, y)]
    blobs += [geomblob.blob(("LINESTRING", [(a, b), (b, a), (a + 1, b - 3)]), O) for a, b in zip(x, y)]
    u8p = C.POINTER(C.c_uint8)
    f32p = C.POINTER(C.c_float)
    for b in blobs:
        arr = np.frombuffer(b, np.uint8).copy()
        o = np.zeros(4, np.float32)
        r = np.zeros(4, np.float32)
        ro = O.lib().orc_cached_bounds(arr.ctypes.data_as(u8p), C.c_uint64(len(arr)), o.ctypes.data_as(f32p))
        rr = R.ref_cached_bounds(arr.ctypes.data_as(u8p), C.c_uint32(len(arr)), r.ctypes.data_as(f32p))
        assert ro == rr
        if ro == 1:
            assert (o == r).all()


def _ref_flat(R, box, node_size=32):
    f32p = C.POINTER(C.c_float)
    u32p = C.POINTER(C.c_uint32)
    box = np.ascontiguousarray(box, np.float32)
    h = R.ref_flat_build(box.ctypes.data_as(f32p), C.c_uint32(len(box)), C.c_uint32(node_size))
    nodes = R.ref_flat_nodes(h)
    nl = R.ref_flat_nlayers(h)
    b = np.zeros((nodes, 4), np.float32)
    idx = np.zeros(nodes, np.uint32)
    lb = np.zeros(nl, np.uint32)
    R.ref_flat_export(h, b.ctypes.data_as(f32p), idx.ctypes.data_as(u32p), lb.ctypes.data_as(u32p))
    return h, b, idx, lb


@pytest.mark.parametrize("n", [1, 2, 31, 32, 33, 1000, 1024, 1025, 40000])
def test_flat_rtree_node_for_node(O, R, n):
    """oracle (sort_mode 0 = the reference's quicksort) == the reference FlatRTree, array for array."""
    rng = np.random.default_rng(n)
    cx, cy = rng.uniform(0, 1000, n), rng.uniform(0, 1000, n)
    w = rng.uniform(0, 20, n)
    if n > 100:  # duplicates -> Hilbert key ties
        cx[: n // 4] = cx[n // 4 : 2 * (n // 4)]
        cy[: n // 4] = cy[n // 4 : 2 * (n // 4)]
    box = O.bbox_from_box2d(cx - w, cy - w, cx + w, cy + w)
    h, rb, ridx, rlb = _ref_flat(R, box)
    t = O.FlatTree(box, 32, sort_mode=0)
    assert list(t.layer_bounds) == list(rlb)
    assert t.nodes == len(ridx)
    assert (t.idx == ridx).all()
    assert (t.boxes.view(np.uint32) == rb.view(np.uint32)).all()
    # Lookup: reference code vs restatement, and the stable-sort tree returns the same SET
    ts = O.FlatTree(box, 32, sort_mode=1)
    f32p = C.POINTER(C.c_float)
    u32p = C.POINTER(C.c_uint32)
    for k in range(50):
        qx, qy, qw = rng.uniform(0, 1000), rng.uniform(0, 1000), rng.uniform(0, 100)
        q = np.array([qx - qw, qy - qw, qx + qw, qy + qw], np.float32)
        out = np.zeros(n, np.uint32)
        cnt = R.ref_flat_lookup(h, q.ctypes.data_as(f32p), out.ctypes.data_as(u32p), C.c_uint64(n))
        got = t.lookup(q)
        assert list(out[:cnt]) == list(got)  # same BFS order
        assert sorted(ts.lookup(q)) == sorted(got)
        brute = np.nonzero(~((q[0] > box[:, 2]) | (q[2] < box[:, 0]) | (q[1] > box[:, 3]) | (q[3] < box[:, 1])))[0]
        assert sorted(got) == list(brute)
    R.ref_flat_free(h)


def test_ref_join_count_matches_oracle(O, R):
    n, m = 20000, 3000
    px, py = O.generate_points(1337, n, (0, 0, 1000, 1000))
    u = O.uniform01(42, 2 * m)
    cx, cy = u[0::2] * 1000, u[1::2] * 1000
    b = [cx - 10, cy - 10, cx + 10, cy + 10]
    box = O.bbox_from_box2d(*b)
    f32p = C.POINTER(C.c_float)
    f64p = C.POINTER(C.c_double)
    h = R.ref_flat_build(box.ctypes.data_as(f32p), C.c_uint32(m), C.c_uint32(32))
    cand = C.c_uint64(0)
    bb = [np.ascontiguousarray(v) for v in b]
    for pred in (0, 1):
        tot = R.ref_flat_join_count(C.c_void_p(h), px.ctypes.data_as(f64p), py.ctypes.data_as(f64p), C.c_uint64(n),
                                    *[v.ctypes.data_as(f64p) for v in bb], C.c_int(pred), C.c_int(2), C.byref(cand))
        on, oc = O.join_points_boxes(px, py, *b, pred=pred, want_pairs=False)
        assert tot == on and cand.value == oc
    R.ref_flat_free(h)


def test_point_linestring_distance_vs_reference_segment_helper(O, RS):
    """orc_distance_point_linestring against the reference's own DistanceToSegmentSquared / ClosestPointOnSegment
    (spatial_functions_scalar.cpp:2158-2184, compiled in oracle/_ref/libref_scalar.so) folded with the loop of
    PointLineStringOperation (:2236-2256: `<` minimum, stop at 0, one sqrt)."""
    from test_native_scalar import make_lines

    rng = np.random.default_rng(17)
    px, py, lists, vx, vy = make_lines(rng, 400, max_len=30)
    got = O.distance_point_linestring(px, py, lists, vx, vy)
    dmax = np.finfo(np.float64).max
    nseg = 0
    for i in range(len(px)):
        o, l = int(lists[i, 0]), int(lists[i, 1])
        best = dmax
        for j in range(max(l - 1, 0)):
            d = RS.ref_dist_to_segment_squared(px[i], py[i], vx[o + j], vy[o + j], vx[o + j + 1], vy[o + j + 1])
            nseg += 1
            if d < best:
                best = d
                if best == 0:
                    break
        assert np.float64(got[i]).view(np.uint64) == np.sqrt(np.float64(best)).view(np.uint64), i
    assert nseg > 5000
