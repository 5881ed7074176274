"""Threading at the boundary: DuckDB calls the probe side concurrently from its worker threads
(spatial_join_physical.cpp:695-945: one OperatorState per thread, one shared global FlatRTree), and index scans
run while other threads work. Here: one built tree, several host threads, each with its own sb_context (= its
own stream and scratch arena), all probing at once through the C-ABI (ctypes drops the GIL during the call);
every thread's result must equal the single-threaded result."""
import threading

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def sort_pairs(p, b):
    a = np.stack([np.asarray(p, np.int64), np.asarray(b, np.int64)], 1) if len(p) else np.zeros((0, 2), np.int64)
    return a[np.lexsort((a[:, 1], a[:, 0]))]


def run_threads(nthreads, body):
    errs, out = [], [None] * nthreads
    start = threading.Barrier(nthreads)

    def work(i):
        try:
            start.wait()
            out[i] = body(i)
        except Exception as e:  # surfaced below
            errs.append((i, repr(e)))

    th = [threading.Thread(target=work, args=(i,)) for i in range(nthreads)]
    [t.start() for t in th]
    [t.join() for t in th]
    assert not errs, errs
    return out


def test_concurrent_probes_share_one_flat_tree(sb, O):
    main = sb.Context(0)
    n, m, T, rounds = 200000, 400000, 4, 3
    u = O.uniform01(42, 2 * n)
    cx, cy = u[0::2] * 10000, u[1::2] * 10000
    bx = np.stack([cx - 10, cy - 10, cx + 10, cy + 10], 1).copy()
    fbox = main.bbox_from_box2d(cx - 10, cy - 10, cx + 10, cy + 10)
    tree = sb.FlatRTree(main, fbox, 32)
    main.sync()
    px, py = O.generate_points(1337, m, (0, 0, 10000, 10000))
    parts = np.array_split(np.arange(m), T)
    ref = [sort_pairs(*tree.probe_points(px[s], py[s], bx, index_base=int(s[0]))) for s in parts]
    assert sum(len(r) for r in ref) > m // 2

    def body(i):
        c = sb.Context(0)
        view = sb.FlatRTree(c, _handle=tree.h, _keepalive=tree)
        try:
            s = parts[i]
            res = [sort_pairs(*view.probe_points(px[s], py[s], bx, index_base=int(s[0]))) for _ in range(rounds)]
        finally:
            view.h = None  # the handle belongs to `tree`
            c.close()
        return res

    got = run_threads(T, body)
    for i in range(T):
        for r in got[i]:
            assert r.shape == ref[i].shape and (r == ref[i]).all()
    tree.close()
    main.close()


def test_concurrent_joins_and_index_scans(sb, O):
    # whole-join calls (build + probe inside one call) and batched index scans from different threads at once
    main = sb.Context(0)
    n = 300000
    x, y = O.generate_points(7, n, (0, 0, 1000, 1000))
    idx = sb.RTreeIndex(main, main.bbox_from_points(x, y), np.arange(n, dtype=np.int64), 128)
    main.sync()
    rng = np.random.default_rng(5)
    qx, qy, qw = rng.uniform(0, 1000, 256), rng.uniform(0, 1000, 256), rng.uniform(0, 40, 256)
    q = main.bbox_from_box2d(qx, qy, qx + qw, qy + qw)
    sq, srow = idx.scan(q)
    m = 20000
    u = O.uniform01(43, 2 * m)
    b = [u[0::2] * 1000 - 2, u[1::2] * 1000 - 2, u[0::2] * 1000 + 2, u[1::2] * 1000 + 2]
    jref = sort_pairs(*main.spatial_join_points_boxes(x, y, *b))
    assert len(jref) > 0 and len(srow) > 0

    def body(i):
        c = sb.Context(0)
        try:
            if i % 2 == 0:
                return [sort_pairs(*c.spatial_join_points_boxes(x, y, *b)) for _ in range(3)]
            view = sb.RTreeIndex.__new__(sb.RTreeIndex)
            view.ctx, view.h, view.cap = c, idx.h, 128
            try:
                return [view.scan(q) for _ in range(5)]
            finally:
                view.h = None
        finally:
            c.close()

    got = run_threads(4, body)
    for i, res in enumerate(got):
        for r in res:
            if i % 2 == 0:
                assert r.shape == jref.shape and (r == jref).all()
            else:
                assert (r[0] == sq).all() and (r[1] == srow).all()
    idx.close()
    main.close()
