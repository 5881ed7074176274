"""GPU parity for the probe-side operator glue (sb_join_op_*): a probe side streamed in 2048-row chunks, probed in device
batches, must produce exactly the reference operator's output -- INNER pairs (spatial_join_physical.cpp:748-865), the
NULL-padded probe rows of LEFT / FULL (:867-938) and the NULL-padded build rows of RIGHT / FULL (:1022-1068) -- as
computed from the oracle's pair set."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx(sb):
    c = sb.Context(0)
    yield c
    c.close()


def sort_pairs(p, b):
    a = np.stack([np.asarray(p, np.int64), np.asarray(b, np.int64)], 1)
    return a[np.lexsort((a[:, 1], a[:, 0]))] if len(a) else a.reshape(0, 2)


def _data(seed, n, m):
    rng = np.random.default_rng(seed)
    cx, cy = rng.uniform(0, 1000, m), rng.uniform(0, 1000, m)
    w = rng.uniform(0.5, 9, m)
    # a corner of the domain without boxes (unmatched probe rows) and boxes outside the points (unmatched build rows)
    cx[: m // 10] += 2000
    px, py = rng.uniform(0, 1000, n), rng.uniform(0, 1000, n)
    px[:7] = cx[m - 7:] - w[m - 7:]  # points exactly on box edges: closed vs strict predicate
    py[:7] = cy[m - 7:]
    return px, py, [cx - w, cy - w, cx + w, cy + w]


@pytest.mark.parametrize("join_type", ["inner", "left", "right", "full"])
@pytest.mark.parametrize("predicate", ["intersects", "within"])
def test_streamed_join_matches_the_operator(ctx, O, sb, join_type, predicate):
    n, m = 70001, 3000
    px, py, b = _data(5, n, m)
    exp, _ = O.join_points_boxes(px, py, *b, pred=0 if predicate == "intersects" else 1)
    e = sort_pairs(exp[:, 0], exp[:, 1])
    op = sb.SpatialJoinOp(ctx, *b, predicate=predicate, join_type=join_type, batch_rows=20000)
    pairs_p, pairs_b, lhs = [], [], []
    flushes = 0
    for beg in range(0, n, 2048):  # DuckDB's STANDARD_VECTOR_SIZE
        r = op.push(px[beg:beg + 2048], py[beg:beg + 2048])
        if r is not None:
            flushes += 1
            pairs_p.append(r[0]), pairs_b.append(r[1]), lhs.append(r[2])
            # the unmatched rows of a batch are ascending
            assert (np.diff(r[2].astype(np.int64)) > 0).all()
    r = op.flush()
    pairs_p.append(r[0]), pairs_b.append(r[1]), lhs.append(r[2])
    rhs = op.finish()
    assert flushes == 70001 // (2048 * 9)  # 9 chunks of 2048 fit a 20000-row batch
    assert ctx.L.sb_join_op_rows_pushed(op.h) == n
    op.close()
    g = sort_pairs(np.concatenate(pairs_p), np.concatenate(pairs_b))
    assert g.shape == e.shape and (g == e).all()
    lhs = np.concatenate(lhs)
    if join_type in ("left", "full"):
        assert (lhs == np.setdiff1d(np.arange(n), e[:, 0])).all()
    else:
        assert len(lhs) == 0
    if join_type in ("right", "full"):
        assert (rhs == np.setdiff1d(np.arange(m), e[:, 1])).all() and len(rhs) >= m // 10
    else:
        assert len(rhs) == 0


def test_join_op_protocol(ctx, O, sb):
    n, m = 5000, 400
    px, py, b = _data(6, n, m)
    exp, _ = O.join_points_boxes(px, py, *b, pred=0)
    op = sb.SpatialJoinOp(ctx, *b, join_type="full", batch_rows=4096)
    L = ctx.L
    pend = C.c_uint64(0)
    assert L.sb_join_op_push(op.h, px.ctypes.data, py.ctypes.data, 4096, C.byref(pend)) == 0 and pend.value == 4096
    # a chunk that does not fit the batch is refused, nothing is lost
    assert L.sb_join_op_push(op.h, px.ctypes.data, py.ctypes.data, 1, C.byref(pend)) == sb.capi.ERR_CAPACITY
    # finish with rows pending is an error
    nr = C.c_uint64(0)
    assert L.sb_join_op_finish(op.h, None, 0, C.byref(nr)) == sb.capi.ERR_ARG
    # pair capacity too small: the batch stays pending and the needed size comes back
    first = int((exp[:, 0] < 4096).sum())
    p1, b1 = np.zeros(8, np.uint32), np.zeros(8, np.uint32)
    lhs = np.zeros(4096, np.uint32)
    npairs, nl = C.c_uint64(0), C.c_uint64(0)
    rc = L.sb_join_op_flush(op.h, p1.ctypes.data, b1.ctypes.data, 8, C.byref(npairs), lhs.ctypes.data, 4096, C.byref(nl))
    assert rc == sb.capi.ERR_CAPACITY and npairs.value == first
    # unmatched-rows capacity too small
    rc = L.sb_join_op_flush(op.h, p1.ctypes.data, b1.ctypes.data, 8, C.byref(npairs), lhs.ctypes.data, 10, C.byref(nl))
    assert rc == sb.capi.ERR_CAPACITY and nl.value == 4096
    op.pending = 4096
    r = op.flush(capacity=first)
    assert len(r[0]) == first
    op.push(px[4096:], py[4096:])
    r2 = op.flush()
    rhs = op.finish()
    g = sort_pairs(np.concatenate([r[0], r2[0]]), np.concatenate([r[1], r2[1]]))
    e = sort_pairs(exp[:, 0], exp[:, 1])
    assert (g == e).all() and (rhs == np.setdiff1d(np.arange(m), e[:, 1])).all()
    # an empty flush is legal
    assert len(op.flush()[0]) == 0
    op.close()
    # empty build side: every probe row is unmatched in a LEFT join
    z = np.zeros(0)
    op = sb.SpatialJoinOp(ctx, z, z, z, z, join_type="left", batch_rows=1000)
    op.push(px[:300], py[:300])
    r = op.flush()
    assert len(r[0]) == 0 and (r[2] == np.arange(300)).all() and len(op.finish()) == 0
    op.close()
