"""N2: the bulk-loaded tree as FixedSizeAllocator block images is loaded into the REFERENCE's own RTree
through the reference's own allocators (oracle/ref_duckdb.cpp rd_rtree_load: FixedSizeAllocator::New() pointer sequence
and whole-buffer bytes must equal the images), then the reference's own code scans it and inserts into / deletes from
it (rtree.cpp:73-592). CPU part: the oracle's images (pins the allocator arithmetic that was "parity unpinned" in
round 1). GPU part: the same with sb_rtree_export_pages images."""
import numpy as np
import pytest


def _invariants(lv, box, rid):
    """every branch entry's box is the union of the entries of its node (checked on the DFS dump)"""
    n = len(lv)
    stack = []  # (index of the branch entry, level)
    acc = {}
    for i in range(n):
        while stack and lv[stack[-1]] >= lv[i]:
            j = stack.pop()
            assert np.array_equal(acc.pop(j), box[j]), j
        if stack:
            p = stack[-1]
            b = acc[p]
            acc[p] = np.array([min(b[0], box[i][0]), min(b[1], box[i][1]), max(b[2], box[i][2]), max(b[3], box[i][3])], np.float32)
        if rid[i] < 0:
            stack.append(i)
            acc[i] = np.array([np.inf, np.inf, -np.inf, -np.inf], np.float32)
    while stack:
        j = stack.pop()
        assert np.array_equal(acc.pop(j), box[j]), j


def _exercise(RDB, O, box, rows, cap, images):
    info, leaf, branch, root = images
    db = RDB.RefDB()
    t = RDB.RefRTree(db, cap, leaf, info["leaf_pages"], branch, info["branch_pages"], np.frombuffer(root.tobytes(), np.uint8))
    assert t.mismatch == 0, "pointer sequence / buffer bytes differ from the reference allocator (%d)" % t.mismatch
    # the reference's scanner over the loaded pages == the oracle's dump of the same tree
    lv, b, rid = t.dump()
    olv, obox, orid = O.StrTree(box, rows, cap).dump()
    assert np.array_equal(lv, olv) and np.array_equal(b, obox.reshape(-1, 4)) and np.array_equal(rid, orid)
    q = np.array([200, 200, 420, 380], np.float32)
    hit = ~((box[:, 0] > q[2]) | (box[:, 2] < q[0]) | (box[:, 1] > q[3]) | (box[:, 3] < q[1]))
    assert np.array_equal(np.sort(t.scan(q)), np.sort(rows[hit]))
    # dynamic maintenance by the reference's code on the bulk-loaded tree: delete a third, insert new rows
    rng = np.random.default_rng(3)
    n = len(rows)
    gone = rng.permutation(n)[: n // 3]
    t.delete(box[gone], rows[gone])
    x, y = rng.uniform(0, 1000, n // 4), rng.uniform(0, 1000, n // 4)
    nb = O.bbox_from_points(x, y)
    nr = np.arange(len(x), dtype=np.int64) + 10 * n
    t.insert(nb, nr)
    keep = np.ones(n, bool)
    keep[gone] = False
    all_box = np.concatenate([box[keep], nb])
    all_row = np.concatenate([rows[keep], nr])
    lv, b, rid = t.dump()
    leaf_rows = rid[rid >= 0]
    assert np.array_equal(np.sort(leaf_rows), np.sort(all_row))
    order = np.argsort(all_row)
    got_order = np.argsort(leaf_rows)
    assert np.array_equal(b[rid >= 0][got_order], all_box[order])
    _invariants(lv, b, rid)
    hit = ~((all_box[:, 0] > q[2]) | (all_box[:, 2] < q[0]) | (all_box[:, 1] > q[3]) | (all_box[:, 3] < q[1]))
    assert np.array_equal(np.sort(t.scan(q)), np.sort(all_row[hit]))
    t.close()
    db.close()


@pytest.mark.parametrize("n,cap", [(1, 128), (100, 128), (5000, 16), (60_000, 128)])
def test_oracle_page_images_load_into_reference_rtree(O, RDB, n, cap):
    x, y = O.generate_points(11, n, (0, 0, 1000, 1000))
    box = O.bbox_from_points(x, y)
    rows = np.arange(n, dtype=np.int64) * 2 + 1
    _exercise(RDB, O, box, rows, cap, O.StrTree(box, rows, cap).pages())


@pytest.mark.gpu
@pytest.mark.parametrize("n,cap", [(5000, 16), (300_000, 128)])
def test_gpu_page_images_load_into_reference_rtree(O, RDB, sb, n, cap):
    ctx = sb.Context(0)
    x, y = O.generate_points(12, n, (0, 0, 1000, 1000))
    box = O.bbox_from_points(x, y)
    rows = np.arange(n, dtype=np.int64) * 2 + 1
    idx = sb.RTreeIndex(ctx, box, rows, cap)
    info, leaf, branch, root = idx.pages()
    _exercise(RDB, O, box, rows, cap, (info, leaf, branch, root))
    idx.close()
    ctx.close()
