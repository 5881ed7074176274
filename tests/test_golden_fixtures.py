"""Parity against tests/golden/ref_vectors.npz -- outputs of the REFERENCE'S OWN code (scripts/make_golden.py ran
oracle/_ref/libref.so in the build container). Works where /root/reference does not exist: the CPU tests pin the
oracle, the -m gpu tests pin the CUDA path directly against the reference's outputs."""
import ctypes as C
import os

import numpy as np
import pytest

G = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_vectors.npz"))


def bits(a):
    return np.ascontiguousarray(a).view(np.uint32 if a.dtype == np.float32 else np.uint64)


def sort_pairs(p, b):
    a = np.stack([np.asarray(p, np.int64), np.asarray(b, np.int64)], 1) if len(p) else np.zeros((0, 2), np.int64)
    return a[np.lexsort((a[:, 1], a[:, 0]))]


def golden_lookup_pairs():
    off, hits = G["flat_q_off"], G["flat_q_hits"]
    q = np.repeat(np.arange(len(off) - 1), np.diff(off))
    return sort_pairs(q, hits)


# ---- CPU: the oracle against the reference's outputs ---------------------------------------------------------------
def test_oracle_rounding(O):
    d = G["round_in"]
    box = O.bbox_from_points(d, d)
    assert (bits(box[:, 0]) == bits(G["round_down"])).all() and (bits(box[:, 2]) == bits(G["round_up"])).all()


def test_oracle_hilbert_and_order(O):
    L = O.lib()
    got = np.array([L.orc_hilbert_encode(16, int(a), int(b)) for a, b in zip(G["hil_x"], G["hil_y"])], np.uint32)
    assert (got == G["hil_out"]).all()
    got = np.array([L.orc_hilbert_f32_to_u32(C.c_float(v)) for v in G["f32_in"]], np.uint32)
    assert (got == G["f32_order"]).all()
    h = O.hilbert_xy(G["hil_x"].astype(np.float64), G["hil_y"].astype(np.float64), np.array([0.0, 0, 65535, 65535]))
    assert (h == G["hil_out"]).all()


def test_oracle_generate_points(O):
    x, y = O.generate_points(1337, len(G["gen_x"]), (0, 0, 10000, 10000))
    assert (x == G["gen_x"]).all() and (y == G["gen_y"]).all()


def test_oracle_flat_tree_and_join(O):
    box = O.bbox_from_box2d(*G["flat_box2d"].T)
    assert (bits(box) == bits(G["flat_in"])).all()
    t = O.FlatTree(box, 32, sort_mode=0)  # the reference's quicksort order
    assert list(t.layer_bounds) == list(G["flat_layers"])
    assert (t.idx == G["flat_idx"]).all() and (bits(t.boxes) == bits(G["flat_boxes"])).all()
    exp = golden_lookup_pairs()
    got = [(i, int(b)) for i, q in enumerate(G["flat_q"]) for b in t.lookup(q)]
    assert (sort_pairs([g[0] for g in got], [g[1] for g in got]) == exp).all()
    b = [np.ascontiguousarray(G["flat_box2d"][:, k]) for k in range(4)]
    for pred in (0, 1):
        n, cand = O.join_points_boxes(G["join_px"], G["join_py"], *b, pred=pred, want_pairs=False)
        assert (n, cand) == tuple(G["join_counts"][pred])


# ---- GPU: the CUDA path against the reference's outputs ------------------------------------------------------------
@pytest.fixture(scope="module")
def ctx(sb):
    c = sb.Context(0)
    yield c
    c.close()


@pytest.mark.gpu
def test_gpu_rounding_hilbert_generator(ctx):
    d = G["round_in"]
    box = ctx.bbox_from_points(d, d)
    assert (bits(box[:, 0]) == bits(G["round_down"])).all() and (bits(box[:, 2]) == bits(G["round_up"])).all()
    h = ctx.st_hilbert_xy(G["hil_x"].astype(np.float64), G["hil_y"].astype(np.float64), np.array([0.0, 0, 65535, 65535]))
    assert (h == G["hil_out"]).all()
    x, y = ctx.st_generate_points(1337, len(G["gen_x"]), (0, 0, 10000, 10000))
    assert (x == G["gen_x"]).all() and (y == G["gen_y"]).all()


@pytest.mark.gpu
def test_gpu_flat_tree_vs_reference_arrays(ctx, sb):
    b2d = G["flat_box2d"]
    box = ctx.bbox_from_box2d(*[np.ascontiguousarray(b2d[:, k]) for k in range(4)])
    assert (bits(box) == bits(G["flat_in"])).all()
    t = sb.FlatRTree(ctx, box, 32)
    assert list(t.layer_bounds) == list(G["flat_layers"]) and t.nodes == len(G["flat_idx"])
    gb, gi = t.export()
    n = len(b2d)
    # leaves: same items, each with its own box (the order inside runs of equal Hilbert keys is the documented
    # difference: stable radix sort here, Hoare quicksort in the reference -- SURVEY D8)
    assert sorted(gi[:n]) == list(range(n))
    assert (bits(gb[:n]) == bits(G["flat_in"][gi[:n]])).all()
    # root box = union of everything, bit for bit
    assert (bits(gb[-1]) == bits(G["flat_boxes"][-1])).all()
    # every internal box is the exact union of its children
    lb = [0] + list(t.layer_bounds)
    for lvl in range(1, len(lb) - 1):
        beg, end = lb[lvl], lb[lvl + 1]
        kids_beg, kids_end = lb[lvl - 1], lb[lvl]
        for j in (0, (end - beg) // 2, end - beg - 1):
            k0 = gi[beg + j]
            k1 = min(k0 + 32, kids_end)
            ch = gb[k0:k1]
            u = np.array([ch[:, 0].min(), ch[:, 1].min(), ch[:, 2].max(), ch[:, 3].max()], np.float32)
            assert kids_beg <= k0 < kids_end and (bits(u) == bits(gb[beg + j])).all()
    # Lookup result sets == the reference's, through the tree walk (box probes)
    p, b = t.probe(G["flat_q"])
    assert (sort_pairs(p, b) == golden_lookup_pairs()).all()


@pytest.mark.gpu
@pytest.mark.parametrize("path", [1, 2, 3])
def test_gpu_join_counts_vs_reference(ctx, sb, path):
    b2d = G["flat_box2d"]
    cols = [np.ascontiguousarray(b2d[:, k]) for k in range(4)]
    ctx.set_option("probe_path", path)
    try:
        for pred, name in ((0, "intersects"), (1, "within")):
            p, b = ctx.spatial_join_points_boxes(G["join_px"], G["join_py"], *cols, predicate=name)
            assert len(p) == G["join_counts"][pred][0]
        # candidates (float-box hits) == the reference's candidate count
        box = ctx.bbox_from_box2d(*cols)
        t = sb.FlatRTree(ctx, box, 32)
        cp, cb = t.probe_points(G["join_px"], G["join_py"])
        assert len(cp) == G["join_counts"][0][1]
    finally:
        ctx.set_option("probe_path", 0)
