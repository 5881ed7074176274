"""Persistent-index export: the STR tree as FixedSizeAllocator block images with IndexPointers.

CPU part: the oracle's sequential emulation of BuildRTreeBottomUp's MakePage order over two fresh FixedSizeAllocators
(oracle.c orc_rtree_export_pages) is read back here the way the reference reads an index: IndexPointer ->
(buffer id, offset, node type) -> segment -> RTreeNode (fixed_size_allocator.hpp:53-61, rtree.cpp:44-52), depth-first like
rtree_index_dump (rtree_index_pragmas.cpp:137-227) -- and must give the oracle's own dump of the tree.
GPU part: sb_rtree_export_pages is byte-identical to the oracle's images.
"""
import numpy as np
import pytest

ENTRY = np.dtype([("ptr", np.uint64), ("box", np.float32, 4)])
ROW_ID, LEAF_PAGE, BRANCH_PAGE = 0, 1, 2


@pytest.fixture(scope="module")
def ctx(sb):
    c = sb.Context(0)
    yield c
    c.close()


def read_node(info, block_size, leaf, branch, ptr):
    """RTree::Ref: pick the allocator by node type, then FixedSizeAllocator::Get."""
    ptr = int(ptr)
    typ, offset, buffer_id = ptr >> 56, (ptr >> 32) & 0xFFFFFF, ptr & 0xFFFFFFFF
    assert typ in (LEAF_PAGE, BRANCH_PAGE)
    img = leaf if typ == LEAF_PAGE else branch
    assert offset < info["segments_per_buffer"]
    base = buffer_id * block_size + info["bitmask_offset"] + offset * info["segment_size"]
    # the segment must be marked taken (validity bit clear) in its buffer's bitmask
    word = img[buffer_id * block_size + 8 * (offset // 64) :][:8].view(np.uint64)[0]
    assert (int(word) >> (offset % 64)) & 1 == 0
    seg = img[base : base + info["segment_size"]]
    count = int(seg[:4].view(np.uint32)[0])
    assert not seg[4:24].any()
    ents = seg[24:].view(ENTRY)
    assert not ents[count:]["ptr"].any()
    return typ, ents[:count]


def dump_pages(info, block_size, leaf, branch, root):
    lvl, box, rid = [], [], []

    def rec(ptr, depth):
        typ, ents = read_node(info, block_size, leaf, branch, ptr)
        for e in ents:
            et = int(e["ptr"]) >> 56
            assert et == (ROW_ID if typ == LEAF_PAGE else et) and (typ == LEAF_PAGE) == (et == ROW_ID)
            lvl.append(depth)
            box.append(e["box"].copy())
            rid.append(int(e["ptr"]) if et == ROW_ID else -1)
            if et != ROW_ID:
                rec(e["ptr"], depth + 1)

    if int(root["ptr"]) != 0:
        rec(root["ptr"], 0)
    return np.array(lvl, np.int32), np.array(box, np.float32).reshape(-1, 4), np.array(rid, np.int64)


def check_bitmasks(info, block_size, img, pages):
    spb = info["segments_per_buffer"]
    nbuf = len(img) // block_size
    assert nbuf == (pages + spb - 1) // spb
    for b in range(nbuf):
        used = min(spb, pages - b * spb)
        words = img[b * block_size : b * block_size + info["bitmask_offset"]].view(np.uint64)
        bits = np.array([(int(words[i // 64]) >> (i % 64)) & 1 for i in range(64 * len(words))])
        assert not bits[:used].any() and bits[used:spb].all() and not bits[spb:].any()


def make_tree_input(O, n, cap):
    rng = np.random.default_rng(n * 11 + cap)
    x, y = O.generate_points(1337, n, (0, 0, 1000, 1000))
    w = rng.uniform(0, 3, n)
    box = O.bbox_from_box2d(x - w, y - w, x + w, y + w)
    rows = rng.permutation(n).astype(np.int64) + 5
    return box, rows


def test_fsa_layout_default_block(O):
    # 128-entry nodes in the default 256 KiB block: 84 segments of 3096 B behind two validity words
    t = O.StrTree(*make_tree_input(O, 10, 128), 128)
    info = t.pages(262136)[0]
    assert (info["segment_size"], info["segments_per_buffer"], info["bitmask_count"], info["bitmask_offset"]) == (
        3096, 84, 2, 16)
    assert 16 + 84 * 3096 <= 262136 < 16 + 85 * 3096


CASES = [(1, 128, 262136), (2, 128, 262136), (128, 128, 262136), (129, 128, 262136), (20000, 128, 262136),
         (5000, 4, 4096), (30000, 9, 1024), (100000, 255, 16384)]


@pytest.mark.parametrize("n,cap,block", CASES)
def test_oracle_pages_read_back_as_the_tree(O, n, cap, block):
    box, rows = make_tree_input(O, n, cap)
    t = O.StrTree(box, rows, cap)
    info, leaf, branch, root = t.pages(block)
    counts = t.level_counts()
    assert info["leaf_pages"] == counts[1] and info["branch_pages"] == sum(counts[2:])
    check_bitmasks(info, block, leaf, info["leaf_pages"])
    check_bitmasks(info, block, branch, info["branch_pages"])
    lvl, bx, rid = dump_pages(info, block, leaf, branch, root)
    elvl, ebx, erid = t.dump()
    assert np.array_equal(lvl, elvl) and np.array_equal(rid, erid)
    assert np.array_equal(bx.view(np.uint32), ebx.view(np.uint32))
    assert np.array_equal(root["box"].view(np.uint32), t.level(t.nlevels - 1)[0]["box"].view(np.uint32))
    assert int(root["ptr"]) >> 56 == (LEAF_PAGE if t.nlevels == 2 else BRANCH_PAGE)


@pytest.mark.gpu
@pytest.mark.parametrize("n,cap,block", CASES)
@pytest.mark.parametrize("device", [False, True])
def test_gpu_pages_match_oracle_images(ctx, sb, O, n, cap, block, device):
    box, rows = make_tree_input(O, n, cap)
    idx = sb.RTreeIndex(ctx, box, rows, cap)
    einfo, eleaf, ebranch, eroot = O.StrTree(box, rows, cap).pages(block)
    info, leaf, branch, root = idx.pages(block, device=device)
    if device:
        leaf, branch = leaf.cpu().numpy(), branch.cpu().numpy()
    assert info == einfo
    assert root.tobytes() == eroot.tobytes()
    assert np.array_equal(leaf, eleaf)
    assert np.array_equal(branch, ebranch)


@pytest.mark.gpu
def test_gpu_pages_empty_tree(ctx, sb):
    idx = sb.RTreeIndex(ctx, np.zeros((0, 4), np.float32), np.zeros(0, np.int64))
    info, leaf, branch, root = idx.pages()
    assert info["leaf_pages"] == 0 and info["branch_pages"] == 0 and len(leaf) == 0 and len(branch) == 0
    assert not root.tobytes().strip(b"\0")


@pytest.mark.gpu
def test_gpu_pages_c1_scale_read_back(ctx, sb, O):
    # 500 k points, default block: the images read back (IndexPointer walk) as the tree the batched scan sees
    n = 500_000
    x, y = O.generate_points(1337, n, (0, 0, 10000, 10000))
    box = ctx.bbox_from_points(x, y)
    idx = sb.RTreeIndex(ctx, box, np.arange(n, dtype=np.int64), 128)
    info, leaf, branch, root = idx.pages()
    assert info["leaf_pages"] == idx.level_counts()[1]
    check_bitmasks(info, 262136, leaf, info["leaf_pages"])
    check_bitmasks(info, 262136, branch, info["branch_pages"])
    lvl, bx, rid = dump_pages(info, 262136, leaf, branch, root)
    leaves = rid >= 0
    assert leaves.sum() == n and np.array_equal(np.sort(rid[leaves]), np.arange(n))
    assert np.array_equal(bx[leaves][np.argsort(rid[leaves])].view(np.uint32), np.asarray(box).view(np.uint32))
    assert (np.bincount(lvl) == np.array(idx.level_counts()[::-1][1:])).all()
