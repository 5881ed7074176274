"""CPU, world_size 2 and 3, gloo: the host logic of the multi-GPU CREATE INDEX (duckdb-spatial_b200/distributed.py
sharded_rtree_build: unique 63-bit keys, exact slice-aligned splitters by bisection, all-to-all, leaf level per rank,
gather). The device work is injected: a numpy restatement of the leaf level stands in for sb_rtree_build_leaf_shard,
and the result is compared with the oracle's STR tree of the concatenated input (levels 0 and 1)."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ENTRY = np.dtype([("ptr", np.uint64), ("box", np.float32, 4)])


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _f32_key(f):
    f = (f + np.float32(0.0)).astype(np.float32)
    r = f.view(np.uint32).astype(np.uint64)
    return np.where(r & 0x80000000, 0xFFFFFFFF - r, r | 0x80000000)


def np_leaf_shard(box, rows, cap, slice_size, node_base, key_before=-1, key_after=-1):
    """STR leaf level of a slice-aligned shard (rtree_index_create_physical.cpp:136-256), rows in ascending global input
    position: slices of the (x key, position) order, inside a slice by (y key, x bin, position) with x bin = #{cut < key}
    + #{cut <= key} over the x keys at the slice boundaries (and at the shard's two ends), `cap` entries per node ordered
    by row id, node box = fp32 union"""
    n = len(rows)
    xk = _f32_key(box[:, 0] + box[:, 2])
    xo = np.argsort(xk, kind="stable")
    cuts = [int(xk[xo[r]]) for r in range(slice_size, n, slice_size)]
    cuts = ([key_before] if key_before >= 0 else []) + cuts + ([key_after] if key_after >= 0 else [])
    cuts = np.array(cuts, np.uint64)
    bins = np.searchsorted(cuts, xk, "left") + np.searchsorted(cuts, xk, "right")
    yk = _f32_key((box[:, 1] + box[:, 3]) / np.float32(2))
    parts = []
    for s in range(0, n, slice_size):
        idx = xo[s:s + slice_size]
        parts.append(idx[np.lexsort((idx, bins[idx], yk[idx]))])
    perm = np.concatenate(parts) if n else xo
    leaf = np.zeros(n, ENTRY)
    nodes = np.zeros((n + cap - 1) // cap, ENTRY)
    for j in range(len(nodes)):
        idx = perm[j * cap:(j + 1) * cap]
        idx = idx[np.argsort(rows[idx], kind="stable")]
        leaf["ptr"][j * cap:j * cap + len(idx)] = rows[idx].astype(np.uint64)
        leaf["box"][j * cap:j * cap + len(idx)] = box[idx]
        b = box[idx]
        nodes["ptr"][j] = (1 << 56) | (node_base + j)
        nodes["box"][j] = [b[:, 0].min(), b[:, 1].min(), b[:, 2].max(), b[:, 3].max()]
    return leaf, nodes


def _worker(rank, world, port, n, cap, out_dir):
    sys.path.insert(0, ROOT)
    import importlib

    import torch
    import torch.distributed as dist

    import oracle as O
    import sb200

    sb200.load()
    D = importlib.import_module("duckdb_spatial_b200.distributed")
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    try:
        x, y = O.generate_points(31, n, (0, 0, 100, 100))
        x = np.round(x, 1)  # many equal x keys: ties must follow the global input position
        if n == 20_000:
            y = np.round(y, 0)  # and many equal y keys inside every slice: (y key, x bin, input position)
        box = O.bbox_from_points(x, y)
        rows = np.arange(n, dtype=np.int64) * 3 + 7
        b, e = D.shard_range(n, rank, world)

        def sort_keys(k):
            s, p = torch.sort(k, stable=True)
            return s, p

        def leaf_shard(bx, rw, cap_, slice_size, node_base, key_before, key_after):
            leaf, nodes = np_leaf_shard(bx.numpy(), rw.numpy(), cap_, slice_size, node_base, key_before, key_after)
            return (torch.from_numpy(leaf.view(np.uint8).reshape(-1, 24).copy()), torch.from_numpy(nodes.view(np.uint8).reshape(-1, 24).copy()))

        def assemble(leaf_all, n_all, nodes_all, cap_):
            return leaf_all.numpy().reshape(-1).view(ENTRY), nodes_all.numpy().reshape(-1).view(ENTRY)

        res = D.sharded_rtree_build(torch.from_numpy(box[b:e].copy()), torch.from_numpy(rows[b:e].copy()), cap, sort_keys, leaf_shard, assemble)
        if rank == 0:
            np.save(os.path.join(out_dir, "leaf.npy"), res[0])
            np.save(os.path.join(out_dir, "nodes.npy"), res[1])
        else:
            assert res is None
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n,cap", [(2, 5001, 16), (3, 20_000, 8), (2, 7, 4), (2, 40_000, 128)])
def test_sharded_rtree_build_gloo(O, tmp_path, world, n, cap):
    import torch.multiprocessing as mp

    mp.spawn(_worker, args=(world, _free_port(), n, cap, str(tmp_path)), nprocs=world, join=True)
    leaf, nodes = np.load(tmp_path / "leaf.npy"), np.load(tmp_path / "nodes.npy")
    x, y = O.generate_points(31, n, (0, 0, 100, 100))
    x = np.round(x, 1)
    if n == 20_000:
        y = np.round(y, 0)
    box = O.bbox_from_points(x, y)
    rows = np.arange(n, dtype=np.int64) * 3 + 7
    t = O.StrTree(box, rows, cap)
    l0 = t.level(0)
    assert len(leaf) == n and np.array_equal(leaf["ptr"], l0["ptr"]) and np.array_equal(leaf["box"], l0["box"])
    # level 1 of the oracle holds the same node entries, re-sorted by y inside the slices: compare as sets
    l1 = t.level(1) if t.nlevels > 2 else None
    if l1 is not None:
        a = np.sort(nodes.view(np.uint8).reshape(-1, 24).copy().view([("b", "V24")]).ravel())
        b = np.sort(np.ascontiguousarray(l1).view(np.uint8).reshape(-1, 24).copy().view([("b", "V24")]).ravel())
        assert np.array_equal(a, b)
