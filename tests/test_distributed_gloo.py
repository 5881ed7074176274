"""CPU, world_size 2, gloo: the multi-GPU host logic of the join (duckdb-spatial_b200/distributed.py) --
probe sharding with global row ids, broadcast of the packed tree + exact boxes, gather of variable-length pair lists.
The device work is injected; here the oracle stands in for the probe kernels (tests may use the oracle)."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n, m, out_dir):
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
        px, py = O.generate_points(1337, n, (0, 0, 1000, 1000))

        def build_tree():
            u = O.uniform01(42, 2 * m)
            cx, cy = u[0::2] * 1000, u[1::2] * 1000
            boxes = np.stack([cx - 10, cy - 10, cx + 10, cy + 10], 1)
            # stand-in for the packed FlatRTree buffer: any byte string only rank 0 knows
            tree = np.frombuffer(O.bbox_from_box2d(*boxes.T).tobytes(), np.uint8).copy()
            return torch.from_numpy(tree), torch.from_numpy(boxes)

        def probe_slice(tree, boxes, b, e):
            bx = boxes.numpy()
            assert tree.numel() == 16 * len(bx)  # the broadcast delivered the whole "tree"
            pairs, _ = O.join_points_boxes(px[b:e], py[b:e], bx[:, 0], bx[:, 1], bx[:, 2], bx[:, 3], pred=0)
            pairs = pairs.astype(np.int64)
            return torch.from_numpy(pairs[:, 0] + b).contiguous(), torch.from_numpy(pairs[:, 1].copy())

        gp, gb, counts = D.sharded_join(n, build_tree, probe_slice, src=0)
        # shard_range covers [0, n) exactly once
        ranges = [D.shard_range(n, r, world) for r in range(world)]
        assert ranges[0][0] == 0 and ranges[-1][1] == n and all(ranges[i][1] == ranges[i + 1][0] for i in range(world - 1))
        if rank == 0:
            np.save(os.path.join(out_dir, "pairs.npy"), np.stack([gp.numpy(), gb.numpy()], 1))
            np.save(os.path.join(out_dir, "counts.npy"), np.array(counts))
        else:
            assert gp is None and gb is None
        # empty-shard edge case: fewer rows than ranks
        gp2, gb2, c2 = D.gather_pairs(torch.arange(rank, dtype=torch.int64), torch.arange(rank, dtype=torch.int64), dst=0)
        if rank == 0:
            assert c2 == list(range(world)) and gp2.numel() == sum(range(world))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n,m", [(20001, 700), (1, 5)])
def test_sharded_join_world2(O, tmp_path, n, m):
    import torch.multiprocessing as mp

    world = 2
    port = _free_port()
    mp.spawn(_worker, args=(world, port, n, m, str(tmp_path)), nprocs=world, join=True)
    got = np.load(tmp_path / "pairs.npy")
    counts = np.load(tmp_path / "counts.npy")
    px, py = O.generate_points(1337, n, (0, 0, 1000, 1000))
    u = O.uniform01(42, 2 * m)
    cx, cy = u[0::2] * 1000, u[1::2] * 1000
    exp, _ = O.join_points_boxes(px, py, cx - 10, cy - 10, cx + 10, cy + 10, pred=0)
    exp = exp.astype(np.int64)
    assert counts.sum() == len(exp) == len(got)
    g = got[np.lexsort((got[:, 1], got[:, 0]))]
    e = exp[np.lexsort((exp[:, 1], exp[:, 0]))]
    assert (g == e).all()
    # rank order: rank 0's rows come first
    half = n // 2 + n % 2
    assert (got[: counts[0], 0] < half).all() and (got[counts[0] :, 0] >= half).all()


def test_shard_range_properties():
    sys.path.insert(0, ROOT)
    import importlib

    import sb200

    sb200.load()
    D = importlib.import_module("duckdb_spatial_b200.distributed")
    for n in (0, 1, 7, 8, 9, 200_000_000):
        for w in (1, 2, 4, 8):
            r = [D.shard_range(n, k, w) for k in range(w)]
            assert r[0][0] == 0 and r[-1][1] == n
            sizes = [e - b for b, e in r]
            assert max(sizes) - min(sizes) <= 1 and all(r[i][1] == r[i + 1][0] for i in range(w - 1))
