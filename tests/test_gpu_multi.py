"""-m gpu, needs >= 2 devices (skipped otherwise): the multi-GPU forms of the join against the oracle --
sb_join_points_boxes_multi (one process drives N devices: tree replicated by peer copies, probe slices, one host pair
list) and the one-process-per-GPU protocol over NCCL (rank 0 builds, tree broadcast, sharded probe, pair gather)."""
import os
import socket
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ndev():
    try:
        import torch

        return torch.cuda.device_count()
    except Exception:
        return 0


def sort_pairs(p, b):
    a = np.stack([np.asarray(p, np.int64), np.asarray(b, np.int64)], 1) if len(p) else np.zeros((0, 2), np.int64)
    return a[np.lexsort((a[:, 1], a[:, 0]))]


def _inputs(O, n, m, side=20.0):
    px, py = O.generate_points(1337, n, (0, 0, 10000, 10000))
    u = O.uniform01(42, 2 * m)
    cx, cy = u[0::2] * 10000, u[1::2] * 10000
    h = side / 2
    return px, py, [cx - h, cy - h, cx + h, cy + h]


@pytest.mark.skipif(_ndev() < 2, reason="needs >= 2 GPUs")
@pytest.mark.parametrize("pred", ["intersects", "within"])
def test_join_points_boxes_multi_vs_oracle(sb, O, pred):
    ndev = min(_ndev(), 4)
    ctxs = [sb.Context(d) for d in range(ndev)]
    for c in ctxs:
        c.set_option("pipe_chunk", 1 << 16)  # many chunks per device at test size
    try:
        px, py, b = _inputs(O, 1_500_001, 120_000)
        gp, gb = sb.spatial_join_points_boxes_multi(ctxs, px, py, *b, predicate=pred)
        exp, _ = O.join_points_boxes(px, py, *b, pred=0 if pred == "intersects" else 1)
        e = sort_pairs(exp[:, 0], exp[:, 1])
        g = sort_pairs(gp, gb)
        assert g.shape == e.shape and (g == e).all()
        # capacity protocol: too small -> SB_ERR_CAPACITY with the needed size
        with pytest.raises(sb.SpatialB200Error):
            sb.spatial_join_points_boxes_multi(ctxs, px, py, *b, predicate=pred, out=(np.empty(1000, np.uint32), np.empty(1000, np.uint32)))
        # one context = the single-device pipeline
        gp1, gb1 = sb.spatial_join_points_boxes_multi(ctxs[:1], px, py, *b, predicate=pred)
        assert (sort_pairs(gp1, gb1) == e).all()
    finally:
        for c in ctxs:
            c.close()


def _nccl_worker(rank, world, port, n, m, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    sys.path.insert(0, ROOT)
    import importlib

    import torch
    import torch.distributed as dist

    import oracle as O
    import sb200

    sb = sb200.load()
    D = importlib.import_module("duckdb_spatial_b200.distributed")
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", device_id=dev)
    ctx = sb.Context(rank)
    ctx.use_torch_stream()
    px, py, b = _inputs(O, n, m)

    def build_tree():
        tb = [torch.from_numpy(v).to(dev) for v in b]
        tree = sb.FlatRTree(ctx, ctx.bbox_from_box2d(*tb), 32)
        p, nb = tree.device_buffer()

        class _Raw:
            __cuda_array_interface__ = {"shape": (nb,), "typestr": "|u1", "data": (p, False), "version": 2}

        buf = torch.empty(nb, dtype=torch.uint8, device=dev)
        buf.copy_(torch.as_tensor(_Raw(), device=dev))
        ctx.sync()
        tree.close()
        return buf, torch.stack(tb, 1).contiguous()

    def probe_slice(tree_buf, boxes, beg, end):
        t = sb.FlatRTree.attach(ctx, tree_buf)
        x = torch.from_numpy(px[beg:end]).to(dev)
        y = torch.from_numpy(py[beg:end]).to(dev)
        pp, pb = t.probe_points(x, y, boxes, "intersects", index_base=beg)
        t.close()
        return pp, pb

    gp, gb, counts = D.sharded_join(n, build_tree, probe_slice, src=0)
    if rank == 0:
        q.put((gp.cpu().numpy(), gb.cpu().numpy(), counts))
    dist.barrier()
    dist.destroy_process_group()
    ctx.close()


@pytest.mark.skipif(_ndev() < 2, reason="needs >= 2 GPUs")
def test_nccl_sharded_join_gathered_pairs_vs_oracle(O):
    import torch.multiprocessing as mp

    world = 2
    n, m = 1_000_003, 80_000
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctxm = mp.get_context("spawn")
    q = ctxm.Queue()
    procs = [ctxm.Process(target=_nccl_worker, args=(r, world, port, n, m, q)) for r in range(world)]
    for p in procs:
        p.start()
    gp, gb, counts = q.get(timeout=600)
    for p in procs:
        p.join(timeout=600)
        assert p.exitcode == 0
    px, py, b = _inputs(O, n, m)
    exp, _ = O.join_points_boxes(px, py, *b, pred=0)
    assert sum(counts) == len(exp)
    assert (sort_pairs(gp, gb) == sort_pairs(exp[:, 0], exp[:, 1])).all()


def _build_worker(rank, world, port, n, cap, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    sys.path.insert(0, ROOT)
    import importlib

    import torch
    import torch.distributed as dist

    import oracle as O
    import sb200

    sb = sb200.load()
    D = importlib.import_module("duckdb_spatial_b200.distributed")
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", device_id=dev)
    ctx = sb.Context(rank)
    ctx.use_torch_stream()
    x, y = O.generate_points(31, n, (0, 0, 1000, 1000))
    x = np.round(x, 2)  # equal x keys: ties follow the global input position
    if cap == 16:
        y = np.round(y, 0)  # equal y keys inside the slices: (y key, x bin, input position)
    box = O.bbox_from_points(x, y)
    rows = np.arange(n, dtype=np.int64) * 3 + 7
    b, e = D.shard_range(n, rank, world)

    def sort_keys(k):
        kk = k.clone()
        vv = torch.arange(len(k), dtype=torch.int32, device=dev)
        ctx.sort_pairs(kk, vv)
        return kk, vv.long()

    tree = D.sharded_rtree_build(torch.from_numpy(box[b:e].copy()).to(dev), torch.from_numpy(rows[b:e].copy()).to(dev), cap, sort_keys,
                                 lambda bx, rw, c, s, nb, kb, ka: sb.RTreeIndex.leaf_shard(ctx, bx, rw, c, s, nb, kb, ka),
                                 lambda leaf, nn, nodes, c: sb.RTreeIndex.from_leaf_level(ctx, leaf, nn, nodes, c))
    if rank == 0:
        ref = sb.RTreeIndex(ctx, box, rows, cap)
        same = tree.nlevels == ref.nlevels and all(np.array_equal(np.asarray(tree.level(l)), np.asarray(ref.level(l))) for l in range(ref.nlevels))
        q.put(bool(same))
        ref.close()
        tree.close()
    dist.barrier()
    dist.destroy_process_group()
    ctx.close()


@pytest.mark.skipif(_ndev() < 2, reason="needs >= 2 GPUs")
@pytest.mark.parametrize("n,cap", [(200_003, 16), (1_500_000, 128)])
def test_nccl_sharded_rtree_build_equals_single_gpu_build(n, cap):
    import torch.multiprocessing as mp

    world = 2
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctxm = mp.get_context("spawn")
    q = ctxm.Queue()
    procs = [ctxm.Process(target=_build_worker, args=(r, world, port, n, cap, q)) for r in range(world)]
    for p in procs:
        p.start()
    same = q.get(timeout=600)
    for p in procs:
        p.join(timeout=600)
        assert p.exitcode == 0
    assert same, "the sharded build differs from the single-GPU build of the concatenated input"
