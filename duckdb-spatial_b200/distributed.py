"""Multi-GPU plumbing of the join probe: one process per GPU over torch.distributed.

The probe shards naturally: rank 0 builds the FlatRTree, the packed tree (one buffer) and the build side's exact boxes
are broadcast, every rank probes a contiguous slice of the probe column with global row ids, and the variable-length
pair lists are gathered. There is no collective on the probe path itself. Everything here is backend-agnostic
(NCCL on CUDA tensors for the product; gloo on CPU tensors for the world-size-2 tests of this logic).
"""
import torch
import torch.distributed as dist


def shard_range(n, rank, world):
    """contiguous slice [begin, end) of n probe rows for `rank`: sizes differ by at most one row."""
    base, rem = divmod(int(n), int(world))
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


def broadcast_bytes(buf, nbytes_on_src, src=0, device=None):
    """broadcast a byte buffer whose size only the source knows. Returns the (received) uint8 tensor."""
    rank = dist.get_rank()
    device = device if device is not None else (buf.device if buf is not None else torch.device("cpu"))
    size = torch.tensor([int(nbytes_on_src) if rank == src else 0], dtype=torch.int64, device=device)
    dist.broadcast(size, src)
    n = int(size.item())
    if rank != src:
        buf = torch.empty(n, dtype=torch.uint8, device=device)
    else:
        assert buf.numel() == n and buf.dtype == torch.uint8
    if n:
        dist.broadcast(buf, src)
    return buf


def gather_pairs(pair_probe, pair_build, dst=0):
    """gather variable-length (probe row, build row) lists to `dst`.

    Returns (probe, build, counts) on dst -- concatenated in rank order -- and (None, None, counts) elsewhere."""
    rank, world = dist.get_rank(), dist.get_world_size()
    device = pair_probe.device
    mine = torch.tensor([pair_probe.numel()], dtype=torch.int64, device=device)
    counts = torch.zeros(world, dtype=torch.int64, device=device)
    dist.all_gather_into_tensor(counts, mine)
    cl = [int(c) for c in counts.tolist()]
    if world == 1:
        return pair_probe, pair_build, cl
    if rank == dst:
        total = sum(cl)
        gp = torch.empty(total, dtype=pair_probe.dtype, device=device)
        gb = torch.empty(total, dtype=pair_build.dtype, device=device)
        ops, o = [], 0
        for r in range(world):
            if r == dst:
                gp[o : o + cl[r]].copy_(pair_probe)
                gb[o : o + cl[r]].copy_(pair_build)
            elif cl[r]:
                ops.append(dist.P2POp(dist.irecv, gp[o : o + cl[r]], r))
                ops.append(dist.P2POp(dist.irecv, gb[o : o + cl[r]], r))
            o += cl[r]
    else:
        gp = gb = None
        ops = []
        if cl[rank]:
            ops = [dist.P2POp(dist.isend, pair_probe, dst), dist.P2POp(dist.isend, pair_build, dst)]
    if ops:
        for w in dist.batch_isend_irecv(ops):
            w.wait()
    return gp, gb, cl


def sharded_join(n_probe, build_tree, probe_slice, src=0):
    """The multi-GPU join protocol with the device work injected:

    build_tree()                      -> (tree bytes uint8 tensor, exact boxes float64 tensor [m,4]) on rank `src`
    probe_slice(tree, boxes, b, e)    -> (pair_probe, pair_build) int32/int64 tensors for probe rows [b, e),
                                         row ids GLOBAL (index base b)
    Returns the gathered pairs on `src` (None elsewhere) and the per-rank counts."""
    rank, world = dist.get_rank(), dist.get_world_size()
    tree = boxes = None
    if rank == src:
        tree, boxes = build_tree()
    dev = tree.device if tree is not None else None
    if dev is None:
        # receivers learn the device from their probe function's first use: keep tensors where the backend wants them
        dev = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend() == "nccl" else torch.device("cpu")
    tree = broadcast_bytes(tree, tree.numel() if rank == src else 0, src, dev)
    bbytes = boxes.contiguous().view(torch.uint8).reshape(-1) if rank == src else None
    bbytes = broadcast_bytes(bbytes, bbytes.numel() if rank == src else 0, src, dev)
    boxes = bbytes.view(torch.float64).reshape(-1, 4)
    b, e = shard_range(n_probe, rank, world)
    pp, pb = probe_slice(tree, boxes, b, e)
    return gather_pairs(pp, pb, src)
