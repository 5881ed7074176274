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


# ---------------------------------------------------------------------------------------------------------------
# Multi-GPU CREATE INDEX: distributed sample sort of the ORDER BY + per-rank leaf level
# ---------------------------------------------------------------------------------------------------------------
def str_slice_size(n, cap):
    """Finalize, rtree_index_create_physical.cpp:349-351: ceil(sqrt(ceil(n / cap))) * cap (= sb_rtree_slice_for)"""
    import math

    return int(math.ceil(math.sqrt(float((n + cap - 1) // cap)))) * cap if n else 0


def str_x_sort_key(box4):
    """order-preserving 32-bit key of the STR x order, as int64: bits of float32(min_x + max_x) (fp32 add,
    ST_Centroid::ExecuteBox<float>, spatial_functions_scalar.cpp:1490-1494; the fp64 x 0.5 is monotone), -0 folded on +0"""
    f = (box4[:, 0] + box4[:, 2]) + 0.0
    r = f.contiguous().view(torch.int32).to(torch.int64) & 0xFFFFFFFF
    neg = r >= 0x80000000
    return torch.where(neg, 0xFFFFFFFF - r, r | 0x80000000)


def exact_splitters(sorted_keys, targets):
    """smallest key values K_j with exactly targets[j] keys (over ALL ranks) below them. sorted_keys: this rank's keys,
    ascending, distinct across ranks, in [0, 2^63). 63 rounds of bisection, one all-reduce of len(targets) counts each."""
    dev = sorted_keys.device
    t = torch.tensor(list(targets), dtype=torch.int64, device=dev)
    lo = torch.zeros_like(t)
    hi = torch.full_like(t, (1 << 63) - 1)
    for _ in range(63):
        mid = lo + (hi - lo) // 2
        cnt = torch.searchsorted(sorted_keys, mid, right=False).to(torch.int64)  # keys < mid on this rank
        dist.all_reduce(cnt)
        ge = cnt >= t
        hi = torch.where(ge, mid, hi)
        lo = torch.where(ge, lo, mid + 1)
    return lo


def gather_rows(t, dst=0):
    """gather variable-length first-dimension slices of a tensor to `dst` (rank order) -> tensor on dst, None elsewhere"""
    rank, world = dist.get_rank(), dist.get_world_size()
    mine = torch.tensor([t.shape[0]], dtype=torch.int64, device=t.device)
    counts = torch.zeros(world, dtype=torch.int64, device=t.device)
    dist.all_gather_into_tensor(counts, mine)
    cl = [int(c) for c in counts.tolist()]
    if rank == dst:
        out = torch.empty((sum(cl),) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
        ops, o = [], 0
        for r in range(world):
            if r == dst:
                out[o : o + cl[r]].copy_(t)
            elif cl[r]:
                ops.append(dist.P2POp(dist.irecv, out[o : o + cl[r]], r))
            o += cl[r]
    else:
        out, ops = None, ([dist.P2POp(dist.isend, t.contiguous(), dst)] if cl[rank] else [])
    if ops:
        for w in dist.batch_isend_irecv(ops):
            w.wait()
    return out


def sharded_rtree_build(box4, row_ids, cap, sort_keys, leaf_shard, assemble, dst=0):
    """CREATE INDEX over a (box, row id) column that is sharded over the ranks, identical -- node for node -- to the
    single-device STR bulk load of the concatenated column (rank order = input order).

    box4 [n_r, 4] float32, row_ids [n_r] int64: this rank's rows.
    sort_keys(keys int64) -> (sorted keys, permutation int64)        device sort of 63-bit keys
    leaf_shard(box4, row_ids, cap, slice, node_base, x_key_before, x_key_after)
                                              -> (leaf entries uint8 [n, 24], node entries uint8 [nodes, 24])
    assemble(leaf entries, n, node entries, cap) -> the tree (called on `dst` only)
    Steps: x keys -> globally unique 63-bit keys (x key, global input position) -> exact, slice-aligned splitters by
    bisection -> all-to-all of the rows -> rows back in global input order + the x keys at the range's two ends -> leaf
    level of the rank's slices -> gather -> upper levels."""
    rank, world = dist.get_rank(), dist.get_world_size()
    dev = box4.device
    n_r = box4.shape[0]
    counts = torch.zeros(world, dtype=torch.int64, device=dev)
    dist.all_gather_into_tensor(counts, torch.tensor([n_r], dtype=torch.int64, device=dev))
    cl = [int(c) for c in counts.tolist()]
    n, g0 = sum(cl), sum(cl[:rank])
    assert 0 < n < (1 << 30), "row count"
    slice_size = str_slice_size(n, cap)
    n_slices = (n + slice_size - 1) // slice_size
    chunk = ((n_slices + world - 1) // world) * slice_size  # rows per rank in the global x order: whole slices
    keys = (str_x_sort_key(box4) << 31) | (g0 + torch.arange(n_r, dtype=torch.int64, device=dev))
    skeys, perm = sort_keys(keys)
    targets = [min(r * chunk, n) for r in range(1, world)]
    cuts = torch.searchsorted(skeys, exact_splitters(skeys, targets), right=False).tolist() if world > 1 else []
    send = [b - a for a, b in zip([0] + cuts, cuts + [n_r])]
    recv_t = torch.zeros(world, dtype=torch.int64, device=dev)
    dist.all_to_all_single(recv_t, torch.tensor(send, dtype=torch.int64, device=dev))
    recv = [int(c) for c in recv_t.tolist()]
    m = sum(recv)

    def exchange(t):
        out = torch.empty((m,) + tuple(t.shape[1:]), dtype=t.dtype, device=dev)
        dist.all_to_all_single(out, t.contiguous(), output_split_sizes=recv, input_split_sizes=send)
        return out

    rkeys = exchange(skeys)
    rbox = exchange(box4[perm])
    rrow = exchange(row_ids[perm])
    # the leaf level wants its rows in ascending global input position (equal y keys of a slice keep that order) and the
    # x keys of the global order at the two ends of the range: the range's own smallest key and the next range's
    _, order = sort_keys(rkeys & ((1 << 31) - 1))
    lo = min(rank * chunk, n)
    assert m == min((rank + 1) * chunk, n) - lo, "exact splitters must give every rank its slice-aligned range"
    first = torch.full((world,), -1, dtype=torch.int64, device=dev)
    dist.all_gather_into_tensor(first, (rkeys.min() >> 31).reshape(1) if m else torch.full((1,), -1, dtype=torch.int64, device=dev))
    fl = [int(v) for v in first.tolist()]
    before = fl[rank] if lo > 0 else -1
    after = next((v for v in fl[rank + 1:] if v >= 0), -1)
    leaf, nodes = leaf_shard(rbox[order].contiguous(), rrow[order].contiguous(), cap, slice_size, lo // cap, before, after)
    leaf_all = gather_rows(leaf, dst)
    nodes_all = gather_rows(nodes, dst)
    if rank == dst:
        return assemble(leaf_all, n, nodes_all, cap)
    return None
