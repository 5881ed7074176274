"""-m gpu, collected last: sb_pairs_checksum_noalloc (static __device__ scratch, no allocation) against
sb_pairs_checksum on the same device-resident pair lists, bit-exact (integer work), on calls in a row that share the
scratch; the size limit is enforced. Passes alone in a fresh process and after the rest of the suite."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_PAIRS = 1 << 26  # SB_NOALLOC_MAX_PAIRS, include/spatial_b200.h


def _noalloc(L, sb, tp, tb, n, lo, hi, out):
    import torch

    rc = L.sb_pairs_checksum_noalloc(0, C.c_void_p(torch.cuda.current_stream().cuda_stream), C.c_void_p(tp.data_ptr()),
                                     C.c_void_p(tb.data_ptr()), n, lo, hi, C.c_void_p(out.data_ptr()))
    sb.capi.check(rc)
    w = out.cpu().numpy().view(np.uint64)  # .cpu() synchronises the stream the kernels ran on
    return int(w[0]), int(w[1])


def test_noalloc_checksum_matches_existing_entry_point(sb, O):
    import torch

    L = sb.capi.lib()
    dev = torch.device("cuda", 0)
    ctx = sb.Context(0)
    try:
        rng = np.random.default_rng(7)
        out = torch.full((2,), -1, dtype=torch.int64, device=dev)  # the caller's output, allocated before the calls
        # sizes: several tiles + ragged tail, then a SHORTER list (stale partials of the first call stay in the
        # scratch behind it), one exact tile, one pair, none, and a longer one again
        for call, n in enumerate((1_000_003, 4_097, 4_096, 1, 0, 3_000_000)):
            p = rng.integers(0, 2**32, max(n, 1), dtype=np.uint64).astype(np.uint32)
            b = rng.integers(0, 2**32, max(n, 1), dtype=np.uint64).astype(np.uint32)
            tp, tb = torch.from_numpy(p.view(np.int32)).to(dev), torch.from_numpy(b.view(np.int32)).to(dev)
            for lo, hi in ((0, 0xFFFFFFFF), (1 << 30, 3 << 30)):
                want = ctx.pairs_checksum(tp, tb, lo, hi, count=n)
                out.fill_(-1)
                got = _noalloc(L, sb, tp, tb, n, lo, hi, out)
                assert got == want, (call, n, lo, hi)
                # twice in a row on the same inputs: the scratch keeps the last call's contents and must not leak in
                assert _noalloc(L, sb, tp, tb, n, lo, hi, out) == want, (call, n, lo, hi)
                sel = (p[:n] >= lo) & (p[:n] < hi)
                assert want == O.pair_checksum(p[:n][sel], b[:n][sel])
    finally:
        ctx.close()


def test_noalloc_does_not_allocate_and_rejects_oversize(sb):
    import torch

    L = sb.capi.lib()
    dev = torch.device("cuda", 0)
    n = 1 << 20
    tp = torch.arange(n, dtype=torch.int32, device=dev)
    tb = torch.arange(n, dtype=torch.int32, device=dev) * 3
    out = torch.zeros(2, dtype=torch.int64, device=dev)
    _noalloc(L, sb, tp, tb, n, 0, 0xFFFFFFFF, out)  # set-up (symbol address) done
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info(0)
    for _ in range(4):
        first = _noalloc(L, sb, tp, tb, n, 0, 0xFFFFFFFF, out)
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info(0)
    assert free1 >= free0  # nothing was taken from the device
    assert first[0] == n
    # beyond the static scratch: a clear error before anything is launched or dereferenced
    rc = L.sb_pairs_checksum_noalloc(0, None, C.c_void_p(tp.data_ptr()), C.c_void_p(tb.data_ptr()), MAX_PAIRS + 1, 0,
                                     0xFFFFFFFF, C.c_void_p(out.data_ptr()))
    assert rc == -2
    with pytest.raises(sb.SpatialB200Error) as e:
        sb.capi.check(rc)
    assert "SB_NOALLOC_MAX_PAIRS" in str(e.value)
    # the largest accepted size runs
    big = 1 << 26
    bp = torch.zeros(big, dtype=torch.int32, device=dev)
    bb = torch.zeros(big, dtype=torch.int32, device=dev)
    c, h = _noalloc(L, sb, bp, bb, big, 0, 1, out)
    assert c == big and h == 0  # mix(0, 0) == 0
    # and a small call right after it sees none of its partials
    assert _noalloc(L, sb, tp, tb, 5, 0, 0xFFFFFFFF, out)[0] == 5
