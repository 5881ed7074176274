"""GPU diagnostic: C1 batched range scans -- event time of the call vs the time of its kernels (profile scope)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, sb200
import bench_extras as X
sb = sb200.load()
ctx = sb.Context(0); ctx.use_torch_stream()
dev = torch.device("cuda", 0); stream = torch.cuda.current_stream()
n = 10_000_000
ix, iy = ctx.st_generate_points(1337, n, (0.0, 0.0, 10000.0, 10000.0), device=True)
rows = torch.arange(n, dtype=torch.int64, device=dev)
idx = sb.RTreeIndex(ctx, ctx.bbox_from_points(ix, iy), rows, 128)
qb = X._range_queries(ctx)
for cap in (1 << 26, 10_000_000):
    ms, (oq, orow) = X.timed(stream, lambda: idx.scan(qb, capacity=cap), 5)
    ctx.profile(True)
    for _ in range(5):
        idx.scan(qb, capacity=cap)
    torch.cuda.synchronize()
    t, k = ctx.profile_read("rtree_scan")
    ctx.profile(False)
    print({"cap": cap, "call_ms": round(ms, 4), "kernels_ms": round(t / 5, 4), "launches": k / 5, "rows": len(orow)}, flush=True)
