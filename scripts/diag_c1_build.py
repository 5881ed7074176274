"""GPU diagnostic: C1 CREATE INDEX (10 M points, cap 128) build time and its profile scopes; checks the level counts and
the 3986-row scan of benchmark/rtree_points_index.benchmark. SPATIAL_B200_LIB selects a variant build."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import sb200

sb = sb200.load()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
cap = int(sys.argv[2]) if len(sys.argv) > 2 else 128
ctx = sb.Context(0)
ctx.use_torch_stream()
for kv in sys.argv[3:]:
    k, v = kv.split("=")
    ctx.set_option(k, int(v))
dev = torch.device("cuda", 0)
stream = torch.cuda.current_stream()
ix, iy = ctx.st_generate_points(1337, n, (0.0, 0.0, 10000.0, 10000.0), device=True)
rows = torch.arange(n, dtype=torch.int64, device=dev)


FUSED = os.environ.get("C1_FROM_POINTS", "1") == "1"


def build():
    if FUSED:
        return sb.RTreeIndex.from_points(ctx, ix, iy, rows, cap)
    return sb.RTreeIndex(ctx, ctx.bbox_from_points(ix, iy), rows, cap)


for _ in range(3):
    build().close()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
reps = 10
e0.record(stream)
for _ in range(reps):
    build().close()
e1.record(stream)
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / reps
ctx.profile(True)
for _ in range(reps):
    build().close()
torch.cuda.synchronize()
scopes = {}
for tag in ("bbox", "str_keys", "str_select", "str_partition", "radix_hist", "radix_onesweep", "str_pack"):
    t, k = ctx.profile_read(tag)
    scopes[tag] = round(t / reps, 4)
ctx.profile(False)
idx = build()
levels = sorted((int(c) for c in idx.level_counts()), reverse=True)
q = ctx.bbox_from_box2d(*[torch.tensor([v], dtype=torch.float64, device=dev) for v in (450.0, 450.0, 650.0, 650.0)])
_, got = idx.scan(q)
brute = int(((ix >= 450) & (ix <= 650) & (iy >= 450) & (iy <= 650)).sum().item())
inside = int(((ix[got.long()] >= 450) & (ix[got.long()] <= 650) & (iy[got.long()] >= 450) & (iy[got.long()] <= 650)).sum().item())
assert inside == brute, (inside, brute)  # the scan returns float-box candidates: every true hit must be among them
print({"lib": os.environ.get("SPATIAL_B200_LIB", "default"), "n": n, "ms": round(ms, 4), "scopes": scopes, "levels": levels,
       "scan": int(len(got)), "brute_force_hits": brute}, flush=True)
