"""Phase breakdown of the C4 polygon join (CUDA-event profile tags of the context)."""
import json, math, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "scripts"))
import torch
import sb200
from bench_configs import star_polygons, DOMAIN

P = int(sys.argv[1]) if len(sys.argv) > 1 else 200_000_000
sb = sb200.load()
ctx = sb.Context(0)
dev = torch.device("cuda", 0)
stream = torch.cuda.Stream(device=dev)
with torch.cuda.stream(stream):
    ctx.use_torch_stream()
    vx, vy, voff = star_polygons(ctx, torch, 99, 20000, 500, 20.0, 60.0, dev)
    poff = torch.arange(20001, dtype=torch.int64, device=dev)
    px, py = ctx.st_generate_points(1337, P, DOMAIN, device=True)
    cap = int(P * 1.3) + (1 << 20)
    out = (torch.empty(cap, dtype=torch.int32, device=dev), torch.empty(cap, dtype=torch.int32, device=dev))
    ctx.join_points_polygons(px, py, vx.reshape(-1), vy.reshape(-1), voff, poff, out=out)
    ctx.profile(True)
    for _ in range(2):
        gp, gb = ctx.join_points_polygons(px, py, vx.reshape(-1), vy.reshape(-1), voff, poff, out=out)
    torch.cuda.synchronize()
    res = {"points": P, "pairs": int(len(gp))}
    for tag in ("pip_prepare", "bp_partition", "bp_cand", "pip_gather", "pip_join"):
        ms, n = ctx.profile_read(tag)
        res[tag + "_ms"] = ms / 2
    print(json.dumps(res))
