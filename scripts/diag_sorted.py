"""GPU experiment: how fast is k_bp_probe when the probes arrive sorted by (bucket, sub-cell)?"""
import ctypes as C, sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, sb200
sb = sb200.load(); capi = sb.capi
P = int(sys.argv[1]) if len(sys.argv) > 1 else 200_000_000
B = 2_000_000
ctx = sb.Context(0); L = ctx.L
D = (0.0, 0.0, 10000.0, 10000.0)
px, py = ctx.st_generate_points(1337, P, D, device=True)
cx, cy = ctx.st_generate_points(42, B, D, device=True)
b = [cx - 10, cy - 10, cx + 10, cy + 10]
minx, miny = float(b[0].min()), float(b[1].min()); w = float(b[2].max()) - minx; h = float(b[3].max()) - miny
def sort_pts(px, py):
    qx = ((px - minx) * (1048576.0 / w)).floor().clamp_(0, 1048575).to(torch.int64)
    qy = ((py - miny) * (1048576.0 / h)).floor().clamp_(0, 1048575).to(torch.int64)
    key = (((qy >> 15) * 32 + (qx >> 15)) << 10) + (((qy >> 10) & 31) * 32 + ((qx >> 10) & 31))
    del qx, qy
    order = torch.argsort(key); del key
    return px[order].contiguous(), py[order].contiguous()
cap = int(P * 8 * 1.15) + (1 << 20)
op = torch.empty(cap, dtype=torch.int32, device="cuda"); ob = torch.empty_like(op)
cnt = C.c_uint64(0)
def join(x, y):
    capi.check(L.sb_join_points_boxes(ctx.h, capi.DEVICE, 0, x.data_ptr(), y.data_ptr(), P, *[v.data_ptr() for v in b], B, op.data_ptr(), ob.data_ptr(), cap, C.byref(cnt)))
for name, (x, y) in (("random", (px, py)), ("sorted", sort_pts(px, py))):
    for i in range(3): join(x, y)
    ctx.set_option("trace", 1)
    print("----", name, flush=True)
    join(x, y)
    ctx.set_option("trace", 0)
    print("pairs", cnt.value, flush=True)
