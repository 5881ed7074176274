"""GPU diagnostic: wall-clock of each phase of the C3 join (device-resident inputs)."""
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
cap = int(P * 8 * 1.15) + (1 << 20)
op = torch.empty(cap, dtype=torch.int32, device="cuda"); ob = torch.empty_like(op)
cnt = C.c_uint64(0)
def T(name, f, n=3):
    for i in range(n):
        torch.cuda.synchronize(); t0 = time.perf_counter(); r = f(); torch.cuda.synchronize()
        print("%-28s %9.3f ms" % (name, 1e3 * (time.perf_counter() - t0)), flush=True)
    return r
T("join(device)", lambda: capi.check(L.sb_join_points_boxes(ctx.h, capi.DEVICE, 0, px.data_ptr(), py.data_ptr(), P, *[v.data_ptr() for v in b], B, op.data_ptr(), ob.data_ptr(), cap, C.byref(cnt))))
print("pairs", cnt.value)
fb = T("bbox_from_box2d", lambda: ctx.bbox_from_box2d(*b))
tree = T("FlatRTree build", lambda: sb.FlatRTree(ctx, fb, 32))
box2d = torch.stack(b, 1).contiguous()
T("probe_points fused", lambda: tree.probe_points(px, py, box2d, out=(op, ob)))
T("probe_points cand", lambda: tree.probe_points(px, py, None, out=(op, ob)))
T("memset out 2x", lambda: (op.zero_(), ob.zero_()))
T("copy px->tmp", lambda: px.clone())
