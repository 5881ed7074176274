"""GPU diagnostic: per-phase wall time of the C3 join with the library's trace option."""
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
def join():
    capi.check(L.sb_join_points_boxes(ctx.h, capi.DEVICE, 0, px.data_ptr(), py.data_ptr(), P, *[v.data_ptr() for v in b], B, op.data_ptr(), ob.data_ptr(), cap, C.byref(cnt)))
for i in range(3):
    join()
for i in range(3):
    torch.cuda.synchronize(); t0 = time.perf_counter(); join(); torch.cuda.synchronize()
    print("join wall %.3f ms" % (1e3 * (time.perf_counter() - t0)), flush=True)
ctx.set_option("trace", 1)
for i in range(2):
    print("---- traced join", i, flush=True)
    join()
ctx.set_option("trace", 0)
def timed(label, n=4):
    for i in range(n):
        torch.cuda.synchronize(); t0 = time.perf_counter(); join(); torch.cuda.synchronize()
        print("%s: join wall %.3f ms" % (label, 1e3 * (time.perf_counter() - t0)), flush=True)
timed("plain")
ctx.profile(True)
timed("profile on")
print(ctx.profile_read("flat_probe"), ctx.profile_read("bp_partition"))
ctx.profile(False)
timed("profile off")
s = torch.cuda.Stream()
with torch.cuda.stream(s):
    ctx.use_torch_stream()
    timed("torch stream")
    ctx.profile(True)
    timed("torch stream + profile")
    ctx.profile(False)
