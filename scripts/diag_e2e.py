"""Timeline of the host-pointer join pipeline (option trace=2) at the headline sizes; run on the GPU box:
   python scripts/diag_e2e.py [pipe_chunk ...] 2> gpurun_out/e2e_timeline.txt"""
import ctypes as C
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sb200  # noqa: E402

sb = sb200.load()
capi = sb.capi

P, B = 200_000_000, 2_000_000
DOMAIN = (0.0, 0.0, 10000.0, 10000.0)
ctx = sb.Context()
L = ctx.L
px, py = ctx.st_generate_points(1337, P, DOMAIN, device=True)
cx, cy = ctx.st_generate_points(42, B, DOMAIN, device=True)
bx0, by0, bx1, by1 = cx - 10.0, cy - 10.0, cx + 10.0, cy + 10.0
hp = [t.cpu().pin_memory() for t in (px, py, bx0, by0, bx1, by1)]
del px, py
cap = int(P * 8 * 1.15) + (1 << 20)
ho_p = torch.empty(cap, dtype=torch.int32).pin_memory()
ho_b = torch.empty(cap, dtype=torch.int32).pin_memory()
cnt = C.c_uint64(0)


def step():
    capi.check(L.sb_join_points_boxes(ctx.h, capi.HOST, capi.PRED_INTERSECTS, hp[0].data_ptr(), hp[1].data_ptr(), P,
                                      hp[2].data_ptr(), hp[3].data_ptr(), hp[4].data_ptr(), hp[5].data_ptr(), B,
                                      ho_p.data_ptr(), ho_b.data_ptr(), cap, C.byref(cnt)))
    return cnt.value


for chunk in [int(x) for x in sys.argv[1:]] or [16 << 20]:
    ctx.set_option("pipe_chunk", chunk)
    ctx.set_option("trace", 0)
    step()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    n = step()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    print("chunk %d: %.2f ms, %d pairs" % (chunk, dt * 1e3, n), flush=True)
    sys.stderr.write("==== chunk %d (untraced %.2f ms)\n" % (chunk, dt * 1e3))
    sys.stderr.flush()
    ctx.set_option("trace", 2)
    step()
    ctx.set_option("trace", 0)

# raw copy-out rate into the same pinned buffers, for comparison with the D lines of the timeline
d = torch.empty(1 << 27, dtype=torch.int32, device="cuda")
for name, h in (("ho_p", ho_p), ("ho_b", ho_b)):
    for rep in range(2):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for k in range(8):
            h[k << 27:(k + 1) << 27].copy_(d, non_blocking=True)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
    print("raw d2h into %s: 8 x 512 MiB at %.1f GB/s" % (name, 8 * 4 * (1 << 27) / dt / 1e9), flush=True)
hx = hp[0]
dd = torch.empty(1 << 26, dtype=torch.float64, device="cuda")
torch.cuda.synchronize()
t0 = time.perf_counter()
dd.copy_(hx[:1 << 26], non_blocking=True)
torch.cuda.synchronize()
print("raw h2d from px: 512 MiB at %.1f GB/s" % (8 * (1 << 26) / (time.perf_counter() - t0) / 1e9), flush=True)
os.system("nvidia-smi topo -m; cat /sys/devices/system/node/node*/cpulist; grep -i 'allowed_list' /proc/self/status; "
          "cat /sys/devices/system/node/node*/meminfo | grep -i 'MemFree'")
for off_h, off_d in ((0, 0), (1, 0), (1, 1), (33, 1), (1056, 0), (1024, 0)):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    n = (1 << 27) - 4096
    for k in range(8):
        ho_p[(k << 27) + off_h:(k << 27) + off_h + n].copy_(d[off_d:off_d + n], non_blocking=True)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    print("raw d2h, host offset %d words, device offset %d words: %.1f GB/s" % (off_h, off_d, 8 * 4 * n / dt / 1e9), flush=True)
