import torch, time
d=torch.empty(1<<30, dtype=torch.int32, device='cuda')   # 4 GB
h=torch.empty(1<<30, dtype=torch.int32).pin_memory()
for name,(a,b) in (("d2h",(h,d)),("h2d",(d,h))):
    for _ in range(2):
        torch.cuda.synchronize(); t=time.perf_counter(); a.copy_(b, non_blocking=True); torch.cuda.synchronize(); dt=time.perf_counter()-t
    print(name, "%.1f GB/s"%(4.294967296/dt))
# duplex
s1,s2=torch.cuda.Stream(),torch.cuda.Stream()
d2=torch.empty(1<<28, dtype=torch.int32, device='cuda'); h2=torch.empty(1<<28, dtype=torch.int32).pin_memory()
torch.cuda.synchronize(); t=time.perf_counter()
with torch.cuda.stream(s1): h.copy_(d, non_blocking=True)
with torch.cuda.stream(s2):
    for _ in range(4): d2.copy_(h2, non_blocking=True)
torch.cuda.synchronize(); dt=time.perf_counter()-t
print("duplex: d2h 4.29 GB + h2d 4.29 GB in %.3f s -> d2h-equivalent %.1f GB/s"%(dt, 4.294967296/dt))
