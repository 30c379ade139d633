import os, sys, time, threading, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, vgpu_b200
from vgpu_b200 import capi, datapath as dp
dev = torch.device("cuda", 0)
nv, per = 4, 8 << 30
bufs = [torch.empty(per, dtype=torch.uint8, device=dev) for _ in range(nv)]
parts = (capi.SmPartition * nv)()
assert capi.lib.vgpu_sm_partition_create(0, 25, nv, parts) == 0, capi.last_error()
streams = [p.stream for p in parts]
def one(i, n):
    torch.cuda.set_device(0)
    for s in range(n):
        for pat in (0, 1):
            t0 = time.perf_counter()
            r = dp.memtest_run(bufs[i], pat, s, 1, stream=streams[i])
            t1 = time.perf_counter()
            if s == n - 1 and i == 0:
                print(f"  vgpu{i} pat{pat}: wall {1e3*(t1-t0):.2f} ms, device fill {r.fill_ms:.2f} + verify {r.verify_ms:.2f} ms")
for nthreads in (1, 4):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    th = [threading.Thread(target=one, args=(i, 3)) for i in range(nthreads)]
    [t.start() for t in th]; [t.join() for t in th]
    torch.cuda.synchronize(); dt = time.perf_counter() - t0
    print(f"{nthreads} thread(s): {nthreads*3*2*2*per/dt/1e9:.0f} GB/s wall")
# same work launched asynchronously from one thread
torch.cuda.synchronize(); t0 = time.perf_counter()
res = [torch.zeros(3, dtype=torch.int64, device=dev) for _ in range(nv)]
for s in range(3):
    for i in range(nv):
        for pat in (0, 1):
            dp.memtest_fill(bufs[i], pat, s, stream=streams[i]); dp.memtest_verify(bufs[i], pat, s, result=res[i], stream=streams[i])
torch.cuda.synchronize(); dt = time.perf_counter() - t0
print(f"async 4 streams: {nv*3*2*2*per/dt/1e9:.0f} GB/s wall")
