"""Round 2 probe: does a large copy-engine transfer issued by one host thread delay kernel launches of another?
Thread A launches a tiny memtest fill in a loop on its own stream and records how long each launch CALL takes on the
host; thread B runs the blocking host-buffer spill (1024 x 2 MiB) with the copy-engine or the LD/ST variant."""
import ctypes as C
import os
import sys
import threading
import time

os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch

import vgpu_b200  # noqa: F401
from vgpu_b200 import capi, datapath as dp

torch.cuda.set_device(0)
dev = torch.device("cuda", 0)
cb = capi.CHUNK_BYTES
n = 1024
hbm = torch.empty(n * cb, dtype=torch.uint8, device=dev)
dp.memtest_fill(hbm, capi.PAT_RANDOM, 1)
host = torch.empty(n * cb, dtype=torch.uint8).pin_memory()
hck = (capi.Cksum * n)()
small = torch.empty(1 << 20, dtype=torch.uint8, device=dev)
sA = torch.cuda.Stream(device=dev)
stop = False
lat = []


def launcher():
    torch.cuda.set_device(0)
    s = C.c_void_p(sA.cuda_stream)
    while not stop:
        t = time.perf_counter()
        capi.lib.vgpu_memtest_fill(small.data_ptr(), 1 << 20, 0, 0, 0, s)
        lat.append(time.perf_counter() - t)
        if len(lat) % 64 == 0:
            sA.synchronize()


for vname, v in (("none", None), ("ldst", capi.XFER_LDST), ("ce", capi.XFER_CE)):
    stop = False
    lat.clear()
    th = threading.Thread(target=launcher)
    th.start()
    time.sleep(0.05)
    t0 = time.perf_counter()
    if v is not None:
        for _ in range(4):
            capi.check(capi.lib.vgpu_spill_to_host(v, 0, hbm.data_ptr(), host.data_ptr(), n, cb, hck))
    else:
        time.sleep(0.2)
    dt = time.perf_counter() - t0
    stop = True
    th.join()
    ls = sorted(lat)
    print(f"{vname:5s}: {len(ls)} launches in {dt * 1e3:.0f} ms; launch call p50 {ls[len(ls) // 2] * 1e6:.1f} us, p99 {ls[int(len(ls) * 0.99)] * 1e6:.1f} us, "
          f"max {ls[-1] * 1e6:.0f} us, >1 ms: {sum(1 for x in ls if x > 1e-3)}", flush=True)
