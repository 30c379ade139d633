"""Round 2: where do small spill batches spend their time?
Per batch size and variant: the time of ONE call between two CUDA events on an idle GPU (what bench.py's sweep
reports), the steady-state time per call when 200 calls are queued back to back (launch overhead hidden), and the
same for a 32-byte "transfer" (pure fixed cost of the path: host code + launch + fold)."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch

import vgpu_b200  # noqa: F401
from vgpu_b200 import capi, datapath as dp

torch.cuda.set_device(0)
dev = torch.device("cuda", 0)
cb = capi.CHUNK_BYTES
nmax = 64
hbm = torch.empty(nmax * cb, dtype=torch.uint8, device=dev)
dst = torch.empty_like(hbm)
dp.memtest_fill(hbm, capi.PAT_RANDOM, 1)
ck = torch.zeros((nmax, 2), dtype=torch.int64, device=dev)
stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
tiers = {"local": (capi.TIER_LOCAL, dst.data_ptr())}
if torch.cuda.device_count() > 1:
    ps = C.c_void_p()
    capi.check(capi.lib.vgpu_slab_alloc(capi.TIER_PEER, 0, 1, nmax * cb, C.byref(ps)))
    tiers["peer"] = (capi.TIER_PEER, ps.value)


def single(fn, reps=30):
    best = 1e9
    for _ in range(reps):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best * 1e3


def steady(fn, k=200):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(k):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / k


for tname, (tier, slab) in tiers.items():
    for vname, v in (("ldst", capi.XFER_LDST), ("tma", capi.XFER_TMA), ("auto", capi.XFER_AUTO)):
        tiny = lambda: capi.lib.vgpu_spill(v, tier, hbm.data_ptr(), slab, None, None, 1, 32, ck.data_ptr(), stream)
        tiny()
        print(f"{tname} {vname}: fixed cost (1 x 32 B): single {single(tiny):6.2f} us, steady {steady(tiny):6.2f} us", flush=True)
        for n in (1, 2, 4, 8, 16, 64):
            f = lambda: capi.lib.vgpu_spill(v, tier, hbm.data_ptr(), slab, None, None, n, cb, ck.data_ptr(), stream)
            f()
            s, t = single(f), steady(f)
            print(f"{tname} {vname} {n:3d} chunks: single {s:7.2f} us = {n * cb / s / 1e3:7.1f} GB/s | steady {t:7.2f} us = {n * cb / t / 1e3:7.1f} GB/s", flush=True)
