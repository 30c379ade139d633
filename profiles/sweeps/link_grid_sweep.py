"""Round 2: how many CTAs does a link-bound transfer need?  For the host tier (and the peer tier when a second GPU
is visible): spill / restore GB/s of the LD/ST and bulk-copy kernels at 512 chunks for a range of grid sizes
(vgpu_policy_set_grid), alone and while ten SM-granular memtest partitions are busy (whose rate is printed too).
Run with CUDA_DEVICE_MAX_CONNECTIONS=32."""
import ctypes as C
import os
import sys

os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch

import vgpu_b200  # noqa: F401
from vgpu_b200 import capi, datapath as dp

dev = torch.device("cuda", 0)
torch.cuda.set_device(0)
cb = capi.CHUNK_BYTES
GIB = 1 << 30
nv, per = 10, 4 * GIB
parts = (capi.SmPartition * nv)()
capi.check(capi.lib.vgpu_sm_partition_create_ex(0, 10, nv, 1, parts))
streams = [torch.cuda.ExternalStream(p.stream, device=dev) for p in parts]
plain = torch.cuda.Stream(device=dev)
bufs = [torch.empty(per, dtype=torch.uint8, device=dev) for _ in range(nv)]
results = [torch.zeros(3, dtype=torch.int64, device=dev) for _ in range(nv)]
n = 512
hbm = torch.empty(n * cb, dtype=torch.uint8, device=dev)
back = torch.empty_like(hbm)
dp.memtest_fill(hbm, capi.PAT_RANDOM, 5)
ck = torch.zeros((n, 2), dtype=torch.int64, device=dev)
res = torch.zeros(2, dtype=torch.int64, device=dev)
main = torch.cuda.current_stream()
tiers = {}
host = C.c_void_p()
capi.check(capi.lib.vgpu_slab_alloc(capi.TIER_HOST, 0, -1, n * cb, C.byref(host)))
tiers["host"] = (capi.TIER_HOST, host.value)
if torch.cuda.device_count() > 1:
    ps = C.c_void_p()
    capi.check(capi.lib.vgpu_slab_alloc(capi.TIER_PEER, 0, 1, n * cb, C.byref(ps)))
    tiers["peer"] = (capi.TIER_PEER, ps.value)


def xfer(tier, slab, v, direction, reps):
    s = C.c_void_p(plain.cuda_stream)
    for _ in range(reps):
        if direction == "spill":
            capi.check(capi.lib.vgpu_spill(v, tier, hbm.data_ptr(), slab, None, None, n, cb, ck.data_ptr(), s))
        else:
            capi.check(capi.lib.vgpu_restore(v, tier, slab, back.data_ptr(), None, None, n, cb, ck.data_ptr(), None, res.data_ptr(), s))


def run(tier, slab, v, direction, reps, do_mt):
    torch.cuda.synchronize()
    e0, m1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    x0, x1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(main)
    if do_mt:
        for s in streams:
            s.wait_event(e0)
        for pat in (0, 1):
            for i, s in enumerate(streams):
                dp.memtest_fill(bufs[i], pat, 1, stream=s)
                dp.memtest_verify(bufs[i], pat, 1, result=results[i], stream=s)
        for s in streams:
            e = torch.cuda.Event(); e.record(s); main.wait_event(e)
        m1.record(main)
    plain.wait_event(e0)
    x0.record(plain)
    xfer(tier, slab, v, direction, reps)
    x1.record(plain)
    torch.cuda.synchronize()
    mt = nv * per * 4 / (e0.elapsed_time(m1) * 1e6) if do_mt else 0.0
    return reps * n * cb / (x0.elapsed_time(x1) * 1e6), mt


print("memtest alone:", f"{run(capi.TIER_HOST, host.value, capi.XFER_CE, 'spill', 1, True)[1]:.0f} GB/s (with one CE spill beside it)", flush=True)
for tname, (tier, slab) in tiers.items():
    reps = 2 if tname == "host" else 24
    for vname, v in (("ldst", capi.XFER_LDST), ("tma", capi.XFER_TMA)):
        for grid in (4, 8, 16, 32, 64, 148, 296, 592, 1184, 4736):
            capi.check(capi.lib.vgpu_policy_set_grid(tier, grid))
            row = []
            for direction in ("spill", "restore"):
                run(tier, slab, v, direction, 1, False)
                alone = max(run(tier, slab, v, direction, reps, False)[0] for _ in range(2))
                busy, mt = run(tier, slab, v, direction, reps, True)
                row.append(f"{direction} alone {alone:7.1f} | beside memtest {busy:7.1f} (memtest {mt:5.0f})")
            print(f"{tname} {vname:4s} grid {grid:5d}: " + " || ".join(row), flush=True)
        capi.check(capi.lib.vgpu_policy_set_grid(tier, 0))
assert res.cpu().tolist() == [0, 0] and torch.equal(back, hbm)
print("restores bit-exact")
