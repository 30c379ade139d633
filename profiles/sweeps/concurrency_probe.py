"""Round 2 probe: what does work in the PRIMARY context cost memtest kernels running in green-context partitions?
Cases: memtest alone; + copy-engine D2H/H2D on a plain stream; + checksum kernel on a plain stream; + LD/ST spill
kernel to pinned host on a plain stream; the same transfers issued on a green-context stream of a spare partition."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch

import vgpu_b200  # noqa: F401
from vgpu_b200 import capi, datapath as dp

dev = torch.device("cuda", 0)
torch.cuda.set_device(0)
cb = capi.CHUNK_BYTES
nv, pct, fine = int(sys.argv[1]) if len(sys.argv) > 1 else 10, int(sys.argv[2]) if len(sys.argv) > 2 else 9, 1
GIB = 1 << 30
per = 4 * GIB
# ask for one partition more than the vGPUs: the spare one hosts the transfer stream in the "green" cases
parts = (capi.SmPartition * (nv + 1))()
capi.check(capi.lib.vgpu_sm_partition_create_ex(0, pct, nv + 1, fine, parts))
print("SMs per partition", [p.sm_count for p in parts], flush=True)
streams = [torch.cuda.ExternalStream(p.stream, device=dev) for p in parts[:nv]]
spare = torch.cuda.ExternalStream(parts[nv].stream, device=dev)
plain = torch.cuda.Stream(device=dev)
bufs = [torch.empty(per, dtype=torch.uint8, device=dev) for _ in range(nv)]
results = [torch.zeros(3, dtype=torch.int64, device=dev) for _ in range(nv)]
n = 512
hbm = torch.empty(n * cb, dtype=torch.uint8, device=dev)
dp.memtest_fill(hbm, capi.PAT_RANDOM, 5)
host = torch.empty(n * cb, dtype=torch.uint8).pin_memory()
ck = torch.zeros((n, 2), dtype=torch.int64, device=dev)
big = torch.empty(16 * GIB, dtype=torch.uint8, device=dev)
ckbig = torch.zeros((16 * GIB // cb, 2), dtype=torch.int64, device=dev)
main = torch.cuda.current_stream()


def xfer(kind, st):
    s = C.c_void_p(st.cuda_stream)
    if kind == "ce":
        with torch.cuda.stream(st):
            for _ in range(4):
                host.copy_(hbm, non_blocking=True)
                hbm.copy_(host, non_blocking=True)
    elif kind == "cksum":  # a whole-GPU read-only kernel: 4 x 16 GiB
        for _ in range(4):
            capi.check(capi.lib.vgpu_checksum(big.data_ptr(), None, 16 * GIB // cb, cb, ckbig.data_ptr(), s))
    elif kind == "ldst":
        for _ in range(4):
            capi.check(capi.lib.vgpu_spill(capi.XFER_LDST, capi.TIER_HOST, hbm.data_ptr(), host.data_ptr(), None, None, n, cb, ck.data_ptr(), s))
    elif kind == "tma":
        for _ in range(4):
            capi.check(capi.lib.vgpu_spill(capi.XFER_TMA, capi.TIER_HOST, hbm.data_ptr(), host.data_ptr(), None, None, n, cb, ck.data_ptr(), s))


def run(kind, st, do_mt=True):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    x0, x1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    m1 = torch.cuda.Event(enable_timing=True)
    e0.record(main)
    ends = []
    if do_mt:
        for i, s in enumerate(streams):
            s.wait_event(e0)
            for p in range(2):
                for pat in (0, 1):
                    dp.memtest_fill(bufs[i], pat, p, stream=s)
                    dp.memtest_verify(bufs[i], pat, p, result=results[i], stream=s)
            e = torch.cuda.Event(); e.record(s); ends.append(e)
    if kind:
        st.wait_event(e0)
        x0.record(st)
        xfer(kind, st)
        x1.record(st)
        ends.append(x1)
    for e in ends[: nv if do_mt else 0]:
        main.wait_event(e)
    m1.record(main)
    for e in ends:
        main.wait_event(e)
    e1.record(main)
    torch.cuda.synchronize()
    mt_bytes = nv * per * 8
    mt_ms = e0.elapsed_time(m1)
    return (mt_bytes / (mt_ms * 1e6) if do_mt else 0.0), (x0.elapsed_time(x1) if kind else 0.0)


run(None, plain)
print(f"memtest alone: {run(None, plain)[0]:.0f} GB/s", flush=True)
for kind in ("ce", "cksum", "ldst", "tma"):
    for where, st in (("plain stream (primary ctx)", plain), ("spare green partition", spare)):
        alone = run(kind, st, do_mt=False)[1]
        mt, xms = run(kind, st)
        print(f"{kind:6s} on {where:28s}: transfer alone {alone:8.2f} ms | with memtest: memtest {mt:6.0f} GB/s, transfer {xms:8.2f} ms", flush=True)
