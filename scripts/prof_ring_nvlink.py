"""Where does the peer-tier restore lose bandwidth when the reverse direction is busy?  (VERDICT r01, next #9.)

ncu serialises kernels, so the ring cannot be watched with it.  This script drives BOTH GPUs from one process —
GPU 0's tier slab lives in GPU 1's HBM and GPU 1's in GPU 0's, as in the ring — and reads the NVLink byte counters
NVML keeps per GPU (NVML_FI_DEV_NVLINK_THROUGHPUT_{DATA,RAW}_{TX,RX}, KiB, summed over links) around long runs of:
restore on GPU 0 alone, restore on both at once, spill on GPU 0 alone, spill on both at once.  RAW counts what the
link carried including packet overhead; the rates it prints are per GPU.  Needs 2 GPUs.

    python scripts/prof_ring_nvlink.py > profiles/r02_ring_nvlink.txt
"""
import ctypes as C
import os
import sys
import time

os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
import pynvml
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import vgpu_b200  # noqa: E402,F401
from vgpu_b200 import capi, datapath as dp  # noqa: E402

n, cb = 2048, capi.CHUNK_BYTES
pynvml.nvmlInit()
G = []
for g in (0, 1):
    dev = torch.device("cuda", g)
    with torch.cuda.device(g):
        hbm = torch.empty(n * cb, dtype=torch.uint8, device=dev)
        back = torch.zeros_like(hbm)
        dp.memtest_fill(hbm, capi.PAT_RANDOM, 2 + g)
        slab = C.c_void_p()
        capi.check(capi.lib.vgpu_slab_alloc(capi.TIER_PEER, g, 1 - g, n * cb, C.byref(slab)))
        ck = torch.zeros((n, 2), dtype=torch.int64, device=dev)
        res = torch.zeros(2, dtype=torch.int64, device=dev)
        st = torch.cuda.Stream(device=dev)
        uuid = torch.cuda.get_device_properties(g).uuid
        h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + str(uuid)).encode())
        G.append(dict(g=g, hbm=hbm, back=back, slab=slab, ck=ck, res=res, st=st, h=h))

FIELDS = [("data_tx", pynvml.NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_TX), ("data_rx", pynvml.NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_RX),
          ("raw_tx", pynvml.NVML_FI_DEV_NVLINK_THROUGHPUT_RAW_TX), ("raw_rx", pynvml.NVML_FI_DEV_NVLINK_THROUGHPUT_RAW_RX)]


def counters(h):
    """Sum over the GPU's links; the aggregate scope (UINT_MAX) is tried first, then link by link."""
    out = {}
    for name, f in FIELDS:
        v = pynvml.nvmlDeviceGetFieldValues(h, [(f, 0xFFFFFFFF)])[0]
        if v.nvmlReturn == 0:
            out[name] = int(v.value.ullVal) * 1024  # KiB -> bytes
            continue
        tot, ok = 0, False
        for link in range(18):
            w = pynvml.nvmlDeviceGetFieldValues(h, [(f, link)])[0]
            if w.nvmlReturn == 0:
                tot += int(w.value.ullVal) * 1024
                ok = True
        out[name] = tot if ok else None
    return out


def launch(x, what, variant, reps):
    s = C.c_void_p(x["st"].cuda_stream)
    with torch.cuda.device(x["g"]):
        for _ in range(reps):
            if what == "spill":
                capi.check(capi.lib.vgpu_spill(variant, capi.TIER_PEER, x["hbm"].data_ptr(), x["slab"], None, None, n, cb, x["ck"].data_ptr(), s))
            else:
                capi.check(capi.lib.vgpu_restore(variant, capi.TIER_PEER, x["slab"], x["back"].data_ptr(), None, None, n, cb, x["ck"].data_ptr(), None, x["res"].data_ptr(), s))


def run(what, who, variant, reps=60):
    for x in who:
        launch(x, what, variant, 2)
    for x in G:
        torch.cuda.synchronize(x["g"])
    time.sleep(0.3)
    c0 = [counters(x["h"]) for x in G]
    ev = []
    for x in who:
        with torch.cuda.device(x["g"]):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(x["st"])
            ev.append((e0, e1))
    for x in who:  # interleave the launches so that both GPUs start together
        launch(x, what, variant, reps)
    for x, (e0, e1) in zip(who, ev):
        with torch.cuda.device(x["g"]):
            e1.record(x["st"])
    for x in G:
        torch.cuda.synchronize(x["g"])
    time.sleep(0.3)
    c1 = [counters(x["h"]) for x in G]
    for x, (e0, e1) in zip(who, ev):
        ms = e0.elapsed_time(e1)
        pay = reps * n * cb
        d = {k: (c1[x["g"]][k] - c0[x["g"]][k]) if c0[x["g"]][k] is not None else None for k in c0[x["g"]]}
        f = lambda v: "   n/a" if v is None else f"{v / (ms * 1e6):6.1f}"
        r = lambda v: "  n/a" if v is None else f"{v / pay:5.3f}"
        print(f"{what:7s} {'both GPUs' if len(who) == 2 else 'GPU 0 only':10s} GPU {x['g']}: payload {pay / (ms * 1e6):6.1f} GB/s | wire GB/s raw tx {f(d['raw_tx'])} rx {f(d['raw_rx'])}"
              f" data tx {f(d['data_tx'])} rx {f(d['data_rx'])} | raw bytes per payload byte tx {r(d['raw_tx'])} rx {r(d['raw_rx'])}", flush=True)


for vname, v in (("bulk-copy kernel", capi.XFER_TMA), ("LD/ST kernel", capi.XFER_LDST)):
    print(f"# {vname}, {n} x 2 MiB per launch, 60 launches per row")
    run("restore", G[:1], v)
    run("restore", G, v)
    run("spill", G[:1], v)
    run("spill", G, v)
for x in G:
    assert x["res"].cpu().tolist() == [0, 0] and torch.equal(x["back"], x["hbm"])
print("round trips bit-exact")
