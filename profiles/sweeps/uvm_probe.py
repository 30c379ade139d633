#!/usr/bin/env python
"""Does the reference's oversubscription route (cuMemAllocManaged + UVM page faults) work on this box at all?
No shim involved: cudaMallocManaged through ctypes, touched from the GPU with cudaMemset, sizes below and above the HBM."""
import ctypes as C
import json
import sys
import time

rt = C.CDLL("libcudart.so.12") if len(sys.argv) < 2 else C.CDLL(sys.argv[1])
rt.cudaGetErrorString.restype = C.c_char_p
out = []


def err(rc):
    return None if rc == 0 else rt.cudaGetErrorString(rc).decode()


attr = C.c_int(0)
rt.cudaDeviceGetAttribute(C.byref(attr), 83, 0)   # cudaDevAttrManagedMemory
out.append({"managed_memory_attr": attr.value})
rt.cudaDeviceGetAttribute(C.byref(attr), 89, 0)   # cudaDevAttrConcurrentManagedAccess
out.append({"concurrent_managed_access_attr": attr.value})
rt.cudaDeviceGetAttribute(C.byref(attr), 88, 0)   # cudaDevAttrPageableMemoryAccess
out.append({"pageable_memory_access_attr": attr.value})
if len(sys.argv) > 2 and sys.argv[2] == "cumulative":     # 4 GiB managed allocations, untouched, until one is refused
    held = 0
    while held < 400:
        p = C.c_void_p(0)
        rc = rt.cudaMallocManaged(C.byref(p), C.c_size_t(4 << 30), 1)
        if rc != 0:
            print(json.dumps({"managed_GiB_granted_before_refusal": held, "refusal": err(rc)}))
            break
        held += 4
    else:
        print(json.dumps({"managed_GiB_granted_before_refusal": None, "held_GiB": held}))
    sys.exit(0)
for gib in [float(x) for x in (sys.argv[2:] or ["1", "64", "200"])]:
    p = C.c_void_p(0)
    n = int(gib * (1 << 30))
    r = {"GiB": gib}
    r["alloc"] = err(rt.cudaMallocManaged(C.byref(p), C.c_size_t(n), 1))
    if r["alloc"] is None:
        t0 = time.perf_counter()
        r["memset_on_gpu"] = err(rt.cudaMemset(p, 1, C.c_size_t(n)))
        r["sync"] = err(rt.cudaDeviceSynchronize())
        r["first_pass_s"] = round(time.perf_counter() - t0, 2)
        if r["sync"] is None and r["memset_on_gpu"] is None:
            t0 = time.perf_counter()
            r["memset_again"] = err(rt.cudaMemset(p, 2, C.c_size_t(n)))
            r["sync_again"] = err(rt.cudaDeviceSynchronize())
            r["second_pass_s"] = round(time.perf_counter() - t0, 2)
            r["second_pass_GBps"] = round(n / 1e9 / max(r["second_pass_s"], 1e-9), 1)
            b = (C.c_ubyte * 1).from_address(p.value + n - 1)
            r["last_byte_read_on_host"] = int(b[0])
        r["free"] = err(rt.cudaFree(p))
    out.append(r)
    print(json.dumps(r), flush=True)
    if any(v for k, v in r.items() if k in ("memset_on_gpu", "sync", "sync_again") and v):
        break
print(json.dumps({"uvm_probe": out}))
