#!/usr/bin/env python
"""Config 1 side figure: quota-hook overhead in ops/s, reference libvgpu.so vs the new shim, both
over the same mock driver (CPU only; the reference binary exists only in the build container).
One op = cuMemAlloc_v2 + cuMemGetInfo_v2 + cuMemFree_v2 of 1 MiB, timed inside the runner."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import quota_harness as H  # noqa: E402

H.build()
env = {"CUDA_DEVICE_MEMORY_LIMIT_0": "18432m", "CUDA_DEVICE_SM_LIMIT": "10", "NVIDIA_DEVICE_MAP": "0:" + H.UUID0, "MOCK_CUDA_LOG": ""}
N = 200000
out = {}
for name, lib in (("reference", H.REFERENCE_LIB), ("new_shim", H.NEW_LIB)):
    if not os.path.exists(lib):
        out[name] = None
        continue
    best = None
    for _ in range(3):
        r = H.run_trace(lib, [(0, f"loop {N} 1048576")], env, keep_region=False)
        rc, ns = r["replies"][0].split()
        assert rc == "0"
        best = min(best, int(ns)) if best else int(ns)
    out[name] = {"ns_per_alloc_info_free": round(best / N, 1), "triples_per_s": round(N / (best * 1e-9))}
if out.get("reference") and out.get("new_shim"):
    out["speedup"] = round(out["reference"]["ns_per_alloc_info_free"] / out["new_shim"]["ns_per_alloc_info_free"], 2)
print(json.dumps(out))
