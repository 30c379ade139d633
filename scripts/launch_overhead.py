#!/usr/bin/env python
"""What the shim's launch path costs: PyTorch launches of a tiny kernel (x.add_(1) on 1024 floats), microseconds per
launch, (a) without the shim, (b) under the shim with the default managed backend, (c) with the VMM backend while
everything fits (unpaged: a shared lock), (d) with the VMM backend once the process is oversubscribed (paged: argument
scan + residency check under the residency lock, 200 managed ranges alive).  One JSON line."""
import json
import os
import subprocess
import sys
import tempfile
import textwrap

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SHIM = os.path.join(ROOT, "4paradigm-k8s-device-plugin_b200", "libvgpu.so")
WORKER = textwrap.dedent("""
    import ctypes, json, sys, time, torch
    torch.cuda.init()
    keep = [torch.empty(12 << 20, dtype=torch.uint8, device="cuda") for _ in range(200)]  # 200 segments of their own (>= 10 MiB)
    big = []
    if sys.argv[1] == "paged":                       # more than may be resident: the process goes paged
        big = [torch.empty(512 << 20, dtype=torch.uint8, device="cuda") for _ in range(3)]
        for b in big:
            b.fill_(1)
    x = torch.zeros(1024, device="cuda")
    y = keep[100][:4096].view(torch.float32)         # a tensor inside one of the managed ranges
    for t in (x, y):
        for _ in range(2000):
            t.add_(1)
    torch.cuda.synchronize()
    out = {}
    for name, t in (("small_tensor", x), ("tensor_in_a_managed_range", y)):
        n = 20000
        t0 = time.perf_counter()
        for _ in range(n):
            t.add_(1)
        t1 = time.perf_counter()
        torch.cuda.synchronize()
        out[name + "_us_per_launch"] = round((t1 - t0) / n * 1e6, 2)
    st = (ctypes.c_uint64 * 8)()
    try:
        ctypes.CDLL(None).cuVGPUOversubStats(st)
        out["paged"] = int(st[6])
    except AttributeError:
        out["paged"] = None
    print(json.dumps(out))
""")


def run(mode):
    env = dict(os.environ)
    if mode != "no_shim":
        td = tempfile.mkdtemp(prefix="vgpulo")
        env.update({"LD_PRELOAD": SHIM, "CUDA_DEVICE_MEMORY_SHARED_CACHE": os.path.join(td, "c.cache"), "LIBCUDA_LOG_LEVEL": "0",
                    "CUDA_DEVICE_MEMORY_LIMIT_0": "16g"})
    if mode in ("vmm_unpaged", "vmm_paged"):
        env.update({"CUDA_OVERSUBSCRIBE": "true", "VGPU_OVERSUB_BACKEND": "vmm"})
    if mode == "vmm_paged":
        env["VGPU_OVERSUB_RESIDENT_MB"] = "3000"      # the 200 x 12 MiB ranges fit, the three 512 MiB ones on top do not
    p = subprocess.run([sys.executable, "-c", WORKER, "paged" if mode == "vmm_paged" else "x"], env=env, capture_output=True, text=True, timeout=600)
    if p.returncode != 0:
        return {"error": p.stderr[-500:]}
    return json.loads(p.stdout.strip().splitlines()[-1])


if __name__ == "__main__":
    print(json.dumps({m: run(m) for m in ("no_shim", "shim_managed_backend", "vmm_unpaged", "vmm_paged")}))
