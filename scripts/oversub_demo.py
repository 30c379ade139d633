#!/usr/bin/env python
"""Two vGPUs whose quotas add up to more than the GPU holds, time-shared by a monitor:
A (80 GiB) and B (110 GiB) on one 180 GB B200.  A fills its memory; the monitor suspends A
(SIGUSR2: A's data is spilled to pinned host RAM by the spill kernel, its HBM released); B
allocates and works; B exits; the monitor resumes A (SIGUSR1: restore + checksum verify) and A
checks every byte.  Both run under the preload shim with CUDA_OVERSUBSCRIBE=true and
VGPU_OVERSUB_BACKEND=vmm.  Prints one JSON line."""
import json
import os
import subprocess
import sys
import tempfile
import textwrap
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
SHIM = os.path.join(ROOT, "4paradigm-k8s-device-plugin_b200", "libvgpu.so")
GIB = 1 << 30
A_GIB = float(os.environ.get("A_GIB", "80"))
B_GIB = float(os.environ.get("B_GIB", "110"))

WORKER = textwrap.dedent("""
    import json, sys, time, torch
    gib = float(sys.argv[1])
    n = int(gib * (1 << 30)) // 8
    a = torch.empty(n, dtype=torch.int64, device="cuda")
    step = 1 << 27
    for i in range(0, n, step):                      # a[i] = i * 40503, in place, no big temporaries
        m = min(step, n - i)
        a[i:i + m] = torch.arange(i, i + m, device="cuda", dtype=torch.int64) * 40503
    torch.cuda.synchronize()
    def check():
        bad = 0
        for i in range(0, n, step):
            m = min(step, n - i)
            bad += int((a[i:i + m] != torch.arange(i, i + m, device="cuda", dtype=torch.int64) * 40503).sum().item())
        return bad
    print(json.dumps({"ready": True, "bad": check()}), flush=True)
    for line in sys.stdin:
        if line.strip() == "check":
            t0 = time.perf_counter(); b = check(); print(json.dumps({"bad": b, "s": time.perf_counter() - t0}), flush=True)
        elif line.strip() == "stop":
            break
""")


def used_gib():
    env = {k: v for k, v in os.environ.items() if k != "LD_PRELOAD"}
    o = subprocess.run(["nvidia-smi", "--query-gpu=memory.used", "--format=csv,noheader,nounits", "-i", "0"], capture_output=True, text=True, env=env).stdout
    return int(o.strip().splitlines()[0]) / 1024


def spawn(gib, cache, quota):
    env = {**os.environ, "LD_PRELOAD": SHIM, "CUDA_DEVICE_MEMORY_SHARED_CACHE": cache, "CUDA_DEVICE_MEMORY_LIMIT_0": quota,
           "CUDA_OVERSUBSCRIBE": "true", "VGPU_OVERSUB_BACKEND": "vmm", "LIBCUDA_LOG_LEVEL": "0",
           "PYTORCH_NO_CUDA_MEMORY_CACHING": "1"}
    return subprocess.Popen([sys.executable, "-c", WORKER, str(gib)], env=env, stdin=subprocess.PIPE, stdout=subprocess.PIPE, text=True)


def wait_used(pred, timeout=180):
    t0 = time.time()
    while time.time() - t0 < timeout:
        u = used_gib()
        if pred(u):
            return u, time.time() - t0
        time.sleep(0.25)
    raise SystemExit(f"timeout waiting for GPU memory to change (now {used_gib():.1f} GiB)")


def main():
    import vgpu_b200  # noqa: F401
    from vgpu_b200 import monitor
    td = tempfile.mkdtemp(prefix="oversub")
    ca, cb = os.path.join(td, "a.cache"), os.path.join(td, "b.cache")
    out = {"A_GiB": A_GIB, "B_GiB": B_GIB}
    u_idle = used_gib()
    A = spawn(A_GIB, ca, f"{int(A_GIB) + 4}g")
    ra = json.loads(A.stdout.readline())
    assert ra["ready"] and ra["bad"] == 0
    u_a = used_gib()
    out["A_resident_GiB"] = round(u_a - u_idle, 1)
    # first suspend pins the host slab (cudaHostAlloc, ~0.4 s per GiB); later ones reuse it
    t0 = time.time()
    monitor.suspend(ca)
    wait_used(lambda u: u < u_a - A_GIB + 2)
    out["first_suspend_s_incl_pinning_host_slab"] = round(time.time() - t0, 2)
    t0 = time.time()
    monitor.resume(ca)
    A.stdin.write("check\n"); A.stdin.flush()
    r1 = json.loads(A.stdout.readline())
    out["first_resume_plus_check_s"] = round(time.time() - t0, 2)
    assert r1["bad"] == 0
    t0 = time.time()
    monitor.suspend(ca)
    u_s, _ = wait_used(lambda u: u < u_a - A_GIB + 2)
    out["suspend_s"] = round(time.time() - t0, 2)
    out["suspend_GBps"] = round(A_GIB * GIB / 1e9 / out["suspend_s"], 1)
    out["A_suspended_GiB_on_gpu"] = round(u_s - u_idle, 1)
    B = spawn(B_GIB, cb, f"{int(B_GIB) + 4}g")           # would not fit next to A
    rb = json.loads(B.stdout.readline())
    assert rb["ready"] and rb["bad"] == 0
    out["B_resident_GiB"] = round(used_gib() - u_s, 1)
    B.stdin.write("stop\n"); B.stdin.flush(); B.wait(timeout=120)
    t0 = time.time()
    monitor.resume(ca)
    A.stdin.write("check\n"); A.stdin.flush()            # parks at the gate until the data is back
    rc = json.loads(A.stdout.readline())
    out["resume_plus_check_s"] = round(time.time() - t0, 2)
    out["A_bad_words_after_resume"] = rc["bad"]
    out["A_back_on_gpu_GiB"] = round(used_gib() - u_idle, 1)
    A.stdin.write("stop\n"); A.stdin.flush(); A.wait(timeout=120)
    out["ok"] = rc["bad"] == 0 and out["A_suspended_GiB_on_gpu"] < 4
    print(json.dumps(out))
    return 0 if out["ok"] else 1


if __name__ == "__main__":
    sys.exit(main())
