#!/usr/bin/env python
"""BASELINE config 4 with an UNMODIFIED program: one vGPU whose limit (300 GB) exceeds the GPU (180 GB).

A plain PyTorch script allocates N tensors that add up to ~292 GB, writes every one, then reads all of them back twice
(forwards, backwards) and checks the contents.  It runs under the preload shim with CUDA_OVERSUBSCRIBE=true and
VGPU_OVERSUB_BACKEND=vmm; the shim decides at every launch / copy which ranges must be in HBM and spills the least
recently used ones — to the idle peer GPU named by VGPU_OVERSUB_PEER while the peer budget lasts, then to pinned host
RAM.  Prints one JSON line with the paging statistics and the rates seen by the program.

    python scripts/oversub_transparent_demo.py [tensors] [GiB each] [peer budget GiB] [vmm|managed] [timeout s] [GiB of HBM left free]

With "managed" the same program runs on the REFERENCE's route instead (cuMemoryAllocate@0x32146 -> cuMemAllocManaged,
the driver's UVM page faults do the paging; host RAM only): the like-for-like arm of this comparison.  Where the box
refuses managed allocations beyond a few tens of GiB, the last argument makes the GPU small instead: a second process
(no shim) holds all HBM but that many GiB for as long as the program runs.
"""
import json
import os
import subprocess
import sys
import tempfile
import textwrap
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SHIM = os.path.join(ROOT, "4paradigm-k8s-device-plugin_b200", "libvgpu.so")
N = int(sys.argv[1]) if len(sys.argv) > 1 else 17
GIB_EACH = float(sys.argv[2]) if len(sys.argv) > 2 else 16.0
PEER_GIB = float(sys.argv[3]) if len(sys.argv) > 3 else 64.0
BACKEND = sys.argv[4] if len(sys.argv) > 4 else "vmm"
TIMEOUT = float(sys.argv[5]) if len(sys.argv) > 5 else 3000
LEAVE_FREE_GIB = float(sys.argv[6]) if len(sys.argv) > 6 else 0.0
BALLAST = textwrap.dedent("""
    import sys, time, torch
    free, _ = torch.cuda.mem_get_info()
    t = torch.empty(free - int(float(sys.argv[1]) * (1 << 30)), dtype=torch.uint8, device="cuda")
    print("ready", torch.cuda.mem_get_info()[0], flush=True)
    time.sleep(100000)
""")

WORKER = textwrap.dedent("""
    import ctypes, json, sys, time, torch
    n_t, gib = int(sys.argv[1]), float(sys.argv[2])
    words = int(gib * (1 << 30)) // 8
    torch.cuda.init()
    step = 1 << 27
    def fill(t, k):                                   # t[i] = i * (40503 + k), in place, 1 GiB of temporaries at most
        for i in range(0, words, step):
            m = min(step, words - i)
            t[i:i + m] = torch.arange(i, i + m, device="cuda", dtype=torch.int64) * (40503 + k)
    def bad(t, k):
        b = 0
        for i in range(0, words, step):
            m = min(step, words - i)
            b += int((t[i:i + m] != torch.arange(i, i + m, device="cuda", dtype=torch.int64) * (40503 + k)).sum().item())
        return b
    ts = []
    t0 = time.perf_counter()
    for k in range(n_t):
        t = torch.empty(words, dtype=torch.int64, device="cuda")
        fill(t, k)
        ts.append(t)
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    print("write pass done in %.1f s" % (t1 - t0), flush=True)
    wrong = 0
    for k in range(n_t):
        wrong += bad(ts[k], k)
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    print("forward read pass done in %.1f s" % (t2 - t1), flush=True)
    for k in reversed(range(n_t)):
        wrong += bad(ts[k], k)
    torch.cuda.synchronize()
    t3 = time.perf_counter()
    st = (ctypes.c_uint64 * 8)()
    ctypes.CDLL(None).cuVGPUOversubStats(st)
    free, total = torch.cuda.mem_get_info()
    gb = n_t * words * 8 / 1e9
    print(json.dumps({"tensors": n_t, "GiB_each": gib, "program_GB": round(gb, 1), "limit_GB": round(total / 1e9, 1), "wrong_words": wrong,
                      "chunks_to_peer": st[0], "chunks_to_host": st[1], "chunks_restored": st[2], "checksum_failures": st[3],
                      "resident_GB_at_end": round(st[4] * 2097152 / 1e9, 1), "paged": st[6],
                      "write_pass_s": round(t1 - t0, 2), "read_pass_fwd_s": round(t2 - t1, 2), "read_pass_back_s": round(t3 - t2, 2),
                      "write_pass_GBps": round(gb / (t1 - t0), 1), "read_pass_fwd_GBps": round(gb / (t2 - t1), 1),
                      "read_pass_back_GBps": round(gb / (t3 - t2), 1)}), flush=True)
    del ts, t
    torch.cuda.empty_cache()                          # frees the ranges: with VGPU_VMEM_TRACE=1 each one reports its time split
""")


def main():
    import pynvml
    pynvml.nvmlInit()
    n_gpu = pynvml.nvmlDeviceGetCount()
    uu = [pynvml.nvmlDeviceGetUUID(pynvml.nvmlDeviceGetHandleByIndex(i)) for i in range(n_gpu)]
    uu = [u.decode() if isinstance(u, bytes) else u for u in uu]
    td = tempfile.mkdtemp(prefix="vgpuover")
    env = dict(os.environ)
    env.update({"LD_PRELOAD": SHIM, "CUDA_DEVICE_MEMORY_SHARED_CACHE": os.path.join(td, "c.cache"), "LIBCUDA_LOG_LEVEL": "0",
                "CUDA_OVERSUBSCRIBE": "true", "CUDA_DEVICE_MEMORY_LIMIT_0": "300000m", "NVIDIA_DEVICE_MAP": f"0:{uu[0]}"})
    if BACKEND == "vmm":
        env["VGPU_OVERSUB_BACKEND"] = "vmm"
    if n_gpu > 1 and BACKEND == "vmm" and PEER_GIB > 0:
        env.update({"VGPU_OVERSUB_PEER": uu[1], "VGPU_OVERSUB_PEER_MB": str(int(PEER_GIB * 1024))})
    ballast = None
    if LEAVE_FREE_GIB > 0:
        ballast = subprocess.Popen([sys.executable, "-c", BALLAST, str(LEAVE_FREE_GIB)], stdout=subprocess.PIPE, text=True)
        ready = ballast.stdout.readline().split()
        assert ready and ready[0] == "ready", "the ballast process did not come up"
    try:
        run(env, ballast)
    finally:
        if ballast is not None:
            ballast.kill()
            ballast.wait()


def run(env, ballast):
    t0 = time.perf_counter()
    try:
        p = subprocess.run([sys.executable, "-c", WORKER, str(N), str(GIB_EACH)], env=env, capture_output=True, text=True, timeout=TIMEOUT)
    except subprocess.TimeoutExpired as e:
        out = (e.stdout or b"").decode() if isinstance(e.stdout, bytes) else (e.stdout or "")
        print(json.dumps({"backend": BACKEND, "tensors": N, "GiB_each": GIB_EACH, "timed_out_after_s": TIMEOUT, "progress": out.strip().splitlines()[-3:]}))
        return
    wall = time.perf_counter() - t0
    if p.returncode != 0:
        print(p.stderr[-3000:], file=sys.stderr)
        raise SystemExit(p.returncode)
    o = json.loads(p.stdout.strip().splitlines()[-1])
    trace = [l for l in p.stderr.splitlines() if l.startswith("[vgpu_vmem]")]      # VGPU_VMEM_TRACE=1: the manager's time split per range
    if trace:
        o["vmem_trace"] = trace
    o.update({"backend": BACKEND, "hbm_left_free_GiB": LEAVE_FREE_GIB or None, "wall_s": round(wall, 1), "peer_tier": (f"GPU 1, budget {PEER_GIB} GiB" if "VGPU_OVERSUB_PEER" in env else None),
              "host_mem_available_GB": round(int(next(l for l in open("/proc/meminfo") if l.startswith("MemAvailable")).split()[1]) / 1e6, 1)})
    print(json.dumps(o))


if __name__ == "__main__":
    main()
