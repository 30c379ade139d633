#!/usr/bin/env python
"""Three vGPUs whose quotas over-commit one B200, time-shared by monitor.Scheduler.

Three containers (directories under a shared root, as VGPU_MONITOR_MODE lays them out, server.go:494-501), each one
PyTorch process under the preload shim with a 70 GiB tensor: 210 GiB on a 180 GB GPU.  Every worker fills its tensor,
then verifies every word ROUNDS times.  The scheduler ticks every 0.5 s: it suspends (SIGUSR2 -> the shim spills the
process's ranges to pinned host RAM with the spill kernel and gives the HBM back) and resumes (SIGUSR1 -> restore +
checksum verify) so that the running set fits, rotating the tenants every `quantum` ticks.  This script also watches the
GPU's used memory (NVML, every 20 ms) to time each suspend / resume end to end.  One JSON line.

    python scripts/monitor_demo.py [GiB per tenant] [rounds] [quantum ticks]
"""
import json
import os
import subprocess
import sys
import tempfile
import textwrap
import threading
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
SHIM = os.path.join(ROOT, "4paradigm-k8s-device-plugin_b200", "libvgpu.so")
GIB = 1 << 30
SIZE_GIB = float(sys.argv[1]) if len(sys.argv) > 1 else 70.0
ROUNDS = int(sys.argv[2]) if len(sys.argv) > 2 else 6
QUANTUM = int(sys.argv[3]) if len(sys.argv) > 3 else 6

WORKER = textwrap.dedent("""
    import json, sys, time, torch
    gib, rounds, k = float(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
    n = int(gib * (1 << 30)) // 8
    a = torch.empty(n, dtype=torch.int64, device="cuda")
    step = 1 << 27
    for i in range(0, n, step):
        m = min(step, n - i)
        a[i:i + m] = torch.arange(i, i + m, device="cuda", dtype=torch.int64) * (40503 + k)
    torch.cuda.synchronize()
    print(json.dumps({"filled": True}), flush=True)
    bad, t0 = 0, time.time()
    for r in range(rounds):
        for i in range(0, n, step):
            m = min(step, n - i)
            bad += int((a[i:i + m] != torch.arange(i, i + m, device="cuda", dtype=torch.int64) * (40503 + k)).sum().item())
        time.sleep(0.3)                 # a little host work between rounds
        print(json.dumps({"round": r, "bad": bad, "t": round(time.time() - t0, 2)}), flush=True)
    print(json.dumps({"done": True, "bad": bad, "seconds": round(time.time() - t0, 2)}), flush=True)
""")


def main():
    import pynvml
    import vgpu_b200  # noqa: F401
    from vgpu_b200 import monitor
    pynvml.nvmlInit()
    h = pynvml.nvmlDeviceGetHandleByIndex(0)
    uuid = pynvml.nvmlDeviceGetUUID(h)
    uuid = uuid.decode() if isinstance(uuid, bytes) else uuid
    total = pynvml.nvmlDeviceGetMemoryInfo(h).total
    root = tempfile.mkdtemp(prefix="vgpushared")
    procs = []
    for k, name in enumerate(("pod-a_main", "pod-b_main", "pod-c_main")):
        os.makedirs(os.path.join(root, name))
        env = dict(os.environ)
        env.update({"LD_PRELOAD": SHIM, "CUDA_DEVICE_MEMORY_SHARED_CACHE": os.path.join(root, name, "vgpu.cache"), "LIBCUDA_LOG_LEVEL": "3",
                    "CUDA_OVERSUBSCRIBE": "true", "VGPU_OVERSUB_BACKEND": "vmm", "CUDA_DEVICE_MEMORY_LIMIT_0": f"{int(SIZE_GIB + 6)}g",
                    "NVIDIA_DEVICE_MAP": f"0:{uuid}", "CUDA_TASK_PRIORITY": "1", "VGPU_OVERSUB_WAIT_S": "300"})
        if pynvml.nvmlDeviceGetCount() > 1:   # an idle second GPU: its HBM is the first spill tier (NVLink), pinned host RAM the second
            pu = pynvml.nvmlDeviceGetUUID(pynvml.nvmlDeviceGetHandleByIndex(1))
            env["VGPU_OVERSUB_PEER"] = pu.decode() if isinstance(pu, bytes) else pu
        procs.append((name, subprocess.Popen([sys.executable, "-c", WORKER, str(SIZE_GIB), str(ROUNDS), str(k)], env=env,
                                             stdout=subprocess.PIPE, stderr=open(os.path.join(root, name + ".err"), "w"), text=True)))
        time.sleep(1.0)   # staggered start, as pods arrive

    # used-memory trace of the GPU, for end-to-end suspend / resume rates
    trace, stop = [], threading.Event()

    def sampler():
        while not stop.is_set():
            trace.append((time.perf_counter(), pynvml.nvmlDeviceGetMemoryInfo(h).used))
            time.sleep(0.02)
    threading.Thread(target=sampler, daemon=True).start()

    events = []
    sched = monitor.Scheduler(root, hbm={uuid.lower(): total}, headroom=6 * GIB, quantum=QUANTUM,
                              hbm_free=lambda g: pynvml.nvmlDeviceGetMemoryInfo(h).free,
                              log=lambda what, name: events.append((time.perf_counter(), what, name)))
    results = {}

    def reader(name, p):
        for line in p.stdout:
            try:
                o = json.loads(line)
            except ValueError:
                continue
            if o.get("done"):
                results[name] = o
    th = [threading.Thread(target=reader, args=(n, p), daemon=True) for n, p in procs]
    [t.start() for t in th]
    t0 = time.perf_counter()
    while any(p.poll() is None for _, p in procs) and time.perf_counter() - t0 < 1500:
        sched.tick()
        time.sleep(0.5)
    for _, p in procs:
        if p.poll() is None:
            p.kill()
    [t.join(timeout=5) for t in th]
    stop.set()
    wall = time.perf_counter() - t0
    logs = {n: open(os.path.join(root, n + ".err")).read() for n, _ in procs}
    errs = {n: logs[n][-1500:] for n, p in procs if p.returncode != 0}

    # the shim times every move of a whole vGPU itself (mover thread, INFO log): drain + spill kernels + unmap on the
    # way out, map + restore kernels + checksum verification on the way back
    import re
    moves = []
    for n, text in logs.items():
        for what, mib, ms in re.findall(r"(suspend|resume): (\d+) MiB of managed ranges (?:out|back) in ([0-9.]+) ms", text):
            moves.append({"what": what, "tenant": n, "GB": round(int(mib) * (1 << 20) / 1e9, 1), "seconds": round(float(ms) / 1e3, 3),
                          "GBps_end_to_end": round(int(mib) * (1 << 20) / 1e9 / max(float(ms) / 1e3, 1e-6), 1)})
    _ = trace
    out = {"tenants": 3, "GiB_each": SIZE_GIB, "gpu_GB": round(total / 1e9, 1), "rounds": ROUNDS, "quantum_ticks": QUANTUM,
           "wall_s": round(wall, 1), "results": results, "all_bit_exact": len(results) == 3 and all(r["bad"] == 0 for r in results.values()),
           "suspends": sum(1 for e in events if e[1] == "suspend"), "resumes": sum(1 for e in events if e[1] == "resume"),
           "moves_timed_by_the_shim": moves,
           "median_suspend_GBps_end_to_end": (sorted(m["GBps_end_to_end"] for m in moves if m["what"] == "suspend") or [None])[len([m for m in moves if m["what"] == "suspend"]) // 2],
           "median_resume_GBps_end_to_end": (sorted(m["GBps_end_to_end"] for m in moves if m["what"] == "resume") or [None])[len([m for m in moves if m["what"] == "resume"]) // 2],
           "errors": errs}
    print(json.dumps(out))
    if not out["all_bit_exact"]:
        raise SystemExit(1)


if __name__ == "__main__":
    main()
