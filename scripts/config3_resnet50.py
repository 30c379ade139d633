#!/usr/bin/env python
"""BASELINE.json config 3: ResNet-50 bf16 synthetic inference, batch 64, on 4 vGPUs sharing one
B200, each under the preload shim with CUDA_DEVICE_SM_LIMIT=25 (green-context SM partition) and a
45 GiB quota — the SM-partition acceptance test.  Not a roofline row for
our kernels (cuDNN does the math); what is checked is the cap: each vGPU runs in 32 of 148 SMs,
the four run concurrently on disjoint partitions, and the aggregate vs one uncapped process.

  python scripts/config3_resnet50.py            # parent: launches the worker processes
"""
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SHIM = os.path.join(ROOT, "4paradigm-k8s-device-plugin_b200", "libvgpu.so")


def worker():
    import ctypes
    import torch
    import torchvision
    torch.manual_seed(0)
    m = torchvision.models.resnet50(weights=None).cuda().to(torch.bfloat16).to(memory_format=torch.channels_last).eval()
    x = torch.randn(64, 3, 224, 224, generator=torch.Generator().manual_seed(0)).cuda().to(torch.bfloat16).to(memory_format=torch.channels_last)
    iters = int(os.environ.get("ITERS", "30"))
    with torch.no_grad():
        for _ in range(5):
            y = m(x)
        torch.cuda.synchronize()
        print("ready", flush=True)
        sys.stdin.readline()  # start gate: all vGPUs begin together
        t0 = time.perf_counter()
        for _ in range(iters):
            y = m(x)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
    try:
        sms = ctypes.CDLL(None).cuVGPUPartitionSmCount(0)
    except AttributeError:
        sms = -1  # not under the shim
    free, total = torch.cuda.mem_get_info()
    print(json.dumps({"images_per_s": 64 * iters / dt, "partition_sms": sms, "quota_total": total,
                      "checksum": float(y.float().abs().mean())}), flush=True)


def launch(n, capped, iters=30):
    procs = []
    for i in range(n):
        env = dict(os.environ, ITERS=str(iters))
        if capped:
            td = tempfile.mkdtemp(prefix="vgpu_c3_")
            env.update({"LD_PRELOAD": SHIM, "CUDA_DEVICE_MEMORY_SHARED_CACHE": os.path.join(td, "shr.cache"),
                        "CUDA_DEVICE_MEMORY_LIMIT_0": "46080m", "CUDA_DEVICE_SM_LIMIT": "25", "LIBCUDA_LOG_LEVEL": "0",
                        "VGPU_SM_PARTITION_INDEX": str(i)})
        procs.append(subprocess.Popen([sys.executable, __file__, "--worker"], env=env, stdin=subprocess.PIPE, stdout=subprocess.PIPE, text=True))
    for p in procs:
        assert p.stdout.readline().strip() == "ready"
    for p in procs:
        p.stdin.write("\n")
        p.stdin.flush()
    out = []
    for p in procs:
        out.append(json.loads(p.stdout.readline()))
        p.wait(timeout=120)
    return out


def run_suite():
    one_free = launch(1, capped=False)
    one_cap = launch(1, capped=True)
    four_cap = launch(4, capped=True)
    four_free = launch(4, capped=False)
    res = {
        "config": "ResNet-50 bf16 channels-last inference, batch 64, synthetic input seed 0",
        "one_uncapped_images_per_s": round(one_free[0]["images_per_s"], 1),
        "one_vgpu_25pct_images_per_s": round(one_cap[0]["images_per_s"], 1),
        "four_vgpus_25pct": {"per_vgpu_images_per_s": [round(o["images_per_s"], 1) for o in four_cap],
                             "aggregate_images_per_s": round(sum(o["images_per_s"] for o in four_cap), 1),
                             "partition_sms": [o["partition_sms"] for o in four_cap],
                             "quota_total_bytes": [o["quota_total"] for o in four_cap]},
        "four_uncapped_processes_aggregate_images_per_s": round(sum(o["images_per_s"] for o in four_free), 1),
        "same_output_capped_vs_uncapped": abs(one_cap[0]["checksum"] - one_free[0]["checksum"]) < 1e-6,
    }
    res["cap_ratio_one_vgpu"] = round(one_cap[0]["images_per_s"] / one_free[0]["images_per_s"], 3)
    res["aggregate_vs_one_uncapped"] = round(res["four_vgpus_25pct"]["aggregate_images_per_s"] / res["one_uncapped_images_per_s"], 3)
    return res


class Mps:
    """Spatial sharing ACROSS PROCESSES needs the MPS daemon: without it the driver time-slices
    whole-GPU contexts, so four capped vGPUs take turns on 32 SMs each (measured: 4 x 1135 img/s).
    With MPS the four green-context partitions run at the same time on disjoint SMs."""

    def __enter__(self):
        self.dir = tempfile.mkdtemp(prefix="mps_")
        os.environ["CUDA_MPS_PIPE_DIRECTORY"] = os.path.join(self.dir, "pipe")
        os.environ["CUDA_MPS_LOG_DIRECTORY"] = os.path.join(self.dir, "log")
        os.makedirs(os.environ["CUDA_MPS_PIPE_DIRECTORY"]); os.makedirs(os.environ["CUDA_MPS_LOG_DIRECTORY"])
        self.ok = subprocess.run(["nvidia-cuda-mps-control", "-d"]).returncode == 0
        time.sleep(1)
        return self

    def __exit__(self, *a):
        if self.ok:
            subprocess.run(["nvidia-cuda-mps-control"], input="quit\n", text=True)
        os.environ.pop("CUDA_MPS_PIPE_DIRECTORY", None); os.environ.pop("CUDA_MPS_LOG_DIRECTORY", None)


if __name__ == "__main__":
    if "--worker" in sys.argv:
        worker()
        sys.exit(0)
    out = {"without_mps": run_suite()}
    with Mps() as m:
        out["with_mps"] = run_suite() if m.ok else "nvidia-cuda-mps-control failed to start"
    print(json.dumps(out))
