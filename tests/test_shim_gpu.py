"""The preload shim under the REAL driver: a PyTorch process with LD_PRELOAD=libvgpu.so and the
env vars the plugin's Allocate would set (Contract A).  Checks that the runtime API really lands
in the hooks (cudart resolves everything through dlsym + cuGetProcAddress), that the quota numbers
are the reference's, and that CUDA_DEVICE_SM_LIMIT becomes a spatial green-context partition."""
import json
import os
import subprocess
import sys
import tempfile
import textwrap

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SHIM = os.path.join(ROOT, "4paradigm-k8s-device-plugin_b200", "libvgpu.so")
GiB = 1 << 30


def run_under_shim(code, env_extra, cache=None, timeout=300):
    td = tempfile.mkdtemp(prefix="vgpushim")
    env = dict(os.environ)
    env.update({"LD_PRELOAD": SHIM, "CUDA_DEVICE_MEMORY_SHARED_CACHE": cache or os.path.join(td, "shr.cache"),
                "LIBCUDA_LOG_LEVEL": "0"})
    env.update(env_extra)
    p = subprocess.run([sys.executable, "-c", textwrap.dedent(code)], env=env, capture_output=True, text=True, timeout=timeout)
    return p


def last_json(p):
    assert p.returncode == 0, p.stderr[-3000:]
    return json.loads(p.stdout.strip().splitlines()[-1])


def test_quota_is_enforced_through_the_runtime_api(cuda_dev):
    code = """
        import json, torch
        out = {}
        torch.cuda.init()
        out["total_prop"] = torch.cuda.get_device_properties(0).total_memory
        f0, t0 = torch.cuda.mem_get_info()
        a = torch.empty(3 << 30, dtype=torch.uint8, device="cuda")
        f1, t1 = torch.cuda.mem_get_info()
        try:
            b = torch.empty(2 << 30, dtype=torch.uint8, device="cuda")
            out["second"] = "allocated"
        except Exception as e:
            out["second"] = type(e).__name__
        f2, _ = torch.cuda.mem_get_info()
        del a
        torch.cuda.empty_cache()
        f3, _ = torch.cuda.mem_get_info()
        out.update(f0=f0, t0=t0, f1=f1, t1=t1, f2=f2, f3=f3)
        print(json.dumps(out))
    """
    p = run_under_shim(code, {"CUDA_DEVICE_MEMORY_LIMIT_0": "4096m"})
    o = last_json(p)
    assert o["t0"] == 4 * GiB == o["t1"] == o["total_prop"]         # cuMemGetInfo total and cuDeviceTotalMem are the limit
    assert o["f0"] <= 4 * GiB and o["f0"] - o["f1"] == 3 * GiB      # free = limit - usage
    assert o["second"] != "allocated"                                # 3 GiB + 2 GiB > 4 GiB
    assert "Device 0 OOM" in p.stderr                                # same ERROR line as the reference (allocator.c:53)
    assert o["f2"] == o["f1"] and o["f3"] == o["f0"]


def test_context_is_billed_to_the_vgpu_and_host_pid_recorded(cuda_dev):
    """Reference set_task_pid@0x17de8 + cuDevicePrimaryCtxRetain@0x29c86: the process's context (NVML usedGpuMemory
    right after the first primary-context retain) counts against the quota, and the pid NVML knows the process by
    goes into the slot.  Same numbers from three sides: cuMemGetInfo, the cache file, NVML."""
    code = """
        import importlib, json, os, sys, torch
        torch.cuda.init()
        f, t = torch.cuda.mem_get_info()
        sys.path.insert(0, os.environ["VGPU_PKG"])
        mon = importlib.import_module("monitor")
        r = mon.read_region(os.environ["CUDA_DEVICE_MEMORY_SHARED_CACHE"])
        me = [p for p in r["procs"] if p["pid"] == os.getpid()][0]
        import pynvml
        pynvml.nvmlInit()
        procs = pynvml.nvmlDeviceGetComputeRunningProcesses(pynvml.nvmlDeviceGetHandleByIndex(0))
        print(json.dumps({"free": f, "total": t, "hostpid": me["hostpid"], "ctx": me["used"][0]["context_size"],
                          "slot_total": me["used"][0]["total"], "nvml": {str(p.pid): p.usedGpuMemory for p in procs}}))
    """
    p = run_under_shim(code, {"CUDA_DEVICE_MEMORY_LIMIT_0": "8g", "VGPU_PKG": os.path.join(ROOT, "4paradigm-k8s-device-plugin_b200")})
    o = last_json(p)
    print("context billing:", o)
    assert o["total"] == 8 * GiB
    if o["hostpid"] == 0:                                   # NVML lists no process for us here: nothing may be billed
        assert o["ctx"] == 0 and o["free"] == o["total"] and "SET_TASK_PID FAILED" in p.stderr
        pytest.skip("NVML does not list this container's processes: context size unknown, as in the reference's SET_TASK_PID FAILED")
    assert 32 << 20 <= o["ctx"] <= 4 * GiB                   # a B200 context is hundreds of MiB
    assert o["total"] - o["free"] == o["ctx"] == o["slot_total"]
    assert str(o["hostpid"]) in o["nvml"]                    # the pid NVML knows us by ...
    assert o["nvml"][str(o["hostpid"])] >= o["ctx"]          # ... and its memory only grew since the probe


def _gpu0_uuid():
    o = subprocess.run(["nvidia-smi", "--query-gpu=uuid", "--format=csv,noheader", "-i", "0"], capture_output=True, text=True).stdout
    return o.strip().splitlines()[0].strip()


def test_two_vgpus_on_one_physical_gpu_are_two_cuda_devices(cuda_dev):
    """NVIDIA_DEVICE_MAP names the GPU twice (reference: initial_virtual_devices@0x27f0d, cuDevicePrimaryCtxRetain
    @0x297be, assigning_virtual_pcibusID@0x2c5c1): two CUDA devices, the second with PCI function .1, a context of its
    own and its own quota — here on the real driver, through the driver API, with a kernel-free data check (memset +
    copy back) inside the second device's context."""
    code = """
        import ctypes, json
        cu = ctypes.CDLL("libcuda.so.1")            # resolved through the preloaded shim first
        out = {}
        assert cu.cuInit(0) == 0
        n = ctypes.c_int(); cu.cuDeviceGetCount(ctypes.byref(n)); out["count"] = n.value
        devs = []
        for i in range(n.value):
            d = ctypes.c_int(-1); rc = cu.cuDeviceGet(ctypes.byref(d), i)
            name = ctypes.create_string_buffer(64); cu.cuDeviceGetName(name, 64, d)
            bus = ctypes.create_string_buffer(32); cu.cuDeviceGetPCIBusId(bus, 32, d)
            tot = ctypes.c_size_t(); cu.cuDeviceTotalMem_v2(ctypes.byref(tot), d)
            sms = ctypes.c_int(); cu.cuDeviceGetAttribute(ctypes.byref(sms), 16, d)
            back = ctypes.c_int(-1); cu.cuDeviceGetByPCIBusId(ctypes.byref(back), bus.value)
            devs.append({"rc": rc, "dev": d.value, "name": name.value.decode(), "bus": bus.value.decode(), "total": tot.value,
                         "sms": sms.value, "by_bus": back.value})
        out["devs"] = devs
        d2 = ctypes.c_int(-1); out["get_oob"] = cu.cuDeviceGet(ctypes.byref(d2), n.value)
        ctxs = []
        for i in range(n.value):
            c = ctypes.c_void_p(); rc = cu.cuDevicePrimaryCtxRetain(ctypes.byref(c), i); ctxs.append(c)
            assert rc == 0, rc
        out["distinct_ctx"] = len({c.value for c in ctxs}) == n.value
        res = []
        for i in range(n.value):
            assert cu.cuCtxSetCurrent(ctxs[i]) == 0
            cur = ctypes.c_int(-1); cu.cuCtxGetDevice(ctypes.byref(cur))
            f, t = ctypes.c_size_t(), ctypes.c_size_t(); cu.cuMemGetInfo_v2(ctypes.byref(f), ctypes.byref(t))
            p = ctypes.c_ulonglong(); rc_fit = cu.cuMemAlloc_v2(ctypes.byref(p), ctypes.c_size_t(t.value // 2))
            q = ctypes.c_ulonglong(); rc_over = cu.cuMemAlloc_v2(ctypes.byref(q), ctypes.c_size_t(t.value // 2 + (64 << 20)))
            f2 = ctypes.c_size_t(); cu.cuMemGetInfo_v2(ctypes.byref(f2), ctypes.byref(t))
            # the memory is real and usable from this context
            cu.cuMemsetD32_v2(ctypes.c_ulonglong(p.value), ctypes.c_uint(0xC0FFEE00 + i), ctypes.c_size_t(1 << 20))
            host = (ctypes.c_uint * 4)(); rc_cp = cu.cuMemcpyDtoH_v2(host, ctypes.c_ulonglong(p.value + (1 << 21)), ctypes.c_size_t(16))
            seen = ctypes.c_void_p(); cu.cuCtxGetCurrent(ctypes.byref(seen))
            shim = ctypes.CDLL(None)
            shim.cuVGPUPartitionCtx.restype = ctypes.c_void_p
            res.append({"cur": cur.value, "free": f.value, "total": t.value, "rc_fit": rc_fit, "rc_over": ctypes.c_int(rc_over).value,
                        "free_after": f2.value, "rc_cp": rc_cp, "word": host[0], "same_handle": seen.value == ctxs[i].value,
                        "sms": shim.cuVGPUPartitionSmCount(i), "partition_ctx": shim.cuVGPUPartitionCtx(i) or 0})
        out["ctx"] = res
        print(json.dumps(out))
    """
    uuid = _gpu0_uuid()
    p = run_under_shim(code, {"NVIDIA_DEVICE_MAP": f"0:{uuid} 1:{uuid}", "CUDA_DEVICE_MEMORY_LIMIT_0": "4g", "CUDA_DEVICE_MEMORY_LIMIT_1": "8g",
                              "CUDA_DEVICE_SM_LIMIT_0": "25", "CUDA_DEVICE_SM_LIMIT_1": "50"})
    o = last_json(p)
    print("virtual devices:", o)
    # each vGPU also gets ITS SM share as a green-context partition, bound behind the handle the application holds
    assert [r["same_handle"] for r in o["ctx"]] == [True, True]
    assert 0 < o["ctx"][0]["sms"] <= 37 and 37 < o["ctx"][1]["sms"] <= 74
    assert o["ctx"][0]["partition_ctx"] and o["ctx"][1]["partition_ctx"] and o["ctx"][0]["partition_ctx"] != o["ctx"][1]["partition_ctx"]
    assert o["count"] == 2 and o["get_oob"] == 101 and o["distinct_ctx"]
    a, b = o["devs"]
    assert (a["dev"], b["dev"]) == (0, 1) and a["name"] == b["name"]
    # the SM-count attribute of a capped vGPU is its partition's size (asked before its context exists: the planned size)
    assert a["sms"] == o["ctx"][0]["sms"] and b["sms"] == o["ctx"][1]["sms"]
    assert a["bus"].endswith(".0") and b["bus"] == a["bus"][:-1] + "1" and (a["by_bus"], b["by_bus"]) == (0, 1)
    assert (a["total"], b["total"]) == (4 * GiB, 8 * GiB)
    for i, (r, lim) in enumerate(zip(o["ctx"], (4 * GiB, 8 * GiB))):
        assert r["cur"] == i and r["total"] == lim and r["free"] <= lim
        assert r["rc_fit"] == 0 and r["rc_over"] == -1                       # each vGPU enforces ITS limit
        assert r["free"] - r["free_after"] == lim // 2
        assert r["rc_cp"] == 0 and r["word"] == 0xC0FFEE00 + i
    assert "Device 0 OOM" in p.stderr and "Device 1 OOM" in p.stderr


def test_pytorch_gets_the_vgpus_of_a_shared_gpu_as_one_device(cuda_dev):
    """The CUDA runtime cannot initialise with two devices on one GPU (it cross-checks the public API against private
    driver tables), so a runtime-API process — recognised by its cuGetExportTable call — sees ONE device that owns both
    vGPUs' quota: 4g + 8g, in cudaMemGetInfo, device properties and NVML alike."""
    code = """
        import json, torch, pynvml
        out = {"count": torch.cuda.device_count()}
        f, t = torch.cuda.mem_get_info()
        x = torch.empty(9 << 30, dtype=torch.uint8, device="cuda")      # more than either vGPU alone
        x[-1] = 7
        try:
            y = torch.empty(4 << 30, dtype=torch.uint8, device="cuda"); over = "allocated"
        except Exception as e:
            over = type(e).__name__
        f2, _ = torch.cuda.mem_get_info()
        pynvml.nvmlInit()
        m = pynvml.nvmlDeviceGetMemoryInfo(pynvml.nvmlDeviceGetHandleByIndex(0))
        out.update(total=t, free=f, free2=f2, over=over, last=int(x[-1].item()), prop_total=torch.cuda.get_device_properties(0).total_memory,
                   nvml_count=pynvml.nvmlDeviceGetCount(), nvml_total=m.total, nvml_used=m.used)
        print(json.dumps(out))
    """
    uuid = _gpu0_uuid()
    p = run_under_shim(code, {"NVIDIA_DEVICE_MAP": f"0:{uuid} 1:{uuid}", "CUDA_DEVICE_MEMORY_LIMIT_0": "4g", "CUDA_DEVICE_MEMORY_LIMIT_1": "8g"})
    o = last_json(p)
    print("pytorch on a shared GPU:", o)
    assert o["count"] == 1 == o["nvml_count"]
    assert o["total"] == 12 * GiB == o["prop_total"] == o["nvml_total"]
    assert o["over"] != "allocated" and o["last"] == 7 and o["free"] - o["free2"] == 9 * GiB
    assert o["nvml_used"] == 12 * GiB - o["free2"]


def test_subset_map_hides_the_other_gpu_from_driver_runtime_and_nvml(cuda_dev):
    """Container sees two GPUs, NVIDIA_DEVICE_MAP gives it the second: one device everywhere — driver API, the CUDA
    runtime (PyTorch) and NVML — and it IS the second GPU, with the map's quota (needs a 2-GPU box)."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    o = subprocess.run(["nvidia-smi", "--query-gpu=uuid,pci.bus_id", "--format=csv,noheader"], capture_output=True, text=True).stdout
    gpus = [tuple(x.strip() for x in line.split(",")) for line in o.strip().splitlines()]
    uuid1, bus1 = gpus[1]
    code = """
        import ctypes, json, os, subprocess, torch, pynvml
        out = {"count": torch.cuda.device_count()}
        f, t = torch.cuda.mem_get_info()
        x = torch.ones(1 << 28, dtype=torch.uint8, device="cuda")      # 256 MiB on "device 0"
        torch.cuda.synchronize()
        env = {k: v for k, v in os.environ.items() if k != "LD_PRELOAD"}
        smi = subprocess.run(["nvidia-smi", "--query-compute-apps=gpu_uuid,pid", "--format=csv,noheader"], capture_output=True, text=True, env=env).stdout
        cu = ctypes.CDLL("libcuda.so.1")
        libc_getenv = ctypes.CDLL(None).getenv; libc_getenv.restype = ctypes.c_void_p   # os.environ is Python's start-up copy
        n = ctypes.c_int(); cu.cuDeviceGetCount(ctypes.byref(n))
        bus = ctypes.create_string_buffer(32); cu.cuDeviceGetPCIBusId(bus, 32, 0)
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(0)
        u = pynvml.nvmlDeviceGetUUID(h)
        out.update(total=t, sum=int(x.sum().item()), drv_count=n.value, bus=bus.value.decode(), nvml_count=pynvml.nvmlDeviceGetCount(),
                   nvml_uuid=u.decode() if isinstance(u, bytes) else u, nvml_total=pynvml.nvmlDeviceGetMemoryInfo(h).total,
                   where=[l.split(",")[0].strip() for l in smi.strip().splitlines() if l.strip().endswith(str(os.getpid()))],
                   cvd=ctypes.string_at(ctypes.cast(libc_getenv(b"CUDA_VISIBLE_DEVICES"), ctypes.c_char_p)).decode())
        print(json.dumps(out))
    """
    # (a) nothing preset: the shim hides the other GPU; (b) the host's "0,1" handed down: read as the container's own
    # ordinals (virtual 0 exists, virtual 1 does not) and rewritten; either way the runtime must come up
    saved = os.environ.pop("CUDA_VISIBLE_DEVICES", None)
    try:
        for preset in (None, "0,1"):
            extra = {"NVIDIA_DEVICE_MAP": f"0:{uuid1}", "CUDA_DEVICE_MEMORY_LIMIT_0": "6g"}
            if preset:
                extra["CUDA_VISIBLE_DEVICES"] = preset
            p = run_under_shim(code, extra)
            o = last_json(p)
            print("subset map, preset", preset, ":", o)
            assert o["cvd"] == uuid1
            assert o["count"] == 1 == o["drv_count"] == o["nvml_count"]
            assert o["total"] == 6 * GiB == o["nvml_total"] and o["sum"] == 1 << 28
            assert o["nvml_uuid"] == uuid1 and o["bus"].lower().endswith(bus1.lower()[-10:])
            assert o["where"] == [uuid1]                              # the process lives on the second GPU only
    finally:
        if saved is not None:
            os.environ["CUDA_VISIBLE_DEVICES"] = saved


@pytest.mark.parametrize("conf", ["expandable_segments:True", "backend:cudaMallocAsync"])
def test_strict_quota_closes_the_allocator_paths_the_reference_leaves_open(cuda_dev, conf):
    """PyTorch with expandable segments allocates through cuMemCreate + cuMemMap, with backend:cudaMallocAsync through
    cuMemAllocAsync; neither is metered by the reference (nor here by default): a 4 GiB vGPU hands out 6 GiB.  With
    VGPU_STRICT_QUOTA=1 the same program is held to its quota."""
    code = """
        import json, torch
        out = {}
        torch.cuda.init()
        keep = []
        for name in ("first", "second"):       # 3 GiB each: together more than the 4 GiB vGPU
            try:
                x = torch.empty(3 << 30, dtype=torch.uint8, device="cuda"); x[-1] = 1; keep.append(x); out[name] = "allocated"
            except Exception as e:
                out[name] = type(e).__name__
        f, t = torch.cuda.mem_get_info()
        out.update(free=f, total=t)
        print(json.dumps(out))
    """
    env = {"CUDA_DEVICE_MEMORY_LIMIT_0": "4096m", "PYTORCH_CUDA_ALLOC_CONF": conf}
    loose = last_json(run_under_shim(code, env))
    strict = last_json(run_under_shim(code, {**env, "VGPU_STRICT_QUOTA": "1"}))
    print(conf, "default:", loose, "strict:", strict)
    assert loose["first"] == loose["second"] == "allocated"              # the reference's hole, reproduced: 6 GiB in a 4 GiB vGPU
    assert strict["first"] == "allocated" and strict["second"] != "allocated"
    assert strict["total"] == 4 * GiB and 3 * GiB <= strict["total"] - strict["free"] <= 4 * GiB


def test_nvidia_smi_in_the_pod_shows_the_quota(cuda_dev):
    """nvidia-smi itself under the shim (it is what people run inside the pod): the vGPU's quota as total memory — and,
    with the GPU named twice in the map, still ONE row, carrying the sum."""
    uuid = _gpu0_uuid()
    td = tempfile.mkdtemp(prefix="vgpushim")
    q = ["nvidia-smi", "--query-gpu=uuid,memory.total,memory.used", "--format=csv,noheader,nounits"]
    for extra, total_mib in (({"CUDA_DEVICE_MEMORY_LIMIT_0": "8g", "NVIDIA_DEVICE_MAP": f"0:{uuid}"}, 8192),
                             ({"CUDA_DEVICE_MEMORY_LIMIT_0": "4g", "CUDA_DEVICE_MEMORY_LIMIT_1": "8g", "NVIDIA_DEVICE_MAP": f"0:{uuid} 1:{uuid}"}, 12288)):
        env = {**os.environ, "LD_PRELOAD": SHIM, "CUDA_DEVICE_MEMORY_SHARED_CACHE": os.path.join(td, f"c{total_mib}"), "LIBCUDA_LOG_LEVEL": "0", **extra}
        p = subprocess.run(q, env=env, capture_output=True, text=True, timeout=120)
        assert p.returncode == 0, p.stderr[-2000:]
        rows = [[x.strip() for x in line.split(",")] for line in p.stdout.strip().splitlines()]
        print("nvidia-smi under the shim:", rows)
        assert len(rows) == 1 and rows[0][0] == uuid and int(rows[0][1]) == total_mib and int(rows[0][2]) == 0


def test_outside_monitor_closes_the_async_allocator_hole_with_memory_override(cuda_dev):
    """The monitor plane end to end on the real driver: a tenant allocates 3 GiB through cudaMallocAsync (not metered, as
    in the reference).  A monitor process loads libvgpu.so next to the tenant's cache file with MEMORY_OVERRIDE, reads what
    NVML says the tenant's host pid really uses and reports it (set_gpu_device_memory_monitor): from then on the tenant's
    quota counts it — cuMemGetInfo shrinks and a metered 2 GiB cuMemAlloc is refused."""
    td = tempfile.mkdtemp(prefix="vgpushim")
    cache = os.path.join(td, "tenant.cache")
    tenant = subprocess.Popen([sys.executable, "-c", textwrap.dedent("""
        import ctypes, json, sys, torch
        torch.cuda.init()
        a = torch.empty(3 << 30, dtype=torch.uint8, device="cuda"); a[-1] = 1; torch.cuda.synchronize()
        cu = ctypes.CDLL("libcuda.so.1")
        def info():
            f, t = ctypes.c_size_t(), ctypes.c_size_t(); rc = cu.cuMemGetInfo_v2(ctypes.byref(f), ctypes.byref(t)); return rc, f.value, t.value
        print(json.dumps({"before": info()}), flush=True)
        sys.stdin.readline()                          # the monitor does its round
        p = ctypes.c_ulonglong()
        rc = ctypes.c_int(cu.cuMemAlloc_v2(ctypes.byref(p), ctypes.c_size_t(2 << 30))).value
        print(json.dumps({"after": info(), "alloc_2g": rc}), flush=True)
    """)], env={**os.environ, "LD_PRELOAD": SHIM, "CUDA_DEVICE_MEMORY_SHARED_CACHE": cache, "CUDA_DEVICE_MEMORY_LIMIT_0": "4096m",
                "PYTORCH_CUDA_ALLOC_CONF": "backend:cudaMallocAsync", "LIBCUDA_LOG_LEVEL": "0"},
        stdin=subprocess.PIPE, stdout=subprocess.PIPE, text=True)
    try:
        before = json.loads(tenant.stdout.readline())["before"]
        monitor = subprocess.run([sys.executable, "-c", textwrap.dedent("""
            import ctypes, json, os, sys
            sys.path.insert(0, os.environ["VGPU_PKG"])
            import importlib; mon = importlib.import_module("monitor")
            import pynvml; pynvml.nvmlInit()
            lib = ctypes.CDLL(os.environ["VGPU_SHIM"])
            lib.set_gpu_device_memory_monitor.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_uint64]
            lib.get_gpu_memory_usage.restype = ctypes.c_uint64
            r = mon.read_region(os.environ["CUDA_DEVICE_MEMORY_SHARED_CACHE"])
            tenant = [p for p in r["procs"] if p["used"][0]["total"] > 0][0]
            real = {p.pid: p.usedGpuMemory for p in pynvml.nvmlDeviceGetComputeRunningProcesses(pynvml.nvmlDeviceGetHandleByIndex(0))}
            used = real[tenant["hostpid"]]
            rc = lib.set_gpu_device_memory_monitor(tenant["hostpid"], 0, used)
            print(json.dumps({"hostpid": tenant["hostpid"], "counted": tenant["used"][0]["total"], "nvml": used, "rc": rc,
                              "usage_after": lib.get_gpu_memory_usage(0)}))
        """)], env={**os.environ, "CUDA_DEVICE_MEMORY_SHARED_CACHE": cache, "MEMORY_OVERRIDE": "1", "ACTIVE_OOM_KILLER": "0", "LIBCUDA_LOG_LEVEL": "0",
                    "VGPU_PKG": os.path.join(ROOT, "4paradigm-k8s-device-plugin_b200"), "VGPU_SHIM": SHIM}, capture_output=True, text=True, timeout=120)
        assert monitor.returncode == 0, monitor.stderr[-2000:]
        m = json.loads(monitor.stdout.strip().splitlines()[-1])
        tenant.stdin.write("\n"); tenant.stdin.flush()
        after = json.loads(tenant.stdout.readline())
    finally:
        tenant.kill()
    print("monitor round:", m, "tenant before:", before, "after:", after)
    assert m["rc"] == 1 and m["counted"] < 1 * GiB < 3 * GiB < m["nvml"] and m["usage_after"] == m["nvml"]
    assert before[0] == 0 and before[2] == 4 * GiB and before[1] > 3 * GiB           # the 3 GiB went unnoticed
    assert after["after"][0] == 0 and after["after"][1] == 4 * GiB - m["nvml"]       # now they count
    assert after["alloc_2g"] == -1                                                   # and the quota holds


def test_nvml_memory_info_shows_the_limit(cuda_dev):
    code = """
        import json, torch, pynvml
        torch.cuda.init()
        a = torch.empty(1 << 30, dtype=torch.uint8, device="cuda")
        pynvml.nvmlInit()
        m = pynvml.nvmlDeviceGetMemoryInfo(pynvml.nvmlDeviceGetHandleByIndex(0))
        print(json.dumps({"total": m.total, "free": m.free, "used": m.used}))
    """
    o = last_json(run_under_shim(code, {"CUDA_DEVICE_MEMORY_LIMIT_0": "8g"}))
    assert o["total"] == 8 * GiB and o["used"] >= 1 * GiB and o["free"] == o["total"] - o["used"]


def test_two_processes_share_the_quota_through_the_cache_file(cuda_dev):
    cache = os.path.join(tempfile.mkdtemp(prefix="vgpushim"), "shared.cache")
    holder = subprocess.Popen([sys.executable, "-c", textwrap.dedent("""
        import sys, torch
        a = torch.empty(3 << 30, dtype=torch.uint8, device="cuda")
        print("holding", flush=True)
        sys.stdin.readline()
    """)], env={**os.environ, "LD_PRELOAD": SHIM, "CUDA_DEVICE_MEMORY_SHARED_CACHE": cache, "CUDA_DEVICE_MEMORY_LIMIT_0": "6g",
               "LIBCUDA_LOG_LEVEL": "0"}, stdin=subprocess.PIPE, stdout=subprocess.PIPE, text=True)
    try:
        assert holder.stdout.readline().strip() == "holding"
        code = """
            import json, torch
            f, t = torch.cuda.mem_get_info()
            try:
                b = torch.empty((3 << 30) + (2 << 20), dtype=torch.uint8, device="cuda"); r = "allocated"
            except Exception as e:
                r = type(e).__name__
            print(json.dumps({"free": f, "total": t, "second": r}))
        """
        # 6 GiB shared: the holder's 3 GiB plus two contexts (billed like the reference does) leave at most 3 GiB
        o = last_json(run_under_shim(code, {"CUDA_DEVICE_MEMORY_LIMIT_0": "6g"}, cache=cache))
        assert o["total"] == 6 * GiB and o["free"] <= 3 * GiB and o["second"] != "allocated"
    finally:
        holder.stdin.write("\n")
        holder.stdin.flush()
        holder.wait(timeout=60)
    # after the holder exits its slot is cleared: a new process sees the whole limit again
    code = "import json, torch; f, t = torch.cuda.mem_get_info(); print(json.dumps({'free': f, 'total': t}))"
    o = last_json(run_under_shim(code, {"CUDA_DEVICE_MEMORY_LIMIT_0": "6g"}, cache=cache))
    assert o["total"] == 6 * GiB and o["free"] > 4 * GiB


def test_sm_limit_becomes_a_green_context_partition(cuda_dev):
    code = """
        import ctypes, json, time, torch
        torch.cuda.init()
        x = torch.randn(8192, 8192, device="cuda", dtype=torch.bfloat16)
        for _ in range(10):                      # warm-up: cuBLAS heuristics, clocks
            y = x @ x
        torch.cuda.synchronize()
        best = 1e9
        for rep in range(3):                     # best of 3 x 20 GEMMs, timed on the device
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(20):
                y = x @ x
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1) / 20 * 1e-3)
        dt = best
        ref = (x[:64].float() @ x.float())
        ok = bool(torch.allclose(y[:64].float(), ref, rtol=2e-2, atol=2.0))
        sms = ctypes.CDLL(None).cuVGPUPartitionSmCount(0)
        print(json.dumps({"sms": sms, "tflops": 2 * 8192**3 / dt / 1e12, "ok": ok,
                          "mp": torch.cuda.get_device_properties(0).multi_processor_count}))
    """
    capped = last_json(run_under_shim(code, {"CUDA_DEVICE_SM_LIMIT": "25", "CUDA_DEVICE_MEMORY_LIMIT_0": "16g"}))
    free = last_json(run_under_shim(code, {"CUDA_DEVICE_SM_LIMIT": "100", "CUDA_DEVICE_MEMORY_LIMIT_0": "16g"}))
    assert capped["ok"] and free["ok"]
    assert free["sms"] == 0 and 0 < capped["sms"] <= 37            # 25 % of 148, granule-rounded down (32)
    assert capped["mp"] == capped["sms"] and free["mp"] >= 100     # the SM-count attribute is the partition's inside one
    # spatial cap: a compute-bound GEMM gets roughly sms/148 of the machine, not all of it
    assert capped["tflops"] < 0.6 * free["tflops"], (capped, free)


def test_core_utilization_policy_disable_lifts_the_sm_cap(cuda_dev):
    """GPU_CORE_UTILIZATION_POLICY=DISABLE (reference: set_env_utilization_switch@0x44b3b turns the throttle off): no
    green-context partition; FORCE, like no policy at all, keeps the 25 % cap."""
    code = """
        import ctypes, json, torch
        torch.cuda.init()
        torch.ones(1, device="cuda").sum().item()
        print(json.dumps({"sms": ctypes.CDLL(None).cuVGPUPartitionSmCount(0)}))
    """
    base = {"CUDA_DEVICE_SM_LIMIT": "25", "CUDA_DEVICE_MEMORY_LIMIT_0": "8g"}
    assert last_json(run_under_shim(code, {**base, "GPU_CORE_UTILIZATION_POLICY": "DISABLE"}))["sms"] == 0
    assert last_json(run_under_shim(code, {**base, "GPU_CORE_UTILIZATION_POLICY": "disable"}))["sms"] == 0
    assert 0 < last_json(run_under_shim(code, {**base, "GPU_CORE_UTILIZATION_POLICY": "FORCE"}))["sms"] <= 37
    assert 0 < last_json(run_under_shim(code, base))["sms"] <= 37


def test_vgpuvalidator_cli_memtest_and_quota(cuda_dev):
    exe = os.path.join(ROOT, "4paradigm-k8s-device-plugin_b200", "vgpuvalidator")
    p = subprocess.run([exe, "--memtest", "--gib", "1", "--random-passes", "2"], capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stderr
    o = json.loads(p.stdout.strip().splitlines()[-1])
    assert o["mismatches"] == 0 and o["passes"] == 66 and o["bytes"] == GiB and o["GBps"] > 100
    p = subprocess.run([exe, "--roundtrip", "--chunks", "32"], capture_output=True, text=True, timeout=300)
    assert p.returncode == 0 and json.loads(p.stdout.strip().splitlines()[-1])["bad_chunks"] == 0, p.stderr
    # inside a vGPU (under the shim) the validator cannot test more memory than the quota allows
    td = tempfile.mkdtemp(prefix="vgpushim")
    env = {**os.environ, "LD_PRELOAD": SHIM, "CUDA_DEVICE_MEMORY_SHARED_CACHE": os.path.join(td, "c"), "CUDA_DEVICE_MEMORY_LIMIT_0": "2g", "LIBCUDA_LOG_LEVEL": "0"}
    p = subprocess.run([exe, "--memtest", "--gib", "4", "--patterns", "random", "--random-passes", "1"], capture_output=True, text=True, timeout=300, env=env)
    assert p.returncode == 2 and "Device 0 OOM" in p.stderr
    p = subprocess.run([exe, "--memtest", "--gib", "1", "--patterns", "random", "--random-passes", "1"], capture_output=True, text=True, timeout=300, env=env)
    assert p.returncode == 0, p.stderr


def test_oversubscribe_vmm_backend_suspend_releases_hbm_and_resume_restores(cuda_dev):
    """CUDA_OVERSUBSCRIBE=true + VGPU_OVERSUB_BACKEND=vmm: large allocations are VMM-backed; a
    suspend spills them to pinned host RAM with the spill kernel and gives the HBM back to the
    device, a resume restores them behind the same pointers, checksum-verified."""
    code = """
        import ctypes, json, torch
        shim = ctypes.CDLL(None)
        real = ctypes.CDLL("libcuda.so.1")
        torch.cuda.init()
        a = torch.empty(6 << 30, dtype=torch.uint8, device="cuda")
        a.view(torch.int64).copy_(torch.arange(a.numel() // 8, device="cuda", dtype=torch.int64) * 2654435761)
        s0 = int(a.view(torch.int64).sum().item())
        small = torch.ones(1024, device="cuda")               # <= 2 MiB: plain cuMemAlloc, stays put
        torch.cuda.synchronize()
        torch.cuda.empty_cache()                              # drop the temporaries' cached segments
        import os, subprocess
        def used_mib():
            # physical view: run nvidia-smi WITHOUT the shim (under it, NVML shows the vGPU's quota usage)
            env = {k: v for k, v in os.environ.items() if k != "LD_PRELOAD"}
            o = subprocess.run(["nvidia-smi", "--query-gpu=memory.used", "--format=csv,noheader,nounits", "-i", "0"], capture_output=True, text=True, env=env).stdout
            return int(o.strip().splitlines()[0])
        u0 = used_mib()
        rc1 = shim.cuVGPUSuspend()
        u1 = used_mib()
        rc2 = shim.cuVGPUResume()
        u2 = used_mib()
        s1 = int(a.view(torch.int64).sum().item())
        print(json.dumps({"rc": [rc1, rc2], "used_mib": [u0, u1, u2], "same": s0 == s1, "small": float(small.sum())}))
    """
    p = run_under_shim(code, {"CUDA_DEVICE_MEMORY_LIMIT_0": "40g", "CUDA_OVERSUBSCRIBE": "true", "VGPU_OVERSUB_BACKEND": "vmm"})
    o = last_json(p)
    assert o["rc"] == [0, 0], p.stderr[-2000:]
    u0, u1, u2 = o["used_mib"]
    assert u0 - u1 >= 6 * 1024 - 64, o      # suspend released (at least) the 6 GiB of HBM behind `a`
    assert u2 - u1 >= 6 * 1024 - 64, o      # resume took it back
    assert o["same"] and o["small"] == 1024.0


def test_prefetch_calls_page_vmm_backed_allocations_out_and_in(cuda_dev):
    """With the VMM backend the prefetch calls of a UVM-style program are its paging controls: prefetch to the CPU (device
    -1) spills the allocation to pinned host RAM and gives the HBM back, prefetch to the device brings it back behind the
    same pointer, checksum-verified.  A prefetch of part of an allocation towards the CPU is a hint that is not taken."""
    code = """
        import ctypes, json, os, subprocess, torch
        cu = ctypes.CDLL("libcuda.so.1")
        torch.cuda.init()
        n = 4 << 30
        a = torch.empty(n, dtype=torch.uint8, device="cuda")
        a.view(torch.int64).copy_(torch.arange(n // 8, device="cuda", dtype=torch.int64) * 2654435761)
        s0 = int(a.view(torch.int64).sum().item())
        torch.cuda.synchronize(); torch.cuda.empty_cache()
        def used_mib():
            env = {k: v for k, v in os.environ.items() if k != "LD_PRELOAD"}
            o = subprocess.run(["nvidia-smi", "--query-gpu=memory.used", "--format=csv,noheader,nounits", "-i", "0"], capture_output=True, text=True, env=env).stdout
            return int(o.strip().splitlines()[0])
        def prefetch(ptr, size, dev):
            return cu.cuMemPrefetchAsync(ctypes.c_ulonglong(ptr), ctypes.c_size_t(size), ctypes.c_int(dev), None)
        u0 = used_mib()
        rc_part = prefetch(a.data_ptr() + (1 << 30), 1 << 30, -1); u_part = used_mib()     # part of it: not taken
        rc_out = prefetch(a.data_ptr(), n, -1); u1 = used_mib()
        rc_in = prefetch(a.data_ptr() + (1 << 20), 1 << 20, 0); u2 = used_mib()            # any part towards the device: its blocks come back
        rc_all = prefetch(a.data_ptr(), n, 0); u3 = used_mib()
        s1 = int(a.view(torch.int64).sum().item())
        print(json.dumps({"rc": [rc_part, rc_out, rc_in, rc_all], "used_mib": [u0, u_part, u1, u2, u3], "same": s0 == s1}))
    """
    p = run_under_shim(code, {"CUDA_DEVICE_MEMORY_LIMIT_0": "40g", "CUDA_OVERSUBSCRIBE": "true", "VGPU_OVERSUB_BACKEND": "vmm"})
    o = last_json(p)
    print("prefetch paging:", o)
    assert o["rc"] == [0, 0, 0, 0], p.stderr[-2000:]
    u0, u_part, u1, u2, u3 = o["used_mib"]
    assert abs(u_part - u0) < 64                     # the partial prefetch towards the CPU moved nothing
    assert u0 - u1 >= 4 * 1024 - 64                  # the whole-allocation one released the 4 GiB
    assert 0 < u2 - u1 <= 4 * 1024 + 64              # a block came back
    assert u3 - u1 >= 4 * 1024 - 64 and o["same"]    # all of it, bit-exact


def test_sigusr2_suspends_and_sigusr1_resumes_a_running_vgpu_process(cuda_dev):
    """The reference's monitor protocol (SIGUSR2 -> suspended, SIGUSR1 -> running) now also moves
    the data: while suspended the process holds (almost) no HBM; after resume it carries on."""
    import signal, time
    td = tempfile.mkdtemp(prefix="vgpushim")
    env = {**os.environ, "LD_PRELOAD": SHIM, "CUDA_DEVICE_MEMORY_SHARED_CACHE": os.path.join(td, "c"), "LIBCUDA_LOG_LEVEL": "0",
           "CUDA_DEVICE_MEMORY_LIMIT_0": "40g", "CUDA_OVERSUBSCRIBE": "true", "VGPU_OVERSUB_BACKEND": "vmm"}
    code = textwrap.dedent("""
        import json, sys, torch
        a = torch.empty(4 << 30, dtype=torch.uint8, device="cuda")
        a.view(torch.int64).copy_(torch.arange(a.numel() // 8, device="cuda", dtype=torch.int64) * 40503)
        s0 = int(a.view(torch.int64).sum().item())
        torch.cuda.empty_cache()
        print("ready", flush=True)
        n = 0
        while True:                      # keeps launching kernels; parks at the gate while suspended
            line = sys.stdin.readline().strip()
            if line == "stop":
                break
            s = int(a.view(torch.int64)[-1].item()) if line == "copy" else 0   # plain D2H copy: gated like launches
            s = int(a.view(torch.int64).sum().item()); n += 1
            print(json.dumps({"same": s == s0, "n": n}), flush=True)
    """)
    p = subprocess.Popen([sys.executable, "-c", code], env=env, stdin=subprocess.PIPE, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)

    def used_mib():
        o = subprocess.run(["nvidia-smi", "--query-gpu=memory.used", "--format=csv,noheader,nounits", "-i", "0"], capture_output=True, text=True).stdout
        return int(o.strip().splitlines()[0])

    def poke():
        p.stdin.write("go\n"); p.stdin.flush()
        return json.loads(p.stdout.readline())

    try:
        assert p.stdout.readline().strip() == "ready"
        assert poke()["same"]
        u0 = used_mib()
        os.kill(p.pid, signal.SIGUSR2)
        deadline = time.time() + 30
        while time.time() < deadline and u0 - used_mib() < 4 * 1024 - 64:
            time.sleep(0.2)
        u1 = used_mib()
        assert u0 - u1 >= 4 * 1024 - 64, (u0, u1)     # suspended: the 4 GiB went to pinned host RAM
        p.stdin.write("copy\n"); p.stdin.flush()       # issued WHILE suspended: must park, not fault
        time.sleep(1.0)
        assert p.poll() is None
        os.kill(p.pid, signal.SIGUSR1)
        r = json.loads(p.stdout.readline())            # the parked request completes once the data is back
        assert r["same"]
        r = poke()
        assert r["same"]
        assert used_mib() - u1 >= 4 * 1024 - 64
    finally:
        try:
            p.stdin.write("stop\n"); p.stdin.flush()
            p.wait(timeout=30)
        except Exception:
            p.kill()


def _validator_smids(env_extra):
    td = tempfile.mkdtemp(prefix="vgpusm")
    env = dict(os.environ)
    env.update({"LD_PRELOAD": SHIM, "CUDA_DEVICE_MEMORY_SHARED_CACHE": os.path.join(td, "shr.cache"), "LIBCUDA_LOG_LEVEL": "0",
                "CUDA_DEVICE_SM_LIMIT": "25", "CUDA_DEVICE_MEMORY_LIMIT_0": "8g"})
    env.update(env_extra)
    exe = os.path.join(ROOT, "4paradigm-k8s-device-plugin_b200", "vgpuvalidator")
    p = subprocess.run([exe, "--smids"], env=env, capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stderr[-2000:]
    return json.loads(p.stdout.strip().splitlines()[-1]), p.stderr


def test_partition_index_from_the_plugin_gives_containers_disjoint_sms(cuda_dev):
    """Four containers of one GPU (split 4 -> CUDA_DEVICE_SM_LIMIT=25), each with the VGPU_SM_PARTITION_INDEX_0 the plugin
    emits when asked to (vgpu_plugin_cfg_t.emit_partition_index): the SMs their kernels really run on (%smid, read by
    `vgpuvalidator --smids`) are four disjoint sets of at most 25 % of the GPU.  Without the variable nothing tells a
    container which vGPU of the GPU it is: both land on group 0 and the shim says that placement is not coordinated."""
    sets = []
    for i in range(4):
        o, err = _validator_smids({"VGPU_SM_PARTITION_INDEX_0": str(i)})
        assert 0 < o["sm_count"] <= 37 and "NOT coordinated" not in err
        sets.append(set(o["sm_ids"]))
    for i in range(4):
        for j in range(i):
            assert not (sets[i] & sets[j]), f"vGPU {i} and vGPU {j} share SMs {sorted(sets[i] & sets[j])}"
    assert sum(len(s) for s in sets) >= 128
    # the validator's kernels use no clusters: it takes the SM-granular split (36 of 148 for 25 %) unless told otherwise
    assert len(sets[0]) == 36
    coarse, _ = _validator_smids({"VGPU_SM_PARTITION_INDEX_0": "0", "VGPU_SM_PARTITION_GRANULE": "coarse"})
    assert coarse["sm_count"] == 32
    a, err_a = _validator_smids({})
    b, err_b = _validator_smids({})
    assert set(a["sm_ids"]) == set(b["sm_ids"]) and "NOT coordinated" in err_a and "NOT coordinated" in err_b
    # the legacy single variable still works for every device
    c, err_c = _validator_smids({"VGPU_SM_PARTITION_INDEX": "2"})
    assert set(c["sm_ids"]) == sets[2] and "NOT coordinated" not in err_c


OVERSUB_ENV = {"CUDA_OVERSUBSCRIBE": "true", "VGPU_OVERSUB_BACKEND": "vmm"}
OVERSUB_WALK = """
    import ctypes, json, torch
    torch.cuda.init()
    G = 1 << 30
    N, SIZE = 10, 2 * G                       # 20 GiB of tensors; at most 6 GiB of them may be resident at once
    ts = []
    for i in range(N):
        t = torch.empty(SIZE // 8, dtype=torch.int64, device="cuda")   # one cudaMalloc each (> the allocator's pooling sizes)
        t.copy_(torch.arange(SIZE // 8, device="cuda", dtype=torch.int64) * (2654435761 + i))   # a launch that names t
        ts.append(t)
    want = [int(((torch.arange(1 << 20, dtype=torch.int64) * (2654435761 + i)).sum()).item()) for i in range(N)]
    ok = True
    for rnd in range(2):                      # two passes over everything, the second backwards
        for i in (range(N) if rnd == 0 else reversed(range(N))):
            ok &= int(ts[i][: 1 << 20].sum().item()) == want[i]      # a launch whose arguments point into ts[i]
            ok &= int(ts[i][-1].item()) == (SIZE // 8 - 1) * (2654435761 + i)   # a plain D2H copy of its last bytes
    two = int((ts[0][: 1 << 20] + ts[N - 1][: 1 << 20]).sum().item()) == want[0] + want[N - 1]   # one kernel, two ranges
    free, total = torch.cuda.mem_get_info()
    st = (ctypes.c_uint64 * 8)()
    rc = ctypes.CDLL(None).cuVGPUOversubStats(st)
    print(json.dumps({"ok": bool(ok), "two": bool(two), "rc": rc, "to_peer": st[0], "to_host": st[1], "restored": st[2], "bad": st[3],
                      "resident": st[4], "managed": st[5], "paged": st[6], "free": free, "total": total}))
"""


def test_an_unmodified_program_larger_than_its_resident_share_is_paged_by_the_shim(cuda_dev):
    """Transparent oversubscription (the reference: cuMemoryAllocate@0x32146 -> cuMemAllocManaged, UVM faults).  A PyTorch
    program whose tensors add up to 20 GiB runs with 6 GiB of resident HBM for its managed ranges: nothing in the program
    knows — the shim finds the ranges a launch or copy names (kernel arguments, copy operands), pages them in and spills
    the least recently used ones to pinned host RAM with the spill / restore kernels.  Every byte checked."""
    p = run_under_shim(OVERSUB_WALK, {**OVERSUB_ENV, "CUDA_DEVICE_MEMORY_LIMIT_0": "30g", "VGPU_OVERSUB_RESIDENT_MB": "6144"}, timeout=900)
    o = last_json(p)
    print("transparent oversubscription:", o)
    assert o["ok"] and o["two"] and o["rc"] == 0 and o["bad"] == 0, p.stderr[-3000:]
    assert o["paged"] == 1 and o["to_host"] >= 7 * 1024 and o["restored"] >= 7 * 1024 and o["to_peer"] == 0
    assert o["resident"] * 2 <= 6144 and o["managed"] * 2 >= 20 * 1024      # chunks of 2 MiB
    assert o["total"] == 30 << 30 and o["total"] - o["free"] >= 20 << 30    # the quota still counts every byte, resident or not
    assert "Device 0 OOM" not in p.stderr


def test_oversubscription_spills_to_the_named_peer_gpu_first_then_to_host(cuda_dev):
    """north_star tier order through the shim: VGPU_OVERSUB_PEER names an idle GPU by UUID; the shim keeps it open to the
    driver (P2P needs both ends) and away from the application, spills there while the peer budget lasts, then to host."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("the peer tier needs 2 GPUs")
    import pynvml
    pynvml.nvmlInit()
    uu = [pynvml.nvmlDeviceGetUUID(pynvml.nvmlDeviceGetHandleByIndex(i)) for i in range(2)]
    uu = [u.decode() if isinstance(u, bytes) else u for u in uu]
    env = {**OVERSUB_ENV, "CUDA_DEVICE_MEMORY_LIMIT_0": "30g", "VGPU_OVERSUB_RESIDENT_MB": "6144", "NVIDIA_DEVICE_MAP": f"0:{uu[0]}",
           "VGPU_OVERSUB_PEER": uu[1], "VGPU_OVERSUB_PEER_MB": "4096"}
    p = run_under_shim(OVERSUB_WALK, env, timeout=900)
    o = last_json(p)
    print("oversubscription, peer then host:", o)
    assert o["ok"] and o["two"] and o["bad"] == 0, p.stderr[-3000:]
    assert o["to_peer"] >= 1024 and o["to_host"] >= 1024 and o["restored"] >= 7 * 1024
    # the application was not given the peer: NVML (what frameworks count GPUs with) shows the vGPU only
    q = run_under_shim("import torch, json; print(json.dumps({'n': torch.cuda.device_count()}))", env)
    assert last_json(q)["n"] == 1


def test_vmm_backend_leaves_a_cuCtxCreate_context_current_and_bills_its_device(cuda_dev):
    """Driver-API program with a context of its own (not the primary one): managed allocations, a suspend and a resume
    must leave that context current and the data intact (the manager works in the primary context internally)."""
    code = """
        import ctypes as C, json
        cu = C.CDLL("libcuda.so.1")
        assert cu.cuInit(0) == 0
        dev, ctx, cur = C.c_int(), C.c_void_p(), C.c_void_p()
        assert cu.cuDeviceGet(C.byref(dev), 0) == 0
        assert cu.cuCtxCreate_v2(C.byref(ctx), 0, dev) == 0
        n = 64 << 20
        p = C.c_ulonglong()
        rc_alloc = cu.cuMemAlloc_v2(C.byref(p), C.c_size_t(n))
        cu.cuCtxGetCurrent(C.byref(cur)); same1 = cur.value == ctx.value
        assert cu.cuMemsetD32_v2(p, 0x5EED1234, C.c_size_t(n // 4)) == 0
        shim = C.CDLL(None)                      # the preloaded library's own entry points
        rc_out = shim.cuVGPUSuspend()
        cu.cuCtxGetCurrent(C.byref(cur)); same2 = cur.value == ctx.value
        rc_in = shim.cuVGPUResume()
        cu.cuCtxGetCurrent(C.byref(cur)); same3 = cur.value == ctx.value
        host = (C.c_uint32 * (n // 4))()
        rc_copy = cu.cuMemcpyDtoH_v2(host, p, C.c_size_t(n))
        good = all(host[i] == 0x5EED1234 for i in range(0, n // 4, 4099))
        free, total = C.c_size_t(), C.c_size_t()
        cu.cuMemGetInfo_v2(C.byref(free), C.byref(total))
        rc_free = cu.cuMemFree_v2(p)
        print(json.dumps({"rc": [rc_alloc, rc_out, rc_in, rc_copy, rc_free], "same": [same1, same2, same3], "good": good,
                          "billed": total.value - free.value}))
    """
    p = run_under_shim(code, {**OVERSUB_ENV, "CUDA_DEVICE_MEMORY_LIMIT_0": "8g"})
    o = last_json(p)
    assert o["rc"] == [0, 0, 0, 0, 0] and o["same"] == [True, True, True] and o["good"], (o, p.stderr[-2000:])
    assert o["billed"] >= 64 << 20


def test_cooperative_grid_sized_from_the_sm_count_fits_inside_a_partition(cuda_dev):
    """A cooperative launch must be co-resident: occupancy x SM count CTAs at most.  Inside a 25 % partition the SM-count
    attribute answers with the partition's size, so a grid sized the usual way is accepted; sized from the GPU's full
    148 SMs it is refused (CUDA_ERROR_COOPERATIVE_LAUNCH_TOO_LARGE = 720) — which is what programs hit when the
    attribute still said 148.  (The reference's temporal throttle never had this failure mode.)"""
    code = """
        import ctypes as C, json
        from cuda.bindings import nvrtc
        src = b'extern "C" __global__ void k(int *out) { if (threadIdx.x == 0) atomicAdd(out, 1); }'
        err, prog = nvrtc.nvrtcCreateProgram(src, b"k.cu", 0, [], [])
        (err,) = nvrtc.nvrtcCompileProgram(prog, 1, [b"--gpu-architecture=compute_100a"])
        assert int(err) == 0
        err, n = nvrtc.nvrtcGetPTXSize(prog)
        ptx = b" " * n
        (err,) = nvrtc.nvrtcGetPTX(prog, ptx)
        cu = C.CDLL("libcuda.so.1")
        assert cu.cuInit(0) == 0
        dev, ctx, mod, fn = C.c_int(), C.c_void_p(), C.c_void_p(), C.c_void_p()
        cu.cuDeviceGet(C.byref(dev), 0)
        assert cu.cuDevicePrimaryCtxRetain(C.byref(ctx), dev) == 0 and cu.cuCtxSetCurrent(ctx) == 0
        sms = C.c_int()
        cu.cuDeviceGetAttribute(C.byref(sms), 16, dev)
        assert cu.cuModuleLoadData(C.byref(mod), ptx) == 0 and cu.cuModuleGetFunction(C.byref(fn), mod, b"k") == 0
        occ = C.c_int()
        assert cu.cuOccupancyMaxActiveBlocksPerMultiprocessor(C.byref(occ), fn, 256, C.c_size_t(0)) == 0
        out = C.c_ulonglong()
        assert cu.cuMemAlloc_v2(C.byref(out), C.c_size_t(4)) == 0 and cu.cuMemsetD32_v2(out, 0, C.c_size_t(1)) == 0
        arg = C.c_ulonglong(out.value)
        params = (C.c_void_p * 1)(C.cast(C.byref(arg), C.c_void_p))
        def coop(grid):
            rc = cu.cuLaunchCooperativeKernel(fn, grid, 1, 1, 256, 1, 1, 0, None, params)
            return rc if rc else cu.cuCtxSynchronize()
        rc_fit = coop(occ.value * sms.value)
        rc_big = coop(occ.value * 148)
        got = C.c_int()
        cu.cuMemcpyDtoH_v2(C.byref(got), out, C.c_size_t(4))
        print(json.dumps({"sms": sms.value, "occ": occ.value, "rc_fit": rc_fit, "rc_big": rc_big, "ctas_ran": got.value,
                          "partition": C.CDLL(None).cuVGPUPartitionSmCount(0)}))
    """
    o = last_json(run_under_shim(code, {"CUDA_DEVICE_SM_LIMIT": "25", "CUDA_DEVICE_MEMORY_LIMIT_0": "4g"}))
    assert o["partition"] == o["sms"] == 32 and o["occ"] >= 1
    assert o["rc_fit"] == 0 and o["ctas_ran"] == o["occ"] * 32
    assert o["rc_big"] == 720
    free = last_json(run_under_shim(code, {"CUDA_DEVICE_SM_LIMIT": "100", "CUDA_DEVICE_MEMORY_LIMIT_0": "4g"}))
    assert free["partition"] == 0 and free["sms"] == 148 and free["rc_fit"] == 0 and free["rc_big"] == 0


def test_preferred_location_advice_decides_who_keeps_the_hbm(cuda_dev):
    """cuMemAdvise(SET_PREFERRED_LOCATION) on a managed range: towards the GPU it is spilled last, towards the CPU first —
    ahead of the least-recently-used order.  Four 1 GiB ranges, 2.5 GiB may be resident."""
    code = """
        import ctypes as C, json, torch
        cu, shim = C.CDLL("libcuda.so.1"), C.CDLL(None)
        shim.cuVGPUResidentBytes.restype = C.c_longlong
        torch.cuda.init()
        G = 1 << 30
        ts = [torch.empty(G // 8, dtype=torch.int64, device="cuda") for _ in range(4)]
        def advise(t, advice, dev):
            return cu.cuMemAdvise(C.c_ulonglong(t.data_ptr()), C.c_size_t(G), C.c_int(advice), C.c_int(dev))
        def resident():
            return [shim.cuVGPUResidentBytes(C.c_ulonglong(t.data_ptr())) // (1 << 20) for t in ts]
        rc = [advise(ts[0], 3, 0), advise(ts[3], 3, -1)]      # range 0 prefers the GPU, range 3 the CPU
        for i, t in enumerate(ts):
            t.fill_(i + 1)                                     # touched in order 0, 1, 2, 3: range 0 is the oldest
        torch.cuda.synchronize()
        after_fill = resident()
        s = int(ts[1][:1024].sum().item())                     # needs range 1 back: by age the victim would be range 0, by advice it is range 3
        after_touch = resident()
        ok = all(int(t[-1].item()) == i + 1 for i, t in enumerate(ts))
        print(json.dumps({"rc": rc, "after_fill": after_fill, "after_touch": after_touch, "sum": s, "ok": ok}))
    """
    o = last_json(run_under_shim(code, {**OVERSUB_ENV, "CUDA_DEVICE_MEMORY_LIMIT_0": "8g", "VGPU_OVERSUB_RESIDENT_MB": "2560"}))
    print("advice:", o)
    assert o["rc"] == [0, 0] and o["ok"] and o["sum"] == 2 * 1024
    # after the fills: range 0 (the oldest, but it prefers the GPU) is still there; ranges 1 and 2 went instead
    assert o["after_fill"] == [1024, 0, 0, 1024]
    # bringing range 1 back spills range 3 (just used, but it prefers the CPU) and leaves range 0 (the oldest) alone
    assert o["after_touch"] == [1024, 1024, 0, 0]



@pytest.mark.gpu
def test_range_lookup_with_hundreds_of_ranges_and_a_stream_ordered_free(cuda_dev):
    """The argument scan finds a range by binary search in an address-sorted index: with 300 ranges alive, some freed
    in between and new ones made (index insert / delete), a paged process still reads back what it wrote to every
    range; and cuMemFreeAsync on a managed range (legal for memory from cuMemAlloc) frees it like cuMemFree."""
    code = """
        import ctypes as C, json, torch
        cu, shim = C.CDLL("libcuda.so.1"), C.CDLL(None)
        torch.cuda.init()
        M = 1 << 20
        ts = [torch.full((12 * M // 8,), i, dtype=torch.int64, device="cuda") for i in range(300)]   # 3.5 GiB, 1.5 GiB may be resident
        for i in range(0, 300, 3):
            ts[i] = None
        torch.cuda.empty_cache()                                   # 100 ranges freed: holes in the index
        for i in range(0, 300, 3):
            ts[i] = torch.full((14 * M // 8,), 1000 + i, dtype=torch.int64, device="cuda")
        bad = 0
        for rnd in range(2):
            for i, t in enumerate(ts):
                want = 1000 + i if i % 3 == 0 else i
                bad += int(t[0].item()) != want or int(t[-1].item()) != want
        st = (C.c_uint64 * 8)()
        shim.cuVGPUOversubStats(st)
        p = C.c_ulonglong(0)
        rc_alloc = cu.cuMemAlloc_v2(C.byref(p), C.c_size_t(64 * M))
        free0, tot = C.c_size_t(0), C.c_size_t(0)
        cu.cuMemGetInfo_v2(C.byref(free0), C.byref(tot))
        rc_free = cu.cuMemFreeAsync(p, None)
        free1 = C.c_size_t(0)
        cu.cuMemGetInfo_v2(C.byref(free1), C.byref(tot))
        rc_again = cu.cuMemFree_v2(p)                              # already gone
        print(json.dumps({"bad": bad, "paged": int(st[6]), "rc": [rc_alloc, rc_free], "again": rc_again, "freed_mib": (free1.value - free0.value) // M}))
    """
    o = last_json(run_under_shim(code, {**OVERSUB_ENV, "CUDA_DEVICE_MEMORY_LIMIT_0": "8g", "VGPU_OVERSUB_RESIDENT_MB": "1536"}, timeout=600))
    print("many ranges:", o)
    assert o["bad"] == 0 and o["paged"] == 1
    assert o["rc"] == [0, 0] and o["again"] != 0 and o["freed_mib"] == 64




def test_a_launch_or_graph_that_names_more_than_64_ranges_gets_all_of_them(cuda_dev):
    """A multi-tensor kernel passes a hundred pointers in one argument struct, and a graph replay needs every range of
    the process: 100 ranges of 12 MiB are spilled by touching 1.2 GiB of other tensors (1.6 GiB may be resident), then
    (a) one torch._foreach_add_ over all 100 and (b) a replay of a graph captured over them must find every one back."""
    code = """
        import ctypes as C, json, torch
        torch.cuda.init()
        M = 1 << 20
        shim = C.CDLL(None)
        shim.cuVGPUResidentBytes.restype = C.c_longlong
        ts = [torch.zeros(12 * M // 8, dtype=torch.int64, device="cuda") for _ in range(100)]
        def resident():
            return sum(shim.cuVGPUResidentBytes(C.c_ulonglong(t.data_ptr())) > 0 for t in ts)
        def push_out():
            big = [torch.ones(600 * M // 8, dtype=torch.int64, device="cuda") for _ in range(2)]
            s = sum(int(b[:1024].sum().item()) for b in big)
            del big
            torch.cuda.empty_cache()
            return s
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):                       # warm-up on the capture stream, everything resident
            for t in ts:
                t.add_(1)
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            for t in ts:
                t.add_(1)                                # capture only records: values stay 1
        push_out()
        left_a = resident()
        torch._foreach_add_(ts, 10)                      # multi-tensor apply: pointer lists inside one struct
        torch.cuda.synchronize()
        after_a = resident()
        push_out()
        left_b = resident()
        g.replay()
        torch.cuda.synchronize()
        after_b = resident()
        vals = sorted({int(t[0].item()) for t in ts} | {int(t[-1].item()) for t in ts})
        st = (C.c_uint64 * 8)()
        shim.cuVGPUOversubStats(st)
        print(json.dumps({"left": [left_a, left_b], "after": [after_a, after_b], "vals": vals, "bad": int(st[3]), "paged": int(st[6])}))
    """
    o = last_json(run_under_shim(code, {**OVERSUB_ENV, "CUDA_DEVICE_MEMORY_LIMIT_0": "8g", "VGPU_OVERSUB_RESIDENT_MB": "1600"}, timeout=600))
    print("more than 64 ranges:", o)
    assert o["paged"] == 1 and o["bad"] == 0
    assert max(o["left"]) <= 40                 # most of the 100 really were spilled before each of the two launches
    assert o["after"] == [100, 100]             # ... and every one was back for it
    assert o["vals"] == [12]                    # 1 (warm-up) + 10 (foreach) + 1 (replay), first and last word of each



def test_two_host_threads_whose_working_sets_do_not_fit_together(cuda_dev):
    """Two threads of one paged process, each adding to its own 1 GiB tensor, 1.5 GiB may be resident: every launch of
    one thread has to push the other thread's tensor out — after draining that thread's queued kernels.  Both tensors end
    with exactly the number of additions made, on their first and last words and in total."""
    code = """
        import ctypes as C, json, threading, torch
        torch.cuda.init()
        G = 1 << 30
        ts = [torch.zeros(G // 8, dtype=torch.int64, device="cuda") for _ in range(2)]
        N = 25
        errs = []
        def work(k):
            try:
                st = torch.cuda.Stream()
                with torch.cuda.stream(st):
                    for i in range(N):
                        ts[k].add_(k + 1)
                    st.synchronize()
            except Exception as e:
                errs.append(repr(e))
        th = [threading.Thread(target=work, args=(k,)) for k in range(2)]
        [t.start() for t in th]
        [t.join() for t in th]
        torch.cuda.synchronize()
        out = [[int(t[0].item()), int(t[-1].item()), int(t.sum().item()) // (G // 8)] for t in ts]
        st = (C.c_uint64 * 8)()
        C.CDLL(None).cuVGPUOversubStats(st)
        print(json.dumps({"out": out, "errs": errs, "to_host": int(st[1]), "bad": int(st[3]), "paged": int(st[6])}))
    """
    o = last_json(run_under_shim(code, {**OVERSUB_ENV, "CUDA_DEVICE_MEMORY_LIMIT_0": "8g", "VGPU_OVERSUB_RESIDENT_MB": "1536"}, timeout=600))
    print("two threads:", o)
    assert not o["errs"] and o["paged"] == 1 and o["bad"] == 0
    assert o["out"] == [[25, 25, 25], [50, 50, 50]]
    assert o["to_host"] >= 2 * 512               # the tensors did trade places (chunks of 2 MiB)



def test_pointer_queries_on_a_spilled_range_still_describe_device_memory(cuda_dev):
    """cudaPointerGetAttributes / cuPointerGetAttribute(s) / cuMemGetAddressRange on a managed range that is spilled at
    that moment: device memory of this device, one allocation from its base over its whole size (the driver itself
    holds one mapping per 1 GiB block, and none for a spilled block) — what an array-interface import relies on."""
    code = """
        import ctypes as C, json, torch
        torch.cuda.init()
        cu, shim = C.CDLL("libcuda.so.1"), C.CDLL(None)
        shim.cuVGPUResidentBytes.restype = C.c_longlong
        G = 1 << 30
        t = torch.arange(3 * G // 16, dtype=torch.int64, device="cuda")           # 1.5 GiB: two blocks of the manager
        others = [torch.ones(G // 8, dtype=torch.int64, device="cuda") for _ in range(2)]
        s = sum(int(o.sum().item()) for o in others)                              # 2 GiB may be resident: t goes out
        gone = shim.cuVGPUResidentBytes(C.c_ulonglong(t.data_ptr()))
        inside = t.data_ptr() + G + 4096                                          # a pointer into the second block
        mt = C.c_uint(0)
        rc1 = cu.cuPointerGetAttribute(C.byref(mt), 2, C.c_ulonglong(inside))     # MEMORY_TYPE
        base, size = C.c_ulonglong(0), C.c_size_t(0)
        rc2 = cu.cuMemGetAddressRange_v2(C.byref(base), C.byref(size), C.c_ulonglong(inside))
        vals = (C.c_ulonglong * 4)()
        attrs = (C.c_int * 4)(11, 12, 9, 8)                                       # RANGE_START_ADDR, RANGE_SIZE, DEVICE_ORDINAL, IS_MANAGED
        ptrs = (C.c_void_p * 4)(*[C.addressof(vals) + 8 * i for i in range(4)])
        others = None
        torch.cuda.empty_cache()
        for o in [torch.ones(G // 8, dtype=torch.int64, device="cuda") for _ in range(2)]:
            s += int(o.sum().item())                                              # and out again
        gone2 = shim.cuVGPUResidentBytes(C.c_ulonglong(t.data_ptr()))
        rc3 = cu.cuPointerGetAttributes(4, attrs, ptrs, C.c_ulonglong(inside))
        class Wrap:                                                               # an object that exports device memory by pointer
            def __init__(self, x):
                self.keep = x
                self.__cuda_array_interface__ = x.__cuda_array_interface__
        del o
        torch.cuda.empty_cache()
        for o in [torch.ones(G // 8, dtype=torch.int64, device="cuda") for _ in range(2)]:
            s += int(o.sum().item())
        gone3 = shim.cuVGPUResidentBytes(C.c_ulonglong(t.data_ptr()))
        v = torch.as_tensor(Wrap(t), device="cuda")                               # cudaPointerGetAttributes decides the device
        ok = v.device.index == 0 and int(v[-1].item()) == 3 * G // 16 - 1 and int(v[G // 8 + 7].item()) == G // 8 + 7
        print(json.dumps({"gone": [gone, gone2, gone3], "rc": [rc1, rc2, rc3], "memory_type": mt.value, "base_ok": base.value == t.data_ptr(),
                          "size": size.value, "range": [vals[0] == t.data_ptr(), int(vals[1])], "ordinal": int(vals[2] & 0xFFFFFFFF),
                          "managed": int(vals[3] & 0xFFFFFFFF), "ok": bool(ok), "s": s}))
    """
    o = last_json(run_under_shim(code, {**OVERSUB_ENV, "CUDA_DEVICE_MEMORY_LIMIT_0": "8g", "VGPU_OVERSUB_RESIDENT_MB": "2048"}, timeout=600))
    print("pointer queries:", o)
    assert o["gone"] == [0, 0, 0]                                   # the range was out of HBM before each query
    assert o["rc"] == [0, 0, 0] and o["memory_type"] == 2           # CU_MEMORYTYPE_DEVICE
    assert o["base_ok"] and o["size"] == 3 << 29
    assert o["range"] == [True, 3 << 29] and o["ordinal"] == 0 and o["managed"] == 0
    assert o["ok"] and o["s"] == 6 * (1 << 27)



def test_small_allocations_get_room_when_the_ranges_fill_the_whole_gpu(cuda_dev):
    """Allocations of 2 MiB and less stay plain driver allocations.  With the physical GPU full of resident managed
    ranges (no operator bound: the engine keeps VGPU_OVERSUB_RESERVE_MB free, here 1 GiB), 6 GiB of them in one burst run
    the driver itself out of HBM: the allocation spills ranges and tries again instead of failing."""
    code = """
        import ctypes as C, json, torch
        torch.cuda.init()
        shim = C.CDLL(None)
        G = 1 << 30
        def stats():
            st = (C.c_uint64 * 8)()
            shim.cuVGPUOversubStats(st)
            return st
        big = []
        while not stats()[6] and len(big) < 60:             # 4 GiB ranges until one no longer fits the GPU
            big.append(torch.full((4 * G // 8,), len(big) + 1, dtype=torch.int64, device="cuda"))
        before = int(stats()[1])
        small = [torch.full((1 << 17,), i, dtype=torch.int64, device="cuda") for i in range(6 * 1024)]   # 1 MiB each: 2 MiB plain segments
        torch.cuda.synchronize()
        ok_small = all(int(small[i][-1].item()) == i and int(small[i][0].item()) == i for i in range(0, len(small), 97))
        ok_big = all(int(b[0].item()) == i + 1 and int(b[-1].item()) == i + 1 for i, b in enumerate(big))
        st = stats()
        print(json.dumps({"n_big": len(big), "spilled_for_small_mib": (int(st[1]) - before) * 2, "ok_small": ok_small, "ok_big": ok_big, "bad": int(st[3])}))
    """
    p = run_under_shim(code, {**OVERSUB_ENV, "CUDA_DEVICE_MEMORY_LIMIT_0": "400g"}, timeout=900)
    o = last_json(p)
    print("small allocations on a full GPU:", o)
    assert o["n_big"] >= 30 and o["ok_small"] and o["ok_big"] and o["bad"] == 0, p.stderr[-2000:]
    assert o["spilled_for_small_mib"] >= 4096          # the burst did not fit what was left: at least one range went out for it


RANDOM_PROGRAM = """
    import ctypes, hashlib, json, random, sys, torch
    seed = int(sys.argv[1])
    rng = random.Random(seed)
    torch.manual_seed(seed)
    torch.cuda.init()
    M = 1 << 20
    def make(i):
        n = rng.randrange(8, 97) * M // 8
        return (torch.arange(n, dtype=torch.int64, device="cuda") * (0x9E3779B1 + 2 * i)) ^ (seed << 20)
    pool = [make(i) for i in range(24)]                    # ~1.2 GiB of int64 tensors, 8..96 MiB each
    side = torch.cuda.Stream()
    trace = hashlib.sha256()
    def pick(k):
        return rng.sample(range(len(pool)), k)
    for step in range(220):
        op = rng.randrange(12)
        if op == 0:                                        # device-to-device copy of a slice (cudaMemcpyAsync)
            i, j = pick(2)
            n = min(pool[i].numel(), pool[j].numel(), rng.randrange(1, 8 * M))
            o = rng.randrange(0, pool[i].numel() - n + 1)
            pool[i][o:o + n].copy_(pool[j][:n])
        elif op == 1:                                      # elementwise kernel over two ranges, in place
            i, j = pick(2)
            n = min(pool[i].numel(), pool[j].numel())
            pool[i][:n].add_(pool[j][:n]).bitwise_xor_(pool[j][:n] >> 3)
        elif op == 2:                                      # memset
            i, = pick(1)
            o = rng.randrange(0, pool[i].numel() // 2)
            pool[i][o:o + rng.randrange(1, pool[i].numel() // 2)].zero_()
        elif op == 3:                                      # new tensor from three others: allocation, free of the old one
            i, j, k = pick(3)
            n = min(pool[i].numel(), pool[j].numel(), pool[k].numel())
            pool[i] = torch.cat([pool[j][: n // 3], pool[k][: n // 3], pool[i][: n // 3]])
        elif op == 4:                                      # strided views: the copy kernel with strides in its argument struct
            i, j = pick(2)
            r = min(pool[i].numel(), pool[j].numel()) // 4096
            pool[i][: r * 4096].view(r, 4096)[:, 7:1000:3].copy_(pool[j][: r * 4096].view(r, 4096)[:, 11:1004:3])
        elif op == 5:                                      # gather through an index tensor that itself lives in a managed range
            i, j = pick(2)
            n = min(pool[i].numel(), pool[j].numel(), 4 * M)
            idx = (pool[j][:n].abs() % pool[i].numel())
            pool[i][:n] = pool[i][idx]
        elif op == 6:                                      # device -> host of a window, hashed
            i, = pick(1)
            o = rng.randrange(0, pool[i].numel() - 1024)
            trace.update(pool[i][o:o + rng.randrange(1, 1 << 18)].cpu().numpy().tobytes())
        elif op == 7:                                      # host -> device into the middle of a range
            i, = pick(1)
            n = rng.randrange(1, 1 << 18)
            h = torch.randint(-(1 << 40), 1 << 40, (n,), dtype=torch.int64)
            pool[i][pool[i].numel() - n:].copy_(h)
        elif op == 8:                                      # work on a second stream, ordered by events
            i, j = pick(2)
            n = min(pool[i].numel(), pool[j].numel())
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                pool[i][:n].mul_(3).sub_(pool[j][:n])
            torch.cuda.current_stream().wait_stream(side)
        elif op == 9:                                      # reduction with library temporaries (cub) + scalar read-back
            i, = pick(1)
            trace.update(str(int(pool[i].sum().item())).encode())
            trace.update(str(int(pool[i][: 2 * M].max().item())).encode())
        elif op == 10:                                     # sort of a window (multi-pass radix sort, scratch allocations)
            i, = pick(1)
            n = min(pool[i].numel(), 2 * M)
            pool[i][:n] = torch.sort(pool[i][:n]).values
        else:                                              # float GEMM over reinterpreted ranges (library kernel, pointers in a struct; values small enough to be exact)
            i, j = pick(2)
            a = (pool[i][: 1 << 19] & 0x3F).to(torch.float32).view(512, 1024)
            b = (pool[j][: 1 << 19] & 0x3F).to(torch.float32).view(1024, 512)
            c = (a @ b).to(torch.int64)
            pool[i][: c.numel()].bitwise_xor_(c.view(-1))
    torch.cuda.synchronize()
    for t in pool:
        trace.update(str(int(t.sum().item())).encode())
        trace.update(t[:: max(1, t.numel() // 4096)].cpu().numpy().tobytes())
    out = {"digest": trace.hexdigest()}
    try:
        st = (ctypes.c_uint64 * 8)()
        ctypes.CDLL(None).cuVGPUOversubStats(st)
        out.update({"to_host": int(st[1]), "restored": int(st[2]), "bad": int(st[3]), "paged": int(st[6])})
    except AttributeError:
        pass
    print(json.dumps(out))
"""


@pytest.mark.parametrize("seed", [11, 12])
def test_a_random_program_gives_the_same_bytes_paged_as_without_the_shim(cuda_dev, seed):
    """Differential check of the residency engine: 220 random operations over 24 tensors (copies of slices, in-place
    kernels over two ranges, memsets, cat = allocate + free, strided copies, gathers through an index held in a managed
    range, H2D / D2H windows, a second stream, cub reductions, radix sort, a float GEMM) run (a) with no shim at all and
    (b) under the shim with a third of the bytes allowed to be resident.  Every value read back on the way and a sample
    of every tensor at the end go into one SHA-256: the digests must be equal."""
    code = textwrap.dedent(RANDOM_PROGRAM)
    plain = subprocess.run([sys.executable, "-c", code, str(seed)], capture_output=True, text=True, timeout=600)
    assert plain.returncode == 0, plain.stderr[-3000:]
    want = json.loads(plain.stdout.strip().splitlines()[-1])
    td = tempfile.mkdtemp(prefix="vgpushim")
    env = dict(os.environ)
    env.update({"LD_PRELOAD": SHIM, "CUDA_DEVICE_MEMORY_SHARED_CACHE": os.path.join(td, "shr.cache"), "LIBCUDA_LOG_LEVEL": "0",
                **OVERSUB_ENV, "CUDA_DEVICE_MEMORY_LIMIT_0": "8g", "VGPU_OVERSUB_RESIDENT_MB": "640"})
    paged = subprocess.run([sys.executable, "-c", code, str(seed)], env=env, capture_output=True, text=True, timeout=900)
    assert paged.returncode == 0, paged.stderr[-3000:]
    got = json.loads(paged.stdout.strip().splitlines()[-1])
    print("random program, seed", seed, ":", got)
    assert got["paged"] == 1 and got["restored"] > 200 and got["bad"] == 0       # it really was paged (chunks of 2 MiB)
    assert got["digest"] == want["digest"]


TRAIN = """
    import json, torch
    torch.manual_seed(0)
    torch.backends.cuda.matmul.allow_tf32 = False
    dev = "cuda"
    D, L, B = 8192, 6, 2048                     # 6 layers x 256 MiB of fp32 weights; with grads and momentum ~4.5 GiB + activations
    layers = [torch.nn.Linear(D, D, bias=True, device=dev) for _ in range(L)]
    for l in layers:
        torch.nn.init.kaiming_normal_(l.weight, nonlinearity="relu")         # activations keep their scale, so the loss really moves
    params = [p for l in layers for p in l.parameters()]
    opt = torch.optim.SGD(params, lr=2e-2, momentum=0.9, foreach=True)       # multi-tensor kernels: pointer lists inside the argument struct
    x = torch.randn(B, D, device=dev)
    y = torch.randn(B, D, device=dev)
    losses = []
    for step in range(4):
        h = x
        for l in layers:
            h = torch.relu(l(h))
        loss = torch.nn.functional.mse_loss(h, y)
        opt.zero_grad(set_to_none=False)
        loss.backward()
        opt.step()
        losses.append(loss.item())
    w = sum(float(p.double().sum().item()) for p in params)
    import ctypes
    st = (ctypes.c_uint64 * 8)()
    try:
        ctypes.CDLL(None).cuVGPUOversubStats(st)
    except AttributeError:
        pass
    print(json.dumps({"losses": losses, "w": w, "to_host": st[1], "restored": st[2], "bad": st[3], "paged": st[6]}))
"""


def test_a_training_loop_gives_the_same_numbers_when_it_is_paged(cuda_dev):
    """cuBLAS GEMMs, elementwise / reduction kernels, autograd and a multi-tensor optimizer under the residency engine: a
    6-layer MLP whose weights, gradients and momentum do not fit the 2 GiB its managed ranges may occupy trains for four
    steps; losses and the final weights equal those of the same program with everything resident, bit for bit."""
    base = {**OVERSUB_ENV, "CUDA_DEVICE_MEMORY_LIMIT_0": "24g"}
    ref = last_json(run_under_shim(TRAIN, base, timeout=900))
    p = run_under_shim(TRAIN, {**base, "VGPU_OVERSUB_RESIDENT_MB": "2048"}, timeout=900)
    got = last_json(p)
    print("paged training:", got)
    assert ref["paged"] == 0 and ref["to_host"] == 0
    assert got["paged"] == 1 and got["to_host"] > 1024 and got["restored"] > 1024 and got["bad"] == 0, p.stderr[-2000:]
    assert got["losses"] == ref["losses"] and got["w"] == ref["w"]
    assert len(set(got["losses"])) == 4 and got["losses"][-1] < got["losses"][0]      # it did train
