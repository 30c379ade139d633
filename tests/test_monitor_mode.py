"""VGPU_MONITOR_MODE (server.go:365-406 pod matching, 416-431 container names, 494-501 the mount
and the cache path): Allocate puts each container's cache file into a per-container host directory that an outside
monitor scans.  Hand-derived expectations from the Go source; CPU only."""
import ctypes as C
import os
import subprocess

import pytest

import vgpu_b200 as V
from vgpu_b200 import capi, monitor, plugin as P

UUID = "GPU-11111111-2222-3333-4444-555555555555"
CACHE = "01234567-89ab-4def-8123-456789abcdef"


def pod(name, phase, *ctrs):
    """ctrs: (container name, gpu count or None)"""
    return {"metadata": {"name": name}, "status": {"phase": phase},
            "spec": {"containers": [{"name": n, "resources": ({"limits": {"nvidia.com/gpu": str(g)}} if g is not None else {})}
                                    for n, g in ctrs]}}


def test_match_pending_pod_follows_the_reference_rules():
    req = [["a", "b"], ["c"]]                                   # two GPU containers: 2 devices and 1 device
    pods = [pod("running", "Running", ("x", 2), ("y", 1)),      # right shape, wrong phase
            pod("first", "Pending", ("x", 2), ("y", 1)),
            pod("sidecar", "Pending", ("init", None), ("x", 2), ("log", None), ("y", 1)),   # non-GPU containers are skipped
            pod("wrong", "Pending", ("x", 1), ("y", 1)),
            pod("toomany", "Pending", ("x", 2), ("y", 1), ("z", 1)),                        # Go: index out of range
            pod("nogpu", "Pending", ("x", None))]                                           # Go: matches, then panics
    assert P.matchPendingPod(pods, req)["metadata"]["name"] == "sidecar"    # the LAST match wins (:401-404)
    assert P.matchPendingPod(pods[:2], req)["metadata"]["name"] == "first"
    assert P.matchPendingPod([pods[0], pods[3]], req) is None
    assert P.gpuContainerNames(pods[2], 2) == ["x", "y"]
    with pytest.raises(P.PluginError):
        P.gpuContainerNames(pods[1], 3)


def _plugin(tmp_path, pods):
    flags = P.Flags(deviceSplitCount=2, deviceMemoryScaling=1.0, deviceCoresScaling=1.0)
    rm = P.ResourceManager([P.Device(UUID, Index="0", Paths=["/dev/nvidia0"], memory_mib=40960)])
    p = P.NewNvidiaDevicePlugin(rm, flags)
    p._podLister = lambda: pods
    p._sharedRoot = str(tmp_path / "shared")
    p.initialize()
    return p


def test_allocate_in_monitor_mode_moves_the_cache_file_into_a_shared_directory(tmp_path, monkeypatch):
    monkeypatch.setenv("VGPU_MONITOR_MODE", "true")
    pods = [pod("train-0", "Pending", ("sidecar", None), ("worker", 1))]
    p = _plugin(tmp_path, pods)
    resp = p.Allocate([[UUID + "-1"]], _cache_uuid=CACHE)[0]
    assert resp.Envs["CUDA_DEVICE_MEMORY_SHARED_CACHE"] == f"/train-0_worker/{CACHE}.cache"
    m = [(x.ContainerPath, x.HostPath, x.ReadOnly) for x in resp.Mounts]
    assert m[0] == ("/train-0_worker", "/usr/local/vgpu/shared/train-0_worker", False)        # appended first (:497-499)
    assert [x[0] for x in m[1:]] == ["/usr/local/vgpu/libvgpu.so", "/etc/ld.so.preload", "/usr/local/vgpu/pciinfo.vgpu",
                                     "/usr/bin/vgpuvalidator", "/vgpu"]
    assert os.path.isdir(tmp_path / "shared" / "train-0_worker")                               # os.MkdirAll (:496)
    # everything else is what the plain path answers
    monkeypatch.delenv("VGPU_MONITOR_MODE")
    plain = _plugin(tmp_path, pods).Allocate([[UUID + "-1"]], _cache_uuid=CACHE)[0]
    assert plain.Envs["CUDA_DEVICE_MEMORY_SHARED_CACHE"] == f"/tmp/{CACHE}.cache"
    assert {k: v for k, v in resp.Envs.items() if k != "CUDA_DEVICE_MEMORY_SHARED_CACHE"} == \
           {k: v for k, v in plain.Envs.items() if k != "CUDA_DEVICE_MEMORY_SHARED_CACHE"}
    assert [(x.ContainerPath, x.HostPath, x.ReadOnly) for x in plain.Mounts] == m[1:]


def test_monitor_mode_without_a_matching_pod_is_an_error_not_a_crash(tmp_path, monkeypatch):
    monkeypatch.setenv("VGPU_MONITOR_MODE", "1")
    p = _plugin(tmp_path, [pod("other", "Pending", ("w", 2))])
    with pytest.raises(P.PluginError):
        p.Allocate([[UUID + "-0"]])
    p._podLister = None
    with pytest.raises(P.PluginError):
        p.Allocate([[UUID + "-0"]])


def test_c_abi_rejects_a_shared_name_with_a_slash():
    cfg = capi.PluginCfg(split=2, mem_scaling=1.0, cores_scaling=1.0, id_strategy=0, list_strategy=0, pcibusfile=b"/x")
    descs = (capi.Desc * 2)()
    assert capi.lib.vgpu_split(UUID.encode(), 40960, 2, 1.0, descs, 2) == 2
    req = (C.POINTER(capi.Desc) * 1)(C.pointer(descs[0]))
    envs, mounts = (capi.KV * 64)(), (capi.Mount * 64)()
    ne, nm = C.c_size_t(64), C.c_size_t(64)
    pu, pi = (C.c_char_p * 1)(UUID.encode()), (C.c_char_p * 1)(b"0")
    for bad in (b"a/b", b""):
        rc = capi.lib.vgpu_allocate_env_monitor(C.byref(cfg), req, 1, pu, pi, 1, CACHE.encode(), bad, envs, C.byref(ne), mounts, C.byref(nm))
        assert rc == capi.E_INVAL
    rc = capi.lib.vgpu_allocate_env_monitor(C.byref(cfg), req, 1, pu, pi, 1, CACHE.encode(), b"pod_ctr", envs, C.byref(ne), mounts, C.byref(nm))
    assert rc == 0 and nm.value == 6 and mounts[0].container_path == b"/pod_ctr" and mounts[0].read_only == 0


def test_host_monitor_scans_the_shared_directories(tmp_path):
    """A process under the shim writes its cache file where monitor mode told it to; monitor.scan finds it by
    <pod>_<container> and reads limits and live usage."""
    from oracle import quota_harness as H
    H.build()
    d = tmp_path / "shared" / "train-0_worker"
    d.mkdir(parents=True)
    cache = str(d / f"{CACHE}.cache")
    env = H._env(H.NEW_LIB, cache, {"CUDA_DEVICE_MEMORY_LIMIT_0": "2g", "NVIDIA_DEVICE_MAP": "0:" + H.UUID0}, "")
    pr = subprocess.Popen([os.path.join(H.REF_DIR, "trace_runner")], stdin=subprocess.PIPE, stdout=subprocess.PIPE, env=env, text=True, bufsize=1)
    try:
        assert pr.stdout.readline().startswith("init")
        pr.stdin.write(f"alloc {64 << 20} 1\n")
        pr.stdin.flush()
        assert pr.stdout.readline().strip() == "0"
        seen = monitor.scan(str(tmp_path / "shared"))
        assert list(seen) == ["train-0_worker"]
        r = seen["train-0_worker"]
        assert r["limit"][0] == 2 << 30 and r["usage"][0] == 64 << 20 and r["procs"][0]["pid"] == pr.pid
    finally:
        pr.stdin.write("quit\n")
        pr.stdin.flush()
        pr.wait(timeout=20)
    assert monitor.scan(str(tmp_path / "shared"))["train-0_worker"]["proc_num"] == 0
    (tmp_path / "shared" / "empty_dir").mkdir()
    (tmp_path / "shared" / "junk_x").mkdir()
    (tmp_path / "shared" / "junk_x" / "a.cache").write_bytes(b"not a region")
    assert list(monitor.scan(str(tmp_path / "shared"))) == ["train-0_worker"]
