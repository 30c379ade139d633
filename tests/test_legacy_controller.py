"""the legacy preferred-allocation path (--enable-legacy-preferred) —
VDeviceController + kubelet checkpoint round trip (vdevice-controller.go:32-286, server.go:104-111,
408-457, 480-485).  Checked against known answers derived from the Go source with a synthetic
kubelet_internal_checkpoint file and a fake pod lister."""
import base64
import json
import os
import tempfile

import pytest

UUID = "GPU-00000000-0000-0000-0000-000000000001"


def make(vgpu, pods, ckpt_entries, split=4):
    P, G = vgpu.plugin, __import__("vgpu_b200.grpc_server", fromlist=["x"])
    td = tempfile.mkdtemp(prefix="ckpt")
    entries = []
    for uid, request, using, resource in ckpt_entries:
        r = P.ContainerAllocateResponse(Annotations={P.annRequest: ",".join(request), P.annUsing: ",".join(using)} if using else {})
        entries.append({"PodUID": uid, "ContainerName": "c", "ResourceName": resource, "DeviceIDs": list(request),
                        "AllocResp": base64.b64encode(G.enc_allocate_container(r)).decode()})
    json.dump({"Data": {"PodDeviceEntries": entries, "RegisteredDevices": {"nvidia.com/gpu": []}}, "Checksum": 0},
              open(os.path.join(td, P.kubeletDeviceManagerCheckpoint), "w"))
    p = P.NewNvidiaDevicePlugin(P.ResourceManager([P.Device(UUID, memory_mib=184320)]), P.Flags(deviceSplitCount=split, enableLegacyPreferred=True))
    p._podLister = lambda: pods
    p._checkpointDir = td
    p.initialize()
    return p


def test_controller_bookkeeping(vgpu):
    C = vgpu.plugin.VDeviceController(["a", "b", "c"])
    assert C.available() == ["a", "b", "c"]
    C.acquire(["r1", "r2"], ["b", "c"])
    assert C.available() == ["a"] and C.idMap == {"a": "", "b": "r1", "c": "r2"}
    C.acquire(["r3"], ["a", "zzz", "b"])            # unknown id skipped; more using than request -> "mismatched"
    assert C.idMap == {"a": "r3", "b": "mismatched", "c": "r2"}
    C.releaseByRequest(["r2", "r3"])                 # frees whatever was handed out for those request ids
    assert C.idMap == {"a": "", "b": "mismatched", "c": ""}
    C.release(["b", "nope"])
    assert C.available() == ["a", "b", "c"]


def test_register_option_flips_in_legacy_mode(vgpu):
    p = make(vgpu, [], [])
    assert p.Register()["Options"]["GetPreferredAllocationAvailable"] is False      # server.go:234
    assert p.GetDevicePluginOptions()["GetPreferredAllocationAvailable"] is True    # server.go:248


def test_checkpoint_marks_running_pods_devices_busy_and_frees_finished_ones(vgpu):
    pods = [{"uid": "pod-run", "phase": "Running"}, {"uid": "pod-done", "phase": "Succeeded"}]
    ck = [("pod-run", [f"{UUID}-0"], [f"{UUID}-2"], "nvidia.com/gpu"),
          ("pod-done", [f"{UUID}-1"], [f"{UUID}-3"], "nvidia.com/gpu"),
          ("pod-gone", [f"{UUID}-1"], [f"{UUID}-1"], "nvidia.com/gpu"),      # pod no longer listed -> released
          ("pod-run", ["x"], ["y"], "other.com/thing")]                       # other resource: ignored
    p = make(vgpu, pods, ck)
    p.vDeviceController.acquire([f"{UUID}-1"], [f"{UUID}-3"])                 # stale state from before the restart
    p.vDeviceController.updateFromCheckpoint()
    assert p.vDeviceController.idMap == {f"{UUID}-0": "", f"{UUID}-1": "", f"{UUID}-2": f"{UUID}-0", f"{UUID}-3": ""}


def test_allocate_in_legacy_mode_substitutes_free_vdevices_and_annotates(vgpu):
    A = vgpu.plugin
    pods = [{"uid": "pod-run", "phase": "Running"}]
    # a running pod was handed vGPU 2 for its request id "-0" (an earlier substitution)
    p = make(vgpu, pods, [("pod-run", [f"{UUID}-0"], [f"{UUID}-2"], "nvidia.com/gpu")])
    # kubelet now asks for vGPU 2 (free in ITS books); it is busy -> the plugin hands out another one
    (r,) = p.Allocate([[f"{UUID}-2"]])
    using = r.Annotations[A.annUsing]
    assert r.Annotations[A.annRequest] == f"{UUID}-2"
    assert using == f"{UUID}-0"                       # first free vGPU of the (only) physical GPU
    assert r.Envs["CUDA_DEVICE_MEMORY_LIMIT_0"] == "46080m" and r.Envs["NVIDIA_DEVICE_MAP"] == f"0:{UUID}"
    assert p.vDeviceController.idMap == {f"{UUID}-0": f"{UUID}-2", f"{UUID}-1": "", f"{UUID}-2": f"{UUID}-0", f"{UUID}-3": ""}
    # asking for a request id that is already on the books means its holder is gone ("device loss"): it is freed first
    (r,) = p.Allocate([[f"{UUID}-2"]])
    assert r.Annotations[A.annUsing] == f"{UUID}-0"
    # drain the GPU; the next request fails like the reference
    p.Allocate([[f"{UUID}-1"]])
    p.Allocate([[f"{UUID}-3"]])
    with pytest.raises(A.PluginError) as e:
        p.Allocate([[f"{UUID}-9"]])
    assert str(e.value) == "no enough devices"
