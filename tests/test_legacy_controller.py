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
    data = {"PodDeviceEntries": entries, "RegisteredDevices": {"nvidia.com/gpu": []}}
    json.dump({"Data": data, "Checksum": P.checkpointChecksum(data)},   # as the kubelet writes it (MarshalCheckpoint)
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


def test_checkpoint_checksum_is_the_kubelets_object_hash(vgpu):
    """checksum.New = FNV-1a/32 over go-spew's `%#v` of checkpoint.checkpointData.  Parity unpinned by a reference
    vector (Go cannot run here, the tree holds no checkpoint fixture): the dump is written out by hand from the rules of
    vendor/github.com/davecgh/go-spew/spew/format.go and the hash checked against FNV's published test vectors."""
    P = vgpu.plugin

    def fnv(b):
        h = 0x811C9DC5
        for x in b:
            h = ((h ^ x) * 0x01000193) & 0xFFFFFFFF
        return h
    assert fnv(b"") == 0x811C9DC5 and fnv(b"a") == 0xE40C292C and fnv(b"foobar") == 0xBF9CF968      # FNV-1a 32 known answers
    # a freshly started kubelet: New(make([]PodDevicesEntry, 0), make(map[string][]string)) marshals [] and {}
    empty = {"PodDeviceEntries": [], "RegisteredDevices": {}}
    assert P.checkpointDump(empty) == "(checkpoint.checkpointData){PodDeviceEntries:([]checkpoint.PodDevicesEntry)[] RegisteredDevices:(map[string][]string)map[]}"
    # JSON null unmarshals to nil slices / maps, which spew prints as <nil>
    assert P.checkpointDump({"PodDeviceEntries": None, "RegisteredDevices": None}) == \
        "(checkpoint.checkpointData){PodDeviceEntries:([]checkpoint.PodDevicesEntry)<nil> RegisteredDevices:(map[string][]string)<nil>}"
    data = {"PodDeviceEntries": [{"PodUID": "u1", "ContainerName": "c", "ResourceName": "nvidia.com/gpu", "DeviceIDs": ["g-0", "g-1"],
                                  "AllocResp": base64.b64encode(bytes([10, 2, 255])).decode()},
                                 {"PodUID": "u2", "ContainerName": "d", "ResourceName": "x", "DeviceIDs": None, "AllocResp": None}],
            "RegisteredDevices": {"z.com/b": ["q"], "nvidia.com/gpu": ["g-0", "g-1", "g-2"]}}
    want = ("(checkpoint.checkpointData){PodDeviceEntries:([]checkpoint.PodDevicesEntry)["
            "{PodUID:(string)u1 ContainerName:(string)c ResourceName:(string)nvidia.com/gpu DeviceIDs:([]string)[g-0 g-1] AllocResp:([]uint8)[10 2 255]} "
            "{PodUID:(string)u2 ContainerName:(string)d ResourceName:(string)x DeviceIDs:([]string)<nil> AllocResp:([]uint8)<nil>}] "
            "RegisteredDevices:(map[string][]string)map[nvidia.com/gpu:[g-0 g-1 g-2] z.com/b:[q]]}")      # map keys sorted (SortKeys)
    assert P.checkpointDump(data) == want and P.checkpointChecksum(data) == fnv(want.encode())


def test_a_checkpoint_with_a_wrong_checksum_is_corrupted(vgpu):
    """vdevice-controller.go:64 reads through the checkpoint manager, which verifies (checkpoint.go:75-78)."""
    p = make(vgpu, [], [("pod-run", [f"{UUID}-0"], [f"{UUID}-2"], "nvidia.com/gpu")])
    path = os.path.join(p._checkpointDir, vgpu.plugin.kubeletDeviceManagerCheckpoint)
    cp = json.load(open(path))
    p.vDeviceController.updateFromCheckpoint()                     # intact: accepted
    cp["Data"]["PodDeviceEntries"][0]["PodUID"] = "tampered"
    json.dump(cp, open(path, "w"))
    with pytest.raises(vgpu.plugin.PluginError, match="checkpoint is corrupted"):
        p.vDeviceController.updateFromCheckpoint()
