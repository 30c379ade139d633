"""Contract A (Allocate / ListAndWatch / Register / GetPreferredAllocation) through the
C-ABI-backed plugin mirror, against (i) the hand-derived known answers of SURVEY.md
§8 c.1 (tests/golden/contract_a.json) and (ii) the pure-Python restatement in
oracle/plugin_model.py.  Reference: server.go:221-326, 361-533; vdevice.go:36-90."""
import json
import os
import random
import re

import pytest

from oracle import plugin_model as M

GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "contract_a.json")))
UUID = GOLD["mock"]["uuid"]
CACHE_RE = re.compile(r"^/tmp/[0-9a-f-]{36}\.cache$")


def make_plugin(vgpu, split=10, mem=1.0, cores=1.0, devices=None, **kw):
    P = vgpu.plugin
    devices = devices or [P.Device(UUID, Index="0", Paths=["/dev/nvidia0"], memory_mib=GOLD["mock"]["total_mib"], numa=0)]
    flags = P.Flags(deviceSplitCount=split, deviceMemoryScaling=mem, deviceCoresScaling=cores,
                    pcibusfile=GOLD["mock"]["pcibusfile"], **kw)
    p = P.NewNvidiaDevicePlugin(P.ResourceManager(devices), flags)
    p.initialize()
    return p


def test_list_and_watch_order_and_health(vgpu):
    p = make_plugin(vgpu)
    stream = p.ListAndWatch()
    first = next(stream)
    assert [d["ID"] for d in first] == [f"{UUID}-{i}" for i in range(10)]
    assert all(d["Health"] == "Healthy" for d in first)
    assert first[0]["Topology"] == {"Nodes": [{"ID": 0}]}
    p.markUnhealthy(p.cachedDevices[0])  # XID event -> every vGPU of that GPU goes Unhealthy, never recovers
    second = next(stream)
    assert all(d["Health"] == "Unhealthy" for d in second)


@pytest.mark.parametrize("case", GOLD["cases"], ids=lambda c: c["name"])
def test_allocate_known_answers(vgpu, case):
    p = make_plugin(vgpu, case["split"], case["mem_scaling"], case["cores_scaling"])
    ids = [UUID + s for s in case["ids"]]
    (resp,) = p.Allocate([ids])
    envs = dict(resp.Envs)
    cache = envs.pop("CUDA_DEVICE_MEMORY_SHARED_CACHE")
    assert CACHE_RE.match(cache)
    assert envs == case["envs"]
    assert [(m.ContainerPath, m.HostPath, m.ReadOnly) for m in resp.Mounts] == [tuple(m) for m in GOLD["mounts"]]
    assert resp.Devices == [] and resp.Annotations == {}
    # and against the independent restatement
    cached = [{"ID": UUID, "Index": "0"}]
    vds = M.device2vdevice([{"ID": UUID, "memory_mib": GOLD["mock"]["total_mib"]}], case["split"], case["mem_scaling"])
    o_envs, o_mounts = M.allocate_container(cached, vds, ids, split=case["split"], mem_scaling=case["mem_scaling"],
                                            cores_scaling=case["cores_scaling"])
    o_envs.pop("CUDA_DEVICE_MEMORY_SHARED_CACHE")
    assert envs == o_envs
    assert [(m.ContainerPath, m.HostPath, m.ReadOnly) for m in resp.Mounts] == o_mounts


def test_shared_cache_is_fresh_per_container(vgpu):
    p = make_plugin(vgpu)
    r = p.Allocate([[UUID + "-0"], [UUID + "-1"]])
    a, b = (x.Envs["CUDA_DEVICE_MEMORY_SHARED_CACHE"] for x in r)
    assert a != b and CACHE_RE.match(a) and CACHE_RE.match(b)
    assert a[5 + 14] == "4"  # uuid4 version nibble, as google/uuid NewString() produces


def test_unknown_device(vgpu):
    p = make_plugin(vgpu)
    with pytest.raises(vgpu.plugin.PluginError) as e:
        p.Allocate([[GOLD["unknown"]["id"]]])
    assert str(e.value) == GOLD["unknown"]["error"]
    with pytest.raises(KeyError):
        M.vdevices_by_ids(M.device2vdevice([{"ID": UUID, "memory_mib": 1}], 2, 1.0), [GOLD["unknown"]["id"]])


def test_register_and_options(vgpu):
    p = make_plugin(vgpu)
    assert p.Register() == GOLD["register"] == M.register()
    assert p.GetDevicePluginOptions() == {"GetPreferredAllocationAvailable": True}


def test_preferred_allocation_single_gpu_quirks(vgpu):
    """size 1 -> first available vGPU of the chosen GPU; size>1 with one
    physical GPU -> policy and fallback both yield nothing -> empty DeviceIDs."""
    p = make_plugin(vgpu)
    avail = [f"{UUID}-{i}" for i in (4, 2, 9)]
    assert p.GetPreferredAllocation([{"AvailableDeviceIDs": avail, "AllocationSize": 1}]) == [[avail[0]]]
    assert p.GetPreferredAllocation([{"AvailableDeviceIDs": avail, "AllocationSize": 2}]) == [[]]
    vds = M.device2vdevice([{"ID": UUID, "memory_mib": 184320}], 10, 1.0)
    assert M.preferred_allocation(vds, avail, [], 1, vgpu.plugin.simplePolicy) == [avail[0]]
    assert M.preferred_allocation(vds, avail, [], 2, vgpu.plugin.simplePolicy) == []
    with pytest.raises(vgpu.plugin.PluginError) as e:
        p.GetPreferredAllocation([{"AvailableDeviceIDs": ["nope"], "AllocationSize": 1}])
    assert str(e.value) == "Unable to retrieve list of available vdevices: unknown device: nope"


def test_index_and_volume_mount_strategies(vgpu):
    P = vgpu.plugin
    devs = [P.Device(f"GPU-{i:08d}-0000-0000-0000-000000000000", Index=str(i), Paths=[f"/dev/nvidia{i}"], memory_mib=184320)
            for i in range(3)]
    p = make_plugin(vgpu, split=2, devices=devs, deviceIDStrategy="index")
    ids = [devs[2].ID + "-1", devs[0].ID + "-0", devs[2].ID + "-0"]
    (r,) = p.Allocate([ids])
    # cachedDevices order, not request order (server.go:571-578); duplicates collapse
    assert r.Envs["NVIDIA_VISIBLE_DEVICES"] == "0,2"
    assert r.Envs["NVIDIA_DEVICE_MAP"] == f"0:{devs[2].ID} 1:{devs[0].ID} 2:{devs[2].ID}"
    p = make_plugin(vgpu, split=2, devices=devs, deviceListStrategy="volume-mounts")
    (r,) = p.Allocate([ids])
    assert r.Envs["NVIDIA_VISIBLE_DEVICES"] == "/var/run/nvidia-container-devices"
    assert [(m.ContainerPath, m.HostPath, m.ReadOnly) for m in r.Mounts[:2]] == [
        (f"/var/run/nvidia-container-devices/{devs[2].ID}", "/dev/null", False),
        (f"/var/run/nvidia-container-devices/{devs[0].ID}", "/dev/null", False)]
    assert len(r.Mounts) == 2 + 5


def test_mig_uuid_gets_one_zero_memory_vdevice(vgpu):
    P = vgpu.plugin
    p = make_plugin(vgpu, split=4, devices=[P.Device("MIG-GPU-abc/1/0", memory_mib=40960)])
    assert [(v.ID, v.memory) for v in p.vDevices] == [("MIG-GPU-abc/1/0-0", 0)]


def test_flag_validation_messages(vgpu):
    P = vgpu.plugin
    for kw, msg in [({"deviceSplitCount": 0}, "invalid --device-split-count option: 0"),
                    ({"deviceMemoryScaling": 0.0}, "invalid --device-memory-scaling option: 0.0"),
                    ({"deviceCoresScaling": -1.0}, "invalid --device-core-scaling option: -1.0"),
                    ({"deviceListStrategy": "x"}, "invalid --device-list-strategy option: x"),
                    ({"deviceIDStrategy": "y"}, "invalid --device-id-strategy option: y")]:
        with pytest.raises(ValueError) as e:
            P.validateFlags(P.Flags(**kw))
        assert str(e.value) == msg


def test_command_line_and_environment_of_the_reference(vgpu):
    """main.go:55-133: the eleven flags, their environment variables and defaults; flag beats environment beats default;
    booleans the way strconv.ParseBool reads them; validateFlags runs before anything starts."""
    P = vgpu.plugin
    d = P.parseFlags([], {})
    assert (d.migStrategy, d.failOnInitError, d.passDeviceSpecs, d.deviceListStrategy, d.deviceIDStrategy, d.nvidiaDriverRoot,
            d.deviceSplitCount, d.deviceMemoryScaling, d.deviceCoresScaling, d.enableLegacyPreferred, d.verbose) == \
        ("none", True, False, "envvar", "uuid", "/", 2, 1.0, 1.0, False, 0)
    # the daemonset of the reference: everything through the command line
    f = P.parseFlags(["--fail-on-init-error=false", "--device-split-count=10", "--device-memory-scaling", "1.5", "--device-cores-scaling=1.2",
                      "--pass-device-specs", "-mig-strategy", "mixed", "--enable-legacy-preferred=T", "--verbose=3"], {})
    assert (f.failOnInitError, f.deviceSplitCount, f.deviceMemoryScaling, f.deviceCoresScaling, f.passDeviceSpecs, f.migStrategy,
            f.enableLegacyPreferred, f.verbose) == (False, 10, 1.5, 1.2, True, "mixed", True, 3)
    # environment only, and a flag that overrides its variable
    env = {"DEVICE_SPLIT_COUNT": "4", "DEVICE_LIST_STRATEGY": "volume-mounts", "DEVICE_ID_STRATEGY": "index", "PASS_DEVICE_SPECS": "1",
           "NVIDIA_DRIVER_ROOT": "/run/nvidia/driver", "PCIBUSFILE": "/tmp/pci.vgpu", "MIG_STRATEGY": ""}
    e = P.parseFlags(["--device-split-count", "8"], env)
    assert (e.deviceSplitCount, e.deviceListStrategy, e.deviceIDStrategy, e.passDeviceSpecs, e.nvidiaDriverRoot, e.pcibusfile, e.migStrategy) == \
        (8, "volume-mounts", "index", True, "/run/nvidia/driver", "/tmp/pci.vgpu", "none")
    for argv, env, msg in [(["--device-split-count=0"], {}, "invalid --device-split-count option: 0"),
                           ([], {"DEVICE_MEMORY_SCALING": "0"}, "invalid --device-memory-scaling option: 0.0"),
                           (["--device-id-strategy=serial"], {}, "invalid --device-id-strategy option: serial"),
                           (["--device-split-count=-1"], {}, 'invalid value "-1" for flag -device-split-count'),
                           ([], {"PASS_DEVICE_SPECS": "yes"}, 'invalid value "yes" for environment variable PASS_DEVICE_SPECS'),
                           (["--gpu-count=2"], {}, "flag provided but not defined: -gpu-count"),
                           (["--verbose"], {}, "flag needs an argument: -verbose")]:
        with pytest.raises(ValueError) as x:
            P.parseFlags(argv, env)
        assert str(x.value) == msg


def test_split_arithmetic_matches_model_on_random_flags(vgpu):
    """uint64(float64(MiB)*scaling/float64(split)) and int(100*cores/float64(split)), many inputs."""
    rng = random.Random(0x5EED)
    lib, Desc = vgpu.capi.lib, vgpu.capi.Desc
    for _ in range(2000):
        mib = rng.choice([184320, 183359, 81920, 40960, 16384, rng.randrange(1, 1 << 20)])
        split = rng.randrange(1, 33)
        ms = rng.choice([1.0, 1.8, 1.6667, 0.5, rng.uniform(0.01, 8.0)])
        cs = rng.choice([1.0, 2.0, 0.5, rng.uniform(0.01, 4.0)])
        arr = (Desc * split)()
        assert lib.vgpu_split(UUID.encode(), mib, split, ms, arr, split) == split
        want = M.device2vdevice([{"ID": UUID, "memory_mib": mib}], split, ms)
        assert [(a.id.decode(), a.memory_mib) for a in arr] == [(w["ID"], w["memory"]) for w in want]
        assert lib.vgpu_sm_limit(split, cs) == M.sm_limit(split, cs)


def test_caller_array_too_small_is_reported(vgpu):
    lib, capi = vgpu.capi.lib, vgpu.capi
    arr = (capi.Desc * 2)()
    assert lib.vgpu_split(UUID.encode(), 1024, 4, 1.0, arr, 2) == capi.E_RANGE


def test_check_health_xid_rules(vgpu):
    """nvidia.go:166-237: XIDs 31/43/45 are application errors; an event without UUID takes every GPU
    down; otherwise only the GPU it names; health checks can be disabled; unhealthy vGPUs show up in the
    next ListAndWatch message and never recover."""
    P = vgpu.plugin
    devs = [P.Device(f"GPU-{i}", memory_mib=1024) for i in range(3)]
    ev = lambda xid, uuid=None: {"Etype": P.XidCriticalError, "Edata": xid, "UUID": uuid}  # noqa: E731
    assert P.checkHealth(devs, [ev(31, "GPU-1"), ev(43, "GPU-1"), ev(45, None)]) == []
    assert P.checkHealth(devs, [ev(79, "GPU-1")]) == [devs[1]]
    assert P.checkHealth(devs, [ev(48, None)]) == devs
    assert P.checkHealth(devs, [ev(79, "GPU-9")]) == []
    assert P.checkHealth(devs, [{"Etype": 0, "Edata": 79, "UUID": "GPU-1", "err": True}]) == []   # WaitForEvent timeout
    assert P.checkHealth(devs, [ev(79, "GPU-1")], disable="all") == [] == P.checkHealth(devs, [ev(79, "GPU-1")], disable="XIDS")
    assert P.checkHealth(devs, [{**ev(79, "GPU-1"), "GpuInstanceId": 3, "ComputeInstanceId": 0}]) == []  # a MIG slice's event
    p = P.NewNvidiaDevicePlugin(P.ResourceManager(devs), P.Flags(deviceSplitCount=2))
    p.initialize()
    for d in P.checkHealth(p.cachedDevices, [ev(79, "GPU-2")]):
        p.markUnhealthy(d)
    msgs = list(p.ListAndWatch())
    assert [d["Health"] for d in msgs[-1]] == ["Healthy"] * 4 + ["Unhealthy"] * 2


def test_partition_index_is_opt_in_and_additive(vgpu):
    """No reference counterpart: with smPartitionIndex the response gains VGPU_SM_PARTITION_INDEX_<n> = the i of
    "<UUID>-<i>" per requested vGPU and nothing else changes; without it Contract A is what server.go:486-507 builds."""
    ids = [UUID + "-3", UUID + "-7"]
    (plain,) = make_plugin(vgpu).Allocate([ids], _cache_uuid="0" * 36)
    (extra,) = make_plugin(vgpu, smPartitionIndex=True).Allocate([ids], _cache_uuid="0" * 36)
    assert not any(k.startswith("VGPU_SM_PARTITION_INDEX") for k in plain.Envs)
    added = {k: v for k, v in extra.Envs.items() if k not in plain.Envs}
    assert added == {"VGPU_SM_PARTITION_INDEX_0": "3", "VGPU_SM_PARTITION_INDEX_1": "7"}
    assert {k: v for k, v in extra.Envs.items() if k in plain.Envs} == dict(plain.Envs)
    assert [(m.ContainerPath, m.HostPath, m.ReadOnly) for m in extra.Mounts] == [(m.ContainerPath, m.HostPath, m.ReadOnly) for m in plain.Mounts]


def test_pci_bus_file_is_what_start_writes_from_lspci(vgpu, tmp_path):
    """main.go:164-185: every lspci line longer than one byte that contains "NVIDIA" contributes its first
    space-separated token; the result goes to $PCIBUSFILE, which Allocate mounts at /usr/local/vgpu/pciinfo.vgpu."""
    lspci = ("00:00.0 Host bridge: Intel Corporation Device 09a2 (rev 04)\n"
             "1b:00.0 3D controller: NVIDIA Corporation Device 2901 (rev a1)\n"
             "\n"
             "1c:00.0 Bridge: NVIDIA Corporation Device 22a3 (rev a1)\n"     # NVSwitch lines count too: the Go code only looks for the word
             "43:00.0 3D controller: NVIDIA Corporation Device 2901 (rev a1)\n"
             "x\n"
             "aa:00.0 Ethernet controller: Mellanox Technologies MT2910 Family [ConnectX-7]")
    out = tmp_path / "pciinfo.vgpu"
    s = vgpu.plugin.loadPciInfo(lspci, str(out))
    assert s == "1b:00.0\n1c:00.0\n43:00.0\n" and out.read_text() == s and (out.stat().st_mode & 0o777) == 0o644
    assert vgpu.plugin.loadPciInfo("", "") == ""                       # no devices, no $PCIBUSFILE: nothing written
    assert vgpu.plugin.loadPciInfo("NVIDIA", "") == "NVIDIA\n"         # a line without a space is its own first token (strings.Split)
