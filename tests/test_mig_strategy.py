"""MIG strategies: mig-strategy.go:46-239, mig.go:84-119, nvidia.go:62-137, server.go:328-358.
Hand-derived from the Go source on synthetic node descriptions; CPU only."""
import pytest

from vgpu_b200 import plugin as P


def mig(i, g, c, mb):
    return P.MigDevice(f"MIG-GPU-{i}", g, c, mb, [f"/dev/nvidia-caps/nvidia-cap{i}"])


def test_resource_names_and_validity():
    assert P.migResourceName(mig(0, 1, 1, 22912)) == "mig-1g.23gb"        # ceil(22912 / 1024) = 23
    assert P.migResourceName(mig(0, 7, 7, 183359)) == "mig-7g.180gb"
    assert P.migResourceName(mig(0, 2, 1, 45824)) == "mig-1c.2g.45gb"     # compute slice != gpu slice
    assert P.validMigDevice(mig(0, 3, 3, 1)) and not P.validMigDevice(mig(0, 2, 1, 1))


def test_strategy_none_lists_every_gpu_even_with_mig_enabled():
    node = [P.NodeGpu("GPU-a", "/dev/nvidia0", True, [mig(0, 1, 1, 22912)], 184320), P.NodeGpu("GPU-b", "/dev/nvidia1", memory_mib=184320)]
    (p,) = P.migPlugins("none", node, P.Flags(deviceSplitCount=2))
    p.initialize()
    assert p.resourceName == "nvidia.com/gpu" and [d.ID for d in p.cachedDevices] == ["GPU-a", "GPU-b"]
    assert [d["ID"] for d in p.apiDevices()] == ["GPU-a-0", "GPU-a-1", "GPU-b-0", "GPU-b-1"]
    with pytest.raises(P.PluginError, match="Unknown strategy"):
        P.migPlugins("both", node)


def test_strategy_single():
    same = [P.NodeGpu("GPU-a", "/dev/nvidia0", True, [mig(0, 1, 1, 22912), mig(1, 1, 1, 22912)]),
            P.NodeGpu("GPU-b", "/dev/nvidia1", True, [mig(2, 1, 1, 22912)])]
    (p,) = P.migPlugins("single", same)
    p.initialize()
    assert p.resourceName == "nvidia.com/gpu" and p.allocatePolicy is None and p.migStrategy == "none"   # :122-129
    assert [(d.ID, d.Index) for d in p.cachedDevices] == [("MIG-GPU-0", "0:0"), ("MIG-GPU-1", "0:1"), ("MIG-GPU-2", "1:0")]
    # no MIG-enabled GPU at all: falls back to "none"
    (q,) = P.migPlugins("single", [P.NodeGpu("GPU-a", "/dev/nvidia0")])
    assert q.allocatePolicy is not None
    with pytest.raises(P.PluginError, match="same migEnabled value"):
        P.migPlugins("single", same + [P.NodeGpu("GPU-c", "/dev/nvidia2")])
    with pytest.raises(P.PluginError, match="No MIG devices associated with /dev/nvidia0: GPU-a"):
        P.migPlugins("single", [P.NodeGpu("GPU-a", "/dev/nvidia0", True, [])])
    with pytest.raises(P.PluginError, match="Unsupported MIG device found: mig-1c.2g.45gb"):
        P.migPlugins("single", [P.NodeGpu("GPU-a", "/dev/nvidia0", True, [mig(0, 2, 1, 45824)])])
    with pytest.raises(P.PluginError, match="More than one MIG device type"):
        P.migPlugins("single", [P.NodeGpu("GPU-a", "/dev/nvidia0", True, [mig(0, 1, 1, 22912), mig(1, 2, 2, 45824)])])


def test_strategy_mixed_and_mig_allocate():
    node = [P.NodeGpu("GPU-a", "/dev/nvidia0", True, [mig(0, 1, 1, 22912), mig(1, 2, 2, 45824), mig(2, 2, 1, 45824)]),
            P.NodeGpu("GPU-b", "/dev/nvidia1", False, memory_mib=184320),
            P.NodeGpu("GPU-c", "/dev/nvidia2", True, [mig(3, 1, 1, 22912)])]
    plugins = P.migPlugins("mixed", node, P.Flags(deviceSplitCount=2))
    assert [p.resourceName for p in plugins] == ["nvidia.com/gpu", "nvidia.com/mig-1g.23gb", "nvidia.com/mig-2g.45gb"]   # the 1c.2g one is skipped
    assert [p.socket.rsplit("/", 1)[-1] for p in plugins] == ["nvidia-gpu.sock", "nvidia-mig-1g.23gb.sock", "nvidia-mig-2g.45gb.sock"]
    assert [p.migStrategy for p in plugins] == ["none", "mixed", "mixed"]
    for p in plugins:
        p.initialize()
    assert [d.ID for d in plugins[0].cachedDevices] == ["GPU-b"]                       # MIG-enabled GPUs are skipped (:181)
    assert [(d.ID, d.Index) for d in plugins[1].cachedDevices] == [("MIG-GPU-0", "0:0"), ("MIG-GPU-3", "2:0")]
    assert [d["ID"] for d in plugins[1].apiDevices()] == ["MIG-GPU-0", "MIG-GPU-3"]   # MIG plugins list devices, not vGPUs (server.go:589-594)
    assert plugins[1].GetPreferredAllocation([{"AvailableDeviceIDs": ["MIG-GPU-0"], "AllocationSize": 1}]) is None
    # MIGAllocate (server.go:328-358): the device list only — no quota env, no vGPU mounts
    (r,) = plugins[1].Allocate([["MIG-GPU-3", "MIG-GPU-0"]])
    assert r.Envs == {"NVIDIA_VISIBLE_DEVICES": "MIG-GPU-3,MIG-GPU-0"} and r.Mounts == []
    with pytest.raises(P.PluginError, match=r"invalid allocation request for 'nvidia.com/mig-1g.23gb': unknown device: MIG-GPU-1"):
        plugins[1].Allocate([["MIG-GPU-1"]])
    # the whole-GPU plugin of the same node still hands out vGPUs with quota envs
    (g,) = plugins[0].Allocate([["GPU-b-1"]], _cache_uuid="01234567-89ab-4def-8123-456789abcdef")
    assert g.Envs["CUDA_DEVICE_MEMORY_LIMIT_0"] == "92160m" and g.Envs["NVIDIA_DEVICE_MAP"] == "0:GPU-b"


def test_mig_allocate_index_strategy_and_volume_mounts():
    node = [P.NodeGpu("GPU-a", "/dev/nvidia0", True, [mig(0, 1, 1, 22912), mig(1, 1, 1, 22912)])]
    flags = P.Flags(deviceIDStrategy=P.DeviceIDStrategyIndex, deviceListStrategy=P.DeviceListStrategyVolumeMounts)
    p = P.migPlugins("mixed", node, flags)[1]
    p.initialize()
    (r,) = p.Allocate([["MIG-GPU-1"]])
    assert r.Envs == {"NVIDIA_VISIBLE_DEVICES": "/var/run/nvidia-container-devices"}
    assert [(m.ContainerPath, m.HostPath) for m in r.Mounts] == [("/var/run/nvidia-container-devices/0:1", "/dev/null")]
