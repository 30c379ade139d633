"""Pure-Python restatement of the Go plugin's Contract A.

TEST INFRASTRUCTURE ONLY.  Follows, function by function:
  Device2VDevice          vdevice.go:36-58
  VDevicesByIDs           vdevice.go:61-75
  UniqueDeviceIDs         vdevice.go:78-90
  deviceIDsFromUUIDs      server.go:565-581
  apiEnvs / apiMounts     server.go:598-616
  Allocate (per container) server.go:459-524
  Register / options      server.go:221-251
  GetPreferredAllocation  server.go:271-326 (+ besteffort_policy.go:39-41 size guard)

PARITY UNPINNED BY THE REFERENCE'S OWN TESTS: the reference has no tests and Go is
not installed here, so the Go code cannot be executed.  Pinned by the hand-derived
known answers of SURVEY.md §8(c.1), held in tests/golden/contract_a.json.
"""
import struct


def _go_uint64_of_float(x: float) -> int:
    # Go uint64(float64) truncates toward zero for in-range values.
    return int(x)


def device2vdevice(devices, split, mem_scaling):
    """devices: list of dict(ID, memory_mib, ...) -> list of dict(ID, uuid, memory)."""
    out = []
    for d in devices:
        if "MIG" in d["ID"]:
            out.append({"ID": f"{d['ID']}-0", "uuid": d["ID"], "memory": 0, "dev": d})
            continue
        memory = _go_uint64_of_float(float(d["memory_mib"]) * mem_scaling / float(split))
        for i in range(split):
            out.append({"ID": f"{d['ID']}-{i}", "uuid": d["ID"], "memory": memory, "dev": d})
    return out


def vdevices_by_ids(vdevices, ids):
    res = []
    for i in ids:
        for vd in vdevices:
            if vd["ID"] == i:
                res.append(vd)
                break
        else:
            raise KeyError(f"unknown device: {i}")
    return res


def unique_device_ids(vdevices):
    ids = []
    for vd in vdevices:
        if vd["uuid"] not in ids:
            ids.append(vd["uuid"])
    return ids


def device_ids_from_uuids(cached_devices, uuids, id_strategy):
    if id_strategy == "uuid":
        return list(uuids)
    out = []
    if id_strategy == "index":
        for d in cached_devices:
            for u in uuids:
                if d["ID"] == u:
                    out.append(d["Index"])
    return out


def sm_limit(split, cores_scaling):
    return int(100 * cores_scaling / float(split))


def allocate_container(cached_devices, vdevices, ids, *, split, mem_scaling, cores_scaling, list_strategy="envvar",
                       id_strategy="uuid", pcibusfile="/usr/local/vgpu/pciinfo.vgpu"):
    vds = vdevices_by_ids(vdevices, ids)
    uuids = unique_device_ids(vds)
    device_ids = device_ids_from_uuids(cached_devices, uuids, id_strategy)
    envs, mounts = {}, []
    if list_strategy == "envvar":
        envs["NVIDIA_VISIBLE_DEVICES"] = ",".join(device_ids)
    if list_strategy == "volume-mounts":
        envs["NVIDIA_VISIBLE_DEVICES"] = "/var/run/nvidia-container-devices"
        for i in device_ids:
            mounts.append(("/var/run/nvidia-container-devices/" + i, "/dev/null", False))
    map_envs = []
    for i, vd in enumerate(vds):
        envs[f"CUDA_DEVICE_MEMORY_LIMIT_{i}"] = f"{vd['memory']}m"
        map_envs.append(f"{i}:{vd['uuid']}")
    envs["CUDA_DEVICE_SM_LIMIT"] = str(sm_limit(split, cores_scaling))
    envs["NVIDIA_DEVICE_MAP"] = " ".join(map_envs)
    envs["CUDA_DEVICE_MEMORY_SHARED_CACHE"] = "/tmp/<uuid4>.cache"  # random; compare by regex
    if mem_scaling > 1:
        envs["CUDA_OVERSUBSCRIBE"] = "true"
    mounts += [
        ("/usr/local/vgpu/libvgpu.so", "/usr/local/vgpu/libvgpu.so", True),
        ("/etc/ld.so.preload", "/usr/local/vgpu/ld.so.preload", True),
        ("/usr/local/vgpu/pciinfo.vgpu", pcibusfile, True),
        ("/usr/bin/vgpuvalidator", "/usr/local/vgpu/vgpuvalidator", True),
        ("/vgpu", "/usr/local/vgpu/license", True),
    ]
    return envs, mounts


def register(policy_present=True, legacy=False):
    return {"Version": "v1beta1", "Endpoint": "nvidia-gpu.sock", "ResourceName": "nvidia.com/gpu",
            "Options": {"GetPreferredAllocationAvailable": policy_present and not legacy}}


def preferred_allocation(vdevices, available_ids, must_include_ids, size, policy):
    avail_vd = vdevices_by_ids(vdevices, available_ids)
    available = unique_device_ids(avail_vd)
    required = unique_device_ids(vdevices_by_ids(vdevices, must_include_ids))
    allocated = policy(available, required, size)
    if len(allocated) == 0 and len(available) >= size:
        allocated = available[0:size]
    out = []
    for u in allocated:
        for vd in avail_vd:
            if vd["uuid"] == u:
                out.append(vd["ID"])
                break
    return out


# ---- bestEffortPolicy (vendor/github.com/NVIDIA/go-gpuallocator/gpuallocator/besteffort_policy.go) ----
# Independent restatement (itertools) used to cross-check vgpu_best_effort.  Devices are indices into
# the pair-score matrix; None is the reference's nil padding.
import itertools


def _pair(score, a, b):
    return 0 if a is None or b is None or a == b else score[a][b]


def _set_score(score, s):
    return sum(_pair(score, a, b) for a, b in itertools.combinations(s, 2))


def _partitions(devices, size):
    """iterateGPUPartitions (besteffort_policy.go:216-292), same enumeration order."""
    if size <= 0 or size > len(devices):
        return
    if size == 1:
        for d in devices:
            yield [[d]]
        return
    devices = list(devices)
    while len(devices) % size:
        devices.append(None)
    padding = devices.count(None)

    def rec(devs, accum):
        if size == len(devs):
            yield accum + [devs]
            return
        for tail in itertools.combinations(devs[1:], size - 1):
            st = [devs[0]] + list(tail)
            p = st.count(None)
            if not (p == 0 or p == padding):
                continue
            remaining = [d for d in devs if d not in st]
            yield from rec(remaining, accum + [st])

    yield from rec(devices, [])


def best_effort(available, required, size, score):
    """bestEffortPolicy.Allocate (besteffort_policy.go:34-90)."""
    if size <= 0 or len(available) < size or len(required) > size:
        return []
    best, best_score = None, 0
    for cand in _partitions(available, size):
        if not any(all(r in st for r in required) and st.count(None) == 0 for st in cand):
            continue
        sc = sum(_set_score(score, st) for st in cand)
        if sc > best_score or best is None:
            best, best_score = cand, sc
    filtered = [st for st in (best or []) if all(r in st for r in required)]
    if not filtered:
        return []
    best_set, bs = filtered[0], _set_score(score, filtered[0])
    for st in filtered[1:]:
        sc = _set_score(score, st)
        if sc > bs:
            best_set, bs = st, sc
    return best_set
