"""Host-side mirror of the reference device plugin for the hot path (Contract A).

Same names, argument meaning and error behaviour as the Go code it mirrors
(server.go, vdevice.go, nvidia.go, main.go flags), so that the parity tests read
like tests of the reference.  All arithmetic and string building happens behind
the C-ABI (vgpu_split / vgpu_lookup / vgpu_sm_limit / vgpu_allocate_env in
include/vgpu.h) — this module only marshals.  MIG strategies, monitor mode and the
legacy VDeviceController are out of scope.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import Callable, Dict, Iterator, List, Optional, Sequence

from . import capi

# pluginapi constants (vendor/k8s.io/kubelet/pkg/apis/deviceplugin/v1beta1/constants.go:19-37)
Healthy = "Healthy"
Unhealthy = "Unhealthy"
Version = "v1beta1"
DevicePluginPath = "/var/lib/kubelet/device-plugins/"
KubeletSocket = DevicePluginPath + "kubelet.sock"

# server.go:44-58
DeviceListStrategyEnvvar = "envvar"
DeviceListStrategyVolumeMounts = "volume-mounts"
DeviceIDStrategyUUID = "uuid"
DeviceIDStrategyIndex = "index"


@dataclass
class Flags:
    """CLI flags that shape Contract A, defaults as in main.go:55-133."""
    migStrategy: str = "none"
    passDeviceSpecs: bool = False
    deviceListStrategy: str = DeviceListStrategyEnvvar
    deviceIDStrategy: str = DeviceIDStrategyUUID
    nvidiaDriverRoot: str = "/"
    deviceSplitCount: int = 2
    deviceMemoryScaling: float = 1.0
    deviceCoresScaling: float = 1.0
    enableLegacyPreferred: bool = False
    pcibusfile: str = "/usr/local/vgpu/pciinfo.vgpu"  # $PCIBUSFILE, main.go:164-185
    smPartitionIndex: bool = False  # no reference flag: add VGPU_SM_PARTITION_INDEX_<n> to the response (vgpu.h, emit_partition_index)
    failOnInitError: bool = True    # main.go:63-69: false = block forever instead of failing when NVML cannot be initialised
    verbose: int = 0                # main.go:126-132


def loadPciInfo(lspci_out: Optional[str] = None, pcibusfile: Optional[str] = None) -> str:
    """start() in main.go:164-185: run `lspci`, keep the bus id of every line that names an NVIDIA device, and write
    them, one per line, to $PCIBUSFILE (mode 0644) when that is set.  The file is what Allocate mounts at
    /usr/local/vgpu/pciinfo.vgpu.  `lspci_out` / `pcibusfile` are injectable (tests); -> the string written."""
    import os
    import subprocess
    if lspci_out is None:
        lspci_out = subprocess.run(["lspci"], capture_output=True, text=True, check=True).stdout   # an error here ends start()
    buf = C.create_string_buffer(len(lspci_out) + 2)
    capi.check(capi.lib.vgpu_pcibus_list(lspci_out.encode(), buf, len(buf)))
    pcibusstr = buf.value.decode()
    if pcibusfile is None:
        pcibusfile = os.environ.get("PCIBUSFILE", "")
    if len(pcibusfile) > 0:
        fd = os.open(pcibusfile, os.O_WRONLY | os.O_CREAT | os.O_TRUNC, 0o644)
        with os.fdopen(fd, "w") as f:
            f.write(pcibusstr)
    return pcibusstr


# main.go:55-133: flag name -> (Flags field, kind, environment variable).  Same names, defaults and env vars, so the
# daemonset manifest of the reference starts this plugin unchanged.
_FLAG_TABLE = {
    "mig-strategy": ("migStrategy", "str", "MIG_STRATEGY"),
    "fail-on-init-error": ("failOnInitError", "bool", "FAIL_ON_INIT_ERROR"),
    "pass-device-specs": ("passDeviceSpecs", "bool", "PASS_DEVICE_SPECS"),
    "device-list-strategy": ("deviceListStrategy", "str", "DEVICE_LIST_STRATEGY"),
    "device-id-strategy": ("deviceIDStrategy", "str", "DEVICE_ID_STRATEGY"),
    "nvidia-driver-root": ("nvidiaDriverRoot", "str", "NVIDIA_DRIVER_ROOT"),
    "device-split-count": ("deviceSplitCount", "uint", "DEVICE_SPLIT_COUNT"),
    "device-memory-scaling": ("deviceMemoryScaling", "float", "DEVICE_MEMORY_SCALING"),
    "device-cores-scaling": ("deviceCoresScaling", "float", "DEVICE_CORES_SCALING"),
    "enable-legacy-preferred": ("enableLegacyPreferred", "bool", "ENABLE_LEGACY_PREFERRED"),
    "verbose": ("verbose", "int", "VERBOSE"),
}
_GO_BOOLS = {"1": True, "t": True, "T": True, "TRUE": True, "true": True, "True": True,
             "0": False, "f": False, "F": False, "FALSE": False, "false": False, "False": False}  # strconv.ParseBool


def _flagValue(name: str, kind: str, text: str, origin: str):
    try:
        if kind == "str":
            return text
        if kind == "bool":
            return _GO_BOOLS[text]
        if kind == "float":
            return float(text)
        v = int(text, 10)
        if kind == "uint" and v < 0:
            raise ValueError(text)
        return v
    except (KeyError, ValueError):
        raise ValueError(f'invalid value "{text}" for {origin}') from None


def parseFlags(argv: List[str], env: Optional[dict] = None) -> Flags:
    """The command line of main.go (urfave/cli v2): `--name value`, `--name=value`, a bare `--name` for a boolean, one or
    two dashes; a flag that is not given takes its environment variable, then its default.  validateFlags runs on the
    result as `c.Before` does.  `argv` without the program name."""
    import os
    env = os.environ if env is None else env
    f = Flags(pcibusfile=env.get("PCIBUSFILE", "") or Flags.pcibusfile)
    given = set()
    i = 0
    while i < len(argv):
        a = argv[i]
        i += 1
        if a == "--":
            break
        if not a.startswith("-") or a == "-":
            raise ValueError(f"unexpected argument: {a}")
        name, eq, val = a.lstrip("-").partition("=")
        if name not in _FLAG_TABLE:
            raise ValueError(f"flag provided but not defined: -{name}")
        fld, kind, _ = _FLAG_TABLE[name]
        if not eq:
            if kind == "bool":
                val = "true"
            else:
                if i >= len(argv):
                    raise ValueError(f"flag needs an argument: -{name}")
                val = argv[i]
                i += 1
        setattr(f, fld, _flagValue(name, kind, val, f"flag -{name}"))
        given.add(name)
    for name, (fld, kind, var) in _FLAG_TABLE.items():
        if name not in given and env.get(var, "") != "":
            setattr(f, fld, _flagValue(name, kind, env[var], f"environment variable {var}"))
    validateFlags(f)
    return f


def validateFlags(f: Flags) -> None:
    """main.go:143-161 — same messages."""
    if f.deviceListStrategy not in (DeviceListStrategyEnvvar, DeviceListStrategyVolumeMounts):
        raise ValueError(f"invalid --device-list-strategy option: {f.deviceListStrategy}")
    if f.deviceIDStrategy not in (DeviceIDStrategyUUID, DeviceIDStrategyIndex):
        raise ValueError(f"invalid --device-id-strategy option: {f.deviceIDStrategy}")
    if f.deviceSplitCount < 1:
        raise ValueError(f"invalid --device-split-count option: {f.deviceSplitCount}")
    if f.deviceMemoryScaling <= 0:
        raise ValueError(f"invalid --device-memory-scaling option: {f.deviceMemoryScaling}")
    if f.deviceCoresScaling <= 0:
        raise ValueError(f"invalid --device-core-scaling option: {f.deviceCoresScaling}")


@dataclass
class Device:
    """nvidia.go:36-41 + buildDevice (148-164).  `memory_mib` stands in for the NVML
    lookup Device2VDevice does (vdevice.go:47; nvml bindings.go:504-507, MiB)."""
    ID: str
    Index: str = "0"
    Paths: List[str] = field(default_factory=list)
    Health: str = Healthy
    numa: Optional[int] = None
    memory_mib: int = 0


@dataclass
class VDevice:
    """vdevice.go:29-33."""
    ID: str
    dev: Device
    memory: int
    Health: str = Healthy
    _desc: Optional[capi.Desc] = None

    @property
    def Topology(self):
        return None if self.dev.numa is None else {"Nodes": [{"ID": self.dev.numa}]}


@dataclass
class Mount:
    ContainerPath: str
    HostPath: str
    ReadOnly: bool = False


@dataclass
class DeviceSpec:
    ContainerPath: str
    HostPath: str
    Permissions: str


@dataclass
class ContainerAllocateResponse:
    Envs: Dict[str, str] = field(default_factory=dict)
    Mounts: List[Mount] = field(default_factory=list)
    Devices: List[DeviceSpec] = field(default_factory=list)
    Annotations: Dict[str, str] = field(default_factory=dict)


class PluginError(Exception):
    """Stands in for the gRPC error a failing RPC returns; str() is the Go error text."""


def Device2VDevice(devices: Sequence[Device], flags: Flags) -> List[VDevice]:
    """vdevice.go:36-58 via vgpu_split."""
    out: List[VDevice] = []
    for d in devices:
        arr = (capi.Desc * max(1, flags.deviceSplitCount))()
        n = capi.check(capi.lib.vgpu_split(d.ID.encode(), d.memory_mib, flags.deviceSplitCount,
                                           flags.deviceMemoryScaling, arr, len(arr)))
        for i in range(n):
            desc = capi.Desc.from_buffer_copy(arr[i])
            out.append(VDevice(ID=desc.id.decode(), dev=d, memory=desc.memory_mib, Health=d.Health, _desc=desc))
    return out


def _desc_array(vdevices: Sequence[VDevice]):
    arr = (capi.Desc * max(1, len(vdevices)))()
    for i, vd in enumerate(vdevices):
        arr[i] = vd._desc
    return arr


def VDevicesByIDs(vdevices: Sequence[VDevice], ids: Sequence[str]) -> List[VDevice]:
    """vdevice.go:61-75 via vgpu_lookup; raises PluginError('unknown device: <id>')."""
    arr = _desc_array(vdevices)
    cids = (C.c_char_p * max(1, len(ids)))(*[s.encode() for s in ids])
    out = (C.POINTER(capi.Desc) * max(1, len(ids)))()
    rc = capi.lib.vgpu_lookup(arr, len(vdevices), cids, len(ids), out)
    if rc == capi.E_UNKNOWN_ID:
        raise PluginError(capi.last_error())
    capi.check(rc)
    base = C.addressof(arr)
    res = []
    for i in range(len(ids)):
        idx = (C.addressof(out[i].contents) - base) // C.sizeof(capi.Desc)
        res.append(vdevices[idx])
    return res


def UniqueDeviceIDs(vdevices: Sequence[VDevice]) -> List[str]:
    """vdevice.go:78-90 (first-seen order)."""
    ids: List[str] = []
    for vd in vdevices:
        if vd.dev.ID not in ids:
            ids.append(vd.dev.ID)
    return ids


def simplePolicy(available: List[str], required: List[str], size: int) -> List[str]:
    """Stand-in for gpuallocator.NewBestEffortPolicy (besteffort_policy.go:34-90) over
    *physical* UUIDs.  The NVLink-topology scoring is out of scope (SURVEY.md §2 row
    20); the edge cases the reference's callers rely on are kept: size<=0, fewer
    available than size, or required not within available -> empty result."""
    if size <= 0 or len(available) < size or len(required) > size:
        return []
    if any(r not in available for r in required):
        return []
    chosen = list(required)
    for a in available:
        if len(chosen) >= size:
            break
        if a not in chosen:
            chosen.append(a)
    return chosen


# vdevice-controller.go:25-30
annRequest = "4paradigm.com/vgpu-request"
annUsing = "4paradigm.com/vgpu-using"
annSep = ","
kubeletDeviceManagerCheckpoint = "kubelet_internal_checkpoint"


def checkpointDump(data: Optional[dict]) -> str:
    """What hashutil.DeepHashObject feeds the hash for a device-manager checkpoint
    (vendor/k8s.io/kubernetes/pkg/util/hash/hash.go:27-36): go-spew's `%#v` of checkpoint.checkpointData
    (vendor/.../devicemanager/checkpoint/checkpoint.go:35-48) with DisableMethods, SortKeys.  The rules, from
    vendor/github.com/davecgh/go-spew/spew/format.go:201-371: a value is preceded by `(<Go type>)` unless it is an
    element of a slice / a key or value of a map; a struct is `{Name:value Name:value}`, a slice `[a b]`, a map
    `map[k:v k:v]` (keys sorted), a nil slice or map `<nil>`, a string its bytes, a byte its decimal value.
    `data` is the JSON object under "Data" as encoding/json would have filled the Go struct: JSON null -> nil,
    AllocResp base64 -> []byte."""
    import base64
    data = data or {}

    def strs(v) -> str:          # []string
        return "<nil>" if v is None else "[" + " ".join(v) + "]"

    def entry(e: dict) -> str:   # checkpoint.PodDevicesEntry, as a slice element (no leading type)
        ar = e.get("AllocResp")
        raw = None if ar is None else base64.b64decode(ar)
        return ("{PodUID:(string)%s ContainerName:(string)%s ResourceName:(string)%s DeviceIDs:([]string)%s AllocResp:([]uint8)%s}" %
                (e.get("PodUID") or "", e.get("ContainerName") or "", e.get("ResourceName") or "", strs(e.get("DeviceIDs")),
                 "<nil>" if raw is None else "[" + " ".join(str(b) for b in raw) + "]"))

    pde, reg = data.get("PodDeviceEntries"), data.get("RegisteredDevices")
    pde_s = "<nil>" if pde is None else "[" + " ".join(entry(e) for e in pde) + "]"
    reg_s = "<nil>" if reg is None else "map[" + " ".join(f"{k}:{strs(reg[k])}" for k in sorted(reg)) + "]"
    return ("(checkpoint.checkpointData){PodDeviceEntries:([]checkpoint.PodDevicesEntry)%s "
            "RegisteredDevices:(map[string][]string)%s}" % (pde_s, reg_s))


def checkpointChecksum(data: Optional[dict]) -> int:
    """checksum.New (vendor/k8s.io/kubernetes/pkg/kubelet/checkpointmanager/checksum/checksum.go:36-45): FNV-1a, 32 bit,
    over checkpointDump, widened to uint64."""
    h = 0x811C9DC5
    for b in checkpointDump(data).encode():
        h = ((h ^ b) * 0x01000193) & 0xFFFFFFFF
    return h


class VDeviceController:
    """vdevice-controller.go:32-286 — the legacy preferred-allocation bookkeeping used with
    --enable-legacy-preferred: idMap[vdevID] -> the request id it was handed out for ("" = free).
    The pod lister and the kubelet checkpoint are injected (a fake lister and a checkpoint file in
    tests).  The checkpoint is read the way the kubelet's checkpoint manager reads it: its checksum is verified
    (checkpointChecksum) and a mismatch is "checkpoint is corrupted" (checkpointmanager/errors)."""

    def __init__(self, deviceIDs: Sequence[str], podLister: Optional[Callable[[], List[dict]]] = None,
                 checkpointDir: str = DevicePluginPath):
        self.idMap: Dict[str, str] = {v: "" for v in deviceIDs}   # insertion-ordered; Go's map order is random
        self.podLister = podLister or (lambda: [])
        self.checkpointDir = checkpointDir

    # vdevice-controller.go:231-243
    def available(self) -> List[str]:
        return [k for k, v in self.idMap.items() if v == ""]

    # vdevice-controller.go:246-261
    def acquire(self, request: Sequence[str], using: Sequence[str]) -> None:
        for i, v in enumerate(using):
            if v not in self.idMap:
                continue  # "Error: device %s unknown"
            self.idMap[v] = request[i] if i < len(request) else "mismatched"

    # vdevice-controller.go:264-274
    def release(self, using: Sequence[str]) -> None:
        for v in using:
            if v in self.idMap:
                self.idMap[v] = ""

    # vdevice-controller.go:277-286
    def releaseByRequest(self, request: Sequence[str]) -> None:
        for k, v in self.idMap.items():
            for r in request:
                if v == r:          # "Error: device %s[%s] loss."
                    self.idMap[k] = ""

    # vdevice-controller.go:60-111
    def updateFromCheckpoint(self) -> None:
        import base64
        import json
        import os
        from . import grpc_server as G
        path = os.path.join(self.checkpointDir, kubeletDeviceManagerCheckpoint)
        try:
            cp = json.load(open(path))
        except (OSError, ValueError) as e:
            raise PluginError(f"read checkpoint error, {e}")
        # GetCheckpoint -> VerifyChecksum (vdevice-controller.go:64; checkpoint.go:75-78)
        if not isinstance(cp, dict) or int(cp.get("Checksum") or 0) != checkpointChecksum(cp.get("Data")):
            raise PluginError("read checkpoint error, checkpoint is corrupted")
        pods = self.podLister()
        for pde in (cp.get("Data") or {}).get("PodDeviceEntries") or []:
            if pde.get("ResourceName") != "nvidia.com/gpu":
                continue
            try:
                resp = G.dec_allocate_container(base64.b64decode(pde.get("AllocResp") or ""))
            except Exception:
                continue  # "unmarshal container allocate response failed"
            requestStr = resp["Annotations"].get(annRequest, "")
            usingStr = resp["Annotations"].get(annUsing, "")
            if requestStr == "" and usingStr == "":
                continue
            request, using = requestStr.split(annSep), usingStr.split(annSep)
            pod = next((p for p in pods if p.get("uid") == pde.get("PodUID")), None)
            if pod is not None and pod.get("phase") in ("Pending", "Running"):
                self.acquire(request, using)
            else:
                self.release(using)

    # vdevice-controller.go:114-159 (informer callbacks; dormant in the reference, kept for the mirror)
    def onAddPod(self, pod: dict) -> None:
        ann = pod.get("annotations", {})
        requestStr, usingStr = ann.get(annRequest, ""), ann.get(annUsing, "")
        if requestStr == "" and usingStr == "":
            return
        self.acquire(requestStr.split(annSep), usingStr.split(annSep))

    def onDeletePod(self, pod: dict) -> None:
        usingStr = pod.get("annotations", {}).get(annUsing, "")
        if usingStr:
            self.release(usingStr.split(annSep))


XidCriticalError = 8  # nvml.XidCriticalError (nvmlEventTypeXidCriticalError)


def checkHealth(devices: Sequence[Device], events: Sequence[dict], disable: str = "") -> List[Device]:
    """nvidia.go:166-237 over an injected event source (NVML's WaitForEvent loop stays with the Go
    host).  Returns, in order, the devices that would be sent down the `unhealthy` channel.
    event = {"Etype": int, "Edata": xid, "UUID": str|None, "GpuInstanceId": int, "ComputeInstanceId": int,
             "err": bool}; `disable` = $DP_DISABLE_HEALTHCHECKS."""
    disable = disable.lower()
    if disable == "all":
        disable = "xids"
    if "xids" in disable:
        return []
    out: List[Device] = []
    for e in events:
        if e.get("err") and e.get("Etype") != XidCriticalError:
            continue
        if e.get("Edata") in (31, 43, 45):  # application errors: the GPU is still healthy
            continue
        uuid = e.get("UUID")
        if not uuid:
            out.extend(devices)             # "All devices will go unhealthy."
            continue
        for d in devices:
            # a full-GPU UUID (not a MIG device): gi = ci = 0xFFFFFFFF
            if d.ID == uuid and e.get("GpuInstanceId", 0xFFFFFFFF) == 0xFFFFFFFF and e.get("ComputeInstanceId", 0xFFFFFFFF) == 0xFFFFFFFF:
                out.append(d)
    return out


def bestEffortPolicy(pair_score: Dict[str, Dict[str, int]]) -> Callable[[List[str], List[str], int], List[str]]:
    """gpuallocator.NewBestEffortPolicy() over physical UUIDs (besteffort_policy.go:34-90) through
    vgpu_best_effort.  `pair_score[a][b]` is calculateGPUPairScore(a, b) (:296-352): 100 per NVLink
    between the two GPUs, 10..60 for PCIe proximity — what gpuallocator.NewDevicesFrom would read from
    NVML; a mock supplies it in tests."""
    uuids = sorted(pair_score)
    index = {u: i for i, u in enumerate(uuids)}
    n = len(uuids)
    flat = (C.c_int * max(1, n * n))()
    for a in uuids:
        for b, v in pair_score[a].items():
            flat[index[a] * n + index[b]] = int(v)

    def allocate(available: List[str], required: List[str], size: int) -> List[str]:
        if any(u not in index for u in list(available) + list(required)):
            return []
        av = (C.c_int * max(1, len(available)))(*[index[u] for u in available])
        rq = (C.c_int * max(1, len(required)))(*[index[u] for u in required])
        out = (C.c_int * max(1, size if size > 0 else 1))()
        k = capi.check(capi.lib.vgpu_best_effort(av, len(available), rq, len(required), size, flat, n, out, len(out)))
        return [uuids[out[i]] for i in range(k) if out[i] >= 0]

    return allocate


# NVML_TOPOLOGY_* -> points (besteffort_policy.go:313-325 through nvml.go:592-615 GetP2PLink)
_TOPOLOGY_POINTS = {0: 60,    # INTERNAL   -> P2PLinkSameBoard
                    10: 50,   # SINGLE     -> P2PLinkSingleSwitch
                    20: 40,   # MULTIPLE   -> P2PLinkMultiSwitch
                    30: 30,   # HOSTBRIDGE -> P2PLinkHostBridge
                    40: 20,   # NODE / CPU -> P2PLinkSameCPU
                    50: 10}   # SYSTEM     -> P2PLinkCrossCPU


def pairScoresFromNvml(uuids: Sequence[str], nvml=None, max_links: int = 18) -> Dict[str, Dict[str, int]]:
    """gpuallocator.NewDevicesFrom (device.go:33-95) + calculateGPUPairScore (besteffort_policy.go:296-352) for the
    GPUs `uuids`: per ordered pair, the PCIe-proximity points of nvmlDeviceGetTopologyCommonAncestor plus 100 per
    NVLink whose remote end is the other GPU (nvml.go:617-660 GetNVLink compares nvmlDeviceGetNvLinkRemotePciInfo bus
    ids; capped at 12 links as there).  Behind an NVSwitch every remote end is a switch, so — as in the reference,
    whose GetNVLink carries a TODO for it — no NVLink points are scored and the PCIe tree decides.
    `nvml` defaults to pynvml (already initialised by the caller or initialised here)."""
    if nvml is None:
        import pynvml as nvml
        nvml.nvmlInit()
    handles = {}
    for u in uuids:
        try:
            handles[u] = nvml.nvmlDeviceGetHandleByUUID(u.encode() if isinstance(u, str) else u)
        except Exception as e:  # NewDevicesFrom: "no device with uuid"
            raise PluginError(f"no device with uuid: {u}") from e

    def bus_id(h):
        b = nvml.nvmlDeviceGetPciInfo(h).busId
        return (b.decode() if isinstance(b, bytes) else b).lower()

    def remote_bus_ids(h):
        out = []
        for link in range(max_links):
            try:
                b = nvml.nvmlDeviceGetNvLinkRemotePciInfo(h, link).busId
            except Exception:  # link not present / not supported: skipped, as deviceGetAllNvLinkRemotePciInfo does
                continue
            out.append((b.decode() if isinstance(b, bytes) else b).lower())
        return out

    bus = {u: bus_id(h) for u, h in handles.items()}
    remotes = {u: remote_bus_ids(h) for u, h in handles.items()}
    score: Dict[str, Dict[str, int]] = {u: {} for u in uuids}
    for a in uuids:
        for b in uuids:
            if a == b:
                continue
            pts = 0
            try:
                pts += _TOPOLOGY_POINTS.get(int(nvml.nvmlDeviceGetTopologyCommonAncestor(handles[a], handles[b])), 0)
            except Exception:
                pass  # P2PLinkUnknown: no points
            pts += 100 * min(12, sum(1 for r in remotes[a] if r == bus[b]))
            score[a][b] = pts
    return score


def matchPendingPod(pods: Sequence[dict], containerRequests: Sequence[Sequence[str]], resourceName: str = "nvidia.com/gpu") -> Optional[dict]:
    """server.go:380-405 (VGPU_MONITOR_MODE): the pod this AllocateRequest is for = the LAST pending pod whose
    GPU-requesting containers, in order, ask for exactly as many devices as the request's container entries
    (containers without the resource are skipped).  Pods are plain dicts in client-go's JSON shape.  The reference
    indexes past the end of ContainerRequests (a Go panic) when a pending pod has more GPU containers than the
    request, and when the last "matching" pod has no GPU container at all; here such pods simply do not match."""
    target = None
    for pod in pods:
        if pod.get("status", {}).get("phase") != "Pending":
            continue
        match, minus, gpu_ctrs = True, 0, 0
        for ctridx, ctr in enumerate(pod.get("spec", {}).get("containers", [])):
            limits = (ctr.get("resources") or {}).get("limits") or {}
            if resourceName not in limits:
                minus += 1
                continue
            k = ctridx - minus
            gpu_ctrs += 1
            if k >= len(containerRequests) or int(limits[resourceName]) != len(containerRequests[k]):
                match = False
                break
        if match and gpu_ctrs > 0:  # the reference also "matches" pending pods without any GPU container, then panics at :419
            target = pod
    return target


def gpuContainerNames(pod: dict, n: int, resourceName: str = "nvidia.com/gpu") -> List[str]:
    """server.go:416-431: the names of the pod's first n containers that request the resource."""
    names = [c["name"] for c in pod.get("spec", {}).get("containers", [])
             if resourceName in ((c.get("resources") or {}).get("limits") or {})]
    if len(names) < n:
        raise PluginError("pod has fewer GPU containers than the allocate request")  # Go: index out of range
    return names[:n]


class ResourceManager:
    """nvidia.go:43-46; here a list supplied by the caller (a mock NVML in tests)."""

    def __init__(self, devices: Sequence[Device]):
        self._devices = list(devices)

    def Devices(self) -> List[Device]:
        return list(self._devices)


class NvidiaDevicePlugin:
    """server.go:62-655, non-MIG path (legacy controller and VGPU_MONITOR_MODE included)."""

    def __init__(self, resourceName: str, resourceManager: ResourceManager, deviceListEnvvar: str,
                 allocatePolicy: Optional[Callable[[List[str], List[str], int], List[str]]], socket: str,
                 flags: Optional[Flags] = None):
        self.resourceName = resourceName
        self.resourceManager = resourceManager
        self.deviceListEnvvar = deviceListEnvvar
        self.allocatePolicy = allocatePolicy
        self.socket = socket
        self.flags = flags or Flags()
        validateFlags(self.flags)
        self.migStrategy = self.flags.migStrategy
        self.vDeviceController: Optional["VDeviceController"] = None  # --enable-legacy-preferred
        self._podLister: Optional[Callable[[], List[dict]]] = None     # injected: client-go lister stand-in
        self._checkpointDir = DevicePluginPath
        self._sharedRoot = "/usr/local/vgpu/shared"  # server.go:496, host side of the monitor-mode mount
        self.cachedDevices: List[Device] = []
        self.vDevices: List[VDevice] = []
        self._health_events: List[Device] = []

    # server.go:98-115
    def initialize(self) -> None:
        self.cachedDevices = self.resourceManager.Devices()
        self.vDevices = Device2VDevice(self.cachedDevices, self.flags)
        if self.flags.enableLegacyPreferred:  # server.go:104-111
            self.vDeviceController = VDeviceController([vd.ID for vd in self.vDevices], self._podLister, self._checkpointDir)

    # server.go:221-243 — what Register sends to kubelet
    def Register(self) -> dict:
        return {
            "Version": Version,
            "Endpoint": self.socket.rsplit("/", 1)[-1],
            "ResourceName": self.resourceName,
            "Options": {"GetPreferredAllocationAvailable": self.allocatePolicy is not None and self.vDeviceController is None},
        }

    # server.go:246-251
    def GetDevicePluginOptions(self) -> dict:
        return {"GetPreferredAllocationAvailable": self.allocatePolicy is not None}

    # server.go:583-596
    def apiDevices(self) -> List[dict]:
        out = []
        if self.migStrategy == "none":
            for d in self.vDevices:
                d.Health = d.dev.Health
                out.append({"ID": d.ID, "Health": d.Health, "Topology": d.Topology})
        else:
            for d in self.cachedDevices:
                out.append({"ID": d.ID, "Health": d.Health, "Topology": None})
        return out

    def markUnhealthy(self, dev: Device) -> None:
        """What checkHealth does through the m.health channel (nvidia.go:166-237)."""
        self._health_events.append(dev)

    # server.go:254-268 — yields the initial list, then one list per health event
    def ListAndWatch(self) -> Iterator[List[dict]]:
        yield self.apiDevices()
        while self._health_events:
            d = self._health_events.pop(0)
            d.Health = Unhealthy  # "no way to recover from the Unhealthy state"
            yield self.apiDevices()

    # server.go:271-326
    def GetPreferredAllocation(self, containerRequests: Sequence[dict]) -> Optional[List[List[str]]]:
        if self.migStrategy == "mixed":
            return None
        responses = []
        for req in containerRequests:
            try:
                availableVDev = VDevicesByIDs(self.vDevices, req.get("AvailableDeviceIDs", []))
            except PluginError as e:
                raise PluginError(f"Unable to retrieve list of available vdevices: {e}")
            available = UniqueDeviceIDs(availableVDev)
            try:
                requiredVDev = VDevicesByIDs(self.vDevices, req.get("MustIncludeDeviceIDs", []))
            except PluginError as e:
                raise PluginError(f"Unable to retrieve list of available vdevices: {e}")
            required = UniqueDeviceIDs(requiredVDev)
            size = int(req.get("AllocationSize", 0))
            allocated = self.allocatePolicy(available, required, size)
            if len(allocated) == 0 and len(available) >= size:
                allocated = available[0:size]
            deviceIds = []
            for uuid in allocated:
                for vd in availableVDev:
                    if vd.dev.ID == uuid:
                        deviceIds.append(vd.ID)
                        break
            responses.append(deviceIds)
        return responses

    # server.go:361-533
    # server.go:328-358
    def MIGAllocate(self, containerRequests: Sequence[Sequence[str]]) -> List[ContainerAllocateResponse]:
        f = self.flags
        known = {d.ID for d in self.cachedDevices}
        responses = []
        for ids in containerRequests:
            for i in ids:
                if i not in known:  # deviceExists, server.go:333
                    raise PluginError(f"invalid allocation request for '{self.resourceName}': unknown device: {i}")
            resp = ContainerAllocateResponse()
            if f.deviceIDStrategy == DeviceIDStrategyUUID:  # deviceIDsFromUUIDs, server.go:565-581
                deviceIDs = list(ids)
            else:
                deviceIDs = [d.Index for d in self.cachedDevices if d.ID in ids]
            root = "/var/run/nvidia-container-devices"
            if f.deviceListStrategy == DeviceListStrategyEnvvar:
                resp.Envs[self.deviceListEnvvar] = ",".join(deviceIDs)
            else:
                resp.Envs[self.deviceListEnvvar] = root
                resp.Mounts = [Mount(f"{root}/{i}", "/dev/null", False) for i in deviceIDs]
            if f.passDeviceSpecs:
                resp.Devices = self.apiDeviceSpecs(f.nvidiaDriverRoot, list(ids))
            responses.append(resp)
        return responses

    def Allocate(self, containerRequests: Sequence[Sequence[str]], _cache_uuid: Optional[str] = None) -> List[ContainerAllocateResponse]:
        if self.migStrategy == "mixed":  # server.go:362-364 (set on the per-MIG-resource plugins only, mig-strategy.go:205)
            return self.MIGAllocate(containerRequests)
        f = self.flags
        cfg = capi.PluginCfg(split=f.deviceSplitCount, mem_scaling=f.deviceMemoryScaling,
                             cores_scaling=f.deviceCoresScaling,
                             id_strategy=0 if f.deviceIDStrategy == DeviceIDStrategyUUID else 1,
                             list_strategy=0 if f.deviceListStrategy == DeviceListStrategyEnvvar else 1,
                             pcibusfile=f.pcibusfile.encode(), emit_partition_index=1 if f.smPartitionIndex else 0)
        n_phys = len(self.cachedDevices)
        phys_uuid = (C.c_char_p * max(1, n_phys))(*[d.ID.encode() for d in self.cachedDevices])
        phys_idx = (C.c_char_p * max(1, n_phys))(*[d.Index.encode() for d in self.cachedDevices])
        responses = []
        import os
        monitorMode = os.environ.get("VGPU_MONITOR_MODE", "")
        ctrNames: List[str] = []
        targetPod = None
        if monitorMode:  # server.go:365-406
            if self._podLister is None:
                raise PluginError("VGPU_MONITOR_MODE needs a pod lister")  # the reference panics without in-cluster config
            targetPod = matchPendingPod(self._podLister(), containerRequests, self.resourceName)
            if targetPod is None:
                raise PluginError("no pending pod matches the allocate request")  # Go: nil-slice index panic at :419
            ctrNames = gpuContainerNames(targetPod, len(containerRequests), self.resourceName)
        if self.vDeviceController is not None:
            self.vDeviceController.updateFromCheckpoint()  # server.go:408-413: release devices of finished pods
        for reqidx, ids in enumerate(containerRequests):
            reqDeviceIDs = list(ids)
            if self.vDeviceController is not None:  # server.go:433-457
                c = self.vDeviceController
                c.releaseByRequest(ids)  # "fix kubelet shutdown after Allocate"
                availableIds = c.available()
                if len(availableIds) < len(ids):
                    raise PluginError("no enough devices")
                pref = self.GetPreferredAllocation([{"AllocationSize": len(ids), "AvailableDeviceIDs": availableIds}])
                if pref is not None and len(pref[0]) == len(ids):
                    reqDeviceIDs = pref[0]
                else:
                    reqDeviceIDs = availableIds[0:len(ids)]  # "Warn: get preferred failed"
                c.acquire(ids, reqDeviceIDs)
            vdevices = VDevicesByIDs(self.vDevices, reqDeviceIDs)
            req = (C.POINTER(capi.Desc) * max(1, len(vdevices)))(*[C.pointer(vd._desc) for vd in vdevices])
            envs = (capi.KV * 64)()
            mounts = (capi.Mount * 64)()
            n_envs, n_mounts = C.c_size_t(64), C.c_size_t(64)
            shared = None
            if monitorMode:  # server.go:494-501
                shared = targetPod["metadata"]["name"] + "_" + ctrNames[reqidx]
                os.makedirs(os.path.join(self._sharedRoot, shared), exist_ok=True)
            capi.check(capi.lib.vgpu_allocate_env_monitor(C.byref(cfg), req, len(vdevices), phys_uuid, phys_idx, n_phys,
                                                          _cache_uuid.encode() if _cache_uuid else None,
                                                          shared.encode() if shared else None,
                                                          envs, C.byref(n_envs), mounts, C.byref(n_mounts)))
            resp = ContainerAllocateResponse()
            for i in range(n_envs.value):
                resp.Envs[envs[i].key.decode()] = envs[i].val.decode()
            for i in range(n_mounts.value):
                resp.Mounts.append(Mount(mounts[i].container_path.decode(), mounts[i].host_path.decode(), bool(mounts[i].read_only)))
            if f.passDeviceSpecs:
                resp.Devices = self.apiDeviceSpecs(f.nvidiaDriverRoot, UniqueDeviceIDs(vdevices))
            if self.vDeviceController is not None:  # server.go:480-485
                resp.Annotations[annRequest] = annSep.join(ids)
                resp.Annotations[annUsing] = annSep.join(reqDeviceIDs)
                self.vDeviceController.acquire(ids, reqDeviceIDs)
            responses.append(resp)
        return responses

    # server.go:618-655 (the four control nodes are included only if they exist on the host)
    def apiDeviceSpecs(self, driverRoot: str, uuids: Sequence[str], _exists=None) -> List[DeviceSpec]:
        import os
        exists = _exists or os.path.exists
        specs = []
        for p in ("/dev/nvidiactl", "/dev/nvidia-uvm", "/dev/nvidia-uvm-tools", "/dev/nvidia-modeset"):
            if exists(p):
                specs.append(DeviceSpec(p, os.path.join(driverRoot, p.lstrip("/")) if driverRoot != "/" else p, "rw"))
        for d in self.cachedDevices:
            for u in uuids:
                if d.ID == u:
                    for p in d.Paths:
                        specs.append(DeviceSpec(p, os.path.join(driverRoot, p.lstrip("/")) if driverRoot != "/" else p, "rw"))
        return specs


def NewNvidiaDevicePlugin(resourceManager: ResourceManager, flags: Optional[Flags] = None,
                          pair_score: Optional[Dict[str, Dict[str, int]]] = None) -> NvidiaDevicePlugin:
    """migStrategyNone.GetPlugins (mig-strategy.go:62-71): the best-effort policy when the caller
    supplies the NVML link scores, else the topology-blind stand-in."""
    policy = bestEffortPolicy(pair_score) if pair_score else simplePolicy
    return NvidiaDevicePlugin("nvidia.com/gpu", resourceManager, "NVIDIA_VISIBLE_DEVICES", policy,
                              DevicePluginPath + "nvidia-gpu.sock", flags)


# ------------------------------------------------------------------ MIG strategies (mig-strategy.go, mig.go, nvidia.go:62-137)
@dataclass
class MigDevice:
    """What the strategies read from a MIG device: nvml.Device.GetAttributes() and its node paths."""
    ID: str
    GpuInstanceSliceCount: int
    ComputeInstanceSliceCount: int
    MemorySizeMB: int
    Paths: List[str] = field(default_factory=list)


@dataclass
class NodeGpu:
    """One physical GPU as NVML lists it: index order is NVML's."""
    ID: str
    Path: str
    migEnabled: bool = False
    migs: List[MigDevice] = field(default_factory=list)
    memory_mib: int = 0


def migResourceName(mig: MigDevice) -> str:
    """getResourceName (mig-strategy.go:139-154 / 218-233): mig-<g>g.<gb>gb, or mig-<c>c.<g>g.<gb>gb when the compute
    slice count differs; gb = ceil(MemorySizeMB / 1024)."""
    g, c = mig.GpuInstanceSliceCount, mig.ComputeInstanceSliceCount
    gb = (mig.MemorySizeMB + 1024 - 1) // 1024
    return f"mig-{g}g.{gb}gb" if g == c else f"mig-{c}c.{g}g.{gb}gb"


def validMigDevice(mig: MigDevice) -> bool:
    """mig-strategy.go:132-137 / 211-216."""
    return mig.GpuInstanceSliceCount == mig.ComputeInstanceSliceCount


class GpuDeviceManager(ResourceManager):
    """nvidia.go:66-102: every GPU, optionally skipping the MIG-enabled ones; Index = NVML index."""

    def __init__(self, node: Sequence[NodeGpu], skipMigEnabledGPUs: bool):
        super().__init__([Device(g.ID, Index=str(i), Paths=[g.Path], memory_mib=g.memory_mib)
                          for i, g in enumerate(node) if not (g.migEnabled and skipMigEnabledGPUs)])


class MigDeviceManager(ResourceManager):
    """nvidia.go:104-137: the MIG devices of MIG-enabled GPUs that match `resource`; Index = "<gpu>:<mig>"."""

    def __init__(self, node: Sequence[NodeGpu], matches: Callable[[MigDevice, str], bool], resource: str):
        super().__init__([Device(m.ID, Index=f"{i}:{j}", Paths=list(m.Paths), memory_mib=m.MemorySizeMB)
                          for i, g in enumerate(node) if g.migEnabled for j, m in enumerate(g.migs) if matches(m, resource)])


def migPlugins(strategy: str, node: Sequence[NodeGpu], flags: Optional[Flags] = None,
               pair_score: Optional[Dict[str, Dict[str, int]]] = None) -> List[NvidiaDevicePlugin]:
    """NewMigStrategy(strategy).GetPlugins() (mig-strategy.go:46-209) over an explicit node description instead of live
    NVML.  Go panics become PluginError with the reference's message."""
    policy = bestEffortPolicy(pair_score) if pair_score else simplePolicy
    sock = DevicePluginPath
    if strategy == "none":  # :62-71 — "Enumerate device even if MIG enabled"
        return [NvidiaDevicePlugin("nvidia.com/gpu", GpuDeviceManager(node, False), "NVIDIA_VISIBLE_DEVICES", policy,
                                   sock + "nvidia-gpu.sock", flags)]
    if strategy not in ("single", "mixed"):
        raise PluginError(f"Unknown strategy: {strategy}")
    enabled = [g for g in node if g.migEnabled]
    if strategy == "single":  # :78-130
        if not enabled:
            return migPlugins("none", node, flags, pair_score)  # "No MIG devices found. Falling back to mig.strategy=none"
        if any(not g.migEnabled for g in node):
            raise PluginError("For mig.strategy=single all devices on the node must all be configured with the same migEnabled value")
    for g in enabled:  # AssertAllMigEnabledDevicesAreValid, mig.go:84-100
        if not g.migs:
            raise PluginError(f"At least one device with migEnabled=true was not configured correctly: No MIG devices associated with {g.Path}: {g.ID}")
    resources: List[str] = []
    for g in enabled:
        for m in g.migs:
            r = migResourceName(m)
            if not validMigDevice(m):
                if strategy == "single":
                    raise PluginError("Unsupported MIG device found: " + r)
                continue  # mixed: "Skipping unsupported MIG device"
            if r not in resources:
                resources.append(r)
    if strategy == "single":
        if not resources:
            raise PluginError("No MIG devices present on node")
        if len(resources) != 1:
            raise PluginError("More than one MIG device type present on node")
        return [NvidiaDevicePlugin("nvidia.com/gpu", MigDeviceManager(node, lambda m, r: True, "gpu"), "NVIDIA_VISIBLE_DEVICES", None,
                                   sock + "nvidia-gpu.sock", flags)]
    plugins = [NvidiaDevicePlugin("nvidia.com/gpu", GpuDeviceManager(node, True), "NVIDIA_VISIBLE_DEVICES", policy,
                                  sock + "nvidia-gpu.sock", flags)]
    for r in sorted(resources):  # Go iterates a map: order unspecified; sorted here
        p = NvidiaDevicePlugin("nvidia.com/" + r, MigDeviceManager(node, lambda m, res: migResourceName(m) == res, r),
                               "NVIDIA_VISIBLE_DEVICES", None, sock + "nvidia-" + r + ".sock", flags)
        p.migStrategy = "mixed"
        plugins.append(p)
    return plugins
