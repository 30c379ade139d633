"""External reader of the shared-cache file (Contract C) and the suspend / resume side of the
monitor protocol — SURVEY.md §8(f) row 3.

The reference exposes the file to an outside agent in VGPU_MONITOR_MODE (server.go:494-501) and its
shim reacts to SIGUSR2 (status 2, suspended) / SIGUSR1 (status 1, running) — sig_swap_stub /
sig_restore_stub — but ships no agent.  This is a minimal one: it decodes the file layout of
SURVEY §8a a9 and signals the processes listed in it.  With the new shim's VMM backend a suspend
also moves the process's data out of HBM and a resume brings it back."""
from __future__ import annotations

import os
import signal
import struct
from typing import Dict, List

REGION_SIZE = 0xC4748
MAGIC = 19920718
SLOT0, STRIDE, MAX_PROCS, MAX_DEV = 0x738, 0x310, 1024, 16


def read_region(path: str) -> Dict:
    b = open(path, "rb").read(REGION_SIZE)
    if len(b) < REGION_SIZE or struct.unpack_from("<i", b, 0)[0] != MAGIC:
        raise ValueError(f"{path}: not an initialised vGPU shared cache")
    ndev = min(struct.unpack_from("<Q", b, 0x30)[0], MAX_DEV)
    out = {
        "owner_pid": struct.unpack_from("<Q", b, 8)[0],
        "device_num": ndev,
        "uuids": [b[0x38 + 96 * i: 0x38 + 96 * (i + 1)].split(b"\0")[0].decode() for i in range(ndev)],
        "limit": list(struct.unpack_from("<16Q", b, 0x638))[:max(ndev, 1)],
        "sm_limit": list(struct.unpack_from("<16Q", b, 0x6B8))[:max(ndev, 1)],
        "proc_num": struct.unpack_from("<i", b, 0xC4738)[0],
        "utilization_switch": struct.unpack_from("<i", b, 0xC473C)[0],
        "recent_kernel": struct.unpack_from("<i", b, 0xC4740)[0],
        "priority": struct.unpack_from("<i", b, 0xC4744)[0],
        "procs": [],
    }
    for s in range(min(out["proc_num"], MAX_PROCS)):
        o = SLOT0 + s * STRIDE
        pid, hostpid = struct.unpack_from("<ii", b, o)
        used = []
        for d in range(max(ndev, 1)):
            ctx, mod, data, off, total = struct.unpack_from("<5Q", b, o + 8 + 40 * d)
            used.append({"context_size": ctx, "module_size": mod, "data_size": data, "total": total})
        out["procs"].append({"pid": pid, "hostpid": hostpid, "status": struct.unpack_from("<i", b, o + 0x308)[0], "used": used})
    out["usage"] = [sum(p["used"][d]["total"] for p in out["procs"]) for d in range(max(ndev, 1))]
    return out


def scan(shared_root: str = "/usr/local/vgpu/shared") -> Dict[str, Dict]:
    """VGPU_MONITOR_MODE (server.go:494-501): every container's cache file sits in <shared_root>/<pod>_<container>/
    on the host.  -> {"<pod>_<container>": region dict} for every directory holding an initialised cache file."""
    import glob
    import os
    out = {}
    for d in sorted(glob.glob(os.path.join(shared_root, "*"))):
        for f in sorted(glob.glob(os.path.join(d, "*.cache"))):
            try:
                out[os.path.basename(d)] = read_region(f)
            except (OSError, ValueError):
                continue
    return out


def pids(path: str) -> List[int]:
    """The HOST pids of the container's processes.  The slot's `pid` is the pid inside the container's pid namespace —
    on the host, where this agent runs, that number names an unrelated process (and SIGUSR2's default action is to
    terminate it) — so only `hostpid`, which the shim fills in for this purpose (set_host_pid@0x4497c, from NVML's
    process list), is ever signalled.  Slots whose host pid is not known yet are skipped, see unreachable()."""
    return [p["hostpid"] for p in read_region(path)["procs"] if p["pid"] > 0 and p["hostpid"] > 0]


def unreachable(path: str) -> List[int]:
    """Container pids of live slots without a host pid (no CUDA context yet, or NVML could not see the process)."""
    return [p["pid"] for p in read_region(path)["procs"] if p["pid"] > 0 and p["hostpid"] <= 0]


def _signal_all(path: str, sig: int, kill) -> List[int]:
    done = []
    for p in pids(path):
        try:
            kill(p, sig)
            done.append(p)
        except ProcessLookupError:   # gone between the scan and the signal: its slot is swept by the next process that starts
            pass
    return done


def suspend(path: str, kill=os.kill) -> List[int]:
    """SIGUSR2 to every process of the container: status -> 2, launches park; with the VMM
    backend their data is spilled to the tier and the HBM released."""
    return _signal_all(path, signal.SIGUSR2, kill)


def resume(path: str, kill=os.kill) -> List[int]:
    return _signal_all(path, signal.SIGUSR1, kill)


def _poke_i32(path: str, offset: int, value: int) -> None:
    with open(path, "r+b") as f:
        f.seek(offset)
        f.write(struct.pack("<i", value))


def set_recent_kernel(path: str, value: int) -> None:
    """The monitor's lever in the file (rate_limiter@0x463b1 reads it at every launch): < 0 holds every launch of the
    container, 0 = idle, the shim writes 2 at a launch and the monitor counts it down."""
    _poke_i32(path, 0xC4740, value)


def publish_memory(path: str, by_hostpid: Dict[int, List[int]]) -> int:
    """What set_gpu_device_memory_monitor@0x42e4c does from a linked monitor: store the HBM this agent measured for a
    process (NVML usedGpuMemory, bytes per device column) in its slot's `monitorused`, where MEMORY_OVERRIDE /
    ACTIVE_OOM_KILLER in the shim read it.  -> slots written."""
    r = read_region(path)
    n = 0
    with open(path, "r+b") as f:
        for s, p in enumerate(r["procs"]):
            vals = by_hostpid.get(p["hostpid"])
            if p["hostpid"] <= 0 or vals is None:
                continue
            for d, v in enumerate(vals[:MAX_DEV]):
                f.seek(SLOT0 + s * STRIDE + 0x288 + 8 * d)
                f.write(struct.pack("<Q", int(v)))
            n += 1
    return n


def nvml_process_memory() -> Dict[int, Dict[str, int]]:
    """{host pid: {GPU UUID (lower case): bytes of HBM NVML attributes to the process}} — what a tenant really holds, as
    opposed to what it has allocated against its quota (a tenant that pages itself holds less than its usage; a
    suspended one next to nothing).  Feed it to Scheduler(measured=...) and to publish_memory."""
    import pynvml
    pynvml.nvmlInit()
    out: Dict[int, Dict[str, int]] = {}
    for i in range(pynvml.nvmlDeviceGetCount()):
        h = pynvml.nvmlDeviceGetHandleByIndex(i)
        u = pynvml.nvmlDeviceGetUUID(h)
        u = (u.decode() if isinstance(u, bytes) else u).lower()
        for p in pynvml.nvmlDeviceGetComputeRunningProcesses(h):
            out.setdefault(int(p.pid), {})[u] = int(p.usedGpuMemory or 0)
    return out


# ----------------------------------------------------------------------------- the scheduler
class Tenant:
    """One container (one cache file) as the scheduler sees it."""

    def __init__(self, name: str, path: str):
        self.name, self.path = name, path
        self.region: Dict = {}
        self.suspended = False       # by this scheduler
        self.idle_ticks = 0          # ticks since its last kernel launch
        self.since = 0               # tick of the last suspend / resume
        self.held = False

    @property
    def priority(self) -> int:       # CUDA_TASK_PRIORITY: 0 = most important, larger = yields first (default 1)
        return self.region.get("priority", 1)

    def usage(self, uuid: str) -> int:
        return sum(u for g, u in zip(self.region.get("uuids", []), self.region.get("usage", [])) if g.lower() == uuid.lower())

    def limit(self, uuid: str) -> int:
        return sum(u for g, u in zip(self.region.get("uuids", []), self.region.get("limit", [])) if g.lower() == uuid.lower())

    def alive(self) -> bool:
        return any(p["pid"] > 0 for p in self.region.get("procs", []))


class Scheduler:
    """Decides WHICH vGPUs of an over-committed GPU hold its HBM — SURVEY.md §8(f) row 3: what turns the reference's
    dormant hooks (suspend_all@0x45ddf / resume_all@0x45eea, `status`, `recent_kernel`, CUDA_TASK_PRIORITY; the shared
    directory of server.go:494-501) into an oversubscription scheduler.  A container alone larger than the GPU pages
    itself (the shim's residency engine); this agent arbitrates BETWEEN containers whose quotas add up to more than
    the GPU: per GPU it keeps the running set within `hbm - headroom`, suspending (SIGUSR2: the shim spills the
    process's managed ranges to the tier and gives the HBM back) the least important, longest idle tenants first,
    resuming (SIGUSR1: restore + checksum verify) the most important, longest waiting ones when there is room, and
    rotating tenants of equal importance every `quantum` ticks when not all of them fit (time sharing).

    tick() is one decision round; run() calls it every `period` seconds.  `hbm` maps GPU UUID -> bytes (NVML by
    default); `kill` and `hbm_of` are injectable for tests."""

    def __init__(self, shared_root: str = "/usr/local/vgpu/shared", hbm: Dict[str, int] = None, headroom: int = 2 << 30,
                 quantum: int = 10, kill=os.kill, log=None, hbm_free=None, measured=None):
        self.root, self.hbm, self.headroom, self.quantum, self.kill = shared_root, hbm, headroom, quantum, kill
        # measured() -> {host pid: {uuid: bytes}} (nvml_process_memory): when given, a RUNNING tenant's demand is what it
        # really holds in HBM rather than what it has allocated (a self-paging tenant allocates more than it holds); a
        # tenant about to be resumed is still counted with its allocation — that is what it will want back.
        self.measured = measured
        # hbm_free(uuid) -> bytes really free on the GPU right now, or None to trust the accounting.  A suspend takes
        # time (the tenant's data is on its way to the tier); nobody is resumed into HBM that is not there yet.
        self.hbm_free = hbm_free
        self.tenants: Dict[str, Tenant] = {}
        self.now = 0
        self.log = log or (lambda *a: None)
        self.actions: List[tuple] = []

    # -- inputs
    def _hbm(self, uuid: str) -> int:
        if self.hbm is None:
            import pynvml
            pynvml.nvmlInit()
            self.hbm = {}
            for i in range(pynvml.nvmlDeviceGetCount()):
                h = pynvml.nvmlDeviceGetHandleByIndex(i)
                u = pynvml.nvmlDeviceGetUUID(h)
                self.hbm[(u.decode() if isinstance(u, bytes) else u).lower()] = pynvml.nvmlDeviceGetMemoryInfo(h).total
        return self.hbm.get(uuid.lower(), 0) or next((v for k, v in self.hbm.items() if k.lower() == uuid.lower()), 0)

    def _refresh(self) -> None:
        import glob
        seen = set()
        for d in sorted(glob.glob(os.path.join(self.root, "*"))):
            for f in sorted(glob.glob(os.path.join(d, "*.cache"))):
                try:
                    r = read_region(f)
                except (OSError, ValueError):
                    continue
                name = os.path.basename(d)
                t = self.tenants.get(name) or Tenant(name, f)
                t.path, t.region = f, r
                self.tenants[name] = t
                seen.add(name)
        for name in list(self.tenants):
            if name not in seen or not self.tenants[name].alive():
                self.tenants.pop(name)   # container gone (its exit handler emptied the slots)

    # -- outputs
    def _do(self, what: str, t: Tenant) -> None:
        self.actions.append((what, t.name))
        self.log(what, t.name)
        if what == "suspend":
            set_recent_kernel(t.path, -1)   # brake first: no new launch gets past rate_limiter's check
            t.held = True
            gone = unreachable(t.path)
            if gone:
                self.actions.append(("unreachable", t.name, tuple(gone)))
            suspend(t.path, self.kill)
            t.suspended, t.since = True, self.now
        elif what == "resume":
            resume(t.path, self.kill)
            set_recent_kernel(t.path, 0)
            t.held = False
            t.suspended, t.since, t.idle_ticks = False, self.now, 0

    # -- one round
    def tick(self) -> List[tuple]:
        self.now += 1
        self.actions = []
        self._refresh()
        for t in self.tenants.values():   # activity: the shim marks 2 at a launch, we count it down
            rk = t.region.get("recent_kernel", 0)
            if rk > 0:
                t.idle_ticks = 0
                set_recent_kernel(t.path, rk - 1)
            elif rk == 0 and not t.suspended:
                t.idle_ticks += 1
        gpus = sorted({u.lower() for t in self.tenants.values() for u in t.region.get("uuids", []) if u})
        for g in gpus:
            self._balance(g)
        return self.actions

    def _held(self, t: Tenant, g: str, seen: Dict[int, Dict[str, int]]) -> int:
        """HBM a running tenant occupies on GPU g: measured when every one of its processes is known to NVML, else its
        allocation."""
        if seen is None:
            return t.usage(g)
        pids_ = [p["hostpid"] for p in t.region.get("procs", []) if p["pid"] > 0]
        if not pids_ or any(p <= 0 or p not in seen for p in pids_):
            return t.usage(g)
        return min(t.usage(g), sum(seen[p].get(g, 0) for p in pids_))

    def _balance(self, g: str) -> None:
        cap = self._hbm(g) - self.headroom
        if cap <= 0:
            return
        here = [t for t in self.tenants.values() if t.usage(g) > 0 or t.limit(g) > 0]
        running = [t for t in here if not t.suspended]
        waiting = [t for t in here if t.suspended]
        seen = self.measured() if self.measured else None
        held = {t.name: self._held(t, g, seen) for t in running}
        demand = sum(held.values())
        # 1. over-committed: the least important, longest idle, largest go first; one tenant always stays
        if demand > cap and len(running) > 1:
            for t in sorted(running, key=lambda t: (-t.priority, -t.idle_ticks, -held[t.name], t.name)):
                if demand <= cap or len(running) <= 1:
                    break
                self._do("suspend", t)
                running.remove(t)
                waiting.append(t)
                demand -= held[t.name]
            return
        # 2. room: the most important, longest waiting come back while they fit
        free_now = self.hbm_free(g) if self.hbm_free else None
        for t in sorted(waiting, key=lambda t: (t.priority, t.since, t.name)):
            if demand + t.usage(g) <= cap and (free_now is None or free_now >= t.usage(g) + self.headroom):
                if free_now is not None:
                    free_now -= t.usage(g)
                self._do("resume", t)
                waiting.remove(t)
                running.append(t)
                demand += t.usage(g)
        # 3. time sharing: a tenant that has waited a quantum takes the place of running tenants that are no more
        #    important than it and have had their quantum, longest running first
        for t in sorted(waiting, key=lambda t: (t.priority, t.since, t.name)):
            if self.now - t.since < self.quantum:
                continue
            victims, freed = [], 0
            for r in sorted(running, key=lambda r: (-r.priority, r.since, r.name)):
                if r.priority < t.priority or self.now - r.since < self.quantum:
                    continue
                victims.append(r)
                freed += held.get(r.name, r.usage(g))
                if demand - freed + t.usage(g) <= cap:
                    break
            if victims and demand - freed + t.usage(g) <= cap:
                for r in victims:
                    self._do("suspend", r)
                    running.remove(r)
                    demand -= held.get(r.name, r.usage(g))
                break   # the waiting tenant is resumed by rule 2 on the next tick, once the HBM is really free

    def run(self, period: float = 1.0, ticks: int = 0) -> None:
        import time
        n = 0
        while not ticks or n < ticks:
            self.tick()
            n += 1
            time.sleep(period)


if __name__ == "__main__":  # python monitor.py list|suspend|resume <cache file> | run [shared root] [period s]
    import json
    import sys
    cmd = sys.argv[1] if len(sys.argv) > 1 else ""
    if cmd == "run":  # on a node: demand and free HBM as NVML measures them
        def _free(uuid):
            import pynvml
            pynvml.nvmlInit()
            for i in range(pynvml.nvmlDeviceGetCount()):
                h = pynvml.nvmlDeviceGetHandleByIndex(i)
                u = pynvml.nvmlDeviceGetUUID(h)
                if (u.decode() if isinstance(u, bytes) else u).lower() == uuid.lower():
                    return pynvml.nvmlDeviceGetMemoryInfo(h).free
            return None
        Scheduler(sys.argv[2] if len(sys.argv) > 2 else "/usr/local/vgpu/shared", hbm_free=_free, measured=nvml_process_memory,
                  log=lambda *a: print(*a, flush=True)).run(float(sys.argv[3]) if len(sys.argv) > 3 else 1.0)
    elif cmd == "list":
        print(json.dumps(read_region(sys.argv[2]), indent=1))
    elif cmd == "suspend":
        print(suspend(sys.argv[2]))
    elif cmd == "resume":
        print(resume(sys.argv[2]))
    else:
        sys.exit("usage: monitor.py list|suspend|resume <cache file> | run [shared root] [period seconds]")
