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
    return [p["pid"] for p in read_region(path)["procs"] if p["pid"] > 0]


def suspend(path: str) -> List[int]:
    """SIGUSR2 to every process of the container: status -> 2, launches park; with the VMM
    backend their data is spilled to the host tier and the HBM released."""
    ps = pids(path)
    for p in ps:
        os.kill(p, signal.SIGUSR2)
    return ps


def resume(path: str) -> List[int]:
    ps = pids(path)
    for p in ps:
        os.kill(p, signal.SIGUSR1)
    return ps


if __name__ == "__main__":  # python monitor.py list|suspend|resume <cache file>
    import json
    import sys
    cmd, path = sys.argv[1], sys.argv[2]
    if cmd == "list":
        print(json.dumps(read_region(path), indent=1))
    elif cmd == "suspend":
        print(suspend(path))
    elif cmd == "resume":
        print(resume(path))
    else:
        sys.exit("usage: monitor.py list|suspend|resume <cache file>")
