"""Pure-Python restatement of the reference's quota semantics (SURVEY.md Appendix B, §8a
a10-a23; libvgpu.so oom_check@0x40528, add_chunk@0x40b73, cuMemGetInfo_v2@0x37348,
cuMemFree_v2@0x32e5d, cuDeviceTotalMem_v2@0x2df64, nvmlDeviceGetMemoryInfo@0x25052).
TEST INFRASTRUCTURE ONLY.  Pinned by golden vectors captured from the reference binary
(tests/golden/quota_*.json.gz); it plays trace_runner's language."""
import re

U64 = (1 << 64) - 1
IPCSIZE = 0x200000


def limit_from_env(v):
    """get_limit_from_env@0x4186c"""
    if not v:
        return 0
    m = re.match(r"\s*(0[xX][0-9a-fA-F]+|0[0-7]*|[1-9][0-9]*)", v)
    if not m:
        return 0
    n = int(m.group(1), 0) if not re.fullmatch(r"0[0-7]+", m.group(1)) else int(m.group(1), 8)
    rest = v[m.end():m.end() + 1]
    scale = {"G": 1 << 30, "g": 1 << 30, "M": 1 << 20, "m": 1 << 20, "K": 1 << 10, "k": 1 << 10}.get(rest, 1)
    res = (n * scale) & U64
    if n and res // scale != n:
        return 0
    return res


class Model:
    def __init__(self, env, phys_total, nprocs=1):
        base = limit_from_env(env.get("CUDA_DEVICE_MEMORY_LIMIT"))
        self.limit = limit_from_env(env.get("CUDA_DEVICE_MEMORY_LIMIT_0")) or base
        self.oversub = env.get("CUDA_OVERSUBSCRIBE") == "true"
        self.phys_total = phys_total
        self.used = [dict() for _ in range(nprocs)]   # per process: tag -> bytes (tracked allocations)
        self.freed = [set() for _ in range(nprocs)]
        self.real_used = 0                            # what the mock driver has handed out
        self.calls = []

    def usage(self):
        return sum(n for u in self.used for n, _real in u.values())

    def _oom(self, n):
        return self.limit != 0 and ((self.usage() + n) & U64) > self.limit

    def op(self, p, line):
        t = line.split()
        k = t[0]
        if k == "total":
            return f"0 {self.limit}"
        if k == "info":
            u = self.usage()
            self.calls.append("cuMemGetInfo_v2") if not (self.limit and self.limit < u) else None
            if self.limit == 0:
                return f"0 {(self.phys_total - u) & U64} {self.phys_total}"
            if self.limit < u:
                return "1 0 0"
            return f"0 {self.limit - u} {self.limit}"
        if k == "ptrattrs":
            self.calls.append("cuPointerGetAttributes n=2")
            return "0 2 0"   # IS_MANAGED is always reported as 0 (cuPointerGetAttributes@0x33cf3)
        if k == "nvmlmem":
            u = self.usage()
            if self.limit == 0:
                return f"0 {self.phys_total} {self.phys_total} {u}"
            return f"0 {self.limit} {(self.limit - u) & U64} {u}"
        if k in ("alloc", "managed", "pitch"):
            if k == "pitch":
                w, h, e, tag = int(t[1], 0), int(t[2], 0), int(t[3], 0), int(t[4], 0)
                n = ((w - 1) // e + 1) * e * h
                real = ((w + 511) // 512) * 512 * h          # the mock's pitch
            else:
                n, tag = int(t[1], 0), int(t[2], 0)
                real = n
            if self._oom(n):
                return "-1" if k == "alloc" else "2"
            if k == "alloc":
                name, flags = ("cuMemAllocManaged", 1) if (self.oversub and n > IPCSIZE) else ("cuMemAlloc_v2", 0)
            elif k == "managed":
                name, flags = "cuMemAllocManaged", 1
            else:
                name, flags = "cuMemAllocPitch_v2", e
            self.calls.append(f"{name} size={real} flags={flags} -> rc=0")
            self.used[p][tag] = (n, real)
            self.real_used += real
            return "0"
        if k == "free":
            tag = int(t[1], 0)
            if tag not in self.used[p]:
                # never allocated (or allocation failed): the runner holds pointer 0 -> cuMemFree_v2(0) -> 0;
                # already freed: the runner still holds the old pointer -> unknown pointer -> -1
                return "-1" if tag in self.freed[p] else "0"
            n, real = self.used[p].pop(tag)
            self.freed[p].add(tag)
            self.calls.append(f"cuMemFree_v2 size={real} -> rc=0")
            self.real_used -= real
            return "0"
        if k == "freeptr":
            return "0" if int(t[1], 0) == 0 else "-1"
        if k in ("hostalloc", "hostalloc2"):
            self.calls.append(f"{'cuMemAllocHost_v2' if k == 'hostalloc' else 'cuMemHostAlloc'} size={int(t[1], 0)}")
            if self._oom(0):
                self.calls.append("cuMemFreeHost")
                return "2"
            return "0"
        if k == "hostreg":
            return "2" if self._oom(0) else "0"   # the mock does not log register / unregister
        raise ValueError(line)
