"""Pure-Python restatement of the reference's quota semantics (SURVEY.md Appendix B, §8a
a10-a23; libvgpu.so oom_check@0x40528, add_chunk@0x40b73, cuMemGetInfo_v2@0x37348,
cuMemFree_v2@0x32e5d, cuDeviceTotalMem_v2@0x2df64, nvmlDeviceGetMemoryInfo@0x25052; context billing:
cuDevicePrimaryCtxRetain@0x294bd, cuDevicePrimaryCtxRelease_v2@0x29f08, cuCtxCreate_v2@0x2a4dd).
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
        self.ndev = int(env.get("MOCK_GPU_COUNT", "1"))
        self.limits = [limit_from_env(env.get(f"CUDA_DEVICE_MEMORY_LIMIT_{d}")) or base for d in range(self.ndev)]
        self.oversub = env.get("CUDA_OVERSUBSCRIBE") == "true"
        self.phys_total = phys_total
        self.used = [dict() for _ in range(nprocs)]   # per process: tag -> (bytes, real bytes, dev)
        self.freed = [set() for _ in range(nprocs)]
        self.calls = []
        # contexts: the mock hands out one handle per device for the primary context and a fresh one per
        # cuCtxCreate; the runner numbers handles in order of first appearance
        self.ctx_size = int(env.get("MOCK_CTX_MIB", "0")) << 20
        self.ctx_billed = [[0] * self.ndev for _ in range(nprocs)]      # bytes billed as kind 0
        self.ctx_flag = [[False] * self.ndev for _ in range(nprocs)]    # vdevices[d].flag
        self.cur = [("primary", 0)] * nprocs                            # current context handle
        self.stack = [[] for _ in range(nprocs)]
        self.ctags = [dict() for _ in range(nprocs)]
        self.seen = [[] for _ in range(nprocs)]
        self.ncreated = [0] * nprocs
        self.dead = [set() for _ in range(nprocs)]

    def usage(self, dev=0):
        return (sum(n for u in self.used for n, _real, d in u.values() if d == dev) +
                sum(b[dev] for b in self.ctx_billed))

    def _oom(self, n, dev):
        lim = self.limits[dev]
        return lim != 0 and ((self.usage(dev) + n) & U64) > lim

    def _ctx_no(self, p, h):
        if h is None:
            return 0
        if h not in self.seen[p]:
            self.seen[p].append(h)
        return self.seen[p].index(h) + 1

    def _bill(self, p, dev):
        if not self.ctx_flag[p][dev]:
            self.ctx_billed[p][dev] += self.ctx_size
            self.ctx_flag[p][dev] = True

    def _ctx_op(self, p, k, t):
        if k == "retain":
            dev, tag = int(t[1]), int(t[2])
            self._bill(p, dev)
            self.ctags[p][tag] = ("primary", dev)
            return f"0 {self._ctx_no(p, ('primary', dev))}"
        if k == "release":
            self.ctx_flag[p][int(t[1])] = False      # the flag goes, the bytes stay billed
            return "0"
        if k == "ctxcreate":
            dev, tag = int(t[1]), int(t[2])
            self._bill(p, dev)                        # a second create is logged as a duplicate, not billed
            h = ("created", self.ncreated[p], dev)
            self.ncreated[p] += 1
            self.calls.append(f"cuCtxCreate_v2 dev={dev}")
            self.stack[p].append(self.cur[p])
            self.cur[p] = self.ctags[p][tag] = h
            return f"0 {self._ctx_no(p, h)}"
        if k == "ctxdestroy":
            h = self.ctags[p].get(int(t[1]))
            if h is None or h[0] != "created" or h in self.dead[p]:
                return "201"
            self.dead[p].add(h)
            self.calls.append("cuCtxDestroy_v2")
            if self.cur[p] == h:
                self.cur[p] = self.stack[p].pop() if self.stack[p] else ("primary", 0)
            return "0"
        if k == "ctxset":
            h = self.ctags[p].get(int(t[1]))
            if h is not None:
                self.cur[p] = h
            return "0"
        if k == "ctxpush":
            self.stack[p].append(self.cur[p])
            self.cur[p] = self.ctags[p].get(int(t[1]))
            return "0"
        if k == "ctxpop":
            h = self.cur[p]
            if self.stack[p]:
                self.cur[p] = self.stack[p].pop()
            return f"0 {self._ctx_no(p, h)}"
        if k == "ctxget":
            return f"0 {self._ctx_no(p, self.cur[p])}"
        if k == "ctxdev":
            return f"0 {self._dev(p)}"
        raise ValueError(k)

    def _dev(self, p):
        h = self.cur[p]
        return 0 if h is None else h[-1] if h[0] == "created" else h[1]

    def _slots(self):
        toks = []
        for p in range(len(self.used)):
            tok = f"p{p}:1:1"
            for d in range(self.ndev):
                data = sum(n for n, _r, dd in self.used[p].values() if dd == d)
                c = self.ctx_billed[p][d]
                if c or data:
                    tok += f" d{d}={c}/0/{data}/{c + data}"
            toks.append(tok)
        return f"n={len(toks)} " + " | ".join(toks)

    def op(self, p, line):
        t = line.split()
        k = t[0]
        if k == "@slots":
            return self._slots()
        if k in ("retain", "release") or k.startswith("ctx"):
            return self._ctx_op(p, k, t)
        dev = self._dev(p)
        limit = self.limits[dev]
        if k == "total":
            return f"0 {self.limits[int(t[1]) if len(t) > 1 else 0]}"
        if k == "info":
            u = self.usage(dev)
            self.calls.append("cuMemGetInfo_v2") if not (limit and limit < u) else None
            if limit == 0:
                return f"0 {(self.phys_total - u) & U64} {self.phys_total}"
            if limit < u:
                return "1 0 0"
            return f"0 {limit - u} {limit}"
        if k == "ptrattrs":
            self.calls.append("cuPointerGetAttributes n=2")
            return "0 2 0"   # IS_MANAGED is always reported as 0 (cuPointerGetAttributes@0x33cf3)
        if k == "nvmlmem":   # the runner asks about NVML device 0
            u, lim0 = self.usage(0), self.limits[0]
            if lim0 == 0:
                return f"0 {self.phys_total} {self.phys_total} {u}"
            return f"0 {lim0} {(lim0 - u) & U64} {u}"
        if k in ("alloc", "managed", "pitch"):
            if k == "pitch":
                w, h, e, tag = int(t[1], 0), int(t[2], 0), int(t[3], 0), int(t[4], 0)
                n = ((w - 1) // e + 1) * e * h
                real = ((w + 511) // 512) * 512 * h          # the mock's pitch
            else:
                n, tag = int(t[1], 0), int(t[2], 0)
                real = n
            if self._oom(n, dev):
                return "-1" if k == "alloc" else "2"
            if k == "alloc":
                name, flags = ("cuMemAllocManaged", 1) if (self.oversub and n > IPCSIZE) else ("cuMemAlloc_v2", 0)
            elif k == "managed":
                name, flags = "cuMemAllocManaged", 1
            else:
                name, flags = "cuMemAllocPitch_v2", e
            self.calls.append(f"{name} size={real} flags={flags} -> rc=0")
            self.used[p][tag] = (n, real, dev)
            return "0"
        if k == "free":
            tag = int(t[1], 0)
            if tag not in self.used[p]:
                # never allocated (or allocation failed): the runner holds pointer 0 -> cuMemFree_v2(0) -> 0;
                # already freed: the runner still holds the old pointer -> unknown pointer -> -1
                return "-1" if tag in self.freed[p] else "0"
            n, real, _d = self.used[p].pop(tag)
            self.freed[p].add(tag)
            self.calls.append(f"cuMemFree_v2 size={real} -> rc=0")
            return "0"
        if k == "freeptr":
            return "0" if int(t[1], 0) == 0 else "-1"
        if k in ("hostalloc", "hostalloc2"):
            self.calls.append(f"{'cuMemAllocHost_v2' if k == 'hostalloc' else 'cuMemHostAlloc'} size={int(t[1], 0)}")
            if self._oom(0, dev):
                self.calls.append("cuMemFreeHost")
                return "2"
            return "0"
        if k == "hostreg":
            return "2" if self._oom(0, dev) else "0"   # the mock does not log register / unregister
        raise ValueError(line)
