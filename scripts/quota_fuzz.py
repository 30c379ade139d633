#!/usr/bin/env python
"""Kitchen-sink differential fuzzer for the quota path: random traces mixing every op the runner knows (device,
managed, pitch and pinned-host allocations, frees, queries, context retain / release, the monitor's helpers, fork,
SIGKILL, launches, live quota resizes) are played through the reference binary and the new shim; replies, OOM log
lines, the mock driver's call log and the final cache file must be identical.  Needs /root/reference.

  python scripts/quota_fuzz.py [first_seed] [last_seed]
"""
import math
import os
import random
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import quota_harness as H  # noqa: E402

M = 1 << 20
ENVS = ({}, {"MEMORY_OVERRIDE": "1"}, {"CUDA_OVERSUBSCRIBE": "true", "CUDA_DEVICE_SM_LIMIT": "40", "CUDA_TASK_PRIORITY": "0"})


def gen(seed, nprocs=4, nops=600, ndev=1, nlate=0):
    """`nlate` of the processes (the highest numbers) join at random points: they claim a slot in a file with history."""
    rng = random.Random(seed)
    early = nprocs - nlate
    alive = set(range(early))
    live = [[] for _ in range(nprocs)]
    tag = [1] * nprocs

    def prologue(p):
        out = [(p, "retain 0 1")]
        if ndev == 2:  # context tag d+1 = primary context of device d; every process starts on device 0
            out += [(p, "retain 1 2"), (p, "ctxset 1")]
        return out

    tr = [op for p in range(early) for op in prologue(p)]
    joins = {p: rng.randrange(nops // 5, 4 * nops // 5) for p in range(early, nprocs)}
    cur = [0] * nprocs   # current device of each process
    where = [dict() for _ in range(nprocs)]
    for step in range(nops):
        for q, when in joins.items():
            if when == step:
                tr += [(q, "@start")] + prologue(q)
                alive.add(q)
        p = rng.choice(sorted(alive))
        r = rng.random()
        if ndev == 2 and rng.random() < 0.08:
            cur[p] = rng.randrange(2)
            tr.append((p, f"ctxset {cur[p] + 1}"))
            continue
        if r < 0.35:
            size = int(math.exp(rng.uniform(math.log(4096), math.log(300 * M))))
            t = tag[p]
            tag[p] += 1
            k = rng.random()
            if k < 0.7:
                tr.append((p, f"alloc {size} {t}"))
            elif k < 0.8:
                tr.append((p, f"managed {size} {t}"))
            elif k < 0.9:
                tr.append((p, f"pitch {rng.randrange(1, 60000)} {rng.randrange(1, 500)} {rng.choice([4, 8, 16])} {t}"))
            else:
                tr.append((p, f"hostalloc {min(size, 1 << 20)} {t}"))
                continue
            live[p].append(t)
            where[p][t] = cur[p]
        elif r < 0.55 and live[p]:
            t = live[p].pop(rng.randrange(len(live[p])))
            if where[p][t] != cur[p]:
                # freed with the owning device current: the reference un-bills the CURRENT device (remove_chunk@0x413dd), which
                # corrupts both columns when it is another one; the shim un-bills the device the allocation was billed to
                # (tests/test_quota_parity.py::test_free_with_another_device_current)
                cur[p] = where[p][t]
                tr.append((p, f"ctxset {cur[p] + 1}"))
            tr.append((p, f"free {t}"))
        elif r < 0.70:
            q = rng.choice(["info", "total 0", "nvmlmem", "@slots", "ctxdev", "ctxget", "devcount", "ptr", "host", "async", "vmm", "ctx", "sig"])
            t = tag[p]
            if q == "ptr":
                q = rng.choice(["ptrattrs", "ptrdev"]) + f" {rng.choice(live[p]) if live[p] else 0}"
            elif q == "host":       # pinned-host calls: checked after the fact with addon 0
                tag[p] += 1
                tr.append((p, rng.choice([f"hostalloc2 {rng.randrange(1, 1 << 20)} {t}", f"hostreg {rng.randrange(1, 1 << 20)}"])))
                continue
            elif q == "async":      # stream-ordered allocator: outside the quota in both libraries
                tag[p] += 1
                tr += [(p, f"allocasync {rng.randrange(1, 64) * M} {t}"), (p, rng.choice([f"freeasync {t}", f"free {t}"]))]
                continue
            elif q == "vmm":        # VMM API: outside the quota in both libraries
                tag[p] += 2
                tr += [(p, f"vcreate {rng.randrange(1, 32) * 2 * M} {t}"), (p, f"vmap {t} {t + 1}"), (p, f"vrelease {t}"), (p, f"vunmap {t + 1}")]
                continue
            elif q == "sig":        # suspended for a moment: only calls that do not park at the gate in between
                tr += [(p, "@usr2"), (p, "info"), (p, "@slots"), (p, "total 0"), (p, "nvmlmem"), (p, "@usr1"), (p, "@slots")]
                continue
            elif q == "ctx":        # a context of the application's own on the current device, entered and left again
                tag[p] += 1
                tr += [(p, f"ctxcreate {cur[p]} {t % 20 + 40}"), (p, "ctxdev"), (p, f"ctxdestroy {t % 20 + 40}"), (p, f"ctxset {cur[p] + 1}")]
                continue
            tr.append((p, q))
        elif r < 0.76:
            tr.append((p, rng.choice(["release 0", "retain 0 1", "retain 0 1"])))
        elif r < 0.84:
            q = rng.choice(sorted(alive))
            tr.append((p, f"libcall $pid{q} 0 {rng.randrange(0, 900) * M} set_gpu_device_memory_monitor"))
        elif r < 0.90:
            tr.append((p, "libcall 0 0 0 " + rng.choice(["get_gpu_memory_usage", "get_gpu_memory_monitor", "get_current_device_memory_limit",
                                                          "get_recent_kernel"])))
        elif r < 0.94:
            tr.append((p, f"forkalloc {rng.randrange(1, 100) * M} {rng.randrange(2)}"))
        elif r < 0.96 and len(alive) > 1 and p != 0:
            tr.append((p, "@kill"))
            alive.discard(p)
        elif r < 0.98:
            tr.append((p, "launch 8 128"))
        else:
            tr.append((p, f"libcall 0 {rng.randrange(500, 3000) * M} 0 set_current_device_memory_limit"))
    return tr


def differences(seed, extra, nprocs=4, ndev=1, nlate=0):
    """-> list of human-readable differences between the two libraries on this seed (empty = identical)."""
    env = {"CUDA_DEVICE_MEMORY_LIMIT_0": "2g", "NVIDIA_DEVICE_MAP": "0:" + H.UUID0, "MOCK_CTX_MIB": "200", "ACTIVE_OOM_KILLER": "0", **extra}
    if ndev == 2:
        env.update({"MOCK_GPU_COUNT": "2", "CUDA_DEVICE_MEMORY_LIMIT_1": "1g", "NVIDIA_DEVICE_MAP": f"0:{H.UUID0} 1:{H.UUID0[:-1]}2"})
    tr = gen(seed, nprocs, ndev=ndev, nlate=nlate)
    a = H.run_trace(H.REFERENCE_LIB, tr, env, nprocs)
    b = H.run_trace(H.NEW_LIB, tr, env, nprocs)
    out = [f"op {i} {tr[i]}: reference {a['replies'][i]!r}, new {b['replies'][i]!r}" for i in range(len(tr)) if a["replies"][i] != b["replies"][i]]
    if a["oom"] != b["oom"]:
        out.append("OOM log lines differ")
    if a["calls"] != b["calls"]:
        out.append("driver call logs differ")
    ra, rb = H.mask_region(a["region"]), H.mask_region(b["region"])
    if ndev == 2:  # under the mock the reference writes device 0's UUID into every uuids[] line: left out (gen_quota_golden.py)
        ra, rb = ra[:0x38] + ra[0x638:], rb[:0x38] + rb[0x638:]
    if ra != rb:
        out.append("final cache files differ")
    return out


def main():
    first = int(sys.argv[1]) if len(sys.argv) > 1 else 0
    last = int(sys.argv[2]) if len(sys.argv) > 2 else first + 10
    H.build()
    bad = 0
    for seed in range(first, last):
        for ndev, nprocs, nlate in ((1, 4, 0), (2, 4, 0), (1, 6, 2)):
            for extra in ENVS:
                d = differences(seed, extra, nprocs=nprocs, ndev=ndev, nlate=nlate)
                if d:
                    bad += 1
                    print(seed, (ndev, nprocs, nlate), extra, d[:3])
    print(f"seeds {first}..{last - 1} x {len(ENVS)} environments x (1 device, 2 devices, 2 late joiners): {bad} with differences")
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
