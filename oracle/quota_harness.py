"""Runs quota traces through a libvgpu (the reference binary or the new shim) over the mock
driver built by oracle/ref_harness.  TEST INFRASTRUCTURE ONLY.

A trace is a list of (proc, op) pairs; `op` is a line of trace_runner's little language.
Runners are driven in lock step (one reply per op), so multi-process interleavings are
deterministic and identical for both libraries."""
import os
import random
import re
import struct
import subprocess
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")
REFERENCE_LIB = "/root/reference/vgpu/libvgpu.so"
NEW_LIB = os.path.join(os.path.dirname(HERE), "4paradigm-k8s-device-plugin_b200", "libvgpu.so")
UUID0 = "GPU-00000000-0000-0000-0000-000000000001"

REGION_SIZE = 0xC4748
SLOT0, STRIDE, NSLOT = 0x738, 0x310, 1024
OOM_RE = re.compile(r"Device (\d+) OOM (\d+) / (\d+)")


def build():
    subprocess.check_call(["make", "-s", "-C", os.path.join(HERE, "ref_harness")])


def have_reference():
    return os.path.exists(REFERENCE_LIB)


def _env(lib, cache, extra, mock_log):
    env = {
        "PATH": os.environ.get("PATH", "/usr/bin:/bin"),
        "LD_LIBRARY_PATH": REF_DIR,
        "LD_PRELOAD": f"{os.path.join(REF_DIR, 'libdlsymshim.so')} {lib}",
        "CUDA_DEVICE_MEMORY_SHARED_CACHE": cache,
        "LIBCUDA_LOG_LEVEL": "0",
        "MOCK_CUDA_LOG": mock_log,
    }
    env.update(extra)
    return env


def run_trace(lib, trace, env_extra, nprocs=1, keep_region=True, parallel=False, pre_cache=None):
    """-> dict(replies=[str...] in trace order, oom=[(dev,need,limit)...], calls=[str...], region=bytes)
    `lib` may be a list with one library per process (old and new library sharing one cache file), and so may
    `env_extra` (processes of one cache file that were started with different limits).
    parallel=True sends every op before reading any reply (processes really run concurrently)."""
    os.makedirs("/tmp/vgpulock", exist_ok=True)  # the reference spins forever without this dir
    try:
        os.unlink("/tmp/vgpulock/lock")
    except FileNotFoundError:
        pass
    with tempfile.TemporaryDirectory(prefix="vgpuq") as td:
        cache = os.path.join(td, "shr.cache")
        if pre_cache is not None:  # a cache file with history (or garbage) that the first process finds
            with open(cache, "wb") as f:
                f.write(pre_cache)
        mock_log = os.path.join(td, "mock.log")
        errs = [open(os.path.join(td, f"err{p}"), "w+") for p in range(nprocs)]
        procs = []
        replies = []
        try:
            def start(p):
                plib = lib[p] if isinstance(lib, (list, tuple)) else lib
                pr = subprocess.Popen([os.path.join(REF_DIR, "trace_runner")], stdin=subprocess.PIPE, stdout=subprocess.PIPE,
                                      stderr=errs[p], env=_env(plib, cache, env_extra[p] if isinstance(env_extra, (list, tuple)) else env_extra, mock_log),
                                      text=True, bufsize=1)
                first = pr.stdout.readline().strip()  # "init <rc>"
                if not first.startswith("init"):
                    raise RuntimeError(f"runner {p} did not start: {first!r}")
                return pr

            late = {p for p, op in trace if op == "@start"}  # these come up when the trace says so, the others now
            procs = [None if p in late else start(p) for p in range(nprocs)]
            if parallel:
                # replies are drained by one reader thread per process while the ops are being written: with the pipes
                # only, a long trace fills the runner's stdout pipe and both sides block
                import threading
                per_proc = [[] for _ in procs]
                want = [sum(1 for p, _ in trace if p == i) for i in range(nprocs)]

                def drain(i):
                    for _ in range(want[i]):
                        line = procs[i].stdout.readline()
                        if not line:
                            break
                        per_proc[i].append(line.strip())

                readers = [threading.Thread(target=drain, args=(i,), daemon=True) for i in range(nprocs)]
                for t in readers:
                    t.start()
                for p, op in trace:
                    procs[p].stdin.write(op + "\n")
                for pr in procs:
                    pr.stdin.flush()
                for t in readers:
                    t.join(timeout=300)
                nxt = [0] * nprocs
                for p, op in trace:
                    if nxt[p] >= len(per_proc[p]):
                        raise RuntimeError(f"runner {p} died on {op!r}")
                    replies.append(per_proc[p][nxt[p]])
                    nxt[p] += 1
                trace = []
            for p, op in trace:
                if op == "@slots":  # answered by the harness: the live process slots of the cache file
                    replies.append(live_slots(open(cache, "rb").read(), [pr.pid if pr else -1 for pr in procs]))
                    continue
                if op == "@start":  # a process that joins later: it claims a slot in a file that already has history
                    procs[p] = start(p)
                    replies.append("started")
                    continue
                if op in ("@usr1", "@usr2"):  # the monitor's suspend (SIGUSR2) / resume (SIGUSR1) protocol
                    import signal
                    import time
                    procs[p].send_signal(signal.SIGUSR2 if op == "@usr2" else signal.SIGUSR1)
                    time.sleep(0.2)  # handlers only write the status word; generous for a busy machine
                    replies.append("signalled")
                    continue
                if op.startswith("@send "):  # an op whose reply is not awaited (it may park at the suspend gate)
                    procs[p].stdin.write(op[6:] + "\n")
                    procs[p].stdin.flush()
                    replies.append("sent")
                    continue
                if op in ("@poll", "@pollready"):  # is the reply of the last @send there yet?  0.4 s of patience; @pollready
                    import select                  # (used where a reply is expected) waits up to 5 s, so a busy machine is not a failure
                    ready, _, _ = select.select([procs[p].stdout], [], [], 0.4 if op == "@poll" else 5.0)
                    replies.append("ready" if ready else "blocked")
                    continue
                if op == "@recv":
                    replies.append(procs[p].stdout.readline().strip())
                    continue
                if op == "@sleep":  # lets signals sent to OTHER processes be delivered before the file is looked at
                    import time
                    time.sleep(0.3)
                    replies.append("slept")
                    continue
                if op == "@kill":   # the process dies without running its exit handler (SIGKILL): its slot stays behind
                    procs[p].kill()
                    procs[p].wait(timeout=20)
                    replies.append("killed")
                    continue
                if "$pid" in op:  # "$pid<k>" = the pid of runner k (what an outside monitor would pass)
                    for k, pr in enumerate(procs):
                        if pr is not None:
                            op = op.replace(f"$pid{k}", str(pr.pid))
                procs[p].stdin.write(op + "\n")
                procs[p].stdin.flush()
                line = procs[p].stdout.readline()
                if not line:
                    raise RuntimeError(f"runner {p} died on {op!r}")
                replies.append(line.strip())
            for pr in procs:  # deterministic exit order
                if pr is None or pr.poll() is not None:
                    continue
                pr.stdin.write("quit\n")
                pr.stdin.flush()
                pr.wait(timeout=20)
        finally:
            for pr in procs:
                if pr is not None and pr.poll() is None:
                    pr.kill()
        oom = []
        for f in errs:
            f.seek(0)
            oom += [tuple(int(x) for x in m.groups()) for m in OOM_RE.finditer(f.read())]
            f.close()
        calls = []
        if os.path.exists(mock_log):
            for line in open(mock_log):
                line = line.strip()
                # the reference probes NVML/primary-context at init to find its host pid; not part of the quota path
                if line.startswith(("cuInit", "cuDevicePrimaryCtxRetain")):
                    continue
                calls.append(line)
        region = open(cache, "rb").read() if keep_region and os.path.exists(cache) else b""
    return {"replies": replies, "oom": sorted(oom), "calls": calls, "region": region}


def live_slots(b: bytes, pids) -> str:
    """One token per live slot, in slot order: runner#:hostpid-is-its-pid:status then per device with any usage
    d<dev>=context/module/data/total.  Pids become runner numbers so that runs compare."""
    n = struct.unpack_from("<i", b, 0xC4738)[0]
    out = []
    for s in range(n):
        o = SLOT0 + s * STRIDE
        pid, hostpid = struct.unpack_from("<ii", b, o)
        who = pids.index(pid) if pid in pids else -1
        tok = f"p{who}:{int(hostpid == pid)}:{struct.unpack_from('<i', b, o + 0x308)[0]}"
        for d in range(16):
            u = struct.unpack_from("<5Q", b, o + 8 + d * 0x28)
            if any(u):
                tok += f" d{d}={u[0]}/{u[1]}/{u[2]}/{u[4]}" + (f"+off{u[3]}" if u[3] else "")  # offset: MEMORY_OVERRIDE only
        mon = struct.unpack_from("<16Q", b, o + 0x288)
        if any(mon):
            tok += " mon=" + ",".join(str(m) for m in mon[:max(i for i, m in enumerate(mon) if m) + 1])
        out.append(tok)
    return f"n={n} " + " | ".join(out)


def mask_region(b: bytes) -> bytes:
    """Zero what legitimately differs between runs: pids (slot.pid, slot.hostpid)."""
    if not b:
        return b
    a = bytearray(b[:REGION_SIZE])
    for s in range(-1, NSLOT):  # -1: where both libraries put their slot when a garbage file says proc_num = -1
        o = SLOT0 + s * STRIDE
        a[o:o + 8] = b"\0" * 8
    return bytes(a)


def region_summary(b: bytes) -> dict:
    if not b:
        return {}
    return {
        "size": len(b),
        "magic": struct.unpack_from("<i", b, 0)[0],
        "owner_pid": struct.unpack_from("<Q", b, 8)[0],
        "device_num": struct.unpack_from("<Q", b, 0x30)[0],
        "uuid0": b[0x38:0x38 + 96].split(b"\0")[0].decode(),
        "limit": list(struct.unpack_from("<16Q", b, 0x638)),
        "sm_limit": list(struct.unpack_from("<16Q", b, 0x6B8)),
        "proc_num": struct.unpack_from("<i", b, 0xC4738)[0],
        "utilization_switch": struct.unpack_from("<i", b, 0xC473C)[0],
        "recent_kernel": struct.unpack_from("<i", b, 0xC4740)[0],
        "priority": struct.unpack_from("<i", b, 0xC4744)[0],
        "masked_crc": __import__("zlib").crc32(mask_region(b)),
    }


# ------------------------------------------------------------------------------ trace generators
def golden_trace():
    """alloc 1, 3, 500, 400, 200 MiB; free the 500 MiB; info after each."""
    MiB = 1 << 20
    ops = ["total 0", "info"]
    for tag, mib in enumerate([1, 3, 500, 400, 200], 1):
        ops += [f"alloc {mib * MiB} {tag}", "info"]
    ops += ["free 3", "info", "nvmlmem"]
    return [(0, o) for o in ops]


def appendix_a_trace():
    """SURVEY Appendix A op script."""
    ops = ["total 0", "info", f"alloc {100 << 20} 1", "info", "freeptr 0xdeadbeef000", "freeptr 0", "free 1", "free 1",
           "alloc 0 2", "info"]
    return [(0, o) for o in ops]


def ctx_trace():
    """Context billing on one device (reference: cuDevicePrimaryCtxRetain@0x294bd, Release_v2@0x29f08,
    cuCtxCreate_v2@0x2a4dd): first retain billed, second not, release forgets the flag but keeps the bytes,
    so the next retain is billed again; cuCtxCreate after a retain is a logged duplicate; destroy refunds nothing."""
    MiB = 1 << 20
    ops = ["info", "retain 0 1", "@slots", "info", "retain 0 2", "info", "ctxset 1", "ctxget", "ctxdev", f"alloc {MiB} 1", "@slots",
           "release 0", "info", "release 0", "info", "retain 0 3", "@slots", "info", "ctxcreate 0 4", "info", "ctxget", "ctxdev",
           "ctxdestroy 4", "info", "ctxget", "release 0", "ctxcreate 0 5", "info", "ctxcreate 0 6", "@slots", "total 0", "nvmlmem",
           f"alloc {200 * MiB} 2", f"alloc {100 * MiB} 3", "info", "free 1", "free 2", "@slots"]
    return [(0, o) for o in ops]


def ctx_create_first_trace():
    """A driver-API program whose first context comes from cuCtxCreate.  (No primary retain after it: the
    reference then hands back the created — here already destroyed — context as the primary one, a remap
    artefact that is not reproduced.)"""
    MiB = 1 << 20
    ops = ["ctxcreate 0 1", "@slots", "info", f"alloc {600 * MiB} 1", f"alloc {200 * MiB} 2", "info", "ctxdestroy 1", "info",
           "ctxcreate 0 2", "info", "ctxget", "ctxdev", "free 1", f"alloc {200 * MiB} 3", "info", "@slots"]
    return [(0, o) for o in ops]


def ctx_two_devices_trace():
    """Two devices: each billed on its own first retain; the current context decides whose quota an allocation hits."""
    MiB = 1 << 20
    ops = ["retain 1 1", "@slots", "info", "ctxset 1", "ctxdev", "info", f"alloc {MiB} 1", "info", "retain 0 2", "ctxset 2", "info",
           "ctxpush 1", "ctxdev", "info", "ctxpop", "ctxdev", "info", "total 0", "total 1", f"alloc {800 * MiB} 2", f"alloc {800 * MiB} 3",
           "ctxset 1", f"alloc {800 * MiB} 4", f"alloc {800 * MiB} 5", f"alloc {800 * MiB} 6", "@slots", "release 1", "retain 1 3", "info",
           "@slots"]
    return [(0, o) for o in ops]


def ctx_procs_trace(nprocs=4):
    """Several processes of one vGPU: every one pays for its context; the sum can pass the limit without any
    allocation (retain is never refused), after which cuMemGetInfo answers 1 and allocations -1."""
    MiB = 1 << 20
    out = []
    for p in range(nprocs):
        out += [(p, "retain 0 1"), (p, "info"), (p, f"alloc {50 * MiB} 1")]
    out.append((0, "@slots"))  # after every process has a context: the reference knows its host pid from cuInit on, the shim from the first context on
    for p in range(nprocs):
        out += [(p, "free 1"), (p, "info"), (p, "hostalloc 4096 2")]
    out.append((0, "@slots"))
    return out


def remap_shared_gpu_trace():
    """NVIDIA_DEVICE_MAP names one physical GPU twice (a pod holding two vGPUs of one split): two CUDA devices,
    the second with PCI function .1, its own limit and usage column and a context of its own.  Allocations stay
    inside the limits here: past them the reference stops enforcing in this mode (its oom_check@0x40567 compares the
    REAL device count with the map's and gives up) — tests/test_quota_parity.py pins that difference separately."""
    MiB = 1 << 20
    ops = ["devcount", "devget 0", "devget 1", "total 0", "total 1", "retain 0 1", "ctxset 1", "ctxdev", "info", f"alloc {MiB} 1", "info",
           "retain 1 2", "ctxset 2", "ctxget", "ctxdev", "info", f"alloc {4 * MiB} 2", "info", "@slots", "ctxset 1", "ctxdev", "info",
           f"alloc {500 * MiB} 3", "info", "ctxset 2", f"alloc {800 * MiB} 4", f"alloc {800 * MiB} 5", "info", "@slots",
           "retain 1 3", "release 1", "retain 1 4", "info", "free 4", "info", "ctxset 1", "free 3", "free 1", "info", "ctxset 2", "free 2", "free 5",
           "info", "@slots"]
    return [(0, o) for o in ops]


def remap_three_on_two_trace():
    """Map "0:A 1:B 2:A" on GPUs A, B: enumerated in physical order — A, A', B — whatever the order of the map."""
    MiB = 1 << 20
    ops = ["devcount", "devget 0", "devget 1", "devget 2", "total 0", "total 1", "total 2", "retain 2 1", "ctxset 1", "ctxdev", "info",
           f"alloc {MiB} 1", "retain 1 2", "ctxset 2", "ctxdev", "info", f"alloc {64 * MiB} 2", "info", "retain 0 3", "ctxset 3", "ctxdev",
           f"alloc {32 * MiB} 3", "info", "@slots", "ctxset 1", "free 1", "info", "ctxset 2", "free 2", "info", "@slots"]
    return [(0, o) for o in ops]


def remap_queries_trace(n):
    """Device-level queries only (no context): count, handle, UUID byte, bus id, total memory of every virtual device."""
    return [(0, "devcount")] + [(0, f"devget {i}") for i in range(n)] + [(0, f"total {i}") for i in range(n)]


def dead_process_trace():
    """A process is killed while it holds memory (no exit handler runs).  Its bytes stay in the cache file until some
    survivor runs into the limit: oom_check@0x40528 then looks for dead pids (rm_quitted_process@0x425b6), clears their
    slots and tries again — so the allocation that first looked like an OOM succeeds."""
    M = 1 << 20
    return [(0, f"alloc {600 * M} 1"), (1, f"alloc {300 * M} 1"), (2, "info"), (0, "@kill"), (2, "info"), (2, "@slots"),
            (1, f"alloc {100 * M} 2"), (1, "info"),                    # fits anyway: nobody looks for the dead yet
            (2, f"alloc {500 * M} 1"),                                  # 600 + 400 + 500 > 1 GiB: reap, retry, fits
            (2, "info"), (2, "@slots"), (1, f"alloc {200 * M} 3"), (1, "info"), (1, "@kill"), (2, f"alloc {500 * M} 2"), (2, "info"),
            (2, "@slots")]


def late_joiner_trace():
    """A process starts when the cache file already holds a dead process's slot: init_proc_slot_withlock@0x44820 claims a
    slot, then clear_proc_slot_nolock@0x4468b kicks the dead (the LAST slot is copied over the dead one — another
    compaction than the OOM path's).  Then the monitor protocol: SIGUSR2 / SIGUSR1 flip the status word of a slot."""
    M = 1 << 20
    return [(0, f"alloc {300 * M} 1"), (1, f"alloc {200 * M} 1"), (2, f"alloc {100 * M} 1"), (0, "@slots"), (0, "@kill"), (1, "@slots"),
            (3, "@start"), (3, "@slots"), (3, "info"), (3, f"alloc {50 * M} 1"), (3, "@slots"), (1, "@kill"), (4, "@start"), (4, "@slots"),
            (4, "info"), (2, "@usr2"), (2, "@slots"), (2, "@usr1"), (2, "@slots"), (2, "info")]


def fork_trace():
    """fork(): the child is a process of its own in the cache file (child_reinit_flag@0x449eb: it claims a slot at its
    first hooked call), its exit handler gives the slot back, `_exit` leaves it behind until some OOM sweep finds it."""
    M = 1 << 20
    ops = [f"alloc {100 * M} 1", "@slots", f"forkalloc {50 * M} 0", "@slots", "info", f"forkalloc {60 * M} 1", "@slots", "info",
           f"alloc {900 * M} 2", "@slots", "info", f"forkalloc {2000 * M} 0", "@slots", f"forkalloc {10 * M} 1", f"forkalloc {10 * M} 1", "@slots",
           "free 2", "info", f"forkalloc {900 * M} 0", "@slots"]
    return [(0, o) for o in ops]


def with_kills(trace, seed, nprocs, nkill):
    """`nkill` of the processes die by SIGKILL at random points of the trace; what they were going to do afterwards is dropped."""
    rng = random.Random(seed)
    victims = rng.sample(range(1, nprocs), min(nkill, nprocs - 1))   # process 0 survives to the end
    when = {v: rng.randrange(len(trace) // 4, 3 * len(trace) // 4) for v in victims}
    out, dead = [], set()
    for i, (p, op) in enumerate(trace):
        for v in victims:
            if when[v] == i and v not in dead:
                out.append((v, "@kill"))
                dead.add(v)
        if p not in dead:
            out.append((p, op))
    return out


def with_contexts(trace, nprocs):
    """Prefix: every process retains the primary context of device 0 (what any CUDA runtime program does)."""
    return [(p, "retain 0 1") for p in range(nprocs)] + list(trace)


def random_trace(seed, nops, nprocs=1, max_bytes=1 << 30, kinds=True):
    """60 % alloc / 40 % free-of-random-live, sizes log-uniform in [256 B, 1 GiB];
    sprinkled with info/total/managed/pitch/host ops."""
    rng = random.Random(seed)
    live = [[] for _ in range(nprocs)]
    next_tag = [1] * nprocs
    out = []
    import math
    lo, hi = math.log(256), math.log(max_bytes)
    for _ in range(nops):
        p = rng.randrange(nprocs)
        r = rng.random()
        if kinds and r < 0.08:
            q = rng.choice(["info", "info", "total 0", "nvmlmem", "ptrattrs"])
            out.append((p, q if q != "ptrattrs" else f"ptrattrs {rng.choice(live[p]) if live[p] else 0}"))
            continue
        if r < 0.63 or not live[p]:
            size = int(math.exp(rng.uniform(lo, hi)))
            tag = next_tag[p]
            next_tag[p] += 1
            k = rng.random() if kinds else 1.0
            if k < 0.08:
                out.append((p, f"managed {size} {tag}"))
            elif k < 0.12:
                w = rng.randrange(1, 1 << 16)
                h = rng.randrange(1, 1 << 10)
                out.append((p, f"pitch {w} {h} {rng.choice([4, 8, 16])} {tag}"))
            elif k < 0.15:
                hop = rng.choice(["hostalloc", "hostalloc2", "hostreg"])
                out.append((p, f"{hop} {min(size, 1 << 20)}" + ("" if hop == "hostreg" else f" {tag}")))
                continue  # host memory: not tracked as live device memory
            else:
                out.append((p, f"alloc {size} {tag}"))
            live[p].append(tag)  # may have failed (OOM): then the later free exercises the unknown-pointer path
        else:
            tag = live[p].pop(rng.randrange(len(live[p])))
            out.append((p, f"free {tag}"))
    return out
