"""The monitor's scheduling policy over crafted cache files: three containers share one GPU whose
HBM their quotas over-commit.  The files are written with the layout of region.h (the reference's shared_region_t);
`kill` is a stand-in that records the signal and flips the slot's status the way the shim's handlers do
(sig_swap_stub / sig_restore_stub), so no GPU and no real process is involved."""
import os
import signal
import struct

from vgpu_b200 import monitor as M

GiB = 1 << 30
UUID = "GPU-00000000-0000-0000-0000-000000000001"
HBM = {UUID.lower(): 180 * GiB}


def write_region(path, usage, limit, pid, hostpid, priority=1, recent_kernel=0, status=1, uuid=UUID):
    b = bytearray(M.REGION_SIZE + 1)
    struct.pack_into("<i", b, 0, M.MAGIC)
    struct.pack_into("<Q", b, 0x30, 1)
    b[0x38:0x38 + len(uuid)] = uuid.encode()
    struct.pack_into("<Q", b, 0x638, limit)
    struct.pack_into("<Q", b, 0x6B8, 100)
    struct.pack_into("<ii", b, M.SLOT0, pid, hostpid)
    struct.pack_into("<5Q", b, M.SLOT0 + 8, 0, 0, usage, 0, usage)
    struct.pack_into("<i", b, M.SLOT0 + 0x308, status)
    struct.pack_into("<i", b, 0xC4738, 1)       # proc_num
    struct.pack_into("<i", b, 0xC473C, 1)
    struct.pack_into("<i", b, 0xC4740, recent_kernel)
    struct.pack_into("<i", b, 0xC4744, priority)
    os.makedirs(os.path.dirname(path), exist_ok=True)
    open(path, "wb").write(b)


class Node:
    """A fake node: containers = directories under the shared root, processes = entries of `by_hostpid`."""

    def __init__(self, root):
        self.root, self.by_hostpid, self.signals = str(root), {}, []

    def add(self, name, usage_gib, limit_gib, pid, hostpid, **kw):
        path = os.path.join(self.root, name, "x.cache")
        write_region(path, usage_gib * GiB, limit_gib * GiB, pid, hostpid, **kw)
        if hostpid > 0:
            self.by_hostpid[hostpid] = path
        return path

    def kill(self, pid, sig):
        self.signals.append((pid, sig))
        path = self.by_hostpid.get(pid)
        assert path is not None, f"signal to pid {pid}, which is not a vGPU process of this node"
        with open(path, "r+b") as f:   # what the shim's handler does: status 2 on SIGUSR2, 1 on SIGUSR1
            f.seek(M.SLOT0 + 0x308)
            f.write(struct.pack("<i", 2 if sig == signal.SIGUSR2 else 1))

    def set(self, name, offset, value):
        M._poke_i32(os.path.join(self.root, name, "x.cache"), offset, value)

    def exit(self, name):
        self.set(name, M.SLOT0, 0)    # pid 0: the exit handler emptied the slot


def sched(node, **kw):
    return M.Scheduler(node.root, hbm=dict(HBM), headroom=4 * GiB, kill=node.kill, **kw)


def test_nothing_happens_while_the_tenants_fit(tmp_path):
    n = Node(tmp_path)
    n.add("a", 60, 100, 11, 1011)
    n.add("b", 60, 100, 12, 1012)
    s = sched(n)
    assert s.tick() == [] and s.tick() == [] and n.signals == []


def test_overcommitted_gpu_suspends_the_least_important_then_the_longest_idle(tmp_path):
    """60 + 70 + 80 GiB of usage on a 180 GiB GPU (176 usable): one tenant has to go.  Priority decides first (larger
    CUDA_TASK_PRIORITY yields), idleness second; the signal goes to the HOST pid (the container pids 11..13 name other
    processes on the host), after the launch brake (recent_kernel = -1) is set."""
    n = Node(tmp_path)
    pa = n.add("a", 60, 100, 11, 1011, priority=0, recent_kernel=2)
    pb = n.add("b", 70, 100, 12, 1012, priority=1, recent_kernel=2)     # active
    pc = n.add("c", 80, 100, 13, 1013, priority=1, recent_kernel=0)     # same importance as b, idle
    s = sched(n)
    assert s.tick() == [("suspend", "c")]
    assert n.signals == [(1013, signal.SIGUSR2)]
    assert M.read_region(pc)["recent_kernel"] == -1 and M.read_region(pc)["procs"][0]["status"] == 2
    assert M.read_region(pa)["recent_kernel"] == 1 and M.read_region(pb)["recent_kernel"] == 1   # activity counted down
    assert s.tick() == []                                               # 130 GiB running: fits, c does not
    # b finishes: its container goes away, and c comes back
    n.exit("b")
    assert s.tick() == [("resume", "c")]
    assert n.signals[-1] == (1013, signal.SIGUSR1) and M.read_region(pc)["recent_kernel"] == 0
    assert M.read_region(pc)["procs"][0]["status"] == 1


def test_priority_beats_idleness_and_a_lone_tenant_is_never_suspended(tmp_path):
    n = Node(tmp_path)
    n.add("hi", 100, 120, 11, 1011, priority=0, recent_kernel=0)        # important but idle
    n.add("lo", 100, 120, 12, 1012, priority=3, recent_kernel=2)        # busy but unimportant
    s = sched(n)
    assert s.tick() == [("suspend", "lo")]
    n2 = Node(tmp_path / "solo")
    n2.add("big", 250, 300, 21, 2021)                                   # alone larger than the GPU: it pages itself (shim)
    assert sched(n2).tick() == []


def test_tenants_of_equal_importance_take_turns(tmp_path):
    """Two 100 GiB tenants on 176 GiB: one runs, the other waits; after `quantum` ticks they swap (suspend first, resume
    once the HBM is free on the next tick), and again a quantum later."""
    n = Node(tmp_path)
    n.add("a", 100, 120, 11, 1011, recent_kernel=2)
    n.add("b", 100, 120, 12, 1012, recent_kernel=2)
    s = sched(n, quantum=3)
    first = s.tick()
    assert len(first) == 1 and first[0][0] == "suspend"
    out, run = first[0][1], ("a" if first[0][1] == "b" else "b")
    log = []
    for _ in range(9):
        n.set(run, 0xC4740, 2)                                          # the running one keeps launching
        acts = s.tick()
        log.append(acts)
        for a in acts:
            if a[0] == "resume":
                run = a[1]
    flat = [a for acts in log for a in acts]
    assert flat[:2] == [("suspend", ("a" if out == "b" else "b")), ("resume", out)]      # the swap after the first quantum
    assert flat[2:4] == [("suspend", out), ("resume", ("a" if out == "b" else "b"))]    # and back after the second
    assert all(len(acts) <= 1 for acts in log)                           # never both moving in one tick


def test_slots_without_a_host_pid_are_reported_not_signalled(tmp_path):
    n = Node(tmp_path)
    n.add("a", 100, 120, 11, 1011, priority=0)
    n.add("b", 100, 120, 4242, 0, priority=2)                           # no CUDA context yet: host pid unknown
    s = sched(n)
    acts = s.tick()
    assert ("suspend", "b") in acts and ("unreachable", "b", (4242,)) in acts
    assert n.signals == []                                              # pid 4242 on the host is somebody else's process


def test_publish_memory_writes_the_monitor_figure_by_host_pid(tmp_path):
    n = Node(tmp_path)
    p = n.add("a", 10, 100, 11, 1011)
    assert M.publish_memory(p, {1011: [7 * GiB], 999: [1]}) == 1
    raw = open(p, "rb").read()
    assert struct.unpack_from("<Q", raw, M.SLOT0 + 0x288)[0] == 7 * GiB


def test_nobody_is_resumed_into_hbm_that_is_not_free_yet(tmp_path):
    """A suspend takes time: the accounting says the victim's HBM is free the moment it is signalled, the GPU says it
    when the data has left.  With a measured-free provider the waiting tenant comes back only then."""
    n = Node(tmp_path)
    n.add("a", 100, 120, 11, 1011, priority=0)
    n.add("b", 100, 120, 12, 1012, priority=1)
    free = {"v": 0}
    s = M.Scheduler(n.root, hbm=dict(HBM), headroom=4 * GiB, kill=n.kill, hbm_free=lambda g: free["v"])
    assert s.tick() == [("suspend", "b")]
    n.exit("a")                                   # a leaves, but its memory (and b's spill) is still draining
    assert s.tick() == [] and s.tick() == []
    free["v"] = 150 * GiB
    assert s.tick() == [("resume", "b")]


def test_measured_hbm_counts_a_self_paging_tenant_by_what_it_holds(tmp_path):
    """A tenant larger than its resident share pages itself (the shim): it has ALLOCATED 150 GiB but holds 40.  With the
    NVML figures the scheduler sees that it fits next to a 100 GiB tenant and leaves both alone; by allocation alone it
    would suspend one of them."""
    n = Node(tmp_path)
    n.add("pager", 150, 300, 11, 1011)
    n.add("plain", 100, 120, 12, 1012)
    by_alloc = sched(n)
    assert [a[0] for a in by_alloc.tick()] == ["suspend"]
    n2 = Node(tmp_path / "again")
    n2.add("pager", 150, 300, 21, 2021)
    n2.add("plain", 100, 120, 22, 2022)
    seen = {2021: {UUID.lower(): 40 * GiB}, 2022: {UUID.lower(): 100 * GiB}}
    s = M.Scheduler(n2.root, hbm=dict(HBM), headroom=4 * GiB, kill=n2.kill, measured=lambda: seen)
    assert s.tick() == [] and n2.signals == []
    seen[2021][UUID.lower()] = 90 * GiB            # the pager grows: 190 GiB held on a 176 GiB budget
    assert [a[0] for a in s.tick()] == ["suspend"]
