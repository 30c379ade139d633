"""gpuallocator's bestEffortPolicy behind the C-ABI (vgpu_best_effort), SURVEY §8(f) row 4: checked
against hand-derived answers on small topologies and, on random score matrices, against the
independent restatement in oracle/plugin_model.py (same enumeration order, same tie-breaking).
Reference: vendor/github.com/NVIDIA/go-gpuallocator/gpuallocator/besteffort_policy.go:34-380."""
import ctypes as C
import random

import pytest

from oracle import plugin_model as M


def call(vgpu, available, required, size, score):
    n = len(score)
    flat = (C.c_int * (n * n))(*[score[i][j] for i in range(n) for j in range(n)])
    av = (C.c_int * max(1, len(available)))(*available)
    rq = (C.c_int * max(1, len(required)))(*required)
    out = (C.c_int * 32)()
    k = vgpu.capi.check(vgpu.capi.lib.vgpu_best_effort(av, len(available), rq, len(required), size, flat, n, out, 32))
    return [out[i] for i in range(k)]


def sym(n, pairs, default=0):
    s = [[default if i != j else 0 for j in range(n)] for i in range(n)]
    for (a, b), v in pairs.items():
        s[a][b] = s[b][a] = v
    return s


def test_edge_cases_of_the_go_code(vgpu):
    s = sym(4, {})
    assert call(vgpu, [0, 1, 2, 3], [], 0, s) == []            # size <= 0
    assert call(vgpu, [0, 1], [], 3, s) == []                   # fewer available than size
    assert call(vgpu, [0, 1, 2], [0, 1, 2], 2, s) == []         # more required than size
    assert call(vgpu, [0, 1, 2, 3], [], 1, s) == [0]            # all scores 0: first candidate wins
    assert call(vgpu, [2, 0, 3], [], 2, s) == [2, 0]            # order of `available` is the enumeration order
    assert call(vgpu, [0, 1, 2], [2], 2, s) == [0, 2]           # the required device must be in a padding-free set


def test_nvlink_pairs_are_kept_together(vgpu):
    # 4 GPUs: (0,1) and (2,3) share 2 NVLinks (200), everything else is same-CPU PCIe (20)
    s = sym(4, {(0, 1): 200, (2, 3): 200}, default=20)
    assert call(vgpu, [0, 1, 2, 3], [], 2, s) == [0, 1]
    assert call(vgpu, [0, 1, 2, 3], [3], 2, s) == [2, 3]
    assert call(vgpu, [1, 2, 3], [], 2, s) == [2, 3]            # GPU 0 busy: take the intact pair, leave 1 alone
    assert call(vgpu, [0, 2, 3], [0], 2, s) == [0, 2]           # 0 must be in: its partner is gone, any will do
    # the policy maximises the WHOLE division, not the single best set: here taking (0,1) would strand
    # 2 and 3 on PCIe, while [0,2]+[1,3] keeps two NVLink pairs
    s2 = sym(4, {(0, 1): 300, (0, 2): 200, (1, 3): 200}, default=0)
    assert call(vgpu, [0, 1, 2, 3], [], 2, s2) == [0, 2]


def test_uniform_nvswitch_box(vgpu):
    """8 B200s behind NVSwitch: every pair scores the same, so the first sets in enumeration order win."""
    s = sym(8, {}, default=1800)
    assert call(vgpu, list(range(8)), [], 4, s) == [0, 1, 2, 3]
    assert call(vgpu, list(range(8)), [5], 4, s) == [4, 5, 6, 7]   # first division with a full set around GPU 5
    assert call(vgpu, [1, 3, 4, 6, 7], [], 3, s) == [1, 3, 4]   # 5 available, size 3: padded to 6


@pytest.mark.parametrize("seed", range(40))
def test_matches_independent_restatement_on_random_topologies(vgpu, seed):
    rng = random.Random(seed)
    n = rng.randrange(2, 9)
    s = sym(n, {(a, b): rng.choice([0, 10, 20, 50, 100, 200, 400]) for a in range(n) for b in range(a + 1, n)})
    available = rng.sample(range(n), rng.randrange(1, n + 1))
    size = rng.randrange(1, min(len(available), 4) + 2)
    required = rng.sample(available, rng.randrange(0, min(size, len(available)) + 1)) if rng.random() < 0.6 else []
    want = M.best_effort(available, required, size, s)
    assert call(vgpu, available, required, size, s) == [(-1 if d is None else d) for d in want]


def test_preferred_allocation_with_the_real_policy(vgpu):
    """GetPreferredAllocation (server.go:271-326) on a 4-GPU node with NVLink pairs: the kubelet gets the
    first available vGPU of each GPU of the best physical set."""
    P = vgpu.plugin
    uu = [f"GPU-{i:08d}-0000-0000-0000-000000000000" for i in range(4)]
    score = {a: {b: (0 if a == b else 20) for b in uu} for a in uu}
    for a, b in ((0, 1), (2, 3)):
        score[uu[a]][uu[b]] = score[uu[b]][uu[a]] = 200
    p = P.NvidiaDevicePlugin("nvidia.com/gpu", P.ResourceManager([P.Device(u, Index=str(i), memory_mib=184320) for i, u in enumerate(uu)]),
                             "NVIDIA_VISIBLE_DEVICES", P.bestEffortPolicy(score), P.DevicePluginPath + "nvidia-gpu.sock", P.Flags(deviceSplitCount=2))
    p.initialize()
    avail = [f"{uu[1]}-1", f"{uu[2]}-0", f"{uu[3]}-1", f"{uu[2]}-1"]        # GPU 0 fully taken
    assert p.GetPreferredAllocation([{"AvailableDeviceIDs": avail, "AllocationSize": 2}]) == [[f"{uu[2]}-0", f"{uu[3]}-1"]]
    assert p.GetPreferredAllocation([{"AvailableDeviceIDs": avail, "MustIncludeDeviceIDs": [f"{uu[1]}-1"], "AllocationSize": 2}]) == [[f"{uu[1]}-1", f"{uu[2]}-0"]]


class _FakeNvml:
    """Four GPUs: 0-1 and 2-3 joined by two NVLinks each, 0/1 under one PCIe switch, 2/3 under another CPU."""
    class _Pci:
        def __init__(self, bus):
            self.busId = bus.encode()

    def __init__(self):
        self.bus = {f"GPU-{i}": f"00000000:{0x10 + i:02X}:00.0" for i in range(4)}
        self.links = {"GPU-0": ["GPU-1", "GPU-1"], "GPU-1": ["GPU-0", "GPU-0"], "GPU-2": ["GPU-3", "GPU-3"], "GPU-3": ["GPU-2", "GPU-2"]}

    def nvmlDeviceGetHandleByUUID(self, u):
        u = u.decode()
        if u not in self.bus:
            raise RuntimeError("not found")
        return u

    def nvmlDeviceGetPciInfo(self, h):
        return self._Pci(self.bus[h])

    def nvmlDeviceGetNvLinkRemotePciInfo(self, h, link):
        if link >= len(self.links[h]):
            raise RuntimeError("NVML_ERROR_INVALID_ARGUMENT")
        return self._Pci(self.bus[self.links[h][link]].lower())      # case differs from GetPciInfo on purpose

    def nvmlDeviceGetTopologyCommonAncestor(self, a, b):
        same_side = (a in ("GPU-0", "GPU-1")) == (b in ("GPU-0", "GPU-1"))
        return 10 if same_side else 50                               # SINGLE switch / SYSTEM


def test_pair_scores_from_nvml_follow_calculateGPUPairScore(vgpu):
    P = vgpu.plugin
    uuids = [f"GPU-{i}" for i in range(4)]
    s = P.pairScoresFromNvml(uuids, nvml=_FakeNvml())
    assert s["GPU-0"]["GPU-1"] == s["GPU-1"]["GPU-0"] == 50 + 200    # single switch + two NVLinks
    assert s["GPU-2"]["GPU-3"] == 250 and s["GPU-0"]["GPU-2"] == s["GPU-3"]["GPU-1"] == 10
    assert "GPU-0" not in s["GPU-0"]
    alloc = P.bestEffortPolicy(s)
    assert alloc(uuids, [], 2) == ["GPU-0", "GPU-1"]
    assert alloc(uuids, ["GPU-3"], 2) == ["GPU-2", "GPU-3"]
    assert alloc(["GPU-0", "GPU-2", "GPU-3"], [], 2) == ["GPU-2", "GPU-3"]
    with pytest.raises(P.PluginError):
        P.pairScoresFromNvml(["GPU-9"], nvml=_FakeNvml())


@pytest.mark.gpu
def test_pair_scores_on_the_real_topology(vgpu):
    """Real NVML: symmetric scores, PCIe points in the reference's table, preferred allocation picks the best pair."""
    import pynvml
    pynvml.nvmlInit()
    n = pynvml.nvmlDeviceGetCount()
    uuids = []
    for i in range(n):
        u = pynvml.nvmlDeviceGetUUID(pynvml.nvmlDeviceGetHandleByIndex(i))
        uuids.append(u.decode() if isinstance(u, bytes) else u)
    s = vgpu.plugin.pairScoresFromNvml(uuids)
    print("pair scores:", s)
    for a in uuids:
        for b in uuids:
            if a != b:
                assert s[a][b] == s[b][a] and s[a][b] % 10 == 0 and 0 <= s[a][b] <= 1260
    if n >= 2:
        best = max(s[a][b] for a in uuids for b in uuids if a != b)
        got = vgpu.plugin.bestEffortPolicy(s)(uuids, [], 2)
        assert len(got) == 2 and s[got[0]][got[1]] == best
