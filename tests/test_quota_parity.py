"""Quota / accounting parity of the new preload shim (libvgpu.so) with the reference.

Three layers, all on the mock driver of oracle/ref_harness (CPU only):
 1. golden replay — tests/golden/quota_*.json.gz were captured from the reference binary
    (oracle/gen_quota_golden.py); the new shim must reproduce every reply, every OOM log line,
    the real-driver call log and the final shared-cache file (pids masked);
 2. the Python restatement oracle/quota_model.py must reproduce the same goldens (pins it);
 3. where the reference binary is present (/root/reference, build container only), fresh
    random traces run through both libraries and are compared live.
The vectors of SURVEY.md §8 c.3 and Appendix A are asserted literally as well."""
import glob
import gzip
import json
import os
import struct

import pytest

from oracle import quota_harness as H
from oracle.quota_model import Model, limit_from_env

GOLD = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "quota_*.json.gz")))
MAP = {"NVIDIA_DEVICE_MAP": "0:" + H.UUID0}


@pytest.fixture(scope="module", autouse=True)
def harness():
    H.build()
    assert os.path.exists(H.NEW_LIB), "libvgpu.so not built: run `make`"


def load(path):
    return json.loads(gzip.open(path).read())


@pytest.mark.parametrize("path", GOLD, ids=lambda p: os.path.basename(p)[6:-8])
def test_new_shim_reproduces_reference_golden(path):
    g = load(path)
    trace = [tuple(x) for x in g["trace"]]
    r = H.run_trace(H.NEW_LIB, trace, g["env"], g["nprocs"])
    bad = [(i, trace[i], g["replies"][i], r["replies"][i]) for i in range(len(trace)) if g["replies"][i] != r["replies"][i]]
    assert not bad, bad[:5]
    assert [list(x) for x in r["oom"]] == g["oom"], "OOM log lines (Device d OOM need / limit) differ"
    assert r["calls"] == g["calls"], "real-driver call trace differs"
    got = H.region_summary(r["region"])
    assert {k: got[k] for k in g["region"]} == g["region"], "final shared-cache file differs"


@pytest.mark.parametrize("path", GOLD, ids=lambda p: os.path.basename(p)[6:-8])
def test_python_model_reproduces_reference_golden(path):
    g = load(path)
    if g["name"].startswith(("remap_", "dead_", "late_", "fork_")):
        pytest.skip("the Python model restates the quota arithmetic, not the virtual-device table / the dead-process sweep")
    m = Model(g["env"], int(g["env"].get("MOCK_TOTAL_MIB", "16384")) << 20, g["nprocs"])
    for i, (p, op) in enumerate(g["trace"]):
        got = m.op(p, op)
        assert got == g["replies"][i], (i, op, got, g["replies"][i])
    assert m.calls == g["calls"]


def test_survey_c3_vector_literally():
    """SURVEY §8 c.3, observed from the reference binary: limit 1024m."""
    env = {"CUDA_DEVICE_MEMORY_LIMIT_0": "1024m", "CUDA_DEVICE_SM_LIMIT": "10", **MAP}
    r = H.run_trace(H.NEW_LIB, H.golden_trace(), env)
    rep = r["replies"]
    assert rep[0] == "0 1073741824"                          # cuDeviceTotalMem_v2 -> the limit
    frees = [int(x.split()[1]) for x in rep if x.count(" ") == 2 and x.startswith("0 ")]
    assert frees == [1073741824, 1072693248, 1069547520, 545259520, 125829120, 125829120, 650117120]
    assert rep[10] == "-1"                                   # 5th alloc: (CUresult)-1, not 2
    assert r["oom"] == [(0, 1157627904, 1073741824)]
    assert rep[-1] == "0 1073741824 650117120 423624704"     # nvmlDeviceGetMemoryInfo: total=limit, free, used
    b = r["region"]
    assert len(b) == 804681 and struct.unpack_from("<i", b, 0)[0] == 19920718
    assert struct.unpack_from("<16Q", b, 0x6B8) == (10,) * 16
    assert struct.unpack_from("<4i", b, 0xC4738) == (0, 1, 2, 1)  # proc_num, utilization_switch, recent_kernel, priority
    # the reference leaves its (zeroed) slot behind with status 1 after exit
    assert struct.unpack_from("<i", b, 0x738 + 0x308)[0] == 1
    # oversubscribe: same decisions, allocations > 2 MiB take the managed route
    r2 = H.run_trace(H.NEW_LIB, H.golden_trace(), {**env, "CUDA_OVERSUBSCRIBE": "true"})
    assert r2["replies"] == rep
    allocs = [c.split(" ->")[0] for c in r2["calls"] if c.startswith("cuMemAlloc")]
    assert allocs == ["cuMemAlloc_v2 size=1048576 flags=0", "cuMemAllocManaged size=3145728 flags=1",
                      "cuMemAllocManaged size=524288000 flags=1", "cuMemAllocManaged size=419430400 flags=1"]


@pytest.mark.parametrize("env,total,free2,alloc_rc,free_p", [
    ({}, 0, 17179869184 - 104857600, "0", "0"),                                     # quirk a18: TotalMem -> 0 when unlimited
    ({"CUDA_DEVICE_MEMORY_LIMIT": "1g"}, 1073741824, 968884224, "0", "0"),
    ({"CUDA_DEVICE_MEMORY_LIMIT_0": "1048576k", "CUDA_DEVICE_MEMORY_LIMIT": "2g"}, 1073741824, 968884224, "0", "0"),
    ({"CUDA_DEVICE_MEMORY_LIMIT_0": "0x40000000"}, 1073741824, 968884224, "0", "0"),  # strtoul base 0
    ({"CUDA_DEVICE_MEMORY_LIMIT_0": "50m"}, 52428800, 52428800, "-1", "0"),
])
def test_survey_appendix_a_rows(env, total, free2, alloc_rc, free_p):
    r = H.run_trace(H.NEW_LIB, H.appendix_a_trace(), env)["replies"]
    assert r[0] == f"0 {total}"
    assert r[2] == alloc_rc
    assert int(r[3].split()[1]) == free2
    assert r[4] == "-1" and r[5] == "0"            # unknown pointer -> -1, NULL -> 0
    assert r[6] == free_p                           # free(p); after a failed alloc the runner holds 0 -> free(0) -> 0
    assert r[7] == ("0" if alloc_rc == "-1" else "-1")  # free(p) again: no longer tracked -> -1
    assert r[8] == "0"                              # alloc 0 bytes succeeds


def test_context_is_billed_like_on_a_real_node():
    """SURVEY "usage includes context_size": with the process visible to NVML (MOCK_CTX_MIB) the first retain
    bills the context, a second does not, release keeps the bytes but forgets the flag, and the host pid is
    recorded in the slot.  Numbers as captured from the reference (tests/golden/quota_ctx_one_device)."""
    env = {"CUDA_DEVICE_MEMORY_LIMIT_0": "1g", "MOCK_CTX_MIB": "300", **MAP}
    ops = ["info", "retain 0 1", "info", "@slots", "retain 0 2", "info", "release 0", "info", "retain 0 3", "info", "@slots"]
    r = H.run_trace(H.NEW_LIB, [(0, o) for o in ops], env)["replies"]
    GiB, C = 1 << 30, 300 << 20
    assert r[0] == f"0 {GiB} {GiB}"
    assert r[2] == f"0 {GiB - C} {GiB}"
    assert r[3] == f"n=1 p0:1:1 d0={C}/0/0/{C}"          # hostpid == pid under the mock, kind 0 = context
    assert r[5] == r[7] == f"0 {GiB - C} {GiB}"
    assert r[9] == f"0 {GiB - 2 * C} {GiB}"
    assert r[10] == f"n=1 p0:1:1 d0={2 * C}/0/0/{2 * C}"


def test_without_nvml_process_listing_nothing_is_billed():
    """The reference's "SET_TASK_PID FAILED" case: NVML shows no process, context_size stays 0."""
    r = H.run_trace(H.NEW_LIB, [(0, "retain 0 1"), (0, "info"), (0, "@slots")], {"CUDA_DEVICE_MEMORY_LIMIT_0": "1g", **MAP})["replies"]
    assert r[1] == f"0 {1 << 30} {1 << 30}" and r[2] == "n=1 p0:0:1"


SHARED = {"NVIDIA_DEVICE_MAP": f"0:{H.UUID0} 1:{H.UUID0}", "CUDA_DEVICE_MEMORY_LIMIT_0": "1g", "CUDA_DEVICE_MEMORY_LIMIT_1": "2g",
          "MOCK_CTX_MIB": "300"}
SHARED_OOM_OPS = ["retain 0 1", "retain 1 2", "ctxset 1", f"alloc {800 << 20} 1", "info", "ctxset 2", f"alloc {800 << 20} 2",
                  f"alloc {800 << 20} 3", f"alloc {800 << 20} 4", "info", "@slots"]


def test_two_vgpus_on_one_gpu_each_keep_their_own_quota():
    """A GPU named twice in NVIDIA_DEVICE_MAP: every virtual device enforces ITS limit (context included)."""
    r = H.run_trace(H.NEW_LIB, [(0, o) for o in SHARED_OOM_OPS], SHARED)
    rep = r["replies"]
    C, M = 300 << 20, 800 << 20
    assert rep[3] == "-1" and rep[4] == f"0 {(1 << 30) - C} {1 << 30}"          # 300 + 800 MiB > 1 GiB on vGPU 0
    assert rep[6:9] == ["0", "0", "-1"] and rep[9] == f"0 {(2 << 30) - C - 2 * M} {2 << 30}"   # vGPU 1: two fit, the third does not
    assert rep[10] == f"n=1 p0:1:1 d0={C}/0/0/{C} d1={C}/0/{2 * M}/{C + 2 * M}"
    assert sorted(r["oom"]) == [(0, C + M, 1 << 30), (1, C + 3 * M, 2 << 30)]


@pytest.mark.skipif(not H.have_reference(), reason="reference binary only exists in the build container")
def test_reference_stops_enforcing_when_a_gpu_is_shared():
    """Evidence for the one deliberate difference in this mode: the reference lets every one of these allocations
    through (oom_check@0x40567 compares the real device count, 1, with the map's, 2, and returns 'no OOM'), after
    which its own cuMemGetInfo answers 1 (limit < usage)."""
    rep = H.run_trace(H.REFERENCE_LIB, [(0, o) for o in SHARED_OOM_OPS], SHARED)["replies"]
    assert rep[3] == "0" and rep[6:9] == ["0", "0", "0"] and rep[4] == rep[9] == "1 0 0"


def test_virtual_device_ordinals_out_of_range_are_invalid_device():
    """The reference dereferences past its table (segfault on `devget 1` with a one-entry map); here: 101."""
    env = {"MOCK_GPU_COUNT": "2", "NVIDIA_DEVICE_MAP": f"0:{H.UUID0[:-1]}2", "CUDA_DEVICE_MEMORY_LIMIT_0": "1g"}
    rep = H.run_trace(H.NEW_LIB, [(0, "devcount"), (0, "devget 1"), (0, "retain 1 1"), (0, "devget 0"), (0, "retain 0 1"), (0, "ctxset 1"), (0, "ctxdev")], env)["replies"]
    assert rep[0] == "0 1" and rep[1].startswith("101 ") and rep[2].startswith("101 ")
    assert rep[3] == "0 0 0 2 0 0000:11:00.0" and rep[4] == "0 1" and rep[6] == "0 0"   # virtual 0 is the SECOND GPU


def test_subset_map_enforces_the_quota_on_the_mapped_gpu():
    """Container sees two GPUs, the map gives it the second: its limit applies to allocations made there (the
    reference gives up on the quota whenever the real device count differs from the map's)."""
    env = {"MOCK_GPU_COUNT": "2", "NVIDIA_DEVICE_MAP": f"0:{H.UUID0[:-1]}2", "CUDA_DEVICE_MEMORY_LIMIT_0": "1g"}
    ops = ["retain 0 1", "ctxset 1", f"alloc {600 << 20} 1", f"alloc {600 << 20} 2", "info", "total 0", "nvmlmem"]
    r = H.run_trace(H.NEW_LIB, [(0, o) for o in ops], env)
    assert r["replies"][2:6] == ["0", "-1", f"0 {(1 << 30) - (600 << 20)} {1 << 30}", f"0 {1 << 30}"]
    # NVML index 0 is the mapped (second) GPU too, and shows column 0's quota
    assert r["replies"][6] == f"0 {1 << 30} {(1 << 30) - (600 << 20)} {600 << 20}"
    assert any(c.startswith("cuMemAlloc_v2 size=629145600") for c in r["calls"])


def test_pointer_device_ordinal_is_the_virtual_device_it_was_allocated_on():
    """Map "A B A" on two GPUs = virtual devices A, A', B.  cuPointerGetAttribute(DEVICE_ORDINAL) names the virtual device
    the application allocated on — consistent with cuCtxGetDevice.  (The reference passes the driver's physical answer
    through: 0, 0, 1 — which is not even a valid index for the third device's neighbours.)"""
    u2 = H.UUID0[:-1] + "2"
    env = {"MOCK_GPU_COUNT": "2", "NVIDIA_DEVICE_MAP": f"0:{H.UUID0} 1:{u2} 2:{H.UUID0}", "CUDA_DEVICE_MEMORY_LIMIT_0": "1g",
           "CUDA_DEVICE_MEMORY_LIMIT_1": "1g", "CUDA_DEVICE_MEMORY_LIMIT_2": "1g", "MOCK_CTX_MIB": "300"}
    ops = ["retain 0 1", "retain 1 2", "retain 2 3", "ctxset 1", f"alloc {1 << 20} 1", "ctxset 2", f"alloc {1 << 20} 2", "ctxset 3", f"alloc {1 << 20} 3",
           "ptrdev 1", "ptrdev 2", "ptrdev 3"]
    assert H.run_trace(H.NEW_LIB, [(0, o) for o in ops], env)["replies"][-3:] == ["0 0", "0 1", "0 2"]
    if H.have_reference():
        assert H.run_trace(H.REFERENCE_LIB, [(0, o) for o in ops], env)["replies"][-3:] == ["0 0", "0 0", "0 1"]


def test_free_with_another_device_current():
    """cuMemFree of a device-0 allocation while device 1's context is current: the bytes go back to device 0, where they
    were billed.  (The reference un-bills the CURRENT device — remove_chunk@0x413dd asks cuCtxGetDevice at free time — so
    device 0 stays charged for ever and device 1's counter wraps below zero; shown when the reference can run.)"""
    u2 = H.UUID0[:-1] + "2"
    env = {"MOCK_GPU_COUNT": "2", "NVIDIA_DEVICE_MAP": f"0:{H.UUID0} 1:{u2}", "CUDA_DEVICE_MEMORY_LIMIT_0": "1g", "CUDA_DEVICE_MEMORY_LIMIT_1": "1g",
           "MOCK_CTX_MIB": "200"}
    C2 = 200 << 20
    ops = ["retain 0 1", "retain 1 2", "ctxset 1", f"alloc {100 << 20} 1", "ctxset 2", "free 1", "@slots", "info", "ctxset 1", "info"]
    rep = H.run_trace(H.NEW_LIB, [(0, o) for o in ops], env)["replies"]
    assert rep[6] == f"n=1 p0:1:1 d0={C2}/0/0/{C2} d1={C2}/0/0/{C2}"
    assert rep[7] == rep[9] == f"0 {(1 << 30) - C2} {1 << 30}"
    if H.have_reference():
        ref = H.run_trace(H.REFERENCE_LIB, [(0, o) for o in ops], env)["replies"]
        wrapped = (1 << 64) - (100 << 20)
        assert ref[6] == f"n=1 p0:1:1 d0={C2}/0/{100 << 20}/{C2 + (100 << 20)} d1={C2}/0/{wrapped}/{C2 - (100 << 20)}"


def test_nvml_shows_a_shared_gpu_as_the_sum_of_its_vgpus():
    ops = ["retain 0 1", "retain 1 2", "ctxset 2", f"alloc {100 << 20} 1", "nvmlmem"]
    rep = H.run_trace(H.NEW_LIB, [(0, o) for o in ops], SHARED)["replies"]
    used = 2 * (300 << 20) + (100 << 20)
    assert rep[4] == f"0 {3 << 30} {(3 << 30) - used} {used}"


def test_runtime_api_clients_get_the_vgpus_of_a_shared_gpu_merged():
    """A process that fetches a driver export table before enumerating devices is a CUDA-runtime program; the runtime
    cannot initialise with two devices on one GPU, so it gets ONE device owning both vGPUs' columns: limit 1g + 2g,
    one context billed, allocations larger than either vGPU alone.  VGPU_SHARED_GPU=split keeps them apart."""
    ops = ["exporttable", "devcount", "devget 1", "total 0", "retain 0 1", "ctxset 1", "info", f"alloc {2000 << 20} 1", f"alloc {800 << 20} 2",
           "info", "nvmlmem", "@slots"]
    rep = H.run_trace(H.NEW_LIB, [(0, o) for o in ops], SHARED)["replies"]
    C, A = 300 << 20, 2000 << 20
    assert rep[1] == "0 1" and rep[2].startswith("101 ") and rep[3] == f"0 {3 << 30}"
    assert rep[6] == f"0 {(3 << 30) - C} {3 << 30}" and rep[7:9] == ["0", "-1"]
    assert rep[9] == f"0 {(3 << 30) - C - A} {3 << 30}" and rep[10] == f"0 {3 << 30} {(3 << 30) - C - A} {C + A}"
    assert rep[11] == f"n=1 p0:1:1 d0={C}/0/{A}/{C + A}"
    rep = H.run_trace(H.NEW_LIB, [(0, o) for o in ops], {**SHARED, "VGPU_SHARED_GPU": "split"})["replies"]
    assert rep[1] == "0 2" and rep[3] == f"0 {1 << 30}" and rep[7] == "-1"
    rep = H.run_trace(H.NEW_LIB, [(0, o) for o in ops[1:]], {**SHARED, "VGPU_SHARED_GPU": "merge"})["replies"]
    assert rep[0] == "0 1" and rep[2] == f"0 {3 << 30}"


def test_unlisted_gpus_are_hidden_from_the_driver_too():
    """Map names the second of two visible GPUs: CUDA_VISIBLE_DEVICES is set (before the real cuInit) so that the
    driver's private paths agree with the hooks; a value the user set is left alone."""
    u2 = H.UUID0[:-1] + "2"
    env = {"MOCK_GPU_COUNT": "2", "NVIDIA_DEVICE_MAP": f"0:{u2}", "CUDA_DEVICE_MEMORY_LIMIT_0": "1g"}
    assert H.run_trace(H.NEW_LIB, [(0, "cvd")], env)["replies"] == [f"0 {u2}"]
    # a preset list counts the container's OWN devices (the map's) and is rewritten as UUIDs; it ends at the first bad entry
    def cvd(user, extra=env):
        return H.run_trace(H.NEW_LIB, [(0, "cvd")], {**extra, "CUDA_VISIBLE_DEVICES": user})["replies"][0]
    assert cvd("0,1") == f"0 {u2}"            # what a 2-GPU host hands down: virtual 0 exists, virtual 1 does not
    assert cvd("0") == f"0 {u2}" and cvd(u2) == f"0 {u2}" and cvd(u2[:12]) == f"0 {u2}"
    assert cvd("1") == "0" and cvd(H.UUID0) == "0"              # names nothing of ours: no device, as the driver would answer
    assert cvd("MIG-GPU-abc/1/0") == "0 MIG-GPU-abc/1/0"      # not ours to interpret
    full = {"MOCK_GPU_COUNT": "2", "NVIDIA_DEVICE_MAP": f"0:{H.UUID0} 1:{u2}"}
    assert H.run_trace(H.NEW_LIB, [(0, "cvd")], full)["replies"] == ["0 -"]       # nothing to hide
    assert cvd("1,0", full) == f"0 {u2},{H.UUID0}"            # the user's order is kept


M = 1 << 20
UNMETERED_OPS = ["info", f"allocasync {600 * M} 1", "info", f"allocasync {600 * M} 2", "freeasync 1", "info", f"vcreate {512 * M} 1", "vmap 1 10", "info",
                 f"vcreate {600 * M} 2", "vrelease 1", "info", "vunmap 10", "info", f"alloc {900 * M} 3", f"allocasync {100 * M} 4", "free 4", "info",
                 "@slots"]


@pytest.mark.skipif(not H.have_reference(), reason="reference binary only exists in the build container")
def test_stream_ordered_and_vmm_allocations_are_outside_the_quota_as_in_the_reference():
    """cuMemAllocAsync / cuMemCreate are not metered by the reference: same replies, same driver
    calls, 1.2 GiB of async allocations inside a 1 GiB vGPU."""
    env = {"CUDA_DEVICE_MEMORY_LIMIT_0": "1g", **MAP}
    a = H.run_trace(H.REFERENCE_LIB, [(0, o) for o in UNMETERED_OPS], env)
    b = H.run_trace(H.NEW_LIB, [(0, o) for o in UNMETERED_OPS], env)
    assert a["replies"] == b["replies"] and a["calls"] == b["calls"]
    assert a["replies"][1] == a["replies"][3] == "0" and a["replies"][2] == f"0 {1 << 30} {1 << 30}"
    assert a["replies"][16] == "-1"                     # cuMemFree of a pointer the library never tracked


def test_strict_quota_also_meters_the_async_allocator_and_the_vmm_api():
    """VGPU_STRICT_QUOTA=1 (no reference counterpart): the same trace with the holes closed."""
    env = {"CUDA_DEVICE_MEMORY_LIMIT_0": "1g", "VGPU_STRICT_QUOTA": "1", **MAP}
    r = H.run_trace(H.NEW_LIB, [(0, o) for o in UNMETERED_OPS], env)
    G = 1 << 30
    assert r["replies"] == [
        f"0 {G} {G}", "0", f"0 {G - 600 * M} {G}",     # the first async allocation is billed
        "2", "0", f"0 {G} {G}",                         # the second does not fit (OUT_OF_MEMORY); freeing gives the bytes back
        "0", "0", f"0 {G - 512 * M} {G}",               # cuMemCreate is billed to the vGPU it is located on
        "2", "0", f"0 {G - 512 * M} {G}",               # a second handle does not fit; a released handle stays billed while mapped
        "0", f"0 {G} {G}",                              # unmapped: the physical memory is gone, so is the bill
        "0", "0", "0", f"0 {G - 900 * M} {G}",          # cuMemFree of a stream-ordered allocation is known now
        f"n=1 p0:0:1 d0=0/0/{900 * M}/{900 * M}"]
    assert r["oom"] == [(0, 1112 * M, G), (0, 1200 * M, G)]     # sorted by the harness: the handle, then the async allocation
    assert "cuMemCreate size=629145600 dev=0 -> rc=0" not in r["calls"]      # refused before the driver was asked


def test_two_processes_share_one_cache_file():
    """SURVEY c.3: A holds 500 MiB, B allocates 100 MiB and reads free = 1024-500-100 MiB;
    managed 900 MiB -> 2 in both; after both exit a third process sees everything free."""
    MiB = 1 << 20
    env = {"CUDA_DEVICE_MEMORY_LIMIT_0": "1024m", **MAP}
    trace = [(0, f"alloc {500 * MiB} 1"), (1, f"alloc {100 * MiB} 1"), (1, "info"), (0, "info"),
             (0, f"managed {900 * MiB} 2"), (1, f"managed {900 * MiB} 2")]
    r = H.run_trace(H.NEW_LIB, trace, env, nprocs=2)
    assert r["replies"] == ["0", "0", "0 444596224 1073741824", "0 444596224 1073741824", "2", "2"]


def test_limit_parser_matches_reference_rules():
    for v, want in [("1024m", 1 << 30), ("1g", 1 << 30), ("1G", 1 << 30), ("1048576k", 1 << 30), ("1073741824", 1 << 30),
                    ("0x40000000", 1 << 30), ("0", 0), ("", 0), (None, 0), ("abc", 0), ("17179869184g", 0), ("010", 8)]:
        assert limit_from_env(v) == want, v


@pytest.mark.skipif(not H.have_reference(), reason="reference binary only exists in the build container")
@pytest.mark.parametrize("seed,nprocs,env", [
    (101, 1, {"CUDA_DEVICE_MEMORY_LIMIT_0": "18432m"}),
    (102, 1, {"CUDA_DEVICE_MEMORY_LIMIT_0": "2g", "CUDA_OVERSUBSCRIBE": "true"}),
    (103, 4, {"CUDA_DEVICE_MEMORY_LIMIT_0": "18432m", "CUDA_OVERSUBSCRIBE": "true"}),
    (104, 3, {"CUDA_DEVICE_MEMORY_LIMIT": "512m"}),
    (105, 2, {}),
    # context billing live (mock NVML lists the process): every process first retains its primary context
    (106, 3, {"CUDA_DEVICE_MEMORY_LIMIT_0": "6g", "MOCK_CTX_MIB": "450"}),
    (107, 4, {"CUDA_DEVICE_MEMORY_LIMIT_0": "2g", "MOCK_CTX_MIB": "517", "CUDA_OVERSUBSCRIBE": "true"}),
    (108, 2, {"MOCK_CTX_MIB": "300"}),
])
def test_live_differential_against_reference_binary(seed, nprocs, env):
    trace = H.random_trace(seed, 2000, nprocs, 1 << 28)
    if "MOCK_CTX_MIB" in env:
        trace = H.with_contexts(trace, nprocs)
        rng = __import__("random").Random(seed)
        for _ in range(40):  # retain / release pairs sprinkled in: each re-retain after a release is billed again
            p = rng.randrange(nprocs)
            trace.insert(rng.randrange(nprocs, len(trace)), (p, rng.choice(["release 0", "retain 0 1", "retain 0 1", "@slots", "ctxget", "ctxdev"])))
    env = {**env, **MAP}
    a = H.run_trace(H.REFERENCE_LIB, trace, env, nprocs)
    b = H.run_trace(H.NEW_LIB, trace, env, nprocs)
    assert a["replies"] == b["replies"]
    assert a["oom"] == b["oom"]
    assert a["calls"] == b["calls"]
    assert H.mask_region(a["region"]) == H.mask_region(b["region"]), "final shared-cache bytes differ"


@pytest.mark.skipif(not H.have_reference(), reason="reference binary only exists in the build container")
@pytest.mark.parametrize("seed,nprocs,nkill,env", [
    (201, 4, 2, {"CUDA_DEVICE_MEMORY_LIMIT_0": "1g"}),
    (202, 6, 4, {"CUDA_DEVICE_MEMORY_LIMIT_0": "3g", "CUDA_OVERSUBSCRIBE": "true"}),
    (203, 5, 3, {"CUDA_DEVICE_MEMORY_LIMIT_0": "4g", "MOCK_CTX_MIB": "300"}),
])
def test_killed_processes_are_swept_like_the_reference_does(seed, nprocs, nkill, env):
    """SIGKILLed processes leave their slots behind; the survivors' OOM path sweeps them.  Replies, OOM lines, driver
    calls and the final cache file (which slot each survivor ends up in, stale bytes included) must match."""
    trace = H.random_trace(seed, 2500, nprocs, 1 << 27)
    if "MOCK_CTX_MIB" in env:
        trace = H.with_contexts(trace, nprocs)
    trace = H.with_kills(trace, seed, nprocs, nkill)
    env = {**env, **MAP}
    a = H.run_trace(H.REFERENCE_LIB, trace, env, nprocs)
    b = H.run_trace(H.NEW_LIB, trace, env, nprocs)
    assert a["replies"] == b["replies"]
    assert a["oom"] == b["oom"] and a["calls"] == b["calls"]
    assert H.mask_region(a["region"]) == H.mask_region(b["region"]), "final shared-cache bytes differ"


@pytest.mark.skipif(not H.have_reference(), reason="reference binary only exists in the build container")
def test_first_process_decides_the_limits_of_a_cache_file():
    """Processes of one cache file started with different limits / SM limits / priority: the file keeps what its first
    process wrote (the others log "Limit inconsistency detected" and live with it)."""
    envs = [{**MAP, "CUDA_DEVICE_MEMORY_LIMIT_0": "1g", "CUDA_DEVICE_SM_LIMIT": "30"},
            {**MAP, "CUDA_DEVICE_MEMORY_LIMIT_0": "2g", "CUDA_DEVICE_SM_LIMIT": "50", "CUDA_TASK_PRIORITY": "0"}, {**MAP}]
    tr = [(0, "info"), (1, "info"), (2, "info"), (1, f"alloc {1500 * M} 1"), (1, "total 0"), (2, f"alloc {900 * M} 1"), (2, "total 0"),
          (0, "nvmlmem"), (0, "@slots")]
    a = H.run_trace(H.REFERENCE_LIB, tr, envs, 3)
    b = H.run_trace(H.NEW_LIB, tr, envs, 3)
    assert a["replies"] == b["replies"] and a["oom"] == b["oom"] and a["calls"] == b["calls"]
    assert b["replies"][3] == "-1" and b["replies"][4] == f"0 {1 << 30}"          # the 2g process is held to 1g
    assert H.mask_region(a["region"]) == H.mask_region(b["region"])
    s = H.region_summary(b["region"])
    assert s["limit"][0] == 1 << 30 and s["sm_limit"][0] == 30 and s["priority"] == 1


def test_threads_of_one_process_hammer_the_hooks():
    """16 threads x 500 x (alloc, info, free) in each of two processes sharing a cache file: every call succeeds, nothing is
    left billed, the allocation list is empty again (thread safety of the list, the region lock and the mover gate)."""
    tr = [(0, "info"), (0, "mtloop 16 500 1048576"), (1, "mtloop 16 500 65536"), (0, "info"), (1, "info"), (0, "@slots"),
          (0, f"alloc {1000 * M} 1"), (1, "mtloop 8 50 33554432"), (1, "@slots")]      # 8 x 32 MiB does not always fit next to 1000 MiB
    r = H.run_trace(H.NEW_LIB, tr, {"CUDA_DEVICE_MEMORY_LIMIT_0": "1g", "MOCK_CUDA_LOG": "", **MAP}, 2)
    rep = r["replies"]
    assert rep[1] == rep[2] == "0" and rep[0] == rep[3] == rep[4] == f"0 {1 << 30} {1 << 30}"
    assert rep[5] == "n=2 p0:0:1 | p1:0:1"
    assert rep[8] == f"n=2 p0:0:1 d0=0/0/{1000 * M}/{1000 * M} | p1:0:1"             # refused or freed: nothing of process 1 stays billed


def _planted(clean, pids, used, limit0=None):
    import struct
    b = bytearray(clean)
    if limit0 is not None:
        struct.pack_into("<Q", b, 0x638, limit0)
    for i, (pid, u) in enumerate(zip(pids, used)):
        o = 0x738 + i * 0x310
        struct.pack_into("<ii", b, o, pid, pid)
        struct.pack_into("<5Q", b, o + 8, 0, 0, u, 0, u)
        struct.pack_into("<i", b, o + 0x308, 1)
    struct.pack_into("<i", b, 0xC4738, len(pids))
    return bytes(b)


@pytest.mark.skipif(not H.have_reference(), reason="reference binary only exists in the build container")
@pytest.mark.parametrize("case", ["zeros", "short", "garbage", "stale_slots", "stale_limit"])
def test_a_cache_file_with_history_is_taken_over_like_the_reference_does(case):
    """The cache file may outlive its container (monitor mode keeps it in a host directory): an all-zero file, a
    truncated one, 0xFF garbage (the reference never resets proc_num and claims slot -1 — so does this library),
    slots of processes that are gone, a limit that is not the one in the environment."""
    env = {"CUDA_DEVICE_MEMORY_LIMIT_0": "1g", **MAP}
    clean = H.run_trace(H.NEW_LIB, [(0, "info")], env)["region"]
    pre = {"zeros": b"\0" * (H.REGION_SIZE + 1), "short": b"abc" * 33, "garbage": b"\xff" * (H.REGION_SIZE + 1),
           "stale_slots": _planted(clean, [4000001, 4000002, 4000003], [300 * M, 200 * M, 100 * M]),
           "stale_limit": _planted(clean, [4000001], [100 * M], limit0=512 * M)}[case]
    tr = [(0, "info"), (0, "@slots"), (0, f"alloc {900 * M} 1"), (0, "info"), (0, "@slots")]
    a = H.run_trace(H.REFERENCE_LIB, tr, env, pre_cache=pre)
    b = H.run_trace(H.NEW_LIB, tr, env, pre_cache=pre)
    assert a["replies"] == b["replies"] and a["oom"] == b["oom"] and a["calls"] == b["calls"]
    assert H.mask_region(a["region"]) == H.mask_region(b["region"])
    if case == "stale_slots":
        assert b["replies"][1] == "n=1 p0:0:1" and b["replies"][2] == "0"          # the dead are kicked at start-up
    if case == "stale_limit":
        assert b["replies"][0] == f"0 {512 * M} {512 * M}" and b["replies"][2] == "-1"   # the file's limit wins


def test_an_unusable_cache_path_fails_closed():
    """CUDA_DEVICE_MEMORY_SHARED_CACHE is an environment variable the tenant can change: pointing it somewhere unusable
    must not buy a process without a quota.  (The reference crashes in its init in that case.)"""
    import subprocess
    env = H._env(H.NEW_LIB, "/nonexistent_dir/x.cache", {"CUDA_DEVICE_MEMORY_LIMIT_0": "1g", **MAP}, "")
    p = subprocess.run([os.path.join(H.REF_DIR, "trace_runner")], input="info\nalloc 2097152000 1\nquit\n", env=env, capture_output=True, text=True, timeout=30)
    assert p.returncode != 0 and "init 0" not in p.stdout and "alloc" not in p.stdout
    assert "refusing to run CUDA without the vGPU quota" in p.stderr


def test_cuGetProcAddress_hands_out_the_hook_with_the_abi_the_caller_asked_for():
    """An unsuffixed name asks for the newest ABI the given CUDA version knows: cuCtxCreate is _v2 / _v3 / _v4 depending
    on it, cuMemPrefetchAsync and cuMemAdvise change signature at 12.2, cuGetProcAddress itself at 12.0; per-thread
    default stream callers get the shim's _ptsz / _ptds twins of the copy / launch entry points and of the async allocator."""
    q = [("11030 0 cuCtxCreate", "cuCtxCreate_v2"), ("11040 0 cuCtxCreate", "cuCtxCreate_v3"), ("12090 0 cuCtxCreate", "cuCtxCreate_v4"),
         ("12010 0 cuMemPrefetchAsync", "cuMemPrefetchAsync"), ("12090 0 cuMemPrefetchAsync", "cuMemPrefetchAsync_v2"),
         ("12090 0 cuMemAdvise", "cuMemAdvise_v2"), ("11080 0 cuGetProcAddress", "cuGetProcAddress"), ("12000 0 cuGetProcAddress", "cuGetProcAddress_v2"),
         ("11030 0 cuDeviceGetUuid", "cuDeviceGetUuid"), ("12090 0 cuDeviceGetUuid", "cuDeviceGetUuid_v2"),
         ("12090 0 cuMemAlloc", "cuMemAlloc_v2"), ("12090 0 cuMemAllocAsync", "cuMemAllocAsync"), ("12090 2 cuMemAllocAsync", "cuMemAllocAsync_ptsz"),
         ("12090 2 cuMemFreeAsync", "cuMemFreeAsync_ptsz"), ("12090 0 cuLaunchKernel", "cuLaunchKernel"), ("12090 0 cuMemcpy2DAsync", "cuMemcpy2DAsync_v2")]
    rep = H.run_trace(H.NEW_LIB, [(0, "procaddr " + a) for a, _ in q], {"CUDA_DEVICE_MEMORY_LIMIT_0": "1g", **MAP})["replies"]
    assert rep == [f"0 {sym}@libvgpu.so" for _, sym in q]
    # per-thread-default-stream flavours of copies and launches are the shim's own twins (they gate, then call the
    # driver's twin); anything not hooked comes straight from the driver
    rep = H.run_trace(H.NEW_LIB, [(0, "procaddr 12090 2 cuLaunchKernel"), (0, "procaddr 12090 2 cuMemcpyHtoDAsync"),
                                  (0, "procaddr 12090 2 cuMemcpyDtoH"), (0, "procaddr 12090 2 cuGraphLaunch"),
                                  (0, "procaddr 12090 0 cuDriverGetVersion")],
                      {"CUDA_DEVICE_MEMORY_LIMIT_0": "1g", **MAP})["replies"]
    assert rep[:4] == ["0 cuLaunchKernel_ptsz@libvgpu.so", "0 cuMemcpyHtoDAsync_v2_ptsz@libvgpu.so", "0 cuMemcpyDtoH_v2_ptds@libvgpu.so",
                       "0 cuGraphLaunch_ptsz@libvgpu.so"], rep
    assert rep[4].endswith("@libcuda.so.1"), rep


GATE_PROBES = [("launch 4 128", True), (f"alloc {M} 9", True), (f"managed {M} 9", True), ("pitch 64 4 4 9", True), ("hostalloc 4096 9", True),
               ("hostalloc2 4096 9", True), ("hostreg 4096", True), ("ptrattrs 1", True), ("info", False), ("total 0", False), ("free 1", False),
               ("freehost 2", False), ("nvmlmem", False), ("devcount", False), ("ctxget", False)]


@pytest.mark.parametrize("probe,parks", GATE_PROBES, ids=[p.split()[0] for p, _ in GATE_PROBES])
def test_which_calls_park_while_the_vgpu_is_suspended(probe, parks):
    """SIGUSR2 sets the status word to 2 (suspended): launches and allocations (device, managed, pitch, pinned host,
    pointer queries) wait at the gate until SIGUSR1, frees and read-only queries go through — call by call as the
    reference does (wait_status_self@0x45ff5 in its hooks)."""
    env = {"CUDA_DEVICE_MEMORY_LIMIT_0": "1g", **MAP}
    tr = [(0, f"alloc {M} 1"), (0, "hostalloc 4096 2"), (0, "@usr2"), (0, "@send " + probe), (0, "@poll" if parks else "@pollready"), (0, "@usr1"),
          (0, "@recv"), (0, "info")]
    libs = [H.NEW_LIB] + ([H.REFERENCE_LIB] if H.have_reference() else [])
    seen = []
    for lib in libs:
        r = H.run_trace(lib, tr, env)["replies"]
        assert r[4] == ("blocked" if parks else "ready"), (lib, probe, r)
        seen.append(r[6:])
    assert all(x == seen[0] for x in seen)                     # and the parked call completes with the same answer afterwards


ODD_OPS = ["total 5", "total 15", "total 16", "total 100", "total -1", "info", "pitch 1 1 1 1", "pitch 4097 3 4 2", "pitch 100 0 8 3", "info",
           "managed 0 4", "info", "alloc 1 5", "alloc 2097152 6", "alloc 2097153 7", "info", "hostreg 0", "hostalloc 0 8", "freehost 8",
           "ptrattrs 5", "ptrattrs 0", "nvmlmem", "free 6", "free 7", "free 5", "info"]


@pytest.mark.skipif(not H.have_reference(), reason="reference binary only exists in the build container")
@pytest.mark.parametrize("env", [
    {"CUDA_DEVICE_MEMORY_LIMIT": "1g", "CUDA_DEVICE_MEMORY_LIMIT_5": "2g"},
    {"CUDA_DEVICE_MEMORY_LIMIT_0": "16m", "CUDA_OVERSUBSCRIBE": "true"},
])
def test_odd_arguments_get_the_reference_answers(env):
    """Device ordinals nobody has (5, 15, 16 — where the reference reads one past limit[] into sm_limit[0] —, 100, -1),
    zero-sized and one-byte allocations, the 2 MiB boundary of the oversubscribe route, zero-height pitch."""
    env = {**env, **MAP}
    tr = [(0, o) for o in ODD_OPS]
    a = H.run_trace(H.REFERENCE_LIB, tr, env)
    b = H.run_trace(H.NEW_LIB, tr, env)
    assert a["replies"] == b["replies"] and a["oom"] == b["oom"] and a["calls"] == b["calls"]
    assert a["replies"][2] == "0 100"                        # limit[16] is sm_limit[0]
    assert H.mask_region(a["region"]) == H.mask_region(b["region"])


@pytest.mark.skipif(not H.have_reference(), reason="reference binary only exists in the build container")
@pytest.mark.parametrize("seed,ndev,nprocs,nlate", [(3, 1, 4, 0), (7, 2, 4, 0), (9, 1, 6, 2)])
def test_kitchen_sink_fuzz_against_the_reference(seed, ndev, nprocs, nlate):
    """scripts/quota_fuzz.py: every op the runner knows in one random trace (allocations of all kinds, contexts, the
    monitor's helpers, fork, SIGKILL, launches, live quota resizes), three environments, four processes."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("quota_fuzz", os.path.join(os.path.dirname(os.path.dirname(__file__)), "scripts", "quota_fuzz.py"))
    fuzz = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(fuzz)
    for extra in fuzz.ENVS:
        assert fuzz.differences(seed, extra, nprocs=nprocs, ndev=ndev, nlate=nlate) == []


def _slots_clean(region):
    """(proc_num, no live slot holds memory).  Slots beyond proc_num may hold stale bytes: the exit
    handler moves the last slot into the leaver's and does not wipe the source (as in the reference)."""
    proc_num = struct.unpack_from("<i", region, 0xC4738)[0]
    used = [struct.unpack_from("<80Q", region, 0x738 + 0x310 * s + 8) for s in range(proc_num)]
    return proc_num, all(not any(u) for u in used)


def test_eight_processes_hammer_one_cache_file_concurrently():
    """Not lock-step: 8 processes x 20 000 alloc/info/free triples at once on one shared cache.
    The semaphore protocol must neither deadlock nor lose a counter update."""
    env = {"CUDA_DEVICE_MEMORY_LIMIT_0": "18432m", "MOCK_CUDA_LOG": "", **MAP}
    r = H.run_trace(H.NEW_LIB, [(p, "loop 20000 1048576") for p in range(8)], env, nprocs=8, parallel=True)
    assert all(x.split()[0] == "0" for x in r["replies"])
    assert _slots_clean(r["region"]) == (0, True)


@pytest.mark.skipif(not H.have_reference(), reason="reference binary only exists in the build container")
def test_old_and_new_library_share_one_cache_file():
    """Contract C interop: processes preloading the reference binary and processes preloading the new
    shim attach to the same file, see each other's usage, and leave it clean."""
    MiB = 1 << 20
    env = {"CUDA_DEVICE_MEMORY_LIMIT_0": "1024m", **MAP}
    libs = [H.REFERENCE_LIB, H.NEW_LIB, H.REFERENCE_LIB, H.NEW_LIB]
    trace = [(0, f"alloc {300 * MiB} 1"), (1, f"alloc {200 * MiB} 1"), (2, "info"), (3, "info"),
             (2, f"alloc {400 * MiB} 1"), (3, f"alloc {200 * MiB} 1"),      # 300+200+400 = 900; +200 > 1024 -> OOM in the new lib
             (1, f"managed {200 * MiB} 2"), (0, "free 1"), (1, "info"), (3, f"alloc {200 * MiB} 2"), (0, "info")]
    r = H.run_trace(libs, trace, env, nprocs=4)
    free_after_two = str((1024 - 500) * MiB)
    assert r["replies"][2].split()[1] == free_after_two and r["replies"][3].split()[1] == free_after_two
    assert r["replies"][4] == "0" and r["replies"][5] == "-1" and r["replies"][6] == "2"
    assert r["replies"][8].split()[1] == str((1024 - 600) * MiB)
    assert r["replies"][9] == "0" and r["replies"][10].split()[1] == str((1024 - 800) * MiB)
    assert _slots_clean(r["region"]) == (0, True)
    # and a concurrent mixed hammer
    r = H.run_trace(libs, [(p, "loop 5000 1048576") for p in range(4)], {**env, "MOCK_CUDA_LOG": ""}, nprocs=4, parallel=True)
    assert all(x.split()[0] == "0" for x in r["replies"]) and _slots_clean(r["region"]) == (0, True)


def test_two_devices_in_one_container_have_separate_limits_and_usage():
    """A container holding vGPUs of two physical GPUs: CUDA_DEVICE_MEMORY_LIMIT_0/_1 and
    NVIDIA_DEVICE_MAP "0:<A> 1:<B>" (server.go:487-493).  New shim only: the reference's device
    remap layer rejects contexts under the mock ("Ctx Not Found"), so this row has no live oracle;
    expectations follow Appendix B (limit[d], usage(d) per device index)."""
    MiB = 1 << 20
    U2 = "GPU-00000000-0000-0000-0000-000000000002"
    env = {"MOCK_GPU_COUNT": "2", "CUDA_DEVICE_MEMORY_LIMIT_0": "1024m", "CUDA_DEVICE_MEMORY_LIMIT_1": "512m",
           "NVIDIA_DEVICE_MAP": f"0:{H.UUID0} 1:{U2}"}
    tr = [(0, "setdev 0"), (0, f"alloc {600 * MiB} 1"), (0, "info"), (0, "setdev 1"), (0, "info"), (0, f"alloc {600 * MiB} 2"),
          (0, f"alloc {300 * MiB} 3"), (0, "info"), (0, "total 1"), (0, "total 0"), (0, "setdev 0"), (0, "info"), (0, "free 3"),
          (0, "setdev 1"), (0, "info")]
    r = H.run_trace(H.NEW_LIB, tr, env)
    rep = r["replies"]
    assert rep[2] == f"0 {424 * MiB} {1024 * MiB}"            # device 0: 1024 - 600
    assert rep[4] == f"0 {512 * MiB} {512 * MiB}"             # device 1 untouched
    assert rep[5] == "-1" and r["oom"] == [(1, 600 * MiB, 512 * MiB)]   # 600 MiB does not fit device 1's 512 MiB
    assert rep[6] == "0" and rep[7] == f"0 {212 * MiB} {512 * MiB}"
    assert rep[8] == f"0 {512 * MiB}" and rep[9] == f"0 {1024 * MiB}"  # cuDeviceTotalMem per device
    assert rep[11] == f"0 {424 * MiB} {1024 * MiB}"           # back on device 0: unchanged by device 1's allocation
    assert rep[12] == "0" and rep[14] == f"0 {512 * MiB} {512 * MiB}"
    s = H.region_summary(r["region"])
    assert s["device_num"] == 2 and s["limit"][:3] == [1024 * MiB, 512 * MiB, 0]


def test_monitor_reads_a_live_cache_file(vgpu):
    """The external-agent view (monitor.py) of a cache two live processes are attached to."""
    import subprocess
    import tempfile
    from vgpu_b200 import monitor
    MiB = 1 << 20
    td = tempfile.mkdtemp(prefix="mon")
    cache = os.path.join(td, "shr.cache")
    # MOCK_CTX_MIB: the mock NVML lists the processes, so the shim learns their host pids at the first context (the
    # monitor signals host pids only: the slot's `pid` means nothing outside the container's pid namespace)
    env = H._env(H.NEW_LIB, cache, {"CUDA_DEVICE_MEMORY_LIMIT_0": "1024m", "CUDA_DEVICE_SM_LIMIT": "25", "MOCK_CTX_MIB": "10", **MAP},
                 os.path.join(td, "mock.log"))
    procs = [subprocess.Popen([os.path.join(H.REF_DIR, "trace_runner")], stdin=subprocess.PIPE, stdout=subprocess.PIPE,
                              stderr=subprocess.DEVNULL, env=env, text=True, bufsize=1) for _ in range(2)]
    try:
        for i, p in enumerate(procs):
            assert p.stdout.readline().startswith("init")
            assert monitor.pids(cache) == [q.pid for q in procs[:i]] or i == 0   # no context yet: not reachable
            p.stdin.write("retain 0 1\n"); p.stdin.flush()
            assert p.stdout.readline().strip().startswith("0")
            p.stdin.write(f"alloc {(90 + 50 * i) * MiB} 1\n"); p.stdin.flush()
            assert p.stdout.readline().strip() == "0"
        r = monitor.read_region(cache)
        assert r["device_num"] == 1 and r["uuids"] == [H.UUID0] and r["limit"] == [1024 * MiB] and r["sm_limit"] == [25]
        assert r["proc_num"] == 2 and sorted(p["pid"] for p in r["procs"]) == sorted(p.pid for p in procs)
        assert sorted(p["used"][0]["total"] for p in r["procs"]) == [100 * MiB, 150 * MiB] and r["usage"] == [250 * MiB]
        assert all(p["status"] == 1 for p in r["procs"])
        assert sorted(monitor.pids(cache)) == sorted(p.pid for p in procs) and monitor.unreachable(cache) == []
        monitor.suspend(cache)                      # SIGUSR2 -> status 2 in every slot
        import time
        for _ in range(50):
            if all(p["status"] == 2 for p in monitor.read_region(cache)["procs"]):
                break
            time.sleep(0.05)
        assert all(p["status"] == 2 for p in monitor.read_region(cache)["procs"])
        monitor.resume(cache)
        for _ in range(50):
            if all(p["status"] == 1 for p in monitor.read_region(cache)["procs"]):
                break
            time.sleep(0.05)
        assert all(p["status"] == 1 for p in monitor.read_region(cache)["procs"])
    finally:
        for p in procs:
            p.stdin.write("quit\n"); p.stdin.flush(); p.wait(timeout=20)
    assert monitor.read_region(cache)["proc_num"] == 0


def test_shim_exports_the_hooked_driver_symbols():
    import subprocess
    out = subprocess.run(["nm", "-D", "--defined-only", H.NEW_LIB], capture_output=True, text=True).stdout
    names = {line.split()[-1] for line in out.splitlines() if line.strip()}
    for n in ["cuInit", "cuMemAlloc_v2", "cuMemAllocManaged", "cuMemAllocPitch_v2", "cuMemFree_v2", "cuMemGetInfo_v2",
              "cuDeviceTotalMem_v2", "cuMemAllocHost_v2", "cuMemHostAlloc", "cuMemHostRegister_v2", "cuLaunchKernel",
              "cuGetProcAddress", "cuGetProcAddress_v2", "cuDevicePrimaryCtxRetain", "nvmlDeviceGetMemoryInfo", "dlsym",
              "cuVGPUViewAllocator", "get_gpu_memory_usage", "set_gpu_device_memory_monitor", "get_current_device_memory_limit",
              "set_current_device_memory_limit", "suspend_all", "resume_all", "wait_status_self"]:
        assert n in names, n
    internal = [n for n in names if n.startswith(("region_", "shim_", "oom_", "add_chunk"))]
    assert not internal, f"internal helpers leak into the dynamic symbol table: {internal}"


@pytest.mark.skipif(not H.have_reference(), reason="reference binary only exists in the build container")
def test_device_uuid_of_every_vgpu_is_the_physical_gpus_as_in_the_reference():
    """cuDeviceGetUuid through the reference answers with the physical GPU's UUID for a plain vGPU and for both vGPUs of
    a GPU named twice (all 16 bytes compared, not just the last one the `devget` probe prints)."""
    for env, n in (({"NVIDIA_DEVICE_MAP": "0:" + H.UUID0, "CUDA_DEVICE_MEMORY_LIMIT_0": "1g"}, 1),
                   ({"NVIDIA_DEVICE_MAP": f"0:{H.UUID0} 1:{H.UUID0}", "CUDA_DEVICE_MEMORY_LIMIT_0": "1g", "CUDA_DEVICE_MEMORY_LIMIT_1": "1g"}, 2)):
        tr = [(0, "devcount")] + [(0, f"uuidfull {i}") for i in range(n)]
        a = H.run_trace(H.REFERENCE_LIB, tr, env)["replies"]
        b = H.run_trace(H.NEW_LIB, tr, env)["replies"]
        assert a == b and all(r == "0 " + H.UUID0[4:].replace("-", "") for r in b[1:]), (a, b)
