"""The helpers an outside monitor links from libvgpu.so (MEMORY_OVERRIDE, ACTIVE_OOM_KILLER, the
suspend / resume protocol, live quota resize).  Expectations were read off the reference binary under the mock driver
(multiprocess_memory_limit.c: set_gpu_device_memory_monitor@0x42e29, get_gpu_memory_usage@0x42be5, suspend_all@0x45ddf
...); where /root/reference exists every trace is also replayed through it and compared reply by reply."""
import pytest

from oracle import quota_harness as H

M = 1 << 20
G = 1 << 30
C = 300 * M
ENV = {"CUDA_DEVICE_MEMORY_LIMIT_0": "1g", "CUDA_DEVICE_SM_LIMIT": "30", "NVIDIA_DEVICE_MAP": "0:" + H.UUID0, "MOCK_CTX_MIB": "300"}


@pytest.fixture(scope="module", autouse=True)
def harness():
    H.build()


def both(trace, env, nprocs):
    """Replies of the new library; compared with the reference's where it can run."""
    new = H.run_trace(H.NEW_LIB, trace, env, nprocs)
    if H.have_reference():
        ref = H.run_trace(H.REFERENCE_LIB, trace, env, nprocs)
        bad = [(i, trace[i], ref["replies"][i], new["replies"][i]) for i in range(len(trace)) if ref["replies"][i] != new["replies"][i]]
        assert not bad, bad[:5]
        assert H.mask_region(ref["region"]) == H.mask_region(new["region"])
    return new["replies"]


def test_getters_answer_from_the_cache_file():
    tr = [(0, "retain 0 1"), (1, "retain 0 1"), (0, f"alloc {100 * M} 1"), (1, f"alloc {50 * M} 1"),
          (0, "libcall 0 0 0 get_gpu_memory_usage"), (1, "libcall 0 0 0 get_current_device_memory_usage"), (1, "libcall 1 0 0 get_gpu_memory_usage"),
          (0, "libcall 0 0 0 get_current_device_memory_limit"), (0, "libcall 16 0 0 get_current_device_memory_limit"),
          (0, "libcall 0 0 0 get_current_device_sm_limit"), (0, "libcall 0 0 0 get_current_priority"), (0, "libcall 0 0 0 get_recent_kernel"),
          (0, "libcall 5 0 0 set_recent_kernel"), (1, "libcall 0 0 0 get_recent_kernel"), (0, "libcall 0 0 0 get_utilization_switch"),
          (0, "libcall 0 0 0 get_gpu_memory_monitor")]
    rep = both(tr, ENV, 2)
    total = 2 * C + 150 * M
    assert rep[4:] == [f"0 {total}", f"0 {total}", "0 0", f"0 {G}", "0 30", "0 30", "0 1", "0 2", "0 0", "0 5", "0 1", "0 0"]


def test_memory_monitor_is_recorded_per_host_pid():
    tr = [(0, "retain 0 1"), (1, "retain 0 1"), (0, f"alloc {100 * M} 1"),
          (1, f"libcall $pid0 0 {500 * M} set_gpu_device_memory_monitor"), (0, "@slots"), (1, "libcall 0 0 0 get_gpu_memory_monitor"),
          (0, "libcall 0 0 0 get_current_device_memory_monitor"), (0, "info"), (0, "nvmlmem"),
          (1, "libcall 424242 0 5 set_gpu_device_memory_monitor"),                      # nobody's host pid: answered 1, nothing changes
          (1, f"libcall $pid1 3 {7 * M} set_gpu_device_memory_monitor"), (1, "libcall 3 0 0 get_gpu_memory_monitor"), (0, "@slots")]
    rep = both(tr, ENV, 2)
    assert rep[3] == "0 1" and rep[5] == rep[6] == f"0 {500 * M}"
    assert rep[4] == f"n=2 p0:1:1 d0={C}/0/{100 * M}/{C + 100 * M} mon={500 * M} | p1:1:1 d0={C}/0/0/{C}"
    assert rep[7] == f"0 {G - 2 * C - 100 * M} {G}"                                    # without MEMORY_OVERRIDE the quota ignores it
    assert rep[9] == "0 1" and rep[11] == f"0 {7 * M}"
    assert rep[12].endswith(f"| p1:1:1 d0={C}/0/0/{C} mon=0,0,0,{7 * M}")


@pytest.mark.parametrize("value,on", [("1", True), ("true", True), ("", True), ("on", True), ("0", False), ("false", False), (None, False)])
def test_memory_override_makes_the_monitor_figure_the_usage(value, on):
    """Set to anything but "0" / "false": total := what the monitor measured (offset = measured - total), so memory the
    hooks never saw is billed; later allocations and frees move the total from there."""
    env = dict(ENV) if value is None else {**ENV, "MEMORY_OVERRIDE": value}
    tr = [(0, "retain 0 1"), (1, "retain 0 1"), (0, f"alloc {100 * M} 1"),
          (1, f"libcall $pid0 0 {900 * M} set_gpu_device_memory_monitor"), (0, "@slots"), (0, "info"), (0, f"alloc {200 * M} 2"),
          (1, f"libcall $pid0 0 {350 * M} set_gpu_device_memory_monitor"), (0, "@slots"), (0, "free 1"), (0, "@slots"), (0, "info"), (0, "nvmlmem")]
    rep = both(tr, env, 2)
    if on:
        assert rep[4].startswith(f"n=2 p0:1:1 d0={C}/0/{100 * M}/{900 * M}+off{500 * M} mon={900 * M} |")
        assert rep[5] == "1 0 0" and rep[6] == "-1"                                   # 900 + 300 MiB > 1 GiB: over the limit
        assert rep[8].startswith(f"n=2 p0:1:1 d0={C}/0/{100 * M}/{350 * M}+off{(350 * M - 900 * M) % (1 << 64)} mon={350 * M} |")
        assert rep[10].startswith(f"n=2 p0:1:1 d0={C}/0/0/{250 * M}+off")
        assert rep[11] == f"0 {G - 250 * M - C} {G}"
    else:
        assert rep[4].startswith(f"n=2 p0:1:1 d0={C}/0/{100 * M}/{C + 100 * M} mon={900 * M} |")
        assert rep[5] == f"0 {G - 2 * C - 100 * M} {G}" and rep[6] == "0"
        assert rep[11] == f"0 {G - 2 * C - 200 * M} {G}"


@pytest.mark.parametrize("env_extra,bytes_,dies", [
    ({}, 1126 * M, False),                                  # 1.0996 x the limit
    ({}, 1127 * M, True),                                   # above 1.1 x: every process of the vGPU is killed
    ({"ACTIVE_OOM_KILLER": "0"}, 1500 * M, False), ({"ACTIVE_OOM_KILLER": "false"}, 1500 * M, False),
    ({"ACTIVE_OOM_KILLER": "no"}, 1500 * M, True), ({"ACTIVE_OOM_KILLER": "1"}, 1500 * M, True),
    ({"CUDA_DEVICE_MEMORY_LIMIT_0": "0"}, 1500 * M, False),  # an unlimited vGPU is never killed
])
def test_active_oom_killer(env_extra, bytes_, dies):
    tr = [(0, "retain 0 1"), (1, "retain 0 1"), (1, f"libcall $pid0 0 {bytes_} set_gpu_device_memory_monitor"), (1, "libcall 0 0 0 get_gpu_memory_monitor")]
    libs = [H.NEW_LIB] + ([H.REFERENCE_LIB] if H.have_reference() else [])
    for lib in libs:
        try:
            rep = H.run_trace(lib, tr, {**ENV, **env_extra}, 2)["replies"]
            survived = True
        except RuntimeError as e:                           # the runner that made the call is in the file too: it dies with the rest
            assert "died on" in str(e)
            survived = False
        assert survived == (not dies), (lib, env_extra, bytes_)
        if survived:
            assert rep[2:] == ["0 1", f"0 {bytes_}"]


def test_a_gpu_named_twice_is_never_swept_by_the_killer():
    env = {**ENV, "NVIDIA_DEVICE_MAP": f"0:{H.UUID0} 1:{H.UUID0}", "CUDA_DEVICE_MEMORY_LIMIT_1": "1g"}
    tr = [(0, "retain 0 1"), (0, f"libcall $pid0 0 {1500 * M} set_gpu_device_memory_monitor"), (0, "libcall 0 0 0 get_gpu_memory_monitor")]
    assert H.run_trace(H.NEW_LIB, tr, env, 1)["replies"][1:] == ["0 1", f"0 {1500 * M}"]
    if H.have_reference():
        assert H.run_trace(H.REFERENCE_LIB, tr, env, 1)["replies"][1:] == ["0 1", f"0 {1500 * M}"]


def test_quota_resize_and_suspend_resume_from_the_monitor():
    tr = [(0, "retain 0 1"), (1, "retain 0 1"), (0, f"alloc {100 * M} 1"),
          (1, f"libcall 0 {2048 * M} 0 set_current_device_memory_limit"), (0, "info"), (0, "total 0"),   # takes effect at once, for everybody
          (1, "libcall 0 50 0 set_current_device_sm_limit_scale"), (0, "libcall 0 3 0 set_current_device_sm_limit_scale"),   # once per file
          (0, "libcall 0 0 0 get_current_device_sm_limit"),
          (1, "libcall 0 0 0 suspend_all"), (0, "@sleep"), (0, "@slots"), (1, "libcall 0 0 0 resume_all"), (0, "@sleep"), (0, "@slots"),
          (0, "libcall 2 0 0 set_current_gpu_status"), (1, "@slots"), (0, "libcall 1 0 0 set_current_gpu_status"),
          (0, "libcall 1 0 0 wait_status_self"), (1, "libcall 1 0 0 wait_status_all")]
    rep = both(tr, ENV, 2)
    assert rep[3] == "0 0" and rep[4] == f"0 {2 * G - 2 * C - 100 * M} {2 * G}" and rep[5] == f"0 {2 * G}"
    assert rep[8] == "0 1500"                                                        # 30 x 50: a multiplication, and the x 3 is ignored
    assert rep[9] == "0 2" and rep[11].count(":1:2") == 2                            # SIGUSR2 reached both processes
    assert rep[12] == "0 2" and rep[14].count(":1:1") == 2
    assert rep[15] == "0 2" and rep[16].startswith("n=2 p0:1:2") and rep[17] == "0 1" and rep[18] == rep[19] == "0 1"


def test_nvml_shows_the_monitor_figure_when_it_is_larger_and_override_is_on():
    """nvmlDeviceGetMemoryInfo in a process started with MEMORY_OVERRIDE: used = max(hook-counted usage, what the monitor
    reported); a process without it keeps the counted usage.  The quota itself (cuMemGetInfo) is not touched by this."""
    envs = [{**ENV, "MEMORY_OVERRIDE": "1"}, ENV]
    tr = [(0, "retain 0 1"), (1, "retain 0 1"), (0, f"alloc {100 * M} 1"), (0, "nvmlmem"),
          (1, f"libcall $pid0 0 {900 * M} set_gpu_device_memory_monitor"), (0, "nvmlmem"), (1, "nvmlmem"), (0, "info"),
          (1, f"libcall $pid0 0 {200 * M} set_gpu_device_memory_monitor"), (0, "nvmlmem")]
    rep = both(tr, envs, 2)
    counted = 2 * C + 100 * M
    assert rep[3] == rep[6] == rep[9] == f"0 {G} {G - counted} {counted}"
    assert rep[5] == f"0 {G} {G - 900 * M} {900 * M}" and rep[7] == f"0 {G - counted} {G}"


def test_override_env_file_is_parsed_like_the_reference(tmp_path):
    """load_env_from_file@0x420cc ("/overrideEnv" in the container): KEY=VALUE per line, over the existing environment,
    value = everything after the first '=', stops at the first line without '=', a last line without newline counts."""
    e1, e2 = tmp_path / "e1", tmp_path / "e2"
    e1.write_text("A=1\nB=x=y=z\nC=\nSTOPS_AFTER_ME=1\n\nNEVER=reached\n")
    e2.write_text("LAST=nonl")
    ops = ["getenv A", f"libcalls load_env_from_file {e1}", "getenv A", "getenv B", "getenv C", "getenv STOPS_AFTER_ME", "getenv NEVER",
           f"libcalls load_env_from_file {e2}", "getenv LAST", f"libcalls load_env_from_file {tmp_path / 'missing'}"]
    env = {"CUDA_DEVICE_MEMORY_LIMIT_0": "1g", "NVIDIA_DEVICE_MAP": "0:" + H.UUID0, "A": "old"}
    rep = both([(0, o) for o in ops], env, 1)
    assert rep == ["0 old", "0 0", "0 1", "0 x=y=z", "0", "0 1", "0 -", "0 0", "0 nonl", "0 0"]


def test_recent_kernel_is_the_activity_mark_and_the_monitors_brake():
    """rate_limiter@0x463b1, first lines: while recent_kernel < 0 every launch of the vGPU waits (the monitor parks
    low-priority tasks that way), and a launch leaves recent_kernel = 2 — once the host pid is known; a process NVML does
    not list is left alone."""
    tr = [(0, "retain 0 1"), (1, "retain 0 1"), (1, "libcall 0 0 0 set_recent_kernel"), (0, "launch 4 128"), (1, "libcall 0 0 0 get_recent_kernel"),
          (1, "libcall -1 0 0 set_recent_kernel"), (0, "@send launch 4 128"), (0, "@poll"), (1, "libcall 0 0 0 set_recent_kernel"), (0, "@recv"),
          (1, "libcall 0 0 0 get_recent_kernel")]
    rep = both(tr, ENV, 2)
    assert rep[4] == "0 2" and rep[7] == "blocked" and rep[9] == "0" and rep[10] == "0 2"
    unseen = {k: v for k, v in ENV.items() if k != "MOCK_CTX_MIB"}      # NVML lists no process: no host pid, no lever
    rep = both([(0, "libcall 0 0 0 set_recent_kernel"), (0, "launch 4 128"), (0, "libcall 0 0 0 get_recent_kernel"),
                (0, "libcall -1 0 0 set_recent_kernel"), (0, "launch 4 128"), (0, "libcall 0 0 0 get_recent_kernel")], unseen, 1)
    assert rep == ["0 0", "0", "0 0", "0 0", "0", "0 4294967295"]


def test_a_new_tenant_of_a_slot_does_not_inherit_the_monitor_figure():
    """set_host_pid@0x455d8 wipes monitorused[] of the slot it writes the host pid into: a process that takes over the
    slot of one that died starts with nothing reported."""
    tr = [(0, "retain 0 1"), (1, "retain 0 1"), (0, f"libcall $pid1 0 {400 * M} set_gpu_device_memory_monitor"), (0, "@slots"), (1, "@kill"),
          (0, f"alloc {1000 * M} 1"),                       # over the limit: the sweep removes the dead process, then it fits
          (2, "@start"), (2, "retain 0 1"), (0, "@slots"), (0, "libcall 0 0 0 get_gpu_memory_monitor")]
    rep = both(tr, ENV, 3)
    assert rep[3].endswith(f"| p1:1:1 d0={C}/0/0/{C} mon={400 * M}")
    assert rep[8].endswith(f"| p2:1:1 d0={C}/0/0/{C}") and rep[9] == "0 0"
