#!/usr/bin/env python
"""bench.py — BASELINE.json's metric on BASELINE.json's configs.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Headline (`value`): vgpuvalidator memtest HBM GB/s on config 2 — 4 vGPUs x 40 GiB on each
B200, each vGPU in its own 25 % green-context SM partition, one step = one walking-ones
pass + one random pass (fill kernel, then verify kernel) over every vGPU buffer;
algorithmic bytes = 2N per pass.  `spill_restore` carries the other half
of the metric: 2 MiB-chunk spill/restore GB/s per tier (local HBM, pinned host, and the
peer GPU over NVLink: ring g -> g+1 for N>1, GPU 1 as the idle peer when N=1 sees two
GPUs) and per kernel variant; `roofline.nvlink` / `roofline.pcie` quote the largest batch
against the nominal link rate and the copy-engine rate measured in the same run.
`config5` is BASELINE config 5 on every GPU (10 vGPUs x 16 GiB memtest concurrent with
the spill loop).  `e2e` is the step through the blocking C-ABI calls with host buffers.
One JSON line on stdout (rank 0).  See DESIGN.md §Measurement.
"""
import argparse
import ctypes as C
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
# One hardware work queue per vGPU stream.  CUDA's default is 8 connections per context; the ten partition streams of
# config 5 (plus the transfer stream) then alias onto shared queues and kernels of DIFFERENT vGPUs serialise behind
# each other in launch order — measured: 3.9 TB/s instead of 6.5-7.2 TB/s for ten 12-14-SM partitions
# (profiles/r02_concurrency.md).  Must be set before the CUDA context exists.  12 = ten vGPU streams + the transfer
# stream + the default stream: every further connection makes context creation slower (the peer tier's one-time
# set-up: 0.36 s with 8, 0.47 s with 12, 1.8 s with 32) for nothing.
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "12")

METRIC = "vGPU spill/restore GB/s per GPU + 8-GPU aggregate; memtest HBM GB/s vs peak"
GIB = 1 << 30
NVLINK_NOMINAL = 900.0  # GB/s per direction (north_star)
PCIE_NOMINAL = 64.0     # GB/s per direction, Gen5 x16 raw (north_star)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--vgpus", type=int, default=4, help="vGPUs per GPU (config 2: 4)")
    ap.add_argument("--vgpu-gib", type=float, default=40.0, help="memtest buffer per vGPU in GiB (config 2: 40)")
    ap.add_argument("--spill-chunks", type=int, default=4096, help="largest batch of the 2 MiB chunk sweep (config 4: 4096)")
    ap.add_argument("--partition", default="green-fine", choices=["green", "green-fine", "streams"],
                    help="green-fine (default): one SM-granular green-context partition per vGPU (the memtest kernels use no "
                         "clusters, so the validator may take the fine split: 4 x 36, 10 x 14 SMs); green: the 8-SM granule that "
                         "application contexts get (4 x 32, 10 x 8); streams: plain streams (ncu cannot enter green contexts)")
    ap.add_argument("--no-spill", action="store_true", help="skip the spill/restore sweep")
    ap.add_argument("--config4", action="store_true",
                    help="also run BASELINE config 4 through the virtual-device-memory manager: 1 vGPU, 180 GB quota + 120 GB "
                         "oversubscribed, paged in 8 GiB windows (tier: GPU 1's HBM if --gpus >= 2, else pinned host at --config4-scale)")
    ap.add_argument("--config4-scale", type=float, default=None, help="size factor for config 4 (default 1.0 with a peer GPU, 0.2 on the host tier)")
    ap.add_argument("--no-config5", action="store_true", help="skip BASELINE config 5 (10 vGPUs x 16 GiB memtest per GPU concurrent with the spill loop)")
    ap.add_argument("--config5-gib", type=float, default=16.0, help="memtest buffer per vGPU in config 5 (16)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--cpu-gib", type=float, default=4.0, help="host buffer of the CPU memtest sample")
    ap.add_argument("--e2e-chunks", type=int, default=256,
                    help="2 MiB chunks the e2e step spills to / restores from a host buffer (256 = 512 MiB each way per GPU per step: fits inside "
                         "the memtest step's window even where eight GPUs share one host path at 12 GB/s each; the host rates themselves are in "
                         "e2e.components and roofline.pcie)")
    ap.add_argument("--xfer-priority", default="normal", choices=["normal", "high"],
                    help="config 5: stream priority of the spill / restore loop beside the memtest partitions")
    ap.add_argument("--peer-grid", type=int, default=0, help="config 5: CTAs per peer-tier transfer launch (0 = the library's default)")
    ap.add_argument("--e2e-variant", default="auto", choices=["auto", "ldst", "tma", "ce"], help="transfer variant of the e2e step's host-buffer calls")
    return ap.parse_args()


def traffic_per_launch(kernel, launch_bytes):
    """DRAM bytes per launch from the committed ncu --set full capture (profiles/r02_traffic.json),
    scaled from the profiled launch size to this run's launch size."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "r02_traffic.json")))["kernels"][kernel]
        return int(t["traffic_ratio"] * launch_bytes), t["traffic_ratio"]
    except Exception:
        return None, None


def link_traffic(link, direction, payload_bytes):
    """Bytes on the wire per launch from the committed ncu link-counter captures (profiles/r02_link_traffic.json)."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "r02_link_traffic.json")))[link][direction]
        return int(t["wire_ratio"] * payload_bytes)
    except Exception:
        return None


def peaks():
    try:
        p = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(p["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (of measured)"
    except Exception:
        return 6650.0, "B200_PROFILING.md fallback 6.65 TB/s (of fallback)"


# ----------------------------------------------------------------------------- clocks
class ClockSampler:
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100", "-i", str(self.index)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), [x.strip() for x in line.split(",")]))

    def stop(self, t0, t1):
        if self.proc:
            self.proc.terminate()
        rows = [r for t, r in self.rows if t0 <= t <= t1 and len(r) >= 6] or [r for _, r in self.rows if len(r) >= 6]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        mhz = [int(r[0]) for r in rows if r[0].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in rows for n, v in zip(names, r[2:6]) if v.lower().startswith("active")})
        return {"sm_mhz": statistics.median(mhz) if mhz else None, "sm_max_mhz": int(rows[0][1]) if rows[0][1].isdigit() else None,
                "reasons": reasons, "samples": len(rows)}


# ----------------------------------------------------------------------------- CPU legs (oracle)
def cpu_memtest_sample(gib, threads, passes_each=1):
    """The oracle's fill+verify (walking-ones pass + random pass) over a host buffer: the
    same math the GPU runs, on the host cores.  Returns (GB/s counted as 2N per pass, seconds)."""
    import numpy as np
    from oracle import cpu
    nwords = int(gib * GIB) // 8
    buf = np.zeros(nwords, dtype=np.uint64)  # touch pages before timing
    sec = 0.0
    bad = 0
    for pat in (cpu.PAT_WALKING_ONES, cpu.PAT_RANDOM):
        s, b = cpu.memtest_mt(buf, pat, 0, passes_each, 0, threads)
        sec += s
        bad += b
    assert bad == 0
    return 2.0 * nwords * 8 * 2 * passes_each / sec / 1e9, sec


def cpu_spill_sample(nchunks, threads, reps=3):
    """Host-memcpy spill: the oracle's copy + (X,S) checksum of 2 MiB chunks between two host buffers.
    Returns (best GB/s counted as payload once per chunk, seconds of the best repetition)."""
    import numpy as np
    from oracle import cpu
    cb = 2097152
    src = cpu.fill(nchunks * cb // 8, cpu.PAT_RANDOM, 0).copy()
    dst = np.zeros_like(src)
    best = 1e30
    for _ in range(reps):
        sec, _xs = cpu.spill_mt(src, dst, nchunks, cb, threads)
        best = min(best, sec)
    return nchunks * cb / best / 1e9, best


def run_reference(args, rank):
    """--impl reference: the reference has no CPU (or any) implementation of memtest or
    explicit spill — its vgpuvalidator is a license tool and its oversubscription is UVM
    paging inside the NVIDIA driver.  Per the task's tier rules the arm then
    times the oracle port of the path on all host cores, on a bounded sample per step: the
    memtest (walking-ones + random pass over a host buffer) and a host-memcpy spill + restore
    of 2 MiB chunks, the same mix as our arm's `e2e` step."""
    if rank != 0:
        return
    cores = len(os.sched_getaffinity(0))
    gib = args.cpu_gib
    nch = max(1, int(32 * gib / 4.0))
    for _ in range(max(1, min(args.warmup, 1))):
        cpu_memtest_sample(min(gib, 0.25), cores)
    t0 = time.perf_counter()
    vals, mt_vals, sp_vals = [], [], []
    for _ in range(args.steps):
        v_mt, s_mt = cpu_memtest_sample(gib, cores)
        v_sp, s_sp = cpu_spill_sample(nch, cores, reps=2)  # one out (spill), one back (restore): 2 copies of nch chunks
        nbytes = v_mt * s_mt * 1e9 + 2 * nch * 2097152
        vals.append(nbytes / (s_mt + 2 * s_sp) / 1e9)
        mt_vals.append(v_mt)
        sp_vals.append(v_sp)
    dt = time.perf_counter() - t0
    v = statistics.mean(vals)
    line = {
        "impl": "reference", "metric": METRIC, "value": round(v, 3), "unit": "GB/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": round(1e3 * dt / args.steps, 3), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": {"workload": "config 2 memtest (walking-ones + random, fill+verify, 2N bytes/pass) + host-memcpy spill/restore of 2 MiB chunks, CPU oracle port on host cores",
                   "sample": f"{gib} GiB host buffer per step instead of 4 x 40 GiB; {nch} chunks spilled and restored per step instead of {args.e2e_chunks}"},
        "cpu_baseline": {"value": round(v, 3), "unit": "GB/s", "cores": cores, "kind": "port",
                         "sample": f"{gib} GiB buffer, 1 walking-ones + 1 random pass, then {nch} x 2 MiB chunks memcpy+checksum out and back, per step, {args.steps} steps",
                         "memtest_GBps": round(statistics.mean(mt_vals), 3), "host_memcpy_spill_GBps": round(statistics.mean(sp_vals), 3)},
        "e2e": {"value": round(v, 3), "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------- ranks
class Ranks:
    """One process per GPU.  vGPUs are independent, so there is no data-path collective: the
    process group only carries the barrier around timed regions and the max (time) / sum (bytes)
    of per-rank scalars.  Works on `gloo` (CPU tests) and `nccl` alike."""

    def __init__(self, world, device=None):
        self.world, self.device = world, device

    def _sync(self):
        if self.device is not None and self.device.type == "cuda":
            import torch
            torch.cuda.synchronize()

    def barrier(self):
        self._sync()
        if self.world > 1:
            import torch.distributed as dist
            dist.barrier()
        self._sync()

    def _reduce(self, x, op):
        if self.world == 1:
            return float(x)
        import torch
        import torch.distributed as dist
        t = torch.tensor([x], dtype=torch.float64, device=self.device if self.device is not None else "cpu")
        dist.all_reduce(t, op=getattr(dist.ReduceOp, op))
        return float(t.item())

    def max(self, x):
        return self._reduce(x, "MAX")

    def min(self, x):
        return self._reduce(x, "MIN")

    def sum(self, x):
        return self._reduce(x, "SUM")


def ring_peer(local, world):
    """Spill tier pairing of config 5: GPU g spills to GPU (g+1) mod N."""
    return (local + 1) % world


def peer_of(local, world, ndev):
    """The GPU whose HBM is this rank's peer tier: the next one in the ring for N>1; with one rank, GPU 1 if the box
    shows a second (idle) GPU — the north star's "idle peer-GPU HBM" — else none."""
    if world > 1:
        return ring_peer(local, world)
    return 1 if ndev > 1 and local == 0 else None


def whole_job_gbps(bytes_per_rank_per_step, steps, world, max_ms):
    """`value`: units all ranks processed / max-over-ranks device time."""
    return world * bytes_per_rank_per_step * steps / (max_ms * 1e6)


def numpy_cksum(chunk_u8):
    """(X,S) of one chunk straight from its definition: X = xor of the u64 words, S = sum of
    w[i]*(2i+1) mod 2^64.  Written out here so the bench can cross-check the kernels without the oracle."""
    import numpy as np
    w = chunk_u8.view(np.uint64)
    k = 2 * np.arange(w.size, dtype=np.uint64) + np.uint64(1)
    with np.errstate(over="ignore"):
        return int(np.bitwise_xor.reduce(w)), int((w * k).sum(dtype=np.uint64))


def make_partitions(capi, torch, dev, local, nv, pct, kind):
    """-> (parts array or None, streams, description).  A failure to form green contexts is reported, not hidden."""
    parts = (capi.SmPartition * nv)()
    if kind == "green":
        rc = capi.lib.vgpu_sm_partition_create(local, pct, nv, parts)
    elif kind == "green-fine":
        rc = capi.lib.vgpu_sm_partition_create_ex(local, pct, nv, capi.SM_FINE, parts)
    else:
        rc = 1
    if rc == 0:
        streams = [torch.cuda.ExternalStream(p.stream, device=dev) for p in parts]
        desc = {"kind": "green_ctx" if kind == "green" else "green_ctx_sm_granular", "percent": pct,
                "sm_per_vgpu": [p.sm_count for p in parts], "sm_total": parts[0].sm_total}
        return parts, streams, desc
    streams = [torch.cuda.Stream(device=dev) for _ in range(nv)]
    return None, streams, {"kind": "plain_streams", "why": capi.last_error() if kind != "streams" else "--partition streams"}


# ----------------------------------------------------------------------------- GPU legs
def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank)
        return

    import torch
    import torch.distributed as dist

    import vgpu_b200  # noqa: F401
    from vgpu_b200 import capi, datapath as dp

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a GPU: the vGPU data path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    ranks = Ranks(world, dev)
    barrier, max_over_ranks = ranks.barrier, ranks.max
    peer = peer_of(local, world, torch.cuda.device_count())

    hbm_peak, peak_src = peaks()
    out_extra = {}

    # ---------------- phase A: spill / restore sweep (config 4; peer ring of config 5) --------
    sweep = None
    if not args.no_spill:
        sweep = spill_sweep(args, capi, dp, torch, dev, local, world, peer, ranks)
        out_extra["spill_restore"] = sweep

    if args.config4:
        out_extra["config4_vmem"] = config4_phase(args, capi, torch, local, world, ranks)
    if not args.no_config5:
        out_extra["config5"] = config5_phase(args, capi, dp, torch, dev, local, world, peer, ranks, hbm_peak)

    # ---------------- phase B: memtest (config 2) ---------------------------------------
    cb = capi.CHUNK_BYTES
    nv = args.vgpus
    # buffers of the e2e step's host-buffer calls (allocated first so the vGPU buffers are sized around them)
    en = args.e2e_chunks
    e_hbm = torch.empty(en * cb, dtype=torch.uint8, device=dev)
    e_back = torch.empty_like(e_hbm)
    e_host = torch.empty(en * cb, dtype=torch.uint8).pin_memory()
    dp.memtest_fill(e_hbm, capi.PAT_RANDOM, 4242 + local)
    free, total = torch.cuda.mem_get_info(dev)
    per = int(args.vgpu_gib * GIB)
    note = None
    if per * nv > free - 4 * GIB:
        per = max(GIB, ((free - 4 * GIB) // nv) // (2 << 20) * (2 << 20))
        note = f"scaled to {per / GIB:.2f} GiB per vGPU: only {free / GIB:.1f} GiB free"
    bufs = [torch.empty(per, dtype=torch.uint8, device=dev) for _ in range(nv)]

    pct = int(100 / nv)
    parts, streams, partition = make_partitions(capi, torch, dev, local, nv, pct, args.partition)
    results = [torch.zeros(3, dtype=torch.int64, device=dev) for _ in range(nv)]

    main_stream = torch.cuda.current_stream()
    KERNELS = (("fill_walking", capi.PAT_WALKING_ONES, "f"), ("verify_walking", capi.PAT_WALKING_ONES, "v"),
               ("fill_random", capi.PAT_RANDOM, "f"), ("verify_random", capi.PAT_RANDOM, "v"))

    def timed(nsteps, first, on=None, marks=None):
        """One step = walking-ones pass + random pass (fill, verify) on every vGPU stream.  `marks` collects CUDA events
        around every launch on vGPU 0's stream — the per-kernel durations INSIDE the timed region."""
        sts = on or streams
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ends = [torch.cuda.Event() for _ in range(nv)]
        ev0.record(main_stream)
        for st in sts:
            st.wait_event(ev0)
        for s in range(first, first + nsteps):
            for i in range(nv):
                for pat, p in ((capi.PAT_WALKING_ONES, s % 64), (capi.PAT_RANDOM, s)):
                    if marks is not None and i == 0:
                        e = torch.cuda.Event(enable_timing=True); e.record(sts[0]); marks.append(e)
                    dp.memtest_fill(bufs[i], pat, p, stream=sts[i])
                    if marks is not None and i == 0:
                        e = torch.cuda.Event(enable_timing=True); e.record(sts[0]); marks.append(e)
                    dp.memtest_verify(bufs[i], pat, p, result=results[i], stream=sts[i])
            if marks is not None:
                e = torch.cuda.Event(enable_timing=True); e.record(sts[0]); marks.append(e)
        for st, e in zip(sts, ends):
            e.record(st)
            main_stream.wait_event(e)
        ev1.record(main_stream)
        return ev0, ev1

    timed(args.warmup, 0)
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    time.sleep(0.3)
    barrier()
    t0 = time.perf_counter()
    marks = []
    ev0, ev1 = timed(args.steps, args.warmup, marks=marks)
    barrier()
    t1 = time.perf_counter()
    clocks = sampler.stop(t0, t1)
    ms = max_over_ranks(ev0.elapsed_time(ev1))
    bad = sum(int(r[0].item()) for r in results)
    if bad:
        raise SystemExit(f"memtest found {bad} mismatching words: not a valid bench run")
    bytes_per_step_rank = 2 * per * 2 * nv  # (fill+verify) x 2 patterns x vGPUs
    value = whole_job_gbps(bytes_per_step_rank, args.steps, world, ms)
    launches = args.steps * nv * 4
    # per-kernel durations inside the timed region (vGPU 0's stream, the other vGPUs busy beside it): 5 marks per step
    in_step = {}
    sm0 = partition.get("sm_per_vgpu", [None])[0]
    for k, (name, _p, _f) in enumerate(KERNELS):
        d = [marks[5 * s + k].elapsed_time(marks[5 * s + k + 1]) for s in range(args.steps)]
        kms = statistics.mean(d)
        in_step[name] = {"ms": round(kms, 4), "GBps": round(per / (kms * 1e6), 1)}
        if sm0:
            in_step[name]["GBps_per_SM"] = round(per / (kms * 1e6) / sm0, 1)
    # the same step without SM caps (plain streams sharing all SMs), for the record
    uncapped = None
    if parts is not None:
        plain = [torch.cuda.Stream(device=dev) for _ in range(nv)]
        timed(2, 0, plain)
        barrier()
        u0, u1 = timed(args.steps, args.warmup, plain)
        barrier()
        ums = max_over_ranks(u0.elapsed_time(u1))
        uncapped = {"GBps": round(whole_job_gbps(bytes_per_step_rank, args.steps, world, ums), 2), "ms_per_step": round(ums / args.steps, 4),
                    "note": "same buffers and kernels on plain streams: no green-context partition, all SMs shared"}

    # per-kernel durations, each timed alone on vGPU 0's stream (CUDA events on that stream)
    kern = {}
    st0 = streams[0]
    for name, pat, fn in KERNELS:
        dp.memtest_fill(bufs[0], pat, 1, stream=st0)
        e = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        for k in range(3):
            e[k].record(st0)
            if fn == "f":
                dp.memtest_fill(bufs[0], pat, 1, stream=st0)
            else:
                dp.memtest_verify(bufs[0], pat, 1, result=results[0], stream=st0)
        e[3].record(st0)
        torch.cuda.synchronize()
        kms = e[0].elapsed_time(e[3]) / 3
        kern[name] = {"ms": round(kms, 4), "GBps": round(per / (kms * 1e6), 1)}
        if sm0:
            kern[name]["GBps_per_SM"] = round(per / (kms * 1e6) / sm0, 1)
    # whole-GPU (no partition) numbers for the same four kernels, one vGPU buffer at a time
    whole = {}
    for name, pat, fn in KERNELS:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        dp.memtest_fill(bufs[0], pat, 1)
        e0.record()
        for k in range(3):
            if fn == "f":
                dp.memtest_fill(bufs[0], pat, 1)
            else:
                dp.memtest_verify(bufs[0], pat, 1, result=results[0])
        e1.record()
        torch.cuda.synchronize()
        kms = e0.elapsed_time(e1) / 3
        whole[name] = {"ms": round(kms, 4), "GBps": round(per / (kms * 1e6), 1)}
    dom = max(whole, key=lambda k: whole[k]["ms"])
    traffic, traffic_ratio = traffic_per_launch(dom, per)

    # e2e: the step through the blocking C-ABI calls a user makes, wall clock.  One host thread per vGPU runs
    # vgpu_memtest_run (result block read back to the host every pass); beside them one thread spills `en` chunks of
    # a resident HBM slice into a caller's HOST buffer and restores them (vgpu_spill_to_host / vgpu_restore_from_host:
    # payload + checksums cross PCIe inside the timed region, restored bytes are checksum-verified in the kernel).
    hck = (capi.Cksum * en)()
    e2e_bad = C.c_uint64(0)
    xfer_sec = {"spill": [], "restore": []}

    def e2e_mt_worker(i, s0, n):
        torch.cuda.set_device(local)
        for s in range(s0, s0 + n):
            for pat, p in ((capi.PAT_WALKING_ONES, s % 64), (capi.PAT_RANDOM, s)):
                r = dp.memtest_run(bufs[i], pat, p, 1, stream=streams[i])
                assert r.mismatches == 0

    e2e_v = {"auto": capi.XFER_AUTO, "ldst": capi.XFER_LDST, "tma": capi.XFER_TMA, "ce": capi.XFER_CE}[args.e2e_variant]

    def e2e_xfer_worker(n, record):
        torch.cuda.set_device(local)
        for _ in range(n):
            w = time.perf_counter()
            capi.check(capi.lib.vgpu_spill_to_host(e2e_v, local, e_hbm.data_ptr(), e_host.data_ptr(), en, cb, hck))
            w1 = time.perf_counter()
            capi.check(capi.lib.vgpu_restore_from_host(e2e_v, local, e_host.data_ptr(), e_back.data_ptr(), en, cb, hck, C.byref(e2e_bad)))
            w2 = time.perf_counter()
            if record:
                xfer_sec["spill"].append(w1 - w)
                xfer_sec["restore"].append(w2 - w1)

    def e2e_run(s0, n, record):
        th = [threading.Thread(target=e2e_mt_worker, args=(i, s0, n)) for i in range(nv)]
        th.append(threading.Thread(target=e2e_xfer_worker, args=(n, record)))
        [t.start() for t in th]
        [t.join() for t in th]

    e2e_run(0, 1, False)
    barrier()
    w0 = time.perf_counter()
    e2e_run(1, args.steps, True)
    barrier()
    e2e_ms = max_over_ranks((time.perf_counter() - w0) * 1e3)
    torch.cuda.synchronize()
    if e2e_bad.value or not torch.equal(e_back, e_hbm):
        raise SystemExit("e2e: bytes restored from the host buffer differ from what was spilled")
    xfer_bytes_step = 2 * en * cb
    e2e_value = whole_job_gbps(bytes_per_step_rank + xfer_bytes_step, args.steps, world, e2e_ms)
    e2e_components = {
        "memtest_GBps": round(whole_job_gbps(bytes_per_step_rank, args.steps, world, e2e_ms), 2),
        "host_spill_GBps_per_gpu": round(en * cb / statistics.mean(xfer_sec["spill"]) / 1e9, 2),
        "host_restore_GBps_per_gpu": round(en * cb / statistics.mean(xfer_sec["restore"]) / 1e9, 2),
        "note": "memtest rate = its bytes / the whole window; spill and restore = payload / their own blocking call, while the memtest runs",
    }

    line = None
    if rank == 0:
        cpu_base = None
        if not args.no_cpu and world == 1:  # the CPU baseline is a rank-0, N=1 figure
            cores = len(os.sched_getaffinity(0))
            v, sec = cpu_memtest_sample(args.cpu_gib, cores, passes_each=40)
            v1, _ = cpu_memtest_sample(min(args.cpu_gib, 0.5), 1)
            cpu_base = {"value": round(v, 3), "unit": "GB/s", "cores": cores, "kind": "port",
                        "sample": f"oracle memtest, {args.cpu_gib} GiB host buffer, 40 walking-ones + 40 random passes ({sec:.1f} s)",
                        "one_thread_GBps": round(v1, 3),
                        "host_memcpy_spill_GBps": {"1_thread": round(cpu_spill_sample(64, 1)[0], 2), f"{cores}_threads": round(cpu_spill_sample(256, cores)[0], 2)}}
        roof = {"bound": "hbm", "achieved": whole[dom]["GBps"], "peak": hbm_peak, "unit": "GB/s",
                "frac": round(whole[dom]["GBps"] / hbm_peak, 4), "traffic": traffic,
                "traffic_note": f"dram read+write bytes per {per / GIB:.0f} GiB launch = {traffic_ratio} x algorithmic bytes (ncu --set full, profiles/r02_traffic.json)",
                "algorithmic_bytes_per_launch": per, "kernel": f"k_memtest_{dom} (whole GPU, timed alone)",
                "peak_source": peak_src, "whole_gpu": whole, "one_partition": kern, "in_step_vgpu0": in_step,
                "step_aggregate": {"GBps": round(value / world, 1), "frac": round(value / world / hbm_peak, 4),
                                   "frac_of_8TBps_spec": round(value / world / 8000.0, 4)},
                "step_aggregate_uncapped": uncapped}
        if sweep is not None:
            roof["nvlink"] = sweep.get("_roofline_nvlink")
            roof["pcie"] = sweep.get("_roofline_pcie")
            sweep.pop("_roofline_nvlink", None)
            sweep.pop("_roofline_pcie", None)
        line = {
            "metric": METRIC, "value": round(value, 2), "unit": "GB/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": round(ms / args.steps, 4), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u64", "data": "synthetic",
            "config": {"workload": f"config 2: vgpuvalidator memtest, {nv} vGPUs x {per / GIB:.2f} GiB per B200, walking-ones + random, fill+verify, bit-exact"
                                   + ("; config 5 (10 vGPUs x 16 GiB per GPU + spill loop) in `config5`; config 4 chunk sweep in `spill_restore`" if not args.no_config5 else ""),
                       "bytes_counted": "2N per pass (N written by fill + N read by verify)", "l2": "inputs (>=1 GiB per vGPU) larger than the 126 MB L2",
                       "sm_partition": partition, "note": note},
            "clocks": clocks,
            "e2e": {"value": round(e2e_value, 2), "unit": "GB/s",
                    "h2d_bytes_per_step": en * cb + en * 16, "d2h_bytes_per_step": en * cb + en * 16 + 16 + nv * 2 * 24,
                    "call": "per step: vgpu_memtest_run on every vGPU (blocking C-ABI, one host thread per vGPU, result block read back each pass) "
                            f"while vgpu_spill_to_host + vgpu_restore_from_host move {en} x 2 MiB chunks between HBM and a caller's host buffer",
                    "bytes_counted": "memtest 2N per pass + spill payload + restore payload, over the wall-clock window (max over ranks)",
                    "components": e2e_components},
            "gpu_launches": launches,
            "roofline": roof,
            "cpu_baseline": cpu_base,
        }
        line.update(out_extra)
        print(json.dumps(line), flush=True)
    if parts is not None:
        torch.cuda.synchronize()
        capi.lib.vgpu_sm_partition_destroy(parts, nv)
    if world > 1:
        dist.destroy_process_group()


def config4_phase(args, capi, torch, local, world, ranks):
    """Config 4: one vGPU whose virtual range (quota + oversubscription) exceeds what may be resident;
    a write pass and a read-verify pass over the whole range in 8 GiB windows, paged by vgpu_vmem
    (LRU evict -> spill kernel -> unmap; map -> restore kernel -> checksum verify).  Rank 0 only."""
    out = None
    if ranks is not None:
        ranks.barrier()
    if local == 0:
        cb = capi.CHUNK_BYTES
        tier, peer = (capi.TIER_PEER, 1) if world > 1 else (capi.TIER_HOST, -1)
        scale = args.config4_scale if args.config4_scale else (1.0 if world > 1 else 0.2)
        free, _ = torch.cuda.mem_get_info()
        quota = int(180e9 * scale) // cb
        over = int(120e9 * scale) // cb
        resident = min(quota, (free - 6 * GIB) // cb)
        total = quota + over
        win = min(4096, resident // 2)
        total -= total % win
        h = C.c_void_p()
        capi.check(capi.lib.vgpu_vmem_create(local, total * cb, resident * cb, tier, peer, (total - resident + win) * cb, C.byref(h)))
        base = C.c_void_p()
        capi.check(capi.lib.vgpu_vmem_base(h, C.byref(base)))
        res = torch.zeros(3, dtype=torch.int64, device=torch.device("cuda", local))
        try:
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for w in range(total // win):
                capi.check(capi.lib.vgpu_vmem_ensure(h, w * win, win))
                capi.check(capi.lib.vgpu_memtest_fill(base.value + w * win * cb, win * cb, capi.PAT_RANDOM, w, 4, None))
            torch.cuda.synchronize()
            t1 = time.perf_counter()
            for w in range(total // win):
                capi.check(capi.lib.vgpu_vmem_ensure(h, w * win, win))
                capi.check(capi.lib.vgpu_memtest_verify(base.value + w * win * cb, win * cb, capi.PAT_RANDOM, w, 4, res.data_ptr(), None))
            torch.cuda.synchronize()
            t2 = time.perf_counter()
            st = capi.VmemStats()
            capi.check(capi.lib.vgpu_vmem_stats(h, C.byref(st)))
            bad = res.cpu().tolist()
            if bad != [0, 0, 0] or st.checksum_failures:
                raise SystemExit(f"config 4: restored data is not bit-exact: {bad}, checksum failures {st.checksum_failures}")
            out = {"tier": "peer HBM (GPU 1) over NVLink" if world > 1 else "pinned host over PCIe", "scale": scale,
                   "virtual_GB": round(total * cb / 1e9, 1), "resident_limit_GB": round(resident * cb / 1e9, 1), "window_chunks": win,
                   "chunks_spilled_total": st.evictions, "chunks_restored_total": st.fetches,
                   "spill_GBps_incl_unmap": round(st.evictions * cb / ((st.evict_ms - st.tier_setup_ms) * 1e6), 1) if st.evict_ms > st.tier_setup_ms else None,
                   "tier_setup_ms_once": round(st.tier_setup_ms, 1),
                   "spill_GBps_incl_unmap_and_setup": round(st.evictions * cb / (st.evict_ms * 1e6), 1) if st.evict_ms else None,
                   "restore_GBps_incl_map": round(st.fetches * cb / (st.fetch_ms * 1e6), 1) if st.fetch_ms and st.fetches else None,
                   "write_pass_GBps": round(total * cb / (t1 - t0) / 1e9, 1), "verify_pass_GBps": round(total * cb / (t2 - t1) / 1e9, 1),
                   "bit_exact": True}
        finally:
            capi.lib.vgpu_vmem_destroy(h)
            torch.cuda.empty_cache()
    if ranks is not None:
        ranks.barrier()
    return out


def config5_phase(args, capi, dp, torch, dev, local, world, peer, ranks, hbm_peak):
    """Config 5 on every GPU at once: 10 vGPUs (10 % SM partitions) x 16 GiB memtest, concurrent with a
    spill/restore loop on its own stream (whole GPU) to the next GPU in the ring, or to pinned host when there is no peer."""
    cb = capi.CHUNK_BYTES
    nv, per = 10, int(args.config5_gib * GIB)
    free, _ = torch.cuda.mem_get_info(dev)
    if nv * per > free - 12 * GIB:
        per = ((free - 12 * GIB) // nv) // cb * cb
    bufs = [torch.empty(per, dtype=torch.uint8, device=dev) for _ in range(nv)]
    parts, streams, partition = make_partitions(capi, torch, dev, local, nv, 10, args.partition)
    results = [torch.zeros(3, dtype=torch.int64, device=dev) for _ in range(nv)]
    n = min(1024, args.spill_chunks)
    hbm = torch.empty(n * cb, dtype=torch.uint8, device=dev)
    back = torch.empty_like(hbm)
    dp.memtest_fill(hbm, capi.PAT_RANDOM, 77)
    ck = torch.zeros((n, 2), dtype=torch.int64, device=dev)
    rres = torch.zeros(2, dtype=torch.int64, device=dev)
    slab = C.c_void_p()
    tier, pd = (capi.TIER_PEER, peer) if peer is not None else (capi.TIER_HOST, -1)
    capi.check(capi.lib.vgpu_slab_alloc(tier, local, pd, n * cb, C.byref(slab)))
    if args.peer_grid:
        capi.check(capi.lib.vgpu_policy_set_grid(capi.TIER_PEER, args.peer_grid))
    xs = torch.cuda.Stream(device=dev, priority=-1 if args.xfer_priority == "high" else 0)
    mts = torch.cuda.Stream(device=dev)  # carries only the "all memtests done" event
    xstream = C.c_void_p(xs.cuda_stream)
    main = torch.cuda.current_stream()
    reps_mt, reps_x = 2, (48 if tier == capi.TIER_PEER else 3)

    def run(do_mt, do_x):
        ranks.barrier()
        e0, e1, m1 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        ends = []
        e0.record(main)
        if do_mt:
            for i, st in enumerate(streams):
                st.wait_event(e0)
            for s in range(reps_mt):  # pass by pass across the vGPUs, as ten validator processes would run side by side
                for pat, p in ((capi.PAT_WALKING_ONES, s), (capi.PAT_RANDOM, s)):
                    for i, st in enumerate(streams):
                        dp.memtest_fill(bufs[i], pat, p, stream=st)
                        dp.memtest_verify(bufs[i], pat, p, result=results[i], stream=st)
            for st in streams:
                e = torch.cuda.Event(); e.record(st); ends.append(e)
                mts.wait_event(e)
            m1.record(mts)  # the moment the last vGPU's memtest ended, whatever the transfer loop is still doing
        ex0, ex1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        if do_x:
            xs.wait_event(e0)
            ex0.record(xs)
            for _ in range(reps_x):
                capi.check(capi.lib.vgpu_spill(capi.XFER_AUTO, tier, hbm.data_ptr(), slab, None, None, n, cb, ck.data_ptr(), xstream))
                capi.check(capi.lib.vgpu_restore(capi.XFER_AUTO, tier, slab, back.data_ptr(), None, None, n, cb, ck.data_ptr(), None, rres.data_ptr(), xstream))
            ex1.record(xs)
            e = torch.cuda.Event(); e.record(xs); ends.append(e)
        for e in ends:
            main.wait_event(e)
        e1.record(main)
        torch.cuda.synchronize()
        return ((ranks.max(e0.elapsed_time(m1)) if do_mt else None), (ranks.max(ex0.elapsed_time(ex1)) if do_x else None),
                ranks.max(e0.elapsed_time(e1)))

    run(True, True)  # warm-up
    ms_mt_in_both, ms_x_in_both, ms_both = run(True, True)
    ms_mt, _, _ = run(True, False)
    _, ms_x, _ = run(False, True)
    ok = all(r.cpu().tolist() == [0, 0, 0] for r in results) and rres.cpu().tolist() == [0, 0] and torch.equal(back, hbm)
    if not ok:
        raise SystemExit("config 5: memtest mismatch or spill/restore round trip not bit-exact")
    mt_bytes = world * nv * per * 2 * 2 * reps_mt
    x_bytes = world * n * cb * 2 * reps_x
    mt_alone = mt_bytes / (ms_mt * 1e6)
    out = {"workload": f"config 5: {nv} vGPUs x {per / GIB:.2f} GiB memtest per GPU (walking-ones + random, fill+verify) concurrent with a spill/restore loop",
           "vgpus_per_gpu": nv, "vgpus_total": nv * world, "vgpu_GiB": round(per / GIB, 2), "sm_partition": partition,
           "spill_tier": (f"peer GPU over NVLink ({'ring g->g+1' if world > 1 else 'idle GPU 1'})" if tier == capi.TIER_PEER else "pinned host over PCIe"),
           "spill_batch_chunks": n,
           "memtest_alone_GBps": round(mt_alone, 1), "memtest_alone_GBps_per_gpu": round(mt_alone / world, 1),
           "memtest_alone_frac_of_hbm_peak": round(mt_alone / world / hbm_peak, 4), "memtest_alone_frac_of_8TBps_spec": round(mt_alone / world / 8000.0, 4),
           "spill_restore_alone_GBps": round(x_bytes / (ms_x * 1e6), 1), "spill_restore_alone_GBps_per_gpu": round(x_bytes / (ms_x * 1e6) / world, 1),
           "concurrent": {"window_ms": round(ms_both, 2), "memtest_ms": round(ms_mt_in_both, 2), "spill_restore_ms": round(ms_x_in_both, 2),
                          "memtest_GBps": round(mt_bytes / (ms_mt_in_both * 1e6), 1),
                          "memtest_GBps_per_gpu": round(mt_bytes / (ms_mt_in_both * 1e6) / world, 1),
                          "memtest_frac_of_hbm_peak": round(mt_bytes / (ms_mt_in_both * 1e6) / world / hbm_peak, 4),
                          "spill_plus_restore_GBps": round(x_bytes / (ms_x_in_both * 1e6), 1),
                          "spill_plus_restore_GBps_per_gpu": round(x_bytes / (ms_x_in_both * 1e6) / world, 1),
                          "note": "both start together; each rate = its bytes / the time from the common start to its own end (max over ranks)"},
           "timing": "CUDA events bracketing all partition streams, max over ranks; aggregate over all GPUs unless marked per_gpu",
           "bit_exact": True}
    torch.cuda.synchronize()
    if args.peer_grid:
        capi.lib.vgpu_policy_set_grid(capi.TIER_PEER, 0)
    if parts is not None:
        capi.lib.vgpu_sm_partition_destroy(parts, nv)
    capi.lib.vgpu_slab_free(tier, local, pd, slab)
    del bufs, hbm, back
    torch.cuda.empty_cache()
    return out


def spill_sweep(args, capi, dp, torch, dev, local, world, peer, ranks):
    """Config 4 sweep: batches of 1..4096 x 2 MiB chunks, per tier and kernel variant, device-timed
    with CUDA events (max over ranks, summed over GPUs).  Restores verify checksums in-kernel; per tier the
    checksums and the bytes that landed in the tier are checked against the local tier's (which the GPU tests
    pin to the oracle) and against the checksum's definition — `oracle_checked`."""
    import numpy as np
    barrier, max_over_ranks = ranks.barrier, ranks.max
    cb = capi.CHUNK_BYTES
    nmax = args.spill_chunks
    sizes = [n for n in (1, 8, 64, 512, 4096) if n <= nmax]
    variants = {"ldst": capi.XFER_LDST, "tma": capi.XFER_TMA, "ce": capi.XFER_CE}
    hbm = torch.empty(nmax * cb, dtype=torch.uint8, device=dev)
    dp.memtest_fill(hbm, capi.PAT_RANDOM, 1000 + local)
    back = torch.empty_like(hbm)
    ck = torch.zeros((nmax, 2), dtype=torch.int64, device=dev)
    res = torch.zeros(2, dtype=torch.int64, device=dev)
    tiers = {}
    local_slab = torch.empty_like(hbm)
    tiers["local"] = (capi.TIER_LOCAL, local_slab.data_ptr(), None)
    host = torch.empty(min(nmax, 1024) * cb, dtype=torch.uint8).pin_memory()
    tiers["host"] = (capi.TIER_HOST, host.data_ptr(), min(nmax, 1024))
    peer_slab = C.c_void_p()
    if peer is not None:
        rc = capi.lib.vgpu_slab_alloc(capi.TIER_PEER, local, peer, nmax * cb, C.byref(peer_slab))
        if rc != 0:
            raise SystemExit("peer slab: " + capi.last_error())
        tiers["peer"] = (capi.TIER_PEER, peer_slab.value, None)
    out = {"chunk_bytes": cb, "bytes_counted": "payload once per chunk per direction", "timing": "CUDA events on the launch stream, best of 5 (20 for batches of up to 64 chunks) after 2 warm-ups, max over ranks"}
    if "peer" in tiers and world == 1:
        # An idle peer (nobody else's context on it): let the library measure its variants on this box, as a deployment
        # does at start-up — "switch ... when measured bandwidth favours it".  With the peer idle the PULL form of the
        # spill (the bulk-copy kernel runs on the peer and READS our HBM: read responses carry 12.5 % packet overhead,
        # writes 18.75 %) wins from 64 chunks up.  With N>1 every GPU is busy in another process and the built-in
        # table (push) stays.
        capi.check(capi.lib.vgpu_calibrate(local, capi.TIER_PEER, peer, min(nmax, 512)))
    names = {capi.XFER_LDST: "ldst", capi.XFER_TMA: "tma", capi.XFER_CE: "ce", capi.XFER_PULL: "pull"}
    out["auto_policy"] = {t: {str(n): names.get(capi.lib.vgpu_policy_get(tiers[t][0], n), "?") for n in sizes} for t in tiers}
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)

    def time_call(fn, reps=5, warm=2):
        for _ in range(warm):
            fn()
        best = 1e30
        for _ in range(reps):
            barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            best = min(best, max_over_ranks(e0.elapsed_time(e1)))
        return best

    # reference checksums of the source: the local tier's LD/ST spill (pinned to the oracle by tests/test_datapath_gpu.py)
    ck_ref = torch.zeros_like(ck)
    capi.check(capi.lib.vgpu_spill(capi.XFER_LDST, capi.TIER_LOCAL, hbm.data_ptr(), local_slab.data_ptr(), None, None, nmax, cb, ck_ref.data_ptr(), stream))
    torch.cuda.synchronize()
    sample = sorted({0, nmax // 3, nmax - 1})
    hbm_sample = {c: hbm[c * cb:(c + 1) * cb].cpu().numpy() for c in sample}
    ck_ref_host = ck_ref.cpu().numpy().view(np.uint64)
    for c in sample:  # the checksum's definition, evaluated with numpy on a few chunks
        if numpy_cksum(hbm_sample[c]) != (int(ck_ref_host[c, 0]), int(ck_ref_host[c, 1])):
            raise SystemExit(f"chunk {c}: kernel checksum differs from the (X,S) definition")

    best_at_max = {}
    for tname, (tier, slab, cap) in tiers.items():
        tout = {}
        tvariants = dict(variants, pull=capi.XFER_PULL) if tname == "peer" else variants
        for n in sizes:
            if cap is not None and n > cap:
                continue
            row = {}
            hp, bp, ckp, rp = hbm.data_ptr(), back.data_ptr(), ck.data_ptr(), res.data_ptr()   # not inside the timed call
            reps = 20 if n <= 64 else 5   # a small batch is a ~10 us event: more samples for the minimum
            for vname, v in tvariants.items():
                def do_spill():
                    capi.check(capi.lib.vgpu_spill(v, tier, hp, slab, None, None, n, cb, ckp, stream))

                def do_restore():
                    capi.check(capi.lib.vgpu_restore(v, tier, slab, bp, None, None, n, cb, ckp, None, rp, stream))
                ms_s = time_call(do_spill, reps=reps)
                ms_r = time_call(do_restore, reps=reps)
                row[vname] = {"spill_GBps": round(world * n * cb / (ms_s * 1e6), 1), "restore_GBps": round(world * n * cb / (ms_r * 1e6), 1)}
            tout[str(n)] = row
        n = max(int(k) for k in tout)
        best_at_max[tname] = (n, {d: max(tout[str(n)][v][d + "_GBps"] for v in tout[str(n)]) / world for d in ("spill", "restore")})
        # parity of this tier on its largest batch, outside the timed region: every variant's checksums == the local
        # tier's; the bytes that landed in the tier, read back by the copy engine, == the source; restore -> bit-exact
        checked = True
        for vname, v in tvariants.items():
            ck.zero_()
            capi.check(capi.lib.vgpu_spill(v, tier, hbm.data_ptr(), slab, None, None, n, cb, ck.data_ptr(), stream))
            torch.cuda.synchronize()
            checked &= bool(torch.equal(ck[:n], ck_ref[:n]))
            back[: n * cb].zero_()
            res.zero_()
            capi.check(capi.lib.vgpu_restore(capi.XFER_CE, tier, slab, back.data_ptr(), None, None, n, cb, ck_ref.data_ptr(), None, res.data_ptr(), stream))
            torch.cuda.synchronize()
            checked &= res.cpu().tolist() == [0, 0] and bool(torch.equal(back[: n * cb], hbm[: n * cb]))
            if vname != "pull":
                back[: n * cb].zero_()
                capi.check(capi.lib.vgpu_restore(v, tier, slab, back.data_ptr(), None, None, n, cb, ck_ref.data_ptr(), ck.data_ptr(), res.data_ptr(), stream))
                torch.cuda.synchronize()
                checked &= res.cpu().tolist() == [0, 0] and bool(torch.equal(back[: n * cb], hbm[: n * cb])) and bool(torch.equal(ck[:n], ck_ref[:n]))
        if not checked:
            raise SystemExit(f"spill/restore on tier {tname}: checksums or bytes differ from the oracle-pinned local tier")
        tout["round_trip_bit_exact"] = True
        tout["oracle_checked"] = True
        tout["oracle_checked_how"] = (f"{n} chunks, every variant: checksums == local-tier LD/ST checksums (oracle-pinned by tests/test_datapath_gpu.py) "
                                      "and == the (X,S) definition in numpy on sampled chunks; tier bytes read back by the copy engine == source; restore bit-exact")
        out[tname] = tout
    # one direction at a time: only rank 0 moves data over its link, the other GPUs are idle
    if "peer" in tiers:
        tier, slab, _ = tiers["peer"]
        n = max(sz for sz in sizes)
        uni = {}
        for vname, v in dict(variants, pull=capi.XFER_PULL).items():
            res_ms = []
            for direction in ("spill", "restore"):
                best = 1e30
                for rep in range(5):
                    barrier()
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    if local == 0:
                        e0.record()
                        if direction == "spill":
                            capi.check(capi.lib.vgpu_spill(v, tier, hbm.data_ptr(), slab, None, None, n, cb, ck.data_ptr(), stream))
                        else:
                            capi.check(capi.lib.vgpu_restore(v, tier, slab, back.data_ptr(), None, None, n, cb, ck.data_ptr(), None, res.data_ptr(), stream))
                        e1.record()
                        torch.cuda.synchronize()
                        if rep >= 1:
                            best = min(best, e0.elapsed_time(e1))
                    barrier()
                res_ms.append(best)
            if local == 0:
                uni[vname] = {"spill_GBps": round(n * cb / (res_ms[0] * 1e6), 1), "restore_GBps": round(n * cb / (res_ms[1] * 1e6), 1)}
        out["peer"]["one_direction_only_rank0"] = {"chunks": n, **uni}

    # copy-engine rate of each link, measured in this run (rank 0 alone, the other GPUs idle): the practical peak
    link = {}
    for tname in ("peer", "host"):
        if tname not in tiers:
            continue
        barrier()
        o, i = C.c_double(0), C.c_double(0)
        if local == 0:
            capi.check(capi.lib.vgpu_link_rate(local, tiers[tname][0], peer if tname == "peer" else -1, min(nmax, 1024) * cb, C.byref(o), C.byref(i)))
            link[tname] = {"out_GBps": round(o.value, 1), "in_GBps": round(i.value, 1)}
        barrier()
    out["link_rate_copy_engine"] = dict(link, how="vgpu_link_rate: cudaMemcpyAsync of 2 GiB out and back, best of 3, rank 0 alone")

    # roofline entries of the two links (rank 0's view; achieved = best kernel variant at the largest batch, one GPU)
    if local == 0:
        if "peer" in tiers:
            n = out["peer"]["one_direction_only_rank0"]["chunks"]
            uni = {k: v for k, v in out["peer"]["one_direction_only_rank0"].items() if k in ("ldst", "tma", "pull")}
            sp = max(v["spill_GBps"] for v in uni.values())
            rs = max(v["restore_GBps"] for v in uni.values())
            ring = best_at_max["peer"][1]
            pk = link.get("peer", {})
            out["_roofline_nvlink"] = {
                "bound": "nvlink", "unit": "GB/s", "peak": NVLINK_NOMINAL, "peak_source": "900 GB/s per direction nominal (north_star)",
                "kernel": "k_xfer_ldst / k_xfer_tma (pushed from this GPU, or — spill only — pulled by the idle peer), best variant, spill slab in the peer GPU's HBM", "chunks_per_launch": n,
                "algorithmic_bytes_per_launch": n * cb,
                "spill": {"achieved": sp, "frac": round(sp / NVLINK_NOMINAL, 4), "frac_of_copy_engine": round(sp / pk["out_GBps"], 4) if pk.get("out_GBps") else None,
                          "traffic": link_traffic("nvlink", "spill", n * cb)},
                "restore": {"achieved": rs, "frac": round(rs / NVLINK_NOMINAL, 4), "frac_of_copy_engine": round(rs / pk["in_GBps"], 4) if pk.get("in_GBps") else None,
                            "traffic": link_traffic("nvlink", "restore", n * cb)},
                "achieved": sp, "frac": round(sp / NVLINK_NOMINAL, 4),
                "measured_copy_engine_peak": pk or None,
                "all_gpus_at_once_per_gpu": {"spill": round(ring["spill"], 1), "restore": round(ring["restore"], 1)} if world > 1 else None,
                "traffic_note": "wire bytes per launch = payload x the ncu nvltx/nvlrx ratio of profiles/r02_link_traffic.json (packet overhead: 18.75 % on writes, 12.5 % on read responses)"}
        else:
            rec = None
            try:  # what the same code measured when a second GPU was visible, committed with the repository
                src = "profiles/r02_bench_n1_idle_peer.json"
                rec = dict(json.load(open(os.path.join(ROOT, src)))["roofline"]["nvlink"], source=src + " (N=1 on a 2-GPU box, GPU 1 idle)")
            except Exception:
                pass
            out["_roofline_nvlink"] = {"bound": "nvlink", "achieved": None, "peak": NVLINK_NOMINAL, "unit": "GB/s",
                                       "why": "one GPU visible: no peer tier can be measured in this run (the N>1 lines carry it)",
                                       "last_recorded": rec}
        n, h = best_at_max["host"]
        pk = link.get("host", {})
        out["_roofline_pcie"] = {
            "bound": "pcie", "unit": "GB/s", "peak": PCIE_NOMINAL, "peak_source": "PCIe Gen5 x16 raw, 64 GB/s per direction (north_star)",
            "kernel": "best variant at the largest batch (copy engine + checksum kernel, or the LD/ST / bulk-copy kernels storing straight to pinned host memory)",
            "chunks_per_launch": n, "algorithmic_bytes_per_launch": n * cb,
            "spill": {"achieved": round(h["spill"], 1), "frac": round(h["spill"] / PCIE_NOMINAL, 4),
                      "frac_of_copy_engine": round(h["spill"] / pk["out_GBps"], 4) if pk.get("out_GBps") else None, "traffic": link_traffic("pcie", "spill", n * cb)},
            "restore": {"achieved": round(h["restore"], 1), "frac": round(h["restore"] / PCIE_NOMINAL, 4),
                        "frac_of_copy_engine": round(h["restore"] / pk["in_GBps"], 4) if pk.get("in_GBps") else None, "traffic": link_traffic("pcie", "restore", n * cb)},
            "achieved": round(h["spill"], 1), "frac": round(h["spill"] / PCIE_NOMINAL, 4),
            "measured_copy_engine_peak": pk or None,
            "note": "per GPU with every rank spilling at once" if world > 1 else None,
            "traffic_note": "wire bytes per launch = payload x the ncu pcie ratio of profiles/r02_link_traffic.json (TLP overhead 12.5 % / 9.4 %)"}

    # e2e host-buffer entry points (blocking, wall clock): vgpu_spill_to_host / vgpu_restore_from_host
    n = min(nmax, 1024)
    hck = (capi.Cksum * n)()
    bad = C.c_uint64(0)
    e2e = {}
    for vname, v in variants.items():
        for _ in range(2):
            capi.check(capi.lib.vgpu_spill_to_host(v, local, hbm.data_ptr(), host.data_ptr(), n, cb, hck))
        barrier()
        w0 = time.perf_counter()
        for _ in range(3):
            capi.check(capi.lib.vgpu_spill_to_host(v, local, hbm.data_ptr(), host.data_ptr(), n, cb, hck))
        ws = max_over_ranks((time.perf_counter() - w0) / 3)
        barrier()
        w0 = time.perf_counter()
        for _ in range(3):
            capi.check(capi.lib.vgpu_restore_from_host(v, local, host.data_ptr(), back.data_ptr(), n, cb, hck, C.byref(bad)))
        wr = max_over_ranks((time.perf_counter() - w0) / 3)
        e2e[vname] = {"spill_GBps": round(world * n * cb / ws / 1e9, 2), "restore_GBps": round(world * n * cb / wr / 1e9, 2)}
    out["host"]["e2e_host_buffer_calls"] = {"chunks": n, "d2h_bytes_per_call": n * cb + n * 16, "h2d_bytes_per_call": n * cb + n * 16, **e2e}
    if peer_slab.value:
        barrier()
        capi.lib.vgpu_slab_free(capi.TIER_PEER, local, peer, peer_slab)
    del hbm, back, local_slab, host
    torch.cuda.empty_cache()
    return out


if __name__ == "__main__":
    main()
