"""N>1 host logic of bench.py on CPU: two ranks over gloo (world_size 2).  The path shards by GPU
with no data-path collective; what the ranks share is the barrier and the
max-of-times / sum-of-bytes reductions that turn per-rank numbers into the whole-job value."""
import os
import socket
import subprocess
import sys
import textwrap

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = textwrap.dedent("""
    import json, os, sys
    sys.path.insert(0, %r)
    import torch.distributed as dist
    import bench
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dist.init_process_group("gloo", rank=rank, world_size=world)
    r = bench.Ranks(world, None)
    r.barrier()
    ms = 100.0 + 50.0 * rank                      # rank 1 is the slow one
    mx = r.max(ms)
    total = r.sum(1000.0 * (rank + 1))
    v = bench.whole_job_gbps(bytes_per_rank_per_step=2e9, steps=4, world=world, max_ms=mx)
    r.barrier()
    print(json.dumps({"rank": rank, "max_ms": mx, "sum": total, "value": v, "peer": bench.ring_peer(rank, world)}), flush=True)
    dist.destroy_process_group()
""") % ROOT


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_two_ranks_agree_on_max_time_and_whole_job_value():
    import json
    port = free_port()
    procs = []
    for rank in range(2):
        env = {**os.environ, "RANK": str(rank), "WORLD_SIZE": "2", "MASTER_ADDR": "127.0.0.1", "MASTER_PORT": str(port)}
        procs.append(subprocess.Popen([sys.executable, "-c", WORKER], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True))
    outs = []
    for p in procs:
        o, e = p.communicate(timeout=120)
        assert p.returncode == 0, e[-2000:]
        outs.append(json.loads(o.strip().splitlines()[-1]))
    for o in outs:
        assert o["max_ms"] == 150.0 and o["sum"] == 3000.0         # both ranks see the slowest rank's time
        assert abs(o["value"] - 2 * 2e9 * 4 / (150.0 * 1e6)) < 1e-9  # all ranks' bytes / max time
    assert sorted(o["peer"] for o in outs) == [0, 1] and outs[0]["peer"] == 1  # ring g -> g+1


def test_reference_arm_runs_on_rank0_only():
    """bench.py --impl reference: rank 0 prints the line, other ranks exit 0 without work."""
    import json
    env0 = {**os.environ, "RANK": "0", "WORLD_SIZE": "2"}
    env1 = {**os.environ, "RANK": "1", "WORLD_SIZE": "2"}
    args = [sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "1", "--cpu-gib", "0.05"]
    p1 = subprocess.run(args, env=env1, capture_output=True, text=True, timeout=120)
    assert p1.returncode == 0 and p1.stdout.strip() == ""
    p0 = subprocess.run(args, env=env0, capture_output=True, text=True, timeout=120)
    assert p0.returncode == 0, p0.stderr[-2000:]
    line = json.loads(p0.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["n_gpus"] == 2 and line["value"] > 0
    assert line["cpu_baseline"]["kind"] == "port" and line["e2e"]["h2d_bytes_per_step"] == 0
    assert line["unit"] == "GB/s" and line["higher_is_better"] is True
