"""bench.py keeps its contract: one JSON line with the keys the driver reads, at a small size."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_bench_json_line_contract(cuda_dev):
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "2", "--warmup", "3", "--vgpu-gib", "2",
                        "--spill-chunks", "64", "--cpu-gib", "0.25", "--e2e-chunks", "64", "--config5-gib", "1"], capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stderr[-3000:]
    lines = [x for x in p.stdout.strip().splitlines() if x.startswith("{")]
    assert len(lines) == 1, "exactly one JSON line on stdout"
    d = json.loads(lines[0])
    assert d["metric"].startswith("vGPU spill/restore GB/s") and d["unit"] == "GB/s" and d["value"] > 100
    assert d["n_gpus"] == 1 and d["steps"] == 2 and d["warmup"] == 3 and d["ms_per_step"] > 0
    assert d["higher_is_better"] is True and d["scaling"] == "weak" and d["vs_baseline"] is None
    assert d["dtype"] == "u64" and d["data"] == "synthetic" and "workload" in d["config"] and "model" not in d["config"]
    assert d["gpu_launches"] == 2 * 4 * 4
    c = d["clocks"]
    assert set(c) >= {"sm_mhz", "sm_max_mhz", "reasons"} and not ({"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"} & set(c["reasons"]))
    e = d["e2e"]
    n_e2e = 64
    assert e["value"] > 100 and e["unit"] == "GB/s"
    assert e["h2d_bytes_per_step"] == n_e2e * 2097152 + n_e2e * 16  # restore payload + expected checksums
    assert e["d2h_bytes_per_step"] == n_e2e * 2097152 + n_e2e * 16 + 16 + 4 * 2 * 24  # spill payload + checksums + result blocks
    assert e["components"]["host_spill_GBps_per_gpu"] > 1 and e["components"]["host_restore_GBps_per_gpu"] > 1
    r = d["roofline"]
    assert r["bound"] == "hbm" and r["unit"] == "GB/s" and r["peak"] > 1000 and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-3
    assert r["traffic"] and 0.9 < r["traffic"] / r["algorithmic_bytes_per_launch"] < 1.1
    assert set(r["in_step_vgpu0"]) == {"fill_walking", "verify_walking", "fill_random", "verify_random"}
    pc = r["pcie"]
    assert pc["bound"] == "pcie" and pc["peak"] == 64.0 and 0 < pc["spill"]["frac"] < 1.2 and pc["spill"]["traffic"] > pc["algorithmic_bytes_per_launch"]
    assert pc["measured_copy_engine_peak"]["out_GBps"] > 1
    assert r["nvlink"]["bound"] == "nvlink"  # numbers only when a second GPU is visible
    b = d["cpu_baseline"]
    assert b["kind"] == "port" and b["cores"] >= 1 and b["value"] > 0 and b["unit"] == "GB/s" and "sample" in b
    s = d["spill_restore"]
    assert s["local"]["round_trip_bit_exact"] and s["host"]["round_trip_bit_exact"]
    assert s["local"]["oracle_checked"] and s["host"]["oracle_checked"]
    assert s["local"]["64"]["ldst"]["spill_GBps"] > 100 and s["host"]["64"]["ce"]["restore_GBps"] > 5
    c5 = d["config5"]
    assert c5["vgpus_per_gpu"] == 10 and c5["bit_exact"] and c5["memtest_alone_GBps"] > 100 and sum(c5["sm_partition"]["sm_per_vgpu"]) <= c5["sm_partition"]["sm_total"]
