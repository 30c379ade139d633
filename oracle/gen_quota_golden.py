#!/usr/bin/env python
"""Captures golden quota vectors FROM THE REFERENCE BINARY (/root/reference/vgpu/libvgpu.so)
run under the mock driver, and writes them to tests/golden/quota_*.json.gz.  Run in the build
container (the reference is not available on the GPU box); the fixtures are committed.

  python oracle/gen_quota_golden.py
"""
import gzip
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import quota_harness as H  # noqa: E402

MAP = {"NVIDIA_DEVICE_MAP": "0:" + H.UUID0}
MAP2 = {"NVIDIA_DEVICE_MAP": f"0:{H.UUID0} 1:{H.UUID0[:-1]}2", "MOCK_GPU_COUNT": "2"}
CTX = {"MOCK_CTX_MIB": "300"}  # the mock NVML lists the process once it holds a context: context billing is live
OUT = os.path.join(ROOT, "tests", "golden")

CASES = [
    # name, trace, env, nprocs, mock physical MiB
    ("c3_golden", H.golden_trace(), {"CUDA_DEVICE_MEMORY_LIMIT_0": "1024m", "CUDA_DEVICE_SM_LIMIT": "10", **MAP}, 1),
    ("c3_golden_oversub", H.golden_trace(), {"CUDA_DEVICE_MEMORY_LIMIT_0": "1024m", "CUDA_DEVICE_SM_LIMIT": "10", "CUDA_OVERSUBSCRIBE": "true", **MAP}, 1),
    ("appA_nolimit", H.appendix_a_trace(), {}, 1),
    ("appA_1g", H.appendix_a_trace(), {"CUDA_DEVICE_MEMORY_LIMIT": "1g"}, 1),
    ("appA_perdev_wins", H.appendix_a_trace(), {"CUDA_DEVICE_MEMORY_LIMIT_0": "1048576k", "CUDA_DEVICE_MEMORY_LIMIT": "2g"}, 1),
    ("appA_nosuffix", H.appendix_a_trace(), {"CUDA_DEVICE_MEMORY_LIMIT_0": "1073741824"}, 1),
    ("appA_hex", H.appendix_a_trace(), {"CUDA_DEVICE_MEMORY_LIMIT_0": "0x40000000"}, 1),
    ("appA_50m", H.appendix_a_trace(), {"CUDA_DEVICE_MEMORY_LIMIT_0": "50m"}, 1),
    ("appA_overflow", H.appendix_a_trace(), {"CUDA_DEVICE_MEMORY_LIMIT_0": "17179869184g"}, 1),
    ("appA_garbage", H.appendix_a_trace(), {"CUDA_DEVICE_MEMORY_LIMIT_0": "abc"}, 1),
    ("sm_limits_per_device", H.golden_trace(), {"CUDA_DEVICE_MEMORY_LIMIT_0": "1024m", "CUDA_DEVICE_SM_LIMIT": "20", "CUDA_DEVICE_SM_LIMIT_0": "30",
                                                "CUDA_DEVICE_SM_LIMIT_3": "0", "CUDA_TASK_PRIORITY": "0", **MAP}, 1),
    ("random_5EED_1p", H.random_trace(0x5EED, 10000), {"CUDA_DEVICE_MEMORY_LIMIT_0": "18432m", "CUDA_DEVICE_SM_LIMIT": "10", **MAP}, 1),
    ("random_5EED_1p_oversub", H.random_trace(0x5EED, 10000), {"CUDA_DEVICE_MEMORY_LIMIT_0": "18432m", "CUDA_DEVICE_SM_LIMIT": "10", "CUDA_OVERSUBSCRIBE": "true", **MAP}, 1),
    ("random_5EED_4p", H.random_trace(0x5EED + 1, 10000, 4), {"CUDA_DEVICE_MEMORY_LIMIT_0": "18432m", "CUDA_DEVICE_SM_LIMIT": "10", **MAP}, 4),
    ("random_5EED_4p_oversub_tight", H.random_trace(0x5EED + 2, 4000, 4, 1 << 26), {"CUDA_DEVICE_MEMORY_LIMIT_0": "300m", "CUDA_OVERSUBSCRIBE": "true", **MAP}, 4),
    # context billing (SURVEY "usage includes context_size"): what a real node shows and the plain mock hides
    ("ctx_one_device", H.ctx_trace(), {"CUDA_DEVICE_MEMORY_LIMIT_0": "1g", **CTX, **MAP}, 1),
    ("ctx_create_first", H.ctx_create_first_trace(), {"CUDA_DEVICE_MEMORY_LIMIT_0": "1g", **CTX, **MAP}, 1),
    ("ctx_two_devices", H.ctx_two_devices_trace(), {"CUDA_DEVICE_MEMORY_LIMIT_0": "1g", "CUDA_DEVICE_MEMORY_LIMIT_1": "2g", **CTX, **MAP2}, 1),
    ("ctx_four_procs", H.ctx_procs_trace(4), {"CUDA_DEVICE_MEMORY_LIMIT_0": "1g", **CTX, **MAP}, 4),
    ("ctx_random_4p", H.with_contexts(H.random_trace(0x5EED + 3, 3000, 4, 1 << 27), 4),
     {"CUDA_DEVICE_MEMORY_LIMIT_0": "4g", "MOCK_CTX_MIB": "517", **MAP}, 4),
    ("ctx_random_4p_oversub", H.with_contexts(H.random_trace(0x5EED + 4, 3000, 4, 1 << 27), 4),
     {"CUDA_DEVICE_MEMORY_LIMIT_0": "3g", "MOCK_CTX_MIB": "517", "CUDA_OVERSUBSCRIBE": "true", **MAP}, 4),
    # processes that die without their exit handler: swept by the OOM path (rm_quitted_process@0x425b6)
    ("dead_process", H.dead_process_trace(), {"CUDA_DEVICE_MEMORY_LIMIT_0": "1g", **MAP}, 3),
    ("dead_process_ctx", H.with_contexts(H.dead_process_trace(), 3), {"CUDA_DEVICE_MEMORY_LIMIT_0": "2g", **CTX, **MAP}, 3),
    ("dead_random_6p", H.with_kills(H.random_trace(0x5EED + 5, 4000, 6, 1 << 27), 0xD1E, 6, 3), {"CUDA_DEVICE_MEMORY_LIMIT_0": "2g", **MAP}, 6),
    ("fork_children", H.fork_trace(), {"CUDA_DEVICE_MEMORY_LIMIT_0": "1g", **MAP}, 1),
    ("late_joiner", H.late_joiner_trace(), {"CUDA_DEVICE_MEMORY_LIMIT_0": "1g", **MAP}, 5),
    # virtual devices: NVIDIA_DEVICE_MAP with a GPU named twice / a subset / another order
    ("remap_shared_gpu", H.remap_shared_gpu_trace(), {"CUDA_DEVICE_MEMORY_LIMIT_0": "1g", "CUDA_DEVICE_MEMORY_LIMIT_1": "2g", **CTX,
                                                    "NVIDIA_DEVICE_MAP": f"0:{H.UUID0} 1:{H.UUID0}"}, 1),
    ("remap_three_on_two", H.remap_three_on_two_trace(), {"CUDA_DEVICE_MEMORY_LIMIT_0": "1g", "CUDA_DEVICE_MEMORY_LIMIT_1": "2g",
                                                        "CUDA_DEVICE_MEMORY_LIMIT_2": "3g", **CTX, "MOCK_GPU_COUNT": "2",
                                                        "NVIDIA_DEVICE_MAP": f"0:{H.UUID0} 1:{H.UUID0[:-1]}2 2:{H.UUID0}"}, 1),
    ("remap_subset_queries", H.remap_queries_trace(1), {"CUDA_DEVICE_MEMORY_LIMIT_0": "50m", "MOCK_GPU_COUNT": "2",
                                                       "NVIDIA_DEVICE_MAP": f"0:{H.UUID0[:-1]}2"}, 1),
    ("remap_swapped_queries", H.remap_queries_trace(2), {"CUDA_DEVICE_MEMORY_LIMIT_0": "50m", "CUDA_DEVICE_MEMORY_LIMIT_1": "60m",
                                                        "MOCK_GPU_COUNT": "2", "NVIDIA_DEVICE_MAP": f"0:{H.UUID0[:-1]}2 1:{H.UUID0}"}, 1),
    ("remap_shared_queries", H.remap_queries_trace(2), {"CUDA_DEVICE_MEMORY_LIMIT_0": "50m", "CUDA_DEVICE_MEMORY_LIMIT_1": "60m",
                                                       "NVIDIA_DEVICE_MAP": f"0:{H.UUID0} 1:{H.UUID0}"}, 1),
]
# Parts of the final cache file that are left out of a vector.  Under the mock the reference's NVML handle remap
# answers device 0 for every mock handle, so with several GPUs it writes device 0's UUID into the other uuids[] lines
# (and, for a map that names only the second GPU, into uuids[0]); for a GPU named twice it writes a made-up UUID for
# the second entry.  The shim writes the real UUID of each virtual device, so the byte-level checksum is not compared.
REGION_SKIP = {"ctx_two_devices": ("masked_crc",), "remap_shared_gpu": ("masked_crc",), "remap_three_on_two": ("masked_crc",),
               "remap_subset_queries": ("masked_crc", "uuid0"), "remap_swapped_queries": ("masked_crc",),
               "remap_shared_queries": ("masked_crc",)}


def main():
    if not H.have_reference():
        sys.exit("reference binary not present: golden vectors can only be captured where /root/reference exists")
    H.build()
    os.makedirs(OUT, exist_ok=True)
    for name, trace, env, nprocs in CASES:
        env = {"MOCK_TOTAL_MIB": "184320", **env}
        r = H.run_trace(H.REFERENCE_LIB, trace, env, nprocs)
        if name in REGION_SKIP:
            r["region_summary"] = {k: v for k, v in H.region_summary(r["region"]).items() if k not in REGION_SKIP[name]}
        doc = {"name": name, "source": "vgpu/libvgpu.so (sha256 d48e2b4c...c0c1ff1) under oracle/ref_harness mocks",
               "env": env, "nprocs": nprocs, "trace": trace, "replies": r["replies"], "oom": r["oom"], "calls": r["calls"],
               "region": r.get("region_summary") or H.region_summary(r["region"])}
        path = os.path.join(OUT, f"quota_{name}.json.gz")
        with gzip.GzipFile(path, "wb", mtime=0) as f:
            f.write(json.dumps(doc, separators=(",", ":")).encode())
        print(f"{name}: {len(trace)} ops, {len(r['oom'])} OOMs, {os.path.getsize(path)} bytes")


if __name__ == "__main__":
    main()
