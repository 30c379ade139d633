"""One launch (after one warm-up) of each hot kernel on cuda:0, for `ncu --set full`.
Sizes are larger than L2 (126 MB).  Run plain first; see scripts/gpu_profile.sh."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import vgpu_b200  # noqa: E402,F401
from vgpu_b200 import capi, datapath as dp  # noqa: E402

dev = torch.device("cuda", 0)
gib = float(sys.argv[1]) if len(sys.argv) > 1 else 4.0
buf = torch.empty(int(gib * (1 << 30)), dtype=torch.uint8, device=dev)
res = torch.zeros(3, dtype=torch.int64, device=dev)
for rep in range(2):  # rep 0 = warm-up
    for pat in (capi.PAT_WALKING_ONES, capi.PAT_RANDOM):
        dp.memtest_fill(buf, pat, 1)
        dp.memtest_verify(buf, pat, 1, result=res)
n = min(512, buf.numel() // (2 * capi.CHUNK_BYTES))
src, dst = buf[: n * capi.CHUNK_BYTES], buf[n * capi.CHUNK_BYTES: 2 * n * capi.CHUNK_BYTES]
for rep in range(2):
    for v in (capi.XFER_LDST, capi.XFER_TMA):
        ck = dp.spill(src, dst, n, variant=v)
ck2 = dp.checksum(src, n)
torch.cuda.synchronize()
assert res.cpu().tolist() == [0, 0, 0] and torch.equal(ck, ck2)
print("prof_kernels ok")
