"""Host-tier spill and restore (LD/ST variant: the kernel itself stores to / loads from pinned host
memory over PCIe), a few launches, for an ncu pass with the pcie__* metrics."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import vgpu_b200  # noqa: E402,F401
from vgpu_b200 import capi, datapath as dp  # noqa: E402

dev = torch.device("cuda", 0)
n, cb = 256, capi.CHUNK_BYTES
hbm = torch.empty(n * cb, dtype=torch.uint8, device=dev)
back = torch.zeros_like(hbm)
host = torch.empty(n * cb, dtype=torch.uint8).pin_memory()
dp.memtest_fill(hbm, capi.PAT_RANDOM, 2)
for rep in range(3):
    ck = dp.spill(hbm, host, n, variant=capi.XFER_LDST, tier=capi.TIER_HOST)
    res = dp.restore(host, back, n, ck, variant=capi.XFER_LDST, tier=capi.TIER_HOST)
torch.cuda.synchronize()
assert res.cpu().tolist() == [0, 0] and torch.equal(back, hbm)
print("prof_host_tier ok")
