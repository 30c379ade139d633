"""Peer-tier spill and restore from ONE process (kernels run on GPU 0, the slab lives in GPU 1's HBM),
a few launches, for an ncu pass with the nvlink counters.  Needs 2 GPUs."""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import vgpu_b200  # noqa: E402,F401
from vgpu_b200 import capi, datapath as dp  # noqa: E402

dev = torch.device("cuda", 0)
n, cb = 1024, capi.CHUNK_BYTES
hbm = torch.empty(n * cb, dtype=torch.uint8, device=dev)
back = torch.zeros_like(hbm)
dp.memtest_fill(hbm, capi.PAT_RANDOM, 2)
slab = C.c_void_p()
capi.check(capi.lib.vgpu_slab_alloc(capi.TIER_PEER, 0, 1, n * cb, C.byref(slab)))
ck = torch.zeros((n, 2), dtype=torch.int64, device=dev)
res = torch.zeros(2, dtype=torch.int64, device=dev)
with torch.cuda.device(0):
    for rep in range(3):
        for v in (capi.XFER_LDST, capi.XFER_TMA):
            capi.check(capi.lib.vgpu_spill(v, capi.TIER_PEER, hbm.data_ptr(), slab, None, None, n, cb, ck.data_ptr(), None))
            capi.check(capi.lib.vgpu_restore(v, capi.TIER_PEER, slab, back.data_ptr(), None, None, n, cb, ck.data_ptr(), None, res.data_ptr(), None))
    torch.cuda.synchronize()
assert res.cpu().tolist() == [0, 0] and torch.equal(back, hbm)
capi.lib.vgpu_slab_free(capi.TIER_PEER, 0, 1, slab)
print("prof_peer_tier ok")
