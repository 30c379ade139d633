"""vgpu_spill_noalloc — the spill entry point for callers that cannot allocate device memory
(a vGPU whose HBM slice is full; code inside the preload shim's allocation hooks).

Collected last on purpose (file name): if it fails, `pytest -x` has already run every other
test.  It is checked against the existing entry point vgpu_spill(VGPU_XFER_LDST) — same
kernel, scratch obtained the ordinary way — bit for bit (integer work, no tolerance), on two
calls in a row that use different batch sizes and data, so that what the first call leaves in
the shared fixed scratch cannot leak into the second; both calls are also compared directly
with the CPU oracle's checksums and bytes.  Passes alone in a fresh process (the
first call does the per-GPU set-up) and after the rest of the suite."""
import ctypes as C

import numpy as np
import pytest
import torch

from oracle import cpu  # the checker

pytestmark = pytest.mark.gpu


def _oracle_spill(src, n, cb, slots=None):
    """(checksums [n,2] u64, destination bytes) the CPU oracle produces for the same source."""
    words = src.cpu().numpy().view(np.uint64)
    o_dst = np.zeros_like(words)
    want = cpu.spill(words, o_dst, n, cb, *(slots or ()))
    return want, o_dst.view(np.uint8)


def _spill_both_ways(capi, src, n, cb, slots=None):
    dev = src.device
    ss = ds = None
    if slots is not None:
        ss = torch.from_numpy(slots[0].view(np.int32)).to(dev)
        ds = torch.from_numpy(slots[1].view(np.int32)).to(dev)
    nslots = src.numel() // cb
    outs = []
    for fn in ("ref", "noalloc"):
        dst = torch.zeros(nslots * cb, dtype=torch.uint8, device=dev)
        ck = torch.zeros((n, 2), dtype=torch.int64, device=dev)
        sp, dp_ = (ss.data_ptr() if ss is not None else None), (ds.data_ptr() if ds is not None else None)
        if fn == "ref":
            rc = capi.lib.vgpu_spill(capi.XFER_LDST, capi.TIER_LOCAL, src.data_ptr(), dst.data_ptr(), sp, dp_, n, cb, ck.data_ptr(), None)
        else:
            rc = capi.lib.vgpu_spill_noalloc(src.data_ptr(), dst.data_ptr(), sp, dp_, n, cb, ck.data_ptr(), None)
        assert rc == 0, capi.last_error()
        torch.cuda.synchronize()
        outs.append((dst, ck))
    return outs


def test_noalloc_spill_matches_existing_entry_point_twice_in_a_row(vgpu, cuda_dev):
    capi, dp = vgpu.capi, vgpu.datapath
    cb = capi.CHUNK_BYTES
    # call 1: 40 contiguous chunks
    src1 = torch.empty(40 * cb, dtype=torch.uint8, device=cuda_dev)
    dp.memtest_fill(src1, capi.PAT_RANDOM, 11)
    (d_ref, ck_ref), (d_new, ck_new) = _spill_both_ways(capi, src1, 40, cb)
    assert torch.equal(ck_new, ck_ref), "checksums differ from vgpu_spill (call 1)"
    assert torch.equal(d_new, d_ref) and torch.equal(d_new, src1)
    want, o_dst = _oracle_spill(src1, 40, cb)
    assert np.array_equal(ck_new.cpu().numpy().view(np.uint64), want), "checksums differ from the oracle (call 1)"
    assert np.array_equal(d_new.cpu().numpy(), o_dst)
    # call 2, straight after: fewer chunks, other data, ragged chunk size, scattered slots
    cb2 = 65536 + 32
    rng = np.random.default_rng(2)
    slots = (rng.permutation(24)[:9].astype(np.uint32), rng.permutation(24)[:9].astype(np.uint32))
    src2 = torch.empty(24 * cb2, dtype=torch.uint8, device=cuda_dev)
    dp.memtest_fill(src2, capi.PAT_RANDOM, 12)
    (d_ref, ck_ref), (d_new, ck_new) = _spill_both_ways(capi, src2, 9, cb2, slots)
    assert torch.equal(ck_new, ck_ref), "checksums differ from vgpu_spill (call 2)"
    assert torch.equal(d_new, d_ref)
    want, o_dst = _oracle_spill(src2, 9, cb2, slots)
    assert np.array_equal(ck_new.cpu().numpy().view(np.uint64), want), "checksums differ from the oracle (call 2)"
    assert np.array_equal(d_new.cpu().numpy(), o_dst)
    # and the full capacity of the fixed scratch
    n = capi.NOALLOC_MAX_CHUNKS
    src3 = torch.empty(n * 4096, dtype=torch.uint8, device=cuda_dev)
    dp.memtest_fill(src3, capi.PAT_RANDOM, 13)
    (d_ref, ck_ref), (d_new, ck_new) = _spill_both_ways(capi, src3, n, 4096)
    assert torch.equal(ck_new, ck_ref) and torch.equal(d_new, d_ref)


def test_noalloc_spill_refuses_batches_beyond_its_fixed_scratch(vgpu, cuda_dev):
    capi = vgpu.capi
    a = torch.zeros(1 << 20, dtype=torch.uint8, device=cuda_dev)
    ck = torch.zeros((8, 2), dtype=torch.int64, device=cuda_dev)
    rc = capi.lib.vgpu_spill_noalloc(a.data_ptr(), a.data_ptr(), None, None, capi.NOALLOC_MAX_CHUNKS + 1, 32, ck.data_ptr(), None)
    assert rc == capi.E_RANGE
    assert "exceed the fixed scratch of 4096 chunks" in capi.last_error()
    assert capi.lib.vgpu_spill_noalloc(a.data_ptr(), a.data_ptr(), None, None, 0, 32, None, None) == 0  # empty batch: no-op
    assert capi.lib.vgpu_spill_noalloc(a.data_ptr() + 8, a.data_ptr(), None, None, 1, 32, ck.data_ptr(), None) == capi.E_INVAL


def test_noalloc_spill_to_pinned_host_memory(vgpu, cuda_dev):
    """The usual destination of an eviction made from inside a hook is the host tier: pinned, mapped host memory.  The
    entry point then launches the small grid a PCIe-bound transfer needs (it asks where `dst` lives; it still allocates
    nothing).  Bytes and checksums against the oracle."""
    capi, dp = vgpu.capi, vgpu.datapath
    cb, n = capi.CHUNK_BYTES, 24
    src = torch.empty(n * cb, dtype=torch.uint8, device=cuda_dev)
    dp.memtest_fill(src, capi.PAT_RANDOM, 21)
    host = torch.zeros(n * cb, dtype=torch.uint8).pin_memory()
    ck = torch.zeros((n, 2), dtype=torch.int64, device=cuda_dev)
    rc = capi.lib.vgpu_spill_noalloc(src.data_ptr(), host.data_ptr(), None, None, n, cb, ck.data_ptr(), None)
    assert rc == 0, capi.last_error()
    torch.cuda.synchronize()
    want, o_dst = _oracle_spill(src, n, cb)
    assert np.array_equal(ck.cpu().numpy().view(np.uint64), want)
    assert np.array_equal(host.numpy(), o_dst)
