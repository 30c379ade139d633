"""Virtual device memory manager on a real GPU: a virtual range larger than its resident HBM,
paged through the host tier by the spill/restore kernels; every byte must survive (config 4 in
miniature — the full-size run is in bench.py)."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu
CB = 2097152


def _mk(capi, total, resident, tier_slots, tier=None, block_chunks=None):
    import os
    if block_chunks:   # residency is tracked per block (default 128 chunks = 256 MiB); small tests use small blocks
        os.environ["VGPU_VMEM_BLOCK_CHUNKS"] = str(block_chunks)
    else:
        os.environ.pop("VGPU_VMEM_BLOCK_CHUNKS", None)
    h = C.c_void_p()
    rc = capi.lib.vgpu_vmem_create(0, total * CB, resident * CB, capi.TIER_HOST if tier is None else tier, -1, tier_slots * CB, C.byref(h))
    assert rc == 0, capi.last_error()
    base = C.c_void_p()
    assert capi.lib.vgpu_vmem_base(h, C.byref(base)) == 0
    return h, base.value


def _stats(capi, h):
    s = capi.VmemStats()
    assert capi.lib.vgpu_vmem_stats(h, C.byref(s)) == 0
    return s


def test_oversubscribed_range_round_trips_every_byte(vgpu, cuda_dev):
    capi = vgpu.capi
    total, resident, win = 256, 64, 32
    h, base = _mk(capi, total, resident, total, block_chunks=8)
    res = torch.zeros(3, dtype=torch.int64, device=cuda_dev)
    try:
        # write every window; the manager evicts LRU windows to the host tier as we go
        for w in range(total // win):
            assert capi.lib.vgpu_vmem_ensure(h, w * win, win) == 0, capi.last_error()
            assert capi.lib.vgpu_memtest_fill(base + w * win * CB, win * CB, capi.PAT_RANDOM, w, 99, None) == 0
        torch.cuda.synchronize()
        s = _stats(capi, h)
        assert s.chunks_total == total and s.chunks_resident == resident and s.chunks_spilled == total - resident
        assert s.evictions == total - resident and s.fetches == 0
        # read everything back (twice, in a different order the second time): the memtest verify
        # kernel regenerates each window's pattern; zero mismatches == restored bytes are bit-exact
        for order in (range(total // win), reversed(range(total // win))):
            for w in order:
                assert capi.lib.vgpu_vmem_ensure(h, w * win, win) == 0, capi.last_error()
                assert capi.lib.vgpu_memtest_verify(base + w * win * CB, win * CB, capi.PAT_RANDOM, w, 99, res.data_ptr(), None) == 0
        torch.cuda.synchronize()
        assert res.cpu().tolist() == [0, 0, 0]
        s = _stats(capi, h)
        assert s.checksum_failures == 0 and s.fetches >= total - resident and s.chunks_resident <= resident
    finally:
        capi.lib.vgpu_vmem_destroy(h)


def test_explicit_evict_gives_the_hbm_back_and_fetch_restores(vgpu, cuda_dev):
    capi = vgpu.capi
    n = 512  # 1 GiB
    h, base = _mk(capi, n, n, n)
    res = torch.zeros(3, dtype=torch.int64, device=cuda_dev)
    try:
        torch.cuda.synchronize()
        free0, _ = torch.cuda.mem_get_info()
        assert capi.lib.vgpu_vmem_ensure(h, 0, n) == 0, capi.last_error()
        assert capi.lib.vgpu_memtest_fill(base, n * CB, capi.PAT_RANDOM, 5, 0, None) == 0
        torch.cuda.synchronize()
        free1, _ = torch.cuda.mem_get_info()
        assert free0 - free1 >= n * CB                       # the range is backed by real HBM
        assert capi.lib.vgpu_vmem_evict(h, 0, n) == 0, capi.last_error()   # "suspend": everything out
        free2, _ = torch.cuda.mem_get_info()
        assert free2 - free1 >= n * CB - (8 << 20)           # ... and the HBM is back with the driver (minus first-use scratch)
        s = _stats(capi, h)
        assert s.chunks_resident == 0 and s.chunks_spilled == n
        assert capi.lib.vgpu_vmem_ensure(h, 0, n) == 0, capi.last_error()  # "resume"
        assert capi.lib.vgpu_memtest_verify(base, n * CB, capi.PAT_RANDOM, 5, 0, res.data_ptr(), None) == 0
        torch.cuda.synchronize()
        assert res.cpu().tolist() == [0, 0, 0]
        assert _stats(capi, h).checksum_failures == 0
    finally:
        capi.lib.vgpu_vmem_destroy(h)


def test_limits_are_reported_not_papered_over(vgpu, cuda_dev):
    capi = vgpu.capi
    h, base = _mk(capi, 64, 16, 8, block_chunks=8)   # tier smaller than what would have to be spilled
    try:
        assert capi.lib.vgpu_vmem_ensure(h, 0, 17) == capi.E_NOMEM          # more than may be resident
        assert capi.lib.vgpu_vmem_ensure(h, 0, 16) == 0
        assert capi.lib.vgpu_vmem_ensure(h, 16, 8) == 0                      # evicts 8 -> tier now full
        assert capi.lib.vgpu_vmem_ensure(h, 24, 8) == capi.E_NOMEM          # nowhere to put 8 more victims
        assert "tier full" in capi.last_error()
        assert capi.lib.vgpu_vmem_ensure(h, 60, 8) == capi.E_INVAL          # beyond the range
    finally:
        capi.lib.vgpu_vmem_destroy(h)


def test_two_tiers_peer_first_then_host(vgpu, cuda_dev):
    """north_star tier order: victims go to the idle peer GPU's HBM over NVLink while it has room,
    then to pinned host RAM; everything comes back bit-exact from whichever tier holds it."""
    if torch.cuda.device_count() < 2:
        pytest.skip("peer tier needs 2 GPUs")
    import os
    capi = vgpu.capi
    os.environ["VGPU_VMEM_BLOCK_CHUNKS"] = "8"
    total, resident, peer, win = 512, 128, 128, 64     # chunks; host takes what the peer cannot
    h = C.c_void_p()
    rc = capi.lib.vgpu_vmem_create_tiered(0, total * CB, resident * CB, 1, peer * CB, total * CB, C.byref(h))
    os.environ.pop("VGPU_VMEM_BLOCK_CHUNKS", None)
    assert rc == 0, capi.last_error()
    base = C.c_void_p()
    assert capi.lib.vgpu_vmem_base(h, C.byref(base)) == 0
    res = torch.zeros(3, dtype=torch.int64, device=cuda_dev)
    try:
        for w in range(total // win):
            assert capi.lib.vgpu_vmem_ensure(h, w * win, win) == 0, capi.last_error()
            assert capi.lib.vgpu_memtest_fill(base.value + w * win * CB, win * CB, capi.PAT_RANDOM, w, 7, None) == 0
        torch.cuda.synchronize()
        s = _stats(capi, h)
        assert s.evictions_peer == peer and s.evictions_host == total - resident - peer   # peer filled first, rest to host
        for w in reversed(range(total // win)):
            assert capi.lib.vgpu_vmem_ensure(h, w * win, win) == 0, capi.last_error()
            assert capi.lib.vgpu_memtest_verify(base.value + w * win * CB, win * CB, capi.PAT_RANDOM, w, 7, res.data_ptr(), None) == 0
        torch.cuda.synchronize()
        assert res.cpu().tolist() == [0, 0, 0] and _stats(capi, h).checksum_failures == 0
    finally:
        capi.lib.vgpu_vmem_destroy(h)


@pytest.mark.parametrize("pull", [False, True])
def test_peer_tier_spills_pushed_or_pulled_by_the_idle_peer(vgpu, cuda_dev, pull):
    """With an idle peer the policy may select the PULL spill (the kernel runs on the peer and reads our HBM: 788 instead
    of 717 GB/s); the manager then gives the peer access to the range.  A range created under the push policy is never
    pulled, even if the policy changes afterwards (the peer has no access to it)."""
    if torch.cuda.device_count() < 2:
        pytest.skip("peer tier needs 2 GPUs")
    import os
    capi = vgpu.capi
    before = capi.lib.vgpu_policy_get(capi.TIER_PEER, 8)
    os.environ["VGPU_VMEM_BLOCK_CHUNKS"] = "8"
    total = 64
    h = C.c_void_p()
    try:
        if pull:
            assert capi.lib.vgpu_policy_set(capi.TIER_PEER, 0, capi.XFER_PULL) == 0
        rc = capi.lib.vgpu_vmem_create_tiered(0, total * CB, total * CB, 1, total * CB, 0, C.byref(h))
        assert rc == 0, capi.last_error()
        if not pull:                                   # the policy turns to PULL after the range exists
            assert capi.lib.vgpu_policy_set(capi.TIER_PEER, 0, capi.XFER_PULL) == 0
        base = C.c_void_p()
        assert capi.lib.vgpu_vmem_base(h, C.byref(base)) == 0
        res = torch.zeros(3, dtype=torch.int64, device=cuda_dev)
        assert capi.lib.vgpu_vmem_ensure(h, 0, total) == 0, capi.last_error()
        assert capi.lib.vgpu_memtest_fill(base.value, total * CB, capi.PAT_RANDOM, 9, 3, None) == 0
        torch.cuda.synchronize()
        assert capi.lib.vgpu_vmem_evict(h, 0, total) == 0, capi.last_error()
        tier, ptr = C.c_int(-1), C.c_void_p()
        assert capi.lib.vgpu_vmem_locate(h, 5, C.byref(tier), C.byref(ptr)) == 0 and tier.value == capi.TIER_PEER
        assert capi.lib.vgpu_vmem_ensure(h, 0, total) == 0, capi.last_error()
        assert capi.lib.vgpu_memtest_verify(base.value, total * CB, capi.PAT_RANDOM, 9, 3, res.data_ptr(), None) == 0
        torch.cuda.synchronize()
        s = _stats(capi, h)
        assert res.cpu().tolist() == [0, 0, 0] and s.checksum_failures == 0 and s.evictions_peer == total
    finally:
        os.environ.pop("VGPU_VMEM_BLOCK_CHUNKS", None)
        capi.lib.vgpu_policy_set(capi.TIER_PEER, 0, before)
        if h:
            capi.lib.vgpu_vmem_destroy(h)


def test_failed_restore_keeps_the_tier_copy_and_a_later_ensure_succeeds(vgpu, cuda_dev):
    """A chunk that comes back from the tier with the wrong checksum must not cost the data: the block is restored once
    more through the copy engine, and if that fails too it is unmapped again and stays spilled (slot and checksums
    kept).  Here the tier slab itself is corrupted, so both attempts fail; after the slab is repaired a second ensure
    succeeds and every byte is back."""
    capi = vgpu.capi
    total, bc = 64, 8
    h, base = _mk(capi, total, total, total, block_chunks=bc)
    res = torch.zeros(3, dtype=torch.int64, device=cuda_dev)
    try:
        assert capi.lib.vgpu_vmem_ensure(h, 0, total) == 0, capi.last_error()
        assert capi.lib.vgpu_memtest_fill(base, total * CB, capi.PAT_RANDOM, 3, 21, None) == 0
        torch.cuda.synchronize()
        assert capi.lib.vgpu_vmem_evict(h, 0, total) == 0, capi.last_error()
        victim = 2 * bc + 5                                   # a chunk of block 2
        tier, ptr = C.c_int(-1), C.c_void_p()
        assert capi.lib.vgpu_vmem_locate(h, victim, C.byref(tier), C.byref(ptr)) == 0
        assert tier.value == capi.TIER_HOST and ptr.value
        byte = (C.c_ubyte * 1).from_address(ptr.value + 12345)  # the host tier is pinned host memory: reachable from here
        byte[0] ^= 0x10
        rc = capi.lib.vgpu_vmem_ensure(h, 0, total)
        assert rc == capi.E_MISMATCH, capi.last_error()
        assert f"first: chunk {victim}" in capi.last_error() and "stays in the tier" in capi.last_error()
        s = _stats(capi, h)
        assert s.checksum_failures == 1 and s.restore_retries == 1          # one chunk, still bad after the copy-engine retry
        assert s.chunks_resident == total - bc and s.chunks_spilled == bc   # only the failing block stayed out
        t2, p2 = C.c_int(-1), C.c_void_p()
        assert capi.lib.vgpu_vmem_locate(h, victim, C.byref(t2), C.byref(p2)) == 0
        assert t2.value == capi.TIER_HOST and p2.value == ptr.value         # same slot: the only good copy was not given away
        byte[0] ^= 0x10                                                      # repair
        assert capi.lib.vgpu_vmem_ensure(h, 0, total) == 0, capi.last_error()
        assert capi.lib.vgpu_memtest_verify(base, total * CB, capi.PAT_RANDOM, 3, 21, res.data_ptr(), None) == 0
        torch.cuda.synchronize()
        assert res.cpu().tolist() == [0, 0, 0]
        s = _stats(capi, h)
        assert s.chunks_resident == total and s.chunks_spilled == 0 and s.checksum_failures == 1
        # and the freed slots are all usable again
        assert capi.lib.vgpu_vmem_evict(h, 0, total) == 0, capi.last_error()
        assert _stats(capi, h).chunks_spilled == total
    finally:
        capi.lib.vgpu_vmem_destroy(h)


def test_manager_leaves_the_callers_context_current(vgpu, cuda_dev):
    """The manager is called on application threads from inside the shim's hooks; a thread whose current context is not
    the primary one (cuCtxCreate, a green context) must find it current again afterwards."""
    capi = vgpu.capi
    cuda = C.CDLL("libcuda.so.1")
    parts = (capi.SmPartition * 2)()
    assert capi.lib.vgpu_sm_partition_create(0, 25, 2, parts) == 0, capi.last_error()
    h = C.c_void_p()
    try:
        assert cuda.cuCtxPushCurrent_v2(C.c_void_p(parts[1].ctx)) == 0
        cur = C.c_void_p()
        h, base = _mk(capi, 32, 16, 32, block_chunks=8)
        assert cuda.cuCtxGetCurrent(C.byref(cur)) == 0 and cur.value == parts[1].ctx
        assert capi.lib.vgpu_vmem_ensure(h, 0, 16) == 0, capi.last_error()
        assert capi.lib.vgpu_vmem_ensure(h, 16, 16) == 0, capi.last_error()   # evicts
        assert cuda.cuCtxGetCurrent(C.byref(cur)) == 0 and cur.value == parts[1].ctx
        assert capi.lib.vgpu_vmem_evict(h, 0, 32) == 0, capi.last_error()
        assert cuda.cuCtxGetCurrent(C.byref(cur)) == 0 and cur.value == parts[1].ctx
        assert capi.lib.vgpu_vmem_destroy(h) == 0
        h = C.c_void_p()
        assert cuda.cuCtxGetCurrent(C.byref(cur)) == 0 and cur.value == parts[1].ctx
    finally:
        popped = C.c_void_p()
        cuda.cuCtxPopCurrent_v2(C.byref(popped))
        if h:
            capi.lib.vgpu_vmem_destroy(h)
        torch.cuda.synchronize()
        capi.lib.vgpu_sm_partition_destroy(parts, 2)


def test_swap_trades_two_ranges_in_lock_step_and_keeps_every_byte(vgpu, cuda_dev):
    """vgpu_vmem_swap(out, in): A (resident) leaves while B (spilled) arrives, one block each way per step.  Contents
    of both verified after each trade by the memtest verify kernel (bit-exact or it counts mismatches); the HBM held
    during the call never exceeds what was held before it (sampled by a thread); a never-touched in-range is only
    mapped; a smaller `out` brings in only as many blocks as it freed."""
    import threading
    import time
    capi = vgpu.capi
    n = 1024                                                # 2 GiB each, blocks of 128 chunks = 256 MiB: 8 steps
    ha, a = _mk(capi, n, n, n, block_chunks=128)
    hb, b = _mk(capi, n, n, n, block_chunks=128)
    hc, c = _mk(capi, n // 4, n // 4, n // 4, block_chunks=128)
    res = torch.zeros(3, dtype=torch.int64, device=cuda_dev)
    lib = capi.lib

    def verify(base, chunks, seed):
        assert lib.vgpu_memtest_verify(base, chunks * CB, capi.PAT_RANDOM, seed, 0, res.data_ptr(), None) == 0
        torch.cuda.synchronize()
        assert res.cpu().tolist() == [0, 0, 0]

    try:
        assert lib.vgpu_vmem_ensure(hb, 0, n) == 0, capi.last_error()
        assert lib.vgpu_memtest_fill(b, n * CB, capi.PAT_RANDOM, 22, 0, None) == 0
        torch.cuda.synchronize()
        assert lib.vgpu_vmem_evict(hb, 0, n) == 0, capi.last_error()          # B waits in the host tier
        assert lib.vgpu_vmem_ensure(ha, 0, n) == 0, capi.last_error()
        assert lib.vgpu_memtest_fill(a, n * CB, capi.PAT_RANDOM, 11, 0, None) == 0
        torch.cuda.synchronize()
        assert lib.vgpu_vmem_evict(ha, 0, n) == 0 and lib.vgpu_vmem_ensure(ha, 0, n) == 0, capi.last_error()   # pinned pieces for A's trip exist now
        free_before, _ = torch.cuda.mem_get_info()
        low = [free_before]
        stop = threading.Event()

        def sample():
            while not stop.is_set():
                low[0] = min(low[0], torch.cuda.mem_get_info()[0])
                time.sleep(0.002)
        th = threading.Thread(target=sample)
        th.start()
        t0 = time.perf_counter()
        rc = lib.vgpu_vmem_swap(ha, hb, 0, n)
        t_swap = time.perf_counter() - t0
        stop.set()
        th.join()
        assert rc == 0, capi.last_error()
        assert low[0] >= free_before - (64 << 20)            # never more HBM than before the call (first-use scratch aside)
        sa, sb = _stats(capi, ha), _stats(capi, hb)
        assert sa.chunks_resident == 0 and sa.chunks_spilled == n and sb.chunks_resident == n and sb.chunks_spilled == 0
        assert sb.checksum_failures == 0
        verify(b, n, 22)
        # the same trade the other way with the two calls it replaces, for the time
        t0 = time.perf_counter()
        assert lib.vgpu_vmem_evict(hb, 0, n) == 0, capi.last_error()
        assert lib.vgpu_vmem_ensure(ha, 0, n) == 0, capi.last_error()
        t_two = time.perf_counter() - t0
        verify(a, n, 11)
        print(f"2 GiB out + 2 GiB in over the host tier: swap {t_swap * 1e3:.0f} ms, evict then ensure {t_two * 1e3:.0f} ms")
        # a never-touched range comes in by mapping alone, taking what A frees
        assert lib.vgpu_vmem_swap(ha, hc, 0, n // 4) == 0, capi.last_error()
        sc = _stats(capi, hc)
        assert sc.chunks_resident == n // 4 and sc.fetches == 0
        assert lib.vgpu_memtest_fill(c, (n // 4) * CB, capi.PAT_RANDOM, 33, 0, None) == 0
        torch.cuda.synchronize()
        # C (2 blocks) leaves, A (8 blocks) wants in: only as many blocks as C freed arrive, the rest is still missing
        assert lib.vgpu_vmem_swap(hc, ha, 0, n) == 0, capi.last_error()
        miss = C.c_uint64(0)
        assert lib.vgpu_vmem_missing(ha, 0, n, C.byref(miss)) == 0
        assert miss.value == n - n // 4
        assert lib.vgpu_vmem_ensure(ha, 0, n) == 0, capi.last_error()
        verify(a, n, 11)
        assert lib.vgpu_vmem_ensure(hc, 0, n // 4) == 0, capi.last_error()
        verify(c, n // 4, 33)
        # argument errors
        assert lib.vgpu_vmem_swap(ha, ha, 0, n) != 0 and lib.vgpu_vmem_swap(ha, hb, 0, n + 1) != 0
    finally:
        for h in (ha, hb, hc):
            lib.vgpu_vmem_destroy(h)
