"""GPU parity tests of the data path (spill / restore / checksum / memtest) through the
C-ABI against the CPU oracle (oracle/vgpu_oracle.c).  Bit-exact: this is integer/byte work."""
import ctypes as C

import numpy as np
import pytest
import torch

from oracle import cpu

pytestmark = pytest.mark.gpu

VARIANTS = {"ldst": 1, "tma": 2, "ce": 3}


def dev_bytes(t: torch.Tensor) -> np.ndarray:
    return t.view(torch.uint8).cpu().numpy()


def rand_words(n, npass=0, seed=0):
    return cpu.fill(n, cpu.PAT_RANDOM, npass, seed).copy()


def to_dev(a: np.ndarray, dev):
    return torch.from_numpy(a.view(np.uint8).copy()).to(dev)


def as_pairs(t: torch.Tensor) -> np.ndarray:
    return t.cpu().numpy().view(np.uint64).reshape(-1, 2)


# ---------------------------------------------------------------- memtest ----------------
@pytest.mark.parametrize("pattern", [cpu.PAT_WALKING_ONES, cpu.PAT_RANDOM])
@pytest.mark.parametrize("nwords,offset_words", [(1, 0), (3, 1), (4, 0), (1000, 0), (1001, 3), (262144, 0), (262147, 2), (1 << 22, 1)])
def test_memtest_fill_matches_oracle(vgpu, cuda_dev, pattern, nwords, offset_words):
    dp = vgpu.datapath
    for npass, seed in [(0, 0), (5, 0), (63, 0xABCDEF)]:
        backing = torch.zeros(nwords + 8, dtype=torch.int64, device=cuda_dev)
        view = backing[offset_words:offset_words + nwords]  # ragged head/tail around the 32-byte vectors
        dp.memtest_fill(view, pattern, npass, seed)
        torch.cuda.synchronize()
        got = backing.cpu().numpy().view(np.uint64)
        want = cpu.fill(nwords, pattern, npass, seed)
        assert np.array_equal(got[offset_words:offset_words + nwords], want)
        assert not got[:offset_words].any() and not got[offset_words + nwords:].any(), "wrote outside the buffer"


@pytest.mark.parametrize("pattern", [cpu.PAT_WALKING_ONES, cpu.PAT_RANDOM])
def test_memtest_verify_clean_and_faulty(vgpu, cuda_dev, pattern):
    dp = vgpu.datapath
    n = (1 << 20) + 5
    want = cpu.fill(n, pattern, 9, 77).copy()
    buf = torch.from_numpy(want.view(np.int64)).to(cuda_dev)
    res = dp.memtest_verify(buf, pattern, 9, 77)
    assert res.cpu().tolist() == [0, 0, 0]
    # inject stuck-at faults: the kernel's counters must equal the oracle's
    bad = want.copy()
    for idx, bits in [(123456, 1 << 17), (n - 1, 0b11), (7, 1 << 63), (500000, 0xFF00)]:
        bad[idx] ^= np.uint64(bits)
    buf = torch.from_numpy(bad.view(np.int64)).to(cuda_dev)
    res = dp.memtest_verify(buf, pattern, 9, 77).cpu().numpy().view(np.uint64)
    o_n, o_first, o_bits = cpu.verify(bad, pattern, 9, 77)
    assert int(res[0]) == o_n == 4
    assert int(~res[1] & np.uint64(cpu.U64_MAX)) == o_first == 7
    assert int(res[2]) == o_bits
    # wrong pass => every word mismatches (walking ones) / all random words differ
    res = dp.memtest_verify(buf, pattern, 10, 77).cpu().numpy().view(np.uint64)
    assert int(res[0]) == cpu.verify(bad, pattern, 10, 77)[0]


@pytest.mark.parametrize("wrap_at", [5, 1024, 1537, 262144 + 3])
def test_random_verify_where_the_word_index_carries_into_the_high_half(vgpu, cuda_dev, wrap_at):
    """The random verify kernel shares the high 32 bits of (key + golden + index) between the 16 words of a thread and
    falls back to full 64-bit arithmetic for the rare thread whose span crosses a 2^32 boundary.  Seeds chosen so that
    the low half wraps `wrap_at` words into the buffer — inside one thread's span, at a vector boundary, in the ragged
    head — must still agree with the oracle bit for bit, clean and with faults planted around the wrap."""
    dp = vgpu.datapath
    G = 0x9E3779B97F4A7C15
    npass = 3
    key_low = (0x1_0000_0000 - wrap_at - (G & 0xFFFFFFFF)) & 0xFFFFFFFF          # (key + G + wrap_at) low half == 0
    key = (0x5A5A5A5A << 32) | key_low
    seed = key ^ G ^ npass                                                        # key = G ^ pass ^ seed
    n = (1 << 19) + 7
    want = cpu.fill(n, cpu.PAT_RANDOM, npass, seed).copy()
    for off in (0, 1, 3):                                                         # aligned and ragged starts
        backing = torch.zeros(n + 8, dtype=torch.int64, device=cuda_dev)
        view = backing[off:off + n]
        view.copy_(torch.from_numpy(want.view(np.int64)))
        assert dp.memtest_verify(view, cpu.PAT_RANDOM, npass, seed).cpu().tolist() == [0, 0, 0]
        dp.memtest_fill(view, cpu.PAT_RANDOM, npass, seed)                        # and what the fill kernel writes is the same
        torch.cuda.synchronize()
        assert np.array_equal(view.cpu().numpy().view(np.uint64), want)
    bad = want.copy()
    for idx in (wrap_at - 1, wrap_at, wrap_at + 1, n - 1):
        bad[idx] ^= np.uint64(1 << (idx % 64))
    buf = torch.from_numpy(bad.view(np.int64)).to(cuda_dev)
    res = dp.memtest_verify(buf, cpu.PAT_RANDOM, npass, seed).cpu().numpy().view(np.uint64)
    o_n, o_first, o_bits = cpu.verify(bad, cpu.PAT_RANDOM, npass, seed)
    assert (int(res[0]), int(~res[1] & np.uint64(cpu.U64_MAX)), int(res[2])) == (o_n, o_first, o_bits) and o_n == 4


def test_memtest_run_reports_rates_and_faults(vgpu, cuda_dev):
    dp, capi = vgpu.datapath, vgpu.capi
    buf = torch.empty(256 << 20, dtype=torch.uint8, device=cuda_dev)
    r = dp.memtest_run(buf, capi.PAT_WALKING_ONES, 0, 64)
    assert r.passes == 64 and r.mismatches == 0 and r.first_bad_word == capi.U64_MAX and r.first_bad_pattern == -1
    assert r.bytes == buf.numel() and r.gbps > 100 and r.fill_ms > 0 and r.verify_ms > 0
    r = dp.memtest_run(buf, capi.PAT_RANDOM, 0, 3, seed=5)
    assert r.passes == 3 and r.mismatches == 0
    # the last pass left in memory must be exactly the oracle's (first 1 MiB compared)
    got = buf[: 1 << 20].cpu().numpy().view(np.uint64)
    assert np.array_equal(got, cpu.fill((1 << 20) // 8, cpu.PAT_RANDOM, 2, 5))


def test_memtest_self_allocating_entry_point(vgpu, cuda_dev):
    capi = vgpu.capi
    out = capi.MemtestResult()
    rc = capi.lib.vgpu_memtest(0, 64 << 20, 3, 2, 0, C.byref(out))
    assert rc == 0, capi.last_error()
    assert out.passes == 66 and out.mismatches == 0 and out.gbps > 0


def test_memtest_rejects_misaligned(vgpu, cuda_dev):
    capi = vgpu.capi
    buf = torch.empty(1024, dtype=torch.uint8, device=cuda_dev)
    assert capi.lib.vgpu_memtest_fill(C.c_void_p(buf.data_ptr() + 1), 512, 0, 0, 0, None) == capi.E_INVAL
    assert capi.lib.vgpu_memtest_fill(C.c_void_p(buf.data_ptr()), 100, 0, 0, 0, None) == capi.E_INVAL
    assert capi.lib.vgpu_memtest_fill(C.c_void_p(buf.data_ptr()), 512, 9, 0, 0, None) == capi.E_INVAL


def test_kernels_reproduce_committed_golden_vectors(vgpu, cuda_dev):
    """tests/golden/datapath_vectors.json: the kernels against committed known answers (no oracle
    library in the loop), including the last words of a 40 GiB buffer (64-bit word indices)."""
    import json, os
    dp, capi = vgpu.datapath, vgpu.capi
    g = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "datapath_vectors.json")))
    for p in g["patterns"]:
        buf = torch.zeros(8, dtype=torch.int64, device=cuda_dev)
        dp.memtest_fill(buf, p["pattern"], p["pass"], p["seed"])
        assert [hex(int(x)) for x in buf.cpu().numpy().view(np.uint64)] == p["first_words"]
    big = torch.empty(40 << 30, dtype=torch.uint8, device=cuda_dev)          # config 2's vGPU size
    for p in g["patterns"]:
        dp.memtest_fill(big, p["pattern"], p["pass"], p["seed"])
        tail = big[-32:].cpu().numpy().view(np.uint64)
        assert [hex(int(x)) for x in tail] == p["words_before_40GiB"]
        assert dp.memtest_verify(big, p["pattern"], p["pass"], p["seed"]).cpu().tolist() == [0, 0, 0]
    del big
    for c in g["checksums"]:
        src = torch.empty(c["nchunks"] * c["chunk_bytes"], dtype=torch.uint8, device=cuda_dev)
        dp.memtest_fill(src, capi.PAT_RANDOM, c["pass"], 0)
        for v in (capi.XFER_LDST, capi.XFER_TMA):
            ck = dp.spill(src, torch.empty_like(src), c["nchunks"], c["chunk_bytes"], variant=v)
            got = [[hex(int(x)) for x in row] for row in as_pairs(ck)]
            assert got == c["xs"], (v, c["source"])


# ---------------------------------------------------------------- checksum / spill / restore
@pytest.mark.parametrize("chunk_bytes,nchunks", [(32, 1), (32, 7), (4096, 3), (32768, 5), (32768 + 32, 4), (2097152, 1), (2097152, 9), (2097152 - 64, 3), (3 * 2097152 + 96, 2)])
def test_checksum_matches_oracle(vgpu, cuda_dev, chunk_bytes, nchunks):
    dp = vgpu.datapath
    words = rand_words(nchunks * chunk_bytes // 8, npass=chunk_bytes & 63)
    src = to_dev(words, cuda_dev)
    got = as_pairs(dp.checksum(src, nchunks, chunk_bytes))
    w = chunk_bytes // 8
    want = np.array([cpu.checksum(words[c * w:(c + 1) * w]) for c in range(nchunks)], dtype=np.uint64)
    assert np.array_equal(got, want)
    # scratch accumulators are self-cleaning: a second launch gives the same answer
    assert np.array_equal(as_pairs(dp.checksum(src, nchunks, chunk_bytes)), want)


@pytest.mark.parametrize("variant", list(VARIANTS))
@pytest.mark.parametrize("chunk_bytes,nchunks", [(32, 3), (16384, 4), (16384 + 32, 3), (65536 - 32, 5), (2097152, 1), (2097152, 17), (2097152 + 4096, 3)])
def test_spill_matches_oracle(vgpu, cuda_dev, variant, chunk_bytes, nchunks):
    dp = vgpu.datapath
    words = rand_words(nchunks * chunk_bytes // 8, npass=3)
    src = to_dev(words, cuda_dev)
    dst = torch.zeros(nchunks * chunk_bytes + 64, dtype=torch.uint8, device=cuda_dev)
    ck = dp.spill(src, dst, nchunks, chunk_bytes, variant=VARIANTS[variant])
    o_dst = np.zeros_like(words)
    want = cpu.spill(words, o_dst, nchunks, chunk_bytes)
    assert np.array_equal(as_pairs(ck), want)
    got = dev_bytes(dst)
    assert np.array_equal(got[: nchunks * chunk_bytes], o_dst.view(np.uint8))
    assert not got[nchunks * chunk_bytes:].any(), "wrote past the last chunk"


@pytest.mark.parametrize("variant", ["ldst", "tma"])
def test_spill_with_slot_tables(vgpu, cuda_dev, variant):
    dp = vgpu.datapath
    cb, slots, n = 2097152, 12, 7
    rng = np.random.default_rng(5)
    ss = rng.permutation(slots)[:n].astype(np.uint32)
    ds = rng.permutation(slots)[:n].astype(np.uint32)
    words = rand_words(slots * cb // 8, npass=8)
    src = to_dev(words, cuda_dev)
    dst = torch.zeros(slots * cb, dtype=torch.uint8, device=cuda_dev)
    ck = dp.spill(src, dst, n, cb, variant=VARIANTS[variant],
                  src_slot=torch.from_numpy(ss.view(np.int32)).to(cuda_dev), dst_slot=torch.from_numpy(ds.view(np.int32)).to(cuda_dev))
    o_dst = np.zeros_like(words)
    want = cpu.spill(words, o_dst, n, cb, ss, ds)
    assert np.array_equal(as_pairs(ck), want)
    assert np.array_equal(dev_bytes(dst), o_dst.view(np.uint8))


def test_ce_variant_with_host_slot_tables(vgpu, cuda_dev):
    capi = vgpu.capi
    cb, slots, n = 65536, 10, 6
    # slot tables in pinned memory: the host walks them to issue copies, the checksum kernel reads them
    pss = torch.tensor([4, 5, 6, 0, 9, 2], dtype=torch.int32).pin_memory()  # 4,5,6 -> 1,2,3 coalesces into one copy
    pds = torch.tensor([1, 2, 3, 7, 0, 5], dtype=torch.int32).pin_memory()
    ss, ds = pss.numpy().view(np.uint32), pds.numpy().view(np.uint32)
    words = rand_words(slots * cb // 8, npass=1)
    src = to_dev(words, cuda_dev)
    dst = torch.zeros(slots * cb, dtype=torch.uint8, device=cuda_dev)
    ck = torch.zeros((n, 2), dtype=torch.int64, device=cuda_dev)
    rc = capi.lib.vgpu_spill(capi.XFER_CE, capi.TIER_LOCAL, src.data_ptr(), dst.data_ptr(), pss.data_ptr(), pds.data_ptr(),
                             n, cb, ck.data_ptr(), None)
    assert rc == 0, capi.last_error()
    torch.cuda.synchronize()
    o_dst = np.zeros_like(words)
    want = cpu.spill(words, o_dst, n, cb, ss, ds)
    assert np.array_equal(as_pairs(ck), want) and np.array_equal(dev_bytes(dst), o_dst.view(np.uint8))
    # device-resident slot tables cannot be walked by the host for per-chunk copies
    dss = pss.to(cuda_dev)
    rc = capi.lib.vgpu_spill(capi.XFER_CE, capi.TIER_LOCAL, src.data_ptr(), dst.data_ptr(), dss.data_ptr(), None, n, cb, ck.data_ptr(), None)
    assert rc == capi.E_UNSUPPORTED


@pytest.mark.parametrize("variant", list(VARIANTS))
def test_restore_round_trip_and_corruption(vgpu, cuda_dev, variant):
    dp = vgpu.datapath
    cb, n = 2097152, 24
    v = VARIANTS[variant]
    words = rand_words(n * cb // 8, npass=21)
    hbm = to_dev(words, cuda_dev)
    tier = torch.zeros(n * cb, dtype=torch.uint8, device=cuda_dev)
    ck = dp.spill(hbm, tier, n, cb, variant=v)
    hbm.zero_()  # the vGPU's slice is gone
    got_ck = torch.zeros((n, 2), dtype=torch.int64, device=cuda_dev)
    res = dp.restore(tier, hbm, n, ck, cb, variant=v, cksum=got_ck)
    assert res.cpu().tolist() == [0, 0]
    assert np.array_equal(dev_bytes(hbm), words.view(np.uint8)), "restored bytes differ"
    assert torch.equal(got_ck.cpu(), ck.cpu())
    # flip one bit in chunk 13 and one word swap in chunk 5 of the spill tier
    t64 = tier.view(torch.int64)
    t64[13 * (cb // 8) + 999] ^= 1 << 33
    a, b = 5 * (cb // 8) + 10, 5 * (cb // 8) + 20
    tmp = t64[a].clone(); t64[a] = t64[b]; t64[b] = tmp
    res = dp.restore(tier, hbm, n, ck, cb, variant=v).cpu().numpy().view(np.uint64)
    assert int(res[0]) == 2
    assert int(~res[1] & np.uint64(cpu.U64_MAX)) == 5


@pytest.mark.parametrize("variant", list(VARIANTS))
def test_host_tier_round_trip_through_host_buffer_entry_points(vgpu, cuda_dev, variant):
    """vgpu_spill_to_host / vgpu_restore_from_host: the calls bench.py's e2e leg times."""
    capi = vgpu.capi
    cb, n = 2097152, 6
    words = rand_words(n * cb // 8, npass=4)
    hbm = to_dev(words, cuda_dev)
    host = torch.zeros(n * cb, dtype=torch.uint8).pin_memory()
    hck = np.zeros((n, 2), dtype=np.uint64)
    rc = capi.lib.vgpu_spill_to_host(VARIANTS[variant], 0, hbm.data_ptr(), host.data_ptr(), n, cb, hck.ctypes.data)
    assert rc == 0, capi.last_error()
    assert np.array_equal(host.numpy(), words.view(np.uint8))
    assert np.array_equal(hck, cpu.spill(words, np.zeros_like(words), n, cb))
    hbm.zero_()
    bad = C.c_uint64(99)
    rc = capi.lib.vgpu_restore_from_host(VARIANTS[variant], 0, host.data_ptr(), hbm.data_ptr(), n, cb, hck.ctypes.data, C.byref(bad))
    assert rc == 0 and bad.value == 0, capi.last_error()
    assert np.array_equal(dev_bytes(hbm), words.view(np.uint8))
    host[3 * cb + 17] ^= 0x40
    rc = capi.lib.vgpu_restore_from_host(VARIANTS[variant], 0, host.data_ptr(), hbm.data_ptr(), n, cb, hck.ctypes.data, C.byref(bad))
    assert rc == capi.E_MISMATCH and bad.value == 1 and "first 3" in capi.last_error()


def test_host_entry_point_pins_pageable_memory(vgpu, cuda_dev):
    capi = vgpu.capi
    cb, n = 65536, 4
    words = rand_words(n * cb // 8)
    hbm = to_dev(words, cuda_dev)
    raw = np.zeros(n * cb + 64, dtype=np.uint8)  # pageable; the slab base must be 32-byte aligned
    off = (-raw.ctypes.data) % 32
    host = raw[off:off + n * cb]
    hck = np.zeros((n, 2), dtype=np.uint64)
    rc = capi.lib.vgpu_spill_to_host(capi.XFER_LDST, 0, hbm.data_ptr(), host.ctypes.data, n, cb, hck.ctypes.data)
    assert rc == 0, capi.last_error()
    assert np.array_equal(host, words.view(np.uint8))


def test_peer_tier_round_trip(vgpu, cuda_dev):
    if torch.cuda.device_count() < 2:
        pytest.skip("peer tier needs 2 GPUs (covered by the --gpus 2 run)")
    capi, dp = vgpu.capi, vgpu.datapath
    cb, n = 2097152, 16
    words = rand_words(n * cb // 8, npass=2)
    hbm = to_dev(words, cuda_dev)
    slab = C.c_void_p()
    assert capi.lib.vgpu_slab_alloc(capi.TIER_PEER, 0, 1, n * cb, C.byref(slab)) == 0, capi.last_error()
    try:
        for v in (capi.XFER_LDST, capi.XFER_TMA, capi.XFER_CE, capi.XFER_PULL):  # PULL: the peer GPU runs the spill kernel
            ck = torch.zeros((n, 2), dtype=torch.int64, device=cuda_dev)
            res = torch.zeros(2, dtype=torch.int64, device=cuda_dev)
            with torch.cuda.device(0):
                assert capi.lib.vgpu_spill(v, capi.TIER_PEER, hbm.data_ptr(), slab, None, None, n, cb, ck.data_ptr(), None) == 0, capi.last_error()
                back = torch.zeros_like(hbm)
                assert capi.lib.vgpu_restore(v, capi.TIER_PEER, slab, back.data_ptr(), None, None, n, cb, ck.data_ptr(), None, res.data_ptr(), None) == 0, capi.last_error()
                torch.cuda.synchronize()
            assert np.array_equal(as_pairs(ck), cpu.spill(words, np.zeros_like(words), n, cb))
            assert res.cpu().tolist() == [0, 0]
            assert np.array_equal(dev_bytes(back), words.view(np.uint8))
    finally:
        capi.lib.vgpu_slab_free(capi.TIER_PEER, 0, 1, slab)


def test_large_round_trip_properties(vgpu, cuda_dev):
    """BASELINE config 4 sizes (4096 x 2 MiB = 8 GiB): too big for the oracle in seconds, so use
    size-independent properties: encode->erase->decode round trip, checksum-of-checksums stable
    across variants, XOR-linearity of X."""
    dp, capi = vgpu.datapath, vgpu.capi
    cb, n = capi.CHUNK_BYTES, 4096
    hbm = torch.empty(n * cb, dtype=torch.uint8, device=cuda_dev)
    dp.memtest_fill(hbm, capi.PAT_RANDOM, 42)
    tier = torch.empty_like(hbm)
    cks = {}
    for name, v in VARIANTS.items():
        tier.zero_()
        cks[name] = dp.spill(hbm, tier, n, cb, variant=v)
        assert torch.equal(tier, hbm), name
    assert torch.equal(cks["ldst"], cks["tma"]) and torch.equal(cks["ldst"], cks["ce"])
    assert torch.equal(dp.checksum(hbm, n, cb), cks["ldst"])
    # first and last chunk against the oracle
    for c in (0, n - 1):
        w = cpu.fill(cb // 8, capi.PAT_RANDOM, 42, first=c * (cb // 8))
        assert tuple(int(x) for x in as_pairs(cks["ldst"][c:c + 1])[0]) == cpu.checksum(w)
    # erase, restore, verify with the memtest kernel (pattern regenerates on the fly)
    hbm.zero_()
    res = dp.restore(tier, hbm, n, cks["ldst"], cb, variant=capi.XFER_TMA)
    assert res.cpu().tolist() == [0, 0]
    assert dp.memtest_verify(hbm, capi.PAT_RANDOM, 42).cpu().tolist() == [0, 0, 0]
    # linearity of X: X(a ^ b) == X(a) ^ X(b)
    other = torch.empty_like(tier)
    dp.memtest_fill(other, capi.PAT_RANDOM, 43)
    xa = dp.checksum(hbm, n, cb)[:, 0].cpu().numpy()
    xb = dp.checksum(other, n, cb)[:, 0].cpu().numpy()
    other.view(torch.int64).bitwise_xor_(hbm.view(torch.int64))
    xab = dp.checksum(other, n, cb)[:, 0].cpu().numpy()
    assert np.array_equal(xab, xa ^ xb)


def test_first_launch_on_a_fresh_non_blocking_stream(vgpu, cuda_dev):
    """The per-stream scratch is zeroed on the caller's stream: on a non-blocking stream (torch
    streams, green-context streams) a legacy-stream memset would race with the first kernel."""
    dp, capi = vgpu.datapath, vgpu.capi
    cb, n = capi.CHUNK_BYTES, 512
    src = torch.empty(n * cb, dtype=torch.uint8, device=cuda_dev)
    dp.memtest_fill(src, capi.PAT_RANDOM, 31)
    want = dp.checksum(src, n, cb)
    torch.cuda.synchronize()
    for _ in range(4):
        st = torch.cuda.Stream(device=cuda_dev)          # cudaStreamNonBlocking
        dst = torch.empty_like(src)
        ck = dp.spill(src, dst, n, cb, variant=capi.XFER_LDST, stream=st)
        st.synchronize()
        assert torch.equal(ck, want) and torch.equal(dst, src)


def test_calibrate_fills_the_auto_policy_from_measurements(vgpu, cuda_dev):
    """vgpu_calibrate times the three variants on this box and rewrites the AUTO table
    ("switch to the copy engine when measured bandwidth favours it")."""
    capi, dp = vgpu.capi, vgpu.datapath
    for tier in (capi.TIER_LOCAL, capi.TIER_HOST):
        assert capi.lib.vgpu_calibrate(0, tier, -1, 64) == 0, capi.last_error()
        for n in (1, 8, 64, 4096):
            assert capi.lib.vgpu_policy_get(tier, n) in (capi.XFER_LDST, capi.XFER_TMA, capi.XFER_CE)
    # whatever was chosen, AUTO still moves bytes exactly
    cb, n = capi.CHUNK_BYTES, 64
    src = torch.empty(n * cb, dtype=torch.uint8, device=cuda_dev)
    dp.memtest_fill(src, capi.PAT_RANDOM, 3)
    dst = torch.zeros_like(src)
    ck = dp.spill(src, dst, n, cb, variant=capi.XFER_AUTO)
    assert torch.equal(dst, src) and torch.equal(ck, dp.checksum(src, n, cb))
    # explicit rules win over measured ones
    assert capi.lib.vgpu_policy_set(capi.TIER_HOST, 8, capi.XFER_CE) == 0
    assert capi.lib.vgpu_policy_get(capi.TIER_HOST, 100) == capi.XFER_CE


def test_argument_validation(vgpu, cuda_dev):
    capi = vgpu.capi
    a = torch.zeros(1 << 20, dtype=torch.uint8, device=cuda_dev)
    ck = torch.zeros((4, 2), dtype=torch.int64, device=cuda_dev)
    f = capi.lib.vgpu_spill
    assert f(1, 0, a.data_ptr() + 8, a.data_ptr(), None, None, 1, 4096, ck.data_ptr(), None) == capi.E_INVAL  # alignment
    assert f(1, 0, a.data_ptr(), a.data_ptr(), None, None, 1, 4100, ck.data_ptr(), None) == capi.E_INVAL      # chunk size
    assert f(9, 0, a.data_ptr(), a.data_ptr(), None, None, 1, 4096, ck.data_ptr(), None) == capi.E_INVAL      # variant
    assert f(1, 5, a.data_ptr(), a.data_ptr(), None, None, 1, 4096, ck.data_ptr(), None) == capi.E_INVAL      # tier
    assert capi.lib.vgpu_restore(1, 0, a.data_ptr(), a.data_ptr(), None, None, 1, 4096, ck.data_ptr(), None, None, None) == capi.E_INVAL


def test_sm_partitions_are_spatial(vgpu, cuda_dev):
    """CUDA_DEVICE_SM_LIMIT=25 with split 4 (config 3): four disjoint green contexts; a
    kernel on one partition's stream gives the same bytes as anywhere else."""
    capi, dp = vgpu.capi, vgpu.datapath
    parts = (capi.SmPartition * 4)()
    rc = capi.lib.vgpu_sm_partition_create(0, 25, 4, parts)
    assert rc == 0, capi.last_error()
    try:
        total = parts[0].sm_total
        assert total >= 100
        for p in parts:
            assert 0 < p.sm_count <= total * 25 // 100 and p.stream
        assert sum(p.sm_count for p in parts) <= total
        bufs = [torch.empty(64 << 20, dtype=torch.uint8, device=cuda_dev) for _ in parts]
        res = [torch.zeros(3, dtype=torch.int64, device=cuda_dev) for _ in parts]
        torch.cuda.synchronize()
        for i, p in enumerate(parts):
            dp.memtest_fill(bufs[i], capi.PAT_RANDOM, i, stream=p.stream)
            dp.memtest_verify(bufs[i], capi.PAT_RANDOM, i, result=res[i], stream=p.stream)
        torch.cuda.synchronize()
        for i in range(4):
            assert res[i].cpu().tolist() == [0, 0, 0]
            assert np.array_equal(bufs[i][:4096].cpu().numpy().view(np.uint64), cpu.fill(512, cpu.PAT_RANDOM, i))
    finally:
        capi.lib.vgpu_sm_partition_destroy(parts, 4)
