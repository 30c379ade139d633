"""Pins the CPU oracle (oracle/vgpu_oracle.c) to published / hand-computed known answers.

The reference has no spill, checksum or memtest code, so there are no
reference vectors to pin against — "parity unpinned by the reference".  What pins it:
the published SplitMix64 test vector, hand-computed (X,S) values, and an independent
numpy restatement."""
import numpy as np
import pytest

from oracle import cpu

GOLDEN = 0x9E3779B97F4A7C15


def test_splitmix64_published_vector():
    # Vigna's splitmix64.c seeded with 0 yields these first three outputs; our
    # splitmix64(x) is the output function applied to state x + GOLDEN.
    assert cpu.splitmix64(0) == 0xE220A8397B1DCDAF
    assert cpu.splitmix64(GOLDEN) == 0x6E789E6AA1B965F4
    assert cpu.splitmix64((2 * GOLDEN) & cpu.U64_MAX) == 0x06C45D188009454F


def test_walking_ones_definition():
    for p in (0, 1, 63, 64, 130):
        b = cpu.fill(200, cpu.PAT_WALKING_ONES, p)
        assert [int(x) for x in b[:4]] == [1 << ((i + p) & 63) for i in range(4)]
        assert int(b[199]) == 1 << ((199 + p) & 63)
        assert all(bin(int(x)).count("1") == 1 for x in b)


def test_random_pattern_definition():
    b = cpu.fill(16, cpu.PAT_RANDOM, 0)
    assert int(b[0]) == cpu.splitmix64(GOLDEN)  # key = GOLDEN ^ pass ^ seed, word i = splitmix64(key + i)
    assert int(b[5]) == cpu.splitmix64(GOLDEN + 5)
    b7 = cpu.fill(16, cpu.PAT_RANDOM, 7, seed=0x1234)
    assert int(b7[3]) == cpu.splitmix64((GOLDEN ^ 7 ^ 0x1234) + 3)
    # first= offsets give the same words as slicing
    assert np.array_equal(cpu.fill(8, cpu.PAT_RANDOM, 3, first=100), cpu.fill(108, cpu.PAT_RANDOM, 3)[100:])


def test_checksum_hand_computed():
    w = np.array([1, 2, 3], dtype=np.uint64)
    assert cpu.checksum(w) == (1 ^ 2 ^ 3, 1 * 1 + 2 * 3 + 3 * 5)
    w = np.array([cpu.U64_MAX, cpu.U64_MAX], dtype=np.uint64)  # wraps mod 2^64
    assert cpu.checksum(w) == (0, (cpu.U64_MAX * 1 + cpu.U64_MAX * 3) % (1 << 64))
    assert cpu.checksum(np.zeros(0, dtype=np.uint64)) == (0, 0)


def test_checksum_detects_swap_and_bitflip():
    b = cpu.fill(4096, cpu.PAT_RANDOM, 0)
    base = cpu.checksum(b)
    s = b.copy(); s[[10, 11]] = s[[11, 10]]
    assert cpu.checksum(s)[0] == base[0] and cpu.checksum(s)[1] != base[1]  # X is order-blind, S is not
    f = b.copy(); f[77] ^= np.uint64(1 << 40)
    assert cpu.checksum(f)[0] != base[0]


@pytest.mark.parametrize("nbytes", [32, 4096, 2097152, 2097152 + 32])
def test_checksum_matches_numpy(nbytes):
    b = cpu.fill(nbytes // 8, cpu.PAT_RANDOM, 11)
    assert cpu.checksum(b) == cpu.checksum_numpy(b)


def test_verify_reports_first_bad_and_bits():
    b = cpu.fill(1000, cpu.PAT_RANDOM, 2, seed=9).copy()
    assert cpu.verify(b, cpu.PAT_RANDOM, 2, seed=9) == (0, cpu.U64_MAX, 0)
    b[700] ^= np.uint64(0b101); b[300] ^= np.uint64(1 << 63)
    assert cpu.verify(b, cpu.PAT_RANDOM, 2, seed=9) == (2, 300, (1 << 63) | 0b101)
    assert cpu.verify(b, cpu.PAT_RANDOM, 3, seed=9)[0] == 1000  # wrong pass: everything differs


def test_spill_with_slot_tables():
    cb, n = 4096, 5
    src = cpu.fill(8 * cb // 8, cpu.PAT_RANDOM, 1).copy()
    dst = np.zeros(8 * cb // 8, dtype=np.uint64)
    ss, ds = [7, 0, 3, 3, 5], [1, 2, 0, 6, 4]
    xs = cpu.spill(src, dst, n, cb, ss, ds)
    w = cb // 8
    for c in range(n):
        assert np.array_equal(dst[ds[c] * w:(ds[c] + 1) * w], src[ss[c] * w:(ss[c] + 1) * w])
        assert tuple(int(v) for v in xs[c]) == cpu.checksum_numpy(src[ss[c] * w:(ss[c] + 1) * w])


def test_mt_drivers_agree_with_scalar():
    b = np.zeros(1 << 16, dtype=np.uint64)
    sec, bad = cpu.memtest_mt(b, cpu.PAT_RANDOM, 0, 2, 0, 3)
    assert bad == 0 and sec > 0
    assert np.array_equal(b, cpu.fill(1 << 16, cpu.PAT_RANDOM, 1))  # last pass left in place
    src = cpu.fill(16 * 512, cpu.PAT_RANDOM, 5).copy()
    dst = np.zeros_like(src)
    sec, xs = cpu.spill_mt(src, dst, 16, 4096, 3)
    assert np.array_equal(src, dst) and np.array_equal(xs, cpu.spill(src, np.zeros_like(src), 16, 4096))


def test_oracle_reproduces_committed_datapath_vectors():
    """tests/golden/datapath_vectors.json (oracle/gen_datapath_golden.py): pins the definitions."""
    import json, os
    g = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "datapath_vectors.json")))
    for p in g["patterns"]:
        assert [hex(int(x)) for x in cpu.fill(8, p["pattern"], p["pass"], p["seed"])] == p["first_words"]
        far = cpu.fill(4, p["pattern"], p["pass"], p["seed"], first=(40 << 30) // 8 - 4)
        assert [hex(int(x)) for x in far] == p["words_before_40GiB"]
    for c in g["checksums"]:
        n = c["chunk_bytes"] // 8
        words = cpu.fill(c["nchunks"] * n, cpu.PAT_RANDOM, c["pass"], 0)
        got = [[hex(v) for v in cpu.checksum(words[i * n:(i + 1) * n])] for i in range(c["nchunks"])]
        assert got == c["xs"]
    # a hand-checkable one: walking ones, pass 0 -> 1, 2, 4, 8 ...
    assert g["patterns"][0]["first_words"] == [hex(1 << i) for i in range(8)]
