"""ctypes wrapper of oracle/liboracle.so (oracle/vgpu_oracle.c) over numpy buffers.

TEST INFRASTRUCTURE ONLY — see oracle/vgpu_oracle.h.  Parity unpinned by the
reference (it has no spill/checksum/memtest code); pinned by hand-computed known
answers in tests/test_oracle.py.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "liboracle.so")


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "vgpu_oracle.c")
    if force or not os.path.exists(LIB_PATH) or os.path.getmtime(LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["gcc", "-O3", "-fPIC", "-shared", "-Wall", "-Wextra", "-o", LIB_PATH, src, "-lpthread"])
    return LIB_PATH


build()
lib = C.CDLL(LIB_PATH)
_u64, _vp, _sz, _i = C.c_uint64, C.c_void_p, C.c_size_t, C.c_int
lib.oracle_splitmix64.restype = _u64
lib.oracle_splitmix64.argtypes = [_u64]
lib.oracle_pattern_word.restype = _u64
lib.oracle_pattern_word.argtypes = [_i, _u64, _u64, _u64]
lib.oracle_fill.restype = None
lib.oracle_fill.argtypes = [_vp, _u64, _u64, _i, _u64, _u64]
lib.oracle_verify.restype = _u64
lib.oracle_verify.argtypes = [_vp, _u64, _u64, _i, _u64, _u64, C.POINTER(_u64), C.POINTER(_u64)]
lib.oracle_checksum.restype = None
lib.oracle_checksum.argtypes = [_vp, _sz, C.POINTER(_u64), C.POINTER(_u64)]
lib.oracle_spill.restype = None
lib.oracle_spill.argtypes = [_vp, _vp, _vp, _vp, _sz, _sz, _vp]
lib.oracle_memtest_mt.restype = C.c_double
lib.oracle_memtest_mt.argtypes = [_vp, _u64, _i, _u64, _u64, _u64, _i, C.POINTER(_u64)]
lib.oracle_spill_mt.restype = C.c_double
lib.oracle_spill_mt.argtypes = [_vp, _vp, _sz, _sz, _vp, _i]

PAT_WALKING_ONES, PAT_RANDOM = 0, 1
U64_MAX = (1 << 64) - 1


def _p(a: np.ndarray):
    return C.c_void_p(a.ctypes.data)


def splitmix64(x: int) -> int:
    return lib.oracle_splitmix64(x & U64_MAX)


def pattern_word(pattern: int, i: int, npass: int, seed: int = 0) -> int:
    return lib.oracle_pattern_word(pattern, i, npass, seed)


def fill(nwords: int, pattern: int, npass: int, seed: int = 0, first: int = 0) -> np.ndarray:
    """Words [first, first+nwords) of the expected buffer, as a fresh array."""
    buf = np.empty(first + nwords, dtype=np.uint64)
    lib.oracle_fill(_p(buf), first, nwords, pattern, npass, seed)
    return buf[first:]


def verify(buf: np.ndarray, pattern: int, npass: int, seed: int = 0):
    """-> (mismatches, first_bad_word or U64_MAX, bad_bits)"""
    buf = np.ascontiguousarray(buf.view(np.uint64))
    fb, bits = _u64(0), _u64(0)
    n = lib.oracle_verify(_p(buf), 0, buf.size, pattern, npass, seed, C.byref(fb), C.byref(bits))
    return int(n), int(fb.value), int(bits.value)


def checksum(chunk: np.ndarray):
    chunk = np.ascontiguousarray(chunk)
    x, s = _u64(0), _u64(0)
    lib.oracle_checksum(_p(chunk), chunk.nbytes, C.byref(x), C.byref(s))
    return int(x.value), int(s.value)


def spill(src: np.ndarray, dst: np.ndarray, nchunks: int, chunk_bytes: int, src_slot=None, dst_slot=None) -> np.ndarray:
    """memcpy + checksum per chunk; returns uint64 [nchunks, 2]."""
    xs = np.zeros((nchunks, 2), dtype=np.uint64)
    ss = None if src_slot is None else np.ascontiguousarray(src_slot, dtype=np.uint32)
    ds = None if dst_slot is None else np.ascontiguousarray(dst_slot, dtype=np.uint32)
    lib.oracle_spill(_p(src), _p(dst), None if ss is None else _p(ss), None if ds is None else _p(ds), nchunks,
                     chunk_bytes, _p(xs))
    return xs


def checksum_numpy(chunk: np.ndarray):
    """Independent numpy restatement used to cross-check the C oracle."""
    w = np.ascontiguousarray(chunk).view(np.uint64)
    x = int(np.bitwise_xor.reduce(w)) if w.size else 0
    k = (2 * np.arange(w.size, dtype=np.uint64) + np.uint64(1))
    with np.errstate(over="ignore"):
        s = int((w * k).sum(dtype=np.uint64)) if w.size else 0
    return x, s


def memtest_mt(buf: np.ndarray, pattern: int, first_pass: int, passes: int, seed: int, threads: int):
    bad = _u64(0)
    sec = lib.oracle_memtest_mt(_p(buf), buf.size, pattern, first_pass, passes, seed, threads, C.byref(bad))
    return sec, int(bad.value)


def spill_mt(src: np.ndarray, dst: np.ndarray, nchunks: int, chunk_bytes: int, threads: int):
    xs = np.zeros((nchunks, 2), dtype=np.uint64)
    sec = lib.oracle_spill_mt(_p(src), _p(dst), nchunks, chunk_bytes, _p(xs), threads)
    return sec, xs
