"""Thin tensor-level helpers over the data-path C-ABI (include/vgpu.h).

torch is used only for device memory and streams; every function below ends in
one C-ABI call that launches the sm_100a kernels in csrc/vgpu_kernels.cuh.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import capi


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else C.c_void_p(t.data_ptr())


def _stream(stream) -> C.c_void_p:
    if stream is None:
        stream = torch.cuda.current_stream()
    if isinstance(stream, int):
        return C.c_void_p(stream)
    return C.c_void_p(stream.cuda_stream)


def _require_cuda(t: torch.Tensor):
    if not t.is_cuda and not t.is_pinned():
        raise ValueError("tensor must live in device memory or pinned host memory (no CPU fallback)")


def checksum(src: torch.Tensor, nchunks: int, chunk_bytes: int = capi.CHUNK_BYTES, src_slot=None, stream=None) -> torch.Tensor:
    """-> int64 tensor [nchunks, 2] holding the u64 bit patterns of (X, S), on src's device."""
    _require_cuda(src)
    dev = src.device if src.is_cuda else torch.device("cuda", torch.cuda.current_device())
    out = torch.empty((max(nchunks, 1), 2), dtype=torch.int64, device=dev)[:nchunks]
    with torch.cuda.device(dev):
        capi.check(capi.lib.vgpu_checksum(_ptr(src), _ptr(src_slot), nchunks, chunk_bytes, _ptr(out) if nchunks else None, _stream(stream)))
    return out


def spill(src: torch.Tensor, dst: torch.Tensor, nchunks: int, chunk_bytes: int = capi.CHUNK_BYTES, *,
          variant: int = capi.XFER_AUTO, tier: int = capi.TIER_LOCAL, src_slot=None, dst_slot=None,
          cksum: Optional[torch.Tensor] = None, stream=None) -> torch.Tensor:
    """Copy chunks src->dst and return their checksums [nchunks,2] (device uint64).  Runs on src's GPU."""
    _require_cuda(src)
    _require_cuda(dst)
    dev = src.device
    if cksum is None:
        cksum = torch.empty((max(nchunks, 1), 2), dtype=torch.int64, device=dev)[:nchunks]
    with torch.cuda.device(dev):
        capi.check(capi.lib.vgpu_spill(variant, tier, _ptr(src), _ptr(dst), _ptr(src_slot), _ptr(dst_slot), nchunks,
                                       chunk_bytes, _ptr(cksum) if nchunks else None, _stream(stream)))
    return cksum


def restore(src: torch.Tensor, dst: torch.Tensor, nchunks: int, expect: torch.Tensor, chunk_bytes: int = capi.CHUNK_BYTES, *,
            variant: int = capi.XFER_AUTO, tier: int = capi.TIER_LOCAL, src_slot=None, dst_slot=None,
            cksum: Optional[torch.Tensor] = None, result: Optional[torch.Tensor] = None, stream=None) -> torch.Tensor:
    """Copy chunks src->dst, verify against `expect` in-kernel.  Runs on dst's GPU.
    Returns the device result block (uint64[2]: bad_chunks, ~first_bad or 0)."""
    _require_cuda(src)
    _require_cuda(dst)
    dev = dst.device
    if result is None:
        result = torch.zeros(2, dtype=torch.int64, device=dev)
    with torch.cuda.device(dev):
        capi.check(capi.lib.vgpu_restore(variant, tier, _ptr(src), _ptr(dst), _ptr(src_slot), _ptr(dst_slot), nchunks,
                                         chunk_bytes, _ptr(expect), _ptr(cksum), _ptr(result), _stream(stream)))
    return result


def memtest_fill(buf: torch.Tensor, pattern: int, npass: int, seed: int = 0, nbytes: Optional[int] = None, stream=None):
    n = buf.numel() * buf.element_size() if nbytes is None else nbytes
    with torch.cuda.device(buf.device):
        capi.check(capi.lib.vgpu_memtest_fill(_ptr(buf), n, pattern, npass, seed, _stream(stream)))


def memtest_verify(buf: torch.Tensor, pattern: int, npass: int, seed: int = 0, nbytes: Optional[int] = None,
                   result: Optional[torch.Tensor] = None, stream=None) -> torch.Tensor:
    """-> device uint64[3]: mismatches, ~first_bad_word (0 = none), bad_bits (accumulates into `result`)."""
    n = buf.numel() * buf.element_size() if nbytes is None else nbytes
    if result is None:
        result = torch.zeros(3, dtype=torch.int64, device=buf.device)
    with torch.cuda.device(buf.device):
        capi.check(capi.lib.vgpu_memtest_verify(_ptr(buf), n, pattern, npass, seed, _ptr(result), _stream(stream)))
    return result


def memtest_run(buf: torch.Tensor, pattern: int, first_pass: int, npasses: int, seed: int = 0, stream=None) -> capi.MemtestResult:
    out = capi.MemtestResult()
    with torch.cuda.device(buf.device):
        capi.check(capi.lib.vgpu_memtest_run(_ptr(buf), buf.numel() * buf.element_size(), pattern, first_pass, npasses,
                                             seed, _stream(stream), C.byref(out)))
    return out
