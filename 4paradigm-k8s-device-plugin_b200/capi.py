"""ctypes binding of include/vgpu.h (libvgpu_dp.so).

This is the same C-ABI a Go maintainer binds with cgo.  There
is no Python or CPU fallback: if the shared library is missing the import
fails, and every data-path call launches sm_100a kernels.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libvgpu_dp.so")

if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} not found: build it with `make` or `python -c 'import __graft_entry__ as g; g.build()'` "
        "(there is no CPU fallback for the vGPU data path)"
    )
lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)

# ---- constants (mirror vgpu.h) -------------------------------------------------
OK, E_INVAL, E_CUDA, E_NOMEM, E_UNKNOWN_ID, E_RANGE, E_MISMATCH, E_UNSUPPORTED = 0, -1, -2, -3, -4, -5, -6, -7
CHUNK_BYTES = 2097152
NOALLOC_MAX_CHUNKS = 4096
XFER_AUTO, XFER_LDST, XFER_TMA, XFER_CE, XFER_PULL = 0, 1, 2, 3, 4
TIER_LOCAL, TIER_PEER, TIER_HOST = 0, 1, 2
PAT_WALKING_ONES, PAT_RANDOM = 0, 1
SM_FINE = 1
ID_LEN, KEY_LEN, VAL_LEN, PATH_LEN = 128, 64, 1024, 256
U64_MAX = (1 << 64) - 1


class Desc(C.Structure):
    _fields_ = [("id", C.c_char * ID_LEN), ("uuid", C.c_char * ID_LEN), ("memory_mib", C.c_uint64),
                ("index", C.c_uint32), ("healthy", C.c_uint32)]


class KV(C.Structure):
    _fields_ = [("key", C.c_char * KEY_LEN), ("val", C.c_char * VAL_LEN)]


class Mount(C.Structure):
    _fields_ = [("container_path", C.c_char * PATH_LEN), ("host_path", C.c_char * PATH_LEN), ("read_only", C.c_int32)]


class PluginCfg(C.Structure):
    _fields_ = [("split", C.c_uint32), ("mem_scaling", C.c_double), ("cores_scaling", C.c_double),
                ("id_strategy", C.c_int32), ("list_strategy", C.c_int32), ("pcibusfile", C.c_char * PATH_LEN),
                ("emit_partition_index", C.c_uint32)]


class Cksum(C.Structure):
    _fields_ = [("x", C.c_uint64), ("s", C.c_uint64)]


class RestoreResult(C.Structure):
    _fields_ = [("bad_chunks", C.c_uint64), ("first_bad_inv", C.c_uint64)]


class MemtestDevResult(C.Structure):
    _fields_ = [("mismatches", C.c_uint64), ("first_bad_inv", C.c_uint64), ("bad_bits", C.c_uint64)]


class MemtestResult(C.Structure):
    _fields_ = [("bytes", C.c_uint64), ("passes", C.c_uint64), ("mismatches", C.c_uint64),
                ("first_bad_word", C.c_uint64), ("bad_bits", C.c_uint64), ("first_bad_pattern", C.c_int32),
                ("first_bad_pass", C.c_int32), ("fill_ms", C.c_double), ("verify_ms", C.c_double),
                ("gbps", C.c_double)]


class VmemStats(C.Structure):
    _fields_ = [("chunks_total", C.c_uint64), ("chunks_resident", C.c_uint64), ("chunks_spilled", C.c_uint64),
                ("resident_limit", C.c_uint64), ("tier_slots", C.c_uint64), ("evictions", C.c_uint64),
                ("fetches", C.c_uint64), ("checksum_failures", C.c_uint64), ("evict_ms", C.c_double), ("fetch_ms", C.c_double),
                ("evictions_peer", C.c_uint64), ("evictions_host", C.c_uint64), ("restore_retries", C.c_uint64), ("tier_setup_ms", C.c_double)]


class SmPartition(C.Structure):
    _fields_ = [("green_ctx", C.c_void_p), ("ctx", C.c_void_p), ("stream", C.c_void_p),
                ("sm_count", C.c_uint32), ("sm_total", C.c_uint32)]


_vp, _sz, _u64, _i, _u = C.c_void_p, C.c_size_t, C.c_uint64, C.c_int, C.c_uint
_pp = C.POINTER

# name -> (restype, argtypes); every symbol vgpu.h declares (tests check this list against the header)
SIGNATURES = {
    "vgpu_abi_version": (_i, []),
    "vgpu_last_error": (C.c_char_p, []),
    "vgpu_split": (_i, [C.c_char_p, _u64, _u, C.c_double, _pp(Desc), _sz]),
    "vgpu_sm_limit": (_i, [_u, C.c_double]),
    "vgpu_lookup": (_i, [_pp(Desc), _sz, _pp(C.c_char_p), _sz, _pp(_pp(Desc))]),
    "vgpu_best_effort": (_i, [_pp(_i), _sz, _pp(_i), _sz, _i, _pp(_i), _sz, _pp(_i), _sz]),
    "vgpu_allocate_env": (_i, [_pp(PluginCfg), _pp(_pp(Desc)), _sz, _pp(C.c_char_p), _pp(C.c_char_p), _sz,
                               C.c_char_p, _pp(KV), _pp(_sz), _pp(Mount), _pp(_sz)]),
    "vgpu_allocate_env_monitor": (_i, [_pp(PluginCfg), _pp(_pp(Desc)), _sz, _pp(C.c_char_p), _pp(C.c_char_p), _sz,
                                       C.c_char_p, C.c_char_p, _pp(KV), _pp(_sz), _pp(Mount), _pp(_sz)]),
    "vgpu_pcibus_list": (_i, [C.c_char_p, C.c_char_p, _sz]),
    "vgpu_checksum": (_i, [_vp, _vp, _sz, _sz, _vp, _vp]),
    "vgpu_spill": (_i, [_i, _i, _vp, _vp, _vp, _vp, _sz, _sz, _vp, _vp]),
    "vgpu_restore": (_i, [_i, _i, _vp, _vp, _vp, _vp, _sz, _sz, _vp, _vp, _vp, _vp]),
    "vgpu_spill_noalloc": (_i, [_vp, _vp, _vp, _vp, _sz, _sz, _vp, _vp]),
    "vgpu_spill_to_host": (_i, [_i, _i, _vp, _vp, _sz, _sz, _vp]),
    "vgpu_restore_from_host": (_i, [_i, _i, _vp, _vp, _sz, _sz, _vp, _pp(_u64)]),
    "vgpu_slab_alloc": (_i, [_i, _i, _i, _sz, _pp(_vp)]),
    "vgpu_slab_free": (_i, [_i, _i, _i, _vp]),
    "vgpu_calibrate": (_i, [_i, _i, _i, _sz]),
    "vgpu_link_rate": (_i, [_i, _i, _i, _sz, _pp(C.c_double), _pp(C.c_double)]),
    "vgpu_policy_get": (_i, [_i, _sz]),
    "vgpu_policy_set": (_i, [_i, _sz, _i]),
    "vgpu_policy_set_grid": (_i, [_i, _u]),
    "vgpu_vmem_create": (_i, [_i, _u64, _u64, _i, _i, _u64, _pp(_vp)]),
    "vgpu_vmem_create_tiered": (_i, [_i, _u64, _u64, _i, _u64, _u64, _pp(_vp)]),
    "vgpu_vmem_base": (_i, [_vp, _pp(_vp)]),
    "vgpu_vmem_ensure": (_i, [_vp, _u64, _u64]),
    "vgpu_vmem_missing": (_i, [_vp, _u64, _u64, _pp(_u64)]),
    "vgpu_vmem_evict": (_i, [_vp, _u64, _u64]),
    "vgpu_vmem_swap": (_i, [_vp, _vp, _u64, _u64]),
    "vgpu_vmem_locate": (_i, [_vp, _u64, _pp(_i), _pp(_vp)]),
    "vgpu_vmem_stats": (_i, [_vp, _pp(VmemStats)]),
    "vgpu_vmem_destroy": (_i, [_vp]),
    "vgpu_memtest_fill": (_i, [_vp, _u64, _i, _u64, _u64, _vp]),
    "vgpu_memtest_verify": (_i, [_vp, _u64, _i, _u64, _u64, _vp, _vp]),
    "vgpu_memtest_run": (_i, [_vp, _u64, _i, _u64, _u64, _u64, _vp, _pp(MemtestResult)]),
    "vgpu_memtest": (_i, [_i, _u64, _u, _u64, _u64, _pp(MemtestResult)]),
    "vgpu_sm_partition_create": (_i, [_i, _u, _u, _pp(SmPartition)]),
    "vgpu_sm_partition_create_ex": (_i, [_i, _u, _u, _u, _pp(SmPartition)]),
    "vgpu_sm_partition_destroy": (_i, [_pp(SmPartition), _u]),
    "vgpu_sm_probe": (_i, [_vp, _pp(_u64), _pp(_u)]),
}
for _name, (_res, _args) in SIGNATURES.items():
    _fn = getattr(lib, _name)  # AttributeError here == library does not export what the header declares
    _fn.restype = _res
    _fn.argtypes = _args


class VgpuError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"vgpu error {code}: {msg}")
        self.code = code
        self.msg = msg


def last_error() -> str:
    return (lib.vgpu_last_error() or b"").decode()


def check(rc: int) -> int:
    if rc < 0:
        raise VgpuError(rc, last_error())
    return rc
