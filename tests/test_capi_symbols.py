"""The C-ABI library loads without a GPU and exports every symbol include/vgpu.h declares."""
import ctypes as C
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_functions():
    txt = open(os.path.join(ROOT, "include", "vgpu.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(vgpu_[a-z0-9_]+)\s*\(", txt)))


def test_header_declares_something():
    names = header_functions()
    assert "vgpu_spill" in names and "vgpu_memtest" in names and "vgpu_allocate_env" in names


def test_library_exports_every_declared_symbol(vgpu):
    lib = vgpu.capi.lib
    for name in header_functions():
        assert hasattr(lib, name), f"libvgpu_dp.so does not export {name}"
    assert set(header_functions()) == set(vgpu.capi.SIGNATURES), "capi.SIGNATURES out of sync with vgpu.h"
    assert lib.vgpu_abi_version() == 2


def test_library_carries_sm100a_code_only(vgpu):
    out = subprocess.run(["cuobjdump", "--list-elf", vgpu.capi.LIB_PATH], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_\d+a?", out))
    assert archs == {"sm_100a"}, archs


def test_struct_sizes_match_header(vgpu):
    c = vgpu.capi
    assert C.sizeof(c.Cksum) == 16 and C.sizeof(c.RestoreResult) == 16 and C.sizeof(c.MemtestDevResult) == 24
    assert C.sizeof(c.Desc) == 128 + 128 + 8 + 4 + 4
    assert C.sizeof(c.MemtestResult) == 5 * 8 + 2 * 4 + 3 * 8


def test_argument_errors_need_no_gpu(vgpu):
    c = vgpu.capi
    assert c.lib.vgpu_policy_get(7, 1) == c.E_INVAL
    assert c.lib.vgpu_policy_get(c.TIER_LOCAL, 1) in (c.XFER_LDST, c.XFER_TMA, c.XFER_CE)
    assert c.lib.vgpu_spill(1, 0, None, None, None, None, 0, c.CHUNK_BYTES, None, None) == c.OK  # 0 chunks: no-op
    assert c.lib.vgpu_spill(1, 0, None, None, None, None, 4, c.CHUNK_BYTES, None, None) == c.E_INVAL
    assert "cksum" in c.last_error()
    assert c.lib.vgpu_vmem_swap(None, None, 0, 0) == c.E_INVAL and "two different ranges" in c.last_error()
    assert c.lib.vgpu_vmem_evict(None, 0, 1) == c.E_INVAL and c.lib.vgpu_vmem_ensure(None, 0, 1) == c.E_INVAL


def test_header_is_valid_c_and_links_from_a_plain_c_caller(vgpu, tmp_path):
    """What a cgo build does: compile a C translation unit against include/vgpu.h and link libvgpu_dp.so."""
    exe = tmp_path / "c_abi_smoke"
    pkg = os.path.dirname(vgpu.capi.LIB_PATH)
    subprocess.check_call(["gcc", "-std=c11", "-Wall", "-Wextra", "-Werror", "-pedantic", "-I", os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "tests", "c_abi_smoke.c"), "-o", str(exe), "-L", pkg, "-lvgpu_dp", f"-Wl,-rpath,{pkg}"])
    out = subprocess.run([str(exe)], capture_output=True, text=True)
    assert out.returncode == 0 and "c_abi_smoke ok" in out.stdout, out.stdout + out.stderr
