"""CPU-only: the C-ABI library loads and exports exactly what include/*.h declare,
and the reference header compiles verbatim against it."""
import ctypes
import os
import re
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "flash_attention_hip_b200", "lib", "libflash_attn.so")
REF_HEADER = "/root/reference/flash_attn.h"

REFERENCE_SYMBOLS = ["flash_attn_fwd", "flash_attn_varlen_fwd", "flash_attn_bwd", "flash_attn_varlen_bwd",
                     "flash_attn_set_error", "flash_attn_error"]


@pytest.fixture(scope="module")
def built_lib():
    if not os.path.exists(LIB):
        from flash_attention_hip_b200 import build
        build.build()
    return LIB


def declared_symbols(header):
    src = open(os.path.join(ROOT, "include", header)).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return re.findall(r"\b(flash_attn\w*)\s*\(", src)


def exported_symbols(lib):
    out = subprocess.run(["nm", "-D", "--defined-only", lib], check=True, capture_output=True, text=True).stdout
    return {line.split()[-1] for line in out.splitlines() if " T " in line}


def test_exports_match_headers(built_lib):
    exp = exported_symbols(built_lib)
    decl = set(declared_symbols("flash_attn.h")) | set(declared_symbols("flash_attn_b200_ext.h"))
    assert set(declared_symbols("flash_attn.h")) == set(REFERENCE_SYMBOLS)
    assert decl <= exp, decl - exp
    assert exp == decl, f"undeclared exports: {exp - decl}"


def test_library_loads_and_error_string_roundtrip(built_lib):
    l = ctypes.CDLL(built_lib)
    l.flash_attn_error.restype = ctypes.c_char_p
    assert l.flash_attn_error() == b""
    l.flash_attn_set_error(b"boom")
    assert l.flash_attn_error() == b"boom"
    l.flash_attn_set_error(None)
    assert l.flash_attn_error() == b""
    l.flash_attn_b200_version.restype = ctypes.c_char_p
    assert b"sm_100a" in l.flash_attn_b200_version()


def test_validation_errors_need_no_gpu(built_lib):
    """Argument checks run before any CUDA call: false + message, never a crash."""
    from flash_attention_hip_b200 import api
    l = api.lib()
    bad = l.flash_attn_fwd(None, None, None, None, None, None, None, 1, 8, 8, 128, 128, 3, 2, 64, 64,
                           0.0, 0.125, 8.0, False, False, True, None, 0, 0, None, None)
    assert bad is False and b"multiple of num_heads_k" in l.flash_attn_error()
    bad = l.flash_attn_fwd(None, None, None, None, None, None, None, 1, 8, 8, 128, 128, 2, 2, 60, 64,
                           0.0, 0.125, 8.0, False, False, True, None, 0, 0, None, None)
    assert bad is False and b"head_size" in l.flash_attn_error()
    bad = l.flash_attn_fwd(None, None, None, None, None, None, None, 1, 8, 8, 128, 128, 2, 2, 64, 64,
                           1.5, 0.125, 8.0, False, False, True, None, 0, 0, None, None)
    assert bad is False and b"p_dropout" in l.flash_attn_error()
    bad = l.flash_attn_fwd(None, None, None, None, None, None, None, 1, 8, 8, 128, 128, 2, 2, 64, 64,
                           0.0, 0.125, 8.0, False, False, True, None, 0, 0, None, None)
    assert bad is False and b"must not be NULL" in l.flash_attn_error()
    bad = l.flash_attn_varlen_fwd(None, None, None, None, None, None, None, None, None, 1, 8, 8, 128, 128,
                                  2, 2, 64, 64, 0.0, 0.125, 8.0, False, False, True, None, 0, 0, None, None)
    assert bad is False and b"cu_seqlens" in l.flash_attn_error()
    bad = l.flash_attn_bwd(*([None] * 12), 1, 8, 8, 128, 128, 2, 2, 64, 64, 0.0, 0.125, 8.0, False, True,
                           0, None, 0, 0, None, None)
    assert bad is False and b"must not be NULL" in l.flash_attn_error()


CALLER = r"""
#include "flash_attn.h"   /* resolved by -I: the reference header or ours */
#include <stdio.h>
typedef bool (*fwd_t)(const void*, const void*, const void*, void*, void*, void*, void*, int, int, int,
    int, int, int, int, int, int, float, float, float, bool, bool, bool, hipStream_t, uint64_t, uint64_t,
    const void*, const int64_t*);
int main() {
  fwd_t f = &flash_attn_fwd;
  void* addrs[] = {(void*)f, (void*)&flash_attn_varlen_fwd, (void*)&flash_attn_bwd,
                   (void*)&flash_attn_varlen_bwd, (void*)&flash_attn_set_error, (void*)&flash_attn_error};
  flash_attn_set_error("hello");
  printf("%s %d\n", flash_attn_error(), (int)(sizeof(addrs) / sizeof(addrs[0])));
  bool ok = f(0,0,0,0,0,0,0, 1,8,8,128,128, 3,2,64,64, 0.f,1.f,1.f, false,false,true, 0, 0,0, 0,0);
  printf("%d %s\n", (int)ok, flash_attn_error());
  return 0;
}
"""


@pytest.mark.skipif(not os.path.exists(REF_HEADER), reason="reference tree not mounted")
def test_reference_header_links_unchanged(built_lib, tmp_path):
    """A TU that includes /root/reference/flash_attn.h VERBATIM links against our .so."""
    src = tmp_path / "caller.cpp"
    src.write_text(CALLER)
    exe = tmp_path / "caller"
    subprocess.run(["g++", "-std=c++17", "-I/root/reference", "-I" + os.path.join(ROOT, "compat"), str(src),
                    "-o", str(exe), built_lib, "-Wl,-rpath," + os.path.dirname(built_lib)], check=True)
    out = subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.splitlines()
    assert out[0] == "hello 6"
    assert out[1].startswith("0 flash_attn_fwd: num_heads")


def test_own_header_is_valid_c_and_links(built_lib, tmp_path):
    src = tmp_path / "caller.c"
    src.write_text(CALLER.replace("hipStream_t", "fa_stream_t"))
    exe = tmp_path / "caller_c"
    subprocess.run(["gcc", "-std=c11", "-Wall", "-Werror", "-I" + os.path.join(ROOT, "include"), str(src),
                    "-o", str(exe), built_lib, "-Wl,-rpath," + os.path.dirname(built_lib)], check=True)
    out = subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.splitlines()
    assert out[0] == "hello 6" and out[1].startswith("0 flash_attn_fwd")


def test_ring_entry_points_validate_without_gpu(built_lib):
    """flash_attn_b200_ring_* (include/flash_attn_b200_ext.h): workspace sizes are plain arithmetic, and bad arguments
    come back as false + message before any CUDA or NCCL call."""
    l = ctypes.CDLL(built_lib)
    l.flash_attn_error.restype = ctypes.c_char_p
    for name in ("fwd", "bwd"):
        f = getattr(l, f"flash_attn_b200_ring_{name}_workspace_bytes")
        f.argtypes = [ctypes.c_int] * 6
        f.restype = ctypes.c_size_t
    b, s, h, hk, d = 1, 16384, 32, 8, 128
    kv = 2 * b * s * hk * d * 2
    one = b * s * h * d * 2 + b * h * s * 4
    assert l.flash_attn_b200_ring_fwd_workspace_bytes(1, b, s, h, hk, d) == one          # no exchange buffers alone
    assert l.flash_attn_b200_ring_fwd_workspace_bytes(8, b, s, h, hk, d) == 2 * kv + 8 * one
    assert l.flash_attn_b200_ring_bwd_workspace_bytes(8, b, s, h, hk, d) > l.flash_attn_b200_ring_bwd_workspace_bytes(1, b, s, h, hk, d)
    l.flash_attn_b200_ring_comm_create.restype = ctypes.c_bool
    comm = ctypes.c_void_p()
    assert l.flash_attn_b200_ring_comm_create(None, 2, 5, ctypes.byref(comm)) is False     # rank out of range
    assert b"bad argument" in l.flash_attn_error()
    assert l.flash_attn_b200_ring_comm_create(None, 64, 0, ctypes.byref(comm)) is False    # more ranks than the ring supports
    l.flash_attn_b200_ring_fwd.restype = ctypes.c_bool
    assert l.flash_attn_b200_ring_fwd(None, *([None] * 7), 1, 128, 2, 2, 64, ctypes.c_float(0.125), True, True, None) is False
    assert b"comm is NULL" in l.flash_attn_error()
