"""The C-ABI library builds for sm_100a, loads without a GPU, exports every entry point that
include/bd_cuda.h and include/bd_api_decl.inc declare, and refuses to run without a device
(there is no CPU fallback).  CPU only."""
import ctypes
import os
import re
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def engine_lib():
    sys.path.insert(0, ROOT)
    import __graft_entry__ as ge

    if not os.path.exists("/usr/local/cuda/bin/nvcc") and not os.path.exists(ge.ENGINE_LIB):
        pytest.skip("nvcc not available and no prebuilt library")
    if os.path.exists("/usr/local/cuda/bin/nvcc"):
        ge.build_engine()
    return ge.ENGINE_LIB


def declared_functions():
    names = set()
    inc = open(os.path.join(ROOT, "include", "bd_api_decl.inc")).read()
    names.update("bd_" + m for m in re.findall(r"BD_FN\((\w+)\)\s*\(", inc))
    hdr = open(os.path.join(ROOT, "include", "bd_cuda.h")).read()
    names.update(re.findall(r"^int (bd_\w+)\(", hdr, flags=re.M))
    return sorted(names)


def test_header_declares_the_expected_surface():
    names = declared_functions()
    assert len(names) >= 30
    for must in ("bd_create", "bd_step", "bd_compute_forces", "bd_nsearch_pairs", "bd_set_stream"):
        assert must in names


def test_library_exports_every_declared_symbol(engine_lib):
    lib = ctypes.CDLL(engine_lib)
    missing = [n for n in declared_functions() if not hasattr(lib, n)]
    assert not missing, missing


def test_oracle_and_engine_share_one_api():
    """orc_* mirrors bd_* so parity tests drive both through the same calls"""
    subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "oracle")], check=True)
    orc = ctypes.CDLL(os.path.join(ROOT, "oracle", "liboracle.so"))
    inc = open(os.path.join(ROOT, "include", "bd_api_decl.inc")).read()
    for m in re.findall(r"BD_FN\((\w+)\)\s*\(", inc):
        assert hasattr(orc, "orc_" + m), m


def test_sass_is_sm_100a(engine_lib):
    cuobjdump = "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(cuobjdump):
        pytest.skip("cuobjdump not available")
    out = subprocess.run([cuobjdump, "-lelf", engine_lib], capture_output=True, text=True).stdout
    assert "sm_100a" in out


def test_no_cpu_fallback_without_a_device(engine_lib):
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    lib = ctypes.CDLL(engine_lib)
    lib.bd_last_error.restype = ctypes.c_char_p
    h = ctypes.c_void_p()
    rc = lib.bd_create(ctypes.c_int64(10), ctypes.c_int32(1), ctypes.c_int32(0), ctypes.byref(h))
    assert rc != 0 and not h
    assert b"no CUDA device" in lib.bd_last_error()


def test_product_never_links_the_oracle(engine_lib):
    out = subprocess.run(["ldd", engine_lib], capture_output=True, text=True).stdout
    assert "oracle" not in out and "micromd" not in out
    for f in os.listdir(os.path.join(ROOT, "2022a-genome-dynamics_b200", "csrc")):
        src = open(os.path.join(ROOT, "2022a-genome-dynamics_b200", "csrc", f)).read()
        assert "oracle/" not in src and "bd_oracle" not in src, f
